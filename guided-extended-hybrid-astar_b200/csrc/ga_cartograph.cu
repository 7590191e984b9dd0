// ga_cartograph.cu -- the accumulated ("cartographed") map of the unknown-environment loop, kept
// device-resident next to the collision checker's patch.
//
// Replaces Cartographing::resetPatch / cartograph / passLocalMap / loadPrevPatch
// (LIB/src/cartographing_lib/cartographing.cpp:7-129) and util::saveTemp / util::copyPatch2Patch
// (LIB/include/util_lib/util2.hpp:20-67).  In the reference every sensor window is merged into a host
// array cell by cell, a window of that array is memcpy'd out and copied again into
// CollisionChecker::patch_arr_; here only the sensor window crosses PCIe, the merge is one
// coalesced u8 kernel and the hand-over to the collision checker a device-to-device copy.
#include "ga_internal.cuh"

#define CART_UNKNOWN 127  // CollisionChecker::SENSOR_UNKNOWN (collision_checking.hpp:45)

// cartographing.cpp:17-96: only cells still unknown are overwritten, and only with 255 (free) or 0
// (occupied).  One thread = 4 consecutive window cells of one window row.
__global__ void __launch_bounds__(256)
cartograph_kernel(uint8_t* __restrict__ cart, int patch_dim, const uint8_t* __restrict__ win, int ox, int oy, int dim)
{
  const int quads = (dim + 3) / 4;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)quads * dim) return;
  const int y = (int)(t / quads), x0 = (int)(t % quads) * 4;
  const int py = oy + y;
  if (py < 0 || py >= patch_dim) return;
#pragma unroll
  for (int k = 0; k < 4; ++k)
  {
    const int x = x0 + k, px = ox + x;
    if (x >= dim || px < 0 || px >= patch_dim) continue;
    const uint8_t v = win[(size_t)y * dim + x];
    uint8_t* cell = cart + (size_t)py * patch_dim + px;
    if ((v == 255 || v == 0) && *cell == CART_UNKNOWN) *cell = v;
  }
}

// util2.hpp:32-67: overlap of the previous patch copied into the new one (see ga_cart_load_prev for
// the reference's exclusive end point)
__global__ void __launch_bounds__(256)
copy_patch_kernel(const uint8_t* __restrict__ prev, int prev_dim, int sp_x, int sp_y, uint8_t* __restrict__ next,
                  int next_dim, int sn_x, int sn_y, int nx, int ny)
{
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)nx * ny) return;
  const int y = (int)(t / nx), x = (int)(t % nx);
  next[(size_t)(sn_y + y) * next_dim + sn_x + x] = prev[(size_t)(sp_y + y) * prev_dim + sp_x + x];
}

void ga_cart_free(ga_ctx* c)
{
  if (c->d_cart) cudaFree(c->d_cart);
  if (c->d_cart_prev) cudaFree(c->d_cart_prev);
  c->d_cart = c->d_cart_prev = nullptr;
  c->cart_dim = c->cart_prev_dim = 0;
}

// Cartographing::resetPatch (cartographing.cpp:7-15): the old map becomes temp_patch_ (util::saveTemp
// keeps the previous temp when there is no old map), the new one is all SENSOR_UNKNOWN
extern "C" int ga_cart_reset(ga_ctx* c, int patch_dim)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, patch_dim > 0, GA_ERR_ARG, "patch_dim must be positive");
  GA_CUDA(c, cudaSetDevice(c->device));
  if (c->d_cart)
  {
    if (c->d_cart_prev)
    {
      GA_CUDA(c, cudaStreamSynchronize(c->stream));
      GA_CUDA(c, cudaFree(c->d_cart_prev));
    }
    c->d_cart_prev = c->d_cart;
    c->cart_prev_dim = c->cart_dim;
    c->d_cart = nullptr;
  }
  GA_CUDA(c, cudaMalloc(&c->d_cart, (size_t)patch_dim * patch_dim));
  c->cart_dim = patch_dim;
  GA_CUDA(c, cudaMemsetAsync(c->d_cart, CART_UNKNOWN, (size_t)patch_dim * patch_dim, c->stream));
  return GA_OK;
}

// Cartographing::cartograph (cartographing.cpp:17-96); window = dim x dim u8 in host memory
extern "C" int ga_cartograph(ga_ctx* c, const uint8_t* window, int ox, int oy, int dim)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, window && dim > 0, GA_ERR_ARG, "bad window");
  GA_REQUIRE(c, c->d_cart != nullptr, GA_ERR_STATE, "no cartographed patch (call ga_cart_reset)");
  GA_CUDA(c, cudaSetDevice(c->device));
  const size_t bytes = (size_t)dim * dim;
  const int rc = ga_stage_reserve(c, bytes);
  if (rc != GA_OK) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));  // the pinned staging buffer may still feed an earlier copy
  memcpy(c->h_stage, window, bytes);
  GA_CUDA(c, cudaMemcpyAsync(c->d_stage, c->h_stage, bytes, cudaMemcpyHostToDevice, c->stream));
  {
    ga_launch_scope ls(c, GA_K_CARTOGRAPH);
    const long long threads = (long long)((dim + 3) / 4) * dim;
    cartograph_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(c->d_cart, c->cart_dim,
                                                                                static_cast<const uint8_t*>(c->d_stage), ox, oy, dim);
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

// cartograph for a window that is already on the device: the ray-cast result of ga_raytrace
// (simulation.py:262-277 feeds gmSimulation's output straight into Cartographing::cartograph)
extern "C" int ga_cartograph_raytraced(ga_ctx* c, int ox, int oy)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->d_cart != nullptr, GA_ERR_STATE, "no cartographed patch (call ga_cart_reset)");
  GA_REQUIRE(c, c->d_pmap != nullptr, GA_ERR_STATE, "no ray-cast window (call ga_raytrace)");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int dim = 801;  // GM_DIM, gridmap_sim.hpp:17
  {
    ga_launch_scope ls(c, GA_K_CARTOGRAPH);
    const long long threads = (long long)((dim + 3) / 4) * dim;
    cartograph_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(c->d_cart, c->cart_dim, c->d_pmap, ox, oy, dim);
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

// CollisionChecker::passLocalMapData's clipping (collision_checking.cpp:287-301) for a source that
// already lives on the device: the whole window is dropped when it overhangs the collision
// checker's patch in x, rows outside are skipped
static int pass_window_d2d(ga_ctx* c, const uint8_t* src, int src_pitch, int ox, int oy, int dim)
{
  const int pd = c->patch_dim;
  if (ox < 0 || ox + dim > pd) return GA_OK;
  int r0 = 0, r1 = dim;
  if (oy < 0) r0 = -oy;
  if (oy + dim > pd) r1 = pd - oy;
  if (r1 <= r0) return GA_OK;
  GA_CUDA(c, cudaMemcpy2DAsync(c->d_patch + (size_t)(oy + r0) * pd + ox, (size_t)pd, src + (size_t)r0 * src_pitch,
                               (size_t)src_pitch, (size_t)dim, (size_t)(r1 - r0), cudaMemcpyDeviceToDevice, c->stream));
  return GA_OK;
}

// Cartographing::passLocalMap (cartographing.cpp:98-117): the window [ox, ox + dim) x [oy, oy + dim) of
// the accumulated map goes to CollisionChecker::patch_arr_.  The reference memcpy's the rows out of its
// map without a bounds check (undefined behaviour for a window that leaves the map); here such a window
// is an argument error.
extern "C" int ga_cart_pass_local_map(ga_ctx* c, int ox, int oy, int dim)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->d_cart != nullptr && c->patch_dim > 0, GA_ERR_STATE, "no cartographed patch / collision patch");
  GA_REQUIRE(c, dim > 0 && ox >= 0 && oy >= 0 && ox + dim <= c->cart_dim && oy + dim <= c->cart_dim, GA_ERR_ARG,
             "window leaves the cartographed patch");
  GA_CUDA(c, cudaSetDevice(c->device));
  return pass_window_d2d(c, c->d_cart + (size_t)oy * c->cart_dim + ox, c->cart_dim, ox, oy, dim);
}

// Cartographing::loadPrevPatch (cartographing.cpp:119-129).  Origins in gm cells: the caller applies
// (utm * grid_tf::con2gm_).toInt(), which is std::round (data_structures1.hpp:57-60).
extern "C" int ga_cart_load_prev(ga_ctx* c, int prev_ox, int prev_oy, int ox, int oy)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->d_cart != nullptr && c->patch_dim > 0, GA_ERR_STATE, "no cartographed patch / collision patch");
  GA_CUDA(c, cudaSetDevice(c->device));
  if (c->d_cart_prev)  // temp_patch_ is empty until the first resetPatch of a non-empty map: nothing to copy then
  {
    const int prev_dim = c->cart_prev_dim, next_dim = c->cart_dim;
    const int opp_x = prev_ox + prev_dim, opp_y = prev_oy + prev_dim;
    const int sp_x = ox - prev_ox > 0 ? ox - prev_ox : 0, sp_y = oy - prev_oy > 0 ? oy - prev_oy : 0;
    const int sn_x = prev_ox - ox > 0 ? prev_ox - ox : 0, sn_y = prev_oy - oy > 0 ? prev_oy - oy : 0;
    // util2.hpp:50-56: end point clamped to next_dim - 1, loop exclusive -- a clamped overlap loses
    // its last row / column, as in the reference
    const int en_x = opp_x - ox < next_dim - 1 ? opp_x - ox : next_dim - 1;
    const int en_y = opp_y - oy < next_dim - 1 ? opp_y - oy : next_dim - 1;
    const int nx = en_x - sn_x, ny = en_y - sn_y;
    if (nx > 0 && ny > 0)
    {
      ga_launch_scope ls(c, GA_K_CARTOGRAPH);
      const long long threads = (long long)nx * ny;
      copy_patch_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(c->d_cart_prev, prev_dim, sp_x, sp_y, c->d_cart,
                                                                                  next_dim, sn_x, sn_y, nx, ny);
      GA_CUDA(c, cudaGetLastError());
    }
  }
  // :127-128: the whole map goes to the collision checker
  return pass_window_d2d(c, c->d_cart, c->cart_dim, 0, 0, c->cart_dim);
}
