// ga_ctx.cu -- context lifetime, parameters, grid storage, host<->device movement, measurement.
#include <cmath>
#include <new>

#include "ga_internal.cuh"

int ga_fail(ga_ctx* ctx, int code, const char* fmt, ...)
{
  if (ctx)
  {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

ga_launch_scope::ga_launch_scope(ga_ctx* ctx, int kid, int n_kernels) : c(ctx), k(kid)
{
  c->launches += n_kernels;  // kernels launched inside the scope
  c->k_launches[k]++;
  if (c->prof_on && c->prof.size() < (1u << 18))
  {
    cudaEvent_t e0;
    if (c->ev_pool.size() >= 2)
    {
      e0 = c->ev_pool.back();
      c->ev_pool.pop_back();
      e1 = c->ev_pool.back();
      c->ev_pool.pop_back();
    }
    else
    {
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
    }
    cudaEventRecord(e0, c->stream);
    c->prof.push_back({ k, e0, e1 });
  }
}

ga_launch_scope::~ga_launch_scope()
{
  if (e1) cudaEventRecord(e1, c->stream);
}

extern "C" int ga_abi_version(void) { return GA_ABI_VERSION; }

extern "C" int ga_create(int device, ga_ctx** out)
{
  if (!out) return GA_ERR_ARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n)
  {
    cudaGetLastError();
    return GA_ERR_CUDA;  // no CUDA device: there is no CPU path
  }
  ga_ctx* c = new (std::nothrow) ga_ctx();
  if (!c) return GA_ERR_NOMEM;
  c->device = device;
  if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess)
  {
    delete c;
    return GA_ERR_CUDA;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
  *out = c;
  return GA_OK;
}

static void free_grids(ga_ctx* c)
{
  void** ptrs[] = { (void**)&c->d_patch,   (void**)&c->d_safety,
                    (void**)&c->d_astar,   (void**)&c->d_fence,  (void**)&c->d_mc,        (void**)&c->d_prox,
                    (void**)&c->d_mres,    (void**)&c->d_odist,  (void**)&c->d_vdist,     (void**)&c->d_hcost[0],
                    (void**)&c->d_hcost[1], (void**)&c->d_hdist[0], (void**)&c->d_hdist[1], (void**)&c->d_hparent[0],
                    (void**)&c->d_hparent[1], (void**)&c->d_hflags[0], (void**)&c->d_hflags[1], (void**)&c->d_hpacked,
                    (void**)&c->d_hsteps[0], (void**)&c->d_hsteps[1], (void**)&c->d_hrec, (void**)&c->d_rec_ovf };
  for (void** p : ptrs)
  {
    if (*p) cudaFree(*p);
    *p = nullptr;
  }
  ga_heur_free(c);
}

extern "C" int ga_destroy(ga_ctx* c)
{
  if (!c) return GA_ERR_ARG;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  free_grids(c);
  ga_cart_free(c);
  ga_rt_free(c);
  ga_reforder_free(c);
  ga_expand_async_free(c);
  if (c->d_centers) cudaFree(c->d_centers);
  if (c->d_stage) cudaFree(c->d_stage);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->h_ll) cudaFreeHost(c->h_ll);
  if (c->h_rec_ovf) cudaFreeHost(c->h_rec_ovf);
  for (auto& r : c->prof)
  {
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  for (auto e : c->ev_pool) cudaEventDestroy(e);
  cudaStreamDestroy(c->stream);
  delete c;
  return GA_OK;
}

extern "C" const char* ga_last_error(const ga_ctx* c) { return c ? c->err : "null context"; }

static int rec_check(ga_ctx* c);

extern "C" int ga_sync(ga_ctx* c)
{
  if (!c) return GA_ERR_ARG;
  GA_CUDA(c, cudaSetDevice(c->device));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  return rec_check(c);
}

extern "C" void* ga_stream(ga_ctx* c) { return c ? (void*)c->stream : nullptr; }

int ga_stage_reserve(ga_ctx* c, size_t bytes)
{
  if (bytes <= c->h_stage_bytes) return GA_OK;
  size_t nb = bytes + bytes / 2 + 4096;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->h_stage) cudaFreeHost(c->h_stage);
  if (c->d_stage) cudaFree(c->d_stage);
  c->h_stage = nullptr;
  c->d_stage = nullptr;
  c->h_stage_bytes = c->d_stage_bytes = 0;
  GA_CUDA(c, cudaMallocHost(&c->h_stage, nb));
  GA_CUDA(c, cudaMalloc(&c->d_stage, nb));
  c->h_stage_bytes = c->d_stage_bytes = nb;
  return GA_OK;
}

// ---- parameters -------------------------------------------------------------------------------

// cv::getStructuringElement(MORPH_ELLIPSE, (2r+1)^2), OpenCV imgproc morph: row dy has the columns
// [r - dx, r + dx], dx = cvRound(r * sqrt((r^2 - dy^2) / r^2)) (round half to even).
static void build_se(int r, ga_se* se)
{
  memset(se, 0, sizeof(*se));
  se->r = r;
  const double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
  for (int i = 0; i <= 2 * r; ++i)
  {
    const int dy = i - r;
    int dx = (int)lrint(r * sqrt((double)(r * r - dy * dy) * inv_r2));
    if (dx > r) dx = r;
    int v = -1;
    for (int k = 0; k < se->n_var; ++k)
      if (se->var_w[k] == dx) v = k;
    if (v < 0)
    {
      v = se->n_var++;
      se->var_w[v] = dx;
    }
    se->row_var[i] = v;
  }
  // column view: hmax(w) = largest |dy| whose half width is >= w
  int hw_of[GA_MAX_DISK_R + 1];
  for (int d = 0; d <= r; ++d) hw_of[d] = se->var_w[se->row_var[r + d]];
  int hmax[GA_MAX_DISK_R + 1];
  for (int w = 0; w <= r; ++w)
  {
    hmax[w] = 0;
    for (int d = 0; d <= r; ++d)
      if (hw_of[d] >= w) hmax[w] = d;  // half widths are non-increasing in |dy| for an ellipse: the last hit is the largest
  }
  se->n_lvl = 0;
  for (int h = 0; h <= r; ++h)
  {
    bool used = false;
    for (int w = 0; w <= r; ++w) used |= hmax[w] == h;
    if (used) se->lvl_h[se->n_lvl++] = h;
  }
  for (int w = 0; w <= r; ++w)
    for (int l = 0; l < se->n_lvl; ++l)
      if (se->lvl_h[l] == hmax[w]) se->lvl_of_w[w] = l;
}

extern "C" int ga_set_collision_params(ga_ctx* c, const ga_collision_params* p, const int* centers)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, p && centers, GA_ERR_ARG, "null argument");
  GA_REQUIRE(c, p->disk_r_c >= 0 && p->disk_r_c <= GA_MAX_DISK_R, GA_ERR_ARG, "disk_r_c out of range [0,16]");
  GA_REQUIRE(c, p->nb_disks >= 1 && p->nb_disks <= GA_MAX_DISKS, GA_ERR_ARG, "nb_disks out of range [1,16]");
  GA_REQUIRE(c, p->yaw_res_deg >= 1 && 360 % p->yaw_res_deg == 0, GA_ERR_ARG, "yaw_res_deg must divide 360");
  GA_REQUIRE(c, p->pool_win == 4, GA_ERR_ARG, "pool_win must be 4 (GM_RES / ASTAR_RES are compile-time in the reference)");
  GA_REQUIRE(c, p->min_thresh >= 0 && p->min_thresh <= 255 && p->max_thresh >= 0 && p->max_thresh <= 255,
             GA_ERR_ARG, "thresholds must be u8");
  GA_CUDA(c, cudaSetDevice(c->device));
  c->cp = *p;
  c->nb_yaws = 360 / p->yaw_res_deg;
  // yaw_res_coll_rad_ = yaw_res_coll_ * util::TO_RAD (collision_checking.cpp:17), TO_RAD = PI / 180
  c->yaw_res_coll_rad = p->yaw_res_deg * (M_PI / 180);
  build_se(p->disk_r_c, &c->se);
  const size_t bytes = sizeof(int) * 2 * (size_t)c->nb_yaws * p->nb_disks;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (c->d_centers) cudaFree(c->d_centers);
  c->d_centers = nullptr;
  GA_CUDA(c, cudaMalloc(&c->d_centers, bytes));
  GA_CUDA(c, cudaMemcpyAsync(c->d_centers, centers, bytes, cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  c->have_cp = true;
  return GA_OK;
}

extern "C" int ga_set_expand_params(ga_ctx* c, const ga_expand_params* p)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, p, GA_ERR_ARG, "null argument");
  GA_REQUIRE(c, p->nb_steer >= 1 && p->nb_steer <= GA_MAX_STEER && (p->nb_steer & 1), GA_ERR_ARG,
             "nb_steer must be odd and <= 31");
  GA_REQUIRE(c, p->wheel_base > 0 && p->arc_l > 0 && p->yaw_res_deg > 0 && p->turn_on_point_angle_deg > 0,
             GA_ERR_ARG, "non-positive geometry");
  c->ep = *p;
  // HybridAStar::calcSteeringInputs (hybrid_a_star_lib.cpp:115-129)
  const double delta = (2 * p->max_steer) / (double)(p->nb_steer - 1);
  const int side = p->nb_steer / 2;
  for (int j = 0, i = -side; i <= side; ++i, ++j) c->steer[j] = i * delta;
  // the accumulating delta loop of setNeighbors (hybrid_a_star_lib.cpp:519-531)
  const double step = p->turn_on_point_angle_deg * (M_PI / 180);
  c->n_ra = 0;
  for (double da = -2 * M_PI + step; da <= 2 * M_PI - step; da += step)
  {
    if (da == 0) continue;
    GA_REQUIRE(c, c->n_ra < GA_MAX_RA, GA_ERR_ARG, "too many rear-axis deltas (turn angle too small)");
    c->ra_delta[c->n_ra++] = da;
  }
  c->have_ep = true;
  return GA_OK;
}

extern "C" int ga_expand_stride(const ga_ctx* c)
{
  if (!c || !c->have_ep) return GA_ERR_STATE;
  return 2 * c->ep.nb_steer + c->n_ra;
}

extern "C" int ga_set_astar_params(ga_ctx* c, const ga_astar_params* p)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, p, GA_ERR_ARG, "null argument");
  GA_REQUIRE(c, p->vor_dim >= 8 && p->vor_dim <= 200, GA_ERR_ARG, "vor_dim out of range [8,200]");
  GA_REQUIRE(c, p->dt_variant == GA_DT_FIXEDPOINT || p->dt_variant == GA_DT_IPPFLOAT, GA_ERR_ARG, "dt_variant");
  // every edge of the heuristic relaxation must cost more than nothing: with zero-cost edges g(n) == g(p) holds both
  // ways, parent pointers could form cycles and the fixed point would not be unique
  GA_REQUIRE(c, p->astar_movement_cost > 0 && p->astar_movement_cost < INFINITY, GA_ERR_ARG,
             "astar_movement_cost must be positive and finite");
  GA_REQUIRE(c, p->astar_lane_movement_cost > 0 && p->astar_lane_movement_cost < INFINITY, GA_ERR_ARG,
             "astar_lane_movement_cost must be positive and finite");
  GA_REQUIRE(c, p->astar_prox_cost >= 0 && p->unknown_cost_w >= 0, GA_ERR_ARG, "prox / unknown cost weights must be >= 0");
  c->ap = *p;
  c->have_ap = true;
  return GA_OK;
}

// ---- grid storage -----------------------------------------------------------------------------

__global__ void fill_f64_kernel(double* p, size_t n, double v)
{
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

static int alloc_grids(ga_ctx* c, int patch_dim, int astar_dim)
{
  const size_t np = (size_t)patch_dim * patch_dim;
  const size_t na = (size_t)astar_dim * astar_dim;
  c->patch_dim = patch_dim;
  c->astar_dim = astar_dim;
  {
    const size_t nv = (size_t)c->ap.vor_dim * c->ap.vor_dim;
    GA_CUDA(c, cudaMalloc(&c->d_patch, np));
    GA_CUDA(c, cudaMalloc(&c->d_safety, np));
    GA_CUDA(c, cudaMalloc(&c->d_astar, na));
    GA_CUDA(c, cudaMalloc(&c->d_fence, na));
    GA_CUDA(c, cudaMalloc(&c->d_mc, na * 8));
    GA_CUDA(c, cudaMalloc(&c->d_prox, na * 8));
    GA_CUDA(c, cudaMalloc(&c->d_mres, na * 8));
    GA_CUDA(c, cudaMalloc(&c->d_odist, nv * 4));
    GA_CUDA(c, cudaMalloc(&c->d_vdist, nv * 4));
    c->vor_alloc = c->ap.vor_dim;
    GA_CUDA(c, cudaMalloc(&c->d_hpacked, na));
    GA_CUDA(c, cudaMalloc(&c->d_hrec, na * 12));
    GA_CUDA(c, cudaMalloc(&c->d_rec_ovf, sizeof(int)));
    GA_CUDA(c, cudaMemsetAsync(c->d_rec_ovf, 0, sizeof(int), c->stream));
    if (!c->h_rec_ovf) GA_CUDA(c, cudaMallocHost(&c->h_rec_ovf, sizeof(int)));
    *c->h_rec_ovf = 0;
    c->rec_valid[0] = c->rec_valid[1] = false;
    for (int s = 0; s < 2; ++s)
    {
      GA_CUDA(c, cudaMalloc(&c->d_hcost[s], na * 8));
      GA_CUDA(c, cudaMalloc(&c->d_hdist[s], na * 8));
      GA_CUDA(c, cudaMalloc(&c->d_hparent[s], na * 4));
      GA_CUDA(c, cudaMalloc(&c->d_hflags[s], na));
      GA_CUDA(c, cudaMalloc(&c->d_hsteps[s], na * 4));
    }
  }
  return ga_heur_alloc(c);
}

extern "C" int ga_patch_reset(ga_ctx* c, int patch_dim)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->have_cp && c->have_ap, GA_ERR_STATE, "set collision and astar params first");
  GA_REQUIRE(c, patch_dim >= 8 && patch_dim <= 32768, GA_ERR_ARG, "patch_dim out of range");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int astar_dim = patch_dim / c->cp.pool_win;  // floor, a_star.cpp:30,59,74
  const size_t np = (size_t)patch_dim * patch_dim;
  const size_t na = (size_t)astar_dim * astar_dim;
  if (patch_dim != c->patch_dim || c->vor_alloc != c->ap.vor_dim)
  {
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
    free_grids(c);
    const int rc = alloc_grids(c, patch_dim, astar_dim);
    if (rc != GA_OK)
    {
      // a half-built context must not look sized: the next reset allocates from scratch
      free_grids(c);
      c->patch_dim = c->astar_dim = c->vor_alloc = 0;
      return rc;
    }
  }
  const size_t nv = (size_t)c->ap.vor_dim * c->ap.vor_dim;
  GA_CUDA(c, cudaMemsetAsync(c->d_patch, GA_SENSOR_UNKNOWN, np, c->stream));
  GA_CUDA(c, cudaMemsetAsync(c->d_safety, GA_UNKNOWN, np, c->stream));
  // AStar::reinit (a_star.cpp:76-90): astar_grid_ <- UNKNOWN, distance windows <- 0, h_prox <- 0,
  // motion_res <- motion_res_max_, movement cost <- astar_movement_cost_
  GA_CUDA(c, cudaMemsetAsync(c->d_astar, GA_UNKNOWN, na, c->stream));
  GA_CUDA(c, cudaMemsetAsync(c->d_fence, 0, na, c->stream));
  GA_CUDA(c, cudaMemsetAsync(c->d_prox, 0, na * 8, c->stream));
  c->prox_zero = true;
  GA_CUDA(c, cudaMemsetAsync(c->d_odist, 0, nv * 4, c->stream));
  GA_CUDA(c, cudaMemsetAsync(c->d_vdist, 0, nv * 4, c->stream));
  fill_f64_kernel<<<c->sm_count * 2, 256, 0, c->stream>>>(c->d_mc, na, c->ap.astar_movement_cost);
  fill_f64_kernel<<<c->sm_count * 2, 256, 0, c->stream>>>(c->d_mres, na, c->ap.motion_res_max);
  for (int s = 0; s < 2; ++s)
  {
    fill_f64_kernel<<<c->sm_count * 2, 256, 0, c->stream>>>(c->d_hcost[s], na, INFINITY);
    fill_f64_kernel<<<c->sm_count * 2, 256, 0, c->stream>>>(c->d_hdist[s], na, INFINITY);
    GA_CUDA(c, cudaMemsetAsync(c->d_hparent[s], 0xFE, na * 4, c->stream));  // 0xFEFEFEFE, rewritten by any build
    GA_CUDA(c, cudaMemsetAsync(c->d_hflags[s], 0, na, c->stream));
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

extern "C" int ga_patch_dim(const ga_ctx* c) { return c ? c->patch_dim : GA_ERR_ARG; }
extern "C" int ga_astar_dim(const ga_ctx* c) { return c ? c->astar_dim : GA_ERR_ARG; }
extern "C" int ga_cart_dim(const ga_ctx* c) { return c ? c->cart_dim : GA_ERR_ARG; }

// GA_GRID_HEUR_REC: everything the four result planes hold in 12 bytes per cell (three 32-bit words): the cost's 64
// bits, then dir:3 (index i of motion_, parent = cell - motion_[i]) | unknown:1 | straight:14 | diagonal:14 steps to the
// root.  closed <=> cost finite; the root is the closed cell with zero steps; cost_dist = straight * L1 + diagonal * L2
// (the arithmetic of finish C, two products and one sum).  A chain with 16383 or more steps of a kind does not fit:
// the kernel raises *ovf.
__global__ void heur_rec_kernel(const double* __restrict__ cost, const int32_t* __restrict__ parent, const uint8_t* __restrict__ flags,
                                const uint32_t* __restrict__ steps, uint32_t* __restrict__ out, int dim, int* __restrict__ ovf)
{
  const int n = dim * dim;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < n; c += gridDim.x * blockDim.x)
  {
    const uint8_t f = flags[c];
    double g = INFINITY;
    uint32_t w = 0;
    if (f & GA_H_CLOSED)
    {
      g = cost[c];
      const uint32_t st = steps[c];
      const int p = parent[c];
      unsigned dir = 0;
      if (p >= 0)
      {
        const int dx = c % dim - p % dim, dy = c / dim - p / dim;  // motion_[i] = (dx, dy)
        const unsigned lut[9] = { 4, 3, 6, 2, 0, 0, 5, 1, 7 };
        dir = lut[(dy + 1) * 3 + (dx + 1)];
      }
      if ((st & 16383u) == 16383u || (st >> 14) == 16383u) *ovf = 1;
      w = dir | ((f & GA_H_UNKNOWN) ? 8u : 0u) | (st << 4);
    }
    const unsigned long long gb = (unsigned long long)__double_as_longlong(g);
    out[3 * (size_t)c] = (uint32_t)gb;
    out[3 * (size_t)c + 1] = (uint32_t)(gb >> 32);
    out[3 * (size_t)c + 2] = w;
  }
}

static int launch_rec(ga_ctx* c, int slot)
{
  GA_REQUIRE(c, c->rec_valid[slot], GA_ERR_STATE,
             "GA_GRID_HEUR_REC needs a device (fixed-point) build of this slot: after a GA_HEUR_REFORDER build or an upload use the four planes");
  ga_launch_scope ls(c, GA_K_HEUR_FINISH);
  heur_rec_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(c->d_hcost[slot], c->d_hparent[slot], c->d_hflags[slot], c->d_hsteps[slot],
                                                          reinterpret_cast<uint32_t*>(c->d_hrec), c->astar_dim, c->d_rec_ovf);
  GA_CUDA(c, cudaMemcpyAsync(c->h_rec_ovf, c->d_rec_ovf, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  c->rec_check = true;
  return GA_OK;
}

// after the stream has been synchronised: did a record kernel meet a chain it cannot hold?
static int rec_check(ga_ctx* c)
{
  if (!c->rec_check) return GA_OK;
  c->rec_check = false;
  if (*c->h_rec_ovf == 0) return GA_OK;
  *c->h_rec_ovf = 0;
  cudaMemsetAsync(c->d_rec_ovf, 0, sizeof(int), c->stream);
  GA_REQUIRE(c, false, GA_ERR_STATE, "GA_GRID_HEUR_REC: a parent chain has 16383 or more steps of one kind; download the four planes instead");
  return GA_OK;
}

// GA_GRID_HEUR_PACKED: one byte per cell for consumers that only need membership and the parent chain
// (getAstarPath, findValidNeighborIndex, OUT_OF_HEURISTIC tests): bits 0-3 = index i of motion_ (a_star.hpp:50-52)
// with parent = cell - motion_[i], 8 = root, 15 = none; bit 4 = in the closed set; bit 5 = is_unknown_
__global__ void heur_pack_kernel(const int32_t* __restrict__ parent, const uint8_t* __restrict__ flags, uint8_t* __restrict__ out,
                                 int dim)
{
  const int n = dim * dim;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < n; c += gridDim.x * blockDim.x)
  {
    const uint8_t f = flags[c];
    unsigned code = 15;
    if (f & GA_H_CLOSED)
    {
      const int p = parent[c];
      if (p == -1)
        code = 8;
      else if (p >= 0)
      {
        const int dx = c % dim - p % dim, dy = c / dim - p / dim;  // motion_[i] = (dx, dy)
        // { 1, 0 }, { 0, 1 }, { -1, 0 }, { 0, -1 }, { -1, -1 }, { -1, 1 }, { 1, -1 }, { 1, 1 }
        const int key = (dy + 1) * 3 + (dx + 1);
        const unsigned lut[9] = { 4, 3, 6, 2, 15, 0, 5, 1, 7 };
        code = (dx >= -1 && dx <= 1 && dy >= -1 && dy <= 1) ? lut[key] : 15;
      }
    }
    out[c] = (uint8_t)(code | ((f & GA_H_CLOSED) ? 16u : 0u) | ((f & GA_H_UNKNOWN) ? 32u : 0u));
  }
}

static int grid_lookup(const ga_ctx* c, int id, void** p, size_t* bytes)
{
  const size_t np = (size_t)c->patch_dim * c->patch_dim;
  const size_t na = (size_t)c->astar_dim * c->astar_dim;
  const size_t nv = (size_t)c->vor_alloc * c->vor_alloc;
  switch (id)
  {
    case GA_GRID_PATCH: *p = c->d_patch; *bytes = np; return 0;
    case GA_GRID_CART: *p = c->d_cart; *bytes = (size_t)c->cart_dim * c->cart_dim; return 0;
    case GA_GRID_PMAP: *p = c->d_pmap; *bytes = c->d_pmap ? (size_t)801 * 801 : 0; return 0;
    case GA_GRID_SAFETY: *p = c->d_safety; *bytes = np; return 0;
    case GA_GRID_ASTAR: *p = c->d_astar; *bytes = na; return 0;
    case GA_GRID_GEOFENCE: *p = c->d_fence; *bytes = na; return 0;
    case GA_GRID_MOVEMENT_COST: *p = c->d_mc; *bytes = na * 8; return 0;
    case GA_GRID_H_PROX: *p = c->d_prox; *bytes = na * 8; return 0;
    case GA_GRID_MOTION_RES: *p = c->d_mres; *bytes = na * 8; return 0;
    case GA_GRID_OBSTACLE_DIST: *p = c->d_odist; *bytes = nv * 4; return 0;
    case GA_GRID_VORONOI_DIST: *p = c->d_vdist; *bytes = nv * 4; return 0;
    case GA_GRID_HEUR_COST: case GA_GRID_HEUR_COST + 1: *p = c->d_hcost[id - GA_GRID_HEUR_COST]; *bytes = na * 8; return 0;
    case GA_GRID_HEUR_DIST: case GA_GRID_HEUR_DIST + 1: *p = c->d_hdist[id - GA_GRID_HEUR_DIST]; *bytes = na * 8; return 0;
    case GA_GRID_HEUR_PARENT: case GA_GRID_HEUR_PARENT + 1: *p = c->d_hparent[id - GA_GRID_HEUR_PARENT]; *bytes = na * 4; return 0;
    case GA_GRID_HEUR_FLAGS: case GA_GRID_HEUR_FLAGS + 1: *p = c->d_hflags[id - GA_GRID_HEUR_FLAGS]; *bytes = na; return 0;
    case GA_GRID_HEUR_PACKED: case GA_GRID_HEUR_PACKED + 1: *p = c->d_hpacked; *bytes = na; return 0;
    case GA_GRID_HEUR_REC: case GA_GRID_HEUR_REC + 1: *p = c->d_hrec; *bytes = na * 12; return 0;
    default: return -1;
  }
}

extern "C" int ga_grid_bytes(const ga_ctx* c, int grid_id, size_t* bytes)
{
  if (!c || !bytes) return GA_ERR_ARG;
  void* p;
  if (c->patch_dim == 0) return GA_ERR_STATE;
  return grid_lookup(c, grid_id, &p, bytes) ? GA_ERR_ARG : GA_OK;
}

extern "C" int ga_device_ptr(ga_ctx* c, int grid_id, void** dptr)
{
  if (!c || !dptr) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 || grid_id == GA_GRID_CART || grid_id == GA_GRID_PMAP, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  size_t b;
  GA_REQUIRE(c, grid_lookup(c, grid_id, dptr, &b) == 0, GA_ERR_ARG, "unknown grid id");
  if (grid_id == GA_GRID_H_PROX) c->prox_zero = false;  // the caller may write it
  return GA_OK;
}

extern "C" int ga_upload(ga_ctx* c, int grid_id, const void* src, size_t bytes)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, src, GA_ERR_ARG, "null source");
  GA_REQUIRE(c, c->patch_dim > 0 || grid_id == GA_GRID_CART || grid_id == GA_GRID_PMAP, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  void* p;
  size_t b;
  GA_REQUIRE(c, grid_lookup(c, grid_id, &p, &b) == 0, GA_ERR_ARG, "unknown grid id");
  GA_REQUIRE(c, b == bytes, GA_ERR_ARG, "size mismatch");
  if (grid_id == GA_GRID_MOVEMENT_COST || grid_id == GA_GRID_H_PROX)
  {
    // the heuristic relaxation needs strictly positive movement costs and non-negative proximity costs (NaN is
    // kept: the reference's 0/0 of setCostmapValue, a_star.cpp:492-502)
    const double* v = static_cast<const double*>(src);
    const size_t n = bytes / sizeof(double);
    bool ok = true;
    if (grid_id == GA_GRID_MOVEMENT_COST)
      for (size_t i = 0; i < n && ok; ++i) ok = v[i] > 0 && v[i] < INFINITY;
    else
      for (size_t i = 0; i < n && ok; ++i) ok = !(v[i] < 0);
    GA_REQUIRE(c, ok, GA_ERR_ARG,
               grid_id == GA_GRID_MOVEMENT_COST ? "movement costs must be positive and finite" : "proximity costs must not be negative");
  }
  GA_CUDA(c, cudaSetDevice(c->device));
  if (grid_id == GA_GRID_H_PROX) c->prox_zero = false;
  if (grid_id >= GA_GRID_HEUR_COST && grid_id < GA_GRID_HEUR_PACKED) c->rec_valid[(grid_id - GA_GRID_HEUR_COST) & 1] = false;
  GA_CUDA(c, cudaMemcpyAsync(p, src, bytes, cudaMemcpyHostToDevice, c->stream));
  // pageable sources are staged by the driver before the call returns; pinned ones are the
  // caller's to keep alive until ga_sync
  return GA_OK;
}

extern "C" int ga_download(ga_ctx* c, int grid_id, void* dst, size_t bytes)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, dst, GA_ERR_ARG, "null destination");
  GA_REQUIRE(c, c->patch_dim > 0 || grid_id == GA_GRID_CART || grid_id == GA_GRID_PMAP, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  void* p;
  size_t b;
  GA_REQUIRE(c, grid_lookup(c, grid_id, &p, &b) == 0, GA_ERR_ARG, "unknown grid id");
  GA_REQUIRE(c, b == bytes, GA_ERR_ARG, "size mismatch");
  GA_CUDA(c, cudaSetDevice(c->device));
  if (grid_id == GA_GRID_HEUR_PACKED || grid_id == GA_GRID_HEUR_PACKED + 1)
  {
    const int slot = grid_id - GA_GRID_HEUR_PACKED;
    ga_launch_scope ls(c, GA_K_HEUR_FINISH);
    heur_pack_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(c->d_hparent[slot], c->d_hflags[slot], c->d_hpacked, c->astar_dim);
  }
  if (grid_id == GA_GRID_HEUR_REC || grid_id == GA_GRID_HEUR_REC + 1)
  {
    const int rc = launch_rec(c, grid_id - GA_GRID_HEUR_REC);
    if (rc) return rc;
  }
  GA_CUDA(c, cudaMemcpyAsync(dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  return rec_check(c);
}

// a w x h window of an astar-sized grid (row pitch = astar_dim) into a dense host array
extern "C" int ga_download_window(ga_ctx* c, int grid_id, int x0, int y0, int w, int h, void* dst)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, dst, GA_ERR_ARG, "null destination");
  GA_REQUIRE(c, c->patch_dim > 0, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  void* p;
  size_t b;
  GA_REQUIRE(c, grid_lookup(c, grid_id, &p, &b) == 0, GA_ERR_ARG, "unknown grid id");
  const size_t na = (size_t)c->astar_dim * c->astar_dim;
  GA_REQUIRE(c, na > 0 && b % na == 0 && grid_id != GA_GRID_PATCH && grid_id != GA_GRID_SAFETY && grid_id != GA_GRID_CART &&
                    grid_id != GA_GRID_PMAP && grid_id != GA_GRID_OBSTACLE_DIST && grid_id != GA_GRID_VORONOI_DIST,
             GA_ERR_ARG, "not an astar-sized grid");
  GA_REQUIRE(c, x0 >= 0 && y0 >= 0 && w > 0 && h > 0 && x0 + w <= c->astar_dim && y0 + h <= c->astar_dim, GA_ERR_ARG,
             "window outside the grid");
  const size_t el = b / na;
  GA_CUDA(c, cudaSetDevice(c->device));
  GA_CUDA(c, cudaMemcpy2DAsync(dst, (size_t)w * el, static_cast<char*>(p) + ((size_t)y0 * c->astar_dim + x0) * el,
                               (size_t)c->astar_dim * el, (size_t)w * el, (size_t)h, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  return GA_OK;
}

// ga_download without the wait: dst (pinned host memory) is valid after ga_sync.  Lets a caller with
// several contexts keep uploads, builds and downloads of different contexts in flight together.
extern "C" int ga_download_async(ga_ctx* c, int grid_id, void* dst, size_t bytes)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, dst, GA_ERR_ARG, "null destination");
  GA_REQUIRE(c, c->patch_dim > 0 || grid_id == GA_GRID_CART || grid_id == GA_GRID_PMAP, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  void* p;
  size_t b;
  GA_REQUIRE(c, grid_lookup(c, grid_id, &p, &b) == 0, GA_ERR_ARG, "unknown grid id");
  GA_REQUIRE(c, b == bytes, GA_ERR_ARG, "size mismatch");
  GA_CUDA(c, cudaSetDevice(c->device));
  if (grid_id == GA_GRID_HEUR_PACKED || grid_id == GA_GRID_HEUR_PACKED + 1)
  {
    const int slot = grid_id - GA_GRID_HEUR_PACKED;
    ga_launch_scope ls(c, GA_K_HEUR_FINISH);
    heur_pack_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(c->d_hparent[slot], c->d_hflags[slot], c->d_hpacked, c->astar_dim);
  }
  if (grid_id == GA_GRID_HEUR_REC || grid_id == GA_GRID_HEUR_REC + 1)
  {
    const int rc = launch_rec(c, grid_id - GA_GRID_HEUR_REC);  // an over-long chain is reported by the next ga_sync
    if (rc) return rc;
  }
  GA_CUDA(c, cudaMemcpyAsync(dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
  return GA_OK;
}

// ---- measurement --------------------------------------------------------------------------------

extern "C" int ga_prof_enable(ga_ctx* c, int on)
{
  if (!c) return GA_ERR_ARG;
  c->prof_on = on != 0;
  return GA_OK;
}

extern "C" int ga_prof_reset(ga_ctx* c)
{
  if (!c) return GA_ERR_ARG;
  GA_CUDA(c, cudaSetDevice(c->device));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  for (auto& r : c->prof)
  {
    c->ev_pool.push_back(r.e0);
    c->ev_pool.push_back(r.e1);
  }
  c->prof.clear();
  c->launches = 0;
  for (auto& v : c->k_launches) v = 0;
  return GA_OK;
}

extern "C" int ga_prof_get(ga_ctx* c, int k, double* ms_total, long long* launches)
{
  if (!c || k < 0 || k >= GA_K_COUNT) return GA_ERR_ARG;
  GA_CUDA(c, cudaSetDevice(c->device));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  double ms = 0;
  for (auto& r : c->prof)
  {
    if (r.k != k) continue;
    float t = 0;
    GA_CUDA(c, cudaEventElapsedTime(&t, r.e0, r.e1));
    ms += t;
  }
  if (ms_total) *ms_total = ms;
  if (launches) *launches = c->k_launches[k];
  return GA_OK;
}

extern "C" long long ga_launch_count(const ga_ctx* c) { return c ? c->launches : 0; }
