// ga_potential.cu -- K3 + K4: the two chamfer distance transforms and the Voronoi potential field
// on the VOR_DIM x VOR_DIM window around ego.
//
// Replaces AStar::calcVoronoiPotentialField (a_star.cpp:440-479), createDistanceTransform
// (:426-433, cv::distanceTransform(DIST_L2, DIST_MASK_5) of OpenCV 4.8.0) and setCostmapValue
// (:486-508).  The Fortune sweep between the two transforms
// (deps_lib/VoronoiDiagramGenerator.cpp:59-86) is serial fp64 geometry and stays on the host:
// ga_potential_sites hands it the OCC cells in cv::findNonZero order, ga_build_potential takes the
// edge end points it marks.
//
// Both OpenCV variants of the 5x5 chamfer are two-pass raster algorithms whose result depends on
// the visiting order (exactly so for the IPP float variant), so the kernel keeps that order: the
// whole window lives in shared memory (204^2 x 4 B = 166 KB), row y is owned by thread y, and the
// rows run skewed by three columns -- cell (y, x) is visited at step x + 3y, after
// (y, x-1), (y-1, x+2) and (y-2, x+1) -- 797 barrier steps per pass instead of 40 000 serial
// cells.  One CTA per transform; the two transforms of a call run side by side.
#include <cmath>

#include "ga_internal.cuh"

#define PB 2            // border rows above / below the staged window
#ifndef DT_K
#define DT_K 5          // cells per lane and step
#endif
#define PX (DT_K + 2)   // border columns on either side
// consecutive lanes' block starts are pitch - DT_K - 2 words apart: the pitch makes that odd (distinct banks)
#define DT_PITCH(V) (((((V) + 2 * PX) - DT_K) & 1) ? (V) + 2 * PX : (V) + 2 * PX + 1)
#define P_THREADS 256   // >= vor_dim

template <bool IPP>
struct dt_traits;
template <>
struct dt_traits<true>  // ippiDistanceTransform_5x5_8u32f_C1R: float32, metrics 1, 1.4, 2.1969
{
  typedef float T;
  __device__ static T hv() { return 1.0f; }
  __device__ static T dg() { return 1.4f; }
  __device__ static T lg() { return 2.1969f; }
  __device__ static T big() { return 3.4e38f; }
  __device__ static float out(T v) { return v; }
};
template <>
struct dt_traits<false>  // OpenCV distanceTransform_5x5: 16.16 fixed point, cvRound(metric * 65536)
{
  typedef unsigned int T;
  __device__ static T hv() { return 65536u; }
  __device__ static T dg() { return 91750u; }
  __device__ static T lg() { return 143976u; }
  __device__ static T big() { return 0x7fffffffu >> 2; }
  __device__ static float out(T v)
  {
    const T c = v > big() ? big() : v;
    return (float)c * (1.f / 65536.f);
  }
};

struct pot_args
{
  const uint8_t* grid;   // astar_grid_
  float* odist;          // obstacle_dist_
  float* vdist;          // voronoi_dist_ (persistent, in/out)
  double* prox;          // h_prox_arr_
  double* mres;          // motion_res_map_
  const int* vor_xy;     // end points, window coordinates
  int* flag;             // [0] = number of OCC cells in the window
  int n_vor, dim, V, ox, oy;
  double alpha, do_max, mr_min, mr_max, dv_min, dv_max;
};

template <bool IPP>
__global__ void __launch_bounds__(P_THREADS) dt_kernel(pot_args a)
{
  typedef dt_traits<IPP> TR;
  typedef typename TR::T T;
  extern __shared__ unsigned char dt_smem[];
  T* s = reinterpret_cast<T*>(dt_smem);
  const int V = a.V, S = DT_PITCH(V), R = V + 2 * PB;
  const int which = blockIdx.x;  // 0: obstacle transform, 1: Voronoi transform

  // Prologue: a warp per window row, lanes along x, the loads of DT_RU rows in flight together (row-wise coalesced
  // accesses and no division per cell: the scattered one-cell-per-thread version took a quarter of the kernel).
  const int warp = threadIdx.x >> 5, ln = threadIdx.x & 31;
  constexpr int NWARP = P_THREADS / 32, DT_RU = 4, DT_XN = (P_THREADS + 31) / 32;  // V <= P_THREADS
  // everything "far": the borders stay that way, the zero cells of the source image are written below
  for (int i = threadIdx.x; i < R * S; i += P_THREADS) s[i] = TR::big();
  __syncthreads();
  // OCC cells of the window (both CTAs look: the Voronoi one must know about the early return); the obstacle transform's
  // source image 1 - (astar_grid == OCC) is zero exactly there
  int occ = 0;
  for (int yb = warp * DT_RU; yb < V; yb += NWARP * DT_RU)
  {
    uint8_t b[DT_RU][DT_XN];
#pragma unroll
    for (int r = 0; r < DT_RU; ++r)
#pragma unroll
      for (int k = 0; k < DT_XN; ++k)
      {
        const int y = yb + r, x = ln + 32 * k;
        b[r][k] = (y < V && x < V) ? __ldg(a.grid + (size_t)(a.oy + y) * a.dim + a.ox + x) : (uint8_t)0;
      }
#pragma unroll
    for (int r = 0; r < DT_RU; ++r)
#pragma unroll
      for (int k = 0; k < DT_XN; ++k)
        if (b[r][k] == GA_OCC)
        {
          occ = 1;
          if (which == 0) s[(size_t)(yb + r + PB) * S + ln + 32 * k + PX] = (T)0;
        }
  }
  occ = __syncthreads_count(occ > 0) ? 1 : 0;
  if (which == 0 && threadIdx.x == 0) a.flag[0] = occ;
  if (which == 1)
  {
    if (!occ) return;  // a_star.cpp:462-465: no obstacle, the old field stays
    // generateVoronoi_wrapper marks the edge end points with OCC (VoronoiDiagramGenerator.cpp:83-84)
    for (int i = threadIdx.x; i < a.n_vor; i += P_THREADS)
    {
      const int x = a.vor_xy[2 * i], y = a.vor_xy[2 * i + 1];
      if (x >= 0 && y >= 0 && x < V && y < V) a.vdist[(size_t)y * V + x] = 2.0f;
    }
    __syncthreads();
    // source image 1 - (voronoi_dist_ == 2): zero at the marked cells (and at whatever held exactly 2.0 before)
    for (int yb = warp * DT_RU; yb < V; yb += NWARP * DT_RU)
    {
      float f[DT_RU][DT_XN];
#pragma unroll
      for (int r = 0; r < DT_RU; ++r)
#pragma unroll
        for (int k = 0; k < DT_XN; ++k)
        {
          const int y = yb + r, x = ln + 32 * k;
          f[r][k] = (y < V && x < V) ? __ldcg(a.vdist + (size_t)y * V + x) : 0.0f;
        }
#pragma unroll
      for (int r = 0; r < DT_RU; ++r)
#pragma unroll
        for (int k = 0; k < DT_XN; ++k)
          if (f[r][k] == 2.0f) s[(size_t)(yb + r + PB) * S + ln + 32 * k + PX] = (T)0;
    }
  }
  __syncthreads();

  // Both passes are shortest-path computations over a DAG (every cell depends on cells earlier in raster order only),
  // so ANY schedule that respects the dependencies gives the raster result bit for bit.  Rows work on blocks of DT_K
  // consecutive cells; row y's blocks are shifted two cells to the left of row y-1's, which makes everything block u of
  // row y needs -- cells up to x + 2 of row y-1, x + 1 of row y-2 -- part of block u of those rows, finished one step
  // earlier.  Row y is therefore busy during steps [y (1 + 2 / DT_K), + V / DT_K): at any step only ~32 rows have work,
  // and with DT_K = 5 the windows of rows y and y + 32 do not overlap.  So ONE WARP runs the whole pass: lane l takes
  // rows l, l + 32, l + 64 ... one after the other, the lanes meet at a __syncwarp per step -- no CTA barrier on the
  // chain (it cost more than the arithmetic: 54 % of the stall samples of the 256-thread version).
  const int lane = threadIdx.x & 31;
  const int steps = (V - 1) + (V - 1 + 2 * (V - 1)) / DT_K + 1;
  if (threadIdx.x < 32)
  {
    // ---- forward pass
    int y = lane;
    for (int t = 0; t < steps; ++t)
    {
      // branch-free step: a lane without work (row not begun, finished, or no row left) computes on clamped addresses
      // and stores nothing -- the divergent branches cost a quarter of the single warp's issue time
      if (y < V && DT_K * (t - y) - 2 * y >= V) y += 32;  // this row is finished: on to the lane's next one (a select)
      const int x0r = DT_K * (t - y) - 2 * y;
      const bool act = y < V && x0r + DT_K - 1 >= 0 && x0r < V;
      const int x0 = min(max(x0r, -(DT_K - 1)), V - 1);
      {
        T* row = s + (size_t)(min(y, V - 1) + PB) * S + PX;
        T u1[DT_K + 4], u2[DT_K + 2], own[DT_K];
#pragma unroll
        for (int j = 0; j < DT_K + 4; ++j) u1[j] = row[x0 - 2 + j - S];
#pragma unroll
        for (int j = 0; j < DT_K + 2; ++j) u2[j] = row[x0 - 1 + j - 2 * S];
#pragma unroll
        for (int j = 0; j < DT_K; ++j) own[j] = row[x0 + j];
        T left = row[x0 - 1];
        // the seven candidates from the rows above do not depend on this block's own results: all DT_K minima are
        // formed first (independent instructions), only min(., left + 1) runs as a chain; selects instead of branches
        T up[DT_K];
#pragma unroll
        for (int j = 0; j < DT_K; ++j)
        {
          const T a0 = min(u2[j] + TR::lg(), u2[j + 2] + TR::lg());        // (y-2, x-1), (y-2, x+1)
          const T a1 = min(u1[j] + TR::lg(), u1[j + 1] + TR::dg());        // (y-1, x-2), (y-1, x-1)
          const T a2 = min(u1[j + 2] + TR::hv(), u1[j + 3] + TR::dg());    // (y-1, x), (y-1, x+1)
          up[j] = min(min(a0, a1), min(a2, u1[j + 4] + TR::lg()));         // (y-1, x+2)
        }
#pragma unroll
        for (int j = 0; j < DT_K; ++j)
        {
          const int x = x0 + j;
          T t0 = min(up[j], left + TR::hv());                              // (y, x-1)
          if (IPP) t0 = min(t0, TR::big());
          const T v = own[j] != (T)0 ? t0 : own[j];
          if (act && x >= 0 && x < V) row[x] = v;
          left = v;
        }
      }
      __syncwarp();
    }
    // ---- backward pass (mirrored: lane's row V-1-y, cells from the right)
    y = lane;
    for (int t = 0; t < steps; ++t)
    {
      if (y < V && DT_K * (t - y) - 2 * y >= V) y += 32;
      const int m0r = DT_K * (t - y) - 2 * y;  // first mirrored column of the block; real column = V - 1 - m
      const bool act = y < V && m0r + DT_K - 1 >= 0 && m0r < V;
      const int m0 = min(max(m0r, -(DT_K - 1)), V - 1);
      {
        T* row = s + (size_t)(V - 1 - min(y, V - 1) + PB) * S + PX;
        const int xr = V - 1 - m0;            // real column of the block's first cell (blocks run right to left)
        T d1[DT_K + 4], d2[DT_K + 2], own[DT_K];
#pragma unroll
        for (int j = 0; j < DT_K + 4; ++j) d1[j] = row[xr + 2 - j + S];
#pragma unroll
        for (int j = 0; j < DT_K + 2; ++j) d2[j] = row[xr + 1 - j + 2 * S];
#pragma unroll
        for (int j = 0; j < DT_K; ++j) own[j] = row[xr - j];
        T right = row[xr + 1];
        T dn[DT_K];
#pragma unroll
        for (int j = 0; j < DT_K; ++j)
        {
          const T a0 = min(d2[j] + TR::lg(), d2[j + 2] + TR::lg());        // (y+2, x+1), (y+2, x-1)
          const T a1 = min(d1[j] + TR::lg(), d1[j + 1] + TR::dg());        // (y+1, x+2), (y+1, x+1)
          const T a2 = min(d1[j + 2] + TR::hv(), d1[j + 3] + TR::dg());    // (y+1, x), (y+1, x-1)
          dn[j] = min(min(a0, a1), min(a2, d1[j + 4] + TR::lg()));         // (y+1, x-2)
        }
#pragma unroll
        for (int j = 0; j < DT_K; ++j)
        {
          const int x = xr - j;
          const T t0 = min(min(own[j], dn[j]), right + TR::hv());          // (y, x+1)
          const T v = (IPP || own[j] > TR::hv()) ? t0 : own[j];
          if (act && x >= 0 && x < V) row[x] = v;
          right = v;
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  // dist_map *= grid_tf::star2con_ (a_star.cpp:432): float element times 0.625
  float* dst = which == 0 ? a.odist : a.vdist;
  for (int y = warp; y < V; y += NWARP)
    for (int x = ln; x < V; x += 32) dst[(size_t)y * V + x] = TR::out(s[(size_t)(y + PB) * S + x + PX]) * 0.625f;
}

// K4: AStar::setCostmapValue (a_star.cpp:486-508)
__global__ void __launch_bounds__(256) costmap_kernel(pot_args a)
{
  if (a.flag[0] == 0) return;  // early return of calcVoronoiPotentialField
  const int V = a.V;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V * V) return;
  const int x = i % V, y = i / V;
  const float d_o = a.odist[i];
  const float d_v = a.vdist[i];
  const size_t g = (size_t)(y + a.oy) * a.dim + x + a.ox;
  double h = 0.0;
  if ((double)d_o < a.do_max)
  {
    // std::pow(v, 2) is v * v (gcc expands pow(x, 2.0) to a multiplication at every -O level > 0)
    const double q = (double)d_o - a.do_max;
    h = (a.alpha / (a.alpha + (double)d_o)) * (double)(d_v / (d_o + d_v)) * (q * q) / (a.do_max * a.do_max);
  }
  a.prox[g] = h;
  const float lo = (float)a.dv_min, hi = (float)a.dv_max;
  const float cl = d_o < lo ? lo : (hi < d_o ? hi : d_o);  // std::clamp
  const double mr = a.mr_min + ((a.mr_max - a.mr_min) / (a.dv_max - a.dv_min)) * ((double)cl - a.dv_min);
  a.mres[g] = mr > a.mr_min ? mr : a.mr_min;
}

// OCC cells of the window in row-major (cv::findNonZero) order
__global__ void __launch_bounds__(1024) sites_kernel(const uint8_t* __restrict__ grid, int dim, int V, int ox, int oy,
                                                     int* __restrict__ xy, int cap, int* __restrict__ count)
{
  __shared__ int s_warp[32];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i0 = 0; i0 < V * V; i0 += 1024)
  {
    const int i = i0 + threadIdx.x;
    const int x = i % V, y = i / V;
    const bool hit = i < V * V && grid[(size_t)(oy + y) * dim + ox + x] == GA_OCC;
    const unsigned m = __ballot_sync(0xFFFFFFFFu, hit);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    int before = s_base;
    for (int w = 0; w < warp; ++w) before += s_warp[w];
    const int pos = before + __popc(m & ((1u << lane) - 1));
    if (hit && pos < cap)
    {
      xy[2 * pos] = x;
      xy[2 * pos + 1] = y;
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_base = pos + (hit ? 1 : 0);
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = s_base;
}

static int window_origin(const ga_ctx* c, int ego_x, int ego_y, int* ox, int* oy)
{
  // getCurrentMapOrigin (a_star.cpp:527-531): clamped at 0 only; the reference's ROI would leave
  // the matrix (OpenCV assertion) when the window overhangs the far border -- an argument error here
  const int V = c->ap.vor_dim;
  const int off = V / 2;
  *ox = ego_x - off > 0 ? ego_x - off : 0;
  *oy = ego_y - off > 0 ? ego_y - off : 0;
  return (*ox + V <= c->astar_dim && *oy + V <= c->astar_dim) ? 0 : -1;
}

extern "C" int ga_potential_sites(ga_ctx* c, int ego_x, int ego_y, int* xy, int cap, int* n_sites)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_ap, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, n_sites && cap >= 0 && (cap == 0 || xy), GA_ERR_ARG, "bad arguments");
  GA_REQUIRE(c, c->ap.vor_dim == c->vor_alloc, GA_ERR_STATE, "vor_dim changed after the last ga_patch_reset: reset the patch first");
  int ox, oy;
  GA_REQUIRE(c, window_origin(c, ego_x, ego_y, &ox, &oy) == 0, GA_ERR_ARG, "potential window leaves the astar grid");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int V = c->ap.vor_dim;
  const size_t b_xy = sizeof(int) * 2 * (size_t)V * V;
  int rc = ga_stage_reserve(c, b_xy + 64);
  if (rc) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  int* d_cnt = (int*)c->d_stage;
  int* d_xy = d_cnt + 16;
  {
    ga_launch_scope ls(c, GA_K_POTENTIAL);
    sites_kernel<<<1, 1024, 0, c->stream>>>(c->d_astar, c->astar_dim, V, ox, oy, d_xy, V * V, d_cnt);
  }
  GA_CUDA(c, cudaGetLastError());
  int* h = (int*)c->h_stage;
  GA_CUDA(c, cudaMemcpyAsync(h, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  const int n = h[0];
  *n_sites = n;
  const int take = n < cap ? n : cap;
  if (take > 0)
  {
    GA_CUDA(c, cudaMemcpyAsync(h + 16, d_xy, sizeof(int) * 2 * (size_t)take, cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
    memcpy(xy, h + 16, sizeof(int) * 2 * (size_t)take);
  }
  return GA_OK;
}

extern "C" int ga_build_potential(ga_ctx* c, int ego_x, int ego_y, const int* vor_xy, int n_vor, int* had_occ)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_ap, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, n_vor >= 0 && (n_vor == 0 || vor_xy), GA_ERR_ARG, "bad arguments");
  GA_REQUIRE(c, c->ap.vor_dim == c->vor_alloc, GA_ERR_STATE, "vor_dim changed after the last ga_patch_reset: reset the patch first");
  int ox, oy;
  GA_REQUIRE(c, window_origin(c, ego_x, ego_y, &ox, &oy) == 0, GA_ERR_ARG, "potential window leaves the astar grid");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int V = c->ap.vor_dim;
  GA_REQUIRE(c, V <= P_THREADS, GA_ERR_ARG, "vor_dim too large");
  const size_t b_xy = sizeof(int) * 2 * (size_t)n_vor;
  int rc = ga_stage_reserve(c, b_xy + 128);
  if (rc) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  int* d_flag = (int*)c->d_stage;
  int* d_xy = d_flag + 16;
  int* h = (int*)c->h_stage;
  if (n_vor)
  {
    memcpy(h + 16, vor_xy, b_xy);
    GA_CUDA(c, cudaMemcpyAsync(d_xy, h + 16, b_xy, cudaMemcpyHostToDevice, c->stream));
  }
  pot_args a;
  a.grid = c->d_astar;
  a.odist = c->d_odist;
  a.vdist = c->d_vdist;
  c->prox_zero = false;
  a.prox = c->d_prox;
  a.mres = c->d_mres;
  a.vor_xy = d_xy;
  a.flag = d_flag;
  a.n_vor = n_vor;
  a.dim = c->astar_dim;
  a.V = V;
  a.ox = ox;
  a.oy = oy;
  a.alpha = c->ap.alpha;
  a.do_max = c->ap.do_max;
  a.mr_min = c->ap.motion_res_min;
  a.mr_max = c->ap.motion_res_max;
  a.dv_min = c->ap.dist_val_min;
  a.dv_max = c->ap.dist_val_max;
  const size_t smem = sizeof(float) * (size_t)(V + 2 * PB) * DT_PITCH(V) + 64;
  {
    ga_launch_scope ls(c, GA_K_DT);
    if (c->ap.dt_variant == GA_DT_IPPFLOAT)
    {
      GA_CUDA(c, cudaFuncSetAttribute(dt_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      dt_kernel<true><<<2, P_THREADS, smem, c->stream>>>(a);
    }
    else
    {
      GA_CUDA(c, cudaFuncSetAttribute(dt_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      dt_kernel<false><<<2, P_THREADS, smem, c->stream>>>(a);
    }
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_POTENTIAL);
    costmap_kernel<<<(V * V + 255) / 256, 256, 0, c->stream>>>(a);
  }
  GA_CUDA(c, cudaGetLastError());
  if (had_occ)
  {
    GA_CUDA(c, cudaMemcpyAsync(h, d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
    *had_occ = h[0];
  }
  return GA_OK;
}
