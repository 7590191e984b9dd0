// ga_heuristic.cu -- K5: the obstacle-aware 2D holonomic heuristic grid.
//
// Replaces AStar::calcDistanceHeuristic (a_star.cpp:110-295; verifyNode :325-343, calcIndex
// :303-316, NodeDisc data_structures2.hpp:47-67).  The reference expands an (priority, id)-ordered
// heap with hash-map open/closed sets; here the same edge relaxation
//     cand(p -> n) = ((g(p) + len_i * mc(n)) + prox(p) * w_prox) + unk(p)       (a_star.cpp:244-257)
// is iterated to its least fixed point as a tiled wavefront: one persistent cooperative kernel,
// 32x32-cell tiles (+1 halo) staged in shared memory, chaotic in-tile relaxation until the tile is
// stable, dirty-tile propagation between grid-wide barriers.  fl(a + w) is monotone in a and every
// w >= 0, so the fixed point is unique and independent of the relaxation order: results are
// bit-identical run to run and to a serial Dijkstra using the same operand order.
//
// HBM/L2 layout: inputs astar_grid_ (u8), restr_geofence_grid_ (u8), movement_cost_map_ (f64),
// h_prox_arr_ (f64); outputs cost (f64), cost_dist (f64), parent (i32), flags (u8), all dense
// astar_dim^2 row-major.  Algorithmic bytes per cell: 18 read + 21 written = 39 B.
#include <cooperative_groups.h>

#include <cmath>

#include "ga_internal.cuh"

namespace cg = cooperative_groups;

#define HT GA_HEUR_TILE
#define HS (HT + 2)          // staged edge incl. halo
#define H_THREADS (HT * HT)  // one thread per cell

struct heur_args
{
  const uint8_t* grid;
  const uint8_t* fence;
  const double* mc;
  const double* prox;
  double* g;
  int32_t* parent;
  double* dist;
  uint8_t* flags;
  int* active;   // [2][n_tiles]
  int* ctl;      // [0..1] change stamps, [2] closed-cell count, [3] outer iterations, [4..5] jump stamps
  int32_t* jump; // [2][ncell]
  int32_t* cnt;  // [2][ncell][2]  (straight, diagonal) steps to jump target
  int dim, tiles_x, n_tiles, goal;
  double w_prox, unk_w, L1, L2;
};

// predecessor offsets: n = p + motion_[i]  =>  p = n - motion_[i]   (a_star.hpp:52-54)
__constant__ int c_mdx[8] = { 1, 0, -1, 0, -1, -1, 1, 1 };
__constant__ int c_mdy[8] = { 0, 1, 0, -1, -1, 1, -1, 1 };

__global__ void heur_init_kernel(heur_args a)
{
  const size_t n = (size_t)a.dim * a.dim;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    a.g[i] = (i == (size_t)a.goal) ? 0.0 : INFINITY;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * a.n_tiles; i += gridDim.x * blockDim.x)
  {
    const int gx = a.goal % a.dim, gy = a.goal / a.dim;
    a.active[i] = (i == (gy / HT) * a.tiles_x + gx / HT) ? 1 : 0;
  }
  if (blockIdx.x == 0 && threadIdx.x < 8) a.ctl[threadIdx.x] = 0;
}

template <bool PROX, bool UNK>
__device__ __forceinline__ int relax_tile_loop(double* sg, const double* sb, const double* sc, int* srow, int me,
                                               int ly, bool blocked, double ln1, double ln2)
{
  // chaotic (in-place) relaxation until no cell of the tile improves.  srow[r] = last inner
  // iteration in which staged row r changed; a row is re-evaluated only if it or a vertical
  // neighbour changed in the previous or current iteration.
  const int lane = threadIdx.x & 31;
  int any_change = 0;
  for (int it = 1;; ++it)
  {
    int changed = 0;
    const int st = max(srow[ly], max(srow[ly + 1], srow[ly + 2]));
    if (st >= it - 1)
    {
      if (!blocked)
      {
        double best = sg[me];
        const double g0 = best;
#pragma unroll
        for (int i = 0; i < 8; ++i)
        {
          const int p = me - c_mdy[i] * HS - c_mdx[i];
          double cand = sg[p] + (i < 4 ? ln1 : ln2);
          if (PROX) cand = cand + sb[p];
          if (UNK) cand = cand + sc[p];
          if (cand < best) best = cand;
        }
        if (best < g0)
        {
          sg[me] = best;
          changed = 1;
        }
      }
      if (__any_sync(0xFFFFFFFFu, changed) && lane == 0) srow[ly + 1] = it;
    }
    const int any = __syncthreads_or(changed);
    if (!any) break;
    any_change = 1;
  }
  return any_change;
}

__global__ void __launch_bounds__(H_THREADS, 1) heur_relax_kernel(heur_args a)
{
  cg::grid_group grid = cg::this_grid();
  __shared__ double sg[HS * HS];
  __shared__ double sb[HS * HS];
  __shared__ double sc[HS * HS];
  __shared__ int srow[HS];
  __shared__ int s_mark[8];
  __shared__ int s_is_active;

  const int lx = threadIdx.x % HT, ly = threadIdx.x / HT;
  const int me = (ly + 1) * HS + lx + 1;

  for (int outer = 0;; ++outer)
  {
    const int cur = outer & 1, nxt = cur ^ 1;
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x)
    {
      __syncthreads();
      if (threadIdx.x == 0)
      {
        s_is_active = a.active[cur * a.n_tiles + tile];
        if (s_is_active) a.active[cur * a.n_tiles + tile] = 0;
      }
      if (threadIdx.x < 8) s_mark[threadIdx.x] = 0;
      if (threadIdx.x < HS) srow[threadIdx.x] = 0;
      __syncthreads();
      if (!s_is_active) continue;

      const int tx = tile % a.tiles_x, ty = tile / a.tiles_x;
      const int ox = tx * HT - 1, oy = ty * HT - 1;
      // ---- stage g, prox * w and unknown cost of the tile + halo
      int has_prox = 0;
      for (int i = threadIdx.x; i < HS * HS; i += H_THREADS)
      {
        const int x = ox + i % HS, y = oy + i / HS;
        double gv = INFINITY, bv = 0.0, cv = 0.0;
        if (x >= 0 && y >= 0 && x < a.dim && y < a.dim)
        {
          const size_t id = (size_t)y * a.dim + x;
          gv = __ldcg(a.g + id);
          bv = __ldg(a.prox + id) * a.w_prox;                   // prox_cost, a_star.cpp:251
          cv = (__ldg(a.grid + id) == GA_UNKNOWN) ? a.unk_w : 0; // unknown_cost, :253-254
          has_prox |= (bv != 0.0);
        }
        sg[i] = gv;
        sb[i] = bv;
        sc[i] = cv;
      }
      const int gx = ox + 1 + lx, gy = oy + 1 + ly;
      bool blocked = true;
      double ln1 = 0, ln2 = 0, g_in = INFINITY;
      size_t id = 0;
      if (gx < a.dim && gy < a.dim)
      {
        id = (size_t)gy * a.dim + gx;
        // verifyNode, a_star.cpp:325-343; the root keeps cost 0
        blocked = (__ldg(a.grid + id) == GA_OCC) || (__ldg(a.fence + id) == 1) || ((int)id == a.goal);
        const double m = __ldg(a.mc + id);
        ln1 = a.L1 * m;  // movement_costs = movement_distances_[i] * movement_weigth, :244-245
        ln2 = a.L2 * m;
      }
      has_prox = __syncthreads_or(has_prox);
      g_in = sg[me];

      int any;
      if (has_prox)
      {
        if (a.unk_w != 0) any = relax_tile_loop<true, true>(sg, sb, sc, srow, me, ly, blocked, ln1, ln2);
        else any = relax_tile_loop<true, false>(sg, sb, sc, srow, me, ly, blocked, ln1, ln2);
      }
      else
      {
        if (a.unk_w != 0) any = relax_tile_loop<false, true>(sg, sb, sc, srow, me, ly, blocked, ln1, ln2);
        else any = relax_tile_loop<false, false>(sg, sb, sc, srow, me, ly, blocked, ln1, ln2);
      }
      if (!any) continue;

      // ---- write back improved cells, note which borders moved
      const double g_out = sg[me];
      if (!blocked && g_out < g_in)
      {
        __stcg(a.g + id, g_out);
        const bool w = lx == 0, e = lx == HT - 1, n = ly == 0, s = ly == HT - 1;
        if (w) s_mark[0] = 1;
        if (e) s_mark[1] = 1;
        if (n) s_mark[2] = 1;
        if (s) s_mark[3] = 1;
        if (w && n) s_mark[4] = 1;
        if (e && n) s_mark[5] = 1;
        if (w && s) s_mark[6] = 1;
        if (e && s) s_mark[7] = 1;
      }
      __syncthreads();
      if (threadIdx.x < 8 && s_mark[threadIdx.x])
      {
        const int ddx[8] = { -1, 1, 0, 0, -1, 1, -1, 1 };
        const int ddy[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };
        const int nx = tx + ddx[threadIdx.x], ny = ty + ddy[threadIdx.x];
        if (nx >= 0 && ny >= 0 && nx < a.tiles_x && ny < a.tiles_x)
        {
          a.active[nxt * a.n_tiles + ny * a.tiles_x + nx] = 1;
          a.ctl[nxt] = outer + 1;  // stamp: iteration outer+1 has work
        }
      }
    }
    __threadfence();
    grid.sync();
    const int more = (__ldcg(a.ctl + nxt) == outer + 1);
    if (!more)
    {
      if (blockIdx.x == 0 && threadIdx.x == 0) a.ctl[3] = outer + 1;
      break;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// finish: parent / is_unknown / closed flags from the converged cost plane, cost_dist by pointer
// jumping over the parent forest (step counts are integers, so the result is order-free).
// ---------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(256) heur_finish_kernel(heur_args a)
{
  cg::grid_group grid = cg::this_grid();
  const int ncell = a.dim * a.dim;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  const int nthr = gridDim.x * blockDim.x;
  int32_t* jump0 = a.jump;
  int32_t* jump1 = a.jump + ncell;
  int2* cnt0 = reinterpret_cast<int2*>(a.cnt);
  int2* cnt1 = cnt0 + ncell;

  // ---- phase 1: canonical parent = smallest motion index whose candidate reproduces g(n)
  int closed = 0;
  for (int n = tid; n < ncell; n += nthr)
  {
    const double gn = __ldcg(a.g + n);
    int32_t par = -2;
    uint8_t fl = 0;
    int2 c = make_int2(0, 0);
    if (gn < INFINITY)
    {
      ++closed;
      fl = GA_H_CLOSED;
      if (n == a.goal)
      {
        par = -1;
        if (__ldg(a.grid + n) == GA_UNKNOWN) fl |= GA_H_UNKNOWN;  // goal_unknown, a_star.cpp:116
      }
      else
      {
        const int nx = n % a.dim, ny = n / a.dim;
        const double m = __ldg(a.mc + n);
        const double ln1 = a.L1 * m, ln2 = a.L2 * m;
#pragma unroll
        for (int i = 0; i < 8; ++i)
        {
          const int px = nx - c_mdx[i], py = ny - c_mdy[i];
          if (px < 0 || py < 0 || px >= a.dim || py >= a.dim) continue;
          const int q = py * a.dim + px;
          const double gq = __ldcg(a.g + q);
          if (!(gq < INFINITY)) continue;
          const bool unk = __ldg(a.grid + q) == GA_UNKNOWN;
          double cand = gq + (i < 4 ? ln1 : ln2);
          cand = cand + __ldg(a.prox + q) * a.w_prox;
          cand = cand + (unk ? a.unk_w : 0.0);
          if (cand == gn)
          {
            par = q;
            if (unk) fl |= GA_H_UNKNOWN;  // NodeDisc::is_unknown_ = unknown-ness of the parent, :253,268
            c = (i < 4) ? make_int2(1, 0) : make_int2(0, 1);
            break;
          }
        }
      }
    }
    a.parent[n] = par;
    a.flags[n] = fl;
    jump0[n] = par >= 0 ? par : -1;
    cnt0[n] = c;
  }
  // closed-cell count
  closed = __reduce_add_sync(0xFFFFFFFFu, closed);
  if ((threadIdx.x & 31) == 0 && closed) atomicAdd(a.ctl + 2, closed);

  // ---- phase 2: pointer jumping; ctl[4 + (r & 1)] = r + 1 when round r still moved a pointer
  int32_t* jc = jump0;
  int32_t* jn = jump1;
  int2* cc = cnt0;
  int2* cn = cnt1;
  __threadfence();
  grid.sync();
  for (int r = 0;; ++r)
  {
    int moved = 0;
    for (int n = tid; n < ncell; n += nthr)
    {
      const int32_t j = __ldcg(jc + n);
      int2 c = __ldcg(cc + n);
      int32_t jj = -1;
      if (j >= 0)
      {
        jj = __ldcg(jc + j);
        const int2 cj = __ldcg(cc + j);
        c.x += cj.x;
        c.y += cj.y;
        moved |= (jj >= 0);
      }
      jn[n] = jj;
      cn[n] = c;
    }
    if (__any_sync(0xFFFFFFFFu, moved) && (threadIdx.x & 31) == 0) a.ctl[4 + (r & 1)] = r + 1;
    __threadfence();
    grid.sync();
    int32_t* tj = jc; jc = jn; jn = tj;
    int2* tc = cc; cc = cn; cn = tc;
    if (__ldcg(a.ctl + 4 + (r & 1)) != r + 1) break;
  }

  // ---- phase 3: NodeDisc::cost_dist_
  for (int n = tid; n < ncell; n += nthr)
  {
    const int2 c = __ldcg(cc + n);
    const bool in = a.flags[n] & GA_H_CLOSED;
    a.dist[n] = in ? (double)c.x * a.L1 + (double)c.y * a.L2 : INFINITY;
  }
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------

void ga_heur_free(ga_ctx* c)
{
  void** ptrs[] = { (void**)&c->d_tile_active, (void**)&c->d_hctl, (void**)&c->d_jump, (void**)&c->d_cnt };
  for (void** p : ptrs)
  {
    if (*p) cudaFree(*p);
    *p = nullptr;
  }
}

int ga_heur_alloc(ga_ctx* c)
{
  const size_t na = (size_t)c->astar_dim * c->astar_dim;
  c->tiles_x = (c->astar_dim + HT - 1) / HT;
  c->n_tiles = c->tiles_x * c->tiles_x;
  GA_CUDA(c, cudaMalloc(&c->d_tile_active, sizeof(int) * 2 * (size_t)c->n_tiles));
  GA_CUDA(c, cudaMalloc(&c->d_hctl, sizeof(int) * 16));
  GA_CUDA(c, cudaMalloc(&c->d_jump, sizeof(int32_t) * 2 * na));
  GA_CUDA(c, cudaMalloc(&c->d_cnt, sizeof(int32_t) * 4 * na));
  int per_sm = 0;
  GA_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, heur_relax_kernel, H_THREADS, 0));
  GA_REQUIRE(c, per_sm >= 1, GA_ERR_CUDA, "heuristic relax kernel does not fit an SM");
  c->relax_grid = per_sm * c->sm_count;
  int per_sm_f = 0;
  GA_CUDA(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_f, heur_finish_kernel, 256, 0));
  GA_REQUIRE(c, per_sm_f >= 1, GA_ERR_CUDA, "heuristic finish kernel does not fit an SM");
  c->finish_grid = per_sm_f * c->sm_count;
  return GA_OK;
}

extern "C" int ga_build_heuristic(ga_ctx* c, int slot, int goal_x, int goal_y, int* n_closed)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_ap, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 (guidance) or 1 (path)");
  const int dim = c->astar_dim;
  GA_REQUIRE(c, goal_x >= 0 && goal_y >= 0 && goal_x < dim && goal_y < dim, GA_ERR_ARG, "goal outside the astar grid");
  GA_CUDA(c, cudaSetDevice(c->device));

  heur_args a;
  a.grid = c->d_astar;
  a.fence = c->d_fence;
  a.mc = c->d_mc;
  a.prox = c->d_prox;
  a.g = c->d_hcost[slot];
  a.parent = c->d_hparent[slot];
  a.dist = c->d_hdist[slot];
  a.flags = c->d_hflags[slot];
  a.active = c->d_tile_active;
  a.ctl = c->d_hctl;
  a.jump = c->d_jump;
  a.cnt = c->d_cnt;
  a.dim = dim;
  a.tiles_x = c->tiles_x;
  a.n_tiles = c->n_tiles;
  a.goal = goal_y * dim + goal_x;
  a.w_prox = c->ap.astar_prox_cost;
  a.unk_w = c->ap.unknown_cost_w;
  // movement_distances_ (a_star.hpp:54-57): ASTAR_RES and ASTAR_RES * sqrt(2)
  a.L1 = 0.625;
  a.L2 = 0.625 * sqrt(2.0);

  {
    ga_launch_scope ls(c, GA_K_HEUR_INIT);
    heur_init_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(a);
  }
  GA_CUDA(c, cudaGetLastError());
  void* args[] = { &a };
  {
    ga_launch_scope ls(c, GA_K_HEUR_RELAX);
    const int g = c->relax_grid < c->n_tiles ? c->relax_grid : c->n_tiles;
    GA_CUDA(c, cudaLaunchCooperativeKernel((const void*)heur_relax_kernel, dim3(g), dim3(H_THREADS), args, 0, c->stream));
  }
  {
    ga_launch_scope ls(c, GA_K_HEUR_FINISH);
    const int want = (int)(((size_t)dim * dim + 255) / 256);
    const int g = c->finish_grid < want ? c->finish_grid : want;
    GA_CUDA(c, cudaLaunchCooperativeKernel((const void*)heur_finish_kernel, dim3(g), dim3(256), args, 0, c->stream));
  }
  if (n_closed)
  {
    GA_CUDA(c, cudaMemcpyAsync(n_closed, c->d_hctl + 2, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return GA_OK;
}
