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
