// ga_heuristic.cu -- K5: the obstacle-aware 2D holonomic heuristic grid.
//
// Replaces AStar::calcDistanceHeuristic (a_star.cpp:110-295; verifyNode :325-343, calcIndex
// :303-316, NodeDisc data_structures2.hpp:47-67).  The reference expands an (priority, id)-ordered
// heap with hash-map open/closed sets; here the same edge relaxation
//     cand(p -> n) = ((g(p) + len_i * mc(n)) + prox(p) * w_prox) + unk(p)       (a_star.cpp:244-257)
// is iterated to its least fixed point as a tiled wavefront.  fl(a + w) is monotone in a and every
// w >= 0, so the fixed point is unique and independent of the relaxation order: results are
// bit-identical run to run and to a serial Dijkstra using the same operand order.
//
// Execution model (no grid-wide barrier anywhere in the relaxation):
//   - the astar grid is cut into 32x32-cell tiles; a tile is relaxed by one CTA: tile + 1-cell halo
//     staged in shared memory (cost plane g, prox * w, unknown cost), chaotic in-place relaxation
//     with per-row activity stamps until no cell of the tile improves, improved cells written back;
//   - tiles whose facing border (or corner) cell improved are activated; activation is dataflow: a
//     per-tile state word (idle / queued / running / running+dirty) and one device-wide ring buffer
//     of tile ids.  Persistent CTAs pop the ring, so the wavefront is balanced dynamically over the
//     148 SMs and a CTA never waits for another one;
//   - termination: a counter of queued-or-running tiles reaches zero;
//   - the same kernel serves a BATCH of independent grids (one descriptor per grid, ring entries
//     carry the grid index): many queries' wavefronts fill the SMs where one grid cannot.
//
// HBM/L2 layout: inputs astar_grid_ (u8), restr_geofence_grid_ (u8), movement_cost_map_ (f64),
// h_prox_arr_ (f64); outputs cost (f64), cost_dist (f64), parent (i32), flags (u8), all dense
// astar_dim^2 row-major.  Algorithmic bytes per cell: 18 read + 21 written = 39 B.
#include <cooperative_groups.h>

#include <cmath>
#include <cstdlib>

#include "ga_internal.cuh"

namespace cg = cooperative_groups;

#define HT GA_HEUR_TILE
#define HS (HT + 2)            // staged edge incl. halo
#define H_ROWS_PER_THREAD 2    // cells per thread in the relax kernel (rows ly and ly + HT/2)
#define H_THREADS (HT * HT / H_ROWS_PER_THREAD)
#define H_WARPS (H_THREADS / 32)

enum { T_IDLE = 0, T_QUEUED = 1, T_RUNNING = 2, T_DIRTY = 3 };

// control block words (ga_ctx::d_hctl)
enum
{
  HC_HEAD = 0,      // ring consumer cursor
  HC_TAIL = 1,      // ring producer cursor
  HC_PENDING = 2,   // tiles queued or running
  HC_CLOSED = 3,    // closed-cell count (finish kernel)
  HC_JUMP0 = 4,     // pointer-jumping round stamps
  HC_JUMP1 = 5,
  HC_VISITS = 8,    // statistics: tile visits, inner iterations, visits that improved a cell
  HC_INNER = 9,
  HC_IMPROVING = 10,
  HC_CYC = 16,      // u64 x 4: SM cycles spent by thread 0 in pop / stage / relax / write-back+activate
  HC_WORDS = 24
};

// one heuristic build = one astar grid
struct heur_grid
{
  const uint8_t* grid;
  const uint8_t* fence;
  const double* mc;
  const double* prox;
  double* g;
  int* state;  // per tile
  int dim, tiles_x, n_tiles, goal;
  double w_prox, unk_w;
};

struct heur_queue
{
  int* ring;      // entries: ((grid index << 20) | tile) + 1, 0 = empty slot
  unsigned mask;  // capacity - 1 (power of two >= total tiles)
  int* ctl;
};

struct heur_args  // finish kernel
{
  const uint8_t* grid;
  const double* mc;
  const double* prox;
  const double* g;
  int32_t* parent;
  double* dist;
  uint8_t* flags;
  int* ctl;
  int32_t* jump;  // [2][ncell]
  int32_t* cnt;   // [2][ncell][2]  (straight, diagonal) steps to jump target
  int dim, goal;
  double w_prox, unk_w, L1, L2;
};

// predecessor offsets: n = p + motion_[i]  =>  p = n - motion_[i]   (a_star.hpp:52-54)
__constant__ int c_mdx[8] = { 1, 0, -1, 0, -1, -1, 1, 1 };
__constant__ int c_mdy[8] = { 0, 1, 0, -1, -1, 1, -1, 1 };

// neighbour tile of border mark k: W, E, N, S, NW, NE, SW, SE
__constant__ int c_ndx[8] = { -1, 1, 0, 0, -1, 1, -1, 1 };
__constant__ int c_ndy[8] = { 0, 0, -1, 1, -1, -1, 1, 1 };

__device__ __forceinline__ int ld_volatile(const int* p) { return *reinterpret_cast<const volatile int*>(p); }

// The ring is a ticket queue: a consumer takes the next ticket (atomicAdd on the head cursor) and
// spins on ITS slot only, a producer takes the next position (atomicAdd on the tail cursor) and
// stores the entry there.  Idle CTAs therefore poll distinct addresses and a pushed tile is picked
// up after one L2 round trip.  Every tile is in the ring at most once (state word), the ring holds
// >= 4x the tile count, so a slot is never overwritten before its previous entry was taken.
__device__ __forceinline__ void q_push(const heur_queue& q, int entry)
{
  const unsigned pos = atomicAdd(reinterpret_cast<unsigned*>(q.ctl + HC_TAIL), 1u);
  *reinterpret_cast<volatile int*>(q.ring + (pos & q.mask)) = entry;
}

// mark tile t of grid gi as needing a (re)run; the caller's writes to g are already fenced.
// The caller has already counted the tile in HC_PENDING; returns 1 if the count was NOT consumed.
__device__ __forceinline__ int tile_activate(const heur_queue& q, int* state, int gi, int t)
{
  for (;;)
  {
    const int old = atomicCAS(state + t, T_IDLE, T_QUEUED);
    if (old == T_IDLE)
    {
      q_push(q, ((gi << 20) | t) + 1);
      return 0;
    }
    if (old == T_QUEUED || old == T_DIRTY) return 1;
    if (atomicCAS(state + t, T_RUNNING, T_DIRTY) == T_RUNNING) return 1;  // else it just finished: retry
  }
}

// blocks until this CTA's ticket is served (returns the entry) or all work is done (returns -1)
__device__ __forceinline__ int q_pop(const heur_queue& q)
{
  const unsigned h = atomicAdd(reinterpret_cast<unsigned*>(q.ctl + HC_HEAD), 1u);
  int* slot = q.ring + (h & q.mask);
  for (unsigned spin = 0;; ++spin)
  {
    const int e = ld_volatile(slot);
    if (e != 0)
    {
      *reinterpret_cast<volatile int*>(slot) = 0;
      return e;
    }
    if ((spin & 3) == 3 && ld_volatile(q.ctl + HC_PENDING) == 0) return -1;
    __nanosleep(20);
  }
}

__global__ void heur_init_kernel(heur_grid G, heur_queue q, int gi, int first)
{
  const size_t n = (size_t)G.dim * G.dim;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    G.g[i] = (i == (size_t)G.goal) ? 0.0 : INFINITY;
  const int gx = G.goal % G.dim, gy = G.goal / G.dim;
  const int goal_tile = (gy / HT) * G.tiles_x + gx / HT;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < G.n_tiles; i += gridDim.x * blockDim.x)
    G.state[i] = (i == goal_tile) ? T_QUEUED : T_IDLE;
  if (blockIdx.x == 0 && threadIdx.x == 0)
  {
    // the ring is empty between builds (every entry popped is cleared): seed it with the goal tile.
    // first == 1: this launch also resets the cursors (one batch = one ring epoch)
    if (first)
    {
      for (int k = 0; k < HC_WORDS; ++k) q.ctl[k] = 0;
    }
    __threadfence();
    atomicAdd(q.ctl + HC_PENDING, 1);
    q_push(q, ((gi << 20) | goal_tile) + 1);
  }
}

// one cell: min over the 8 predecessors of ((g(p) + len_i * mc(n)) + prox(p) * w) + unk(p), operand
// order of a_star.cpp:257.  The candidates are independent; the minimum is taken as a tree (min is
// exact, so its association does not matter).  Without prox / unknown terms every straight
// (diagonal) candidate adds the same constant, and fl(x + c) is monotone in x, so
// min_i fl(g_i + c) == fl(min_i g_i + c) bit for bit: two additions instead of eight.
// A NaN candidate (0/0 in h_prox_arr_, a_star.cpp:492-497) never wins: x < NaN, NaN < x are false.
#define MIN2(a, b) ((b) < (a) ? (b) : (a))
template <bool PROX, bool UNK>
__device__ __forceinline__ double relax_cell(const double* sg, const double* sb, const double* sc, int me, double ln1,
                                             double ln2)
{
  // staged offsets of the predecessors p = n - motion_[i], i = 0..7 (straight 0..3, diagonal 4..7)
  constexpr int off[8] = { -1, -HS, 1, HS, HS + 1, -HS + 1, HS - 1, -HS - 1 };
  if (!PROX && !UNK)
  {
    const double s01 = MIN2(sg[me + off[0]], sg[me + off[1]]), s23 = MIN2(sg[me + off[2]], sg[me + off[3]]);
    const double d01 = MIN2(sg[me + off[4]], sg[me + off[5]]), d23 = MIN2(sg[me + off[6]], sg[me + off[7]]);
    const double cs = MIN2(s01, s23) + ln1, cd = MIN2(d01, d23) + ln2;
    return MIN2(cs, cd);
  }
  double c[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
  {
    const int p = me + off[i];
    double cand = sg[p] + (i < 4 ? ln1 : ln2);
    if (PROX) cand = cand + sb[p];
    if (UNK) cand = cand + sc[p];
    c[i] = cand;
  }
  const double m01 = MIN2(c[0], c[1]), m23 = MIN2(c[2], c[3]), m45 = MIN2(c[4], c[5]), m67 = MIN2(c[6], c[7]);
  const double m0 = MIN2(m01, m23), m1 = MIN2(m45, m67);
  return MIN2(m0, m1);
}
#undef MIN2

// chaotic (in-place) relaxation of the staged tile until no cell improves.  srow[r] = last inner
// iteration in which staged row r changed; a thread's rows are re-evaluated only if one of them or
// a vertical neighbour changed in the previous or current iteration.  Returns improved ? 1 : 0.
template <bool PROX, bool UNK>
__device__ __forceinline__ int relax_tile_loop(double* sg, const double* sb, const double* sc, int* srow, int lx, int wy,
                                               const bool* blocked, const double* ln1, const double* ln2, int* iters)
{
  const int lane = threadIdx.x & 31;
  int any_change = 0;
  int it = 1;
  for (;; ++it)
  {
    int changed = 0;
    int st = 0;
#pragma unroll
    for (int r = 0; r < H_ROWS_PER_THREAD; ++r)
    {
      const int ly = wy + r * H_WARPS;
      st = max(st, max(srow[ly], max(srow[ly + 1], srow[ly + 2])));
    }
    if (st >= it - 1)
    {
      double best[H_ROWS_PER_THREAD], g0[H_ROWS_PER_THREAD];
#pragma unroll
      for (int r = 0; r < H_ROWS_PER_THREAD; ++r)
      {
        const int me = (wy + r * H_WARPS + 1) * HS + lx + 1;
        g0[r] = sg[me];
        best[r] = relax_cell<PROX, UNK>(sg, sb, sc, me, ln1[r], ln2[r]);
      }
#pragma unroll
      for (int r = 0; r < H_ROWS_PER_THREAD; ++r)
      {
        const int ly = wy + r * H_WARPS;
        const int ch = (!blocked[r] && best[r] < g0[r]) ? 1 : 0;
        if (ch) sg[(ly + 1) * HS + lx + 1] = best[r];
        if (__any_sync(0xFFFFFFFFu, ch) && lane == 0) srow[ly + 1] = it;
        changed |= ch;
      }
    }
    const int any = __syncthreads_or(changed);
    if (!any) break;
    any_change = 1;
  }
  *iters = it;
  return any_change;
}

__global__ void __launch_bounds__(H_THREADS, 2) heur_relax_kernel(heur_grid single, const heur_grid* many, heur_queue q)
{
  __shared__ double sg[HS * HS];
  __shared__ double sb[HS * HS];
  __shared__ double sc[HS * HS];
  __shared__ int srow[HS];
  __shared__ int s_mark[8];
  __shared__ int s_entry, s_cont;
  __shared__ unsigned s_key[8];  // per direction: min improved border value (float bits, rounded down)

  if (threadIdx.x == 0) s_cont = 0;
  const int lx = threadIdx.x % HT, wy = threadIdx.x / HT;
  const double L1 = 0.625, L2 = 0.625 * 1.4142135623730951;  // movement_distances_ (a_star.hpp:54-57)

  for (;;)
  {
    __syncthreads();
    long long t0 = 0, t1 = 0, t2 = 0, t3 = 0;
    if (threadIdx.x == 0)
    {
      t0 = clock64();
      s_entry = s_cont ? s_cont : q_pop(q);
      s_cont = 0;
      t1 = clock64();
    }
    __syncthreads();
    const int e = s_entry;
    if (e < 0) return;
    const int gi = (e - 1) >> 20, tile = (e - 1) & 0xFFFFF;
    // field-wise select: taking the address of the by-value descriptor would spill it to local memory
#define HG(f) (many ? many[gi].f : single.f)
    heur_grid G;
    G.grid = HG(grid);
    G.fence = HG(fence);
    G.mc = HG(mc);
    G.prox = HG(prox);
    G.g = HG(g);
    G.state = HG(state);
    G.dim = HG(dim);
    G.tiles_x = HG(tiles_x);
    G.goal = HG(goal);
    G.w_prox = HG(w_prox);
    G.unk_w = HG(unk_w);
#undef HG
    const int dim = G.dim, tiles_x = G.tiles_x;
    double* __restrict__ gp = G.g;
    if (threadIdx.x == 0)
    {
      atomicExch(G.state + tile, T_RUNNING);  // ring tiles: QUEUED -> RUNNING; continued tiles already are
      // reserve the pending counts of the (at most 8) neighbours this visit may activate NOW, so the
      // increment is performed long before any of them can finish; the unused ones are returned at
      // release.  HC_PENDING can then never touch zero while work exists.
      atomicAdd(q.ctl + HC_PENDING, 8);
      __threadfence();
    }
    if (threadIdx.x < 8)
    {
      s_mark[threadIdx.x] = 0;
      s_key[threadIdx.x] = 0xFFFFFFFFu;
    }
    if (threadIdx.x < HS) srow[threadIdx.x] = 0;
    __syncthreads();

    const int tx = tile % tiles_x, ty = tile / tiles_x;
    const int ox = tx * HT - 1, oy = ty * HT - 1;
    const double w_prox = G.w_prox, unk_w = G.unk_w;
    // ---- stage g, prox * w and unknown cost of the tile + halo: own cells first, halo ring by the
    //      first 4 * HT + 4 threads, all loads of a thread in flight together
    int has_prox = 0;
    bool blocked[H_ROWS_PER_THREAD];
    double ln1[H_ROWS_PER_THREAD], ln2[H_ROWS_PER_THREAD], g_in[H_ROWS_PER_THREAD];
    size_t id[H_ROWS_PER_THREAD];
#pragma unroll
    for (int r = 0; r < H_ROWS_PER_THREAD; ++r)
    {
      const int ly = wy + r * H_WARPS;
      const int gx = ox + 1 + lx, gy = oy + 1 + ly;
      double gv = INFINITY, bv = 0.0, cv = 0.0;
      blocked[r] = true;
      ln1[r] = ln2[r] = 0.0;
      id[r] = 0;
      if (gx < dim && gy < dim)
      {
        id[r] = (size_t)gy * dim + gx;
        gv = __ldcg(gp + id[r]);
        const uint8_t cell = __ldg(G.grid + id[r]);
        bv = __ldg(G.prox + id[r]) * w_prox;                 // prox_cost, a_star.cpp:251
        cv = (cell == GA_UNKNOWN) ? unk_w : 0.0;             // unknown_cost, :253-254
        // verifyNode, a_star.cpp:325-343; the root keeps cost 0
        blocked[r] = (cell == GA_OCC) || (__ldg(G.fence + id[r]) == 1) || ((int)id[r] == G.goal);
        const double m = __ldg(G.mc + id[r]);
        ln1[r] = L1 * m;  // movement_costs = movement_distances_[i] * movement_weigth, :244-245
        ln2[r] = L2 * m;
        has_prox |= (bv != 0.0);
      }
      const int me = (ly + 1) * HS + lx + 1;
      sg[me] = gv;
      sb[me] = bv;
      sc[me] = cv;
      g_in[r] = gv;
    }
    if (threadIdx.x < 4 * HT + 4)
    {
      // halo ring: rows 0 and HS-1 (HS cells each), then columns 0 and HS-1 of rows 1..HT
      const int k = threadIdx.x;
      int hx, hy;
      if (k < HS) { hx = k; hy = 0; }
      else if (k < 2 * HS) { hx = k - HS; hy = HS - 1; }
      else if (k < 2 * HS + HT) { hx = 0; hy = k - 2 * HS + 1; }
      else { hx = HS - 1; hy = k - 2 * HS - HT + 1; }
      const int x = ox + hx, y = oy + hy;
      double gv = INFINITY, bv = 0.0, cv = 0.0;
      if (x >= 0 && y >= 0 && x < dim && y < dim)
      {
        const size_t hid = (size_t)y * dim + x;
        gv = __ldcg(gp + hid);
        bv = __ldg(G.prox + hid) * w_prox;
        cv = (__ldg(G.grid + hid) == GA_UNKNOWN) ? unk_w : 0.0;
        has_prox |= (bv != 0.0);
      }
      sg[hy * HS + hx] = gv;
      sb[hy * HS + hx] = bv;
      sc[hy * HS + hx] = cv;
    }
    has_prox = __syncthreads_or(has_prox);
    if (threadIdx.x == 0) t2 = clock64();

    int any, iters = 0;
    if (has_prox)
    {
      if (unk_w != 0) any = relax_tile_loop<true, true>(sg, sb, sc, srow, lx, wy, blocked, ln1, ln2, &iters);
      else any = relax_tile_loop<true, false>(sg, sb, sc, srow, lx, wy, blocked, ln1, ln2, &iters);
    }
    else
    {
      if (unk_w != 0) any = relax_tile_loop<false, true>(sg, sb, sc, srow, lx, wy, blocked, ln1, ln2, &iters);
      else any = relax_tile_loop<false, false>(sg, sb, sc, srow, lx, wy, blocked, ln1, ln2, &iters);
    }

    if (threadIdx.x == 0) t3 = clock64();
    // ---- write back improved cells, note which borders moved
    if (any)
    {
#pragma unroll
      for (int r = 0; r < H_ROWS_PER_THREAD; ++r)
      {
        const int ly = wy + r * H_WARPS;
        const double g_out = sg[(ly + 1) * HS + lx + 1];
        if (!blocked[r] && g_out < g_in[r])
        {
          __stcg(gp + id[r], g_out);
          const bool w = lx == 0, ee = lx == HT - 1, n = ly == 0, s = ly == HT - 1;
          if (w | ee | n | s)
          {
            const unsigned kb = __float_as_uint(__double2float_rd(g_out));
            if (w) { s_mark[0] = 1; atomicMin(s_key + 0, kb); }
            if (ee) { s_mark[1] = 1; atomicMin(s_key + 1, kb); }
            if (n) { s_mark[2] = 1; atomicMin(s_key + 2, kb); }
            if (s) { s_mark[3] = 1; atomicMin(s_key + 3, kb); }
            if (w && n) { s_mark[4] = 1; atomicMin(s_key + 4, kb); }
            if (ee && n) { s_mark[5] = 1; atomicMin(s_key + 5, kb); }
            if (w && s) { s_mark[6] = 1; atomicMin(s_key + 6, kb); }
            if (ee && s) { s_mark[7] = 1; atomicMin(s_key + 7, kb); }
          }
        }
      }
      __threadfence();
    }
    __syncthreads();
    if (threadIdx.x < 32)
    {
      // ---- activate the neighbours whose halo changed.  The neighbour with the smallest improved
      //      border value (the direction the wavefront moves) is continued by THIS CTA when it is
      //      idle -- no ring hand-off on the critical chain -- the others go to the ring.
      const int k = threadIdx.x;
      int nt = -1;
      unsigned key = 0xFFFFFFFFu;
      if (k < 8 && s_mark[k])
      {
        const int nx = tx + c_ndx[k], ny = ty + c_ndy[k];
        if (nx >= 0 && ny >= 0 && nx < tiles_x && ny < tiles_x)
        {
          nt = ny * tiles_x + nx;
          key = ((unsigned)s_key[k] & ~7u) | (unsigned)k;
        }
      }
      const unsigned want = __ballot_sync(0xFFFFFFFFu, nt >= 0);
      const unsigned best_key = __reduce_min_sync(0xFFFFFFFFu, key);
      const int n_act = __popc(want);
      int unused = 0, cont = 0;
      if (nt >= 0)
      {
        if (key == best_key && atomicCAS(G.state + nt, T_IDLE, T_RUNNING) == T_IDLE)
          cont = ((gi << 20) | nt) + 1;
        else
          unused = tile_activate(q, G.state, gi, nt);
      }
      const int n_unused = __reduce_add_sync(0xFFFFFFFFu, unused);
      const int cont_entry = __reduce_max_sync(0xFFFFFFFFu, cont);
      if (k == 0)
      {
        atomicAdd(q.ctl + HC_VISITS, 1);
        atomicAdd(q.ctl + HC_INNER, iters);
        if (any) atomicAdd(q.ctl + HC_IMPROVING, 1);
        // release the tile: re-queue it if a neighbour wrote into our halo meanwhile
        int dec = (8 - n_act) + n_unused + 1;
        if (atomicCAS(G.state + tile, T_RUNNING, T_IDLE) == T_DIRTY)
        {
          atomicExch(G.state + tile, T_QUEUED);
          q_push(q, e);
          dec -= 1;
        }
        if (dec) atomicSub(q.ctl + HC_PENDING, dec);
        s_cont = cont_entry;
        unsigned long long* cyc = reinterpret_cast<unsigned long long*>(q.ctl + HC_CYC);
        const long long t4 = clock64();
        atomicAdd(cyc + 0, (unsigned long long)(t1 - t0));
        atomicAdd(cyc + 1, (unsigned long long)(t2 - t1));
        atomicAdd(cyc + 2, (unsigned long long)(t3 - t2));
        atomicAdd(cyc + 3, (unsigned long long)(t4 - t3));
      }
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
  if ((threadIdx.x & 31) == 0 && closed) atomicAdd(a.ctl + HC_CLOSED, closed);

  // ---- phase 2: pointer jumping; ctl[HC_JUMP0 + (r & 1)] = r + 1 when round r still moved a pointer
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
    if (__any_sync(0xFFFFFFFFu, moved) && (threadIdx.x & 31) == 0) a.ctl[HC_JUMP0 + (r & 1)] = r + 1;
    __threadfence();
    grid.sync();
    int32_t* tj = jc; jc = jn; jn = tj;
    int2* tc = cc; cc = cn; cn = tc;
    if (__ldcg(a.ctl + HC_JUMP0 + (r & 1)) != r + 1) break;
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
  void** ptrs[] = { (void**)&c->d_tile_active, (void**)&c->d_hctl, (void**)&c->d_jump, (void**)&c->d_cnt,
                    (void**)&c->d_hring };
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
  GA_REQUIRE(c, c->n_tiles < (1 << 20), GA_ERR_ARG, "astar grid too large for the tile ring encoding");
  unsigned cap = 64;
  while (cap < 4u * (unsigned)c->n_tiles + 1024u) cap <<= 1;  // + one ticket per persistent CTA
  c->hring_cap = (int)cap;
  GA_CUDA(c, cudaMalloc(&c->d_tile_active, sizeof(int) * (size_t)c->n_tiles));  // per-tile state words
  GA_CUDA(c, cudaMalloc(&c->d_hring, sizeof(int) * (size_t)cap));
  GA_CUDA(c, cudaMemsetAsync(c->d_hring, 0, sizeof(int) * (size_t)cap, c->stream));
  GA_CUDA(c, cudaMalloc(&c->d_hctl, sizeof(int) * HC_WORDS));
  GA_CUDA(c, cudaMemsetAsync(c->d_hctl, 0, sizeof(int) * HC_WORDS, c->stream));
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

static heur_grid make_grid_desc(ga_ctx* c, int slot, int goal_x, int goal_y)
{
  heur_grid G;
  G.grid = c->d_astar;
  G.fence = c->d_fence;
  G.mc = c->d_mc;
  G.prox = c->d_prox;
  G.g = c->d_hcost[slot];
  G.state = c->d_tile_active;
  G.dim = c->astar_dim;
  G.tiles_x = c->tiles_x;
  G.n_tiles = c->n_tiles;
  G.goal = goal_y * c->astar_dim + goal_x;
  G.w_prox = c->ap.astar_prox_cost;
  G.unk_w = c->ap.unknown_cost_w;
  return G;
}

extern "C" int ga_build_heuristic(ga_ctx* c, int slot, int goal_x, int goal_y, int* n_closed)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_ap, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 (guidance) or 1 (path)");
  const int dim = c->astar_dim;
  GA_REQUIRE(c, goal_x >= 0 && goal_y >= 0 && goal_x < dim && goal_y < dim, GA_ERR_ARG, "goal outside the astar grid");
  GA_CUDA(c, cudaSetDevice(c->device));

  const heur_grid G = make_grid_desc(c, slot, goal_x, goal_y);
  heur_queue q;
  q.ring = c->d_hring;
  q.mask = (unsigned)c->hring_cap - 1;
  q.ctl = c->d_hctl;
  {
    ga_launch_scope ls(c, GA_K_HEUR_INIT);
    heur_init_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(G, q, 0, 1);
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_HEUR_RELAX);
    const int g = c->relax_grid < c->n_tiles ? c->relax_grid : c->n_tiles;
    heur_relax_kernel<<<g, H_THREADS, 0, c->stream>>>(G, nullptr, q);
  }
  GA_CUDA(c, cudaGetLastError());

  heur_args a;
  a.grid = c->d_astar;
  a.mc = c->d_mc;
  a.prox = c->d_prox;
  a.g = c->d_hcost[slot];
  a.parent = c->d_hparent[slot];
  a.dist = c->d_hdist[slot];
  a.flags = c->d_hflags[slot];
  a.ctl = c->d_hctl;
  a.jump = c->d_jump;
  a.cnt = c->d_cnt;
  a.dim = dim;
  a.goal = G.goal;
  a.w_prox = G.w_prox;
  a.unk_w = G.unk_w;
  // movement_distances_ (a_star.hpp:54-57): ASTAR_RES and ASTAR_RES * sqrt(2)
  a.L1 = 0.625;
  a.L2 = 0.625 * sqrt(2.0);
  void* args[] = { &a };
  {
    ga_launch_scope ls(c, GA_K_HEUR_FINISH);
    const int want = (int)(((size_t)dim * dim + 255) / 256);
    const int g = c->finish_grid < want ? c->finish_grid : want;
    GA_CUDA(c, cudaLaunchCooperativeKernel((const void*)heur_finish_kernel, dim3(g), dim3(256), args, 0, c->stream));
  }
  if (n_closed)
  {
    GA_CUDA(c, cudaMemcpyAsync(n_closed, c->d_hctl + HC_CLOSED, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return GA_OK;
}

// tile visits, inner relaxation iterations and improving visits of the last heuristic build
extern "C" int ga_heur_stats(ga_ctx* c, long long* out7)
{
  if (!c || !out7) return GA_ERR_ARG;
  GA_REQUIRE(c, c->d_hctl != nullptr, GA_ERR_STATE, "no patch");
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  int w[HC_WORDS];
  GA_CUDA(c, cudaMemcpy(w, c->d_hctl, sizeof w, cudaMemcpyDeviceToHost));
  out7[0] = w[HC_VISITS];
  out7[1] = w[HC_INNER];
  out7[2] = w[HC_IMPROVING];
  memcpy(out7 + 3, w + HC_CYC, 4 * sizeof(long long));
  return GA_OK;
}
