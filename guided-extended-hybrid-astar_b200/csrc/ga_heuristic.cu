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
#include <cmath>
#include <cstdlib>

#include "ga_internal.cuh"

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
  int edge;  // owned cells per tile edge (HT for the Jacobi kernel, SW_O for the sweep kernel)
  double w_prox, unk_w;
};

struct heur_queue
{
  int* ring;      // entries: ((grid index << 20) | tile) + 1, 0 = empty slot
  unsigned mask;  // capacity - 1 (power of two >= total tiles)
  int* ctl;
  unsigned poll_ns;  // back-off of an idle worker between two looks at its ring slot
  unsigned early;    // sweep kernel: publish once after this many changing sweeps (0 = only when converged)
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
  unsigned* scratch;  // [ncell] finish A: block-local terminal (10 bits) | straight steps (10) | diagonal steps (10)
  int2* tot;          // [ncell] finish B: (straight, diagonal) steps of a terminal cell to the root
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
    __nanosleep(q.poll_ns);
  }
}

// warp-wide variant: every lane executes the same control flow (lane 0 takes the ticket, all lanes
// poll the same word -> one request), so the warp stays converged and the sweeps' shuffles compile
// to plain SHFL
__device__ __forceinline__ int q_pop_warp(const heur_queue& q, const int lane)
{
  unsigned h = 0;
  if (lane == 0) h = atomicAdd(reinterpret_cast<unsigned*>(q.ctl + HC_HEAD), 1u);
  h = __shfl_sync(0xFFFFFFFFu, h, 0);
  int* slot = q.ring + (h & q.mask);
  int e;
  for (unsigned spin = 0;; ++spin)
  {
    e = ld_volatile(slot);
    if (e != 0) break;
    if ((spin & 3) == 3 && ld_volatile(q.ctl + HC_PENDING) == 0)
    {
      e = -1;
      break;
    }
    __nanosleep(q.poll_ns);
  }
  __syncwarp();
  if (e > 0 && lane == 0) *reinterpret_cast<volatile int*>(slot) = 0;
  return e;
}

__global__ void heur_init_kernel(heur_grid G, heur_queue q, int gi, int first)
{
  const size_t n = (size_t)G.dim * G.dim;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    G.g[i] = (i == (size_t)G.goal) ? 0.0 : INFINITY;
  const int gx = G.goal % G.dim, gy = G.goal / G.dim;
  const int goal_tile = (gy / G.edge) * G.tiles_x + gx / G.edge;
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
// Sweep relaxation: one WARP per tile, Gauss-Seidel line sweeps.
//
// A tile is a 32x32 staged block = 30x30 owned cells + a 1-cell halo ring (halo cells are sources:
// read, never updated).  The block lives in the warp's private shared-memory slice (pitch 33
// doubles: row- and column-wise lane access are both conflict-free).  One sweep visits the 30 owned
// lines in order; lane l updates cell l of the line from the three cells of the PREVIOUS line
// (lanes l-1, l, l+1: their freshly updated values arrive by warp shuffle), so a single sweep
// carries a value across the whole tile along any path inside its 3-move cone ({S,SE,SW} for the
// downward sweep, ...).  The four sweeps (down, right, up, left) cover all eight moves; the tile is
// converged when three sweeps in a row after the last changing one (or four from the start) change
// nothing (a sweep is idempotent).  Compared with the Jacobi iteration of heur_relax_kernel this
// moves a front 30 cells per ~1 us with 32 threads instead of 1 cell per 0.46 us with 512.
// Warps are independent workers on the ticket ring (no __syncthreads anywhere).
// ---------------------------------------------------------------------------------------------
#define SW_O 30                    // owned cells per tile edge
#define SW_P 33                    // shared-memory pitch (doubles)
#define SW_WARPS 6                 // workers per CTA (one CTA per SM: 6 x 33.5 KB of shared memory)
#define SW_FULL 0xFFFFFFFFu

struct sweep_tile
{
  double g[32 * SW_P];   // cost plane, relaxed in place
  double w[32 * SW_P];   // value last read from / written to global memory
  double mc[32 * SW_P];  // movement_cost_map_ of the cell (edge length factor of the DESTINATION); +inf = blocked
  double pb[32 * SW_P];  // h_prox_arr_ * astar_prox_cost_ of the cell (term of the PREDECESSOR)
  unsigned urow[32], ucol[32];  // UNKNOWN masks: urow[y] bit x / ucol[x] bit y
};

// NaN-safe minimum (a NaN candidate, 0/0 in h_prox_arr_, never wins -- a_star.cpp:492-497)
__device__ __forceinline__ double dmin(double a, double b) { return fmin(a, b); }

// One line step of a sweep: the new value of cell `lane` of line s from the three cells of line sp
// (a = this lane's value there; the neighbours' arrive by shuffle).  ORIENT 0: lane = x, lines are
// rows; ORIENT 1: lane = y, lines are columns.  Blocked cells (OCC, geofence, root, off-grid, halo)
// carry mc = +inf, so their candidates are +inf / NaN and they keep their value; lanes 0 / 31 are
// halo cells, so what their missing neighbour "sends" does not matter.
template <int ORIENT, bool PROX, bool UNK>
__device__ __forceinline__ double sweep_step(const sweep_tile& S, const int lane, const int sp, const int s, const double a,
                                             const double unk_w, const double L1, const double L2, double* old_out)
{
  constexpr int dl = ORIENT == 0 ? 1 : SW_P;  // index stride of the lane
  constexpr int ds = ORIENT == 0 ? SW_P : 1;  // index stride of the line
  const int idx = s * ds + lane * dl;
  const double al = __shfl_up_sync(SW_FULL, a, 1), ar = __shfl_down_sync(SW_FULL, a, 1);
  const double m = S.mc[idx];
  const double ln1 = L1 * m, ln2 = L2 * m;  // movement_distances_[i] * movement_weigth, a_star.cpp:244-245
  const double old = S.g[idx];
  *old_out = old;
  if (!PROX && !UNK)
  {
    // equal addends: min_i fl(g_i + c) == fl(min_i g_i + c); no NaN on this path
    const double c1 = a + ln1;
    const double m0 = c1 < old ? c1 : old;  // overlaps the shuffle latency
    const double m2 = ar < al ? ar : al;
    const double c2 = m2 + ln2;
    return c2 < m0 ? c2 : m0;
  }
  const int lm = lane > 0 ? lane - 1 : 0, lp = lane < 31 ? lane + 1 : 31;
  double c0 = a + ln1, c1 = al + ln2, c2 = ar + ln2;
  if (PROX)
  {
    const int pi = sp * ds;
    c0 = c0 + S.pb[pi + lane * dl];
    c1 = c1 + S.pb[pi + lm * dl];
    c2 = c2 + S.pb[pi + lp * dl];
  }
  if (UNK)
  {
    const unsigned um = (ORIENT == 0 ? S.urow : S.ucol)[sp];
    c0 = c0 + (((um >> lane) & 1u) ? unk_w : 0.0);
    c1 = c1 + (((um >> lm) & 1u) ? unk_w : 0.0);
    c2 = c2 + (((um >> lp) & 1u) ? unk_w : 0.0);
  }
  return dmin(dmin(old, c0), dmin(c1, c2));
}

// one sweep over the 30 owned lines; dir = +1 / -1: line order.  Returns 1 if this lane improved a cell.
template <int ORIENT, bool PROX, bool UNK>
__device__ __forceinline__ int sweep_lines(sweep_tile& S, const int lane, const int dir, const double unk_w,
                                           const double L1, const double L2)
{
  constexpr int dl = ORIENT == 0 ? 1 : SW_P, ds = ORIENT == 0 ? SW_P : 1;
  int s = dir > 0 ? 0 : 31;
  double a = S.g[s * ds + lane * dl];
  int changed = 0;
#pragma unroll 5
  for (int k = 0; k < SW_O; ++k)
  {
    const int sp = s;
    s += dir;
    double old;
    const double nv = sweep_step<ORIENT, PROX, UNK>(S, lane, sp, s, a, unk_w, L1, L2, &old);
    const bool imp = nv < old;
    if (imp) S.g[s * ds + lane * dl] = nv;
    changed |= imp;
    a = nv;
  }
  return changed;
}

// can the halo improve a border cell?  (first line step of each of the four sweeps, nothing stored)
template <bool PROX, bool UNK>
__device__ __forceinline__ int halo_improves(const sweep_tile& S, const int lane, const double unk_w, const double L1,
                                             const double L2)
{
  double old;
  int imp = 0;
  imp |= sweep_step<0, PROX, UNK>(S, lane, 0, 1, S.g[lane], unk_w, L1, L2, &old) < old;
  imp |= sweep_step<0, PROX, UNK>(S, lane, 31, 30, S.g[31 * SW_P + lane], unk_w, L1, L2, &old) < old;
  imp |= sweep_step<1, PROX, UNK>(S, lane, 0, 1, S.g[lane * SW_P], unk_w, L1, L2, &old) < old;
  imp |= sweep_step<1, PROX, UNK>(S, lane, 31, 30, S.g[lane * SW_P + 31], unk_w, L1, L2, &old) < old;
  return __any_sync(SW_FULL, imp);
}

template <bool PROX, bool UNK>
__device__ __forceinline__ int sweep_once(sweep_tile& S, const int lane, const int d, const double unk_w, const double L1,
                                          const double L2)
{
  int ch;
  if (d == 0) ch = sweep_lines<0, PROX, UNK>(S, lane, +1, unk_w, L1, L2);
  else if (d == 1) ch = sweep_lines<1, PROX, UNK>(S, lane, +1, unk_w, L1, L2);
  else if (d == 2) ch = sweep_lines<0, PROX, UNK>(S, lane, -1, unk_w, L1, L2);
  else ch = sweep_lines<1, PROX, UNK>(S, lane, -1, unk_w, L1, L2);
  ch = __any_sync(SW_FULL, ch);
  __syncwarp();
  return ch;
}

struct sweep_visit  // what a visit needs to publish its results
{
  double* gp;
  int* state;
  int dim, tiles_x, gi, tx, ty, ox, oy;
};

// write back the owned cells that improved since the last publication, activate the neighbour tiles
// whose halo they are.  FINAL: the neighbour in front (smallest improved border value) is taken
// over by this worker when it is idle -- no ring hand-off on the critical chain; returns its entry.
// n_act / n_unused: pending-count bookkeeping (reserved at the start of the visit).
template <bool FINAL>
__device__ __forceinline__ int sweep_publish(sweep_tile& S, const sweep_visit& V, const heur_queue& q, const int lane,
                                             int* n_act, int* n_unused)
{
  unsigned col_key = 0xFFFFFFFFu, kn = 0xFFFFFFFFu, ks = 0xFFFFFFFFu;
  const bool own_x = lane >= 1 && lane <= SW_O;
  const int gx = V.ox + lane;
  double* grow = V.gp + (size_t)(V.oy + 1) * V.dim + gx;
#pragma unroll 10
  for (int r = 1; r <= SW_O; ++r, grow += V.dim)
  {
    const int idx = r * SW_P + lane;
    const double v = S.g[idx];
    const bool st = own_x & (v < S.w[idx]);
    const unsigned kb = st ? __float_as_uint(__double2float_rd(v)) : 0xFFFFFFFFu;
    if (st) __stcg(grow, v);
    if (st) S.w[idx] = v;
    col_key = min(col_key, kb);
    if (r == 1) kn = kb;
    if (r == SW_O) ks = kb;
  }
  // per direction (W, E, N, S, NW, NE, SW, SE) the smallest improved border value
  unsigned kk = 0xFFFFFFFFu;
  {
    const unsigned k_w = __shfl_sync(SW_FULL, col_key, 1), k_e = __shfl_sync(SW_FULL, col_key, SW_O);
    const unsigned k_n = __reduce_min_sync(SW_FULL, kn), k_s = __reduce_min_sync(SW_FULL, ks);
    const unsigned k_nw = __shfl_sync(SW_FULL, kn, 1), k_ne = __shfl_sync(SW_FULL, kn, SW_O);
    const unsigned k_sw = __shfl_sync(SW_FULL, ks, 1), k_se = __shfl_sync(SW_FULL, ks, SW_O);
    kk = lane == 0 ? k_w : lane == 1 ? k_e : lane == 2 ? k_n : lane == 3 ? k_s : lane == 4 ? k_nw : lane == 5 ? k_ne
       : lane == 6 ? k_sw : lane == 7 ? k_se : 0xFFFFFFFFu;
  }
  // the stores above are visible before any neighbour is told to look and before the tile is released
  if (__any_sync(SW_FULL, col_key != 0xFFFFFFFFu)) __threadfence();
  __syncwarp();
  if (!__any_sync(SW_FULL, kk != 0xFFFFFFFFu)) return 0;
  int nt = -1;
  unsigned mykey = 0xFFFFFFFFu;
  if (kk != 0xFFFFFFFFu)
  {
    const int nx = V.tx + c_ndx[lane], ny = V.ty + c_ndy[lane];
    if (nx >= 0 && ny >= 0 && nx < V.tiles_x && ny < V.tiles_x)
    {
      nt = ny * V.tiles_x + nx;
      mykey = (kk & ~7u) | (unsigned)lane;
    }
  }
  const unsigned want = __ballot_sync(SW_FULL, nt >= 0);
  const unsigned best_key = __reduce_min_sync(SW_FULL, mykey);
  int unused = 0, c = 0;
  if (nt >= 0)
  {
    if (FINAL && mykey == best_key && atomicCAS(V.state + nt, T_IDLE, T_RUNNING) == T_IDLE)
      c = ((V.gi << 20) | nt) + 1;
    else
      unused = tile_activate(q, V.state, V.gi, nt);
  }
  __syncwarp();
  *n_act += __popc(want);
  *n_unused += __reduce_add_sync(SW_FULL, unused);
  return __reduce_max_sync(SW_FULL, c);
}

// relax the staged tile to its fixed point; publishes early once (after `early` changing sweeps,
// 0 = never) so that the tiles downstream start while this one is still confirming convergence
template <bool PROX, bool UNK>
__device__ __forceinline__ int sweep_tile_loop(sweep_tile& S, const sweep_visit& V, const heur_queue& q, const int lane,
                                               int d, const bool must_run, const double unk_w, const double L1,
                                               const double L2, int* n_sweeps, int* n_act, int* n_unused)
{
  *n_sweeps = 0;
  if (!must_run && !halo_improves<PROX, UNK>(S, lane, unk_w, L1, L2)) return 0;
  int streak = 0, n = 0, n_changing = 0;
  while (streak < 4)
  {
    const int ch = sweep_once<PROX, UNK>(S, lane, d, unk_w, L1, L2);
    streak = ch ? 1 : streak + 1;  // a sweep is idempotent: after a changing one, three clean others suffice
    n_changing += ch;
    d = (d + 1) & 3;
    ++n;
    if (ch && n_changing == (int)q.early) sweep_publish<false>(S, V, q, lane, n_act, n_unused);
  }
  *n_sweeps = n;
  return n_changing > 0;
}

__device__ __forceinline__ void cp_async8(const unsigned dst_smem, const void* src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst_smem), "l"(src));
}

#define SW_RESERVE 16  // pending counts reserved per visit: two publications x 8 neighbours

__global__ void __launch_bounds__(SW_WARPS * 32, 1) heur_sweep_kernel(heur_grid single, const heur_grid* many, heur_queue q)
{
  extern __shared__ __align__(16) unsigned char sw_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  sweep_tile& S = reinterpret_cast<sweep_tile*>(sw_smem)[warp];
  const double L1 = 0.625, L2 = 0.625 * 1.4142135623730951;  // movement_distances_ (a_star.hpp:54-57)
  int cont = 0;

  for (;;)
  {
    const long long t0 = clock64();
    int e = cont;
    if (!e) e = q_pop_warp(q, lane);
    const long long t1 = clock64();
    cont = 0;
    if (e < 0) return;
    sweep_visit V;
    V.gi = (e - 1) >> 20;
    const int gi = V.gi, tile = (e - 1) & 0xFFFFF;
#define HG(f) (many ? many[gi].f : single.f)
    const uint8_t* __restrict__ p_grid = HG(grid);
    const uint8_t* __restrict__ p_fence = HG(fence);
    const double* __restrict__ p_mc = HG(mc);
    const double* __restrict__ p_prox = HG(prox);
    double* __restrict__ gp = HG(g);
    int* state = HG(state);
    const int dim = HG(dim), tiles_x = HG(tiles_x), goal = HG(goal);
    const double w_prox = HG(w_prox), unk_w = HG(unk_w);
#undef HG
    if (lane == 0)
    {
      atomicExch(state + tile, T_RUNNING);
      // reserve the pending counts of the activations this visit may make NOW, long before any of
      // them can finish; the unused ones are returned at release (see heur_relax_kernel)
      atomicAdd(q.ctl + HC_PENDING, SW_RESERVE);
      __threadfence();
    }
    __syncwarp();

    const int tx = tile % tiles_x, ty = tile / tiles_x;
    const int ox = tx * SW_O - 1, oy = ty * SW_O - 1;
    V.gp = gp; V.state = state; V.dim = dim; V.tiles_x = tiles_x; V.tx = tx; V.ty = ty; V.ox = ox; V.oy = oy;
    // ---- stage the 32x32 block (lane = x); every address is clamped into the grid so that all
    //      loads are unconditional and in flight together; out-of-grid cells are fixed up after.
    const int gx = ox + lane;
    const bool x_in = gx >= 0 && gx < dim;
    const int cx = min(max(gx, 0), dim - 1);
    // (a) the static planes go straight to shared memory (read-only during the kernel: L1 is fine)
    {
      const unsigned s_mc = (unsigned)__cvta_generic_to_shared(S.mc + lane), s_pb = (unsigned)__cvta_generic_to_shared(S.pb + lane);
#pragma unroll 8
      for (int r = 0; r < 32; ++r)
      {
        const int cy = min(max(oy + r, 0), dim - 1);
        const size_t id = (size_t)cy * dim + cx;
        cp_async8(s_mc + r * SW_P * 8, p_mc + id);
        cp_async8(s_pb + r * SW_P * 8, p_prox + id);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
    // (b) the cost plane (L2: other tiles write it), cell and fence bytes through registers
    const int r_lo = max(0, -oy), r_hi = min(32, dim - oy);  // staged rows inside the grid
    const unsigned vm = x_in ? ((r_hi >= 32 ? 0xFFFFFFFFu : ((1u << r_hi) - 1u)) & ~((1u << r_lo) - 1u)) : 0u;
    unsigned uc = 0, bc = 0;
    for (int r0 = 0; r0 < 32; r0 += 16)
    {
      double v_g[16];
      uint8_t v_c[16], v_f[16];
#pragma unroll
      for (int j = 0; j < 16; ++j)
      {
        const int cy = min(max(oy + r0 + j, 0), dim - 1);
        const size_t id = (size_t)cy * dim + cx;
        v_g[j] = __ldcg(gp + id);
        v_c[j] = __ldg(p_grid + id);
        v_f[j] = __ldg(p_fence + id);
      }
#pragma unroll
      for (int j = 0; j < 16; ++j)
      {
        const int r = r0 + j;
        const unsigned in = (vm >> r) & 1u;
        const double gv = in ? v_g[j] : INFINITY;
        const int idx = r * SW_P + lane;
        S.g[idx] = gv;
        S.w[idx] = gv;
        uc |= (v_c[j] == GA_UNKNOWN ? 1u : 0u) << r;                   // unknown_cost, a_star.cpp:253-254
        bc |= ((v_c[j] == GA_OCC) | (v_f[j] == 1) ? 1u : 0u) << r;     // verifyNode, a_star.cpp:325-343
      }
    }
    uc &= vm;
    // blocked: off-grid, OCC / geofence, the halo ring (it belongs to the neighbour tiles), the root (keeps cost 0)
    bc |= ~vm | 0x80000001u | ((lane == 0 || lane == 31) ? 0xFFFFFFFFu : 0u);
    const int glx = goal % dim - ox, gly = goal / dim - oy;
    if (lane == glx && gly >= 0 && gly < 32) bc |= 1u << gly;
    const bool must_run = glx >= 1 && glx <= SW_O && gly >= 1 && gly <= SW_O;  // the root's tile is not a fixed point of its old halo
    if (unk_w != 0)
    {
      S.ucol[lane] = uc;
#pragma unroll
      for (int r = 0; r < 32; ++r)  // row masks = transpose of the column masks
      {
        const unsigned um = __ballot_sync(SW_FULL, (uc >> r) & 1u);
        if (lane == r) S.urow[r] = um;
      }
    }
    // (c) static planes arrived: blocked cells get mc = +inf, prox becomes prox_cost (a_star.cpp:251)
    asm volatile("cp.async.wait_all;" ::: "memory");
    int has_prox = 0;
#pragma unroll 8
    for (int r = 0; r < 32; ++r)
    {
      const int idx = r * SW_P + lane;
      if ((bc >> r) & 1u) S.mc[idx] = INFINITY;
      const double bv = ((vm >> r) & 1u) ? S.pb[idx] * w_prox : 0.0;
      S.pb[idx] = bv;
      has_prox |= (bv != 0.0);
    }
    has_prox = __any_sync(SW_FULL, has_prox);
    __syncwarp();
    // first sweep: away from the side whose halo holds the smallest value (where the front comes from)
    int d0 = 0;
    {
      const unsigned k_t = __float_as_uint(__double2float_rd(S.g[lane])), k_l = __float_as_uint(__double2float_rd(S.g[lane * SW_P]));
      const unsigned k_b = __float_as_uint(__double2float_rd(S.g[31 * SW_P + lane]));
      const unsigned k_r = __float_as_uint(__double2float_rd(S.g[lane * SW_P + 31]));
      const unsigned m_t = __reduce_min_sync(SW_FULL, k_t), m_l = __reduce_min_sync(SW_FULL, k_l);
      const unsigned m_b = __reduce_min_sync(SW_FULL, k_b), m_r = __reduce_min_sync(SW_FULL, k_r);
      unsigned best = m_t;
      if (m_l < best) { best = m_l; d0 = 1; }
      if (m_b < best) { best = m_b; d0 = 2; }
      if (m_r < best) { best = m_r; d0 = 3; }
    }
    __syncwarp();
    const long long t2 = clock64();

    int any, n_sweeps = 0, n_act = 0, n_unused = 0;
    if (has_prox)
    {
      if (unk_w != 0) any = sweep_tile_loop<true, true>(S, V, q, lane, d0, must_run, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused);
      else any = sweep_tile_loop<true, false>(S, V, q, lane, d0, must_run, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused);
    }
    else
    {
      if (unk_w != 0) any = sweep_tile_loop<false, true>(S, V, q, lane, d0, must_run, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused);
      else any = sweep_tile_loop<false, false>(S, V, q, lane, d0, must_run, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused);
    }
    const long long t3 = clock64();

    if (any) cont = sweep_publish<true>(S, V, q, lane, &n_act, &n_unused);
    if (lane == 0)
    {
      atomicAdd(q.ctl + HC_VISITS, 1);
      atomicAdd(q.ctl + HC_INNER, n_sweeps);
      if (any) atomicAdd(q.ctl + HC_IMPROVING, 1);
      // release the tile: re-queue it if a neighbour wrote into our halo meanwhile
      int dec = (SW_RESERVE - n_act) + n_unused + 1;
      if (atomicCAS(state + tile, T_RUNNING, T_IDLE) == T_DIRTY)
      {
        atomicExch(state + tile, T_QUEUED);
        q_push(q, e);
        dec -= 1;
      }
      if (dec) atomicSub(q.ctl + HC_PENDING, dec);
      unsigned long long* cyc = reinterpret_cast<unsigned long long*>(q.ctl + HC_CYC);
      const long long t4 = clock64();
      atomicAdd(cyc + 0, (unsigned long long)(t1 - t0));
      atomicAdd(cyc + 1, (unsigned long long)(t2 - t1));
      atomicAdd(cyc + 2, (unsigned long long)(t3 - t2));
      atomicAdd(cyc + 3, (unsigned long long)(t4 - t3));
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// finish: parent / is_unknown / closed flags from the converged cost plane, and cost_dist =
// (straight steps) * L1 + (diagonal steps) * L2 along the canonical parent chain (integers, so the
// result is order-free).  Three plain launches, no grid-wide barrier:
//   A  one CTA per 32x32 block: canonical parent of every cell (smallest motion index whose
//      candidate reproduces g(n) bit for bit), then every cell walks its chain INSIDE the block in
//      shared memory down to the block's terminal cell (the one whose parent lies outside, or the
//      root) and keeps (terminal, steps to it);
//   B  only terminal cells: walk terminal -> parent -> that block's terminal -> ... to the root,
//      one L2 read pair per block crossed (a few dozen), summing the block-local step counts;
//   C  every cell: its local counts + its terminal's total.
// ---------------------------------------------------------------------------------------------
#define FB 32            // finish block edge
#define FS (FB + 2)      // staged edge incl. halo

__device__ __forceinline__ unsigned fin_pack(int t_local, int cs, int cd) { return (unsigned)t_local | ((unsigned)cs << 10) | ((unsigned)cd << 20); }

__global__ void __launch_bounds__(FB * FB) heur_finish_a_kernel(heur_args single, const heur_args* many)
{
  __shared__ double sg[FS * FS];
  __shared__ double sb[FS * FS];
  __shared__ uint8_t su[FS * FS];
  __shared__ short lp[FB * FB];
  __shared__ uint8_t lt[FB * FB];
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, nbx = (dim + FB - 1) / FB;
  if ((int)blockIdx.x >= nbx * nbx) return;  // batches of unequal grids
  const int bx = blockIdx.x % nbx, by = blockIdx.x / nbx;
  const int ox = bx * FB - 1, oy = by * FB - 1;
  const double w_prox = a.w_prox, unk_w = a.unk_w;
  for (int i = threadIdx.x; i < FS * FS; i += FB * FB)
  {
    const int x = ox + i % FS, y = oy + i / FS;
    double gv = INFINITY, bv = 0.0;
    uint8_t u = 0;
    if (x >= 0 && y >= 0 && x < dim && y < dim)
    {
      const size_t id = (size_t)y * dim + x;
      gv = __ldcg(a.g + id);
      bv = __ldg(a.prox + id) * w_prox;
      u = __ldg(a.grid + id) == GA_UNKNOWN;
    }
    sg[i] = gv;
    sb[i] = bv;
    su[i] = u;
  }
  __syncthreads();
  const int lx = threadIdx.x % FB, ly = threadIdx.x / FB;
  const int nx = ox + 1 + lx, ny = oy + 1 + ly;
  const int me = (ly + 1) * FS + lx + 1;
  const bool in = nx < dim && ny < dim;
  const int n = ny * dim + nx;
  const double gn = sg[me];
  int32_t par = -2;
  uint8_t fl = 0;
  int lpar = -1, typ = 0;
  if (in && gn < INFINITY)
  {
    fl = GA_H_CLOSED;
    if (n == a.goal)
    {
      par = -1;
      if (su[me]) fl |= GA_H_UNKNOWN;  // goal_unknown, a_star.cpp:116
    }
    else
    {
      const double m = __ldg(a.mc + n);
      const double ln1 = a.L1 * m, ln2 = a.L2 * m;
#pragma unroll
      for (int i = 0; i < 8; ++i)
      {
        const int dx = -c_mdx[i], dy = -c_mdy[i];  // predecessor p = n - motion_[i]
        const int q = me + dy * FS + dx;
        const double gq = sg[q];
        if (!(gq < INFINITY)) continue;  // off-grid halo cells are +inf as well
        double cand = gq + (i < 4 ? ln1 : ln2);
        cand = cand + sb[q];
        cand = cand + (su[q] ? unk_w : 0.0);
        if (cand == gn)
        {
          par = (ny + dy) * dim + (nx + dx);
          if (su[q]) fl |= GA_H_UNKNOWN;  // NodeDisc::is_unknown_ = unknown-ness of the parent, :253,268
          typ = i < 4 ? 0 : 1;
          const int plx = lx + dx, ply = ly + dy;
          if (plx >= 0 && ply >= 0 && plx < FB && ply < FB) lpar = ply * FB + plx;
          break;
        }
      }
    }
  }
  lp[threadIdx.x] = (short)lpar;
  lt[threadIdx.x] = (uint8_t)typ;
  if (in)
  {
    a.parent[n] = par;
    a.flags[n] = fl;
  }
  const int closed = __syncthreads_count(fl != 0);
  if (threadIdx.x == 0 && closed) atomicAdd(a.ctl + HC_CLOSED, closed);
  // chain walk inside the block
  int cur = threadIdx.x, cs = 0, cd = 0;
  for (int p = lpar; p >= 0; p = lp[cur])
  {
    if (lt[cur]) ++cd; else ++cs;
    cur = p;
  }
  if (in) a.scratch[n] = fin_pack(cur, cs, cd);
}

__global__ void __launch_bounds__(256) heur_finish_b_kernel(heur_args single, const heur_args* many)
{
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, ncell = dim * dim;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  int p = __ldg(a.parent + c);
  if (p < 0) return;
  int cx = c % dim, cy = c / dim;
  int px = p % dim, py = p / dim;
  if (px / FB == cx / FB && py / FB == cy / FB) return;  // not a terminal cell
  int cs = 0, cd = 0;
  for (;;)
  {
    if (px != cx && py != cy) ++cd; else ++cs;  // the step terminal -> parent
    if (p == a.goal) break;
    const unsigned s = __ldg(a.scratch + p);    // parent -> its block's terminal
    cs += (s >> 10) & 1023;
    cd += (s >> 20) & 1023;
    const int t = ((py / FB) * FB + (int)((s & 1023) / FB)) * dim + (px / FB) * FB + (int)((s & 1023) % FB);
    if (t == a.goal) break;
    cx = t % dim;
    cy = t / dim;
    p = __ldg(a.parent + t);
    px = p % dim;
    py = p / dim;
  }
  a.tot[c] = make_int2(cs, cd);
}

__global__ void __launch_bounds__(256) heur_finish_c_kernel(heur_args single, const heur_args* many)
{
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, ncell = dim * dim;
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  double d = INFINITY;
  if (a.flags[c] & GA_H_CLOSED)
  {
    const unsigned s = a.scratch[c];
    int cs = (s >> 10) & 1023, cd = (s >> 20) & 1023;
    const int cx = c % dim, cy = c / dim;
    const int t = ((cy / FB) * FB + (int)((s & 1023) / FB)) * dim + (cx / FB) * FB + (int)((s & 1023) % FB);
    if (t != a.goal)
    {
      const int2 tt = a.tot[t];
      cs += tt.x;
      cd += tt.y;
    }
    d = (double)cs * a.L1 + (double)cd * a.L2;  // NodeDisc::cost_dist_
  }
  a.dist[c] = d;
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
  const char* impl = getenv("GA_HEUR_IMPL");  // development switch: "jacobi" = CTA-per-tile Jacobi kernel
  c->heur_impl = (impl && !strcmp(impl, "jacobi")) ? 0 : 1;
  const int edge = c->heur_impl ? SW_O : HT;
  c->tiles_x = (c->astar_dim + edge - 1) / edge;
  c->n_tiles = c->tiles_x * c->tiles_x;
  GA_REQUIRE(c, c->n_tiles < (1 << 20), GA_ERR_ARG, "astar grid too large for the tile ring encoding");
  unsigned cap = 64;
  while (cap < 4u * (unsigned)c->n_tiles + 4096u) cap <<= 1;  // + one ticket per persistent worker
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
  if (c->heur_impl)
  {
    GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)(SW_WARPS * sizeof(sweep_tile))));
    c->relax_grid = c->sm_count;
  }
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
  G.edge = c->heur_impl ? SW_O : HT;
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
  q.poll_ns = 20;
  q.early = 1;
  if (const char* s = getenv("GA_SW_EARLY")) q.early = (unsigned)atoi(s);
  if (const char* s = getenv("GA_POLL_NS")) q.poll_ns = (unsigned)atoi(s);
  {
    ga_launch_scope ls(c, GA_K_HEUR_INIT);
    heur_init_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(G, q, 0, 1);
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_HEUR_RELAX);
    if (c->heur_impl)
    {
      int nw = SW_WARPS;
      if (const char* s = getenv("GA_SW_WARPS")) nw = atoi(s);
      const int want = (c->n_tiles + nw - 1) / nw;
      int g = c->relax_grid < want ? c->relax_grid : want;
      if (const char* s = getenv("GA_SW_CTAS")) g = atoi(s);
      heur_sweep_kernel<<<g, nw * 32, nw * sizeof(sweep_tile), c->stream>>>(G, nullptr, q);
    }
    else
    {
      const int g = c->relax_grid < c->n_tiles ? c->relax_grid : c->n_tiles;
      heur_relax_kernel<<<g, H_THREADS, 0, c->stream>>>(G, nullptr, q);
    }
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
  a.scratch = reinterpret_cast<unsigned*>(c->d_jump);
  a.tot = reinterpret_cast<int2*>(c->d_cnt);
  a.dim = dim;
  a.goal = G.goal;
  a.w_prox = G.w_prox;
  a.unk_w = G.unk_w;
  // movement_distances_ (a_star.hpp:54-57): ASTAR_RES and ASTAR_RES * sqrt(2)
  a.L1 = 0.625;
  a.L2 = 0.625 * sqrt(2.0);
  {
    ga_launch_scope ls(c, GA_K_HEUR_FINISH);
    const int nb = (dim + FB - 1) / FB, cells = (dim * dim + 255) / 256;
    heur_finish_a_kernel<<<nb * nb, FB * FB, 0, c->stream>>>(a, nullptr);
    heur_finish_b_kernel<<<cells, 256, 0, c->stream>>>(a, nullptr);
    heur_finish_c_kernel<<<cells, 256, 0, c->stream>>>(a, nullptr);
  }
  GA_CUDA(c, cudaGetLastError());
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
