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
//   - the astar grid is cut into tiles of 30x30 owned cells; a tile is relaxed by ONE WARP: the
//     32x32 block (tile + 1-cell halo ring) lives in the warp's slice of shared memory and is
//     relaxed by Gauss-Seidel line sweeps (see the sweep section below), improved cells written back;
//   - tiles whose facing border (or corner) cell improved are activated; activation is dataflow: a
//     per-tile state word (idle / queued / running / running+dirty) and one device-wide ticket ring
//     of tile ids.  Persistent warps (8 or 12 per SM) take tickets, so the wavefront is balanced
//     dynamically over the 148 SMs and a worker never waits for another one;
//   - single builds (latency mode) put FOUR warps on a tile, one sweep direction each: the leader runs the
//     protocol, three helpers sweep on command between two named barriers (GROUP template parameter);
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

enum { T_IDLE = 0, T_QUEUED = 1, T_RUNNING = 2, T_DIRTY = 3 };

// control block words (ga_ctx::d_hctl)
enum
{
  HC_HEAD = 0,      // ring consumer cursor
  HC_TAIL = 1,      // ring producer cursor
  HC_PENDING = 2,   // tiles queued or running
  HC_CLOSED = 3,    // closed-cell count (finish kernel A)
  HC_VISITS = 8,    // statistics: tile visits, sweeps, visits that improved a cell
  HC_INNER = 9,
  HC_IMPROVING = 10,
  HC_CYC = 16,      // u64 x 4: SM cycles spent by the workers in pop / stage / sweeps (+ early publication) / final publication
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
  uint8_t* cls;  // per cell, written by the init kernel: bit0 = UNKNOWN, bit1 = blocked (OCC or geofence)
  int* state;    // per tile
  int* tprox;    // per tile: == epoch when some cell of the staged block carries a prox cost (set by the init kernel)
  int epoch;     // build counter of the context (stamps need no clearing between builds)
  int dim, tiles_x, n_tiles, goal;
  int edge;  // owned cells per tile edge (SW_O)
  double w_prox, unk_w;
};

struct heur_queue
{
  int* ring;      // entries: ((grid index << 20) | tile) + 1, 0 = empty slot
  unsigned mask;  // capacity - 1 (power of two >= total tiles)
  int* ctl;
  unsigned poll_ns;  // back-off of an idle worker between two looks at its ring slot
  unsigned early;    // publish once after this many changing sweeps (0 = only when converged)
  unsigned linger;   // SM cycles a converged tile stays staged waiting for the next publication (0 = release at once)
  unsigned prefetch; // 1: a tile the wave reaches for the first time queues its idle neighbours before sweeping
  unsigned refresh;  // a running visit picks up new halo values between sweeps: 1 = state word read during the sweep, 2 = waited for before it
  unsigned round;    // 1: the four sweeps run in lock step (sweep_round), 0: one after the other
  unsigned pair;     // 0 / 2 / 4 (needs round == 0): warps per tile in single builds, sweeping different directions at the same time
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
  int* closed;        // closed-cell counter of this grid
  int* tlist;         // [blocks][FIN_SLOTS] finish A: the terminal cells of each 32x32 block (parent in another block)
  int* tcount;        // [blocks] how many
  unsigned* scratch;  // [ncell] finish A: displacement of the block's terminal cell and (straight, diagonal) steps to it
  int2* tot;          // [ncell] finish B: (straight, diagonal) steps of a terminal cell to the root
  uint32_t* steps;    // [ncell] finish C: straight | diagonal << 14 steps of a closed cell to the root (saturating; GA_GRID_HEUR_REC)
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

#ifdef SW_TRACE  // development: per-tile event log (tools/trace_wave.py); never defined in the shipped build
__device__ unsigned long long g_trace[1 << 20];
__device__ unsigned g_trace_n;
__device__ __forceinline__ void trace_ev(int lane, int tile, int ev, int arg)
{
  if (lane != 0) return;
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  const unsigned i = atomicAdd(&g_trace_n, 1u);
  if (i < (1u << 20)) g_trace[i] = (t << 24) | ((unsigned long long)(tile & 0xFFF) << 12) | ((unsigned long long)(ev & 15) << 8) | (arg & 255);
}
#define TRACE(lane, tile, ev, arg) trace_ev(lane, tile, ev, arg)
#else
#define TRACE(lane, tile, ev, arg) ((void)0)
#endif

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
// The state word is two bits -- T_QUEUED = "news", T_RUNNING = "a worker has it", T_DIRTY = both -- so ONE
// atomic OR does every transition: idle -> queued (and we push it), running -> dirty, the rest unchanged.
// The caller has already counted the tile in HC_PENDING; returns 1 if the count was NOT consumed.
__device__ __forceinline__ int tile_activate(const heur_queue& q, int* state, int gi, int t)
{
  if (atomicOr(state + t, (int)T_QUEUED) != T_IDLE) return 1;
  q_push(q, ((gi << 20) | t) + 1);
  return 0;
}

// blocks until this worker's ticket is served (returns the entry) or all work is done (returns -1).
// Warp-wide: every lane executes the same control flow (lane 0 takes the ticket, all lanes
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

// every tile whose staged block [t * edge - 1, t * edge + edge] holds cell i carries a prox cost (at most 2 x 2 tiles)
__device__ __forceinline__ void init_stamp_prox(const heur_grid& G, size_t i)
{
  const int x = (int)(i % G.dim), y = (int)(i / G.dim);
  for (int ty = y / G.edge - 1; ty <= y / G.edge + 1; ++ty)
    for (int tx = x / G.edge - 1; tx <= x / G.edge + 1; ++tx)
      if (tx >= 0 && ty >= 0 && tx < G.tiles_x && ty < G.tiles_x && x >= tx * G.edge - 1 && x <= tx * G.edge + G.edge &&
          y >= ty * G.edge - 1 && y <= ty * G.edge + G.edge)
        G.tprox[ty * G.tiles_x + tx] = G.epoch;
}

// verifyNode (a_star.cpp:325-343) and unknown_cost (:253-254) inputs, folded once per build
__device__ __forceinline__ uint8_t init_cls(uint8_t cell, uint8_t fence)
{
  return (uint8_t)((cell == GA_UNKNOWN ? 1 : 0) | ((cell == GA_OCC || fence == 1) ? 2 : 0));
}

// one grid per blockIdx.y.  The control block (cursors, pending count) was zeroed by the host.  Four cells per thread
// and step (4-byte loads of the byte planes, 16-byte stores of the cost plane); G.prox == nullptr: the context's
// proximity plane is known to be zero (prox_free), it is not read at all.
__global__ void heur_init_kernel(heur_grid single, const heur_grid* many, heur_queue q)
{
  const int gi = blockIdx.y;
  const heur_grid& G = many ? many[gi] : single;
  const size_t n = (size_t)G.dim * G.dim, n4 = n / 4;
  const size_t goal = (size_t)G.goal;
  const double w_prox = G.w_prox;
  const bool with_prox = G.prox != nullptr && w_prox != 0.0;
  for (size_t i4 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i4 < n4; i4 += (size_t)gridDim.x * blockDim.x)
  {
    const size_t i = 4 * i4;
    const uchar4 cell = *reinterpret_cast<const uchar4*>(G.grid + i), fc = *reinterpret_cast<const uchar4*>(G.fence + i);
    double2 p0 = make_double2(0.0, 0.0), p1 = p0;
    if (with_prox)
    {
      p0 = *reinterpret_cast<const double2*>(G.prox + i);
      p1 = *reinterpret_cast<const double2*>(G.prox + i + 2);
    }
    double2 g0 = make_double2(INFINITY, INFINITY), g1 = g0;
    if (goal - i < 4)  // the root (unsigned difference: goal in [i, i + 4))
    {
      if (goal == i) g0.x = 0.0;
      if (goal == i + 1) g0.y = 0.0;
      if (goal == i + 2) g1.x = 0.0;
      if (goal == i + 3) g1.y = 0.0;
    }
    *reinterpret_cast<double2*>(G.g + i) = g0;
    *reinterpret_cast<double2*>(G.g + i + 2) = g1;
    *reinterpret_cast<uchar4*>(G.cls + i) = make_uchar4(init_cls(cell.x, fc.x), init_cls(cell.y, fc.y), init_cls(cell.z, fc.z), init_cls(cell.w, fc.w));
    if (with_prox)
    {
      if (p0.x * w_prox != 0.0) init_stamp_prox(G, i);
      if (p0.y * w_prox != 0.0) init_stamp_prox(G, i + 1);
      if (p1.x * w_prox != 0.0) init_stamp_prox(G, i + 2);
      if (p1.y * w_prox != 0.0) init_stamp_prox(G, i + 3);
    }
  }
  // the last n mod 4 cells
  if (blockIdx.x == 0 && threadIdx.x < (unsigned)(n - 4 * n4))
  {
    const size_t i = 4 * n4 + threadIdx.x;
    G.g[i] = (i == goal) ? 0.0 : INFINITY;
    G.cls[i] = init_cls(G.grid[i], G.fence[i]);
    if (with_prox && G.prox[i] * w_prox != 0.0) init_stamp_prox(G, i);
  }
  const int gx = G.goal % G.dim, gy = G.goal / G.dim;
  const int goal_tile = (gy / G.edge) * G.tiles_x + gx / G.edge;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < G.n_tiles; i += gridDim.x * blockDim.x)
    G.state[i] = (i == goal_tile) ? T_QUEUED : T_IDLE;
  if (blockIdx.x == 0 && threadIdx.x == 0)
  {
    // the ring is empty between builds (every entry popped is cleared): seed it with the root's tile
    atomicAdd(q.ctl + HC_PENDING, 1);
    q_push(q, ((gi << 20) | goal_tile) + 1);
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
// nothing (a sweep is idempotent).  A CTA-per-tile Jacobi iteration (round 1's first kernel) moved a
// front 1 cell per 0.46 us with 512 threads; a sweep moves it 30 cells per 1.7 us with 32.
// Warps are independent workers on the ticket ring (no __syncthreads anywhere).
// ---------------------------------------------------------------------------------------------
#define SW_O 30                    // owned cells per tile edge
#define SW_P 33                    // shared-memory pitch (doubles)
#define SW_WARPS 8                 // workers per CTA (one CTA per SM: 8 x 25 KB of shared memory)
#define SW_WARPS_LEAN 12           // the same without the proximity plane (12 x 17 KB)
#define SW_FULL 0xFFFFFFFFu
#define SW_EARLY_ALL 255           // heur_queue::early: publish after every changing sweep

template <bool HAS_PB>  // false: a launch whose grids carry no proximity cost at all -- no pb plane, 12 workers per SM instead of 8
struct sweep_tile_t
{
  double g[32 * SW_P];   // cost plane, relaxed in place
  double mc[32 * SW_P];  // movement_cost_map_ of the cell (edge length factor of the DESTINATION); +inf = blocked
  double pb[HAS_PB ? 32 * SW_P : 1];  // prox_cost of the cell = h_prox_arr_ x astar_prox_cost_ (term of the PREDECESSOR)
  unsigned urow[32], ucol[32];  // UNKNOWN masks: urow[y] bit x / ucol[x] bit y
  unsigned dirty[32];           // dirty[y] bit x: cell improved since it was last written to global memory
  // pair mode (latency): a helper warp sweeps the perpendicular direction at the same time; its share of the dirty map,
  // the command word (direction | PROX << 2 | UNK << 3, -1 = leave), its "changed" answer and the two weights
  unsigned dirty2[3][32];
  int cmd[4], chh[4];
  double par_w, par_u;
};

// NaN-safe minimum (a NaN candidate, 0/0 in h_prox_arr_, never wins -- a_star.cpp:492-497)
__device__ __forceinline__ double dmin(double a, double b) { return fmin(a, b); }

__device__ __forceinline__ double umin_f64(double a, double b)
{
#ifdef SW_DMIN
  return b < a ? b : a;
#else
  const unsigned long long ua = (unsigned long long)__double_as_longlong(a), ub = (unsigned long long)__double_as_longlong(b);
  return __longlong_as_double((long long)(ub < ua ? ub : ua));
#endif
}

// One line step of a sweep: the new value of cell `lane` of line s from the three cells of line sp
// (a = this lane's value there; the neighbours' arrive by shuffle).  ORIENT 0: lane = x, lines are
// rows; ORIENT 1: lane = y, lines are columns.  Blocked cells (OCC, geofence, root, off-grid, halo)
// carry mc = +inf, so their candidates are +inf / NaN and they keep their value; lanes 0 / 31 are
// halo cells, so what their missing neighbour "sends" does not matter.
template <int ORIENT, bool PROX, bool UNK, class TILE>
__device__ __forceinline__ double sweep_step(const TILE& S, const int lane, const int sp, const int s, const double a,
                                             const double w_prox, const double unk_w, const double L1, const double L2, double* old_out)
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
    // values are >= 0 or +inf here, so the IEEE order is the order of the bit patterns: the minima
    // run on the integer pipe (shorter latency than DSETP + select on the dependent chain)
    const double c1 = a + ln1;
    const double m0 = umin_f64(c1, old);  // overlaps the shuffle latency
    const double m2 = umin_f64(ar, al);
    const double c2 = m2 + ln2;
    return umin_f64(c2, m0);
  }
  const int lm = lane > 0 ? lane - 1 : 0, lp = lane < 31 ? lane + 1 : 31;
  double c0 = a + ln1, c1 = al + ln2, c2 = ar + ln2;
  if (PROX)
  {
    const int pi = sp * ds;
    c0 = c0 + S.pb[pi + lane * dl];  // prox_cost of the predecessor (h_prox x astar_prox_cost_, multiplied once at staging), a_star.cpp:251
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
  // candidates are >= 0, +inf or NaN (0/0 in h_prox_arr_, a_star.cpp:492-497); a NaN's bit pattern is
  // above +inf's as an unsigned integer whatever its sign, so it never wins: fmin's result on the integer pipe
  return umin_f64(umin_f64(old, c0), umin_f64(c1, c2));
}

// one sweep over the 30 owned lines; dir = +1 / -1: line order.  Returns 1 if this lane improved a cell.
template <int ORIENT, bool PROX, bool UNK, bool HELPER, class TILE>
__device__ __forceinline__ int sweep_lines(TILE& S, const int lane, const int dir, const double w_prox, const double unk_w,
                                           const double L1, const double L2, const int hidx)
{
  unsigned* const dirty = HELPER ? S.dirty2[hidx - 1] : S.dirty;  // (a member of the shared tile either way)
  constexpr int dl = ORIENT == 0 ? 1 : SW_P, ds = ORIENT == 0 ? SW_P : 1;
  int s = dir > 0 ? 0 : 31;
  double a = S.g[s * ds + lane * dl];
  int changed = 0;
  unsigned dm = dirty[lane];  // this lane keeps row `lane` of the dirty map, whichever way the sweep runs
#pragma unroll 5
  for (int k = 0; k < SW_O; ++k)
  {
    const int sp = s;
    s += dir;
    double old;
    const double nv = sweep_step<ORIENT, PROX, UNK>(S, lane, sp, s, a, w_prox, unk_w, L1, L2, &old);
    const bool imp = nv < old;
    if (imp) S.g[s * ds + lane * dl] = nv;
    changed |= imp;
    if (ORIENT == 0)
    {
      const unsigned b = __ballot_sync(SW_FULL, imp);  // lane = x, line = row s: the ballot IS row s's news
      if (lane == s) dm |= b;
    }
    else dm |= (unsigned)imp << s;                     // lane = y, line = column s
    a = nv;
  }
  dirty[lane] = dm;
  return changed;
}

// can the halo improve a border cell?  (first line step of each of the four sweeps, nothing stored)
template <bool PROX, bool UNK, class TILE>
__device__ __forceinline__ int halo_improves(const TILE& S, const int lane, const double w_prox, const double unk_w, const double L1,
                                             const double L2)
{
  double old;
  int imp = 0;
  imp |= sweep_step<0, PROX, UNK>(S, lane, 0, 1, S.g[lane], w_prox, unk_w, L1, L2, &old) < old;
  imp |= sweep_step<0, PROX, UNK>(S, lane, 31, 30, S.g[31 * SW_P + lane], w_prox, unk_w, L1, L2, &old) < old;
  imp |= sweep_step<1, PROX, UNK>(S, lane, 0, 1, S.g[lane * SW_P], w_prox, unk_w, L1, L2, &old) < old;
  imp |= sweep_step<1, PROX, UNK>(S, lane, 31, 30, S.g[lane * SW_P + 31], w_prox, unk_w, L1, L2, &old) < old;
  return __any_sync(SW_FULL, imp);
}

template <bool PROX, bool UNK, bool HELPER, class TILE>
__device__ __forceinline__ int sweep_once(TILE& S, const int lane, const int d, const double w_prox, const double unk_w, const double L1,
                                          const double L2, const int hidx = 0)
{
  int ch;
  if (d == 0) ch = sweep_lines<0, PROX, UNK, HELPER>(S, lane, +1, w_prox, unk_w, L1, L2, hidx);
  else if (d == 1) ch = sweep_lines<1, PROX, UNK, HELPER>(S, lane, +1, w_prox, unk_w, L1, L2, hidx);
  else if (d == 2) ch = sweep_lines<0, PROX, UNK, HELPER>(S, lane, -1, w_prox, unk_w, L1, L2, hidx);
  else ch = sweep_lines<1, PROX, UNK, HELPER>(S, lane, -1, w_prox, unk_w, L1, L2, hidx);
  ch = __any_sync(SW_FULL, ch);
  __syncwarp();
  return ch;
}

// One ROUND: the four sweeps in lock step, one line step of each per iteration.  The four chains are
// independent instruction streams (each carries its own line in registers), so their shuffle / fp64
// latencies overlap: a round costs about 1.5 sweeps instead of 4.  Any interleaving of relaxations
// is admissible -- every stored value is the exact candidate of a value some neighbour held, values
// only fall, and the loop ends on a round that stores nothing, i.e. on the fixed point.  Where two
// chains meet in one cell in the same step (row s x column s) the later store may put the larger of
// two improvements back; the next round repairs it (and reports a change now).
template <bool PROX, bool UNK, class TILE>
__device__ __forceinline__ int sweep_round(TILE& S, const int lane, const double w_prox, const double unk_w, const double L1,
                                           const double L2)
{
  double a0 = S.g[lane], a1 = S.g[lane * SW_P], a2 = S.g[31 * SW_P + lane], a3 = S.g[lane * SW_P + 31];
  int changed = 0;
  unsigned dm = S.dirty[lane];
#pragma unroll 2
  for (int k = 0; k < SW_O; ++k)
  {
    const int sf = k + 1, sb = SW_O - k;
    double o0, o1, o2, o3;
    const double n0 = sweep_step<0, PROX, UNK>(S, lane, sf - 1, sf, a0, w_prox, unk_w, L1, L2, &o0);  // rows, downwards
    const double n1 = sweep_step<1, PROX, UNK>(S, lane, sf - 1, sf, a1, w_prox, unk_w, L1, L2, &o1);  // columns, rightwards
    const double n2 = sweep_step<0, PROX, UNK>(S, lane, sb + 1, sb, a2, w_prox, unk_w, L1, L2, &o2);  // rows, upwards
    const double n3 = sweep_step<1, PROX, UNK>(S, lane, sb + 1, sb, a3, w_prox, unk_w, L1, L2, &o3);  // columns, leftwards
    const bool i0 = n0 < o0, i1 = n1 < o1, i2 = n2 < o2, i3 = n3 < o3;
    if (i0) S.g[sf * SW_P + lane] = n0;
    if (i1) S.g[lane * SW_P + sf] = n1;
    if (i2) S.g[sb * SW_P + lane] = n2;
    if (i3) S.g[lane * SW_P + sb] = n3;
    changed |= (int)(i0 | i1 | i2 | i3);
    const unsigned b0 = __ballot_sync(SW_FULL, i0), b2 = __ballot_sync(SW_FULL, i2);
    if (lane == sf) dm |= b0;
    if (lane == sb) dm |= b2;
    dm |= ((unsigned)i1 << sf) | ((unsigned)i3 << sb);
    a0 = n0; a1 = n1; a2 = n2; a3 = n3;
  }
  S.dirty[lane] = dm;
  changed = __any_sync(SW_FULL, changed);
  __syncwarp();
  return changed;
}

struct sweep_visit  // what a visit needs to publish its results
{
  double* gp;
  int* state;
  int dim, tiles_x, gi, tile, tx, ty, ox, oy;
};

// second half of a publication: kk = (lanes 0..7: W, E, N, S, NW, NE, SW, SE) the smallest improved
// border value towards that neighbour, 0xFFFFFFFF = nothing; `stored` = this lane wrote to the cost
// plane.  FINAL: the neighbour in front (smallest improved border value) is taken over by this
// worker when it is idle -- no ring hand-off on the critical chain; returns its entry.
// n_act / n_unused: pending-count bookkeeping (reserved at the start of the visit).
template <bool FINAL>
__device__ __forceinline__ int sweep_notify(const sweep_visit& V, const heur_queue& q, const int lane, const unsigned kk,
                                            const bool stored, int* n_act, int* n_unused)
{
  // the stores are visible before any neighbour is told to look and before the tile is released
  if (__any_sync(SW_FULL, stored)) asm volatile("fence.acq_rel.gpu;" ::: "memory");
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

__device__ __forceinline__ unsigned sweep_dir_keys(const int lane, const unsigned k_w, const unsigned k_e, const unsigned kn, const unsigned ks)
{
  const unsigned k_n = __reduce_min_sync(SW_FULL, kn), k_s = __reduce_min_sync(SW_FULL, ks);
  const unsigned k_nw = __shfl_sync(SW_FULL, kn, 1), k_ne = __shfl_sync(SW_FULL, kn, SW_O);
  const unsigned k_sw = __shfl_sync(SW_FULL, ks, 1), k_se = __shfl_sync(SW_FULL, ks, SW_O);
  return lane == 0 ? k_w : lane == 1 ? k_e : lane == 2 ? k_n : lane == 3 ? k_s : lane == 4 ? k_nw : lane == 5 ? k_ne
       : lane == 6 ? k_sw : lane == 7 ? k_se : 0xFFFFFFFFu;
}

// write back the owned cells that improved since the last publication, activate the neighbour tiles
// whose halo they are
template <bool FINAL, class TILE>
__device__ __forceinline__ int sweep_publish(TILE& S, const sweep_visit& V, const heur_queue& q, const int lane,
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
    const bool st = own_x & ((S.dirty[r] >> lane) & 1u);
    const unsigned kb = st ? __float_as_uint(__double2float_rd(v)) : 0xFFFFFFFFu;
    if (st) __stcg(grow, v);
    col_key = min(col_key, kb);
    if (r == 1) kn = kb;
    if (r == SW_O) ks = kb;
  }
  __syncwarp();
  S.dirty[lane] = 0;
  __syncwarp();
  const unsigned k_w = __shfl_sync(SW_FULL, col_key, 1), k_e = __shfl_sync(SW_FULL, col_key, SW_O);
  const unsigned kk = sweep_dir_keys(lane, k_w, k_e, kn, ks);
  return sweep_notify<FINAL>(V, q, lane, kk, col_key != 0xFFFFFFFFu, n_act, n_unused);
}

// the border cells only (rows 1 / 30 with lane = x, columns 1 / 30 with lane = y): all a neighbour
// needs, a third of the latency.  The interior goes out with the visit's final publication.
template <class TILE>
__device__ __forceinline__ void sweep_publish_border(TILE& S, const sweep_visit& V, const heur_queue& q, const int lane,
                                                     int* n_act, int* n_unused)
{
  const unsigned none = 0xFFFFFFFFu;
  const bool own = lane >= 1 && lane <= SW_O;
  const int i_n = 1 * SW_P + lane, i_s = SW_O * SW_P + lane, i_w = lane * SW_P + 1, i_e = lane * SW_P + SW_O;
  const double v_n = S.g[i_n], v_s = S.g[i_s], v_w = S.g[i_w], v_e = S.g[i_e];
  const unsigned d_n = S.dirty[1], d_s = S.dirty[SW_O], d_y = S.dirty[lane];  // rows 1 / 30 (lane = x), row `lane` (lane = y)
  const bool s_n = own && ((d_n >> lane) & 1u), s_s = own && ((d_s >> lane) & 1u);
  const bool s_w = own && ((d_y >> 1) & 1u), s_e = own && ((d_y >> SW_O) & 1u);
  __syncwarp();  // every lane has read the map before the published cells are taken out of it
  S.dirty[lane] = (lane == 1 || lane == SW_O) ? 0u : (d_y & ~((1u << 1) | (1u << SW_O)));
  double* g00 = V.gp + (size_t)V.oy * V.dim + V.ox;
  if (s_n) __stcg(g00 + (size_t)1 * V.dim + lane, v_n);
  if (s_s) __stcg(g00 + (size_t)SW_O * V.dim + lane, v_s);
  if (s_w) __stcg(g00 + (size_t)lane * V.dim + 1, v_w);
  if (s_e) __stcg(g00 + (size_t)lane * V.dim + SW_O, v_e);
  const unsigned kn = s_n ? __float_as_uint(__double2float_rd(v_n)) : none, ks = s_s ? __float_as_uint(__double2float_rd(v_s)) : none;
  const unsigned kw = s_w ? __float_as_uint(__double2float_rd(v_w)) : none, ke = s_e ? __float_as_uint(__double2float_rd(v_e)) : none;
  const unsigned k_w = __reduce_min_sync(SW_FULL, kw), k_e = __reduce_min_sync(SW_FULL, ke);
  const unsigned kk = sweep_dir_keys(lane, k_w, k_e, kn, ks);
  sweep_notify<false>(V, q, lane, kk, s_n || s_s || s_w || s_e, n_act, n_unused);
}

// A neighbour published into the halo while this visit was running (state word DIRTY): take the
// mark back and re-read the halo ring only -- the owned cells are written by nobody else and the
// static planes have not changed.  Replaces release + ring + pop + a full re-stage on the chain a
// corrected value travels along.
// the state word taken and reset to RUNNING in one atomic (acquire): DIRTY = a neighbour's mark, consumed by this call
__device__ __forceinline__ int state_take(const sweep_visit& V, const int lane)
{
  int old_state = 0;
  if (lane == 0)
    asm volatile("atom.acquire.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old_state) : "l"(V.state + V.tile), "r"((int)T_RUNNING) : "memory");
  return __shfl_sync(SW_FULL, old_state, 0);
}

template <class TILE>
__device__ __forceinline__ int sweep_refresh_halo(TILE& S, const sweep_visit& V, const int lane, const bool taken = false)
{
  if (lane == 0 && !taken)
  {
    int old_state;  // acquire: the marker's cost-plane writes are visible to the loads below
    asm volatile("atom.acquire.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old_state) : "l"(V.state + V.tile), "r"((int)T_RUNNING) : "memory");
    (void)old_state;
  }
  __syncwarp();
  const int dim = V.dim, g_ = V.ox + lane;   // lane = x for the two rows
  const int h_ = V.oy + lane;                // lane = y for the two columns
  const int cx = min(max(g_, 0), dim - 1), cy = min(max(h_, 0), dim - 1);
  const int y0 = V.oy, y1 = V.oy + 31, x0 = V.ox, x1 = V.ox + 31;
  const double v_t = __ldcg(V.gp + (size_t)max(y0, 0) * dim + cx), v_b = __ldcg(V.gp + (size_t)min(y1, dim - 1) * dim + cx);
  const double v_l = __ldcg(V.gp + (size_t)cy * dim + max(x0, 0)), v_r = __ldcg(V.gp + (size_t)cy * dim + min(x1, dim - 1));
  const bool x_in = g_ >= 0 && g_ < dim, y_in = h_ >= 0 && h_ < dim;
  const double n_t = (x_in && y0 >= 0) ? v_t : INFINITY, n_b = (x_in && y1 < dim) ? v_b : INFINITY;
  const double n_l = (y_in && x0 >= 0) ? v_l : INFINITY, n_r = (y_in && x1 < dim) ? v_r : INFINITY;
  // the side the news came from = the one with the smallest improved halo value; the next sweep
  // runs away from it (same numbering as the first sweep of a visit)
  const unsigned none = 0xFFFFFFFFu;
  const unsigned k_t = n_t < S.g[lane] ? __float_as_uint(__double2float_rd(n_t)) : none;
  const unsigned k_l = n_l < S.g[lane * SW_P] ? __float_as_uint(__double2float_rd(n_l)) : none;
  const unsigned k_b = n_b < S.g[31 * SW_P + lane] ? __float_as_uint(__double2float_rd(n_b)) : none;
  const unsigned k_r = n_r < S.g[lane * SW_P + 31] ? __float_as_uint(__double2float_rd(n_r)) : none;
  __syncwarp();  // corner cells are read by two lanes: all reads before the writes
  S.g[lane] = n_t;
  S.g[31 * SW_P + lane] = n_b;
  S.g[lane * SW_P] = n_l;
  S.g[lane * SW_P + 31] = n_r;
  const unsigned m_t = __reduce_min_sync(SW_FULL, k_t), m_l = __reduce_min_sync(SW_FULL, k_l);
  const unsigned m_b = __reduce_min_sync(SW_FULL, k_b), m_r = __reduce_min_sync(SW_FULL, k_r);
  int d = -1;
  unsigned best = none;
  if (m_t < best) { best = m_t; d = 0; }
  if (m_l < best) { best = m_l; d = 1; }
  if (m_b < best) { best = m_b; d = 2; }
  if (m_r < best) { best = m_r; d = 3; }
  __syncwarp();
  return d;  // -1: no halo cell got smaller
}

// The wave has reached this tile for the first time: queue the idle neighbours NOW, so that idle
// workers stage them (3 us) while this tile runs its first sweep; they then wait a moment for the
// publication (`linger`) instead of being activated, popped and staged after it.
__device__ __forceinline__ void sweep_prefetch(const sweep_visit& V, const heur_queue& q, const int lane, int* n_act, int* n_unused)
{
  int unused = 0;
  if (lane < 8)
  {
    unused = 1;
    const int nx = V.tx + c_ndx[lane], ny = V.ty + c_ndy[lane];
    if (nx >= 0 && ny >= 0 && nx < V.tiles_x && ny < V.tiles_x)
    {
      const int nt = ny * V.tiles_x + nx;
      if (atomicCAS(V.state + nt, T_IDLE, T_QUEUED) == T_IDLE)
      {
        q_push(q, ((V.gi << 20) | nt) + 1);
        unused = 0;
      }
    }
  }
  __syncwarp();
  *n_act += 8;
  *n_unused += __reduce_add_sync(SW_FULL, unused);
}

// named barrier of a leader / helper pair (64 threads); also orders their shared-memory accesses
template <int GROUP>
__device__ __forceinline__ void pair_bar(const int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(32 * GROUP) : "memory"); }

// the helper warp of a pair: sweeps on command, touches nothing else (its dirty bits go to dirty2, merged by the leader)
template <int GROUP, class TILE>
__device__ __forceinline__ void sweep_helper(TILE& S, const int lane, const int pbar, const int hidx, const double L1, const double L2)
{
  for (;;)
  {
    pair_bar<GROUP>(pbar);
    const int cmd = S.cmd[hidx];
    if (cmd < 0) return;
    const int d = cmd & 3;
    const double w_prox = S.par_w, unk_w = S.par_u;
    S.dirty2[hidx - 1][lane] = 0;
    int ch;
    if (cmd & 4)
      ch = (cmd & 8) ? sweep_once<true, true, true>(S, lane, d, w_prox, unk_w, L1, L2, hidx) : sweep_once<true, false, true>(S, lane, d, w_prox, unk_w, L1, L2, hidx);
    else
      ch = (cmd & 8) ? sweep_once<false, true, true>(S, lane, d, w_prox, unk_w, L1, L2, hidx) : sweep_once<false, false, true>(S, lane, d, w_prox, unk_w, L1, L2, hidx);
    if (lane == 0) S.chh[hidx] = ch;
    pair_bar<GROUP>(pbar + 1);
  }
}

// relax the staged tile to its fixed point.  Publication: once after `early` changing sweeps (0 =
// never) plus the final one by the caller, or after every changing sweep (SW_EARLY_ALL).  With
// `refresh` a neighbour's publication is picked up between sweeps; with `linger` (SM cycles; needs
// SW_EARLY_ALL, everything is published then) the converged tile stays staged that long waiting for
// the next one.
template <bool PROX, bool UNK, int GROUP, class TILE>
__device__ __forceinline__ int sweep_tile_loop(TILE& S, const sweep_visit& V, const heur_queue& q, const int lane,
                                               int d, const bool must_run, const bool fresh, const double w_prox, const double unk_w,
                                               const double L1, const double L2, int* n_sweeps, int* n_act, int* n_unused, int* n_reserved,
                                               const int pbar)
{
  int n = 0, n_changing = 0, n_pub = 0;
  const bool run0 = must_run || halo_improves<PROX, UNK>(S, lane, w_prox, unk_w, L1, L2);
  TRACE(lane, V.tile, 2, (int)run0 | ((int)fresh << 1));
  const bool can_linger = q.linger != 0 && q.early == SW_EARLY_ALL;  // everything is published when the tile waits
  if (!run0 && !can_linger) return 0;
  bool prefetch = q.prefetch && fresh;
  bool may_linger = can_linger && !run0;  // armed by a changing sweep, spent by one wait
  // directions swept without a change since the last change (a sweep is idempotent, so a changing
  // one counts for its own direction); 15 = fixed point.  prim = the next sweep's direction: round
  // robin, restarted away from the side the news came from after a halo refresh.
  unsigned mask = run0 ? 0u : 15u;
  int pref_h = (d >> 2) & 3, pref_v = (d >> 4) & 3, help_d = -1;  // pair mode: preferred perpendiculars, the helper's last direction (-1: restart)
  d &= 3;
  int prim = d, last_ch = 0;
  for (;;)
  {
    // refresh 1: looked at after the sweep, 2: waited for now.  Groups look with the exchange itself: a mark is seen
    // and taken back in one round trip instead of a load followed by the exchange
    const bool take = GROUP > 1 && q.refresh == 2;
    int st = take ? state_take(V, lane) : ld_volatile(V.state + V.tile);
    if (q.refresh == 2)
    {
      if (st != T_DIRTY && may_linger && !last_ch)
      {
        // the last sweep was clean (or the tile was staged ahead of the wave): news first, the
        // confirming sweeps later -- nobody waits for those
        TRACE(lane, V.tile, 6, 0);
        const long long t_end = clock64() + (long long)q.linger;
        do st = take ? state_take(V, lane) : ld_volatile(V.state + V.tile);
        while (st != T_DIRTY && __all_sync(SW_FULL, clock64() < t_end));
        may_linger = false;
        TRACE(lane, V.tile, 8, st == T_DIRTY);
      }
      if (st == T_DIRTY)
      {
        const int dn = sweep_refresh_halo(S, V, lane, take);
        const bool news = dn >= 0 && (GROUP == 4 || halo_improves<PROX, UNK>(S, lane, w_prox, unk_w, L1, L2));
        TRACE(lane, V.tile, 5, (dn + 1) | ((int)news << 3));
        if (news)
        {
          mask = 0;
          prim = dn;
          if (dn & 1) pref_h = dn; else pref_v = dn;
          help_d = -1;
        }
        st = T_RUNNING;
      }
    }
    if (mask == 15u) break;
    if (prefetch)
    {
      sweep_prefetch(V, q, lane, n_act, n_unused);
      prefetch = false;
    }
    int ch;
    if (GROUP == 1 && q.round)  // (groups are only formed without lock-step rounds: the path is compiled out of their kernels)
    {
      ch = sweep_round<PROX, UNK>(S, lane, w_prox, unk_w, L1, L2);
      mask = ch ? 0u : 15u;  // a clean round has looked at every cell from all eight sides
      n += 4;
    }
    else if (GROUP == 2)
    {
      // pair mode: the helper warp sweeps the next (perpendicular) direction while this warp sweeps `prim` -- a wave that
      // crosses the tile diagonally needs both, one after the other they cost two sweep times per tile
      d = prim;
      // after news: the perpendicular that leaves the side the wave comes from; then the opposite pair, and so on
      const int d2 = help_d < 0 ? ((d & 1) ? pref_v : pref_h) : ((help_d + 2) & 3);
      help_d = d2;
      if (lane == 0) S.cmd[1] = d2 | (PROX ? 4 : 0) | (UNK ? 8 : 0);
      pair_bar<GROUP>(pbar);
      const int ch_a = sweep_once<PROX, UNK, false>(S, lane, d, w_prox, unk_w, L1, L2);
      pair_bar<GROUP>(pbar + 1);
      const int ch_b = S.chh[1];
      S.dirty[lane] |= S.dirty2[0][lane];
      __syncwarp();
      // a clean sweep only counts when nothing changed while it ran; a changing one counts for itself when the other was clean
      if (ch_a && ch_b) mask = 0u;
      else if (ch_a) mask = 1u << d;
      else if (ch_b) mask = 1u << d2;
      else mask |= (1u << d) | (1u << d2);
      ch = ch_a | ch_b;
      prim = (d + 2) & 3;
      n += 2;
    }
    else if (GROUP == 4)
    {
      // quad mode: all four directions at once, one warp each
      d = prim;
      if (lane >= 1 && lane < 4) S.cmd[lane] = ((d + lane) & 3) | (PROX ? 4 : 0) | (UNK ? 8 : 0);
      pair_bar<GROUP>(pbar);
      const int ch_a = sweep_once<PROX, UNK, false>(S, lane, d, w_prox, unk_w, L1, L2);
      pair_bar<GROUP>(pbar + 1);
      const int c1 = S.chh[1], c2 = S.chh[2], c3 = S.chh[3];
      S.dirty[lane] |= S.dirty2[0][lane] | S.dirty2[1][lane] | S.dirty2[2][lane];
      __syncwarp();
      const int nch = ch_a + c1 + c2 + c3;
      if (nch == 0) mask = 15u;
      else if (nch > 1) mask = 0u;
      else mask = ch_a ? (1u << d) : c1 ? (1u << ((d + 1) & 3)) : c2 ? (1u << ((d + 2) & 3)) : (1u << ((d + 3) & 3));
      ch = nch > 0;
      n += 4;
    }
    else
    {
      d = prim;
      ch = sweep_once<PROX, UNK, false>(S, lane, d, w_prox, unk_w, L1, L2);
      mask = ch ? (1u << d) : (mask | (1u << d));
      prim = (d + 1) & 3;  // round robin (trying the perpendiculars before the way back measured slower)
      ++n;
    }
    TRACE(lane, V.tile, 3, d | (ch << 2));
    n_changing += ch;
    last_ch = ch;
    if (ch) may_linger = can_linger;
    if (ch && (n_changing == (int)q.early || q.early == SW_EARLY_ALL))
    {
      if (n_pub++ > 0)  // publications beyond the reserved two: top the pending counts up first
      {
        if (lane == 0) atomicAdd(q.ctl + HC_PENDING, 8);
        *n_reserved += 8;
      }
      if (q.early == SW_EARLY_ALL) sweep_publish_border(S, V, q, lane, n_act, n_unused);
      else sweep_publish<false>(S, V, q, lane, n_act, n_unused);
      TRACE(lane, V.tile, 4, 0);
    }
    if (q.refresh == 1 && st == T_DIRTY)
    {
      const int dn = sweep_refresh_halo(S, V, lane);
      if (dn >= 0 && halo_improves<PROX, UNK>(S, lane, w_prox, unk_w, L1, L2))
      {
        mask = 0;
        prim = dn;
        if (dn & 1) pref_h = dn; else pref_v = dn;
        help_d = -1;
      }
    }
  }
  *n_sweeps = n;
  return n_changing > 0;
}

__device__ __forceinline__ void cp_async8(const unsigned dst_smem, const void* src)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst_smem), "l"(src));
}

#define SW_RESERVE 24  // pending counts reserved per visit: (two publications + the prefetch) x 8 neighbours

template <bool HAS_PB, int WARPS, int GROUP>
__global__ void __launch_bounds__(WARPS * 32, 1) heur_sweep_kernel(heur_grid single, const heur_grid* many, heur_queue q)
{
  extern __shared__ __align__(16) unsigned char sw_smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // pair mode (q.pair, latency mode only): warps 2p / 2p + 1 are the leader / helper of worker p and share one tile
  const int pbar = GROUP > 1 ? 1 + 2 * (warp / GROUP) : 0;  // named barriers pbar (go) and pbar + 1 (done) of the group
  sweep_tile_t<HAS_PB>& S = reinterpret_cast<sweep_tile_t<HAS_PB>*>(sw_smem)[warp / GROUP];
  const double L1 = 0.625, L2 = 0.625 * 1.4142135623730951;  // movement_distances_ (a_star.hpp:54-57)
  if (GROUP > 1 && (warp % GROUP))
  {
    sweep_helper<GROUP>(S, lane, pbar, warp % GROUP, L1, L2);
    return;
  }
  int cont = 0;

  for (;;)
  {
    const long long t0 = clock64();
    int e = cont;
    if (!e) e = q_pop_warp(q, lane);
    const long long t1 = clock64();
    cont = 0;
    if (e < 0)
    {
      if (GROUP > 1)  // send the helpers home
      {
        if (lane < 4) S.cmd[lane] = -1;
        pair_bar<GROUP>(pbar);
      }
      return;
    }
    sweep_visit V;
    V.gi = (e - 1) >> 20;
    const int gi = V.gi, tile = (e - 1) & 0xFFFFF;
#define HG(f) (many ? many[gi].f : single.f)
    const uint8_t* __restrict__ p_cls = HG(cls);
    const double* __restrict__ p_mc = HG(mc);
    const double* __restrict__ p_prox = HG(prox);
    double* __restrict__ gp = HG(g);
    int* state = HG(state);
    const int dim = HG(dim), tiles_x = HG(tiles_x), goal = HG(goal);
    const double w_prox = HG(w_prox), unk_w = HG(unk_w);
    const int has_prox = HAS_PB && __ldg(HG(tprox) + tile) == HG(epoch);
#undef HG
    const int tx = tile % tiles_x, ty = tile / tiles_x;
    const int ox = tx * SW_O - 1, oy = ty * SW_O - 1;
    TRACE(lane, tile, 1, 0);
    V.gp = gp; V.state = state; V.tile = tile; V.dim = dim; V.tiles_x = tiles_x; V.tx = tx; V.ty = ty; V.ox = ox; V.oy = oy;
    // ---- stage the 32x32 block (lane = x); every address is clamped into the grid so that all
    //      loads are unconditional and in flight together; out-of-grid cells are fixed up after.
    const int gx = ox + lane;
    const bool x_in = gx >= 0 && gx < dim;
    const int cx = min(max(gx, 0), dim - 1);
    // (a) the static planes go straight to shared memory (read-only during the kernel: L1 is fine);
    //     issued before the tile is marked RUNNING -- they do not depend on what neighbours write
    {
      const unsigned s_mc = (unsigned)__cvta_generic_to_shared(S.mc + lane), s_pb = (unsigned)__cvta_generic_to_shared(S.pb + lane);
#pragma unroll 8
      for (int r = 0; r < 32; ++r)
      {
        const int cy = min(max(oy + r, 0), dim - 1);
        const size_t id = (size_t)cy * dim + cx;
        cp_async8(s_mc + r * SW_P * 8, p_mc + id);
        if (HAS_PB && has_prox) cp_async8(s_pb + r * SW_P * 8, p_prox + id);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    }
#ifdef SW_DBG
    const long long ta = clock64();
#endif
    if (lane == 0)
    {
      // reserve the pending counts of the activations this visit may make NOW, long before any of
      // them can finish; the unused ones are returned at release, so HC_PENDING can never touch
      // zero while work exists
      atomicAdd(q.ctl + HC_PENDING, SW_RESERVE);
      // acquire: whoever moved the state word before us (activation = release fence + CAS) has its
      // cost-plane writes visible to the loads below; a writer that comes later sees RUNNING and
      // marks the tile dirty.  One L2 round trip instead of a full fence.sc.
      int old_state;  // not needed: the visit owns the tile whatever it was
      asm volatile("atom.acquire.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old_state) : "l"(state + tile), "r"((int)T_RUNNING) : "memory");
      (void)old_state;  // the visit owns the tile whatever the word held
      (void)old_state;
    }
    __syncwarp();
#ifdef SW_DBG
    const long long tb = clock64();
#endif
    // (b) the cost plane (L2: other tiles write it) and the cell class bytes through registers
    const int r_lo = max(0, -oy), r_hi = min(32, dim - oy);  // staged rows inside the grid
    const unsigned vm = x_in ? ((r_hi >= 32 ? 0xFFFFFFFFu : ((1u << r_hi) - 1u)) & ~((1u << r_lo) - 1u)) : 0u;
    unsigned uc = 0, bc = 0;
    bool reached = false;  // some owned cell already has a cost: the wave has been here before
    {
      double v_g[32];
      uint8_t v_c[32];
#pragma unroll
      for (int r = 0; r < 32; ++r)
      {
        const int cy = min(max(oy + r, 0), dim - 1);
        const size_t id = (size_t)cy * dim + cx;
        v_g[r] = __ldcg(gp + id);
        v_c[r] = __ldg(p_cls + id);
      }
      __syncwarp();  // keeps every load ahead of the first use: one memory round trip, not several
#pragma unroll
      for (int r = 0; r < 32; ++r)
      {
        const double gv = ((vm >> r) & 1u) ? v_g[r] : INFINITY;
        const int idx = r * SW_P + lane;
        S.g[idx] = gv;
        if (r >= 1 && r <= SW_O) reached |= gv < INFINITY;
        uc |= (unsigned)(v_c[r] & 1) << r;
        bc |= (unsigned)((v_c[r] >> 1) & 1) << r;
      }
    }
    uc &= vm;
#ifdef SW_DBG
    const long long tc = clock64();
#endif
    // blocked: off-grid, OCC / geofence, the halo ring (it belongs to the neighbour tiles), the root (keeps cost 0)
    bc |= ~vm | 0x80000001u | ((lane == 0 || lane == 31) ? 0xFFFFFFFFu : 0u);
    const int glx = goal % dim - ox, gly = goal / dim - oy;
    if (lane == glx && gly >= 0 && gly < 32) bc |= 1u << gly;
    const bool must_run = glx >= 1 && glx <= SW_O && gly >= 1 && gly <= SW_O;  // the root's tile is not a fixed point of its old halo
    S.dirty[lane] = 0;
    if (GROUP > 1 && lane == 0)
    {
      S.par_w = w_prox;
      S.par_u = unk_w;
    }
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
    // (c) static planes arrived: blocked cells get mc = +inf (off-grid cells hold clamped copies of
    //     prox: harmless, they are blocked sources with g = +inf)
    asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll 8
    for (int r = 0; r < 32; ++r)
      if ((bc >> r) & 1u) S.mc[r * SW_P + lane] = INFINITY;
    if (HAS_PB && has_prox)
    {
      // prox_cost = h_prox x astar_prox_cost_ once per staged cell instead of three times per line step (the product is
      // rounded before it is added either way)
#pragma unroll 8
      for (int r = 0; r < 32; ++r) S.pb[r * SW_P + lane] = S.pb[r * SW_P + lane] * w_prox;
    }
    __syncwarp();
#ifdef SW_DBG
    const long long td = clock64();
#endif
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
      // pair mode: the helper's perpendicular direction also leaves the side with the smaller halo values
      // (bits 2-3: rightwards 1 / leftwards 3, bits 4-5: downwards 0 / upwards 2)
      d0 |= ((m_l <= m_r) ? 1 : 3) << 2;
      d0 |= ((m_t <= m_b) ? 0 : 2) << 4;
    }
    __syncwarp();
    const long long t2 = clock64();

    const bool fresh = !__any_sync(SW_FULL, reached && lane >= 1 && lane <= SW_O);
    int any, n_sweeps = 0, n_act = 0, n_unused = 0, n_reserved = SW_RESERVE;
    if constexpr (!HAS_PB)
    {
      if (unk_w != 0) any = sweep_tile_loop<false, true, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
      else any = sweep_tile_loop<false, false, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
    }
    else if (has_prox)
    {
      if (unk_w != 0) any = sweep_tile_loop<true, true, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
      else any = sweep_tile_loop<true, false, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
    }
    else
    {
      if (unk_w != 0) any = sweep_tile_loop<false, true, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
      else any = sweep_tile_loop<false, false, GROUP>(S, V, q, lane, d0, must_run, fresh, w_prox, unk_w, L1, L2, &n_sweeps, &n_act, &n_unused, &n_reserved, pbar);
    }
    const long long t3 = clock64();

    if (any) cont = sweep_publish<true>(S, V, q, lane, &n_act, &n_unused);
    TRACE(lane, tile, 7, any);
    if (lane == 0)
    {
      atomicAdd(q.ctl + HC_VISITS, 1);
      atomicAdd(q.ctl + HC_INNER, n_sweeps);
      if (any) atomicAdd(q.ctl + HC_IMPROVING, 1);
      // release the tile: re-queue it if a neighbour wrote into our halo meanwhile
      int dec = (n_reserved - n_act) + n_unused + 1;
      if (atomicCAS(state + tile, T_RUNNING, T_IDLE) == T_DIRTY)
      {
        atomicExch(state + tile, T_QUEUED);
        q_push(q, e);
        dec -= 1;
      }
      if (dec) atomicSub(q.ctl + HC_PENDING, dec);
      unsigned long long* cyc = reinterpret_cast<unsigned long long*>(q.ctl + HC_CYC);
      const long long t4 = clock64();
#ifdef SW_DBG
      atomicAdd(cyc + 0, (unsigned long long)(ta - t1));
      atomicAdd(cyc + 1, (unsigned long long)(tb - ta));
      atomicAdd(cyc + 2, (unsigned long long)(tc - tb));
      atomicAdd(cyc + 3, (unsigned long long)(td - tc));
#else
      atomicAdd(cyc + 0, (unsigned long long)(t1 - t0));
      atomicAdd(cyc + 1, (unsigned long long)(t2 - t1));
      atomicAdd(cyc + 2, (unsigned long long)(t3 - t2));
      atomicAdd(cyc + 3, (unsigned long long)(t4 - t3));
#endif
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
//   B  one thread per terminal cell (A appended them to a list): walk terminal -> parent -> that
//      block's terminal -> ... to the root, one L2 read pair per block crossed (a few dozen),
//      summing the block-local step counts;
//   C  every cell: its local counts + its terminal's total.
// ---------------------------------------------------------------------------------------------
#define FB 32            // finish block edge
#define FS (FB + 2)      // staged edge incl. halo
#define FIN_SLOTS 128    // terminal-list slots per block: only its 124 border cells can have their parent outside

// scratch word of a cell: displacement of its block's terminal cell (dx + 32, dy + 32: 6 bits each),
// straight steps to it (10 bits), diagonal steps (10 bits)
__device__ __forceinline__ unsigned fin_pack(int dx, int dy, int cs, int cd)
{
  return (unsigned)(dx + 32) | ((unsigned)(dy + 32) << 6) | ((unsigned)cs << 12) | ((unsigned)cd << 22);
}
__device__ __forceinline__ int fin_target(unsigned s, int cell, int dim) { return cell + ((int)((s >> 6) & 63) - 32) * dim + (int)(s & 63) - 32; }

// bit i set: candidate of motion index i reproduces g(n) bit for bit.  Predecessor offsets
// p = n - motion_[i] (a_star.hpp:52-54): (dx, dy) = (-1,0) (0,-1) (1,0) (0,1) (1,1) (1,-1) (-1,1) (-1,-1)
template <bool PROX, bool UNK>
__device__ __forceinline__ unsigned fin_match(const double* sg, const double* sb, const uint8_t* su, const int me, const double gn,
                                              const double ln1, const double ln2, const double unk_w)
{
  constexpr int off[8] = { -1, -FS, 1, FS, FS + 1, -FS + 1, FS - 1, -FS - 1 };
  unsigned match = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
  {
    const int q = me + off[i];
    double cand = sg[q] + (i < 4 ? ln1 : ln2);
    if (PROX) cand = cand + sb[q];
    if (UNK) cand = cand + (su[q] ? unk_w : 0.0);
    match |= (cand == gn ? 1u : 0u) << i;
  }
  return match;
}

__global__ void __launch_bounds__(FB * FB, 2) heur_finish_a_kernel(heur_args single, const heur_args* many)
{
  __shared__ double sg[FS * FS];
  __shared__ double sb[FS * FS];
  __shared__ uint8_t su[FS * FS];
  __shared__ int s_cnt[2][FB];   // per warp: closed cells, terminal cells
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, nbx = (dim + FB - 1) / FB;
  if ((int)blockIdx.x >= nbx * nbx) return;  // batches of unequal grids
  const int bx = blockIdx.x % nbx, by = blockIdx.x / nbx;
  const int ox = bx * FB - 1, oy = by * FB - 1;
  const double w_prox = a.w_prox, unk_w = a.unk_w;
  int any_prox = 0;
  // staging without a division: every thread loads its own cell (row-wise coalesced) and the first 4 * FB + 4 threads
  // one cell of the halo ring; the cell's own movement cost goes out with them (one memory round trip for all).
  // a.prox == nullptr: the proximity plane is known to be zero (prox_free) and is not read.
  const bool with_prox = a.prox != nullptr;
  const int own_x = bx * FB + (int)(threadIdx.x % FB), own_y = by * FB + (int)(threadIdx.x / FB);
  const double m_own = (own_x < dim && own_y < dim) ? __ldg(a.mc + (size_t)own_y * dim + own_x) : 0.0;
  {
    const int k = threadIdx.x;
    // ring position of thread k: top row, bottom row, left column, right column (corners belong to the rows)
    int rx = -1, ry = -1;
    if (k < FS) { rx = k; ry = 0; }
    else if (k < 2 * FS) { rx = k - FS; ry = FS - 1; }
    else if (k < 2 * FS + FB) { rx = 0; ry = k - 2 * FS + 1; }
    else if (k < 2 * FS + 2 * FB) { rx = FS - 1; ry = k - 2 * FS - FB + 1; }
    // own cell first, then (warps 0..4 only: a real branch, the other warps skip it) the ring cell; all loads of a
    // thread are issued before the first shared store
    const int px0 = (int)(threadIdx.x % FB) + 1, py0 = (int)(threadIdx.x / FB) + 1;
    double g0 = INFINITY, b0 = 0.0, g1 = INFINITY, b1 = 0.0;
    uint8_t u0 = 0, u1 = 0;
    {
      const int x = ox + px0, y = oy + py0;
      if (x < dim && y < dim)  // own cells have x, y >= 0
      {
        const size_t id = (size_t)y * dim + x;
        g0 = __ldcg(a.g + id);
        if (with_prox) b0 = __ldg(a.prox + id);
        u0 = __ldg(a.grid + id);
      }
    }
    if (rx >= 0)
    {
      const int x = ox + rx, y = oy + ry;
      if (x >= 0 && y >= 0 && x < dim && y < dim)
      {
        const size_t id = (size_t)y * dim + x;
        g1 = __ldcg(a.g + id);
        if (with_prox) b1 = __ldg(a.prox + id);
        u1 = __ldg(a.grid + id);
      }
    }
    {
      const int i = py0 * FS + px0;
      const double b = b0 * w_prox;
      sg[i] = g0;
      if (with_prox) sb[i] = b;
      su[i] = u0 == GA_UNKNOWN;
      any_prox |= (b != 0.0);
    }
    if (rx >= 0)
    {
      const int i = ry * FS + rx;
      const double b = b1 * w_prox;
      sg[i] = g1;
      if (with_prox) sb[i] = b;
      su[i] = u1 == GA_UNKNOWN;
      any_prox |= (b != 0.0);
    }
  }
  const int has_prox = __syncthreads_or(any_prox);
  const int lx = threadIdx.x % FB, ly = threadIdx.x / FB, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nx = ox + 1 + lx, ny = oy + 1 + ly;
  const int me = (ly + 1) * FS + lx + 1;
  const bool in = nx < dim && ny < dim;
  const int n = ny * dim + nx;
  const double gn = sg[me];
  int32_t par = -2;
  uint8_t fl = 0;
  int lpar = -1;
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
      const double ln1 = a.L1 * m_own, ln2 = a.L2 * m_own;
      // all eight candidates, branch-free (a +inf / off-grid predecessor gives +inf or NaN: no match);
      // the prox and unknown terms only where the block has any (adding 0.0 changes nothing)
      unsigned match;
      if (has_prox) match = (unk_w != 0) ? fin_match<true, true>(sg, sb, su, me, gn, ln1, ln2, unk_w)
                                         : fin_match<true, false>(sg, sb, su, me, gn, ln1, ln2, unk_w);
      else match = (unk_w != 0) ? fin_match<false, true>(sg, sb, su, me, gn, ln1, ln2, unk_w)
                                : fin_match<false, false>(sg, sb, su, me, gn, ln1, ln2, unk_w);
      if (match)
      {
        const int i = __ffs(match) - 1;  // canonical parent: smallest motion index
        // predecessor offset p = n - motion_[i]: dx + 1, dy + 1 packed two bits per index
        constexpr unsigned DXC = (0u << 0) | (1u << 2) | (2u << 4) | (1u << 6) | (2u << 8) | (2u << 10) | (0u << 12) | (0u << 14);
        constexpr unsigned DYC = (1u << 0) | (0u << 2) | (1u << 4) | (2u << 6) | (2u << 8) | (0u << 10) | (2u << 12) | (0u << 14);
        const int dx = (int)((DXC >> (2 * i)) & 3u) - 1, dy = (int)((DYC >> (2 * i)) & 3u) - 1;
        par = (ny + dy) * dim + (nx + dx);
        if (su[me + dy * FS + dx]) fl |= GA_H_UNKNOWN;  // NodeDisc::is_unknown_ = unknown-ness of the parent, :253,268
        const int plx = lx + dx, ply = ly + dy;
        if (plx >= 0 && ply >= 0 && plx < FB && ply < FB) lpar = (ply * FB + plx) | (i < 4 ? 0 : 1024);
      }
    }
  }
  if (in)
  {
    a.parent[n] = par;
    a.flags[n] = fl;
  }
  // counters: closed cells, and the block's terminal cells appended to the grid's terminal list
  const bool term = par >= 0 && lpar < 0;
  const unsigned m_closed = __ballot_sync(0xFFFFFFFFu, fl != 0), m_term = __ballot_sync(0xFFFFFFFFu, term);
  if (lane == 0)
  {
    s_cnt[0][warp] = __popc(m_closed);
    s_cnt[1][warp] = __popc(m_term);
  }
  __syncthreads();
  if (warp == 0)
  {
    const int nc = s_cnt[0][lane], nt = s_cnt[1][lane];
    int incl = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
      const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += v;
    }
    const int tot_c = __reduce_add_sync(0xFFFFFFFFu, nc);
    s_cnt[1][lane] = incl - nt;  // exclusive offset of the warp
    if (lane == 31)
    {
      if (tot_c) atomicAdd(a.closed, tot_c);  // no return value: fire and forget
      a.tcount[blockIdx.x] = incl;            // the block's slice of the terminal list (no atomic round trip)
    }
  }
  __syncthreads();
  if (term)
  {
    a.tlist[blockIdx.x * FIN_SLOTS + s_cnt[1][warp] + __popc(m_term & ((1u << lane) - 1u))] = n;
    a.tot[n] = make_int2(-1, -1);  // "not known yet" for finish B's shortcut
  }
  // chain to the block's terminal cell by pointer jumping in shared memory: every cell holds (pointer, straight steps,
  // diagonal steps) and replaces it by the composition with its pointer's entry until every pointer is a terminal --
  // log2(chain length) rounds instead of one hop per step (a warp used to wait for its longest chain: ~360 of the
  // ~560 warp instructions per cell).  Parent pointers cannot form a cycle while every edge weight is positive
  // (ga_set_astar_params and ga_upload enforce that); the round cap keeps a corrupted plane from hanging the kernel.
  unsigned* sj = reinterpret_cast<unsigned*>(sb);  // the prox plane is not needed any more
  __syncthreads();
  {
    const unsigned diag = lpar >= 0 ? (unsigned)(lpar >> 10) : 0u;
    sj[threadIdx.x] = lpar >= 0 ? ((unsigned)(lpar & 1023) | ((1u - diag) << 10) | (diag << 21)) : (unsigned)threadIdx.x;
  }
  __syncthreads();
  for (int round = 0; round < 11; ++round)
  {
    const unsigned w = sj[threadIdx.x];
    const unsigned wp = sj[w & 1023u];
    const bool changed = (wp & 1023u) != (w & 1023u);
    const unsigned nw = (wp & 1023u) | ((w & ~1023u) + (wp & ~1023u));  // steps add field by field (each < 1024: no carry across)
    __syncthreads();
    sj[threadIdx.x] = nw;
    if (!__syncthreads_or(changed)) break;
  }
  const unsigned fin = sj[threadIdx.x];
  const int cur = (int)(fin & 1023u), cs = (int)((fin >> 10) & 2047u), cd = (int)(fin >> 21);
  if (in) a.scratch[n] = fin_pack(cur % FB - lx, cur / FB - ly, cs, cd);
}

// one thread per terminal cell: terminal -> parent -> that block's terminal -> ... -> root.  A walk ends
// early at a terminal whose total another thread has already written (the threads nearer the root
// finish first), so the longest chain costs about half its hops.
__global__ void __launch_bounds__(FIN_SLOTS) heur_finish_b_kernel(heur_args single, const heur_args* many)
{
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, goal = a.goal, nbx = (dim + FB - 1) / FB;
  if ((int)blockIdx.x >= nbx * nbx || (int)threadIdx.x >= a.tcount[blockIdx.x]) return;  // one CTA = one finish-A block's terminals
  const int c = a.tlist[blockIdx.x * FIN_SLOTS + threadIdx.x];
  const long long* tot64 = reinterpret_cast<const long long*>(a.tot);
  int cur = c, cs = 0, cd = 0;
  for (int hops = 0; hops < nbx * nbx * 4 + 8; ++hops)  // a chain crosses every block at most a few times
  {
    const int p = __ldg(a.parent + cur);
    if (p < 0) break;  // no parent (cannot happen for a terminal of a consistent plane)
    const long long known = cur != c ? __ldcg(tot64 + cur) : -1;  // same dependency level as the parent load
    if ((int)(known & 0xFFFFFFFFll) >= 0)
    {
      cs += (int)(known & 0xFFFFFFFFll);
      cd += (int)(known >> 32);
      break;
    }
    const int d = p - cur;
    if (d == 1 || d == -1 || d == dim || d == -dim) ++cs; else ++cd;  // the step terminal -> parent
    if (p == goal) break;
    const unsigned s = __ldg(a.scratch + p);  // parent -> its block's terminal
    cs += (s >> 12) & 1023;
    cd += (s >> 22) & 1023;
    cur = fin_target(s, p, dim);
    if (cur == goal) break;
  }
  __stcg(reinterpret_cast<long long*>(a.tot) + c, ((long long)cd << 32) | (unsigned)cs);  // one 8-byte store: never seen half written
}

__device__ __forceinline__ double fin_dist(const heur_args& a, int c, unsigned flag, unsigned s, int dim, uint32_t* steps)
{
  *steps = 0;
  if (!(flag & GA_H_CLOSED)) return INFINITY;
  int cs = (s >> 12) & 1023, cd = (s >> 22) & 1023;
  const int t = fin_target(s, c, dim);
  if (t != a.goal)
  {
    const int2 tt = a.tot[t];
    cs += tt.x;
    cd += tt.y;
  }
  *steps = (uint32_t)min(cs, 16383) | ((uint32_t)min(cd, 16383) << 14);
  return (double)cs * a.L1 + (double)cd * a.L2;  // NodeDisc::cost_dist_
}

// four cells per thread: 4-byte / 16-byte loads of the flag and scratch planes, 16-byte stores of cost_dist
__global__ void __launch_bounds__(256) heur_finish_c_kernel(heur_args single, const heur_args* many)
{
  const heur_args& a = many ? many[blockIdx.y] : single;
  const int dim = a.dim, ncell = dim * dim;
  const int c = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
  if (c >= ncell) return;
  if (c + 4 <= ncell)
  {
    const uchar4 f = *reinterpret_cast<const uchar4*>(a.flags + c);
    const uint4 s = *reinterpret_cast<const uint4*>(a.scratch + c);
    uint4 st;
    const double d0 = fin_dist(a, c, f.x, s.x, dim, &st.x), d1 = fin_dist(a, c + 1, f.y, s.y, dim, &st.y);
    const double d2 = fin_dist(a, c + 2, f.z, s.z, dim, &st.z), d3 = fin_dist(a, c + 3, f.w, s.w, dim, &st.w);
    *reinterpret_cast<double2*>(a.dist + c) = make_double2(d0, d1);
    *reinterpret_cast<double2*>(a.dist + c + 2) = make_double2(d2, d3);
    *reinterpret_cast<uint4*>(a.steps + c) = st;
  }
  else
    for (int k = c; k < ncell; ++k) a.dist[k] = fin_dist(a, k, a.flags[k], a.scratch[k], dim, a.steps + k);
}

// ---------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------

void ga_heur_free(ga_ctx* c)
{
  void** ptrs[] = { (void**)&c->d_tile_active, (void**)&c->d_hctl, (void**)&c->d_jump, (void**)&c->d_cnt,
                    (void**)&c->d_hring, (void**)&c->d_hcls, (void**)&c->d_bring, (void**)&c->d_bdesc };
  for (void** p : ptrs)
  {
    if (*p) cudaFree(*p);
    *p = nullptr;
  }
  c->bring_cap = c->bdesc_cap = 0;
  if (c->ev_batch) cudaEventDestroy(c->ev_batch);
  c->ev_batch = nullptr;
}

static unsigned ring_capacity(size_t n_tiles)
{
  unsigned cap = 64;
  while (cap < 4u * (unsigned)n_tiles + 4096u) cap <<= 1;  // every tile at most once + one ticket per persistent worker
  return cap;
}

int ga_heur_alloc(ga_ctx* c)
{
  const size_t na = (size_t)c->astar_dim * c->astar_dim;
  c->tiles_x = (c->astar_dim + SW_O - 1) / SW_O;
  c->n_tiles = c->tiles_x * c->tiles_x;
  GA_REQUIRE(c, c->n_tiles < (1 << 20), GA_ERR_ARG, "astar grid too large for the tile ring encoding");
  c->hring_cap = (int)ring_capacity((size_t)c->n_tiles);
  // per-tile state words + prox stamps (stamp == build epoch: the tile's block carries a prox cost)
  GA_CUDA(c, cudaMalloc(&c->d_tile_active, sizeof(int) * 2 * (size_t)c->n_tiles));
  GA_CUDA(c, cudaMemsetAsync(c->d_tile_active, 0, sizeof(int) * 2 * (size_t)c->n_tiles, c->stream));
  c->heur_epoch = 0;
  GA_CUDA(c, cudaMalloc(&c->d_hring, sizeof(int) * (size_t)c->hring_cap));
  GA_CUDA(c, cudaMemsetAsync(c->d_hring, 0, sizeof(int) * (size_t)c->hring_cap, c->stream));
  GA_CUDA(c, cudaMalloc(&c->d_hctl, sizeof(int) * HC_WORDS));
  GA_CUDA(c, cudaMemsetAsync(c->d_hctl, 0, sizeof(int) * HC_WORDS, c->stream));
  GA_CUDA(c, cudaMalloc(&c->d_hcls, na));
  GA_CUDA(c, cudaMalloc(&c->d_jump, sizeof(int32_t) * (na + (size_t)((c->astar_dim + 31) / 32) * ((c->astar_dim + 31) / 32) * 129)));  // scratch words + per-block terminal slots and counts
  GA_CUDA(c, cudaMalloc(&c->d_cnt, sizeof(int32_t) * 2 * na));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<true, SW_WARPS, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS * sizeof(sweep_tile_t<true>))));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<false, SW_WARPS_LEAN, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS_LEAN * sizeof(sweep_tile_t<false>))));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<true, SW_WARPS, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS / 2 * sizeof(sweep_tile_t<true>))));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<false, SW_WARPS_LEAN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS_LEAN / 2 * sizeof(sweep_tile_t<false>))));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<true, SW_WARPS, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS / 4 * sizeof(sweep_tile_t<true>))));
  GA_CUDA(c, cudaFuncSetAttribute(heur_sweep_kernel<false, SW_WARPS_LEAN, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(SW_WARPS_LEAN / 4 * sizeof(sweep_tile_t<false>))));
  return GA_OK;
}

// no cell of the context's h_prox plane can hold a cost: the plane is all zero since ga_patch_reset (nothing was
// uploaded into it, no potential field built, its device pointer not handed out), or the weight is zero
static bool prox_free(const ga_ctx* c) { return c->prox_zero || c->ap.astar_prox_cost == 0.0; }

static void make_descs(ga_ctx* c, int slot, int goal_x, int goal_y, int* closed, heur_grid* G, heur_args* a)
{
  ++c->heur_epoch;
  G->grid = c->d_astar;
  G->fence = c->d_fence;
  G->mc = c->d_mc;
  G->prox = prox_free(c) ? nullptr : c->d_prox;  // nullptr: known to be all zero, never read
  G->g = c->d_hcost[slot];
  G->cls = c->d_hcls;
  G->state = c->d_tile_active;
  G->tprox = c->d_tile_active + c->n_tiles;
  G->epoch = c->heur_epoch;
  G->dim = c->astar_dim;
  G->tiles_x = c->tiles_x;
  G->n_tiles = c->n_tiles;
  G->edge = SW_O;
  G->goal = goal_y * c->astar_dim + goal_x;
  G->w_prox = c->ap.astar_prox_cost;
  G->unk_w = c->ap.unknown_cost_w;
  a->grid = c->d_astar;
  a->mc = c->d_mc;
  a->prox = G->prox;
  a->g = c->d_hcost[slot];
  a->parent = c->d_hparent[slot];
  a->dist = c->d_hdist[slot];
  a->flags = c->d_hflags[slot];
  a->closed = closed;
  a->scratch = reinterpret_cast<unsigned*>(c->d_jump);
  a->tlist = c->d_jump + (size_t)c->astar_dim * c->astar_dim;
  a->tcount = a->tlist + (size_t)((c->astar_dim + FB - 1) / FB) * ((c->astar_dim + FB - 1) / FB) * FIN_SLOTS;
  a->tot = reinterpret_cast<int2*>(c->d_cnt);
  a->steps = c->d_hsteps[slot];
  c->rec_valid[slot] = true;  // a device build follows (a reference-order replay clears it again)
  a->dim = c->astar_dim;
  a->goal = G->goal;
  a->w_prox = G->w_prox;
  a->unk_w = G->unk_w;
  // movement_distances_ (a_star.hpp:54-57): ASTAR_RES and ASTAR_RES * sqrt(2)
  a->L1 = 0.625;
  a->L2 = 0.625 * sqrt(2.0);
}

// init -> sweep relaxation -> finish A/B/C for n grids (descriptors in device memory when n > 1)
// Scheduling mode of the sweep kernel.  Few grids (latency mode): the machine is mostly idle, so
// every changing sweep is published at once, neighbours are staged ahead of the wave and converged
// tiles stay staged for a moment.  Many grids (throughput mode): no speculative work, the four sweeps
// of a tile run in lock step.  GA_SW_* (development) override either.
static void queue_knobs(const ga_ctx* c, int n_grids, heur_queue* q)
{
  const bool lat = n_grids <= 4;
  q->early = c->sw_early >= 0 ? (unsigned)c->sw_early : (lat ? SW_EARLY_ALL : 1u);
  q->round = c->sw_round >= 0 ? (unsigned)c->sw_round : (lat ? 0u : 1u);
  q->refresh = c->sw_refresh >= 0 ? (unsigned)c->sw_refresh : (lat ? 2u : 1u);
  q->linger = c->sw_linger >= 0 ? (unsigned)c->sw_linger : (lat ? 8000u : 0u);
  q->prefetch = c->sw_prefetch >= 0 ? (unsigned)(c->sw_prefetch & 1) : (lat ? 1u : 0u);
  // warps per tile in single builds: 4 (leader + three helpers, one direction each); stage-ahead knob + 4: no helpers, + 8: pairs
  q->pair = (lat && q->round == 0) ? ((c->sw_prefetch >= 0 && (c->sw_prefetch & 4)) ? 0u : (c->sw_prefetch >= 0 && (c->sw_prefetch & 8)) ? 2u : 4u) : 0u;
}

static int launch_build(ga_ctx* c, int n, const heur_grid& G0, const heur_args& a0, const heur_grid* dG, const heur_args* dA,
                        const heur_queue& q, int max_dim, size_t total_tiles, bool lean)
{
  GA_CUDA(c, cudaMemsetAsync(q.ctl, 0, sizeof(int) * HC_WORDS, c->stream));  // ring cursors, pending count, statistics
  {
    ga_launch_scope ls(c, GA_K_HEUR_INIT);
    // two cells per thread for one grid (latency bound otherwise), fewer blocks per grid in large batches
    const int per_grid = n > 4 ? c->sm_count : (max_dim * max_dim / 4 + 255) / 256;  // one step per thread for one grid (latency bound)
    heur_init_kernel<<<dim3(per_grid, n), 256, 0, c->stream>>>(G0, dG, q);
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_HEUR_RELAX);
    // no grid of the launch carries a proximity cost: the tiles need no prox plane and 12 workers fit an SM
    const int nw = lean ? (c->sw_warps == SW_WARPS ? SW_WARPS_LEAN : c->sw_warps) : c->sw_warps;
    heur_queue qk = q;
    if (qk.pair && nw % (int)qk.pair) qk.pair = (nw & 1) ? 0 : 2;  // groups need a multiple of their size
    const int grp = qk.pair ? (int)qk.pair : 1;
    const int workers = nw / grp;
    const size_t want = (total_tiles + workers - 1) / workers;
    const int g = (size_t)c->sm_count < want ? c->sm_count : (int)want;
#define SW_LAUNCH(G)                                                                                                                  \
  do                                                                                                                                  \
  {                                                                                                                                   \
    if (lean) heur_sweep_kernel<false, SW_WARPS_LEAN, G><<<g, nw * 32, workers * sizeof(sweep_tile_t<false>), c->stream>>>(G0, dG, qk); \
    else heur_sweep_kernel<true, SW_WARPS, G><<<g, nw * 32, workers * sizeof(sweep_tile_t<true>), c->stream>>>(G0, dG, qk);           \
  } while (0)
    if (grp == 4) SW_LAUNCH(4);
    else if (grp == 2) SW_LAUNCH(2);
    else SW_LAUNCH(1);
#undef SW_LAUNCH
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_HEUR_FINISH, 3);
    const int nb = (max_dim + FB - 1) / FB, cells = ((max_dim * max_dim + 3) / 4 + 255) / 256;
    heur_finish_a_kernel<<<dim3(nb * nb, n), FB * FB, 0, c->stream>>>(a0, dA);
    heur_finish_b_kernel<<<dim3(nb * nb, n), FIN_SLOTS, 0, c->stream>>>(a0, dA);
    heur_finish_c_kernel<<<dim3(cells, n), 256, 0, c->stream>>>(a0, dA);
  }
  GA_CUDA(c, cudaGetLastError());
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

  heur_grid G;
  heur_args a;
  make_descs(c, slot, goal_x, goal_y, c->d_hctl + HC_CLOSED, &G, &a);
  heur_queue q;
  q.ring = c->d_hring;
  q.mask = (unsigned)c->hring_cap - 1;
  q.ctl = c->d_hctl;
  q.poll_ns = (unsigned)c->sw_poll_ns;
  queue_knobs(c, 1, &q);
  const int rc = launch_build(c, 1, G, a, nullptr, nullptr, q, dim, (size_t)c->n_tiles, prox_free(c));
  if (rc != GA_OK) return rc;
  if (n_closed)
  {
    GA_CUDA(c, cudaMemcpyAsync(n_closed, c->d_hctl + HC_CLOSED, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return GA_OK;
}

// scheduling knobs of the sweep kernel (tests: every schedule gives the same fixed point); -1 = automatic
extern "C" int ga_set_heur_knobs(ga_ctx* c, const int* k7)
{
  if (!c || !k7) return GA_ERR_ARG;
  c->sw_warps = k7[0] >= 1 && k7[0] <= SW_WARPS ? k7[0] : SW_WARPS;
  c->sw_poll_ns = k7[1] >= 0 ? k7[1] : 20;
  c->sw_early = k7[2];
  c->sw_round = k7[3];
  c->sw_refresh = k7[4];
  c->sw_linger = k7[5];
  c->sw_prefetch = k7[6];
  return GA_OK;
}

extern "C" int ga_build_heuristic_ex(ga_ctx* c, int slot, const ga_heur_query* q, int* n_closed)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, q != nullptr, GA_ERR_ARG, "query is NULL");
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_ap, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 (guidance) or 1 (path)");
  const int dim = c->astar_dim;
  GA_REQUIRE(c, q->goal_x >= 0 && q->goal_y >= 0 && q->goal_x < dim && q->goal_y < dim, GA_ERR_ARG, "goal outside the astar grid");
  if (!(q->flags & GA_HEUR_REFORDER))
  {
    GA_REQUIRE(c, (q->flags & GA_HEUR_EARLY_EXIT) == 0, GA_ERR_ARG,
               "GA_HEUR_EARLY_EXIT is a property of the reference's closing order: combine it with GA_HEUR_REFORDER");
    if (c->near_slot == slot) c->near_slot = -1;
    return ga_build_heuristic(c, slot, q->goal_x, q->goal_y, n_closed);
  }
  GA_REQUIRE(c, q->start_x >= 0 && q->start_y >= 0 && q->start_x < dim && q->start_y < dim, GA_ERR_ARG,
             "start outside the astar grid");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int rc = ga_heur_reforder_build(c, slot, q, n_closed);
  if (rc != GA_OK) return rc;
  if (q->flags & GA_HEUR_ONLY_NEAR)
  {
    c->near_slot = slot;
    c->near_start[0] = q->start_x;
    c->near_start[1] = q->start_y;
  }
  else if (c->near_slot == slot)
    c->near_slot = -1;
  return GA_OK;
}

// AStar::nodes_near_goal_ (a_star.cpp:182-193)
extern "C" int ga_heur_near_nodes(ga_ctx* c, int slot, int start_x, int start_y, int* ids, double* dist, int cap, int* n)
{
  if (!c || !n) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0, GA_ERR_STATE, "no patch");
  GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 or 1");
  if (c->near_slot == slot && c->near_start[0] == start_x && c->near_start[1] == start_y)
  {
    const int32_t* rid;
    const double* rd;
    int rn;
    ga_heur_reforder_near(c, &rid, &rd, &rn);
    *n = rn;
    for (int i = 0; i < rn && i < cap; ++i)
    {
      if (ids) ids[i] = rid[i];
      if (dist) dist[i] = rd[i];
    }
    return GA_OK;
  }
  // derive the list from the slot's flag plane: the 19 x 19 box around the start holds every cell with dist < 10
  const int dim = c->astar_dim;
  const int x0 = start_x - 9 < 0 ? 0 : start_x - 9, y0 = start_y - 9 < 0 ? 0 : start_y - 9;
  const int x1 = start_x + 9 >= dim ? dim - 1 : start_x + 9, y1 = start_y + 9 >= dim ? dim - 1 : start_y + 9;
  int cnt = 0;
  if (x1 >= x0 && y1 >= y0)
  {
    const int w = x1 - x0 + 1, hgt = y1 - y0 + 1;
    uint8_t box[19 * 19];
    GA_CUDA(c, cudaSetDevice(c->device));
    GA_CUDA(c, cudaMemcpy2DAsync(box, (size_t)w, c->d_hflags[slot] + (size_t)y0 * dim + x0, (size_t)dim, (size_t)w, (size_t)hgt,
                                 cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
    for (int y = 0; y < hgt; ++y)
      for (int x = 0; x < w; ++x)
      {
        if (!(box[y * w + x] & GA_H_CLOSED)) continue;
        const double xd = (x0 + x) - start_x, yd = (y0 + y) - start_y;
        const double d = fabs(sqrt(xd * xd + yd * yd));
        if (!(d < 10)) continue;
        if (cnt < cap)
        {
          if (ids) ids[cnt] = (y0 + y) * dim + (x0 + x);
          if (dist) dist[cnt] = d;
        }
        ++cnt;
      }
  }
  *n = cnt;
  return GA_OK;
}

extern "C" int ga_heur_closing_order(ga_ctx* c, int* ids, int cap, int* n)
{
  if (!c || !n) return GA_ERR_ARG;
  const int32_t* o;
  int cnt;
  ga_heur_reforder_order(c, &o, &cnt);
  *n = cnt;
  for (int i = 0; i < cnt && i < cap; ++i) ids[i] = o[i];
  return GA_OK;
}

// n independent grids (one context each, same device) as ONE launch chain on the first context's
// stream: the wavefronts of all grids share the ticket ring and fill the SMs together.
extern "C" int ga_build_heuristic_batch(ga_ctx* const* ctxs, int n, int slot, const int* goal_xy, int* n_closed)
{
  if (!ctxs || n < 1 || !ctxs[0]) return GA_ERR_ARG;
  ga_ctx* c = ctxs[0];
  GA_REQUIRE(c, goal_xy != nullptr, GA_ERR_ARG, "goal_xy is NULL");
  GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 (guidance) or 1 (path)");
  GA_REQUIRE(c, n <= 2047, GA_ERR_ARG, "at most 2047 grids per batch (ring entry encoding)");
  size_t total_tiles = 0;
  int max_dim = 0;
  bool lean = true;
  for (int i = 0; i < n; ++i)
  {
    ga_ctx* k = ctxs[i];
    GA_REQUIRE(c, k != nullptr, GA_ERR_ARG, "NULL context in the batch");
    GA_REQUIRE(c, k->device == c->device, GA_ERR_ARG, "all contexts of a batch must live on one device");
    GA_REQUIRE(c, k->patch_dim > 0 && k->have_ap, GA_ERR_STATE, "a context of the batch has no patch / params");
    for (int j = 0; j < i; ++j) GA_REQUIRE(c, ctxs[j] != k, GA_ERR_ARG, "a context appears twice in the batch");
    const int gx = goal_xy[2 * i], gy = goal_xy[2 * i + 1];
    GA_REQUIRE(c, gx >= 0 && gy >= 0 && gx < k->astar_dim && gy < k->astar_dim, GA_ERR_ARG, "goal outside the astar grid");
    total_tiles += (size_t)k->n_tiles;
    max_dim = k->astar_dim > max_dim ? k->astar_dim : max_dim;
    lean = lean && prox_free(k);
  }
  GA_CUDA(c, cudaSetDevice(c->device));
  // batch scratch on the leading context: ring for all tiles, descriptors, closed counters
  const unsigned cap = ring_capacity(total_tiles);
  if ((int)cap > c->bring_cap)
  {
    if (c->d_bring) GA_CUDA(c, cudaFree(c->d_bring));
    c->d_bring = nullptr;
    GA_CUDA(c, cudaMalloc(&c->d_bring, sizeof(int) * (size_t)cap));
    GA_CUDA(c, cudaMemsetAsync(c->d_bring, 0, sizeof(int) * (size_t)cap, c->stream));
    c->bring_cap = (int)cap;
  }
  const size_t desc_bytes = (size_t)n * (sizeof(heur_grid) + sizeof(heur_args) + 2 * sizeof(int));
  if (desc_bytes > c->bdesc_cap)
  {
    if (c->d_bdesc) GA_CUDA(c, cudaFree(c->d_bdesc));
    c->d_bdesc = nullptr;
    GA_CUDA(c, cudaMalloc(&c->d_bdesc, desc_bytes));
    c->bdesc_cap = desc_bytes;
  }
  heur_grid* dG = reinterpret_cast<heur_grid*>(c->d_bdesc);
  heur_args* dA = reinterpret_cast<heur_args*>(dG + n);
  int* dClosed = reinterpret_cast<int*>(dA + n);
  std::vector<heur_grid> hG((size_t)n);
  std::vector<heur_args> hA((size_t)n);
  for (int i = 0; i < n; ++i) make_descs(ctxs[i], slot, goal_xy[2 * i], goal_xy[2 * i + 1], dClosed + 2 * i, &hG[i], &hA[i]);
  // order the leading stream after whatever the other contexts still have in flight
  for (int i = 1; i < n; ++i)
  {
    ga_ctx* k = ctxs[i];
    if (!k->ev_batch) GA_CUDA(c, cudaEventCreateWithFlags(&k->ev_batch, cudaEventDisableTiming));
    GA_CUDA(c, cudaEventRecord(k->ev_batch, k->stream));
    GA_CUDA(c, cudaStreamWaitEvent(c->stream, k->ev_batch, 0));
  }
  GA_CUDA(c, cudaMemcpyAsync(dG, hG.data(), sizeof(heur_grid) * n, cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(dA, hA.data(), sizeof(heur_args) * n, cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemsetAsync(dClosed, 0, sizeof(int) * 2 * n, c->stream));
  heur_queue q;
  q.ring = c->d_bring;
  q.mask = cap - 1;
  q.ctl = c->d_hctl;
  q.poll_ns = (unsigned)c->sw_poll_ns;
  queue_knobs(c, n, &q);
  const int rc = launch_build(c, n, hG[0], hA[0], dG, dA, q, max_dim, total_tiles, lean);
  if (rc != GA_OK) return rc;
  // the other contexts' streams continue after the batch
  if (!c->ev_batch) GA_CUDA(c, cudaEventCreateWithFlags(&c->ev_batch, cudaEventDisableTiming));
  GA_CUDA(c, cudaEventRecord(c->ev_batch, c->stream));
  for (int i = 1; i < n; ++i) GA_CUDA(c, cudaStreamWaitEvent(ctxs[i]->stream, c->ev_batch, 0));
  if (n_closed)
  {
    GA_CUDA(c, cudaMemcpy2DAsync(n_closed, sizeof(int), dClosed, 2 * sizeof(int), sizeof(int), (size_t)n,
                                 cudaMemcpyDeviceToHost, c->stream));
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

#ifdef SW_TRACE
extern "C" __attribute__((visibility("default"))) int ga_heur_trace(ga_ctx* c, unsigned long long* out, unsigned cap, unsigned* n, int reset)
{
  cudaStreamSynchronize(c->stream);
  unsigned cnt = 0;
  cudaMemcpyFromSymbol(&cnt, g_trace_n, sizeof cnt);
  if (cnt > (1u << 20)) cnt = 1u << 20;
  if (cnt > cap) cnt = cap;
  if (out && cnt) cudaMemcpyFromSymbol(out, g_trace, sizeof(unsigned long long) * cnt);
  *n = cnt;
  if (reset)
  {
    const unsigned z = 0;
    cudaMemcpyToSymbol(g_trace_n, &z, sizeof z);
  }
  return 0;
}
#endif
