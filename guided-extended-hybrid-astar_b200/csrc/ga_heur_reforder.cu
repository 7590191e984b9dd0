// ga_heur_reforder.cu -- GA_HEUR_REFORDER: AStar::calcDistanceHeuristic with the reference's own closing
// order (LIB/src/hybridastar_planning_lib/a_star.cpp:110-295), for callers that need the serial
// Hybrid A* search to expand the same nodes and return the same path as on the reference.
//
// Why this part runs on the host.  The reference's result is not a shortest-path map: which cells are
// closed (early exit after the start cell, a_star.cpp:198-219), the few cells whose cost stays above the
// optimum (no re-opening, Euclidean term taken at the parent, :238-241, 260-263) and the parent / cost_dist
// chain all depend on the exact order in which (priority, id) pairs leave ONE binary heap
// (data_structures2.hpp:233-255).  That order is a serial dependency chain of heap operations; replaying it
// is host work like the search itself.  The order-free fixed point of the same relaxation is what the GPU
// computes (ga_heuristic.cu, GA_HEUR_FIXEDPOINT) -- equal to this result wherever the reference is optimal.
// Here the device supplies the inputs (the pooled grid and the potential field are device-resident results of
// K1 / K4) and receives the four result planes back, so every later consumer (ga_download, ga_valid_close_pose,
// the successor epilogue of ga_expand) reads one and the same slot in either mode.
//
// The replay keeps the reference's arithmetic operand for operand (fp64, no contraction) and its heap
// discipline: lazy deletion, one heap entry per strict improvement, an open node leaves with whichever of its
// entries comes first and carries its CURRENT best record.  Dense per-cell records replace the two
// unordered_maps; the pop sequence only depends on the set of (priority, id) pairs in the heap, not on the
// container, so it is the reference's.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "ga_internal.cuh"

namespace
{
struct heap_entry
{
  double prio;
  uint32_t id;
};
// std::greater<std::pair<double, size_t>>: smallest priority first, smallest id among equal priorities
inline bool before(const heap_entry& a, const heap_entry& b) { return a.prio < b.prio || (a.prio == b.prio && a.id < b.id); }

class pair_heap
{
public:
  void clear() { v_.clear(); }
  bool empty() const { return v_.empty(); }
  void push(double prio, uint32_t id)
  {
    size_t i = v_.size();
    v_.push_back({ prio, id });
    const heap_entry e = v_[i];
    while (i > 0)
    {
      const size_t p = (i - 1) >> 2;
      if (!before(e, v_[p])) break;
      v_[i] = v_[p];
      i = p;
    }
    v_[i] = e;
  }
  heap_entry pop()
  {
    const heap_entry top = v_[0];
    const heap_entry last = v_.back();
    v_.pop_back();
    const size_t n = v_.size();
    if (n)
    {
      size_t i = 0;
      for (;;)
      {
        const size_t c0 = 4 * i + 1;
        if (c0 >= n) break;
        size_t m = c0;
        const size_t ce = std::min(c0 + 4, n);
        for (size_t c = c0 + 1; c < ce; ++c)
          if (before(v_[c], v_[m])) m = c;
        if (!before(v_[m], last)) break;
        v_[i] = v_[m];
        i = m;
      }
      v_[i] = last;
    }
    return top;
  }

private:
  std::vector<heap_entry> v_;  // 4-ary: shallower than binary, the order of pops is the same total order
};

struct host_planes  // per-context host mirrors for the replay (pinned once, reused)
{
  uint8_t* grid = nullptr;
  uint8_t* fence = nullptr;
  double* mc = nullptr;
  double* prox = nullptr;
  double* cost = nullptr;
  double* dist = nullptr;
  int32_t* parent = nullptr;
  uint8_t* flags = nullptr;
  uint8_t* state = nullptr;  // 0 unseen, 1 open, 2 closed
  size_t cells = 0;
};
}  // namespace

struct ga_reforder_state
{
  host_planes h;
  pair_heap heap;
  std::vector<int32_t> near_ids;
  std::vector<double> near_dist;
  std::vector<int32_t> order;  // closing order of the last build (diagnostics / tests)
};

void ga_reforder_free(ga_ctx* c)
{
  ga_reforder_state* s = c->reforder;
  if (!s) return;
  cudaFreeHost(s->h.grid);
  cudaFreeHost(s->h.mc);
  cudaFreeHost(s->h.cost);
  delete s;
  c->reforder = nullptr;
}

static int reforder_reserve(ga_ctx* c, size_t cells)
{
  if (!c->reforder) c->reforder = new ga_reforder_state();
  host_planes& h = c->reforder->h;
  if (h.cells == cells) return GA_OK;
  cudaFreeHost(h.grid);
  cudaFreeHost(h.mc);
  cudaFreeHost(h.cost);
  h = host_planes();
  // three pinned arenas: bytes in, doubles in, results out (+ the state bytes)
  GA_CUDA(c, cudaMallocHost(&h.grid, 3 * cells));
  h.fence = h.grid + cells;
  h.state = h.fence + cells;
  GA_CUDA(c, cudaMallocHost(&h.mc, 2 * cells * sizeof(double)));
  h.prox = h.mc + cells;
  GA_CUDA(c, cudaMallocHost(&h.cost, cells * (2 * sizeof(double) + sizeof(int32_t) + 1)));
  h.dist = h.cost + cells;
  h.parent = reinterpret_cast<int32_t*>(h.dist + cells);
  h.flags = reinterpret_cast<uint8_t*>(h.parent + cells);
  h.cells = cells;
  return GA_OK;
}

// a_star.cpp:110-295 on the host mirrors.  Returns the number of closed cells.
static int replay(ga_reforder_state* s, const ga_astar_params& ap, int dim, int goal_x, int goal_y, int start_x, int start_y,
                  bool early_exit, bool only_near, unsigned max_extra_nodes)
{
  host_planes& h = s->h;
  const size_t cells = h.cells;
  const double inf = INFINITY;
  std::fill(h.cost, h.cost + cells, inf);
  std::fill(h.dist, h.dist + cells, inf);
  std::fill(h.parent, h.parent + cells, -2);
  std::memset(h.flags, 0, cells);
  std::memset(h.state, 0, cells);
  s->heap.clear();
  s->near_ids.clear();
  s->near_dist.clear();
  s->order.clear();

  // motion_ and movement_distances_ (a_star.hpp:50-57)
  static const int DX[8] = { 1, 0, -1, 0, -1, -1, 1, 1 };
  static const int DY[8] = { 0, 1, 0, -1, -1, 1, -1, 1 };
  const double L1 = 0.625, L2 = 0.625 * std::sqrt(2.0);
  const double LEN[8] = { L1, L1, L1, L1, L2, L2, L2, L2 };

  const uint32_t goal_id = (uint32_t)goal_y * dim + goal_x;
  const uint32_t start_id = (uint32_t)start_y * dim + start_x;
  h.state[goal_id] = 1;
  h.cost[goal_id] = 0.0;
  h.dist[goal_id] = 0.0;
  h.parent[goal_id] = -1;
  h.flags[goal_id] = h.grid[goal_id] == GA_UNKNOWN ? GA_H_UNKNOWN : 0;
  s->heap.push(0.0, goal_id);

  bool start_found = start_id == goal_id, is_near = false;
  unsigned nr_extra = 0;
  size_t n_closed = 0;
  const size_t limit = (size_t)std::round(0.95 * ((double)dim * dim));  // explosion guard, :159
  bool hit_limit = false;

  for (;;)
  {
    if (n_closed > limit)
    {
      hit_limit = true;
      break;
    }
    if (s->heap.empty()) break;
    const uint32_t c_id = s->heap.pop().id;
    if (h.state[c_id] != 1) continue;  // stale entry of a node that already left (:290-293)
    h.state[c_id] = 2;
    h.flags[c_id] |= GA_H_CLOSED;
    s->order.push_back((int32_t)c_id);
    ++n_closed;
    const int cx = (int)(c_id % (uint32_t)dim), cy = (int)(c_id / (uint32_t)dim);
    const double xd = cx - start_x, yd = cy - start_y;
    const double d_start = std::abs(std::sqrt(xd * xd + yd * yd));  // Point::dist2

    if (only_near && d_start < 10)  // :182-193
    {
      s->near_ids.push_back((int32_t)c_id);
      s->near_dist.push_back(d_start);
      is_near = true;
    }
    if (early_exit)  // :198-207, on flat indices
    {
      const int diff = std::abs((int)c_id - (int)start_id);
      if (diff == 0)
        start_found = true;
      else if (diff == 1 || diff == dim)
        is_near = true;
    }
    if ((early_exit || only_near) && (start_found || is_near))  // :210-219
    {
      if (nr_extra > max_extra_nodes) continue;  // closed, not expanded
      ++nr_extra;
    }

    const double c_cost = h.cost[c_id], c_dist = h.dist[c_id];
    const double prox_cost = h.prox[c_id] * ap.astar_prox_cost;
    const bool c_unknown = h.grid[c_id] == GA_UNKNOWN;
    const double unknown_cost = c_unknown ? ap.unknown_cost_w : 0;
    const double h_euclid = ap.astar_lane_movement_cost * d_start * 0.625;

    for (int i = 0; i < 8; ++i)
    {
      const int nx = cx + DX[i], ny = cy + DY[i];
      if (nx < 0 || ny < 0 || nx >= dim || ny >= dim) continue;  // verifyNode :325-343
      const uint32_t n_id = (uint32_t)ny * dim + nx;
      if (h.grid[n_id] == GA_OCC || h.fence[n_id] == 1) continue;
      const uint8_t st = h.state[n_id];
      if (st == 2) continue;
      const double movement_costs = LEN[i] * h.mc[n_id];
      const double n_cost = c_cost + movement_costs + prox_cost + unknown_cost;
      if (st == 1 && !(h.cost[n_id] > n_cost)) continue;  // strict improvement only (:280)
      h.state[n_id] = 1;
      h.cost[n_id] = n_cost;
      h.dist[n_id] = c_dist + LEN[i];
      h.parent[n_id] = (int32_t)c_id;
      h.flags[n_id] = c_unknown ? GA_H_UNKNOWN : 0;  // NodeDisc::is_unknown_ = the PARENT cell's state (:253, 268)
      s->heap.push(h_euclid + n_cost, n_id);
    }
  }
  if (hit_limit)  // nodes still open are not part of the closed set
    for (size_t i = 0; i < cells; ++i)
      if (h.state[i] == 1)
      {
        h.cost[i] = inf;
        h.dist[i] = inf;
        h.parent[i] = -2;
        h.flags[i] = 0;
      }
  return (int)n_closed;
}

int ga_heur_reforder_build(ga_ctx* c, int slot, const ga_heur_query* q, int* n_closed)
{
  const int dim = c->astar_dim;
  const size_t cells = (size_t)dim * dim;
  const int rc = reforder_reserve(c, cells);
  if (rc != GA_OK) return rc;
  ga_reforder_state* s = c->reforder;
  host_planes& h = s->h;
  // inputs are device-resident results of K1 / K4 (+ uploads): bring them over
  GA_CUDA(c, cudaMemcpyAsync(h.grid, c->d_astar, cells, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(h.fence, c->d_fence, cells, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(h.mc, c->d_mc, cells * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(h.prox, c->d_prox, cells * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  const int n = replay(s, c->ap, dim, q->goal_x, q->goal_y, q->start_x, q->start_y, (q->flags & GA_HEUR_EARLY_EXIT) != 0,
                       (q->flags & GA_HEUR_ONLY_NEAR) != 0, q->max_extra_nodes);
  GA_CUDA(c, cudaMemcpyAsync(c->d_hcost[slot], h.cost, cells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(c->d_hdist[slot], h.dist, cells * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(c->d_hparent[slot], h.parent, cells * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemcpyAsync(c->d_hflags[slot], h.flags, cells, cudaMemcpyHostToDevice, c->stream));
  c->rec_valid[slot] = false;  // the planes are the replay's: no steps plane belongs to them
  c->near_slot = slot;
  c->heur_host_valid[slot] = true;
  if (n_closed) *n_closed = n;
  return GA_OK;
}

// host-resident copies of the last REFORDER build of `slot` (valid until the next build / patch reset)
// The replay on caller-owned HOST planes, no context and no device involved: for callers that hold host
// mirrors of the inputs anyway, and for the CPU-side tests that pin the replay on the reference's outputs.
extern "C" int ga_heur_reforder_host(const ga_astar_params* ap, int dim, const uint8_t* grid, const uint8_t* fence,
                                     const double* mc, const double* prox, const ga_heur_query* q, double* cost, double* dist,
                                     int32_t* parent, uint8_t* flags, int32_t* order, int* n_closed)
{
  if (!ap || !grid || !fence || !mc || !prox || !q || !cost || !dist || !parent || !flags || dim < 1) return GA_ERR_ARG;
  if (q->goal_x < 0 || q->goal_y < 0 || q->goal_x >= dim || q->goal_y >= dim) return GA_ERR_ARG;
  if (q->start_x < 0 || q->start_y < 0 || q->start_x >= dim || q->start_y >= dim) return GA_ERR_ARG;
  ga_reforder_state s;
  const size_t cells = (size_t)dim * dim;
  std::vector<uint8_t> state(cells);
  s.h.grid = const_cast<uint8_t*>(grid);
  s.h.fence = const_cast<uint8_t*>(fence);
  s.h.mc = const_cast<double*>(mc);
  s.h.prox = const_cast<double*>(prox);
  s.h.cost = cost;
  s.h.dist = dist;
  s.h.parent = parent;
  s.h.flags = flags;
  s.h.state = state.data();
  s.h.cells = cells;
  const int n = replay(&s, *ap, dim, q->goal_x, q->goal_y, q->start_x, q->start_y, (q->flags & GA_HEUR_EARLY_EXIT) != 0,
                       (q->flags & GA_HEUR_ONLY_NEAR) != 0, q->max_extra_nodes);
  if (order) std::copy(s.order.begin(), s.order.end(), order);
  if (n_closed) *n_closed = n;
  return GA_OK;
}

void ga_heur_reforder_near(const ga_ctx* c, const int32_t** ids, const double** dist, int* n)
{
  const ga_reforder_state* s = c->reforder;
  *n = s ? (int)s->near_ids.size() : 0;
  *ids = s ? s->near_ids.data() : nullptr;
  *dist = s ? s->near_dist.data() : nullptr;
}

void ga_heur_reforder_order(const ga_ctx* c, const int32_t** ids, int* n)
{
  const ga_reforder_state* s = c->reforder;
  *n = s ? (int)s->order.size() : 0;
  *ids = s ? s->order.data() : nullptr;
}
