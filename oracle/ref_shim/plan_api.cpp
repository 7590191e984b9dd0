// TEST / BENCH INFRASTRUCTURE ONLY -- C entry points (gap_*) that drive the reference's OWN serial
// Hybrid A* search (LIB/src/hybridastar_planning_lib/hybrid_a_star_lib.cpp, compiled unmodified where
// it lies under /root/reference) the way planning.py does for one planning query.
//
// The same file is compiled into two libraries (oracle/Makefile):
//   arm A  oracle/_ref/libga_ref.so    -- hybrid_a_star_lib.cpp on top of the reference's own AStar /
//                                        CollisionChecker / Vehicle: the pure reference planner;
//   arm B  oracle/_ref/libga_planb.so  -- the SAME hybrid_a_star_lib.cpp on top of the drop-in static
//                                        classes of guided-extended-hybrid-astar_b200/host/dropin/
//                                        (bodies = calls into libgridaccel.so).
// Both are linked with --wrap on CollisionChecker::checkPathCollision and Vehicle::move_car_some_steps,
// so that the sequence of expanded nodes and of every collision query with its answer can be compared
// between the arms without touching the reference's source (hAstarCore keeps its closed set in an
// unordered_map, the pop order is not stored anywhere).
//
// Call order of one query = planning.py:866-965 (do_planning) reduced to its first cycle:
//   create_patch (:315-345) -> process_meas_grid (:577-588) -> processSafetyPatch (:603)
//   -> recalculateEnv (:775) -> getAstarPath (:777) -> hybridAStarPlanning (:843), stopped before the
//   FITPACK spline interpolation (post-search; no gfortran here): hAstarCore + getFinalPath.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <array>
#include <chrono>
#include <execution>
#include <fstream>
#include <iostream>
#include <queue>
#include <sstream>
#include <unordered_map>

#include "yaml-cpp/yaml.h"
#define private public  // hAstarCore, getFinalPath, closed_set_, timeout_ms_ are private statics
#include "hybridastar_planning_lib/hybrid_a_star_lib.hpp"
#undef private

#include <cmath>
#include <cstdio>
#include <cstring>

// FITPACK (Fortran) cannot be built here; BSpline1D is only used after the search
// (interpolatePath, lane-graph smoothing) -- a link stub that must never run.
namespace fitpack_wrapper
{
BSpline1D::BSpline1D(std::vector<double>&, std::vector<double>&, int, double) : k(0), n(0), nc(0)
{
  throw std::runtime_error("BSpline1D stub: FITPACK is not available in this build");
}
double BSpline1D::operator()(double) { return 0.0; }
}  // namespace fitpack_wrapper

// ---- trace ------------------------------------------------------------------------------------
namespace
{
struct Trace
{
  std::vector<double> expanded;  // parent pose (x, y, yaw) of every node setNeighbors was called on, in order
  uint64_t n_checks = 0;
  uint64_t n_check_poses = 0;
  uint64_t hash = 1469598103934665603ull;  // FNV-1a over (n, last pose bits, answer) of every path check
  double last_parent[3] = { NAN, NAN, NAN };
  bool on = false;
  void reset()
  {
    expanded.clear();
    n_checks = n_check_poses = 0;
    hash = 1469598103934665603ull;
    last_parent[0] = last_parent[1] = last_parent[2] = NAN;
  }
  void mix(const void* p, size_t n)
  {
    const unsigned char* b = static_cast<const unsigned char*>(p);
    for (size_t i = 0; i < n; ++i)
    {
      hash ^= b[i];
      hash *= 1099511628211ull;
    }
  }
} g_trace;
char g_err[512];
std::optional<NodeHybrid> g_final;
Path g_path;
double g_ms_heur = 0, g_ms_search = 0;
}  // namespace

extern "C" {
// the real functions (reference's in arm A, the drop-in's in arm B)
bool __real__ZN16CollisionChecker18checkPathCollisionERKSt6vectorIdSaIdEES4_S4_(const std::vector<double>&,
                                                                                 const std::vector<double>&,
                                                                                 const std::vector<double>&);
MotionPrimitive __real__ZN7Vehicle19move_car_some_stepsERK4PoseIdEddid(const Pose<double>&, double, double, int, double);

bool __wrap__ZN16CollisionChecker18checkPathCollisionERKSt6vectorIdSaIdEES4_S4_(const std::vector<double>& x,
                                                                                 const std::vector<double>& y,
                                                                                 const std::vector<double>& yaw)
{
  const bool ok = __real__ZN16CollisionChecker18checkPathCollisionERKSt6vectorIdSaIdEES4_S4_(x, y, yaw);
  if (g_trace.on)
  {
    const uint64_t n = x.size();
    ++g_trace.n_checks;
    g_trace.n_check_poses += n;
    g_trace.mix(&n, sizeof n);
    if (n)
    {
      g_trace.mix(&x.back(), 8);
      g_trace.mix(&y.back(), 8);
      g_trace.mix(&yaw.back(), 8);
    }
    const unsigned char a = ok ? 1 : 0;
    g_trace.mix(&a, 1);
  }
  return ok;
}

MotionPrimitive __wrap__ZN7Vehicle19move_car_some_stepsERK4PoseIdEddid(const Pose<double>& pose, double arc_l,
                                                                      double motion_res, int direction, double steer)
{
  if (g_trace.on)
  {
    const double p[3] = { pose.x, pose.y, pose.yaw };
    if (std::memcmp(p, g_trace.last_parent, sizeof p) != 0)
    {
      std::memcpy(g_trace.last_parent, p, sizeof p);
      g_trace.expanded.insert(g_trace.expanded.end(), p, p + 3);
    }
  }
  return __real__ZN7Vehicle19move_car_some_stepsERK4PoseIdEddid(pose, arc_l, motion_res, direction, steer);
}
}

#define GAP_TRY(...)                                    \
  try                                                   \
  {                                                     \
    __VA_ARGS__;                                             \
    return 0;                                           \
  }                                                     \
  catch (const std::exception& e)                       \
  {                                                     \
    std::snprintf(g_err, sizeof g_err, "%s", e.what()); \
    return -1;                                          \
  }

static double now_ms()
{
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

extern "C" {
const char* gap_last_error() { return g_err; }

// PathPlanning.__init__ (planning.py:52-125) + Sim.__init__'s parameter push (simulation.py:73-84):
// vehicle, HybridAStar::initialize(1000, (0,0), share_dir), weights of params.yml.
// params12 = { alpha_, do_max_, do_min_, astar_prox_cost_, astar_movement_cost_, astar_lane_movement_cost_,
//              steer_change_cost_, steer_cost_, back_cost_, h_prox_cost_, switch_cost_, h_dist_cost_ }
int gap_init(const char* share_dir, double max_steer, double wb, double lf, double lb, double width, int is_ushift,
             const double* params12, int timeout_ms, int is_sim)
{
  GAP_TRY({
    Vehicle::initialize(max_steer, wb, lf, lb, width, is_ushift != 0);
    HybridAStar::setSim(is_sim != 0);
    // weights first: AStar::init_structs fills movement_cost_map_ with astar_movement_cost_
    AStar::alpha_ = params12[0];
    AStar::do_max_ = params12[1];
    AStar::do_min_ = params12[2];
    AStar::astar_prox_cost_ = params12[3];
    AStar::astar_movement_cost_ = params12[4];
    AStar::astar_lane_movement_cost_ = params12[5];
    HybridAStar::steer_change_cost_ = params12[6];
    HybridAStar::steer_cost_ = params12[7];
    HybridAStar::back_cost_ = params12[8];
    HybridAStar::h_prox_cost_ = params12[9];
    HybridAStar::switch_cost_ = params12[10];
    HybridAStar::h_dist_cost_ = params12[11];
    HybridAStar::initialize(1000, Point<double>(0, 0), share_dir);
    if (timeout_ms > 0) HybridAStar::timeout_ms_ = timeout_ms;
  })
}

// create_patch (planning.py:315-345): square patch around ego and goal, padded; HybridAStar::reinit.
// out3 = { origin_utm.x, origin_utm.y, patch_dim_gm }
int gap_create_patch(double ego_x, double ego_y, double goal_x, double goal_y, double padding, double* out3)
{
  GAP_TRY({
    const double min_x = std::min(ego_x, goal_x), min_y = std::min(ego_y, goal_y);
    const double max_x = std::max(ego_x, goal_x), max_y = std::max(ego_y, goal_y);
    const double dx = std::abs(min_x - max_x), dy = std::abs(min_y - max_y);
    const double patch_dim_utm = std::max(dx, dy) + 2 * padding;
    const int patch_dim_gm = grid_tf::utm2grid_round(patch_dim_utm);
    const Point<double> origin(dx / 2 + min_x - patch_dim_utm / 2, dy / 2 + min_y - patch_dim_utm / 2);
    HybridAStar::reinit(origin, patch_dim_gm);
    out3[0] = origin.x;
    out3[1] = origin.y;
    out3[2] = patch_dim_gm;
  })
}

// a patch that IS the map (synthetic configs): origin (0, 0), dim = map side
int gap_reinit(double origin_x, double origin_y, int patch_dim)
{
  GAP_TRY({ HybridAStar::reinit(Point<double>(origin_x, origin_y), patch_dim); })
}

// process_meas_grid (planning.py:577-588), simulation branch: window around the ego cell -> Cartographing
// -> CollisionChecker::patch_arr_.  ego in patch metres.
int gap_insert_window_sim(const uint8_t* window, int dim, double ego_x_patch, double ego_y_patch)
{
  GAP_TRY({
    const int ex = grid_tf::utm2grid_round(ego_x_patch), ey = grid_tf::utm2grid_round(ego_y_patch);
    const int off = dim / 2;
    const Point<int> origin(std::max(ex - off, 0), std::max(ey - off, 0));
    Vec2DFlat<uint8_t> local;
    local.resize_and_reset(dim, dim, 0);
    std::memcpy(local.getPtr(), window, (size_t)dim * dim);
    Cartographing::cartograph(local, origin, dim);
    Cartographing::passLocalMap(origin, dim);
  })
}

int gap_pass_local_map(const uint8_t* window, int ox, int oy, int dim)
{
  GAP_TRY({ CollisionChecker::passLocalMapData(window, Point<int>(ox, oy), dim); })
}

int gap_process_safety()
{
  GAP_TRY({ CollisionChecker::processSafetyPatch(); })
}

int gap_check_pose(double x, double y, double yaw)
{
  try
  {
    return CollisionChecker::checkPose(Pose<double>(x, y, yaw)) ? 1 : 0;
  }
  catch (const std::exception& e)
  {
    std::snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// env_handling (planning.py:756-783): recalculateEnv(goal, ego) + getAstarPath; poses in patch metres.
// Returns the number of cells of the guidance path (or -1 on error).
int gap_recalc_env(const double* ego3, const double* goal3, double* ms_out)
{
  try
  {
    const NodeHybrid ego = HybridAStar::createNode(Pose<double>(ego3[0], ego3[1], ego3[2]), 0.0);
    const NodeHybrid goal = HybridAStar::createNode(Pose<double>(goal3[0], goal3[1], goal3[2]), 0.0);
    const double t0 = now_ms();
    HybridAStar::recalculateEnv(goal, ego);
    if (ms_out) *ms_out = now_ms() - t0;
    const auto path = AStar::getAstarPath(ego.x_index, ego.y_index);
    return (int)path.first.size();
  }
  catch (const std::exception& e)
  {
    std::snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// plan_path (planning.py:823-864) up to the spline interpolation: hAstarCore + getFinalPath
// (hybrid_a_star_lib.cpp:1356-1547).  start = ego on the first cycle (get_shifted_planning_nodes).
// out8 = { found, n_closed, n_open, n_path, path.cost, idx_analytic, final.is_analytic, ms }
int gap_plan(const double* ego3, const double* start3, const double* goal3, int do_analytic, int trace, double* out8)
{
  GAP_TRY({
    const NodeHybrid ego = HybridAStar::createNode(Pose<double>(ego3[0], ego3[1], ego3[2]), 0.0);
    const NodeHybrid start = HybridAStar::createNode(Pose<double>(start3[0], start3[1], start3[2]), 0.0);
    const NodeHybrid goal = HybridAStar::createNode(Pose<double>(goal3[0], goal3[1], goal3[2]), 0.0);
    g_trace.reset();
    g_trace.on = trace != 0;
    g_path = Path();
    const double t0 = now_ms();
    g_final = HybridAStar::hAstarCore(ego, start, goal, do_analytic != 0);
    const double t1 = now_ms();
    g_trace.on = false;
    if (g_final) g_path = HybridAStar::getFinalPath(*g_final, HybridAStar::closed_set_);
    out8[0] = g_final ? 1 : 0;
    out8[1] = (double)HybridAStar::closed_set_.size();
    out8[2] = (double)HybridAStar::open_set_.size();
    out8[3] = (double)g_path.x_list.size();
    out8[4] = g_path.cost;
    out8[5] = g_path.idx_analytic;
    out8[6] = g_final ? (g_final->is_analytic ? 1 : 0) : 0;
    out8[7] = t1 - t0;
  })
}

// final path of the last gap_plan: x, y, yaw, delta (n doubles each), dir, type (n ints each)
int gap_get_path(double* x, double* y, double* yaw, double* delta, int* dir, int* type)
{
  const size_t n = g_path.x_list.size();
  for (size_t i = 0; i < n; ++i)
  {
    x[i] = g_path.x_list[i];
    y[i] = g_path.y_list[i];
    yaw[i] = g_path.yaw_list[i];
    delta[i] = g_path.delta_list[i];
    dir[i] = g_path.direction_list[i];
    type[i] = (int)g_path.types[i];
  }
  return (int)n;
}

// trace of the last gap_plan: out3 = { n_expanded, n_path_checks, n_checked_poses }, hash of all checks
int gap_get_trace(uint64_t* out3, uint64_t* hash)
{
  out3[0] = g_trace.expanded.size() / 3;
  out3[1] = g_trace.n_checks;
  out3[2] = g_trace.n_check_poses;
  *hash = g_trace.hash;
  return 0;
}
int gap_get_expanded(double* xyyaw)
{
  std::memcpy(xyyaw, g_trace.expanded.data(), g_trace.expanded.size() * sizeof(double));
  return (int)(g_trace.expanded.size() / 3);
}

// the closed set of the search as sorted rows: key, x_index, y_index, yaw_index, parent, cost, dist
int gap_get_closed(int64_t* key5, double* cost2, int cap)
{
  std::vector<size_t> keys;
  keys.reserve(HybridAStar::closed_set_.size());
  for (const auto& kv : HybridAStar::closed_set_) keys.push_back(kv.first);
  std::sort(keys.begin(), keys.end());
  int n = 0;
  for (const size_t k : keys)
  {
    if (n >= cap) break;
    const NodeHybrid& nd = HybridAStar::closed_set_.at(k);
    key5[5 * n + 0] = (int64_t)k;
    key5[5 * n + 1] = nd.x_index;
    key5[5 * n + 2] = nd.y_index;
    key5[5 * n + 3] = nd.yaw_index;
    key5[5 * n + 4] = nd.parent_index;
    cost2[2 * n + 0] = nd.cost;
    cost2[2 * n + 1] = nd.dist;
    ++n;
  }
  return n;
}

// heuristic planes of closed_set_path_ (which = 1) / closed_set_guidance_ (0) as dense arrays
int gap_get_heuristic(int which, double* cost, double* dist, int32_t* parent, uint8_t* flags)
{
  const auto& cs = which ? AStar::closed_set_path_ : AStar::closed_set_guidance_;
  const size_t n = (size_t)AStar::astar_dim_ * AStar::astar_dim_;
  for (size_t i = 0; i < n; ++i)
  {
    cost[i] = INFINITY;
    dist[i] = INFINITY;
    parent[i] = -2;
    flags[i] = 0;
  }
  size_t cnt = 0;
  for (const auto& kv : cs)
  {
    const size_t i = kv.first;
    const NodeDisc& nd = kv.second;
    cost[i] = nd.cost_;
    dist[i] = nd.cost_dist_;
    parent[i] = nd.parent_index_;
    flags[i] = (uint8_t)(1 | (nd.is_unknown_ ? 2 : 0));
    ++cnt;
  }
  return (int)cnt;
}

int gap_astar_dim() { return AStar::astar_dim_; }

#ifdef GA_DROPIN_HPP
// arm B switches: heuristic mode (0 = the reference's closing order, 1 = GPU fixed point), DT arithmetic, counters
int gap_b200_set_mode(int fixedpoint)
{
  ga_dropin::setHeuristicMode(fixedpoint ? GA_HEUR_FIXEDPOINT : GA_HEUR_REFORDER);
  return 0;
}
int gap_set_dt_variant(int v)
{
  GAP_TRY({ ga_dropin::setDistanceTransformVariant(v); })
}
int gap_b200_stats(unsigned long long* out8, unsigned long long* ns3)
{
  const ga_dropin::Stats& s = ga_dropin::stats();
  out8[0] = s.expansions_prefetched;
  out8[1] = s.path_checks_cached;
  out8[2] = s.path_checks_direct;
  out8[3] = s.poses_checked;
  out8[4] = s.lookahead_launches;
  out8[5] = s.lookahead_hits;
  out8[6] = s.lookahead_rechecks;
  out8[7] = (unsigned long long)ga_launch_count(ga_dropin::ctx());
  ns3[0] = s.ns_submit;
  ns3[1] = s.ns_collect;
  ns3[2] = s.ns_harvest;
  return 0;
}
int gap_b200_set_lookahead(int on)
{
  GAP_TRY({ ga_dropin::setLookahead(on != 0); })
}
long long gap_b200_launches() { return ga_launch_count(ga_dropin::ctx()); }
// the DEVICE copy of a map (1 h_prox_arr_, 2 motion_res_map_, 3 movement_cost_map_): what the kernels read
int gap_b200_device_map(int which, double* dst)
{
  GAP_TRY({
    const size_t n = (size_t)AStar::astar_dim_ * AStar::astar_dim_;
    const int id = which == 1 ? GA_GRID_H_PROX : which == 2 ? GA_GRID_MOTION_RES : GA_GRID_MOVEMENT_COST;
    ga_dropin::check(ga_download(ga_dropin::ctx(), id, dst, n * sizeof(double)));
  })
}
// the DEVICE's four result planes of a heuristic slot (0 guidance, 1 path): what the closed set's accessors must reproduce
int gap_b200_heur_planes(int which, double* cost, double* dist, int32_t* parent, uint8_t* flags)
{
  GAP_TRY({
    const size_t n = (size_t)AStar::astar_dim_ * AStar::astar_dim_;
    ga_ctx* c = ga_dropin::ctx();
    ga_dropin::check(ga_download(c, GA_GRID_HEUR_COST + which, cost, n * 8));
    ga_dropin::check(ga_download(c, GA_GRID_HEUR_DIST + which, dist, n * 8));
    ga_dropin::check(ga_download(c, GA_GRID_HEUR_PARENT + which, parent, n * 4));
    ga_dropin::check(ga_download(c, GA_GRID_HEUR_FLAGS + which, flags, n));
  })
}
#else
int gap_set_dt_variant(int v)
{
  cv::shim_dt_variant = v;
  return 0;
}
#endif

// dense copies of the planner-side maps: 0 astar_grid_ (u8), 1 h_prox_arr_ (f64), 2 motion_res_map_ (f64),
// 3 movement_cost_map_ (f64), 4 patch_safety_arr_ (u8, patch_dim^2)
int gap_get_map(int which, void* dst)
{
  GAP_TRY({
    const size_t n = (size_t)AStar::astar_dim_ * AStar::astar_dim_;
    switch (which)
    {
      case 0: std::memcpy(dst, AStar::astar_grid_.getPtr(), n); break;
      case 1: std::memcpy(dst, AStar::h_prox_arr_.getPtr(), n * 8); break;
      case 2: std::memcpy(dst, AStar::motion_res_map_.getPtr(), n * 8); break;
      case 3: std::memcpy(dst, AStar::movement_cost_map_.getPtr(), n * 8); break;
      default: throw std::runtime_error("gap_get_map: unknown map");
    }
  })
}
}
