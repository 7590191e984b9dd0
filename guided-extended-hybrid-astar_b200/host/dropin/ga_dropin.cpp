// ga_dropin.cpp -- bodies of the drop-in static classes (ga_dropin.hpp): every grid computation is a call into
// libgridaccel.so.  Host code here is what the reference also runs on the host around its grids: the footprint
// model (fp64 trig, once), the motion-primitive rollout (glibc, it defines the search state), the Fortune sweep
// (the reference's own VoronoiDiagramGenerator), lane-graph and geofence rasterisation.
#include "ga_dropin.hpp"

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>

namespace ga_dropin
{
namespace
{
ga_ctx* g_ctx = nullptr;
unsigned g_heur_mode = GA_HEUR_REFORDER;
Stats g_stats;
}  // namespace

ga_ctx* ctx()
{
  if (!g_ctx)
  {
    if (ga_abi_version() != GA_ABI_VERSION) throw std::runtime_error("gridaccel: ABI version mismatch");
    int dev = 0;
    if (const char* s = std::getenv("GA_DEVICE")) dev = std::atoi(s);
    if (ga_create(dev, &g_ctx) != GA_OK)
      throw std::runtime_error("gridaccel: ga_create failed -- no usable CUDA device (there is no CPU path)");
  }
  return g_ctx;
}
void check(int rc)
{
  if (rc != GA_OK) throw std::runtime_error(std::string("gridaccel: ") + ga_last_error(g_ctx));
}
void setHeuristicMode(unsigned flags) { g_heur_mode = flags & GA_HEUR_REFORDER; }
unsigned heuristicMode() { return g_heur_mode; }
Stats& stats() { return g_stats; }

namespace
{
constexpr int NB_STEER = 13;  // HybridAStar::NB_STEER (hybrid_a_star_lib.hpp:48)
double g_turn_on_point_angle = 90.0;  // config.yml TURN_ON_POINT_ANGLE (read in CollisionChecker::initialize)

// All primitives of one expanded node, rolled out on the host and checked on the device in one launch.  The poses live
// in flat arrays that are reused from expansion to expansion (no allocation per primitive); a MotionPrimitive is only
// built for the one the caller asks for.
struct Fan
{
  bool valid = false;
  double parent[3] = { 0, 0, 0 }, arc_l = 0, motion_res = 0;
  std::vector<double> x, y, yaw;  // poses of all primitives, primitive k = [off[k], off[k + 1])
  std::vector<int> off;
  std::vector<double> key;        // steer (move fan) or delta_angle (rear-axis fan)
  std::vector<int> dir;
  std::vector<int> first_hit;
  int last = -1;  // primitive handed out last (checkPathCollision is asked about it next)
  int size() const { return (int)key.size(); }
  int count(int k) const { return off[(size_t)k + 1] - off[(size_t)k]; }
  void clear()
  {
    x.clear();
    y.clear();
    yaw.clear();
    off.assign(1, 0);
    key.clear();
    dir.clear();
    last = -1;
  }
  void close_prim(double k, int d)
  {
    off.push_back((int)x.size());
    key.push_back(k);
    dir.push_back(d);
  }
};
Fan g_move_fan, g_ra_fan;
std::vector<double> g_x, g_y, g_yaw;
std::vector<int> g_seg;

void check_fan(Fan& f)
{
  f.first_hit.assign((size_t)f.size(), -1);
  if (!f.x.empty())
    check(ga_check_poses(ctx(), f.x.data(), f.y.data(), f.yaw.data(), (int)f.x.size(), f.off.data(), f.size(), 0,
                         f.first_hit.data()));
  g_stats.poses_checked += f.x.size();
  ++g_stats.expansions_prefetched;
}

// index of the fan primitive whose pose lists are exactly (x, y, yaw), or -1
int match_fan(const Fan& f, const std::vector<double>& x, const std::vector<double>& y, const std::vector<double>& yaw)
{
  if (!f.valid) return -1;
  auto is = [&](int k) {
    const size_t n = (size_t)f.count(k), o = (size_t)f.off[(size_t)k];
    return x.size() == n && y.size() == n && yaw.size() == n &&
           (n == 0 || (std::memcmp(f.x.data() + o, x.data(), n * 8) == 0 && std::memcmp(f.y.data() + o, y.data(), n * 8) == 0 &&
                       std::memcmp(f.yaw.data() + o, yaw.data(), n * 8) == 0));
  };
  if (f.last >= 0 && is(f.last)) return f.last;
  for (int k = 0; k < f.size(); ++k)
    if (is(k)) return k;
  return -1;
}

// ---- look-ahead: the fans of the nodes the search will expand next ----------------------------------------------
// Every node the search expands (except the first) is a successor it created earlier.  When a node's fan has been
// answered, its collision-free successors are handed to ga_expand_submit: the device rolls out and checks THEIR fans
// while the host goes on with the search; the answers are waiting in host memory when one of them is popped.  The
// device rollout uses its own libm (last-bit differences), so an answer is only trusted when no pose came within
// 1e-9 cells / yaw bins of a boundary (ga_succ::near_cell_edge); the others are re-checked from the host's poses.
struct PoseKey
{
  uint64_t x, y, yaw;
  bool operator==(const PoseKey& o) const { return x == o.x && y == o.y && yaw == o.yaw; }
};
struct PoseKeyHash
{
  size_t operator()(const PoseKey& k) const
  {
    uint64_t h = k.x * 0x9E3779B97F4A7C15ull;
    h ^= (k.y + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.yaw + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    return (size_t)h;
  }
};
// answers of one fan, filed under the pose bits of the node it belongs to: open addressing in one preallocated array
// (no allocation per entry), validity by epoch (forgetting everything = one increment)
struct FanSlot
{
  PoseKey key;
  double motion_res;
  uint32_t epoch;
  int8_t ans[2 * NB_STEER];  // 0 collides, 1 free, 2 ask again (near a cell / yaw-bin edge)
};
class FanTable
{
public:
  FanTable() : slots_(size_t(1) << 19), mask_((size_t(1) << 19) - 1) {}
  void reset()
  {
    ++epoch_;
    count_ = 0;
  }
  FanSlot* find(const PoseKey& k)
  {
    for (size_t i = PoseKeyHash()(k) & mask_;; i = (i + 1) & mask_)
    {
      FanSlot& s = slots_[i];
      if (s.epoch != epoch_) return nullptr;
      if (s.key == k) return &s;
    }
  }
  FanSlot* insert(const PoseKey& k)
  {
    if (count_ * 2 > slots_.size()) reset();  // cannot happen within one search of the reference's size; stay bounded
    for (size_t i = PoseKeyHash()(k) & mask_;; i = (i + 1) & mask_)
    {
      FanSlot& s = slots_[i];
      if (s.epoch != epoch_)
      {
        s.epoch = epoch_;
        s.key = k;
        ++count_;
        return &s;
      }
      if (s.key == k) return &s;
    }
  }

private:
  std::vector<FanSlot> slots_;
  size_t mask_, count_ = 0;
  uint32_t epoch_ = 1;
};
FanTable g_lookahead;
struct Pending
{
  int ticket = 0, n = 0;
  PoseKey key[2 * NB_STEER];
  double motion_res[2 * NB_STEER];
};
constexpr int MAX_PENDING = 3;  // ga_expand_submit keeps four submissions alive
Pending g_pend[MAX_PENDING];
int g_pend_head = 0, g_pend_count = 0;  // oldest submission, number in flight
bool g_lookahead_on = true;
bool g_expand_params_set = false;
double g_expand_arc_l = 0, g_yaw_res_deg = 10.0;
std::vector<uint8_t> g_flags;

inline PoseKey key_of(double x, double y, double yaw)
{
  PoseKey k;
  std::memcpy(&k.x, &x, 8);
  std::memcpy(&k.y, &y, 8);
  std::memcpy(&k.yaw, &yaw, 8);
  return k;
}

void lookahead_reset()
{
  while (g_pend_count > 0)
  {
    g_flags.resize((size_t)32 * (size_t)ga_expand_stride(ctx()));
    ga_expand_collect_flags(ctx(), g_pend[g_pend_head].ticket, g_flags.data(), nullptr);  // drain; results are dropped
    g_pend_head = (g_pend_head + 1) % MAX_PENDING;
    --g_pend_count;
  }
  g_lookahead.reset();
}

inline unsigned long long now_ns()
{
  return (unsigned long long)std::chrono::duration_cast<std::chrono::nanoseconds>(
             std::chrono::steady_clock::now().time_since_epoch()).count();
}

// files the answers of the OLDEST submission still in flight (waits for it if it has not finished)
void lookahead_harvest()
{
  if (g_pend_count == 0) return;
  Pending& g_pending = g_pend[g_pend_head];
  g_pend_head = (g_pend_head + 1) % MAX_PENDING;
  --g_pend_count;
  const unsigned long long t0 = now_ns();
  const int stride = ga_expand_stride(ctx());
  g_flags.resize((size_t)32 * (size_t)stride);
  int n = 0;
  check(ga_expand_collect_flags(ctx(), g_pending.ticket, g_flags.data(), &n));
  const unsigned long long t1 = now_ns();
  g_stats.ns_collect += t1 - t0;
  for (int j = 0; j < n && j < g_pending.n; ++j)
  {
    FanSlot* a = g_lookahead.insert(g_pending.key[j]);
    a->motion_res = g_pending.motion_res[j];
    const uint8_t* f = g_flags.data() + (size_t)j * stride;
    for (int k = 0; k < 2 * NB_STEER; ++k) a->ans[k] = (f[k] & GA_SUCC_AMBIGUOUS) ? 2 : ((f[k] & GA_SUCC_FREE) ? 1 : 0);
  }
  g_stats.ns_harvest += now_ns() - t1;
  g_stats.lookahead_launches++;
}

// successors of the fan just answered -> device
void lookahead_submit(const Fan& f, double arc_l)
{
  if (!g_lookahead_on) return;
  ga_ctx* c = ctx();
  if (!g_expand_params_set || g_expand_arc_l != arc_l)
  {
    const ga_expand_params ep{ Vehicle::max_steer_, Vehicle::w_b_, arc_l, g_yaw_res_deg, g_turn_on_point_angle, NB_STEER };
    check(ga_set_expand_params(c, &ep));
    g_expand_params_set = true;
    g_expand_arc_l = arc_l;
  }
  if (g_pend_count == MAX_PENDING) lookahead_harvest();
  Pending& g_pending = g_pend[(g_pend_head + g_pend_count) % MAX_PENDING];
  ga_node nodes[2 * NB_STEER];
  int n = 0;
  const int ad = AStar::astar_dim_;
  for (int k = 0; k < f.size() && n < 2 * NB_STEER; ++k)
  {
    if (f.first_hit[(size_t)k] >= 0 || f.count(k) == 0) continue;
    const size_t e = (size_t)f.off[(size_t)k + 1] - 1;
    const double x = f.x[e], y = f.y[e], yaw = f.yaw[e];
    const int xi = static_cast<int>(x * grid_tf::con2star_), yi = static_cast<int>(y * grid_tf::con2star_);  // mapCont2Disc
    if (!(-1 < xi && xi < ad && -1 < yi && yi < ad)) continue;
    const PoseKey key = key_of(x, y, yaw);
    if (g_lookahead.find(key) != nullptr) continue;
    const double mr = AStar::motion_res_map_(yi, xi);  // the resolution hAstarCore will roll this node out with (:1483)
    nodes[n] = ga_node{ x, y, yaw, mr, 0, 0 };
    g_pending.key[n] = key;
    g_pending.motion_res[n] = mr;
    ++n;
  }
  if (n == 0) return;
  const unsigned long long t0 = now_ns();
  check(ga_expand_submit(c, nodes, n, &g_pending.ticket));
  g_stats.ns_submit += now_ns() - t0;
  g_pending.n = n;
  ++g_pend_count;
}
}  // namespace
void setLookahead(bool on)
{
  if (!on) lookahead_reset();
  g_lookahead_on = on;
}
}  // namespace ga_dropin

using namespace ga_dropin;

// =====================================================================================================
// Vehicle (LIB/src/collision_checker_lib/vehicle.cpp)
// =====================================================================================================
void Vehicle::setPose(const Pose<double>& pose) { pose_utm_ = pose; }

void Vehicle::initialize(double max_steer, double w_b, double l_f, double l_b, double width, bool is_ushift)  // :11-32
{
  max_steer_ = max_steer;
  w_b_ = w_b;
  lf_ = l_f;
  lb_ = l_b;
  length_ = l_f + l_b;
  width_ = width;
  is_ushift_ = is_ushift;
  max_curvature_ = tan(max_steer_) / w_b_;
  f_r_corner_ = Point<double>(lf_, -width_ / 2);
  f_l_corner_ = Point<double>(lf_, width_ / 2);
  r_r_corner_ = Point<double>(-lb_, -width_ / 2);
  r_l_corner_ = Point<double>(-lb_, width_ / 2);
  geo_center_ = (lb_ + lf_) / 2 - lb_;
  vehicle_vertices_ = { f_l_corner_, f_r_corner_, r_r_corner_, r_l_corner_, f_l_corner_ };
  vis_vehicle_vertices_ = vehicle_vertices_;
  g_move_fan.valid = g_ra_fan.valid = false;
}

void Vehicle::move(Pose<double>& pose, double distance, double steer)  // :34-43, one Euler step of the bicycle model
{
  pose.x += distance * cos(pose.yaw);
  pose.y += distance * sin(pose.yaw);
  pose.yaw += distance * tan(steer) / w_b_;
  pose.yaw = util::constrainAngleMinPIPlusPi(pose.yaw);
}

namespace
{
// Vehicle::move_car_some_steps (:100-122) into the fan's flat arrays
void roll_out(Fan& f, const Pose<double>& pose, double arc_l, double motion_res, int direction, double steer)
{
  const auto n = static_cast<size_t>(round(arc_l / motion_res));
  Pose<double> p = pose;
  for (size_t i = 0; i < n; ++i)
  {
    Vehicle::move(p, motion_res * direction, steer);
    f.x.push_back(p.x);
    f.y.push_back(p.y);
    f.yaw.push_back(p.yaw);
  }
  f.close_prim(steer, direction);
}

// Vehicle::turn_on_rear_axis (:45-98) into the fan's flat arrays
void rotate_on_axis(Fan& f, const Pose<double>& pose, double delta_angle, double yaw_res_coll)
{
  double yaw = pose.yaw;
  const double goal_yaw = util::constrainAngleMinPIPlusPi(yaw + delta_angle);
  const double step = yaw_res_coll * util::TO_RAD;
  double margin;
  do
  {
    yaw = util::constrainAngleMinPIPlusPi(yaw + util::sgn(delta_angle) * step);
    f.x.push_back(pose.x);
    f.y.push_back(pose.y);
    f.yaw.push_back(yaw);
    margin = std::abs(util::getAngleDiff(yaw, goal_yaw));
  } while (margin >= step);
  if (f.yaw.back() != goal_yaw) f.yaw.back() = goal_yaw;
  f.close_prim(delta_angle, 1);
}

// the MotionPrimitive the reference's function returns, from primitive k of a fan
MotionPrimitive make_primitive(const Fan& f, int k, bool rear_axis)
{
  const size_t n = (size_t)f.count(k), o = (size_t)f.off[(size_t)k];
  MotionPrimitive mp;
  mp.x_list_.assign(f.x.begin() + (long)o, f.x.begin() + (long)(o + n));
  mp.y_list_.assign(f.y.begin() + (long)o, f.y.begin() + (long)(o + n));
  mp.yaw_list_.assign(f.yaw.begin() + (long)o, f.yaw.begin() + (long)(o + n));
  mp.nb_elements_ = n;
  if (rear_axis)
  {
    mp.dir_list_.assign(n, 1);
    mp.delta_list_.assign(n, f.key[(size_t)k] > 0 ? util::PI / 2.0 : -util::PI / 2.0);
  }
  else
  {
    mp.dir_list_.assign(n, f.dir[(size_t)k]);
    mp.delta_list_.assign(n, f.key[(size_t)k]);
  }
  return mp;
}
}  // namespace

MotionPrimitive Vehicle::move_car_some_steps(const Pose<double>& pose, double arc_l, double motion_res, int direction,
                                             double steer)
{
  Fan& f = g_move_fan;
  const bool same = f.valid && f.parent[0] == pose.x && f.parent[1] == pose.y && f.parent[2] == pose.yaw &&
                    f.arc_l == arc_l && f.motion_res == motion_res;
  if (!same)
  {
    // the fan HybridAStar::setNeighbors is about to ask for (hybrid_a_star_lib.cpp:492-510): calcSteeringInputs'
    // 13 steers (:115-129) x directions {+1, -1}
    f.clear();
    const double delta_steer = (2 * max_steer_) / static_cast<double>(NB_STEER - 1);
    const int side = NB_STEER / 2;
    for (int i = -side; i <= side; ++i)
      for (const int d : { 1, -1 }) roll_out(f, pose, arc_l, motion_res, d, i * delta_steer);
    f.parent[0] = pose.x;
    f.parent[1] = pose.y;
    f.parent[2] = pose.yaw;
    f.arc_l = arc_l;
    f.motion_res = motion_res;
    f.valid = true;
    g_ra_fan.valid = false;
    // answers: from the look-ahead when this node's fan was checked while its parent was expanded, else now
    // (submissions still in flight are only waited for when this node's answers are not on file yet)
    const PoseKey me = key_of(pose.x, pose.y, pose.yaw);
    const FanSlot* hit = g_lookahead_on ? g_lookahead.find(me) : nullptr;
    while (hit == nullptr && g_pend_count > 0)
    {
      lookahead_harvest();
      hit = g_lookahead.find(me);
    }
    if (hit != nullptr && hit->motion_res == motion_res)
    {
      f.first_hit.assign((size_t)f.size(), -1);
      g_x.clear();
      g_y.clear();
      g_yaw.clear();
      g_seg.assign(1, 0);
      std::vector<int> again;
      for (int k = 0; k < f.size(); ++k)
      {
        const int8_t a = hit->ans[k];
        if (a == 0)
          f.first_hit[(size_t)k] = 0;
        else if (a == 2)
        {
          const long o = f.off[(size_t)k], e = f.off[(size_t)k + 1];
          g_x.insert(g_x.end(), f.x.begin() + o, f.x.begin() + e);
          g_y.insert(g_y.end(), f.y.begin() + o, f.y.begin() + e);
          g_yaw.insert(g_yaw.end(), f.yaw.begin() + o, f.yaw.begin() + e);
          g_seg.push_back((int)g_x.size());
          again.push_back(k);
        }
      }
      if (!again.empty())
      {
        std::vector<int> h(again.size(), -1);
        check(ga_check_poses(ctx(), g_x.data(), g_y.data(), g_yaw.data(), (int)g_x.size(), g_seg.data(), (int)again.size(), 0,
                             h.data()));
        for (size_t i = 0; i < again.size(); ++i) f.first_hit[(size_t)again[i]] = h[i];
        g_stats.lookahead_rechecks += again.size();
      }
      ++g_stats.lookahead_hits;
    }
    else
      check_fan(f);
    lookahead_submit(f, arc_l);
  }
  for (int k = 0; k < f.size(); ++k)
    if (f.dir[(size_t)k] == direction && f.key[(size_t)k] == steer)
    {
      f.last = k;
      return make_primitive(f, k, false);
    }
  f.last = -1;  // a primitive outside the fan: rolled out alone, checked by its own launch
  Fan one;
  one.clear();
  roll_out(one, pose, arc_l, motion_res, direction, steer);
  return make_primitive(one, 0, false);
}

MotionPrimitive Vehicle::turn_on_rear_axis(const Pose<double>& pose, double delta_angle, double yaw_res_coll)
{
  Fan& f = g_ra_fan;
  const bool same = f.valid && f.parent[0] == pose.x && f.parent[1] == pose.y && f.parent[2] == pose.yaw &&
                    f.motion_res == yaw_res_coll;
  if (!same)
  {
    // the rotations setNeighbors tries (hybrid_a_star_lib.cpp:512-533), same loop arithmetic
    f.clear();
    const double a = g_turn_on_point_angle * util::TO_RAD;
    if (a > 1e-6)
      for (double d = -2 * util::PI + a; d <= 2 * util::PI - a; d += a)
      {
        if (d == 0) continue;
        rotate_on_axis(f, pose, d, yaw_res_coll);
      }
    f.parent[0] = pose.x;
    f.parent[1] = pose.y;
    f.parent[2] = pose.yaw;
    f.motion_res = yaw_res_coll;
    f.valid = true;
    check_fan(f);
  }
  for (int k = 0; k < f.size(); ++k)
    if (f.key[(size_t)k] == delta_angle)
    {
      f.last = k;
      return make_primitive(f, k, true);
    }
  f.last = -1;
  Fan one;
  one.clear();
  rotate_on_axis(one, pose, delta_angle, yaw_res_coll);
  return make_primitive(one, 0, true);
}

// ---- host mirrors of the astar-side maps that still have to reach the device ------------------------------
namespace
{
bool g_maps_dirty = false;  // host movement cost / geofence / carried-over h_prox, motion_res not yet on the device
int g_vor_dt_variant = GA_DT_IPPFLOAT;  // the reference's Docker OpenCV is built WITH_IPP

void flush_maps()
{
  if (!g_maps_dirty) return;
  ga_ctx* c = ctx();
  const int d = ga_astar_dim(c);
  if (d != AStar::astar_dim_ || d <= 0) return;  // device grids get their size with CollisionChecker::resetPatch
  const size_t n = (size_t)d * d;
  auto full = [d](const auto& grid) { return grid.getDims() == std::pair<int, int>(d, d); };
  if (full(AStar::movement_cost_map_)) check(ga_upload(c, GA_GRID_MOVEMENT_COST, AStar::movement_cost_map_.getPtr(), n * 8));
  if (full(AStar::restr_geofence_grid_)) check(ga_upload(c, GA_GRID_GEOFENCE, AStar::restr_geofence_grid_.getPtr(), n));
  if (full(AStar::h_prox_arr_)) check(ga_upload(c, GA_GRID_H_PROX, AStar::h_prox_arr_.getPtr(), n * 8));
  if (full(AStar::motion_res_map_)) check(ga_upload(c, GA_GRID_MOTION_RES, AStar::motion_res_map_.getPtr(), n * 8));
  check(ga_sync(c));
  g_maps_dirty = false;
}
}  // namespace

// =====================================================================================================
// CollisionChecker (LIB/src/collision_checker_lib/collision_checking.cpp)
// =====================================================================================================
void CollisionChecker::initialize(size_t patch_dim, const std::string& path2config)  // :12-31
{
  YAML::Node config = YAML::LoadFile(path2config);
  yaw_res_coll_ = config["YAW_RES_COLL"].as<int>();
  yaw_res_coll_rad_ = yaw_res_coll_ * util::TO_RAD;
  nb_disc_yaws_ = 360 / yaw_res_coll_;
  safety_distance_m_ = config["SAFETY_DISTANCE_M"].as<double>();
  search_dist_ = config["SEARCH_DIST"].as<double>();
  search_dist_cells_ = search_dist_ * grid_tf::con2gm_;
  min_thresh_ = config["MIN_THRESH"].as<int>();
  max_thresh_ = config["MAX_THRESH"].as<int>();
  len_per_disk_ = config["LEN_PER_DISK"].as<double>();
  double_disk_rows_ = config["DOUBLE_DISK_ROWS"].as<bool>();
  max_patch_ins_dist_ = config["MAX_PATCH_INS_DIST"].as<double>();
  try
  {
    g_turn_on_point_angle = config["TURN_ON_POINT_ANGLE"].as<double>();
    g_yaw_res_deg = config["YAW_RES"].as<double>();
  }
  catch (const std::exception&)
  {
  }
  calculateDisks();  // the device needs the footprint model before the first reset sizes the grids
  resetPatch(patch_dim);
}

double CollisionChecker::getDiskRadius(double length, double width, unsigned int nb_disks)  // :44-47
{
  return sqrt(((length * length) / (nb_disks * nb_disks) + (width * width)) / 4);
}

std::vector<double> CollisionChecker::getDiskPositions(double radius, unsigned int nb_disks, double width, double lb)  // :57-70
{
  std::vector<double> positions;
  const double dist = 2 * sqrt((radius * radius) - (width * width) / 4);
  const double first = dist / 2 - lb;
  for (unsigned int i = 0; i < nb_disks; ++i) positions.push_back(first + i * dist);
  return positions;
}

// :75-145 -- footprint disks and the per-yaw centre LUT in fp64 on the host, exactly as the reference fills
// disk_centers_ (Point::rotate, * con2gm_, toInt = std::round); the LUT and the dilation radius go to the device
void CollisionChecker::calculateDisks()
{
  const unsigned int eff_nb = ceil(Vehicle::length_ / len_per_disk_);
  const int rows = double_disk_rows_ ? 2 : 1;
  nb_disks_ = rows * eff_nb;
  const double eff_width = Vehicle::width_ / rows;
  const double radius = getDiskRadius(Vehicle::length_, eff_width, eff_nb);
  const std::vector<double> x_pos = getDiskPositions(radius, eff_nb, eff_width, Vehicle::lb_);
  disk_r_ = radius + safety_distance_m_;
  disk_r_c_ = static_cast<int>(ceil(disk_r_ * grid_tf::con2gm_));
  disk_diameter_c_ = 2 * disk_r_c_ + 1;

  disk_centers_.resize_and_reset(nb_disks_, nb_disc_yaws_, Point<int>(0, 0));
  std::vector<int> lut((size_t)nb_disc_yaws_ * nb_disks_ * 2);
  for (unsigned int yi = 0; yi < nb_disc_yaws_; ++yi)
  {
    const double yaw = yi * yaw_res_coll_rad_;
    unsigned int k = 0;
    for (const double xp : x_pos)
    {
      for (int r = 0; r < rows; ++r)
      {
        const double y_off = double_disk_rows_ ? (r == 0 ? eff_width / 2.0 : -eff_width / 2.0) : 0.0;
        const Point<double> centre = Point<double>(xp, y_off).rotate(yaw);
        const Point<int> cell = (centre * grid_tf::con2gm_).toInt();
        disk_centers_(yi, k) = cell;
        lut[((size_t)yi * nb_disks_ + k) * 2] = cell.x;
        lut[((size_t)yi * nb_disks_ + k) * 2 + 1] = cell.y;
        ++k;
      }
    }
  }
  const ga_collision_params p{ (int)min_thresh_, (int)max_thresh_, disk_r_c_, (int)yaw_res_coll_, (int)nb_disks_,
                               (int)ceil(grid_tf::star2gm_) };
  lookahead_reset();
  check(ga_set_collision_params(ctx(), &p, lut.data()));
  g_move_fan.valid = g_ra_fan.valid = false;
  g_expand_params_set = false;
}

std::vector<Point<double>> CollisionChecker::returnDiskPositions(double yaw)  // :320-331
{
  std::vector<Point<double>> points;
  points.reserve(nb_disks_);
  const int yaw_idx = getYawIdx(yaw);
  for (unsigned int d = 0; d < nb_disks_; ++d) points.push_back(grid_tf::grid2utm(disk_centers_(yaw_idx, d).toDouble()));
  return points;
}

void CollisionChecker::resetPatch(size_t patch_dim)  // :37-42; sizes the astar-side device grids too (ga_patch_reset)
{
  patch_dim_ = patch_dim;
  AStar::pushParams();
  check(ga_patch_reset(ctx(), (int)patch_dim));  // every device grid back to its initial value
  host_patch_stale_ = host_safety_stale_ = true;
  g_move_fan.valid = g_ra_fan.valid = false;
  lookahead_reset();
  g_maps_dirty = true;  // the host mirrors (lane costs, geofence, carried-over potential field) are authoritative
  flush_maps();
}

void CollisionChecker::processSafetyPatch()  // :193-224 (+ the 4x4 pooling of AStar::calcAstarGridCuda, fused)
{
  lookahead_reset();
  check(ga_build_safety(ctx()));
  host_safety_stale_ = true;
  g_move_fan.valid = g_ra_fan.valid = false;
}

void CollisionChecker::syncPatch()
{
  patch_arr_.resize_and_reset(patch_dim_, patch_dim_, SENSOR_UNKNOWN);
  check(ga_download(ctx(), GA_GRID_PATCH, patch_arr_.getPtr(), patch_dim_ * patch_dim_));
  host_patch_stale_ = false;
}
void CollisionChecker::syncSafetyPatch()
{
  patch_safety_arr_.resize_and_reset(patch_dim_, patch_dim_, UNKNOWN);
  check(ga_download(ctx(), GA_GRID_SAFETY, patch_safety_arr_.getPtr(), patch_dim_ * patch_dim_));
  host_safety_stale_ = false;
}

void CollisionChecker::insertMinipatches(const std::map<std::pair<int, int>, Minipatch>& minipatches,
                                         const Point<double>& ego_utm, bool only_nearest, bool only_new)  // :147-174
{
  for (const auto& [patch_idx, minipatch] : minipatches)
  {
    if (only_new and !minipatch.is_new_) continue;
    if (only_nearest && minipatch.center_.dist2(ego_utm) > max_patch_ins_dist_) continue;
    const auto origin_grid = grid_tf::utm2grid_round(grid_tf::utm2patch_utm(minipatch.origin_));
    passLocalMap(minipatch.patch_, origin_grid, minipatch.width_);
  }
}

void CollisionChecker::passLocalMap(const py::array_t<uint8_t>& local_map, const Point<int>& origin, int dim)
{
  passLocalMapData(local_map.data(), origin, dim);
}
void CollisionChecker::passLocalMap(const Vec2DFlat<uint8_t>& local_map, const Point<int>& origin, int dim)
{
  passLocalMapData(const_cast<Vec2DFlat<uint8_t>&>(local_map).getPtr(), origin, dim);  // the const getPtr() of Vec2DFlat does not compile
}
// :269-313 -- only the window crosses PCIe; the clipping rules are applied by ga_patch_insert
void CollisionChecker::passLocalMapData(const unsigned char* local_map, const Point<int>& origin, int dim)
{
  check(ga_patch_insert(ctx(), local_map, origin.x, origin.y, dim));
  check(ga_sync(ctx()));
  host_patch_stale_ = true;
}

int CollisionChecker::getYawIdx(double yaw)  // :333-337
{
  const int yaw_idx = static_cast<int>(util::constrainAngleZero2Pi(yaw) / yaw_res_coll_rad_);
  return std::min(yaw_idx, static_cast<int>(nb_disc_yaws_ - 1));
}

namespace
{
int first_hit(const double* x, const double* y, const double* yaw, int n, int variant)
{
  const int seg[2] = { 0, n };
  int hit = -1;
  check(ga_check_poses(ctx(), x, y, yaw, n, seg, 1, variant, &hit));
  g_stats.poses_checked += (unsigned long long)n;
  return hit;
}
}  // namespace

// :344-387 take CELL coordinates; the centre of the cell maps back to the same cell and yaw bin on the device
bool CollisionChecker::checkGrid(const Pose<int>& pose)
{
  if (pose.x < 0 || pose.y < 0) return false;  // every disk lookup of :351-354 would be off-map or the pose is
  const double x = (pose.x + 0.5) * grid_tf::gm2con_, y = (pose.y + 0.5) * grid_tf::gm2con_;
  const double yaw = (pose.yaw + 0.5) * yaw_res_coll_rad_;
  return first_hit(&x, &y, &yaw, 1, 0) < 0;
}
bool CollisionChecker::checkGridVariant(const Pose<int>& pose)
{
  if (pose.x < 0 || pose.y < 0)
  {
    // a negative cell cannot be expressed in metres (truncation toward zero): resolve the few lookups on a host copy
    if (host_safety_stale_) syncSafetyPatch();
    for (unsigned int d = 0; d < nb_disks_; ++d)
    {
      const Point<int> c = disk_centers_(pose.yaw, d) + Point<int>(pose.x, pose.y);
      if (c.x < 0 || c.y < 0 || c.x >= (int)patch_dim_ || c.y >= (int)patch_dim_) continue;
      if (patch_safety_arr_(c.y, c.x) == OCC) return false;
    }
    return true;
  }
  const double x = (pose.x + 0.5) * grid_tf::gm2con_, y = (pose.y + 0.5) * grid_tf::gm2con_;
  const double yaw = (pose.yaw + 0.5) * yaw_res_coll_rad_;
  return first_hit(&x, &y, &yaw, 1, 1) < 0;
}
bool CollisionChecker::checkPose(const Pose<double>& pose) { return first_hit(&pose.x, &pose.y, &pose.yaw, 1, 0) < 0; }
bool CollisionChecker::checkPoseVariant(const Pose<double>& pose) { return first_hit(&pose.x, &pose.y, &pose.yaw, 1, 1) < 0; }

int CollisionChecker::getPathCollisionIndex(const std::vector<double>& x_list, const std::vector<double>& y_list,
                                            const std::vector<double>& yaw_list)  // :440-452
{
  if (x_list.empty()) return -1;
  ++g_stats.path_checks_direct;
  return first_hit(x_list.data(), y_list.data(), yaw_list.data(), (int)x_list.size(), 0);
}

bool CollisionChecker::checkPathCollision(const std::vector<double>& x_list, const std::vector<double>& y_list,
                                          const std::vector<double>& yaw_list)  // :418-431
{
  // a primitive of the fan the last expansion rolled out: its answer came back with that launch
  for (Fan* f : { &g_move_fan, &g_ra_fan })
  {
    const int k = f->last >= 0 ? match_fan(*f, x_list, y_list, yaw_list) : -1;
    if (k >= 0)
    {
      ++g_stats.path_checks_cached;
      return f->first_hit[(size_t)k] < 0;
    }
  }
  return getPathCollisionIndex(x_list, y_list, yaw_list) < 0;
}

// =====================================================================================================
// AStar (LIB/src/hybridastar_planning_lib/a_star.cpp)
// =====================================================================================================

namespace ga_dropin
{
void setDistanceTransformVariant(int ga_dt_variant)
{
  g_vor_dt_variant = ga_dt_variant;
  AStar::pushParams();
}
}  // namespace ga_dropin

void AStar::pushParams()
{
  ga_astar_params p{};
  p.astar_movement_cost = astar_movement_cost_;
  p.astar_prox_cost = astar_prox_cost_;
  p.astar_lane_movement_cost = astar_lane_movement_cost_;
  p.unknown_cost_w = unknown_cost_w_;
  p.alpha = alpha_;
  p.do_max = do_max_;
  p.do_min = do_min_;
  p.motion_res_min = motion_res_min_;
  p.motion_res_max = motion_res_max_;
  p.dist_val_min = dist_val_min_;
  p.dist_val_max = dist_val_max_;
  p.vor_dim = (int)VOR_DIM;
  p.dt_variant = g_vor_dt_variant;
  check(ga_set_astar_params(ctx(), &p));
  params_pushed_ = true;
}

void AStar::initialize(int patch_dim, const Point<double>& patch_origin_utm, const std::string& path2config)  // :11-53
{
  path2config_ = path2config;
  YAML::Node config = YAML::LoadFile(path2config_);
  patch_origin_astar_ = (patch_origin_utm * grid_tf::con2star_).toInt();
  patch_dim_ = patch_dim;
  unknown_cost_w_ = config["ASTAR_UNKNOWN_COST"].as<double>();
  astar_dim_ = std::floor(static_cast<double>(patch_dim_) * GM_RES / ASTAR_RES);
  heuristic_early_exit_ = config["HEURISTIC_EARLY_EXIT"].as<bool>();
  max_extra_nodes_ = config["MAX_EXTRA_NODES_ASTAR"].as<unsigned int>();
  motion_res_min_ = config["MOTION_RES_MIN"].as<double>();
  motion_res_max_ = config["MOTION_RES_MAX"].as<double>();
  dist_val_min_ = config["MIN_DIST_VAL"].as<double>();
  dist_val_max_ = config["MAX_DIST_VAL"].as<double>();
  init_structs(patch_dim);
  // Pooling::init: the 4 x 4 max is part of the fused safety kernel, nothing to set up
}

void AStar::init_structs(int patch_dim)  // :55-68
{
  patch_dim_ = patch_dim;
  astar_dim_ = std::floor(static_cast<double>(patch_dim_) * GM_RES / ASTAR_RES);
  astar_grid_.resize_and_reset(astar_dim_, astar_dim_, CollisionChecker::UNKNOWN);
  h_prox_arr_.resize_and_reset(astar_dim_, astar_dim_, 0.0);
  motion_res_map_.resize_and_reset(astar_dim_, astar_dim_, motion_res_max_);
  restr_geofence_grid_.resize_and_reset(astar_dim_, astar_dim_, 0);
  movement_cost_map_.resize_and_reset(astar_dim_, astar_dim_, astar_movement_cost_);
  astar_grid_stale_ = true;
  g_maps_dirty = true;
}

// :70-99 -- incl. the carry-over of h_prox_arr_ / motion_res_map_ across a patch move (saveTemp + copyPatch2Patch on
// the host mirrors; they reach the device once CollisionChecker::resetPatch has resized its grids)
void AStar::reinit(const Point<double>& patch_origin_utm, int patch_dim)
{
  patch_dim_ = patch_dim;
  astar_dim_ = std::floor(static_cast<double>(patch_dim_) * GM_RES / ASTAR_RES);
  patch_origin_utm_ = patch_origin_utm;
  astar_grid_.resize_and_reset(astar_dim_, astar_dim_, CollisionChecker::UNKNOWN);
  astar_grid_stale_ = true;
  obstacle_dist_.resize_and_reset(VOR_DIM, VOR_DIM, CollisionChecker::FREE);
  voronoi_dist_.resize_and_reset(VOR_DIM, VOR_DIM, CollisionChecker::FREE);

  util::saveTemp(h_prox_arr_, temp_h_prox_arr_, 0.0);
  h_prox_arr_.resize_and_reset(astar_dim_, astar_dim_, 0.0);
  util::saveTemp(motion_res_map_, temp_motion_res_map_, motion_res_max_);
  motion_res_map_.resize_and_reset(astar_dim_, astar_dim_, motion_res_max_);
  movement_cost_map_.resize_and_reset(astar_dim_, astar_dim_, astar_movement_cost_);

  processGeofence();

  const Point<int> next_origin_astar = (patch_origin_utm * grid_tf::con2star_).toInt();
  util::copyPatch2Patch(patch_origin_astar_, next_origin_astar, h_prox_arr_, temp_h_prox_arr_);
  util::copyPatch2Patch(patch_origin_astar_, next_origin_astar, motion_res_map_, temp_motion_res_map_);
  patch_origin_astar_ = next_origin_astar;

  closed_set_path_.clear();
  closed_set_guidance_.clear();
  g_maps_dirty = true;
}

size_t AStar::calcIndex(size_t x_ind, size_t y_ind) { return y_ind * astar_dim_ + x_ind; }  // :303-316
size_t AStar::calcIndex(const NodeDisc& node) { return calcIndex(node.pos.x, node.pos.y); }

Point<int> AStar::reverse2DIndex(int idx, int dim)  // :515-519
{
  const int first = idx % dim;
  return { first, (idx - first) / dim };
}

Point<int> AStar::getCurrentMapOrigin(const Point<int>& ego_pos, size_t dim)  // :527-531
{
  const int offset = std::floor(dim / 2);
  return { std::max(ego_pos.x - offset, 0), std::max(ego_pos.y - offset, 0) };
}

// :536-542 -- Pooling::execute is fused into processSafetyPatch's kernel: astar_grid_ on the device is current; the
// host copy the reference's callers (GUI) read follows here (astar_dim^2 bytes)
void AStar::calcAstarGridCuda() { syncAstarGrid(); }

void AStar::syncAstarGrid()
{
  astar_grid_.resize_and_reset(astar_dim_, astar_dim_, CollisionChecker::UNKNOWN);
  check(ga_download(ctx(), GA_GRID_ASTAR, astar_grid_.getPtr(), (size_t)astar_dim_ * astar_dim_));
  astar_grid_stale_ = false;
}
void AStar::syncDistanceGrids()
{
  obstacle_dist_.resize_and_reset(VOR_DIM, VOR_DIM, 0.0f);
  voronoi_dist_.resize_and_reset(VOR_DIM, VOR_DIM, 0.0f);
  check(ga_download(ctx(), GA_GRID_OBSTACLE_DIST, obstacle_dist_.getPtr(), VOR_DIM * VOR_DIM * sizeof(float)));
  check(ga_download(ctx(), GA_GRID_VORONOI_DIST, voronoi_dist_.getPtr(), VOR_DIM * VOR_DIM * sizeof(float)));
}

// :440-479 -- both distance transforms and setCostmapValue run on the device; the Fortune sweep between them is the
// reference's own serial fp64 code (deps_lib/VoronoiDiagramGenerator.cpp) on the host
void AStar::calcVoronoiPotentialField(const Point<int>& ego_index)
{
  pushParams();  // alpha_, do_max_, do_min_ are plain statics the caller may have changed
  flush_maps();
  ga_ctx* c = ctx();
  static std::vector<int> sites(2 * VOR_DIM * VOR_DIM);
  int n_sites = 0;
  check(ga_potential_sites(c, ego_index.x, ego_index.y, sites.data(), (int)(VOR_DIM * VOR_DIM), &n_sites));
  std::vector<int> vor;
  if (n_sites > 0)
  {
    std::vector<cv::Point> indices((size_t)n_sites);  // cv::findNonZero order (row-major), a_star.cpp:457-459
    for (int i = 0; i < n_sites; ++i) indices[(size_t)i] = cv::Point(sites[2 * (size_t)i], sites[2 * (size_t)i + 1]);
    static Vec2DFlat<float> marks;
    marks.resize_and_reset(VOR_DIM, VOR_DIM, 0.0f);
    vdg::VoronoiDiagramGenerator::generateVoronoi_wrapper(indices, marks, VOR_DIM - 1, VOR_DIM - 1, do_min_, CollisionChecker::OCC);
    const float* m = marks.getPtr();
    for (int y = 0; y < (int)VOR_DIM; ++y)
      for (int x = 0; x < (int)VOR_DIM; ++x)
        if (m[(size_t)y * VOR_DIM + x] == (float)CollisionChecker::OCC)
        {
          vor.push_back(x);
          vor.push_back(y);
        }
  }
  int had = 0;
  check(ga_build_potential(c, ego_index.x, ego_index.y, vor.empty() ? nullptr : vor.data(), (int)(vor.size() / 2), &had));
  if (!had) return;  // no obstacle in the window: the old field stays (:462-465)
  // the window of h_prox_arr_ / motion_res_map_ that changed, back into the host mirrors the search reads
  const Point<int> o = getCurrentMapOrigin(ego_index, VOR_DIM);
  static std::vector<double> win(VOR_DIM * VOR_DIM);
  for (const int id : { (int)GA_GRID_H_PROX, (int)GA_GRID_MOTION_RES })
  {
    check(ga_download_window(c, id, o.x, o.y, (int)VOR_DIM, (int)VOR_DIM, win.data()));
    Vec2DFlat<double>& dst = id == GA_GRID_H_PROX ? h_prox_arr_ : motion_res_map_;
    for (size_t y = 0; y < VOR_DIM; ++y)
      std::memcpy(dst.getPtr() + (size_t)(o.y + (int)y) * astar_dim_ + o.x, win.data() + y * VOR_DIM, VOR_DIM * sizeof(double));
  }
}

// :110-295
void AStar::calcDistanceHeuristic(const Point<int>& goal_pos, const Point<int>& start_pos, bool for_path, bool get_only_near)
{
  pushParams();
  flush_maps();
  ga_ctx* c = ctx();
  const int slot = for_path ? 1 : 0;
  std::unordered_map<size_t, NodeDisc>& cs = for_path ? closed_set_path_ : closed_set_guidance_;
  ga_heur_query q{};
  q.goal_x = goal_pos.x;
  q.goal_y = goal_pos.y;
  q.start_x = start_pos.x;
  q.start_y = start_pos.y;
  q.max_extra_nodes = max_extra_nodes_;
  q.flags = GA_HEUR_FIXEDPOINT;
  if (heuristicMode() & GA_HEUR_REFORDER)
    q.flags = GA_HEUR_REFORDER | (heuristic_early_exit_ ? GA_HEUR_EARLY_EXIT : 0) | (get_only_near ? GA_HEUR_ONLY_NEAR : 0);
  int n = 0;
  check(ga_build_heuristic_ex(c, slot, &q, &n));
  const size_t cells = (size_t)astar_dim_ * astar_dim_;
  ClosedPlanes& p = cs.planes;
  p.dim = astar_dim_;
  p.cells = cells;
  if (q.flags & GA_HEUR_REFORDER)
  {
    p.packed = false;
    p.cost.resize(cells);
    p.dist.resize(cells);
    p.parent.resize(cells);
    p.flags.resize(cells);
    check(ga_download(c, GA_GRID_HEUR_COST + slot, p.cost.data(), cells * 8));
    check(ga_download(c, GA_GRID_HEUR_DIST + slot, p.dist.data(), cells * 8));
    check(ga_download(c, GA_GRID_HEUR_PARENT + slot, p.parent.data(), cells * 4));
    check(ga_download(c, GA_GRID_HEUR_FLAGS + slot, p.flags.data(), cells));
  }
  else
  {
    // device build: the planes come as 12-byte records and are read in place by the closed set's accessors
    p.rec.resize(3 * cells);
    if (ga_download(c, GA_GRID_HEUR_REC + slot, p.rec.data(), cells * 12) == GA_OK)
      p.packed = true;
    else  // a parent chain too long for the record format: the four planes
    {
      p.packed = false;
      p.cost.resize(cells);
      p.dist.resize(cells);
      p.parent.resize(cells);
      p.flags.resize(cells);
      check(ga_download(c, GA_GRID_HEUR_COST + slot, p.cost.data(), cells * 8));
      check(ga_download(c, GA_GRID_HEUR_DIST + slot, p.dist.data(), cells * 8));
      check(ga_download(c, GA_GRID_HEUR_PARENT + slot, p.parent.data(), cells * 4));
      check(ga_download(c, GA_GRID_HEUR_FLAGS + slot, p.flags.data(), cells));
    }
  }
  p.n_closed = (size_t)n;

  nodes_near_goal_.clear();
  if (get_only_near)
  {
    int ids[400];
    double dist[400];
    int cnt = 0;
    check(ga_heur_near_nodes(c, slot, start_pos.x, start_pos.y, ids, dist, 400, &cnt));
    for (int i = 0; i < cnt && i < 400; ++i) nodes_near_goal_.emplace_back(dist[i], cs.at((size_t)ids[i]));
  }
}

int AStar::findValidNeighborIndex(int start_idx, const std::unordered_map<size_t, NodeDisc>& heuristic)  // :351-378
{
  for (const int cand : { start_idx + 1, start_idx - 1, start_idx + astar_dim_, start_idx - astar_dim_ })
    if (cand >= 0 && heuristic.contains((size_t)cand)) return cand;
  std::cout << "No node around the ego/start position was in heuristic, this should never be the case" << std::endl;
  return -1;
}

std::pair<std::vector<int>, std::vector<int>> AStar::getAstarPath(int x_index, int y_index)  // :386-421
{
  std::vector<int> x_list, y_list;
  size_t idx = calcIndex(x_index, y_index);
  if (!closed_set_guidance_.contains(idx))
  {
    const int index = findValidNeighborIndex(static_cast<int>(idx), closed_set_guidance_);
    if (index == -1) return { x_list, y_list };
    idx = static_cast<size_t>(index);
  }
  const ClosedPlanes& p = closed_set_guidance_.planes;
  for (int cur = (int)idx; cur != -1; cur = p.parent_of((size_t)cur))
  {
    x_list.push_back(cur % astar_dim_);
    y_list.push_back(cur / astar_dim_);
  }
  return { x_list, y_list };
}

void AStar::resetMovementMap()  // :544-547
{
  movement_cost_map_.resize_and_reset(astar_dim_, astar_dim_, astar_movement_cost_);
  g_maps_dirty = true;
}

// :553-602 -- cells along the lane-graph edges (Bresenham, 3 x 3 brush) get the lane movement cost
void AStar::setMovementMap(const LaneGraph::edges_t& edges)
{
  auto brush = [](const Point<int>& from, const Point<int>& to) {
    for (const auto& point : util::drawline(from, to))
      for (const auto& p_diff : patch_coords_)
      {
        const Point<int> q = point + p_diff;
        if (q.x < 0 or q.x >= astar_dim_ or q.y < 0 or q.y >= astar_dim_) continue;
        movement_cost_map_(q) = astar_lane_movement_cost_;
      }
  };
  for (const auto& edge : edges)
  {
    const auto& [n1_opt, node_center, n3_opt] = edge;
    const Point<int> p_center = (node_center.point_utm_patch * grid_tf::con2star_).toInt();
    if (n1_opt) brush(p_center, ((*n1_opt).point_utm_patch * grid_tf::con2star_).toInt());
    if (n3_opt) brush(p_center, ((*n3_opt).point_utm_patch * grid_tf::con2star_).toInt());
  }
  g_maps_dirty = true;
}

// :607-652 -- restricted-area polygon rasterised into restr_geofence_grid_ (the reference's loop bounds, including
// its inner loop starting at min_x)
void AStar::processGeofence()
{
  restr_geofence_grid_.resize_and_reset(astar_dim_, astar_dim_, 0);
  int min_x = astar_dim_, min_y = astar_dim_, max_x = 0, max_y = 0;
  for (auto& point : restr_geofence_.vertices)
  {
    const Point<int> pa = ((point - patch_origin_utm_) * grid_tf::con2star_).toInt();
    min_x = std::min(min_x, pa.x);
    min_y = std::min(min_y, pa.y);
    max_x = std::max(max_x, pa.x);
    max_y = std::max(max_y, pa.y);
  }
  for (int x_idx = min_x; x_idx < max_x; ++x_idx)
    for (int y_idx = min_x; y_idx < max_y; ++y_idx)
    {
      if (x_idx < 0 || y_idx < 0 || x_idx >= astar_dim_ || y_idx >= astar_dim_) continue;  // the reference writes out of bounds here
      const Point<int> pa(x_idx, y_idx);
      const Point<double> p_utm = pa.toDouble() * grid_tf::star2con_ + patch_origin_utm_;
      if (util::point_in_poly(restr_geofence_, p_utm)) restr_geofence_grid_(pa) = 1;
    }
  g_maps_dirty = true;
}
