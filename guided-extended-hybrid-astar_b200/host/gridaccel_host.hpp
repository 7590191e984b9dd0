// gridaccel_host.hpp -- C++ host mirror of the reference's grid classes on top of the C ABI
// (include/gridaccel.h, libgridaccel.so).  Header-only, C++17, no CUDA headers needed.
//
// The reference keeps this path behind two all-static classes (LIB = src/freespace_planner/library):
//   CollisionChecker  LIB/include/collision_checker_lib/collision_checking.hpp:36-120
//   AStar             LIB/include/hybridastar_planning_lib/a_star.hpp:32-140
// plus Vehicle (collision_checker_lib/vehicle.hpp:15-45) and the containers Point / Pose /
// Vec2DFlat / NodeDisc (util_lib/data_structures1.hpp:39-514, data_structures2.hpp:47-67).
// The classes below keep the reference's member names, argument meaning and call order, so the
// serial Hybrid A* search (hybrid_a_star_lib.cpp) reads the same; state lives in a ga_ctx (one per
// planner instance) instead of inline statics.  Every grid computation runs in libgridaccel.so on
// the GPU: there is no CPU path here, a missing device or library is a std::runtime_error.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "gridaccel.h"

namespace gridaccel
{
constexpr double ASTAR_RES = 0.625;   // LIB/include/params.hpp:7
constexpr double GM_RES = 0.15625;    // params.hpp:8
constexpr double CON2GM = 1.0 / GM_RES;
constexpr double CON2STAR = 1.0 / ASTAR_RES;
constexpr double PI = 3.14159265358979323846;
constexpr double TO_RAD = PI / 180.0;

template <typename T>
struct Point
{
  T x{}, y{};
  Point() = default;
  Point(T x_, T y_) : x(x_), y(y_) {}
};
template <typename T>
struct Pose
{
  T x{}, y{}, yaw{};
  Pose() = default;
  Pose(T x_, T y_, T yaw_) : x(x_), y(y_), yaw(yaw_) {}
};

// Vec2DFlat<T> (data_structures1.hpp:338-514): dense row-major, (y, x) -> vec_[y * xDim + x]
template <typename T>
class Vec2DFlat
{
public:
  void resize_and_reset(size_t x_dim, size_t y_dim, T val)
  {
    xDim_ = x_dim;
    yDim_ = y_dim;
    vec_.assign(x_dim * y_dim, val);
  }
  T& operator()(size_t y, size_t x) { return vec_[y * xDim_ + x]; }
  const T& operator()(size_t y, size_t x) const { return vec_[y * xDim_ + x]; }
  T& operator()(const Point<int>& p) { return vec_[(size_t)p.y * xDim_ + p.x]; }
  const T& operator()(const Point<int>& p) const { return vec_[(size_t)p.y * xDim_ + p.x]; }
  T* getPtr() { return vec_.data(); }
  const T* getPtr() const { return vec_.data(); }
  size_t size() const { return vec_.size(); }
  size_t xDim() const { return xDim_; }
  size_t yDim() const { return yDim_; }

private:
  std::vector<T> vec_;
  size_t xDim_ = 0, yDim_ = 0;
};

// NodeDisc (data_structures2.hpp:47-67)
struct NodeDisc
{
  Point<int> pos;
  double cost_ = 0, cost_dist_ = 0;
  int parent_index_ = -1;
  bool is_unknown_ = false;
};

// flat "KEY: value  # comment" reader for LIB/config/config.yml and params.yml (yaml-cpp stand-in)
class Config
{
public:
  std::map<std::string, std::string> kv;
  void load(const std::string& path)
  {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("gridaccel: cannot open config " + path);
    std::string line;
    while (std::getline(f, line))
    {
      const size_t hash = line.find('#');
      if (hash != std::string::npos) line.erase(hash);
      const size_t colon = line.find(':');
      if (colon == std::string::npos) continue;
      const std::string k = trim(line.substr(0, colon)), v = trim(line.substr(colon + 1));
      if (!k.empty() && !v.empty()) kv[k] = v;
    }
  }
  double num(const std::string& k, double dflt) const
  {
    auto it = kv.find(k);
    return it == kv.end() ? dflt : std::stod(it->second);
  }
  bool flag(const std::string& k, bool dflt) const
  {
    auto it = kv.find(k);
    if (it == kv.end()) return dflt;
    return it->second == "True" || it->second == "true" || it->second == "1";
  }

private:
  static std::string trim(const std::string& s)
  {
    const size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
  }
};

// one planner instance = one ga_ctx (device arenas + one CUDA stream)
class Context
{
public:
  explicit Context(int device = 0)
  {
    if (ga_abi_version() != GA_ABI_VERSION) throw std::runtime_error("gridaccel: ABI version mismatch");
    const int rc = ga_create(device, &h_);
    if (rc != GA_OK) throw std::runtime_error("gridaccel: ga_create failed (no usable CUDA device; there is no CPU path)");
  }
  ~Context()
  {
    if (h_) ga_destroy(h_);
  }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  ga_ctx* handle() const { return h_; }
  void check(int rc) const
  {
    if (rc != GA_OK) throw std::runtime_error(std::string("gridaccel: ") + ga_last_error(h_));
  }
  void sync() const { check(ga_sync(h_)); }
  // ga_patch_reset sizes the astar-side grids too: a CollisionChecker used without an AStar (the
  // reference's gtest does that) gets the shipped config.yml / params.yml values
  bool astar_params_set = false;
  void ensureAstarParams()
  {
    if (astar_params_set) return;
    ga_astar_params p{ 5.0, 5.0, 3.0, 0.0, 100.0, 7.0, 3.0, 0.1, 1.0, 3.0, 10.0, 200, GA_DT_IPPFLOAT };
    check(ga_set_astar_params(h_, &p));
    astar_params_set = true;
  }
  template <typename T>
  void download(int grid_id, Vec2DFlat<T>& out, size_t dim) const
  {
    if (out.xDim() != dim || out.yDim() != dim) out.resize_and_reset(dim, dim, T());
    check(ga_download(h_, grid_id, out.getPtr(), dim * dim * sizeof(T)));
  }
  template <typename T>
  void upload(int grid_id, const Vec2DFlat<T>& in) const
  {
    check(ga_upload(h_, grid_id, in.getPtr(), in.size() * sizeof(T)));
    sync();
  }

private:
  ga_ctx* h_ = nullptr;
};

// Vehicle (vehicle.hpp:15-45, vehicle.cpp:11-32)
struct Vehicle
{
  double max_steer_ = 0.55, w_b_ = 3.165, lf_ = 3.993, lb_ = 1.17, length_ = 5.163, width_ = 2.13;
  bool is_ushift_ = false;
  void initialize(double max_steer, double w_b, double l_f, double l_b, double width, bool is_ushift)
  {
    max_steer_ = max_steer;
    w_b_ = w_b;
    lf_ = l_f;
    lb_ = l_b;
    length_ = l_f + l_b;
    width_ = width;
    is_ushift_ = is_ushift;
  }
};

// CollisionChecker (collision_checking.hpp:36-120)
class CollisionChecker
{
public:
  enum GRID_VAL { FREE = 0, UNKNOWN = 1, OCC = 2, SENSOR_FREE = 255, SENSOR_UNKNOWN = 127, SENSOR_OCC = 0 };

  double max_patch_ins_dist_ = 40, disk_r_ = 0;
  int disk_r_c_ = 0;
  Vec2DFlat<uint8_t> patch_arr_, patch_safety_arr_;  // host copies, refreshed by fetchPatch / fetchSafetyPatch

  CollisionChecker(Context& ctx, const Vehicle& veh) : ctx_(ctx), veh_(veh) {}

  // collision_checking.cpp:12-31
  void initialize(size_t patch_dim, const std::string& path2config)
  {
    Config cfg;
    cfg.load(path2config);
    configure(cfg);
    resetPatch(patch_dim);
  }
  void configure(const Config& cfg)
  {
    yaw_res_coll_ = (int)cfg.num("YAW_RES_COLL", 3);
    safety_distance_m_ = cfg.num("SAFETY_DISTANCE_M", 0.1);
    min_thresh_ = (int)cfg.num("MIN_THRESH", 100);
    max_thresh_ = (int)cfg.num("MAX_THRESH", 130);
    len_per_disk_ = cfg.num("LEN_PER_DISK", 2.0);
    double_disk_rows_ = cfg.flag("DOUBLE_DISK_ROWS", false);
    max_patch_ins_dist_ = cfg.num("MAX_PATCH_INS_DIST", 40);
    yaw_res_coll_rad_ = yaw_res_coll_ * TO_RAD;
    nb_disc_yaws_ = 360 / yaw_res_coll_;
    calculateDisks();
  }

  static double getDiskRadius(double length, double width, unsigned nb_disks)  // :44-47
  {
    return std::sqrt(((length * length) / (nb_disks * nb_disks) + (width * width)) / 4);
  }
  static std::vector<double> getDiskPositions(double radius, unsigned nb_disks, double width, double lb)  // :57-70
  {
    std::vector<double> pos;
    const double dist = 2 * std::sqrt((radius * radius) - (width * width) / 4);
    const double first = dist / 2 - lb;
    for (unsigned i = 0; i < nb_disks; ++i) pos.push_back(first + i * dist);
    return pos;
  }

  // collision_checking.cpp:75-145: footprint disks + per-yaw centre LUT, fp64 on the host exactly as
  // the reference initialises them (Point::rotate, then * con2gm_, then std::round)
  void calculateDisks()
  {
    const unsigned eff_nb = (unsigned)std::ceil(veh_.length_ / len_per_disk_);
    const int mult = double_disk_rows_ ? 2 : 1;
    nb_disks_ = mult * eff_nb;
    const double eff_width = veh_.width_ / mult;
    const double radius = getDiskRadius(veh_.length_, eff_width, eff_nb);
    const std::vector<double> positions = getDiskPositions(radius, eff_nb, eff_width, veh_.lb_);
    disk_r_ = radius + safety_distance_m_;
    disk_r_c_ = (int)std::ceil(disk_r_ * CON2GM);
    disk_centers_.assign((size_t)nb_disc_yaws_ * nb_disks_ * 2, 0);
    for (int yi = 0; yi < nb_disc_yaws_; ++yi)
    {
      const double yaw = yi * yaw_res_coll_rad_;
      const double c = std::cos(yaw), s = std::sin(yaw);
      int k = 0;
      for (const double pos : positions)
      {
        const double offs[2] = { double_disk_rows_ ? eff_width / 2.0 : 0.0, -eff_width / 2.0 };
        for (int o = 0; o < mult; ++o)
        {
          const double rx = pos * c - offs[o] * s, ry = pos * s + offs[o] * c;
          disk_centers_[((size_t)yi * nb_disks_ + k) * 2] = (int)std::round(rx * CON2GM);
          disk_centers_[((size_t)yi * nb_disks_ + k) * 2 + 1] = (int)std::round(ry * CON2GM);
          ++k;
        }
      }
    }
    ga_collision_params p{ min_thresh_, max_thresh_, disk_r_c_, yaw_res_coll_, (int)nb_disks_, 4 };
    ctx_.check(ga_set_collision_params(ctx_.handle(), &p, disk_centers_.data()));
  }

  void resetPatch(size_t patch_dim)  // :37-42
  {
    patch_dim_ = patch_dim;
    ctx_.ensureAstarParams();
    ctx_.check(ga_patch_reset(ctx_.handle(), (int)patch_dim));
  }

  // :269-313 -- only the window crosses PCIe, the patch stays on the device
  void passLocalMapData(const unsigned char* local_map, const Point<int>& origin, int dim)
  {
    ctx_.check(ga_patch_insert(ctx_.handle(), local_map, origin.x, origin.y, dim));
    ctx_.sync();
  }
  void passLocalMap(const Vec2DFlat<uint8_t>& local_map, const Point<int>& origin, int dim)
  {
    passLocalMapData(local_map.getPtr(), origin, dim);
  }

  // :193-224 (+ Pooling::execute, fused: AStar::astar_grid_ is current afterwards)
  void processSafetyPatch() { ctx_.check(ga_build_safety(ctx_.handle())); }

  void fetchPatch() { ctx_.download(GA_GRID_PATCH, patch_arr_, patch_dim_); }
  void fetchSafetyPatch() { ctx_.download(GA_GRID_SAFETY, patch_safety_arr_, patch_dim_); }

  int getYawIdx(double yaw) const  // :333-337, util1.cpp:8-15
  {
    while (yaw < 0) yaw += 2 * PI;
    yaw = std::fmod(yaw, 2 * PI);
    return std::min((int)(yaw / yaw_res_coll_rad_), nb_disc_yaws_ - 1);
  }

  std::vector<Point<double>> returnDiskPositions(double yaw) const  // :320-331 (patch frame, metres)
  {
    std::vector<Point<double>> pts;
    const int yi = getYawIdx(yaw);
    for (unsigned d = 0; d < nb_disks_; ++d)
      pts.emplace_back(disk_centers_[((size_t)yi * nb_disks_ + d) * 2] * GM_RES,
                       disk_centers_[((size_t)yi * nb_disks_ + d) * 2 + 1] * GM_RES);
    return pts;
  }

  // :344-409.  checkGrid takes cell coordinates; the cell centre maps back to the same cell and bin.
  bool checkGrid(const Pose<int>& pose) { return gridCheck(pose, 0); }
  bool checkGridVariant(const Pose<int>& pose) { return gridCheck(pose, 1); }
  bool checkPose(const Pose<double>& pose) { return firstHit(&pose.x, &pose.y, &pose.yaw, 1, 0) < 0; }
  bool checkPoseVariant(const Pose<double>& pose) { return firstHit(&pose.x, &pose.y, &pose.yaw, 1, 1) < 0; }

  // :418-452
  int getPathCollisionIndex(const std::vector<double>& x, const std::vector<double>& y, const std::vector<double>& yaw)
  {
    if (x.empty()) return -1;
    return firstHit(x.data(), y.data(), yaw.data(), (int)x.size(), 0);
  }
  bool checkPathCollision(const std::vector<double>& x, const std::vector<double>& y, const std::vector<double>& yaw)
  {
    return getPathCollisionIndex(x, y, yaw) < 0;
  }

  // batched form: many pose lists (segments) in ONE launch; first_hit[s] = index in segment s or -1
  void checkSegments(const double* x, const double* y, const double* yaw, int n_poses, const int* seg_off, int n_seg,
                     int variant, int* first_hit)
  {
    ctx_.check(ga_check_poses(ctx_.handle(), x, y, yaw, n_poses, seg_off, n_seg, variant, first_hit));
  }

  // HybridAStar::setNeighbors' rollout + collision part for a batch of nodes (hybrid_a_star_lib.cpp:492-534)
  void setExpandParams(double arc_l, double yaw_res_deg, double turn_on_point_angle_deg, int nb_steer = 13)
  {
    ga_expand_params p{ veh_.max_steer_, veh_.w_b_, arc_l, yaw_res_deg, turn_on_point_angle_deg, nb_steer };
    ctx_.check(ga_set_expand_params(ctx_.handle(), &p));
  }
  int expandStride() const { return ga_expand_stride(ctx_.handle()); }
  void expandNodes(const std::vector<ga_node>& nodes, std::vector<ga_succ>& out)
  {
    out.assign(nodes.size() * (size_t)expandStride(), ga_succ{});
    ctx_.check(ga_expand(ctx_.handle(), nodes.data(), (int)nodes.size(), out.data()));
  }

  unsigned nb_disks() const { return nb_disks_; }
  const std::vector<int>& disk_centers() const { return disk_centers_; }
  size_t patch_dim() const { return patch_dim_; }

private:
  int firstHit(const double* x, const double* y, const double* yaw, int n, int variant)
  {
    const int seg[2] = { 0, n };
    int hit = -1;
    ctx_.check(ga_check_poses(ctx_.handle(), x, y, yaw, n, seg, 1, variant, &hit));
    return hit;
  }
  bool gridCheck(const Pose<int>& p, int variant)
  {
    if (p.x < 0 || p.y < 0) return variant != 0 ? gridCheckNegative(p) : false;
    const Pose<double> c((p.x + 0.5) * GM_RES, (p.y + 0.5) * GM_RES, (p.yaw + 0.5) * yaw_res_coll_rad_);
    return firstHit(&c.x, &c.y, &c.yaw, 1, variant) < 0;
  }
  bool gridCheckNegative(const Pose<int>& p)
  {
    // checkGridVariant with a negative cell: truncation toward zero cannot express it as metres,
    // so resolve the (at most nb_disks) lookups against a host copy of the grid
    fetchSafetyPatch();
    for (unsigned d = 0; d < nb_disks_; ++d)
    {
      const int px = disk_centers_[((size_t)p.yaw * nb_disks_ + d) * 2] + p.x;
      const int py = disk_centers_[((size_t)p.yaw * nb_disks_ + d) * 2 + 1] + p.y;
      if (px < 0 || py < 0 || px >= (int)patch_dim_ || py >= (int)patch_dim_) continue;
      if (patch_safety_arr_(py, px) == OCC) return false;
    }
    return true;
  }

  Context& ctx_;
  const Vehicle& veh_;
  size_t patch_dim_ = 0;
  int min_thresh_ = 100, max_thresh_ = 130, yaw_res_coll_ = 3, nb_disc_yaws_ = 120;
  double yaw_res_coll_rad_ = 3 * TO_RAD, safety_distance_m_ = 0.1, len_per_disk_ = 2.0;
  bool double_disk_rows_ = false;
  unsigned nb_disks_ = 0;
  std::vector<int> disk_centers_;
};

// GridMapSim (LIB/include/gridmap_sim_lib/gridmap_sim.hpp:92-112): simulated lidar ray-cast on an
// 801 x 801 ground-truth window; the result also stays on the device for Cartographing::cartographRaytraced
class GridMapSim
{
public:
  static constexpr size_t GM_DIM = 801;
  explicit GridMapSim(Context& ctx) : ctx_(ctx) {}
  void gmSimulation(const uint8_t* grid_map_gt, uint8_t* pmap_out) { ctx_.check(ga_raytrace(ctx_.handle(), grid_map_gt, pmap_out)); }

private:
  Context& ctx_;
};

// The grid reads of HybridAStar that score sampled poses (hybrid_a_star_lib.cpp:265-289, 588-594,
// 670-676): AStar::h_prox_arr_ stays on the device, whole batches of analytic-expansion candidates
// are scored per call.
class HybridAStar
{
public:
  HybridAStar(Context& ctx, const Vehicle& veh, double h_prox_cost = 5.0, double interp_res = 0.1) : ctx_(ctx)
  {
    const ga_prox_params p{ veh.lf_, veh.lb_, veh.width_, 1.0 / 0.625 /* grid_tf::con2star_ */, h_prox_cost, interp_res };
    ctx_.check(ga_set_prox_params(ctx_.handle(), &p));
  }
  // getProxOfCorners(pose * grid_tf::con2star_) for n poses, xyyaw = n x (x, y, yaw) in metres / rad
  void getProxOfCorners(const double* xyyaw, int n, double* out) { ctx_.check(ga_prox_of_corners(ctx_.handle(), xyyaw, n, out)); }
  double getProxOfCorners(const Pose<double>& pose)  // hybrid_a_star_lib.cpp:265, pose in metres
  {
    const double p[3] = { pose.x, pose.y, pose.yaw };
    double v = 0;
    getProxOfCorners(p, 1, &v);
    return v;
  }
  // step 3 of getValidClosePose (hybrid_a_star_lib.cpp:1604-1668): near_cells = AStar::calcIndex of the nodes of
  // nodes_near_goal_; the reference's default (ego_pose) when no candidate qualifies
  Pose<double> getValidClosePose(const Pose<double>& center_pose, const std::vector<int>& near_cells, const Pose<double>& ego_pose)
  {
    const double c[3] = { center_pose.x, center_pose.y, center_pose.yaw };
    double out[3];
    int found = 0;
    ctx_.check(ga_valid_close_pose(ctx_.handle(), c, near_cells.data(), (int)near_cells.size(), out, &found));
    return found ? Pose<double>(out[0], out[1], out[2]) : ego_pose;
  }
  // the prox_cost term of getRSPathCosts / getRAPathCosts for every path of a ragged batch
  std::vector<double> getPathProxCosts(const std::vector<double>& xyyaw, const std::vector<int>& path_offsets)
  {
    std::vector<double> out(path_offsets.empty() ? 0 : path_offsets.size() - 1);
    if (!out.empty()) ctx_.check(ga_path_prox_costs(ctx_.handle(), xyyaw.data(), path_offsets.data(), (int)out.size(), out.data()));
    return out;
  }

private:
  Context& ctx_;
};

// Cartographing (LIB/include/cartographing_lib/cartographing.hpp:19-41): the accumulated map of the
// unknown-environment loop, device-resident.  Same member names; the map itself is fetched on demand.
class Cartographing
{
public:
  explicit Cartographing(Context& ctx) : ctx_(ctx) {}

  void resetPatch(size_t patch_dim) { ctx_.check(ga_cart_reset(ctx_.handle(), (int)patch_dim)); }  // cartographing.cpp:7-15
  void cartograph(const Vec2DFlat<uint8_t>& local_map, const Point<int>& origin, int dim)          // :57-96
  {
    ctx_.check(ga_cartograph(ctx_.handle(), local_map.getPtr(), origin.x, origin.y, dim));
  }
  void cartographRaytraced(const Point<int>& origin) { ctx_.check(ga_cartograph_raytraced(ctx_.handle(), origin.x, origin.y)); }
  void passLocalMap(const Point<int>& origin, int dim)                                            // :98-117
  {
    ctx_.check(ga_cart_pass_local_map(ctx_.handle(), origin.x, origin.y, dim));
  }
  void loadPrevPatch(const Point<double>& prev_origin_utm, const Point<double>& origin_utm)         // :119-129
  {
    const double con2gm = 1.0 / 0.15625;  // grid_tf::con2gm_; Point<double>::toInt = std::round
    ctx_.check(ga_cart_load_prev(ctx_.handle(), (int)std::round(prev_origin_utm.x * con2gm), (int)std::round(prev_origin_utm.y * con2gm),
                                 (int)std::round(origin_utm.x * con2gm), (int)std::round(origin_utm.y * con2gm)));
  }
  void getMap(Vec2DFlat<uint8_t>& out) const { ctx_.download(GA_GRID_CART, out, (size_t)ga_cart_dim(ctx_.handle())); }

private:
  Context& ctx_;
};

// Read API of std::unordered_map<size_t, NodeDisc> (AStar::closed_set_guidance_ / closed_set_path_,
// a_star.hpp:79-80) over the dense result planes the device produces: find / end / at / contains /
// size / clear / iteration -- what hybrid_a_star_lib.cpp:243,784,794 and the GUI use.
class ClosedSet
{
public:
  struct value_type
  {
    size_t first;
    NodeDisc second;
  };
  class const_iterator
  {
  public:
    const_iterator(const ClosedSet* s, size_t i) : s_(s), i_(i) { load(); }
    const value_type& operator*() const { return v_; }
    const value_type* operator->() const { return &v_; }
    const_iterator& operator++()
    {
      i_ = s_->next(i_ + 1);
      load();
      return *this;
    }
    bool operator==(const const_iterator& o) const { return i_ == o.i_; }
    bool operator!=(const const_iterator& o) const { return i_ != o.i_; }

  private:
    void load()
    {
      if (i_ < s_->flags.size()) v_ = value_type{ i_, s_->node(i_) };
    }
    const ClosedSet* s_;
    size_t i_;
    value_type v_;
  };

  Vec2DFlat<double> cost, dist;
  Vec2DFlat<int32_t> parent;
  Vec2DFlat<uint8_t> flags;
  int dim = 0;

  bool contains(size_t id) const { return id < flags.size() && (flags.getPtr()[id] & GA_H_CLOSED); }
  const_iterator find(size_t id) const { return contains(id) ? const_iterator(this, id) : end(); }
  const_iterator begin() const { return const_iterator(this, next(0)); }
  const_iterator end() const { return const_iterator(this, flags.size()); }
  NodeDisc at(size_t id) const
  {
    if (!contains(id)) throw std::out_of_range("ClosedSet::at");
    return node(id);
  }
  size_t size() const { return n_closed; }
  bool empty() const { return n_closed == 0; }
  void clear()
  {
    n_closed = 0;
    if (flags.size()) std::memset(flags.getPtr(), 0, flags.size());
  }
  size_t n_closed = 0;

private:
  friend class const_iterator;
  NodeDisc node(size_t id) const
  {
    NodeDisc n;
    n.pos = Point<int>((int)(id % dim), (int)(id / dim));
    n.cost_ = cost.getPtr()[id];
    n.cost_dist_ = dist.getPtr()[id];
    n.parent_index_ = parent.getPtr()[id];
    n.is_unknown_ = (flags.getPtr()[id] & GA_H_UNKNOWN) != 0;
    return n;
  }
  size_t next(size_t i) const
  {
    const uint8_t* f = flags.getPtr();
    const size_t n = flags.size();
    while (i < n && !(f[i] & GA_H_CLOSED)) ++i;
    return i;
  }
};

// AStar (a_star.hpp:32-140)
class AStar
{
public:
  static constexpr size_t VOR_DIM = 200;  // a_star.hpp:36

  double alpha_ = 100.0, do_max_ = 7.0, do_min_ = 3.0;  // params.yml
  double astar_movement_cost_ = 5.0, astar_prox_cost_ = 5.0, astar_lane_movement_cost_ = 3.0;
  int astar_dim_ = 0;
  int dt_variant_ = GA_DT_IPPFLOAT;  // the reference's Docker OpenCV is built WITH_IPP
  ClosedSet closed_set_path_, closed_set_guidance_;
  Vec2DFlat<uint8_t> astar_grid_, restr_geofence_grid_;
  Vec2DFlat<double> movement_cost_map_, h_prox_arr_, motion_res_map_;
  Vec2DFlat<float> obstacle_dist_, voronoi_dist_;

  explicit AStar(Context& ctx) : ctx_(ctx) {}

  // a_star.cpp:11-53 (config) -- weights come from params.yml as the Python side pushes them
  void initialize(int patch_dim, const Point<double>& patch_origin_utm, const std::string& path2config)
  {
    Config cfg;
    cfg.load(path2config);
    configure(cfg);
    patch_origin_utm_ = patch_origin_utm;
    patch_dim_ = patch_dim;
    astar_dim_ = (int)std::floor((double)patch_dim * GM_RES / ASTAR_RES);
  }
  void configure(const Config& cfg)
  {
    unknown_cost_w_ = cfg.num("ASTAR_UNKNOWN_COST", 0);
    motion_res_min_ = cfg.num("MOTION_RES_MIN", 0.1);
    motion_res_max_ = cfg.num("MOTION_RES_MAX", 1.0);
    dist_val_min_ = cfg.num("MIN_DIST_VAL", 3);
    dist_val_max_ = cfg.num("MAX_DIST_VAL", 10);
    heuristic_early_exit_ = cfg.flag("HEURISTIC_EARLY_EXIT", true);
    max_extra_nodes_ = (unsigned)cfg.num("MAX_EXTRA_NODES_ASTAR", 10000);
    pushParams();
  }
  void pushParams()
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
    p.dt_variant = dt_variant_;
    ctx_.check(ga_set_astar_params(ctx_.handle(), &p));
    ctx_.astar_params_set = true;
  }
  // a_star.cpp:70-99.  Call BEFORE CollisionChecker::resetPatch, as HybridAStar::reinit does (hybrid_a_star_lib.cpp:
  // 100-101): the potential field and the motion-resolution map are saved here and carried over the patch move by
  // finishReinit() once ga_patch_reset has resized the device grids (util::saveTemp + copyPatch2Patch, util2.hpp:20-67).
  void reinit(const Point<double>& patch_origin_utm, int patch_dim)
  {
    if (ga_patch_dim(ctx_.handle()) > 0)
    {
      const size_t d = (size_t)ga_astar_dim(ctx_.handle());
      ctx_.download(GA_GRID_H_PROX, temp_h_prox_arr_, d);
      ctx_.download(GA_GRID_MOTION_RES, temp_motion_res_map_, d);
      have_temp_ = true;
    }
    prev_origin_astar_ = patch_origin_astar_;
    patch_origin_utm_ = patch_origin_utm;
    patch_origin_astar_ = Point<int>((int)std::round(patch_origin_utm.x * CON2STAR), (int)std::round(patch_origin_utm.y * CON2STAR));
    patch_dim_ = patch_dim;
    astar_dim_ = (int)std::floor((double)patch_dim * GM_RES / ASTAR_RES);
  }
  // after CollisionChecker::resetPatch: copy the overlap of the previous maps into the new ones (copyPatch2Patch,
  // including its exclusive end point) and upload them
  void finishReinit()
  {
    astar_dim_ = ga_astar_dim(ctx_.handle());
    if (!have_temp_) return;
    const size_t nd = (size_t)astar_dim_;
    h_prox_arr_.resize_and_reset(nd, nd, 0.0);
    motion_res_map_.resize_and_reset(nd, nd, motion_res_max_);
    const int prev_dim = (int)temp_h_prox_arr_.xDim(), next_dim = astar_dim_;
    const Point<int> po = prev_origin_astar_, no = patch_origin_astar_;
    const Point<int> start_prev(std::max(no.x - po.x, 0), std::max(no.y - po.y, 0));
    const Point<int> start_next(std::max(po.x - no.x, 0), std::max(po.y - no.y, 0));
    const Point<int> end_next(std::min(po.x + prev_dim - no.x, next_dim - 1), std::min(po.y + prev_dim - no.y, next_dim - 1));
    for (int x = 0; x < end_next.x - start_next.x; ++x)
      for (int y = 0; y < end_next.y - start_next.y; ++y)
      {
        h_prox_arr_((size_t)(start_next.y + y), (size_t)(start_next.x + x)) =
            temp_h_prox_arr_((size_t)(start_prev.y + y), (size_t)(start_prev.x + x));
        motion_res_map_((size_t)(start_next.y + y), (size_t)(start_next.x + x)) =
            temp_motion_res_map_((size_t)(start_prev.y + y), (size_t)(start_prev.x + x));
      }
    ctx_.upload(GA_GRID_H_PROX, h_prox_arr_);
    ctx_.upload(GA_GRID_MOTION_RES, motion_res_map_);
    have_temp_ = false;
  }

  size_t calcIndex(size_t x, size_t y) const { return y * (size_t)astar_dim_ + x; }  // a_star.cpp:303-316

  // a_star.cpp:536-542: Pooling::execute is fused into processSafetyPatch's kernel
  void calcAstarGridCuda() {}

  // a_star.cpp:110-295.  reference_order = true (default): the reference's own closing order with its start cell,
  // early exit (config HEURISTIC_EARLY_EXIT / MAX_EXTRA_NODES_ASTAR) and get_only_near -- planes and
  // nodes_near_goal_ bit-identical to the reference's (GA_HEUR_REFORDER).  false: the GPU fixed point, every cell
  // reachable from the goal closed with its optimal cost; the start cell only matters for get_only_near then.
  bool reference_order_ = true;
  bool heuristic_early_exit_ = true;
  unsigned max_extra_nodes_ = 10000;
  std::vector<std::pair<double, NodeDisc>> nodes_near_goal_;
  void calcDistanceHeuristic(const Point<int>& goal_pos, const Point<int>& start_pos, bool for_path = true,
                             bool get_only_near = false)
  {
    const int slot = for_path ? 1 : 0;
    int n = 0;
    ga_heur_query q{};
    q.goal_x = goal_pos.x;
    q.goal_y = goal_pos.y;
    q.start_x = start_pos.x;
    q.start_y = start_pos.y;
    q.max_extra_nodes = max_extra_nodes_;
    q.flags = reference_order_ ? (GA_HEUR_REFORDER | (heuristic_early_exit_ ? GA_HEUR_EARLY_EXIT : 0) |
                                  (get_only_near ? GA_HEUR_ONLY_NEAR : 0))
                               : GA_HEUR_FIXEDPOINT;
    ctx_.check(ga_build_heuristic_ex(ctx_.handle(), slot, &q, &n));
    ClosedSet& cs = for_path ? closed_set_path_ : closed_set_guidance_;
    astar_dim_ = ga_astar_dim(ctx_.handle());
    const size_t d = (size_t)astar_dim_;
    cs.dim = astar_dim_;
    ctx_.download(GA_GRID_HEUR_COST + slot, cs.cost, d);
    ctx_.download(GA_GRID_HEUR_DIST + slot, cs.dist, d);
    ctx_.download(GA_GRID_HEUR_PARENT + slot, cs.parent, d);
    ctx_.download(GA_GRID_HEUR_FLAGS + slot, cs.flags, d);
    cs.n_closed = (size_t)n;
    nodes_near_goal_.clear();  // a_star.cpp:141
    if (get_only_near)
    {
      int ids[400], cnt = 0;
      double dist[400];
      ctx_.check(ga_heur_near_nodes(ctx_.handle(), slot, start_pos.x, start_pos.y, ids, dist, 400, &cnt));
      for (int i = 0; i < cnt && i < 400; ++i) nodes_near_goal_.emplace_back(dist[i], cs.at((size_t)ids[i]));
    }
  }

  // a_star.cpp:351-378
  int findValidNeighborIndex(int start_idx, const ClosedSet& heuristic) const
  {
    const int cand[4] = { start_idx + 1, start_idx - 1, start_idx + astar_dim_, start_idx - astar_dim_ };
    for (int c : cand)
      if (c >= 0 && heuristic.contains((size_t)c)) return c;
    return -1;
  }
  // a_star.cpp:386-421
  std::pair<std::vector<int>, std::vector<int>> getAstarPath(int x_index, int y_index) const
  {
    std::vector<int> xs, ys;
    const ClosedSet& cs = closed_set_guidance_;
    size_t idx = calcIndex((size_t)x_index, (size_t)y_index);
    if (!cs.contains(idx))
    {
      const int i = findValidNeighborIndex((int)idx, cs);
      if (i == -1) return { xs, ys };
      idx = (size_t)i;
    }
    int cur = (int)idx;
    while (cur != -1)
    {
      xs.push_back(cur % astar_dim_);
      ys.push_back(cur / astar_dim_);
      cur = cs.parent.getPtr()[cur];
    }
    return { xs, ys };
  }

  // a_star.cpp:440-479.  The Fortune sweep (deps_lib/VoronoiDiagramGenerator.cpp) is serial fp64
  // geometry and runs on the host: `sweep` receives the OCC cells of the window in cv::findNonZero
  // order and returns the edge end-point cells it marks (generateVoronoi_wrapper :59-86).
  template <typename Sweep>
  bool calcVoronoiPotentialField(const Point<int>& ego_index, Sweep&& sweep)
  {
    std::vector<int> sites(2 * VOR_DIM * VOR_DIM);
    int n_sites = 0;
    ctx_.check(ga_potential_sites(ctx_.handle(), ego_index.x, ego_index.y, sites.data(), (int)(VOR_DIM * VOR_DIM), &n_sites));
    sites.resize(2 * (size_t)n_sites);
    std::vector<int> vor;
    if (n_sites > 0) vor = sweep(sites);
    int had = 0;
    ctx_.check(ga_build_potential(ctx_.handle(), ego_index.x, ego_index.y, vor.empty() ? nullptr : vor.data(),
                                  (int)(vor.size() / 2), &had));
    return had != 0;
  }

  void fetchGrids()
  {
    astar_dim_ = ga_astar_dim(ctx_.handle());
    const size_t d = (size_t)astar_dim_;
    ctx_.download(GA_GRID_ASTAR, astar_grid_, d);
    ctx_.download(GA_GRID_GEOFENCE, restr_geofence_grid_, d);
    ctx_.download(GA_GRID_MOVEMENT_COST, movement_cost_map_, d);
    ctx_.download(GA_GRID_H_PROX, h_prox_arr_, d);
    ctx_.download(GA_GRID_MOTION_RES, motion_res_map_, d);
    ctx_.download(GA_GRID_OBSTACLE_DIST, obstacle_dist_, VOR_DIM);
    ctx_.download(GA_GRID_VORONOI_DIST, voronoi_dist_, VOR_DIM);
  }
  void resetMovementMap()  // a_star.cpp:544-547
  {
    movement_cost_map_.resize_and_reset((size_t)astar_dim_, (size_t)astar_dim_, astar_movement_cost_);
    ctx_.upload(GA_GRID_MOVEMENT_COST, movement_cost_map_);
  }
  void uploadMovementMap() { ctx_.upload(GA_GRID_MOVEMENT_COST, movement_cost_map_); }
  void uploadGeofence() { ctx_.upload(GA_GRID_GEOFENCE, restr_geofence_grid_); }

private:
  Context& ctx_;
  Point<double> patch_origin_utm_;
  Point<int> patch_origin_astar_, prev_origin_astar_;
  Vec2DFlat<double> temp_h_prox_arr_, temp_motion_res_map_;
  bool have_temp_ = false;
  int patch_dim_ = 0;
  double unknown_cost_w_ = 0, motion_res_min_ = 0.1, motion_res_max_ = 1.0, dist_val_min_ = 3, dist_val_max_ = 10;
};
}  // namespace gridaccel
