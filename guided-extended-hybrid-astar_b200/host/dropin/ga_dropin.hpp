// ga_dropin.hpp -- the reference's all-static classes AStar / CollisionChecker / Vehicle, same names, same
// static member functions and public static data, with bodies that call libgridaccel.so (include/gridaccel.h).
//
// This is the drop-in boundary SURVEY.md 8(b) describes: hybrid_a_star_lib.cpp, cartographing.cpp and the
// pybind wrapper compile UNCHANGED against it.  How a reference maintainer uses it: add
//     -include ga_dropin.hpp      (before every TU that includes the three headers below)
// to the compile flags, drop a_star.cpp / collision_checking.cpp / vehicle.cpp / max_pool.cu from the build,
// add ga_dropin.cpp and link libgridaccel.so.  The header defines the include guards of
//     LIB/include/hybridastar_planning_lib/a_star.hpp            (FREESPACE_PLANNER_A_STAR_HPP)
//     LIB/include/collision_checker_lib/collision_checking.hpp   (COLLISION_CHECKING_HPP)
//     LIB/include/collision_checker_lib/vehicle.hpp              (PLANNING_BINDINGS_VEHICLE_HPP)
// so the reference's own definitions of these classes are never parsed; everything else of the reference
// (util_lib containers and transforms, the Fortune sweep, Reeds-Shepp, the search) is used as it is.
//
// State: the reference keeps one planner per process in inline statics; so does this header, on top of ONE
// ga_ctx (ga_dropin::ctx()).  There is no CPU path: without a CUDA device the first call throws.
#ifndef GA_DROPIN_HPP
#define GA_DROPIN_HPP

#define FREESPACE_PLANNER_A_STAR_HPP
#define COLLISION_CHECKING_HPP
#define PLANNING_BINDINGS_VEHICLE_HPP

#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <exception>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

// what the replaced headers pulled in for their includers (hybrid_a_star_lib.cpp reads its config with yaml-cpp,
// VoronoiDiagramGenerator.hpp uses cv::Point)
#include "opencv2/imgproc.hpp"
#include "yaml-cpp/yaml.h"

#include "util_lib/data_structures1.hpp"
#include "util_lib/data_structures2.hpp"
#include "util_lib/transforms.hpp"
#include "util_lib/util1.hpp"
#include "util_lib/util2.hpp"

#include "deps_lib/VoronoiDiagramGenerator.hpp"

#include "gridaccel.h"

namespace py = pybind11;

// ---- closed sets ------------------------------------------------------------------------------------
// AStar::closed_set_path_ / closed_set_guidance_ are std::unordered_map<size_t, NodeDisc> in the reference and
// hybrid_a_star_lib.hpp names that type in its signatures (calcCost, getDistance2goal, check4Expansions).  The
// device produces dense planes; building a hash map from them would cost more than the heuristic itself.  NodeDisc
// is a program-defined type, so the standard permits specialising std::unordered_map for it: this explicit
// specialisation keeps the read API the planner and the GUI use (find / end / at / contains / count / size / empty /
// begin / iteration) over the dense planes.  Mutation other than clear() is not offered: only calcDistanceHeuristic
// writes these sets.
namespace ga_dropin
{
struct ClosedPlanes
{
  // either the four planes (after a reference-order replay: GA_GRID_HEUR_COST / _DIST / _PARENT / _FLAGS) ...
  std::vector<double> cost, dist;
  std::vector<int32_t> parent;
  std::vector<uint8_t> flags;
  // ... or the 12-byte records of a device build (GA_GRID_HEUR_REC: 12 B per cell over PCIe instead of 21), read in place
  std::vector<uint32_t> rec;
  bool packed = false;
  int dim = 0;
  size_t cells = 0, n_closed = 0;

  bool closed(size_t i) const
  {
    return packed ? (rec[3 * i + 1] & 0x7FF00000u) != 0x7FF00000u  // cost_ finite
                  : (flags[i] & GA_H_CLOSED) != 0;
  }
  bool unknown(size_t i) const { return packed ? ((rec[3 * i + 2] >> 3) & 1u) != 0 : (flags[i] & GA_H_UNKNOWN) != 0; }
  double cost_of(size_t i) const
  {
    if (!packed) return cost[i];
    double g;
    std::memcpy(&g, &rec[3 * i], sizeof g);
    return g;
  }
  // cost_dist_ = straight * ASTAR_RES + diagonal * (ASTAR_RES * sqrt 2): two products and one sum, as the device forms it
  __attribute__((optimize("fp-contract=off"))) double dist_of(size_t i) const
  {
    if (!packed) return dist[i];
    const uint32_t w = rec[3 * i + 2];
    const double a = (double)((w >> 4) & 16383u) * 0.625, b = (double)((w >> 18) & 16383u) * (0.625 * std::sqrt(2.0));
    return a + b;
  }
  int32_t parent_of(size_t i) const
  {
    if (!packed) return parent[i];
    const uint32_t w = rec[3 * i + 2];
    if ((w >> 4) == 0) return -1;  // the root: the closed cell with no step
    // motion_ (a_star.hpp:50-52): parent = cell - motion_[dir]
    static constexpr int mdx[8] = { 1, 0, -1, 0, -1, -1, 1, 1 }, mdy[8] = { 0, 1, 0, -1, -1, 1, -1, 1 };
    const int d = (int)(w & 7u);
    return (int32_t)(((int)(i / (size_t)dim) - mdy[d]) * dim + ((int)(i % (size_t)dim) - mdx[d]));
  }
};
}  // namespace ga_dropin

template <>
class std::unordered_map<size_t, NodeDisc, std::hash<size_t>, std::equal_to<size_t>,
                         std::allocator<std::pair<const size_t, NodeDisc>>>
{
public:
  using key_type = size_t;
  using mapped_type = NodeDisc;
  using value_type = std::pair<const size_t, NodeDisc>;
  using size_type = size_t;

  class const_iterator
  {
  public:
    using iterator_category = std::forward_iterator_tag;
    using value_type = std::pair<const size_t, NodeDisc>;
    using difference_type = std::ptrdiff_t;
    using pointer = const value_type*;
    using reference = const value_type&;
    const_iterator() = default;
    const_iterator(const ga_dropin::ClosedPlanes* p, size_t i) : p_(p), i_(i) {}
    reference operator*() const
    {
      load();
      return *reinterpret_cast<const value_type*>(buf_);
    }
    pointer operator->() const { return &**this; }
    const_iterator& operator++()
    {
      ++i_;
      skip();
      return *this;
    }
    const_iterator operator++(int)
    {
      const_iterator t = *this;
      ++*this;
      return t;
    }
    bool operator==(const const_iterator& o) const { return i_ == o.i_; }
    bool operator!=(const const_iterator& o) const { return i_ != o.i_; }
    void skip()
    {
      const size_t n = p_->cells;
      while (i_ < n && !p_->closed(i_)) ++i_;
    }

  private:
    void load() const
    {
      const int d = p_->dim;
      new (buf_) value_type(i_, NodeDisc(Point<int>((int)(i_ % (size_t)d), (int)(i_ / (size_t)d)), p_->cost_of(i_), p_->dist_of(i_),
                                         p_->parent_of(i_), p_->unknown(i_)));
    }
    const ga_dropin::ClosedPlanes* p_ = nullptr;
    size_t i_ = 0;
    alignas(value_type) mutable unsigned char buf_[sizeof(value_type)];
  };
  using iterator = const_iterator;

  bool contains(size_t id) const { return id < planes.cells && planes.closed(id); }
  size_t count(size_t id) const { return contains(id) ? 1 : 0; }
  const_iterator find(size_t id) const { return contains(id) ? const_iterator(&planes, id) : end(); }
  const_iterator begin() const
  {
    const_iterator it(&planes, 0);
    it.skip();
    return it;
  }
  const_iterator end() const { return const_iterator(&planes, planes.cells); }
  const_iterator cbegin() const { return begin(); }
  const_iterator cend() const { return end(); }
  NodeDisc at(size_t id) const
  {
    if (!contains(id)) throw std::out_of_range("closed set: cell not in the heuristic");
    return (*const_iterator(&planes, id)).second;
  }
  size_t size() const { return planes.n_closed; }
  bool empty() const { return planes.n_closed == 0; }
  void clear()
  {
    planes.n_closed = 0;
    planes.packed = false;
    planes.flags.assign(planes.cells, (uint8_t)0);
  }

  ga_dropin::ClosedPlanes planes;  // filled by AStar::calcDistanceHeuristic
};

// ---- context ----------------------------------------------------------------------------------------
namespace ga_dropin
{
// the one planner of this process (the reference's inline statics, as one ga_ctx).  Device = $GA_DEVICE or 0.
ga_ctx* ctx();
void check(int rc);  // throws std::runtime_error(ga_last_error) on failure
// GA_HEUR_REFORDER (default: the reference's closing order, identical search sequences) or GA_HEUR_FIXEDPOINT
// (the GPU fixed point: every reachable cell closed, optimal costs)
void setHeuristicMode(unsigned ga_heur_flags);
unsigned heuristicMode();
// look-ahead fan checks (default on): see ga_dropin.cpp
void setLookahead(bool on);
// which arithmetic of cv::distanceTransform the potential field follows (GA_DT_IPPFLOAT: OpenCV built WITH_IPP, the
// reference's Docker image; GA_DT_FIXEDPOINT: plain OpenCV)
void setDistanceTransformVariant(int ga_dt_variant);

struct Stats
{
  unsigned long long expansions_prefetched = 0;  // nodes whose primitives were rolled out and checked in one launch
  unsigned long long path_checks_cached = 0;     // checkPathCollision calls answered from that launch
  unsigned long long path_checks_direct = 0;     // calls that needed their own launch (analytic expansions ...)
  unsigned long long poses_checked = 0;
  unsigned long long lookahead_launches = 0;  // asynchronous fan checks of the successors just created
  unsigned long long lookahead_hits = 0;      // expansions answered from them (no launch on the critical path)
  unsigned long long lookahead_rechecks = 0;
  unsigned long long ns_submit = 0, ns_collect = 0, ns_harvest = 0;  // host time in ga_expand_submit / waiting + copying in ga_expand_collect / filling the answer table
   // primitives re-checked because a pose lay on a cell / yaw-bin edge
};
Stats& stats();
}  // namespace ga_dropin

// ---- Vehicle (LIB/include/collision_checker_lib/vehicle.hpp:15-45) ------------------------------------------
class Vehicle
{
public:
  inline static Pose<double> pose_utm_;
  inline static bool is_ushift_;
  inline static double max_steer_;
  inline static double max_curvature_;
  inline static double w_b_;
  inline static double lf_;
  inline static double lb_;
  inline static double length_;
  inline static double width_;
  inline static double geo_center_;
  inline static Point<double> f_r_corner_;
  inline static Point<double> f_l_corner_;
  inline static Point<double> r_r_corner_;
  inline static Point<double> r_l_corner_;
  inline static std::vector<Point<double>> vehicle_vertices_;
  inline static std::vector<Point<double>> vis_vehicle_vertices_;

  static void setPose(const Pose<double>& pose);
  static void initialize(double max_steer, double w_b, double l_f, double l_b, double width, bool is_ushift);
  static void move(Pose<double>& pose, double distance, double steer);
  // Rolls out the primitive on the host (glibc, the reference's operand order: node poses must be bit-identical to
  // the reference's, they are the search state).  The FIRST call for a new parent pose rolls out the parent's whole
  // steering x direction fan and sends all poses through ga_check_poses in ONE launch; checkPathCollision then
  // answers the fan's primitives from that result (north_star: one kernel launch per expansion batch).
  static MotionPrimitive
  move_car_some_steps(const Pose<double>& pose, double arc_l, double motion_res, int direction, double steer);
  // same for the U-Shift rear-axis fan (+-90, +-180, +-270 degrees): first call checks all six in one launch
  static MotionPrimitive turn_on_rear_axis(const Pose<double>& pose, double delta_angle, double yaw_res_coll);
};

// ---- CollisionChecker (LIB/include/collision_checker_lib/collision_checking.hpp:36-120) ---------------------
class CollisionChecker
{
public:
  enum GRID_VAL
  {
    FREE,     // 0
    UNKNOWN,  // 1
    OCC,      // 2
    SENSOR_FREE = 255,
    SENSOR_UNKNOWN = 127,
    SENSOR_OCC = 0
  };

private:
  inline static size_t patch_dim_;
  inline static uint8_t min_thresh_;
  inline static uint8_t max_thresh_;
  inline static unsigned int yaw_res_coll_;
  inline static double yaw_res_coll_rad_;
  inline static unsigned int nb_disc_yaws_;
  inline static unsigned int nb_disks_;
  inline static bool double_disk_rows_;
  inline static double len_per_disk_;
  inline static int disk_diameter_c_;
  inline static double safety_distance_m_;
  inline static double search_dist_;
  inline static double search_dist_cells_;
  inline static Vec2DFlat<Point<int>> disk_centers_;
  inline static bool host_patch_stale_ = true, host_safety_stale_ = true;

public:
  inline static double max_patch_ins_dist_;
  inline static double disk_r_;
  inline static int disk_r_c_;

  // Host copies of the device-resident grids.  The search never reads them (its lookups run on the device);
  // GUI-style consumers call syncPatch() / syncSafetyPatch() first.
  inline static Vec2DFlat<uint8_t> patch_arr_;
  inline static Vec2DFlat<uint8_t> patch_safety_arr_;
  static void syncPatch();
  static void syncSafetyPatch();

  static void initialize(size_t patch_dim, const std::string& path2config);
  static void calculateDisks();
  static double getDiskRadius(double length, double width, unsigned int nb_disks);
  static std::vector<double> getDiskPositions(double radius, unsigned int nb_disks, double width, double lb);
  static std::vector<Point<double>> returnDiskPositions(double yaw);
  static void resetPatch(size_t patch_dim);
  static void processSafetyPatch();
  static void insertMinipatches(const std::map<std::pair<int, int>, Minipatch>& minipatches,
                                const Point<double>& ego_utm,
                                bool only_nearest,
                                bool only_new);
  static void passLocalMap(const py::array_t<uint8_t>& local_map, const Point<int>& origin, int dim);
  static void passLocalMap(const Vec2DFlat<uint8_t>& local_map, const Point<int>& origin, int dim);
  static void passLocalMapData(const unsigned char* local_map, const Point<int>& origin, int dim);
  static int getYawIdx(double yaw);
  static bool checkGrid(const Pose<int>& pose);
  static bool checkGridVariant(const Pose<int>& pose);
  static bool checkPose(const Pose<double>& pose);
  static bool checkPoseVariant(const Pose<double>& pose);
  static int getPathCollisionIndex(const std::vector<double>& x_list,
                                   const std::vector<double>& y_list,
                                   const std::vector<double>& yaw_list);
  static bool checkPathCollision(const std::vector<double>& x_list,
                                 const std::vector<double>& y_list,
                                 const std::vector<double>& yaw_list);
  static size_t patchDim() { return patch_dim_; }
};

// ---- AStar (LIB/include/hybridastar_planning_lib/a_star.hpp:32-140) -----------------------------------------
class AStar
{
private:
  inline static constexpr uint8_t NB_GRID_MOTIONS = 8;
  inline static constexpr size_t VOR_DIM = 200;

  inline static Point<double> patch_origin_utm_;
  inline static Point<int> patch_origin_astar_;
  inline static double unknown_cost_w_;
  inline static size_t patch_dim_;
  inline static bool heuristic_early_exit_;
  inline static unsigned int max_extra_nodes_;
  inline static double motion_res_min_;
  inline static double motion_res_max_;
  inline static double dist_val_min_;
  inline static double dist_val_max_;
  inline static constexpr std::array<Point<int>, NB_GRID_MOTIONS + 1> patch_coords_ = {
    { { 0, 0 }, { 1, 0 }, { 0, 1 }, { -1, 0 }, { 0, -1 }, { -1, -1 }, { -1, 1 }, { 1, -1 }, { 1, 1 } }
  };
  inline static bool params_pushed_ = false;
  inline static bool astar_grid_stale_ = true;

public:
  inline static double alpha_;
  inline static double do_max_;
  inline static double do_min_;

  inline static Vec2DFlat<float> obstacle_dist_;  // host copies, refreshed by syncDistanceGrids()
  inline static Vec2DFlat<float> voronoi_dist_;

  inline static double astar_movement_cost_;
  inline static double astar_prox_cost_;
  inline static double astar_lane_movement_cost_;

  inline static int astar_dim_;
  inline static std::string path2config_;

  inline static std::unordered_map<size_t, NodeDisc> closed_set_path_;
  inline static std::unordered_map<size_t, NodeDisc> closed_set_guidance_;

  inline static std::vector<std::pair<double, NodeDisc>> nodes_near_goal_;

  inline static Vec2DFlat<uint8_t> astar_grid_;  // host copy, refreshed by syncAstarGrid()
  inline static Vec2DFlat<double> movement_cost_map_;
  inline static Vec2DFlat<uint8_t> restr_geofence_grid_;
  inline static Polygon restr_geofence_;

  // host mirrors kept current: the serial search reads them cell by cell (hybrid_a_star_lib.cpp:276-279, 324, 1483)
  inline static Vec2DFlat<double> h_prox_arr_;
  inline static Vec2DFlat<double> temp_h_prox_arr_;
  inline static Vec2DFlat<double> motion_res_map_;
  inline static Vec2DFlat<double> temp_motion_res_map_;

  inline static std::vector<size_t> path_indices_;

  static void initialize(int patch_dim, const Point<double>& patch_origin_utm, const std::string& path2config);
  static void init_structs(int patch_dim);
  static void reinit(const Point<double>& patch_origin_utm, int patch_dim);
  static std::pair<std::vector<int>, std::vector<int>> getAstarPath(int x_ind, int y_ind);
  static void calcDistanceHeuristic(const Point<int>& goal_pos,
                                    const Point<int>& start_pos,
                                    bool for_path = true,
                                    bool get_only_near = false);
  static void calcVoronoiPotentialField(const Point<int>& ego_index);
  static Point<int> reverse2DIndex(int idx, int dim);
  static void calcAstarGridCuda();
  static size_t calcIndex(size_t x_ind, size_t y_ind);
  static size_t calcIndex(const NodeDisc& node);
  static void resetMovementMap();
  static void setMovementMap(const LaneGraph::edges_t& edges);
  static void processGeofence();
  static int findValidNeighborIndex(int start_idx, const std::unordered_map<size_t, NodeDisc>& heuristic);

  static void syncAstarGrid();
  static void syncDistanceGrids();
  static void pushParams();  // weights (alpha_, do_max_, astar_*_cost_ ...) are plain statics the caller may change

private:
  static Point<int> getCurrentMapOrigin(const Point<int>& ego_pos, size_t dim);
};

#endif  // GA_DROPIN_HPP
