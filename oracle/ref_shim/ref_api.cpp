// TEST INFRASTRUCTURE ONLY -- C entry points (gar_*) over the reference's OWN translation units
// (compiled unmodified from /root/reference by oracle/Makefile's `ref` target into
// oracle/_ref/libga_ref.so).  Used to pin the plain-C oracle against the real reference code and
// as the "reference" CPU baseline of bench.py.  Nothing here is shipped or linked by the product.
//
// What is NOT the reference's code in this library, and why:
//   - cv::* (OpenCV 4.8.0, absent): cv_shim.cpp, pinned against cv2 goldens;
//   - Pooling::init / execute (LIB/src/cuda_lib/max_pool.cu:3-79, cuDNN int8 max pooling):
//     restated below as the integer max over disjoint windows;
//   - fitpack_wrapper::BSpline1D (Fortran FITPACK, no gfortran): not referenced by these TUs.
// every std / third-party header first, so that the access override below touches only the
// reference's own class definitions
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

#include "opencv2/imgproc.hpp"
#include "yaml-cpp/yaml.h"
#define private public  // disk_centers_, nb_disks_ ... are private statics the checker must read
#include "hybridastar_planning_lib/a_star.hpp"
#include "cartographing_lib/cartographing.hpp"
#undef private

#include <cmath>
#include <cstdio>
#include <cstring>

// ---- Pooling on the CPU (max_pool.cu:21-79: window = stride = pooling_dim, no padding, MAX)
static int g_pool_dim = 4;
void Pooling::init(int pooling_dim) { g_pool_dim = pooling_dim; }
void Pooling::execute(const uint8_t* image, uint8_t* result, int in_dim, int out_dim)
{
  for (int Y = 0; Y < out_dim; ++Y)
    for (int X = 0; X < out_dim; ++X)
    {
      uint8_t m = 0;
      for (int j = 0; j < g_pool_dim; ++j)
        for (int i = 0; i < g_pool_dim; ++i)
        {
          const uint8_t v = image[(size_t)(Y * g_pool_dim + j) * in_dim + (size_t)X * g_pool_dim + i];
          if (v > m) m = v;
        }
      result[(size_t)Y * out_dim + X] = m;
    }
}

static char g_err[512];
#define GAR_TRY(body)                                   \
  try                                                   \
  {                                                     \
    body;                                               \
    return 0;                                           \
  }                                                     \
  catch (const std::exception& e)                       \
  {                                                     \
    std::snprintf(g_err, sizeof g_err, "%s", e.what()); \
    return -1;                                          \
  }

extern "C" {
const char* gar_last_error() { return g_err; }

// HybridAStar::initialize's share for this path (hybrid_a_star_lib.cpp:12-88) + create_patch's
// reinit (planning.py:315-337): vehicle, weights, AStar, CollisionChecker, then reinit.
int gar_init(const char* config_yml, int patch_dim, double max_steer, double wb, double lf, double lb, double width,
             int is_ushift, double astar_movement_cost, double astar_prox_cost, double astar_lane_movement_cost,
             double alpha, double do_max, double do_min)
{
  GAR_TRY({
    Vehicle::initialize(max_steer, wb, lf, lb, width, is_ushift != 0);
    AStar::astar_movement_cost_ = astar_movement_cost;
    AStar::astar_prox_cost_ = astar_prox_cost;
    AStar::astar_lane_movement_cost_ = astar_lane_movement_cost;
    AStar::alpha_ = alpha;
    AStar::do_max_ = do_max;
    AStar::do_min_ = do_min;
    AStar::restr_geofence_.vertices.clear();
    AStar::initialize(patch_dim, Point<double>(0.0, 0.0), config_yml);
    CollisionChecker::initialize((size_t)patch_dim, config_yml);
    AStar::reinit(Point<double>(0.0, 0.0), patch_dim);
  })
}

int gar_reinit(int patch_dim)
{
  GAR_TRY({
    CollisionChecker::resetPatch((size_t)patch_dim);
    AStar::reinit(Point<double>(0.0, 0.0), patch_dim);
  })
}

int gar_patch_dim() { return (int)CollisionChecker::patch_dim_; }
int gar_astar_dim() { return AStar::astar_dim_; }
int gar_nb_disks() { return (int)CollisionChecker::nb_disks_; }
int gar_nb_yaws() { return (int)CollisionChecker::nb_disc_yaws_; }
int gar_disk_r_c() { return CollisionChecker::disk_r_c_; }
double gar_disk_r() { return CollisionChecker::disk_r_; }

// disk_centers_ LUT as [(yaw * nb_disks + disk) * 2 + {x, y}]
void gar_disk_centers(int* out)
{
  const int ny = (int)CollisionChecker::nb_disc_yaws_, nd = (int)CollisionChecker::nb_disks_;
  for (int y = 0; y < ny; ++y)
    for (int d = 0; d < nd; ++d)
    {
      const Point<int> p = CollisionChecker::disk_centers_(y, d);
      out[(y * nd + d) * 2] = p.x;
      out[(y * nd + d) * 2 + 1] = p.y;
    }
}

void gar_dil_kernel(uint8_t* out)  // (2r+1)^2
{
  const cv::Mat& k = CollisionChecker::dil_kernel_;
  for (int y = 0; y < k.rows; ++y) std::memcpy(out + (size_t)y * k.cols, k.ptr<uint8_t>(y), (size_t)k.cols);
}

int gar_pass_local_map(const uint8_t* win, int ox, int oy, int dim)
{
  GAR_TRY(CollisionChecker::passLocalMapData(win, Point<int>(ox, oy), dim))
}
int gar_process_safety_patch() { GAR_TRY(CollisionChecker::processSafetyPatch()) }
int gar_calc_astar_grid() { GAR_TRY(AStar::calcAstarGridCuda()) }

// grid ids follow include/gridaccel.h ga_grid_id (0..8)
static void* grid_ptr(int id, size_t* bytes)
{
  const size_t pd = CollisionChecker::patch_dim_, ad = (size_t)AStar::astar_dim_, vd = AStar::VOR_DIM;
  switch (id)
  {
    case 0: *bytes = pd * pd; return CollisionChecker::patch_arr_.getPtr();
    case 1: *bytes = pd * pd; return CollisionChecker::patch_safety_arr_.getPtr();
    case 2: *bytes = ad * ad; return AStar::astar_grid_.getPtr();
    case 3: *bytes = ad * ad; return AStar::restr_geofence_grid_.getPtr();
    case 4: *bytes = ad * ad * 8; return AStar::movement_cost_map_.getPtr();
    case 5: *bytes = ad * ad * 8; return AStar::h_prox_arr_.getPtr();
    case 6: *bytes = ad * ad * 8; return AStar::motion_res_map_.getPtr();
    case 7: *bytes = vd * vd * 4; return AStar::obstacle_dist_.getPtr();
    case 8: *bytes = vd * vd * 4; return AStar::voronoi_dist_.getPtr();
    default: *bytes = 0; return nullptr;
  }
}
int gar_get_grid(int id, void* dst, size_t bytes)
{
  size_t n = 0;
  void* p = grid_ptr(id, &n);
  if (!p || n != bytes) return -1;
  std::memcpy(dst, p, n);
  return 0;
}
int gar_set_grid(int id, const void* src, size_t bytes)
{
  size_t n = 0;
  void* p = grid_ptr(id, &n);
  if (!p || n != bytes) return -1;
  std::memcpy(p, src, n);
  return 0;
}

int gar_set_heuristic_config(int early_exit, unsigned max_extra_nodes, double unknown_cost_w)
{
  AStar::heuristic_early_exit_ = early_exit != 0;
  AStar::max_extra_nodes_ = max_extra_nodes;
  AStar::unknown_cost_w_ = unknown_cost_w;
  return 0;
}

// AStar::calcDistanceHeuristic (a_star.cpp:110-295); returns the closed-set size
int gar_calc_distance_heuristic(int gx, int gy, int sx, int sy, int for_path, int get_only_near)
{
  try
  {
    AStar::calcDistanceHeuristic(Point<int>(gx, gy), Point<int>(sx, sy), for_path != 0, get_only_near != 0);
    return (int)(for_path ? AStar::closed_set_path_ : AStar::closed_set_guidance_).size();
  }
  catch (const std::exception& e)
  {
    std::snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// dense dump of a closed set: cost (+inf = absent), cost_dist, parent (-2 absent), flags
// (bit0 closed, bit1 is_unknown_) -- the planes GA_GRID_HEUR_* carry
int gar_dump_closed_set(int for_path, double* cost, double* dist, int32_t* parent, uint8_t* flags)
{
  const auto& cs = for_path ? AStar::closed_set_path_ : AStar::closed_set_guidance_;
  const size_t n = (size_t)AStar::astar_dim_ * AStar::astar_dim_;
  for (size_t i = 0; i < n; ++i)
  {
    cost[i] = INFINITY;
    dist[i] = INFINITY;
    parent[i] = -2;
    flags[i] = 0;
  }
  for (const auto& [id, nd] : cs)
  {
    if (id >= n || (size_t)nd.pos.y * AStar::astar_dim_ + nd.pos.x != id) return -1;
    cost[id] = nd.cost_;
    dist[id] = nd.cost_dist_;
    parent[id] = nd.parent_index_;
    flags[id] = 1 | (nd.is_unknown_ ? 2 : 0);
  }
  return (int)cs.size();
}

int gar_nodes_near_goal(double* dist, int* xy, int cap)
{
  int n = 0;
  for (const auto& [d, nd] : AStar::nodes_near_goal_)
  {
    if (n >= cap) break;
    dist[n] = d;
    xy[2 * n] = nd.pos.x;
    xy[2 * n + 1] = nd.pos.y;
    ++n;
  }
  return (int)AStar::nodes_near_goal_.size();
}

int gar_astar_path(int x, int y, int* xs, int* ys, int cap)
{
  try
  {
    const auto [xl, yl] = AStar::getAstarPath(x, y);
    const int n = (int)xl.size();
    for (int i = 0; i < n && i < cap; ++i)
    {
      xs[i] = xl[i];
      ys[i] = yl[i];
    }
    return n;
  }
  catch (const std::exception& e)
  {
    std::snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

// AStar::calcVoronoiPotentialField (a_star.cpp:440-479) incl. the reference's Fortune sweep
int gar_calc_voronoi_potential_field(int ego_x, int ego_y, int dt_variant)
{
  cv::shim_dt_variant = dt_variant;
  GAR_TRY(AStar::calcVoronoiPotentialField(Point<int>(ego_x, ego_y)))
}

// vdg::VoronoiDiagramGenerator::generateVoronoi_wrapper (VoronoiDiagramGenerator.cpp:59-86) on a
// zeroed scratch window: the cells the sweep marks, row-major, as (x, y) pairs.  sites = OCC cells
// of the window in cv::findNonZero order (a_star.cpp:457-459).
int gar_voronoi_endpoints(const int* sites_xy, int n_sites, int vor_dim, double min_dist, int* out_xy, int cap)
{
  try
  {
    std::vector<cv::Point> pts;
    pts.reserve(n_sites);
    for (int i = 0; i < n_sites; ++i) pts.emplace_back(sites_xy[2 * i], sites_xy[2 * i + 1]);
    Vec2DFlat<float> scratch;
    scratch.resize_and_reset(vor_dim, vor_dim, 0.0f);
    if (n_sites > 0)
      vdg::VoronoiDiagramGenerator::generateVoronoi_wrapper(pts, scratch, vor_dim - 1, vor_dim - 1, min_dist,
                                                            CollisionChecker::OCC);
    int n = 0;
    for (int y = 0; y < vor_dim; ++y)
      for (int x = 0; x < vor_dim; ++x)
        if (scratch(y, x) == 2.0f)
        {
          if (n < cap)
          {
            out_xy[2 * n] = x;
            out_xy[2 * n + 1] = y;
          }
          ++n;
        }
    return n;
  }
  catch (const std::exception& e)
  {
    std::snprintf(g_err, sizeof g_err, "%s", e.what());
    return -1;
  }
}

int gar_yaw_idx(double yaw) { return CollisionChecker::getYawIdx(yaw); }
int gar_check_grid(int x, int y, int yaw_idx, int variant)
{
  const Pose<int> p(x, y, yaw_idx);
  return variant ? CollisionChecker::checkGridVariant(p) : CollisionChecker::checkGrid(p);
}
int gar_check_pose(double x, double y, double yaw, int variant)
{
  const Pose<double> p(x, y, yaw);
  return variant ? CollisionChecker::checkPoseVariant(p) : CollisionChecker::checkPose(p);
}
// segments as in ga_check_poses: first colliding index inside each segment, -1 if none
void gar_check_segments(const double* x, const double* y, const double* yaw, const int* seg_off, int n_seg, int variant,
                        int* first_hit)
{
  for (int s = 0; s < n_seg; ++s)
  {
    int hit = -1;
    if (!variant)
    {
      const std::vector<double> xl(x + seg_off[s], x + seg_off[s + 1]), yl(y + seg_off[s], y + seg_off[s + 1]),
          wl(yaw + seg_off[s], yaw + seg_off[s + 1]);
      hit = CollisionChecker::getPathCollisionIndex(xl, yl, wl);
    }
    else
    {
      for (int i = seg_off[s]; i < seg_off[s + 1]; ++i)
        if (!CollisionChecker::checkPoseVariant(Pose<double>(x[i], y[i], yaw[i])))
        {
          hit = i - seg_off[s];
          break;
        }
    }
    first_hit[s] = hit;
  }
}

// Vehicle::move_car_some_steps (vehicle.cpp:100-122); returns the number of poses
int gar_move_car_some_steps(double x, double y, double yaw, double arc_l, double motion_res, int direction, double steer,
                            double* xs, double* ys, double* yaws, int cap)
{
  const MotionPrimitive mp = Vehicle::move_car_some_steps(Pose<double>(x, y, yaw), arc_l, motion_res, direction, steer);
  const int n = (int)mp.x_list_.size();
  for (int i = 0; i < n && i < cap; ++i)
  {
    xs[i] = mp.x_list_[i];
    ys[i] = mp.y_list_[i];
    yaws[i] = mp.yaw_list_[i];
  }
  return n;
}

// Vehicle::turn_on_rear_axis (vehicle.cpp:45-98); returns the number of yaws
int gar_turn_on_rear_axis(double x, double y, double yaw, double delta_angle, double yaw_res_deg, double* yaws, int cap)
{
  const MotionPrimitive mp = Vehicle::turn_on_rear_axis(Pose<double>(x, y, yaw), delta_angle, yaw_res_deg);
  const int n = (int)mp.yaw_list_.size();
  for (int i = 0; i < n && i < cap; ++i) yaws[i] = mp.yaw_list_[i];
  return n;
}

// ---- Cartographing (LIB/src/cartographing_lib/cartographing.cpp): the accumulated map of the
//      unknown-environment loop.  Its patch is a private static: read through the access override.
int gar_cart_reset(int patch_dim) { GAR_TRY(Cartographing::resetPatch((size_t)patch_dim)) }
int gar_cartograph(const uint8_t* win, int ox, int oy, int dim)
{
  GAR_TRY({
    Vec2DFlat<uint8_t> w;
    w.resize_and_reset(dim, dim, 0);
    std::memcpy(w.getPtr(), win, (size_t)dim * dim);
    Cartographing::cartograph(w, Point<int>(ox, oy), dim);
  })
}
int gar_cart_pass_local_map(int ox, int oy, int dim) { GAR_TRY(Cartographing::passLocalMap(Point<int>(ox, oy), dim)) }
int gar_cart_load_prev(double prev_x, double prev_y, double x, double y)
{
  GAR_TRY(Cartographing::loadPrevPatch(Point<double>(prev_x, prev_y), Point<double>(x, y)))
}
int gar_cart_dim() { return (int)Cartographing::patch_dim_; }
int gar_cart_get(uint8_t* dst, size_t bytes)
{
  if (bytes != Cartographing::patch_dim_ * Cartographing::patch_dim_) return -1;
  std::memcpy(dst, Cartographing::patch_arr_.getPtr(), bytes);
  return 0;
}
double gar_con2gm() { return grid_tf::con2gm_; }

// ---- proximity scoring.  util::bilinearInterpolation / util::getBilinInterp,
// Vehicle's corners, Point::rotatePreCalc and AStar::h_prox_arr_ are the reference's own;
// HybridAStar::getProxOfCorners itself (hybrid_a_star_lib.cpp:265-289) lives in a TU that needs the
// Reeds-Shepp / spline side of the planner and is not built here, so its 15 lines are composed
// below from those parts, statement by statement.
double gar_bilinear(double q11, double q12, double q21, double q22, int x1, int x2, int y1, int y2, double x, double y)
{
  return util::bilinearInterpolation(q11, q12, q21, q22, x1, x2, y1, y2, x, y);
}
double gar_bilin_interp(double x, double y) { return util::getBilinInterp(x, y, AStar::h_prox_arr_); }
double gar_prox_of_corners(double x, double y, double yaw)  // pose in astar cells (pose * con2star_)
{
  const double cos_yaw = cos(yaw), sin_yaw = sin(yaw);
  const Point<double> offset(x, y);
  const Point<double> f_r = Vehicle::f_r_corner_.rotatePreCalc(cos_yaw, sin_yaw) + offset;
  const Point<double> f_l = Vehicle::f_l_corner_.rotatePreCalc(cos_yaw, sin_yaw) + offset;
  const Point<double> r_r = Vehicle::r_r_corner_.rotatePreCalc(cos_yaw, sin_yaw) + offset;
  const Point<double> r_l = Vehicle::r_l_corner_.rotatePreCalc(cos_yaw, sin_yaw) + offset;
  const double v_1 = util::getBilinInterp(f_r.x, f_r.y, AStar::h_prox_arr_);
  const double v_2 = util::getBilinInterp(f_l.x, f_l.y, AStar::h_prox_arr_);
  const double v_3 = util::getBilinInterp(r_r.x, r_r.y, AStar::h_prox_arr_);
  const double v_4 = util::getBilinInterp(r_l.x, r_l.y, AStar::h_prox_arr_);
  const double max_value = std::max({ v_1, v_2, v_3, v_4 });
  return std::isnan(max_value) ? 0.0 : max_value;
}
double gar_con2star() { return grid_tf::con2star_; }
// step 3 of HybridAStar::getValidClosePose (hybrid_a_star_lib.cpp:1604-1668), composed like
// getProxOfCorners above: the loop is restated, CollisionChecker::checkPose, AStar::calcIndex, grid_tf and
// Point::toInt are the reference's.  near_cells = calcIndex of the nodes of AStar::nodes_near_goal_.
int gar_valid_close_pose(const double* center, const int* near_cells, int n_near, double yaw_res_coll, double* out_pose)
{
  const Pose<double> center_pose = { center[0], center[1], center[2] };
  const double angle_weight = 20.0, dist_weight = 1.0, prox_weight = 1.0;
  const double dxy_max = 5.0, dphi_max = util::PI;
  const double gm_res = GM_RES;
  const double phi_res = yaw_res_coll * util::TO_RAD;
  double min_cost = std::numeric_limits<double>::max();
  int found_any = 0;
  const int dxy_max_idx = static_cast<int>(std::floor(dxy_max) / gm_res);
  const int dphi_max_idx = static_cast<int>(std::floor(dphi_max) / phi_res);
  for (int dxi = -dxy_max_idx; dxi < dxy_max_idx; ++dxi)
    for (int dyi = -dxy_max_idx; dyi < dxy_max_idx; ++dyi)
      for (int dphii = -dphi_max_idx; dphii < dphi_max_idx; ++dphii)
      {
        const double x_diff = dxi * gm_res, y_diff = dyi * gm_res, phi_diff = dphii * phi_res;
        const double dist2 = x_diff * x_diff + y_diff * y_diff;
        const Pose<double> pose = { center_pose.x + x_diff, center_pose.y + y_diff, center_pose.yaw + phi_diff };
        const Pose<double> ps = grid_tf::utm2astar(pose);
        const double prox_cost = gar_prox_of_corners(ps.x, ps.y, ps.yaw);
        const double cost = dist2 * dist_weight + std::abs(phi_diff) * angle_weight + prox_cost * prox_weight;
        if (cost > min_cost) continue;
        const Point<int> index2d = grid_tf::utm2astar(pose.getPoint()).toInt();
        const size_t index = AStar::calcIndex(index2d.x, index2d.y);
        bool found = false;
        for (int k = 0; k < n_near; ++k)
          if (index == (size_t)near_cells[k])
          {
            found = true;
            break;
          }
        if (not found) continue;
        if (CollisionChecker::checkPose(pose))
        {
          out_pose[0] = pose.x;
          out_pose[1] = pose.y;
          out_pose[2] = pose.yaw;
          min_cost = cost;
          found_any = 1;
        }
      }
  return found_any;
}
// the prox_cost loop of getRSPathCosts / getRAPathCosts (hybrid_a_star_lib.cpp:588-594) over a ragged batch
void gar_path_prox_costs(const double* poses_m, const int* offsets, int n_paths, double h_prox_cost, double interp_res, double* out)
{
  for (int i = 0; i < n_paths; ++i)
  {
    double prox_cost = 0;
    for (int j = offsets[i]; j < offsets[i + 1]; ++j)
      prox_cost += gar_prox_of_corners(poses_m[3 * j] * grid_tf::con2star_, poses_m[3 * j + 1] * grid_tf::con2star_, poses_m[3 * j + 2]) *
                   h_prox_cost * interp_res;
    out[i] = prox_cost;
  }
}

}  // extern "C"
