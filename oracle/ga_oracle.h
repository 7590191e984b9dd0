/*
 * ga_oracle.h -- CPU restatement (plain C) of the grid hot path of
 * uulm-mrm/guided-extended-hybrid-astar's freespace_planner library.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product library
 * (libgridaccel.so) never links, loads or calls anything in oracle/.
 *
 * Parity status: PINNED.  The OpenCV-dependent arithmetic (ellipse structuring element,
 * dilate, distanceTransform) is pinned against cv2 4.13.0 outputs committed under
 * tests/golden/ (generator: tests/golden/make_golden.py); everything else is pinned against
 * the reference's own translation units compiled unmodified into oracle/_ref/ (see
 * oracle/Makefile, oracle/ref_shim/) and against fixtures generated from that build.
 *
 * Every function cites the reference file:line it restates.  LIB =
 * src/freespace_planner/library in the reference tree.
 */
#ifndef GA_ORACLE_H
#define GA_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* LIB/include/collision_checker_lib/collision_checking.hpp:39-47 */
enum { GAO_FREE = 0, GAO_UNKNOWN = 1, GAO_OCC = 2, GAO_SENSOR_UNKNOWN = 127 };

/* LIB/include/params.hpp:7-8 */
#define GAO_ASTAR_RES 0.625
#define GAO_GM_RES 0.15625

/* ---- occupancy: threshold, dilation, pooling ------------------------------------------- */

/* cv::getStructuringElement(MORPH_ELLIPSE, (2r+1)x(2r+1), anchor (r,r)) as called at
 * LIB/src/collision_checker_lib/collision_checking.cpp:106-107 (OpenCV 4.8.0 imgproc/morph).
 * Writes the half-width of every row (dx such that columns [r-dx, r+dx] are set) into
 * half_w[2r+1]; optional dense mask se[(2r+1)^2] (may be NULL). Returns number of taps. */
int gao_ellipse_rows(int r, int* half_w, uint8_t* se);

/* collision_checking.cpp:213-221: v<min -> FREE, v>max -> OCC, else UNKNOWN */
void gao_threshold(const uint8_t* patch, uint8_t* out, size_t n, int min_thresh, int max_thresh);

/* cv::dilate(out, out, dil_kernel_) at collision_checking.cpp:223: max over the in-image part
 * of the structuring element (default BORDER_CONSTANT with morphologyDefaultBorderValue ==
 * "ignore outside pixels" for dilation). in != out. */
void gao_dilate_disk(const uint8_t* in, uint8_t* out, int dim, int r);

/* CollisionChecker::processSafetyPatch, collision_checking.cpp:193-224 (CPU branch) */
void gao_process_safety_patch(const uint8_t* patch, uint8_t* safety, int dim, int min_thresh,
                              int max_thresh, int disk_r_c);

/* Pooling::execute (LIB/src/cuda_lib/max_pool.cu:21-79) via AStar::calcAstarGridCuda
 * (a_star.cpp:536-542): 4x4 / stride 4 max, floor output size */
void gao_pool_max(const uint8_t* safety, uint8_t* astar, int patch_dim, int astar_dim, int win);

/* CollisionChecker::passLocalMapData, collision_checking.cpp:269-313 */
void gao_pass_local_map(uint8_t* patch, int patch_dim, const uint8_t* local_map, int ox, int oy,
                        int dim);

/* ---- footprint disks + collision lookups ------------------------------------------------ */

typedef struct
{
  int nb_disks;
  int nb_yaws;
  int yaw_res_deg;
  double disk_r;   /* disk_r_  (m)     */
  int disk_r_c;    /* disk_r_c_ (cells) */
} gao_disks_info;

/* Cartographing::cartograph (LIB/src/cartographing_lib/cartographing.cpp:17-96): cells of the
 * dim x dim window (row-major, [y][x]) that land inside the cartographed patch overwrite it only
 * where the patch is still SENSOR_UNKNOWN (127) and the window value is 255 or 0. */
void gao_cartograph(uint8_t* cart, int patch_dim, const uint8_t* local_map, int ox, int oy, int dim);

/* util::copyPatch2Patch (LIB/include/util_lib/util2.hpp:32-67) as Cartographing::loadPrevPatch uses
 * it (cartographing.cpp:119-129): copy the overlap of the previous patch (origin prev_o, gm cells)
 * into the new one (origin o).  The reference's end point is clamped to next_dim - 1 and the loop
 * is exclusive, so a clamped overlap loses its last row / column: reproduced. */
void gao_copy_patch_to_patch(const uint8_t* prev, int prev_dim, int prev_ox, int prev_oy, uint8_t* next,
                             int next_dim, int ox, int oy);

/* CollisionChecker::calculateDisks / getDiskRadius / getDiskPositions,
 * collision_checking.cpp:44-145.  centers[(yaw_idx*nb_disks + disk)*2 + {0:x,1:y}] */
int gao_calc_disks(double veh_length, double veh_width, double veh_lb, double len_per_disk,
                   int double_disk_rows, double safety_distance_m, int yaw_res_deg,
                   gao_disks_info* info, int* centers, int centers_cap);

/* util::constrainAngleZero2Pi / constrainAngleMinPIPlusPi, LIB/src/util_lib/util1.cpp:8-24 */
double gao_angle_0_2pi(double a);
double gao_angle_mpi_pi(double a);
/* util::getAngleDiff util1.cpp:26-34 */
double gao_angle_diff(double a1, double a2);

/* CollisionChecker::getYawIdx, collision_checking.cpp:333-337 */
int gao_yaw_idx(double yaw, int yaw_res_deg);

/* checkGrid / checkGridVariant, collision_checking.cpp:344-387; returns 1 = valid */
int gao_check_grid(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks, int x,
                   int y, int yaw_idx, int variant);
/* checkPose / checkPoseVariant, collision_checking.cpp:389-409 */
int gao_check_pose(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                   int yaw_res_deg, double x, double y, double yaw, int variant);
/* getPathCollisionIndex, collision_checking.cpp:440-452 (checkPathCollision == (idx < 0)) */
int gao_path_collision_index(const uint8_t* safety, int patch_dim, const int* centers,
                             int nb_disks, int yaw_res_deg, const double* x, const double* y,
                             const double* yaw, int n, int variant);
/* Batched form used as the checker of ga_check_poses: segment s covers poses
 * [seg_off[s], seg_off[s+1]); first_hit[s] = index inside the segment or -1 */
void gao_check_segments(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                        int yaw_res_deg, const double* x, const double* y, const double* yaw,
                        const int* seg_off, int nseg, int variant, int* first_hit);

/* ---- motion primitives ------------------------------------------------------------------ */

/* Vehicle::move, LIB/src/collision_checker_lib/vehicle.cpp:34-43 */
void gao_vehicle_move(double* x, double* y, double* yaw, double distance, double steer,
                      double wheel_base);
/* Vehicle::move_car_some_steps, vehicle.cpp:100-122; returns number of poses written */
int gao_move_car_some_steps(double x, double y, double yaw, double arc_l, double motion_res,
                            int direction, double steer, double wheel_base, double* xs,
                            double* ys, double* yaws, int cap);
/* Vehicle::turn_on_rear_axis, vehicle.cpp:45-98; returns number of yaws written */
int gao_turn_on_rear_axis(double yaw, double delta_angle, double yaw_res_deg, double* yaws,
                          int cap);
/* HybridAStar::calcSteeringInputs, LIB/src/hybridastar_planning_lib/hybrid_a_star_lib.cpp:115-129 */
void gao_steering_inputs(double max_steer, int nb_steer, double* out);

/* One expansion as HybridAStar::setNeighbors (hybrid_a_star_lib.cpp:492-534) does it for the
 * collision-relevant part: nb_steer x {+1,-1} primitives (calcNextNode :427-483: rollout,
 * mapCont2Disc :410-417, verifyIndex :150-154, checkPathCollision) and, if with_rear_axis,
 * the rear-axis rotations (calcRearAxisNode :350-403) for delta in
 * {-2pi+step, ..., 2pi-step} \ {0}.  Outputs per primitive k (order = reference order:
 * steer-major, direction +1 then -1, then rear-axis deltas):
 *   valid[k] (1 = neighbour created), disc[k*3+{x,y,yaw}], last pose end[k*3+{x,y,yaw}],
 *   n_poses[k].  Returns number of primitives. */
typedef struct
{
  double max_steer, wheel_base, arc_l, yaw_res_astar_deg; /* YAW_RES (deg) of the search */
  int nb_steer;
  int with_rear_axis;
  double turn_on_point_angle_deg; /* TURN_ON_POINT_ANGLE */
  int astar_dim;
} gao_expand_params;

int gao_expand_node(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks,
                    int yaw_res_coll_deg, const gao_expand_params* p, double x, double y,
                    double yaw, double motion_res, uint8_t* valid, int* disc, double* end_pose,
                    int* n_poses, int cap);

/* ---- heuristic grid ---------------------------------------------------------------------- */

typedef struct
{
  int astar_dim;
  double astar_prox_cost;          /* params.yml astar_prox_cost_ */
  double astar_lane_movement_cost; /* params.yml astar_lane_movement_cost_ */
  double unknown_cost_w;           /* config.yml ASTAR_UNKNOWN_COST */
  int heuristic_early_exit;        /* config.yml HEURISTIC_EARLY_EXIT */
  unsigned int max_extra_nodes;    /* config.yml MAX_EXTRA_NODES_ASTAR */
  int get_only_near;
} gao_heur_params;

/* flags bits of the dense result */
enum { GAO_H_CLOSED = 1, GAO_H_UNKNOWN = 2 };

/* AStar::calcDistanceHeuristic in the reference's exact pop order (a_star.cpp:110-295) on
 * dense arrays instead of unordered_maps.  Outputs are dense astar_dim^2 arrays; cells not in
 * the closed set keep cost = +inf, parent = -2, flags = 0.  closed_order (cap astar_dim^2)
 * receives cell ids in closing order; near_ids/near_dist (cap astar_dim^2) the
 * nodes_near_goal_ list.  Returns number of closed cells. */
int gao_heuristic_reforder(const gao_heur_params* p, const uint8_t* astar_grid,
                           const uint8_t* geofence, const double* movement_cost,
                           const double* h_prox, int goal_x, int goal_y, int start_x,
                           int start_y, double* cost, double* cost_dist, int32_t* parent,
                           uint8_t* flags, int32_t* closed_order, int32_t* near_ids,
                           double* near_dist, int* n_near);

/* The fixed point the wavefront kernels compute: least solution of
 *   g(goal) = 0,  g(n) = min_p fl(fl(fl(g(p) + fl(len_i*mc(n))) + prox(p)*w) + unk(p))
 * over valid n (a_star.cpp:229-257 edge costs, no early exit, no closed-set order effects).
 * parent = predecessor with the smallest motion index i whose candidate equals g(n);
 * cost_dist = n_straight*0.625 + n_diag*(0.625*sqrt2) along the parent chain. */
int gao_heuristic_fixedpoint(const gao_heur_params* p, const uint8_t* astar_grid,
                             const uint8_t* geofence, const double* movement_cost,
                             const double* h_prox, int goal_x, int goal_y, double* cost,
                             double* cost_dist, int32_t* parent, uint8_t* flags);

/* AStar::getAstarPath + findValidNeighborIndex (a_star.cpp:351-421) on the dense result.
 * Returns number of path cells written (0 when no start neighbour is in the set). */
int gao_astar_path(int astar_dim, const int32_t* parent, const uint8_t* flags, int x, int y,
                   int* xs, int* ys, int cap);

/* ---- distance transform + potential field ---------------------------------------------- */

enum { GAO_DT_FIXEDPOINT = 0, GAO_DT_IPPFLOAT = 1 };

/* cv::distanceTransform(src, DIST_L2, DIST_MASK_5) (OpenCV 4.8.0 imgproc/distransform.cpp) on
 * a w x h u8 image with row stride `stride`: distance (in cells) to the nearest zero pixel.
 * variant FIXEDPOINT = OpenCV's own 16.16 integer chamfer; IPPFLOAT = the IPP float32 raster
 * chamfer OpenCV dispatches to when built WITH_IPP. */
void gao_distance_transform(const uint8_t* src, int stride, int w, int h, float* dist,
                            int variant);

typedef struct
{
  double alpha, do_max, do_min;                 /* params.yml */
  double motion_res_min, motion_res_max;        /* config.yml MOTION_RES_MIN/MAX */
  double dist_val_min, dist_val_max;            /* config.yml MIN/MAX_DIST_VAL */
  int vor_dim;                                  /* a_star.hpp:36 VOR_DIM */
  int dt_variant;
} gao_pot_params;

/* AStar::calcVoronoiPotentialField (a_star.cpp:440-479) with the Fortune sweep (deps_lib,
 * host-side third-party code) replaced by its output: the list of Voronoi edge endpoint
 * cells (window coordinates) it marks.  obstacle_dist / voronoi_dist are the persistent
 * vor_dim^2 float windows (in/out, a_star.cpp:80-81,472-473); h_prox / motion_res the
 * astar_dim^2 maps (in/out).  Returns 0 if the window holds no OCC cell (early return,
 * a_star.cpp:462-465; obstacle_dist is still rewritten), else 1. */
int gao_voronoi_potential_field(const gao_pot_params* p, const uint8_t* astar_grid,
                                int astar_dim, int ego_x, int ego_y, const int* vor_xy,
                                int n_vor, float* obstacle_dist, float* voronoi_dist,
                                double* h_prox, double* motion_res);

/* AStar::setCostmapValue, a_star.cpp:486-508, for one cell */
void gao_costmap_value(const gao_pot_params* p, float d_obs, float d_vor, double* h_prox,
                       double* motion_res);

/* ---- proximity scoring of sampled poses ---- */
double gao_bilinear(double q11, double q12, double q21, double q22, int x1, int x2, int y1, int y2, double x, double y);
double gao_bilin_at(double x, double y, const double* grid, int dim);
typedef struct { double lf, lb, width, con2star, h_prox_cost, interp_res; } gao_prox_params;
double gao_prox_of_corners(const double* grid, int dim, const gao_prox_params* p, double x_m, double y_m, double yaw);
void gao_path_prox_costs(const double* grid, int dim, const gao_prox_params* p, const double* poses, const int* offsets,
                         int n_paths, double* out);
int gao_valid_close_pose(const uint8_t* safety, int patch_dim, const int* centers, int nb_disks, int yaw_res_deg,
                         const double* prox, int astar_dim, const gao_prox_params* p, const uint8_t* member,
                         const double* center, double* out_pose);

#ifdef __cplusplus
}
#endif
#endif
