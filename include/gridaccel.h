/*
 * gridaccel.h -- C ABI of libgridaccel.so, the B200 (sm_100a) backend for the data-parallel grid
 * work of guided-extended-hybrid-astar's freespace_planner library.
 *
 * The reference has no FFI layer for this path: the work sits behind two all-static C++ classes
 * (CollisionChecker, AStar) and one pybind11 module.  Every entry point below names the reference
 * function(s) it replaces (LIB = src/freespace_planner/library in the reference tree).  The C++
 * mirror classes that keep the reference's names on top of this ABI live in
 * guided-extended-hybrid-astar_b200/host/, the binding a reference maintainer would add is in
 * INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every host buffer, the context owns every
 *     device buffer and one CUDA stream;
 *   - every function returns 0 on success, a negative ga_status on failure, never throws, never
 *     exits; ga_last_error() gives the text of the last failure of that context;
 *   - all work of one context is ordered on its stream; functions that return data to the host
 *     synchronise that stream, the others only enqueue (ga_sync waits);
 *   - one context per planner instance; a context is not thread-safe, different contexts are
 *     independent (this is what lifts the reference's one-planner-per-process restriction);
 *   - there is no CPU fallback: without a CUDA device ga_create fails with GA_ERR_CUDA.
 *   - grids are dense row-major, index = y * dim + x, exactly Vec2DFlat<T>
 *     (LIB/include/util_lib/data_structures1.hpp:338-514).
 */
#ifndef GRIDACCEL_H
#define GRIDACCEL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GA_ABI_VERSION 1

#if defined(GA_BUILD) && defined(__GNUC__)
#define GA_API __attribute__((visibility("default")))
#else
#define GA_API
#endif

typedef struct ga_ctx ga_ctx;

typedef enum
{
  GA_OK = 0,
  GA_ERR_ARG = -1,    /* bad argument / call order */
  GA_ERR_CUDA = -2,   /* CUDA runtime error (text in ga_last_error) */
  GA_ERR_STATE = -3,  /* required state missing (no patch, no params ...) */
  GA_ERR_NOMEM = -4
} ga_status;

/* GRID_VAL, LIB/include/collision_checker_lib/collision_checking.hpp:39-47 */
enum { GA_FREE = 0, GA_UNKNOWN = 1, GA_OCC = 2, GA_SENSOR_UNKNOWN = 127 };

/* device grids that ga_upload / ga_download / ga_device_ptr address */
typedef enum
{
  GA_GRID_PATCH = 0,          /* u8  patch_dim^2   CollisionChecker::patch_arr_            */
  GA_GRID_SAFETY = 1,         /* u8  patch_dim^2   CollisionChecker::patch_safety_arr_     */
  GA_GRID_ASTAR = 2,          /* u8  astar_dim^2   AStar::astar_grid_                      */
  GA_GRID_GEOFENCE = 3,       /* u8  astar_dim^2   AStar::restr_geofence_grid_             */
  GA_GRID_MOVEMENT_COST = 4,  /* f64 astar_dim^2   AStar::movement_cost_map_               */
  GA_GRID_H_PROX = 5,         /* f64 astar_dim^2   AStar::h_prox_arr_                      */
  GA_GRID_MOTION_RES = 6,     /* f64 astar_dim^2   AStar::motion_res_map_                  */
  GA_GRID_OBSTACLE_DIST = 7,  /* f32 vor_dim^2     AStar::obstacle_dist_                   */
  GA_GRID_VORONOI_DIST = 8,   /* f32 vor_dim^2     AStar::voronoi_dist_                    */
  GA_GRID_CART = 9,           /* u8  cart_dim^2    Cartographing::patch_arr_ (accumulated map) */
  GA_GRID_PMAP = 10,          /* u8  801^2         GridMapSim::pmap_ (last ray-cast window)    */
  /* heuristic result planes, slot 0 = closed_set_guidance_, slot 1 = closed_set_path_
   * (dense restatement of std::unordered_map<size_t, NodeDisc>, a_star.hpp:69-72,
   * NodeDisc data_structures2.hpp:47-67): id = base + slot */
  GA_GRID_HEUR_COST = 16,     /* f64 astar_dim^2   NodeDisc::cost_        (+inf = not in set) */
  GA_GRID_HEUR_DIST = 18,     /* f64 astar_dim^2   NodeDisc::cost_dist_                      */
  GA_GRID_HEUR_PARENT = 20,   /* i32 astar_dim^2   NodeDisc::parent_index_ (-1 root, -2 none) */
  GA_GRID_HEUR_FLAGS = 22,    /* u8  astar_dim^2   bit0 in closed set, bit1 is_unknown_       */
  /* download only: one byte per cell for callers that need membership and the parent chain but take costs from
   * the device (getAstarPath a_star.cpp:386-421, findValidNeighborIndex :351-378, calcCost's OUT_OF_HEURISTIC test
   * hybrid_a_star_lib.cpp:242-247): bits 0-3 = i with parent = cell - motion_[i] (a_star.hpp:50-52), 8 = root,
   * 15 = none; bit 4 = closed; bit 5 = is_unknown_.  1 B/cell over PCIe instead of 21. */
  GA_GRID_HEUR_PACKED = 24,
  /* (+slot) download only, after a device (GA_HEUR_FIXEDPOINT) build of the slot: the four result planes in 12 bytes
   * per cell instead of 21 -- three 32-bit words: the 64 bits of cost_ (low word first; +inf = not in the set), then
   * dir:3 | is_unknown_:1 | straight:14 | diagonal:14 with parent = cell - motion_[dir] (a_star.hpp:50-52) and
   * cost_dist_ = straight * 0.625 + diagonal * (0.625 * sqrt 2) (two products, one sum: what the device computes).
   * The root is the closed cell with zero steps.  A chain of 16383 or more steps of one kind does not fit:
   * ga_download (or the ga_sync after ga_download_async) then fails with GA_ERR_STATE; use the four planes. */
  GA_GRID_HEUR_REC = 26
} ga_grid_id;

enum { GA_H_CLOSED = 1, GA_H_UNKNOWN = 2 };

/* ---- lifetime ------------------------------------------------------------------------------ */

GA_API int ga_abi_version(void);
GA_API int ga_create(int device, ga_ctx** out);
GA_API int ga_destroy(ga_ctx* ctx);
GA_API const char* ga_last_error(const ga_ctx* ctx);
GA_API int ga_sync(ga_ctx* ctx);
/* the context's cudaStream_t, for callers that record their own events on it */
GA_API void* ga_stream(ga_ctx* ctx);

/* ---- occupancy: patch, threshold, dilation, pooling --------------------------------------- */

/* CollisionChecker::initialize / calculateDisks (collision_checking.cpp:12-31,75-145): the host
 * computes the footprint model exactly as the reference does and hands over the results.
 * centers = disk_centers_ LUT, [(yaw_idx * nb_disks + disk) * 2 + {0:x, 1:y}] gm cells.
 * disk_r_c = CollisionChecker::disk_r_c_; the structuring element is the
 * cv::getStructuringElement(MORPH_ELLIPSE, (2r+1)^2) raster of collision_checking.cpp:106-107. */
typedef struct
{
  int min_thresh;   /* config.yml MIN_THRESH */
  int max_thresh;   /* config.yml MAX_THRESH */
  int disk_r_c;     /* dilation radius in gm cells (<= 16) */
  int yaw_res_deg;  /* config.yml YAW_RES_COLL */
  int nb_disks;     /* <= 16 */
  int pool_win;     /* ceil(grid_tf::star2gm_) = 4 (a_star.cpp:30-31) */
} ga_collision_params;

GA_API int ga_set_collision_params(ga_ctx* ctx, const ga_collision_params* p, const int* centers);

/* CollisionChecker::resetPatch (collision_checking.cpp:37-42) + AStar::reinit's grid sizing
 * (a_star.cpp:70-99): patch_arr_ <- 127, patch_safety_arr_ <- UNKNOWN, astar_dim =
 * floor(patch_dim / pool_win); astar_grid_ <- UNKNOWN, movement cost <- astar_movement_cost,
 * h_prox <- 0, motion_res <- motion_res_max, geofence <- 0, distance windows <- 0, heuristic
 * slots emptied.  (Carrying h_prox / motion_res over a patch move, a_star.cpp:84-96, is a
 * download / shift / upload the host mirror does.) */
GA_API int ga_patch_reset(ga_ctx* ctx, int patch_dim);

/* CollisionChecker::passLocalMapData (collision_checking.cpp:269-313): copy a dim x dim u8
 * window (host memory) into patch_arr_ at (ox, oy).  Whole window dropped when it overhangs in
 * x, rows skipped when they overhang in y -- as the reference does.  Only the window crosses
 * PCIe; the patch stays device-resident. */
GA_API int ga_patch_insert(ga_ctx* ctx, const uint8_t* window, int ox, int oy, int dim);

/* CollisionChecker::processSafetyPatch (collision_checking.cpp:193-224) followed by
 * AStar::calcAstarGridCuda -> Pooling::execute (a_star.cpp:536-542, LIB/src/cuda_lib/max_pool.cu:21-79):
 * threshold -> disk dilation -> patch_safety_arr_, and its pool_win x pool_win max -> astar_grid_. */
GA_API int ga_build_safety(ga_ctx* ctx);

/* ---- accumulated map of the unknown-environment loop --------------- */

/* Cartographing (LIB/src/cartographing_lib/cartographing.cpp:7-129): the map every sensor window is
 * merged into, device-resident next to the collision checker's patch.
 * ga_cart_reset        = resetPatch (:7-15): the old map is kept as temp_patch_, the new one is all 127;
 * ga_cartograph        = cartograph (:17-96): cells of the dim x dim host window that land inside
 *                        overwrite the map only where it is still 127 and the window holds 255 or 0;
 * ga_cart_pass_local_map = passLocalMap (:98-117): window of the map -> CollisionChecker::patch_arr_
 *                        (device to device, passLocalMapData's clipping); the window must lie inside
 *                        the map (the reference reads out of bounds otherwise);
 * ga_cart_load_prev    = loadPrevPatch (:119-129) + util::copyPatch2Patch (util2.hpp:32-67): overlap
 *                        of the previous map copied in, then the whole map handed to the collision
 *                        checker.  Origins in gm cells = (utm * con2gm).toInt(), i.e. std::round. */
GA_API int ga_cart_reset(ga_ctx* ctx, int patch_dim);
GA_API int ga_cartograph(ga_ctx* ctx, const uint8_t* window, int ox, int oy, int dim);
GA_API int ga_cart_pass_local_map(ga_ctx* ctx, int ox, int oy, int dim);
GA_API int ga_cart_load_prev(ga_ctx* ctx, int prev_ox, int prev_oy, int ox, int oy);
GA_API int ga_cart_dim(const ga_ctx* ctx);

/* GridMapSim::gmSimulation / raytracing (LIB/src/gridmap_sim_lib/gridmap_sim.cpp:42-103, SURVEY.md
 * 8(f) rank 2): the simulated lidar of the unknown-environment loop.  grid_map_gt = 801 x 801 u8
 * ground-truth window around the ego cell in host memory (255 = occupied); result: 0 free up to the
 * first occupied cell of each of the 180 two-degree sectors, 255 that cell, 127 behind it.  It stays
 * on the device (GA_GRID_PMAP) and is also copied to pmap_out when that is not NULL.
 * ga_cartograph_raytraced = Cartographing::cartograph of that device-resident window at (ox, oy)
 * (simulation.py:262-277 -> planning.py): the cycle window -> ray-cast -> merge -> hand-over ->
 * safety grid never leaves the GPU after the ground-truth upload. */
GA_API int ga_raytrace(ga_ctx* ctx, const uint8_t* grid_map_gt, uint8_t* pmap_out);
GA_API int ga_cartograph_raytraced(ga_ctx* ctx, int ox, int oy);

/* ---- footprint collision checks ------------------------------------------------------------ */

/* CollisionChecker::getPathCollisionIndex / checkPathCollision / checkPose(/Variant)
 * (collision_checking.cpp:389-452), batched: segment s covers poses [seg_off[s], seg_off[s+1]);
 * first_hit[s] = index inside the segment of the first colliding pose, -1 if none.
 * variant = 0: an off-map disk invalidates the pose (checkGrid :344-362); 1: it is skipped
 * (checkGridVariant :369-387).  Host pointers; one kernel launch for the whole batch. */
GA_API int ga_check_poses(ga_ctx* ctx, const double* x, const double* y, const double* yaw, int n_poses,
                   const int* seg_off, int n_seg, int variant, int* first_hit);

/* HybridAStar::setNeighbors' collision-relevant part for a batch of expanded nodes
 * (LIB/src/hybridastar_planning_lib/hybrid_a_star_lib.cpp:492-534: calcNextNode :427-483,
 * calcRearAxisNode :350-403, Vehicle::move_car_some_steps / turn_on_rear_axis
 * LIB/src/collision_checker_lib/vehicle.cpp:45-122, mapCont2Disc :410-417): one launch rolls out
 * every primitive of every node and checks every pose. */
typedef struct
{
  double max_steer;               /* Vehicle::max_steer_ */
  double wheel_base;              /* Vehicle::w_b_ */
  double arc_l;                   /* HybridAStar::arc_l_ */
  double yaw_res_deg;             /* config.yml YAW_RES (search yaw discretisation) */
  double turn_on_point_angle_deg; /* config.yml TURN_ON_POINT_ANGLE */
  int nb_steer;                   /* NB_STEER = 13 (hybrid_a_star_lib.hpp:48), <= 31 */
} ga_expand_params;

typedef struct
{
  double x, y, yaw;    /* parent pose, metres / rad, patch frame */
  double motion_res;   /* AStar::motion_res_map_ at the parent (hybrid_a_star_lib.cpp:1483) */
  int32_t with_rear_axis; /* 1: also the U-Shift rear-axis rotations (:512-533) */
  int32_t pad_;
} ga_node;

typedef struct
{
  double x, y, yaw;    /* last pose of the primitive */
  int32_t xi, yi, yawi; /* mapCont2Disc of the last pose */
  int32_t n_poses;     /* poses rolled out */
  uint8_t valid;       /* 1 = neighbour would be created (on grid and collision-free) */
  uint8_t near_cell_edge; /* 1 = some pose lies within 1e-9 cells of a cell / yaw-bin boundary:
                             device libm and glibc may disagree there, re-check on the host */
  uint8_t collision_free; /* 1 = every rolled-out pose passed checkPose (valid = collision_free && on the astar grid) */
  uint8_t collision_ambiguous; /* 1 = some pose lies within 1e-9 of a gm-cell / yaw-bin edge AND the neighbouring cell or
                             bin gives another answer: collision_free may differ for poses rolled out with glibc */
  uint8_t pad_[4];
} ga_succ;

/* number of primitives ga_expand writes per node: 2*nb_steer, + rear-axis deltas when the node
 * asks for them; out must hold n_nodes * ga_expand_stride(ctx) entries, unused entries have
 * n_poses = 0 */
GA_API int ga_set_expand_params(ga_ctx* ctx, const ga_expand_params* p);
GA_API int ga_expand_stride(const ga_ctx* ctx);
GA_API int ga_expand(ga_ctx* ctx, const ga_node* nodes, int n_nodes, ga_succ* out);

/* ga_expand with the successor epilogue: the grid reads the serial search makes for every
 * successor are done in the same launch, so movement_cost_map_, motion_res_map_, h_prox_arr_ and the closed-set
 * planes never have to be copied to the host.  Per (node, primitive), same layout as `out`:
 *   heur_cost     closed_set (slot) cost_ at the successor's astar cell, GA_OUT_OF_HEURISTIC when the cell is not in
 *                 the set (HybridAStar::calcCost, hybrid_a_star_lib.cpp:242-250)
 *   heur_dist     cost_dist_ there (getDistance2goal, :781-798), GA_OUT_OF_HEURISTIC when absent
 *   movement_cost AStar::movement_cost_map_(yi, xi) (getPathCosts, :324)
 *   motion_res    AStar::motion_res_map_(yi, xi): the resolution the successor is rolled out with when it is
 *                 expanded itself (:1483)
 * and per node out_parent_prox = getProxOfCorners(parent pose * con2star_) (getPathCosts :335-336, the same for all
 * primitives of the node; device sin / cos: 1e-15 relative, see ga_prox_of_corners).  Needs ga_set_prox_params. */
#define GA_OUT_OF_HEURISTIC 999999999.0 /* HybridAStar::OUT_OF_HEURISTIC, hybrid_a_star_lib.hpp:103 */
typedef struct
{
  double heur_cost, heur_dist, movement_cost, motion_res;
} ga_succ_cost;
/* Asynchronous ga_expand for the nodes a search will expand NEXT (the successors it has just created; <= 32 per
 * submission, <= 4 submissions in flight): ga_expand_submit launches and returns, ga_expand_collect waits for that
 * submission (usually it is long finished) and copies its n_nodes * ga_expand_stride results.  The device round trip
 * then overlaps the caller's own work instead of sitting on its critical path. */
GA_API int ga_expand_submit(ga_ctx* ctx, const ga_node* nodes, int n_nodes, int* ticket);
GA_API int ga_expand_collect(ga_ctx* ctx, int ticket, ga_succ* out, int* n_nodes);
/* the same wait, but only one byte per primitive comes back (n_nodes * ga_expand_stride bytes) -- all a caller needs
 * that rolls the poses out itself and wants the collision answers (the drop-in's look-ahead) */
enum { GA_SUCC_ROLLED = 1, GA_SUCC_FREE = 2, GA_SUCC_AMBIGUOUS = 4, GA_SUCC_VALID = 8 };
GA_API int ga_expand_collect_flags(ga_ctx* ctx, int ticket, uint8_t* flags, int* n_nodes);
GA_API int ga_expand_ex(ga_ctx* ctx, const ga_node* nodes, int n_nodes, int slot, ga_succ* out, ga_succ_cost* out_cost,
                        double* out_parent_prox);

/* ---- proximity scoring of sampled poses ----------------------------- */

typedef struct
{
  double lf, lb, width; /* Vehicle::lf_, lb_, width_ (vehicle.cpp:16-27): the corner offsets, metres */
  double con2star;      /* grid_tf::con2star_: metres -> astar cells */
  double h_prox_cost;   /* HybridAStar::h_prox_cost_ */
  double interp_res;    /* HybridAStar::interp_res_ (sampling distance of the analytic paths) */
} ga_prox_params;

GA_API int ga_set_prox_params(ga_ctx* ctx, const ga_prox_params* p);

/* HybridAStar::getProxOfCorners(pose * grid_tf::con2star_) (hybrid_a_star_lib.cpp:265-289, with
 * util::getBilinInterp, util2.cpp:41-52): poses = n x (x, y, yaw) in metres / rad, patch frame;
 * out[i] = the largest bilinear sample of h_prox_arr_ under the four vehicle corners.  Cell indices
 * are clamped into the grid (the reference clamps at 0 only and reads past the array beyond the
 * last row / column).  fp64; sin / cos are the device's, so results agree with the host's to ~1e-15
 * relative, not bit for bit. */
GA_API int ga_prox_of_corners(ga_ctx* ctx, const double* poses, int n, double* out);

/* The prox-cost loops of getRAPathCosts / getRSPathCosts (hybrid_a_star_lib.cpp:588-594, 670-676)
 * for n_paths sampled paths at once: path i owns poses[path_offsets[i] .. path_offsets[i+1])
 * (ragged, may be empty; path_offsets[0] = 0), out[i] = sum in list order of
 * getProxOfCorners(pose * con2star_) * h_prox_cost_ * interp_res_. */
GA_API int ga_path_prox_costs(ga_ctx* ctx, const double* poses, const int* path_offsets, int n_paths, double* out);

/* Step 3 of HybridAStar::getValidClosePose (hybrid_a_star_lib.cpp:1604-1668):
 * the collision-free pose of smallest cost around center_pose = (x, y, yaw) in metres / rad, among
 * the 64 x 64 x 114 offsets (gm_res_ steps within +-5 m, yaw_res_coll_ steps within +-3 rad) whose
 * astar cell is one of near_cells (AStar::nodes_near_goal_ as indices y * astar_dim + x; steps 1-2,
 * the heuristic sweep and the closest node, are ga_build_heuristic and host code).  cost = squared
 * offset + 20 |yaw offset| + getProxOfCorners; equal costs resolve to the last candidate in the
 * reference's loop order, as its `cost > min_cost` pruning does.  found = 0 (out_pose untouched)
 * when no candidate qualifies -- the reference then returns the ego pose.  Needs
 * ga_set_collision_params and ga_set_prox_params.  The proximity term carries the device's sin / cos
 * (1e-15 relative), which can only change the winner between candidates whose costs differ by less. */
GA_API int ga_valid_close_pose(ga_ctx* ctx, const double* center_pose, const int* near_cells, int n_near, double* out_pose,
                               int* found);

/* ga_valid_close_pose for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  Same
 * arguments and the same result as ga_valid_close_pose, bit for bit; the call performs no cudaMalloc / cudaMallocHost /
 * cudaFree: its scratch (cell list, membership map, result words) is a static device array of
 * GA_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit, and the candidates are evaluated
 * twice instead of being stored.  Limits: n_near <= GA_NOALLOC_MAX_NEAR and an astar grid of at most
 * GA_NOALLOC_MAX_ASTAR_DIM cells per side (patch_dim <= 4096), else GA_ERR_ARG with a message in ga_last_error.
 * All calls on a device share the array: calls that overlap on the GPU (two contexts, two streams, two threads) race;
 * the array keeps the last call's contents and is cleared per call where it must start clean. */
#define GA_NOALLOC_MAX_NEAR 65536
#define GA_NOALLOC_MAX_ASTAR_DIM 1024
#define GA_NOALLOC_SCRATCH_BYTES (256 + 4 * GA_NOALLOC_MAX_NEAR + GA_NOALLOC_MAX_ASTAR_DIM * GA_NOALLOC_MAX_ASTAR_DIM)
GA_API int ga_valid_close_pose_noalloc(ga_ctx* ctx, const double* center_pose, const int* near_cells, int n_near,
                                       double* out_pose, int* found);

/* ---- heuristic grid ------------------------------------------------------------------------- */

typedef struct
{
  double astar_movement_cost;      /* params.yml astar_movement_cost_ (movement map fill) */
  double astar_prox_cost;          /* params.yml astar_prox_cost_ */
  double astar_lane_movement_cost; /* params.yml astar_lane_movement_cost_ (REFORDER only) */
  double unknown_cost_w;           /* config.yml ASTAR_UNKNOWN_COST */
  /* potential field, a_star.cpp:486-508 */
  double alpha, do_max, do_min;
  double motion_res_min, motion_res_max, dist_val_min, dist_val_max;
  int vor_dim;                     /* VOR_DIM = 200, a_star.hpp:36 */
  int dt_variant;                  /* ga_dt_variant */
} ga_astar_params;

typedef enum { GA_DT_FIXEDPOINT = 0, GA_DT_IPPFLOAT = 1 } ga_dt_variant;

GA_API int ga_set_astar_params(ga_ctx* ctx, const ga_astar_params* p);

/* AStar::calcDistanceHeuristic (a_star.cpp:110-295).  slot 0 = closed_set_guidance_
 * (for_path = false), 1 = closed_set_path_.  Computes the least fixed point of the reference's
 * edge relaxation  g(n) = min_p ((g(p) + len_i * mc(n)) + prox(p) * w_prox) + unk(p)  over the
 * 8-connected grid from the goal, in fp64 with the reference's operand order, plus parent,
 * cost_dist and is_unknown planes.  The reference's pop order, early exit and start cell only
 * decide WHICH reachable cells it closes and, at a few cells, a suboptimal cost;
 * the fixed point needs no start cell.  n_closed (may be NULL; forces a stream sync otherwise)
 * receives the number of reachable cells. */
GA_API int ga_build_heuristic(ga_ctx* ctx, int slot, int goal_x, int goal_y, int* n_closed);

/* AStar::calcDistanceHeuristic(goal_pos, start_pos, for_path, get_only_near) with the reference's full
 * signature (a_star.cpp:110-113) and the two config.yml switches it reads (HEURISTIC_EARLY_EXIT,
 * MAX_EXTRA_NODES_ASTAR, a_star.cpp:31-32).
 *   flags = 0 (GA_HEUR_FIXEDPOINT): the GPU fixed point, exactly ga_build_heuristic; start / early exit /
 *           max_extra_nodes do not apply (every reachable cell is closed);
 *   GA_HEUR_REFORDER: the reference's own closing order -- closed-set membership under early exit, cost_,
 *           cost_dist_, parent_index_, is_unknown_ equal the reference's bit for bit (needed for identical
 *           Hybrid A* node sequences).  The order is one serial chain of heap operations and is
 *           replayed on the host (csrc/ga_heur_reforder.cu); inputs come from and results return to the device.
 *   GA_HEUR_EARLY_EXIT / GA_HEUR_ONLY_NEAR: HEURISTIC_EARLY_EXIT / get_only_near (REFORDER only).
 * ga_heur_near_nodes = AStar::nodes_near_goal_ (a_star.cpp:182-193) of the last build of `slot`: closed cells
 * closer than 10 cells to (start_x, start_y), as ids y * astar_dim + x with their distances; in closing order
 * after a REFORDER build with GA_HEUR_ONLY_NEAR for the same start, in row-major order otherwise (derived from
 * the slot's planes).  ga_heur_closing_order = the cells of the last REFORDER build in the order they were
 * closed (the reference does not expose it; tests and diagnostics do). */
enum { GA_HEUR_FIXEDPOINT = 0, GA_HEUR_REFORDER = 1, GA_HEUR_EARLY_EXIT = 2, GA_HEUR_ONLY_NEAR = 4 };
typedef struct
{
  int goal_x, goal_y;
  int start_x, start_y;
  unsigned flags;
  unsigned max_extra_nodes; /* config.yml MAX_EXTRA_NODES_ASTAR */
} ga_heur_query;
GA_API int ga_build_heuristic_ex(ga_ctx* ctx, int slot, const ga_heur_query* q, int* n_closed);
GA_API int ga_heur_near_nodes(ga_ctx* ctx, int slot, int start_x, int start_y, int* ids, double* dist, int cap, int* n);
GA_API int ga_heur_closing_order(ga_ctx* ctx, int* ids, int cap, int* n);
/* the GA_HEUR_REFORDER replay alone, on caller-owned HOST planes (dim x dim each; no context, no device): for
 * callers that keep host mirrors of astar_grid_ / movement_cost_map_ / h_prox_arr_ anyway.  order (may be NULL)
 * receives the closing order (up to dim * dim ids). */
GA_API int ga_heur_reforder_host(const ga_astar_params* ap, int dim, const uint8_t* grid, const uint8_t* fence,
                                 const double* movement_cost, const double* h_prox, const ga_heur_query* q, double* cost,
                                 double* cost_dist, int32_t* parent, uint8_t* flags, int32_t* order, int* n_closed);

/* The same for n independent queries at once (BASELINE config 5: batches of start/goal queries,
 * one heuristic grid per query; the reference needs one process per planner for this because all
 * of AStar's state is static, a_star.hpp:65-134).  ctxs[i] = the i-th query's context (all on one
 * device, each at most once), goal_xy[2i], goal_xy[2i+1] its goal cell, n_closed[i] (array of n,
 * may be NULL; forces a sync otherwise) its reachable-cell count.  All grids are relaxed by ONE
 * kernel chain on ctxs[0]'s stream -- their wavefronts share the worker warps of the GPU -- and
 * every other context's stream is ordered before and after it, so later calls on any of the
 * contexts see the finished planes.  Errors are reported on ctxs[0]. */
GA_API int ga_build_heuristic_batch(ga_ctx* const* ctxs, int n, int slot, const int* goal_xy, int* n_closed);

/* AStar::calcVoronoiPotentialField (a_star.cpp:440-479) around ego (astar cells): two chamfer
 * distance transforms (createDistanceTransform :426-433) + setCostmapValue (:486-508).  The
 * Fortune sweep (deps_lib/VoronoiDiagramGenerator.cpp, serial fp64 geometry) stays on the host:
 * vor_xy = the n_vor edge end points it marks, window coordinates.  The host asks for the OCC
 * cells of the window with ga_potential_sites first.  Returns GA_OK; *had_occ (may be NULL)
 * tells whether the window held an obstacle (0 = early return of a_star.cpp:462-465). */
GA_API int ga_potential_sites(ga_ctx* ctx, int ego_x, int ego_y, int* xy, int cap, int* n_sites);
GA_API int ga_build_potential(ga_ctx* ctx, int ego_x, int ego_y, const int* vor_xy, int n_vor,
                       int* had_occ);

/* ---- data movement -------------------------------------------------------------------------- */

/* whole-grid host <-> device copies (bytes must equal the grid's size) */
GA_API int ga_upload(ga_ctx* ctx, int grid_id, const void* src, size_t bytes);
GA_API int ga_download(ga_ctx* ctx, int grid_id, void* dst, size_t bytes);
/* ga_download without the wait: dst must be pinned host memory and is valid after ga_sync(ctx).
 * With one context per planner this keeps the copies of one planner overlapped with the kernels of
 * another (bench.py's pipelined e2e leg). */
GA_API int ga_download_async(ga_ctx* ctx, int grid_id, void* dst, size_t bytes);
/* a w x h window at (x0, y0) of an astar_dim^2 grid into a dense w x h host array (e.g. the 200 x 200 window of
 * h_prox_arr_ / motion_res_map_ that calcVoronoiPotentialField just rewrote, a_star.cpp:440-479) */
GA_API int ga_download_window(ga_ctx* ctx, int grid_id, int x0, int y0, int w, int h, void* dst);
GA_API int ga_grid_bytes(const ga_ctx* ctx, int grid_id, size_t* bytes);
/* device address of a grid (for zero-copy consumers such as the bench's resident-input leg) */
GA_API int ga_device_ptr(ga_ctx* ctx, int grid_id, void** dptr);
GA_API int ga_patch_dim(const ga_ctx* ctx);
GA_API int ga_astar_dim(const ga_ctx* ctx);

/* ---- measurement ---------------------------------------------------------------------------- */

/* kernels this library launches; ga_prof_* brackets each launch with CUDA events on the
 * context's stream when profiling is on (used by bench.py for the live roofline figure) */
typedef enum
{
  GA_K_PACK = 0,      /* unused (the pack step lives inside the fused kernel since round 1) */
  GA_K_DILATE = 1,    /* fused threshold + disk dilation + expand + 4x4 pool */
  GA_K_HEUR_INIT = 2,
  GA_K_HEUR_RELAX = 3,
  GA_K_HEUR_FINISH = 4,
  GA_K_COLLISION = 5,
  GA_K_EXPAND = 6,
  GA_K_DT = 7,
  GA_K_POTENTIAL = 8,
  GA_K_CARTOGRAPH = 9, /* window merge / patch-to-patch copy of the accumulated map */
  GA_K_RAYTRACE = 10,  /* simulated lidar ray-cast */
  GA_K_PROX = 11,      /* corner proximity of sampled poses / paths */
  GA_K_COUNT = 12
} ga_kernel_id;

GA_API int ga_prof_enable(ga_ctx* ctx, int on);
GA_API int ga_prof_reset(ga_ctx* ctx);
/* total device milliseconds and launches of kernel k since the last reset (syncs the stream) */
GA_API int ga_prof_get(ga_ctx* ctx, int k, double* ms_total, long long* launches);
/* work counters of the last ga_build_heuristic on this context: out7 = { tile visits, inner
 * line sweeps summed over visits, visits that improved a cell, SM cycles summed over
 * visits spent waiting for a tile / staging / sweeping / publishing } (syncs the stream) */
GA_API int ga_heur_stats(ga_ctx* ctx, long long* out7);
/* scheduling of the sweep kernel, k7 = { worker warps per SM (1..8; single builds group them: four warps per tile), idle back-off ns, early publications, lock-step
 * rounds, in-place halo refresh, linger ns, stage-ahead (+ 4: single builds with one warp per tile, + 8: with two) }; -1 = chosen per launch.  Every setting yields the same
 * fixed point (tests/test_gpu_parity.py::test_heuristic_schedules_agree); this exists for that test and for tuning. */
GA_API int ga_set_heur_knobs(ga_ctx* ctx, const int* k7);
/* how ga_build_safety stages a tile's raw bytes when the patch side is a multiple of 16: 1 (default) = one bulk tensor
 * copy per tile (TMA), 0 = per-thread cp.async.  Same result either way (tests/test_gpu_parity.py::test_safety_staging_modes);
 * this exists for that test and for measurement. */
GA_API int ga_set_safety_staging(ga_ctx* ctx, int use_tma);
/* launches of ALL library kernels since ga_create / ga_prof_reset (always counted) */
GA_API long long ga_launch_count(const ga_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
