// ga_internal.cuh -- context layout and launch helpers shared by the libgridaccel.so sources.
// Product code: nothing here (or in any file of csrc/) includes, links or calls oracle/.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "gridaccel.h"

#define GA_MAX_DISK_R 16   // 32 + 2r bits of one plane row must fit a 64-bit window
#define GA_MAX_DISKS 16
#define GA_MAX_STEER 31
#define GA_MAX_RA 16       // rear-axis deltas per node

struct ga_se  // structuring element, passed to kernels by value
{
  int r;                 // radius in cells
  int n_var;             // distinct half widths
  int var_w[GA_MAX_DISK_R + 1];        // half width of variant v
  int row_var[2 * GA_MAX_DISK_R + 1];  // variant index of SE row dy + r
  // the same element seen by columns: a cell at horizontal distance w is covered by the rows |dy| <= hmax(w)
  // (half widths shrink monotonically with |dy|); levels = the distinct hmax values, ascending
  int n_lvl;
  int lvl_h[GA_MAX_DISK_R + 1];     // hmax of level l
  int lvl_of_w[GA_MAX_DISK_R + 1];  // level whose hmax belongs to distance w (w = 0 .. r)
};

struct ga_prof_rec
{
  int k;
  cudaEvent_t e0, e1;
};

struct ga_ctx
{
  int device = 0;
  cudaStream_t stream = nullptr;
  char err[512] = { 0 };
  int sm_count = 148;

  // ---- parameters
  bool have_cp = false, have_ep = false, have_ap = false, have_pp = false;
  ga_collision_params cp{};
  ga_expand_params ep{};
  ga_astar_params ap{};
  ga_prox_params pp{};
  ga_se se{};
  int nb_yaws = 0;
  double yaw_res_coll_rad = 0.0;
  int n_ra = 0;                 // rear-axis deltas per node
  double ra_delta[GA_MAX_RA] = { 0 };
  double steer[GA_MAX_STEER] = { 0 };
  int* d_centers = nullptr;     // nb_yaws * nb_disks * 2

  // ---- grids
  int patch_dim = 0, astar_dim = 0;
  uint8_t* d_patch = nullptr;
  uint8_t* d_safety = nullptr;
  uint8_t* d_astar = nullptr;
  uint8_t* d_fence = nullptr;
  double* d_mc = nullptr;
  double* d_prox = nullptr;
  // tensor map of d_patch for the safety kernel's bulk copy (a CUtensorMap, kept opaque here), and what it was made for
  alignas(64) unsigned char tm_patch[128];
  const void* tm_ptr = nullptr;
  int tm_dim = 0, tm_rows = 0;
  bool tm_ok = false, use_tma = true;
  bool prox_zero = false;  // d_prox is all zero (set by ga_patch_reset, cleared by whatever may write the plane)
  double* d_mres = nullptr;
  float* d_odist = nullptr;
  float* d_vdist = nullptr;
  int vor_alloc = 0;
  double* d_hcost[2] = { nullptr, nullptr };
  double* d_hdist[2] = { nullptr, nullptr };
  int32_t* d_hparent[2] = { nullptr, nullptr };
  uint8_t* d_hflags[2] = { nullptr, nullptr };
  uint8_t* d_hpacked = nullptr;  // GA_GRID_HEUR_PACKED scratch (filled on demand by ga_download)
  uint32_t* d_hsteps[2] = { nullptr, nullptr };  // (straight | diagonal << 14) steps of every closed cell to the root, written by finish C
  uint8_t* d_hrec = nullptr;     // GA_GRID_HEUR_REC scratch: 12 B per cell (filled on demand by ga_download)
  bool rec_valid[2] = { false, false };  // the slot's planes come from the device fixed point (steps plane is theirs)
  int* d_rec_ovf = nullptr;      // set by the record kernel when a chain is longer than the record holds
  int* h_rec_ovf = nullptr;      // pinned copy, looked at by ga_download / ga_sync
  bool rec_check = false;

  // ---- accumulated map (Cartographing::patch_arr_ / temp_patch_)
  uint8_t* d_cart = nullptr;
  uint8_t* d_cart_prev = nullptr;
  int cart_dim = 0, cart_prev_dim = 0;

  // ---- simulated lidar (GridMapSim): precast lists, ground-truth window, result window
  int* d_rt_entries = nullptr;
  int* d_rt_offsets = nullptr;
  uint8_t* d_rt_gt = nullptr;
  uint8_t* d_pmap = nullptr;

  // ---- heuristic scratch
  int tiles_x = 0, n_tiles = 0;
  int* d_tile_active = nullptr;   // per-tile state word (idle / queued / running / dirty)
  int* d_hctl = nullptr;          // control words (ring cursors, counters)
  int* d_hring = nullptr;         // ring buffer of activated tiles
  int hring_cap = 0;              // power of two >= n_tiles
  uint8_t* d_hcls = nullptr;      // per-cell class byte of the current build (bit0 UNKNOWN, bit1 blocked)
  int32_t* d_jump = nullptr;      // finish scratch: block-local (terminal, steps) per cell
  int32_t* d_cnt = nullptr;       // finish scratch: (n_straight, n_diag) of terminal cells to the root
  int heur_epoch = 0;             // build counter (stamps the per-tile prox flags)
  int sw_warps = 8, sw_poll_ns = 20;  // sweep kernel knobs (ga_set_heur_knobs; 8 = SW_WARPS)
  int sw_early = -1, sw_round = -1, sw_refresh = -1, sw_linger = -1, sw_prefetch = -1;  // -1 = per launch (queue_knobs)
  int* d_bring = nullptr;         // batch ring (leading context of ga_build_heuristic_batch)
  int bring_cap = 0;
  void* d_bdesc = nullptr;        // batch descriptors + closed counters
  size_t bdesc_cap = 0;
  cudaEvent_t ev_batch = nullptr; // stream ordering between the contexts of a batch
  struct ga_reforder_state* reforder = nullptr;  // host side of GA_HEUR_REFORDER (ga_heur_reforder.cu)
  int near_slot = -1;             // slot whose nodes_near_goal_ list the last REFORDER build recorded
  int near_start[2] = { -1, -1 };
  bool heur_host_valid[2] = { false, false };

  // ---- staging (pinned host + device) for pose batches
  void* h_stage = nullptr;
  size_t h_stage_bytes = 0;
  void* d_stage = nullptr;
  size_t d_stage_bytes = 0;
  void* h_ll = nullptr;  // mapped pinned buffer of the low-latency ga_check_poses path (host / device view)
  void* d_ll = nullptr;
  unsigned ll_ticket = 0;
  void* expand_async = nullptr;  // ring of ga_expand_submit (ga_collision.cu)

  // ---- measurement
  bool prof_on = false;
  std::vector<ga_prof_rec> prof;
  std::vector<cudaEvent_t> ev_pool;
  long long launches = 0;
  long long k_launches[GA_K_COUNT] = { 0 };
};

int ga_fail(ga_ctx* ctx, int code, const char* fmt, ...);

#define GA_CUDA(ctx, call)                                                                        \
  do                                                                                              \
  {                                                                                               \
    cudaError_t e__ = (call);                                                                     \
    if (e__ != cudaSuccess)                                                                       \
      return ga_fail((ctx), GA_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,             \
                     cudaGetErrorString(e__));                                                    \
  } while (0)

#define GA_REQUIRE(ctx, cond, code, msg)                                                          \
  do                                                                                              \
  {                                                                                               \
    if (!(cond)) return ga_fail((ctx), (code), "%s: %s", __func__, (msg));                        \
  } while (0)

// Brackets one kernel launch: counts it and, when profiling, records events around it.
struct ga_launch_scope
{
  ga_ctx* c;
  int k;
  cudaEvent_t e1 = nullptr;
  ga_launch_scope(ga_ctx* ctx, int kid, int n_kernels = 1);
  ~ga_launch_scope();
};

int ga_stage_reserve(ga_ctx* ctx, size_t bytes);

// implemented per subsystem
int ga_heur_alloc(ga_ctx* ctx);
void ga_heur_free(ga_ctx* ctx);
int ga_heur_reforder_build(ga_ctx* ctx, int slot, const ga_heur_query* q, int* n_closed);
void ga_heur_reforder_near(const ga_ctx* ctx, const int32_t** ids, const double** dist, int* n);
void ga_heur_reforder_order(const ga_ctx* ctx, const int32_t** ids, int* n);
void ga_reforder_free(ga_ctx* ctx);
void ga_expand_async_free(ga_ctx* ctx);
void ga_cart_free(ga_ctx* ctx);
void ga_rt_free(ga_ctx* ctx);
