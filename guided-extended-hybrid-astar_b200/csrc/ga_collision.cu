// ga_collision.cu -- K6: batched disk-footprint collision checks and motion-primitive rollout.
//
// Replaces CollisionChecker::checkGrid / checkGridVariant / checkPose / getYawIdx /
// getPathCollisionIndex / checkPathCollision (collision_checking.cpp:333-452) and the
// collision-relevant part of HybridAStar::setNeighbors (hybrid_a_star_lib.cpp:492-534) with
// Vehicle::move / move_car_some_steps / turn_on_rear_axis (vehicle.cpp:34-122).
//
// Arithmetic notes: compiled with -fmad=false so every product and sum rounds separately, as the
// reference's baseline x86-64 build does.  fmod, division, sqrt, conversions are IEEE-exact on
// the device; sin/cos/tan are CUDA libm (<= 2 ulp) and may differ from glibc in the last bit, so
// ga_expand flags primitives whose poses come within 1e-9 cells of a discretisation boundary.
#include <climits>
#include <cmath>
#include <mutex>

#include "ga_prox.cuh"

#define GA_CON2GM (1.0 / 0.15625)   // grid_tf::con2gm_  (transforms.hpp:73-74, params.hpp:8)
#define GA_CON2STAR (1.0 / 0.625)   // grid_tf::con2star_ (transforms.hpp:76-77, params.hpp:7)
#define GA_PI 3.14159265358979323846

// util::constrainAngleZero2Pi (util1.cpp:8-15)
__device__ __forceinline__ double angle_0_2pi(double a)
{
  while (a < 0) a += 2 * GA_PI;
  return fmod(a, 2 * GA_PI);
}
// util::constrainAngleMinPIPlusPi (util1.cpp:17-24)
__device__ __forceinline__ double angle_mpi_pi(double a)
{
  while (a < 0) a += 2 * GA_PI;
  return fmod(a + GA_PI, 2.0 * GA_PI) - GA_PI;
}
// util::getAngleDiff (util1.cpp:26-34)
__device__ __forceinline__ double angle_diff(double a1, double a2)
{
  a2 = angle_mpi_pi(a2);
  a1 = angle_mpi_pi(a1);
  return angle_mpi_pi(a2 - a1);
}

struct coll_view
{
  const uint8_t* safety;
  const int* centers;
  int patch_dim;
  int nb_disks;
  int nb_yaws;
  double yaw_res_rad;
};

// checkPose (collision_checking.cpp:389-396) -> checkGrid(/Variant) (:344-387); 1 = valid.
// frac_margin (optional) receives the smallest distance of x*con2gm, y*con2gm, yaw/res to an
// integer boundary.
__device__ __forceinline__ int check_pose_dev(const coll_view& v, double x, double y, double yaw, int variant,
                                              double* frac_margin)
{
  const double fx = x * GA_CON2GM;
  const double fy = y * GA_CON2GM;
  const int gx = (int)fx;
  const int gy = (int)fy;
  // getYawIdx (:333-337)
  const double fa = angle_0_2pi(yaw) / v.yaw_res_rad;
  int yi = (int)fa;
  yi = yi < v.nb_yaws - 1 ? yi : v.nb_yaws - 1;
  if (frac_margin)
  {
    const double mx = fabs(fx - rint(fx)), my = fabs(fy - rint(fy)), ma = fabs(fa - rint(fa));
    *frac_margin = fmin(fmin(mx, my), ma);
  }
  const int* cen = v.centers + (size_t)yi * v.nb_disks * 2;
  for (int k = 0; k < v.nb_disks; ++k)
  {
    const int px = cen[2 * k] + gx;
    const int py = cen[2 * k + 1] + gy;
    if (px >= v.patch_dim || py >= v.patch_dim || px < 0 || py < 0)
    {
      if (variant) continue;
      return 0;
    }
    if (v.safety[(size_t)py * v.patch_dim + px] == GA_OCC) return 0;
  }
  return 1;
}

// The same check for a pose whose coordinates may be off by a rounding error (device libm against glibc in a
// rolled-out primitive): when x * con2gm, y * con2gm or yaw / res lies within eps of an integer the pose could fall
// into either neighbouring cell / yaw bin; every combination is evaluated and *ambiguous is set when they disagree.
__device__ __forceinline__ int check_cell_dev(const coll_view& v, int gx, int gy, int yi, int variant)
{
  const int* cen = v.centers + (size_t)yi * v.nb_disks * 2;
  for (int k = 0; k < v.nb_disks; ++k)
  {
    const int px = cen[2 * k] + gx, py = cen[2 * k + 1] + gy;
    if (px >= v.patch_dim || py >= v.patch_dim || px < 0 || py < 0)
    {
      if (variant) continue;
      return 0;
    }
    if (v.safety[(size_t)py * v.patch_dim + px] == GA_OCC) return 0;
  }
  return 1;
}

__device__ __forceinline__ int check_pose_robust(const coll_view& v, double x, double y, double yaw, int variant, double eps,
                                                 int* ambiguous)
{
  const double fx = x * GA_CON2GM, fy = y * GA_CON2GM;
  const double fa = angle_0_2pi(yaw) / v.yaw_res_rad;
  const int gx = (int)fx, gy = (int)fy;
  int yi = (int)fa;
  yi = yi < v.nb_yaws - 1 ? yi : v.nb_yaws - 1;
  const int r0 = check_cell_dev(v, gx, gy, yi, variant);
  const bool nx = fabs(fx - rint(fx)) < eps, ny = fabs(fy - rint(fy)) < eps, na = fabs(fa - rint(fa)) < eps;
  if (nx | ny | na)
  {
    const int gx2 = (int)rint(fx) == gx ? gx - 1 : (int)rint(fx);
    const int gy2 = (int)rint(fy) == gy ? gy - 1 : (int)rint(fy);
    int ya = (int)rint(fa) == (int)fa ? (int)fa - 1 : (int)rint(fa);
    ya = ya < 0 ? v.nb_yaws - 1 : (ya > v.nb_yaws - 1 ? v.nb_yaws - 1 : ya);
    for (int m = 1; m < 8; ++m)
    {
      if (((m & 1) && !nx) || ((m & 2) && !ny) || ((m & 4) && !na)) continue;
      if (check_cell_dev(v, (m & 1) ? gx2 : gx, (m & 2) ? gy2 : gy, (m & 4) ? ya : yi, variant) != r0) *ambiguous = 1;
    }
  }
  return r0;
}

// ---- ga_check_poses ---------------------------------------------------------------------------

__global__ void __launch_bounds__(256)
check_poses_kernel(coll_view v, const double* __restrict__ x, const double* __restrict__ y,
                   const double* __restrict__ yaw, int n_poses, const int* __restrict__ seg_off, int n_seg,
                   int variant, int* __restrict__ first_hit)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_poses) return;
  if (check_pose_dev(v, x[i], y[i], yaw[i], variant, nullptr)) return;
  // segment of pose i: largest s with seg_off[s] <= i
  int lo = 0, hi = n_seg;
  while (hi - lo > 1)
  {
    const int mid = (lo + hi) >> 1;
    if (seg_off[mid] <= i) lo = mid; else hi = mid;
  }
  atomicMin(&first_hit[lo], i - seg_off[lo]);
}

__global__ void fill_i32_kernel(int* p, int n, int v)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

__global__ void fix_first_hit_kernel(int* p, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && p[i] == INT_MAX) p[i] = -1;
}

// ---- low-latency path of ga_check_poses ----------------------------------------------------------
// The serial search asks about ONE expansion at a time (26-32 primitives, <= a few hundred poses) and waits for the
// answer, so a call costs latency, not throughput.  For such batches
//   - the poses travel INSIDE the kernel's parameter block (up to 25 KB, sm_70+ accepts 32 KB): they reach the SM with
//     the launch itself, no copy-engine operation and no PCIe read round trip;
//   - one CTA does everything: one warp per segment (looping), lanes take 32 poses at a time in list order, the first
//     ballot with a set bit gives the first colliding index -- no init / fix-up kernels, no atomics;
//   - the answers are written to host memory the device maps, followed by a ticket word; the host spins on that word
//     instead of entering the driver (cudaStreamSynchronize costs several microseconds by itself).
template <int CAP>
struct ll_params
{
  double x[CAP], y[CAP], yaw[CAP];
  unsigned short seg[CAP / 4 + 2];
  int n_seg, variant;
  unsigned ticket;
};
struct ll_host  // mapped pinned memory
{
  int hit[1024];
  volatile unsigned ticket;
};

template <int CAP>
__global__ void __launch_bounds__(256) check_fan_kernel(coll_view v, const __grid_constant__ ll_params<CAP> p, ll_host* out)
{
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int seg = warp; seg < p.n_seg; seg += 8)
  {
    const int s0 = p.seg[seg], s1 = p.seg[seg + 1];
    int first = -1;
    for (int base = s0; base < s1 && first < 0; base += 32)
    {
      const int i = base + lane;
      int bad = 0;
      if (i < s1) bad = !check_pose_dev(v, p.x[i], p.y[i], p.yaw[i], p.variant, nullptr);
      const unsigned m = __ballot_sync(0xFFFFFFFFu, bad);
      if (m) first = base - s0 + __ffs(m) - 1;
    }
    if (lane == 0) out->hit[seg] = first;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0)
  {
    __threadfence_system();
    out->ticket = p.ticket;
  }
}

template <int CAP>
static int check_fan_launch(ga_ctx* c, const double* x, const double* y, const double* yaw, int n_poses, const int* seg_off,
                            int n_seg, int variant, int* first_hit);

static coll_view make_view(const ga_ctx* c)
{
  coll_view v;
  v.safety = c->d_safety;
  v.centers = c->d_centers;
  v.patch_dim = c->patch_dim;
  v.nb_disks = c->cp.nb_disks;
  v.nb_yaws = c->nb_yaws;
  v.yaw_res_rad = c->yaw_res_coll_rad;
  return v;
}

template <int CAP>
static int check_fan_launch(ga_ctx* c, const double* x, const double* y, const double* yaw, int n_poses, const int* seg_off,
                            int n_seg, int variant, int* first_hit)
{
  if (!c->h_ll)
  {
    GA_CUDA(c, cudaHostAlloc(&c->h_ll, sizeof(ll_host), cudaHostAllocMapped));
    GA_CUDA(c, cudaHostGetDevicePointer(&c->d_ll, c->h_ll, 0));
    static_cast<ll_host*>(c->h_ll)->ticket = 0;
  }
  static_assert(sizeof(ll_params<CAP>) + sizeof(coll_view) + 16 < 32764, "kernel parameter block too large");
  ll_params<CAP> p;
  memcpy(p.x, x, sizeof(double) * (size_t)n_poses);
  memcpy(p.y, y, sizeof(double) * (size_t)n_poses);
  memcpy(p.yaw, yaw, sizeof(double) * (size_t)n_poses);
  for (int s = 0; s <= n_seg; ++s) p.seg[s] = (unsigned short)seg_off[s];
  p.n_seg = n_seg;
  p.variant = variant;
  p.ticket = ++c->ll_ticket;
  if (p.ticket == 0) p.ticket = ++c->ll_ticket;
  ll_host* hb = static_cast<ll_host*>(c->h_ll);
  {
    ga_launch_scope ls(c, GA_K_COLLISION);
    check_fan_kernel<CAP><<<1, 256, 0, c->stream>>>(make_view(c), p, static_cast<ll_host*>(c->d_ll));
  }
  GA_CUDA(c, cudaGetLastError());
  // wait for the ticket word; look at the stream now and then so that a failed launch cannot hang the caller
  for (unsigned long long spins = 0; hb->ticket != p.ticket; ++spins)
  {
    if ((spins & 0xFFFFF) == 0xFFFFF)
    {
      const cudaError_t q = cudaStreamQuery(c->stream);
      if (q != cudaErrorNotReady && hb->ticket != p.ticket)
      {
        if (q == cudaSuccess) break;  // finished: the store is visible after the stream is idle
        return ga_fail(c, GA_ERR_CUDA, "%s: %s", __func__, cudaGetErrorString(q));
      }
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  memcpy(first_hit, hb->hit, sizeof(int) * (size_t)n_seg);
  return GA_OK;
}

extern "C" int ga_check_poses(ga_ctx* c, const double* x, const double* y, const double* yaw, int n_poses,
                              const int* seg_off, int n_seg, int variant, int* first_hit)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, n_poses >= 0 && n_seg >= 0 && first_hit && seg_off, GA_ERR_ARG, "bad arguments");
  GA_REQUIRE(c, n_poses == 0 || (x && y && yaw), GA_ERR_ARG, "null pose arrays");
  if (n_seg == 0) return GA_OK;
  GA_REQUIRE(c, seg_off[0] == 0 && seg_off[n_seg] == n_poses, GA_ERR_ARG, "seg_off must span [0, n_poses]");
  GA_CUDA(c, cudaSetDevice(c->device));
  if (n_poses <= 256 && n_seg <= 64) return check_fan_launch<256>(c, x, y, yaw, n_poses, seg_off, n_seg, variant, first_hit);
  if (n_poses <= 1024 && n_seg <= 256) return check_fan_launch<1024>(c, x, y, yaw, n_poses, seg_off, n_seg, variant, first_hit);
  // staging layout: x | y | yaw | seg_off | first_hit
  const size_t b_pose = sizeof(double) * (size_t)n_poses;
  const size_t b_seg = sizeof(int) * (size_t)(n_seg + 1);
  const size_t b_hit = sizeof(int) * (size_t)n_seg;
  const size_t total = 3 * b_pose + b_seg + b_hit + 64;
  int rc = ga_stage_reserve(c, total);
  if (rc) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));  // staging buffer reuse
  char* h = (char*)c->h_stage;
  char* d = (char*)c->d_stage;
  memcpy(h, x, b_pose);
  memcpy(h + b_pose, y, b_pose);
  memcpy(h + 2 * b_pose, yaw, b_pose);
  memcpy(h + 3 * b_pose, seg_off, b_seg);
  const size_t off_hit = (3 * b_pose + b_seg + 15) & ~(size_t)15;
  GA_CUDA(c, cudaMemcpyAsync(d, h, 3 * b_pose + b_seg, cudaMemcpyHostToDevice, c->stream));
  int* d_hit = (int*)(d + off_hit);
  fill_i32_kernel<<<(n_seg + 255) / 256, 256, 0, c->stream>>>(d_hit, n_seg, INT_MAX);
  if (n_poses > 0)
  {
    ga_launch_scope ls(c, GA_K_COLLISION);
    check_poses_kernel<<<(n_poses + 255) / 256, 256, 0, c->stream>>>(
        make_view(c), (const double*)d, (const double*)(d + b_pose), (const double*)(d + 2 * b_pose), n_poses,
        (const int*)(d + 3 * b_pose), n_seg, variant, d_hit);
  }
  fix_first_hit_kernel<<<(n_seg + 255) / 256, 256, 0, c->stream>>>(d_hit, n_seg);
  GA_CUDA(c, cudaGetLastError());
  GA_CUDA(c, cudaMemcpyAsync(h + off_hit, d_hit, b_hit, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  memcpy(first_hit, h + off_hit, b_hit);
  return GA_OK;
}

// ---- ga_valid_close_pose ------------------------------------------------------------------------
// Step 3 of HybridAStar::getValidClosePose (hybrid_a_star_lib.cpp:1604-1668): 64 x 64 x 114 candidate
// poses around the centre, cost = squared offset + 20 |yaw offset| + corner proximity, a candidate
// counts when its astar cell is one of nodes_near_goal_ and its footprint is collision free.  The
// reference walks them in order and takes a candidate when its cost is not ABOVE the best so far:
// the result is the smallest cost, and among equal costs the LAST candidate in loop order.  Here:
// one thread per candidate, a 64-bit atomic minimum over the cost bits (costs are >= 0, so the IEEE
// order is the integer order), then an atomic maximum of the loop index over the candidates that hold it.

#define CP_DXY_MAX 5.0        // dxy_max, :1613
#define CP_ANGLE_W 20.0       // angle_weight, :1609
#define CP_GM_RES 0.15625     // HybridAStar::gm_res_ = GM_RES (:34)

struct close_pose_view
{
  coll_view cv;
  prox_args pa;
  const uint8_t* member;  // astar_dim^2 bytes: 1 = cell is in nodes_near_goal_
  double cx, cy, cyaw, phi_res;
  int nxy, nphi;          // dxy_max_idx, dphi_max_idx: loops run [-n, n)
};

__global__ void fill_u8_scatter_kernel(uint8_t* m, const int* __restrict__ cells, int n, int cells_total)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && cells[i] >= 0 && cells[i] < cells_total) m[cells[i]] = 1;
}

__device__ __forceinline__ bool close_pose_eval(const close_pose_view& v, int idx, unsigned long long* key)
{
  const int dphii = idx % (2 * v.nphi) - v.nphi;
  const int r = idx / (2 * v.nphi);
  const int dyi = r % (2 * v.nxy) - v.nxy, dxi = r / (2 * v.nxy) - v.nxy;
  const double x_diff = dxi * CP_GM_RES, y_diff = dyi * CP_GM_RES, phi_diff = dphii * v.phi_res;
  const double dist2 = x_diff * x_diff + y_diff * y_diff;
  const double px = v.cx + x_diff, py = v.cy + y_diff, pyaw = v.cyaw + phi_diff;
  const double prox_cost = prox_of_corners(px, py, pyaw, v.pa);
  const double cost = dist2 * 1.0 + fabs(phi_diff) * CP_ANGLE_W + prox_cost * 1.0;
  *key = (unsigned long long)__double_as_longlong(cost);
  // grid_tf::utm2astar(pose.getPoint()).toInt() (std::round) -> AStar::calcIndex(size_t, size_t): a
  // negative coordinate wraps there, which only matters when the sum lands on a real cell again
  const long long ix = (long long)round(px * v.pa.p.con2star), iy = (long long)round(py * v.pa.p.con2star);
  const long long index = iy * v.pa.dim + ix;
  if (index < 0 || index >= (long long)v.pa.dim * v.pa.dim || !v.member[index]) return false;
  return check_pose_dev(v.cv, px, py, pyaw, 0, nullptr) != 0;
}

__global__ void __launch_bounds__(256) close_pose_min_kernel(close_pose_view v, int n, unsigned long long* __restrict__ keys,
                                                             unsigned long long* __restrict__ best)
{
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long key = ~0ull;
  if (idx < n)
  {
    unsigned long long k;
    if (close_pose_eval(v, idx, &k)) key = k;
    keys[idx] = key;
  }
  // warp minimum first: one atomic per warp
  for (int o = 16; o > 0; o >>= 1)
  {
    const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, key, o);
    key = other < key ? other : key;
  }
  if ((threadIdx.x & 31) == 0 && key != ~0ull) atomicMin(best, key);
}

__global__ void __launch_bounds__(256) close_pose_arg_kernel(int n, const unsigned long long* __restrict__ keys,
                                                             const unsigned long long* __restrict__ best, int* __restrict__ arg)
{
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long b = *best;
  if (idx < n && b != ~0ull && keys[idx] == b) atomicMax(arg, idx);
}

extern "C" int ga_valid_close_pose(ga_ctx* c, const double* center_pose, const int* near_cells, int n_near, double* out_pose,
                                   int* found)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp, GA_ERR_STATE, "no patch / collision params");
  GA_REQUIRE(c, c->have_pp, GA_ERR_STATE, "no prox params (call ga_set_prox_params)");
  GA_REQUIRE(c, center_pose && out_pose && found && n_near >= 0 && (n_near == 0 || near_cells), GA_ERR_ARG, "bad arguments");
  GA_CUDA(c, cudaSetDevice(c->device));
  close_pose_view v;
  v.cv = make_view(c);
  v.pa = prox_args{ c->d_prox, c->astar_dim, c->pp };
  v.cx = center_pose[0];
  v.cy = center_pose[1];
  v.cyaw = center_pose[2];
  v.phi_res = c->yaw_res_coll_rad;  // yaw_res_coll_ * TO_RAD, :1616
  v.nxy = static_cast<int>(std::floor(CP_DXY_MAX) / CP_GM_RES);      // :1623
  v.nphi = static_cast<int>(std::floor(GA_PI) / v.phi_res);          // :1624
  const int n = 2 * v.nxy * 2 * v.nxy * 2 * v.nphi;
  *found = 0;
  if (n <= 0 || n_near == 0) return GA_OK;
  const size_t cells = (size_t)c->astar_dim * c->astar_dim;
  // staging layout: near_cells | member map | keys | best | arg
  const size_t b_near = (sizeof(int) * (size_t)n_near + 255) & ~(size_t)255, b_mem = (cells + 255) & ~(size_t)255;
  const size_t b_keys = sizeof(unsigned long long) * (size_t)n;
  const int rc = ga_stage_reserve(c, b_near + b_mem + b_keys + 64);
  if (rc) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));  // staging buffer reuse
  char* h = (char*)c->h_stage;
  char* d = (char*)c->d_stage;
  memcpy(h, near_cells, sizeof(int) * (size_t)n_near);
  GA_CUDA(c, cudaMemcpyAsync(d, h, sizeof(int) * (size_t)n_near, cudaMemcpyHostToDevice, c->stream));
  uint8_t* d_mem = (uint8_t*)(d + b_near);
  unsigned long long* d_keys = (unsigned long long*)(d + b_near + b_mem);
  unsigned long long* d_best = d_keys + n;
  int* d_arg = (int*)(d_best + 1);
  GA_CUDA(c, cudaMemsetAsync(d_mem, 0, cells, c->stream));
  GA_CUDA(c, cudaMemsetAsync(d_best, 0xFF, 16, c->stream));  // best = all ones, arg = -1
  v.member = d_mem;
  {
    ga_launch_scope ls(c, GA_K_COLLISION, 3);
    fill_u8_scatter_kernel<<<(n_near + 255) / 256, 256, 0, c->stream>>>(d_mem, (const int*)d, n_near, (int)cells);
    close_pose_min_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(v, n, d_keys, d_best);
    close_pose_arg_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(n, d_keys, d_best, d_arg);
  }
  GA_CUDA(c, cudaGetLastError());
  int* h_arg = (int*)(h + b_near + b_mem + b_keys + 8);
  GA_CUDA(c, cudaMemcpyAsync(h_arg, d_arg, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  const int idx = *h_arg;
  if (idx < 0) return GA_OK;
  // the winning pose, with the loop's own arithmetic (:1632-1640)
  const int dphii = idx % (2 * v.nphi) - v.nphi, r = idx / (2 * v.nphi);
  const int dyi = r % (2 * v.nxy) - v.nxy, dxi = r / (2 * v.nxy) - v.nxy;
  out_pose[0] = v.cx + dxi * CP_GM_RES;
  out_pose[1] = v.cy + dyi * CP_GM_RES;
  out_pose[2] = v.cyaw + dphii * v.phi_res;
  *found = 1;
  return GA_OK;
}

// ---- ga_valid_close_pose_noalloc ------------------------------------------------------------------
// The same search for callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): no counterpart in the reference.  ga_valid_close_pose keeps
// its scratch (cell list, membership map, one key per candidate, the two result words) in the context's staging buffer,
// which grows by cudaMalloc / cudaMallocHost; here the scratch is a file-scope device array of fixed size (every GPU
// that loads the library holds GA_NOALLOC_SCRATCH_BYTES until exit) and the per-candidate keys are not stored at all:
// the second pass evaluates the candidates again (same instructions, same bits) instead of reading them back.
//   layout: [0, 256)                       best key (u64), arg (i32)
//           [256, +4 * MAX_NEAR)           near_cells as uploaded
//           [.., +MAX_ASTAR_DIM^2)         membership map, 1 byte per astar cell
// All calls, from every context and stream of a process, share the array on a device: calls that overlap on the GPU
// race.  It keeps the last call's contents; what must start clean (the map, the two result words) is cleared per call.
#define NA_OFF_NEAR 256
#define NA_OFF_MEMBER (NA_OFF_NEAR + 4 * GA_NOALLOC_MAX_NEAR)
static_assert(NA_OFF_MEMBER + GA_NOALLOC_MAX_ASTAR_DIM * GA_NOALLOC_MAX_ASTAR_DIM == GA_NOALLOC_SCRATCH_BYTES, "scratch layout");
static __device__ __align__(256) unsigned char g_noalloc_scratch[GA_NOALLOC_SCRATCH_BYTES];

__global__ void __launch_bounds__(256) close_pose_min_nokeys_kernel(close_pose_view v, int n, unsigned long long* __restrict__ best)
{
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long key = ~0ull;
  if (idx < n)
  {
    unsigned long long k;
    if (close_pose_eval(v, idx, &k)) key = k;
  }
  for (int o = 16; o > 0; o >>= 1)
  {
    const unsigned long long other = __shfl_xor_sync(0xFFFFFFFFu, key, o);
    key = other < key ? other : key;
  }
  if ((threadIdx.x & 31) == 0 && key != ~0ull) atomicMin(best, key);
}

__global__ void __launch_bounds__(256) close_pose_arg_nokeys_kernel(close_pose_view v, int n, const unsigned long long* __restrict__ best,
                                                                    int* __restrict__ arg)
{
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned long long b = *best;
  if (idx >= n || b == ~0ull) return;
  unsigned long long k;
  if (close_pose_eval(v, idx, &k) && k == b) atomicMax(arg, idx);
}

// device address of the scratch array on `device` (which must be current).  In host code the array's name is a
// host-side stand-in, not a device address: it is resolved once per device and the pointer is what kernels get.
static int noalloc_scratch(ga_ctx* c, unsigned char** out)
{
  static std::mutex mu;
  static unsigned char* ptr[64] = { nullptr };
  GA_REQUIRE(c, c->device >= 0 && c->device < 64, GA_ERR_ARG, "device index out of range for the no-alloc scratch table");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[c->device])
  {
    void* p = nullptr;
    GA_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    ptr[c->device] = static_cast<unsigned char*>(p);
  }
  *out = ptr[c->device];
  return GA_OK;
}

extern "C" int ga_valid_close_pose_noalloc(ga_ctx* c, const double* center_pose, const int* near_cells, int n_near, double* out_pose,
                                           int* found)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp, GA_ERR_STATE, "no patch / collision params");
  GA_REQUIRE(c, c->have_pp, GA_ERR_STATE, "no prox params (call ga_set_prox_params)");
  GA_REQUIRE(c, center_pose && out_pose && found && n_near >= 0 && (n_near == 0 || near_cells), GA_ERR_ARG, "bad arguments");
  GA_REQUIRE(c, n_near <= GA_NOALLOC_MAX_NEAR, GA_ERR_ARG, "ga_valid_close_pose_noalloc: n_near exceeds GA_NOALLOC_MAX_NEAR (65536)");
  GA_REQUIRE(c, c->astar_dim <= GA_NOALLOC_MAX_ASTAR_DIM, GA_ERR_ARG,
             "ga_valid_close_pose_noalloc: astar grid larger than GA_NOALLOC_MAX_ASTAR_DIM (1024) cells per side");
  GA_CUDA(c, cudaSetDevice(c->device));
  unsigned char* scratch = nullptr;
  const int rc = noalloc_scratch(c, &scratch);
  if (rc) return rc;
  close_pose_view v;
  v.cv = make_view(c);
  v.pa = prox_args{ c->d_prox, c->astar_dim, c->pp };
  v.cx = center_pose[0];
  v.cy = center_pose[1];
  v.cyaw = center_pose[2];
  v.phi_res = c->yaw_res_coll_rad;
  v.nxy = static_cast<int>(std::floor(CP_DXY_MAX) / CP_GM_RES);
  v.nphi = static_cast<int>(std::floor(GA_PI) / v.phi_res);
  const int n = 2 * v.nxy * 2 * v.nxy * 2 * v.nphi;
  *found = 0;
  if (n <= 0 || n_near == 0) return GA_OK;
  const size_t cells = (size_t)c->astar_dim * c->astar_dim;
  unsigned long long* d_best = reinterpret_cast<unsigned long long*>(scratch);
  int* d_arg = reinterpret_cast<int*>(scratch + 8);
  int* d_near = reinterpret_cast<int*>(scratch + NA_OFF_NEAR);
  uint8_t* d_mem = scratch + NA_OFF_MEMBER;
  // pageable source: the driver has taken the bytes when the call returns
  GA_CUDA(c, cudaMemcpyAsync(d_near, near_cells, sizeof(int) * (size_t)n_near, cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemsetAsync(d_mem, 0, cells, c->stream));
  GA_CUDA(c, cudaMemsetAsync(d_best, 0xFF, 16, c->stream));  // best = all ones, arg = -1
  v.member = d_mem;
  {
    ga_launch_scope ls(c, GA_K_COLLISION, 3);
    fill_u8_scatter_kernel<<<(n_near + 255) / 256, 256, 0, c->stream>>>(d_mem, d_near, n_near, (int)cells);
    close_pose_min_nokeys_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(v, n, d_best);
    close_pose_arg_nokeys_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(v, n, d_best, d_arg);
  }
  GA_CUDA(c, cudaGetLastError());
  int idx = -1;
  GA_CUDA(c, cudaMemcpyAsync(&idx, d_arg, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  if (idx < 0) return GA_OK;
  const int dphii = idx % (2 * v.nphi) - v.nphi, r = idx / (2 * v.nphi);
  const int dyi = r % (2 * v.nxy) - v.nxy, dxi = r / (2 * v.nxy) - v.nxy;
  out_pose[0] = v.cx + dxi * CP_GM_RES;
  out_pose[1] = v.cy + dyi * CP_GM_RES;
  out_pose[2] = v.cyaw + dphii * v.phi_res;
  *found = 1;
  return GA_OK;
}

// ---- ga_expand ----------------------------------------------------------------------------------

struct expand_view
{
  double steer[GA_MAX_STEER];
  double ra_delta[GA_MAX_RA];
  double wheel_base, arc_l;
  double yaw_res_inv;        // a_star_yaw_res_inv_ = 1 / (YAW_RES * TO_RAD), hybrid_a_star_lib.cpp:38-40
  double yaw_res_coll_deg;   // yaw_res_coll_ handed to turn_on_rear_axis (:359)
  int nb_steer, n_ra, astar_dim;
};

#define GA_EDGE_EPS 1e-9

// one thread per (node, primitive); primitives in the reference's order: steer-major,
// direction +1 then -1 (direction_inputs_, hybrid_a_star_lib.hpp:50), then the rear-axis deltas
// successor epilogue (optional): the grid reads the serial search makes per successor, done where the grids live
struct expand_epilogue
{
  const double* hcost;   // closed-set cost_ plane of the chosen slot
  const double* hdist;   // cost_dist_ plane
  const uint8_t* hflags;
  const double* mc;      // movement_cost_map_
  const double* mres;    // motion_res_map_
  ga_succ_cost* out;     // per (node, primitive), nullptr = no epilogue
  double* parent_prox;   // per node
  prox_args pa;          // h_prox_arr_ + vehicle corners
};

// completion signal of an asynchronous launch: the last CTA to finish writes the ticket into mapped host memory
struct expand_done
{
  unsigned* counter;          // device memory, zero between launches
  volatile unsigned* flag;    // mapped host memory (nullptr: synchronous launch, nothing to signal)
  unsigned ticket;
};

__device__ __forceinline__ void expand_one(const coll_view& v, const expand_view& e, const expand_epilogue& ep,
                                           const ga_node* __restrict__ nodes, ga_succ* __restrict__ out, long long t);

__global__ void __launch_bounds__(128)
expand_kernel(coll_view v, expand_view e, expand_epilogue ep, const ga_node* __restrict__ nodes, int n_nodes,
              ga_succ* __restrict__ out, expand_done done)
{
  const int stride = 2 * e.nb_steer + e.n_ra;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t < (long long)n_nodes * stride) expand_one(v, e, ep, nodes, out, t);
  if (done.flag)
  {
    __threadfence_system();  // this thread's results are visible to the host before the CTA reports
    __syncthreads();
    if (threadIdx.x == 0)
    {
      if (atomicAdd(done.counter, 1u) == gridDim.x - 1)
      {
        *done.counter = 0;
        __threadfence_system();
        *done.flag = done.ticket;
      }
    }
  }
}

// kernel-parameter variant for the few nodes of one search step (the nodes travel with the launch)
#define GA_EXPAND_ASYNC_MAX 32
struct expand_nodes_param
{
  ga_node n[GA_EXPAND_ASYNC_MAX];
};
__global__ void __launch_bounds__(128)
expand_param_kernel(coll_view v, expand_view e, const __grid_constant__ expand_nodes_param nodes, int n_nodes,
                    ga_succ* __restrict__ out, uint8_t* __restrict__ flags, expand_done done)
{
  const int stride = 2 * e.nb_steer + e.n_ra;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  expand_epilogue ep{};
  if (t < (long long)n_nodes * stride)
  {
    expand_one(v, e, ep, nodes.n, out, t);
    // one byte per primitive for callers that only need the collision answer (ga_expand_collect_flags)
    const ga_succ& s = out[t];
    flags[t] = (uint8_t)((s.n_poses > 0 ? GA_SUCC_ROLLED : 0) | (s.collision_free ? GA_SUCC_FREE : 0) |
                         (s.collision_ambiguous ? GA_SUCC_AMBIGUOUS : 0) | (s.valid ? GA_SUCC_VALID : 0));
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0)
  {
    if (atomicAdd(done.counter, 1u) == gridDim.x - 1)
    {
      *done.counter = 0;
      __threadfence_system();
      *done.flag = done.ticket;
    }
  }
}

__device__ __forceinline__ void expand_one(const coll_view& v, const expand_view& e, const expand_epilogue& ep,
                                           const ga_node* __restrict__ nodes, ga_succ* __restrict__ out, long long t)
{
  const int stride = 2 * e.nb_steer + e.n_ra;
  const int node = (int)(t / stride);
  const int k = (int)(t % stride);
  const ga_node nd = nodes[node];
  ga_succ s;
  s.x = nd.x; s.y = nd.y; s.yaw = nd.yaw;
  s.xi = s.yi = s.yawi = 0;
  s.n_poses = 0;
  s.valid = 0;
  s.near_cell_edge = 0;
  s.collision_free = 0;
  s.collision_ambiguous = 0;
  for (int i = 0; i < 4; ++i) s.pad_[i] = 0;

  double x = nd.x, y = nd.y, yaw = nd.yaw;
  double margin = 1.0, m;
  int amb = 0;
  if (k < 2 * e.nb_steer)
  {
    // calcNextNode (hybrid_a_star_lib.cpp:427-483)
    const double steer = e.steer[k >> 1];
    const int direction = (k & 1) ? -1 : 1;
    // Vehicle::move_car_some_steps (vehicle.cpp:100-122)
    const long long nb = (long long)round(e.arc_l / nd.motion_res);
    const double dist = nd.motion_res * direction;
    int ok = 1, n = 0;
    for (long long i = 0; i < nb; ++i)
    {
      // Vehicle::move (vehicle.cpp:34-43)
      x += dist * cos(yaw);
      y += dist * sin(yaw);
      yaw += dist * tan(steer) / e.wheel_base;
      yaw = angle_mpi_pi(yaw);
      ++n;
      if (ok)
      {
        ok = check_pose_dev(v, x, y, yaw, 0, &m);
        margin = fmin(margin, m);
        if (m < GA_EDGE_EPS) check_pose_robust(v, x, y, yaw, 0, GA_EDGE_EPS, &amb);
      }
    }
    s.n_poses = n;
    s.collision_free = (uint8_t)ok;
    if (n > 0)
    {
      // mapCont2Disc (:410-417)
      const double fx = x * GA_CON2STAR, fy = y * GA_CON2STAR, fa = yaw * e.yaw_res_inv;
      s.x = x; s.y = y; s.yaw = yaw;
      s.xi = (int)fx;
      s.yi = (int)fy;
      s.yawi = (int)round(fa);
      margin = fmin(margin, fmin(fabs(fx - rint(fx)), fabs(fy - rint(fy))));
      margin = fmin(margin, fabs(fabs(fa - trunc(fa)) - 0.5));
      // verifyIndex (:150-154)
      const int on_grid = (-1 < s.xi && s.xi < e.astar_dim) && (-1 < s.yi && s.yi < e.astar_dim);
      s.valid = (uint8_t)(on_grid && ok);
    }
  }
  else if (nd.with_rear_axis)
  {
    // calcRearAxisNode (:350-403) with Vehicle::turn_on_rear_axis (vehicle.cpp:45-98)
    const double delta_angle = e.ra_delta[k - 2 * e.nb_steer];
    const double goal_yaw = angle_mpi_pi(yaw + delta_angle);
    const double res = e.yaw_res_coll_deg * (GA_PI / 180);
    const double sg = (double)((0.0 < delta_angle) - (delta_angle < 0.0));
    double ang_margin;
    int ok = 1, n = 0;
    double prev_yaw = yaw;
    int have_prev = 0;
    do
    {
      // the previous yaw is final once a next one is produced: check it now
      if (have_prev && ok)
      {
        ok = check_pose_dev(v, x, y, prev_yaw, 0, &m);
        margin = fmin(margin, m);
        if (m < GA_EDGE_EPS) check_pose_robust(v, x, y, prev_yaw, 0, GA_EDGE_EPS, &amb);
      }
      yaw = angle_mpi_pi(yaw + sg * res);
      prev_yaw = yaw;
      have_prev = 1;
      ++n;
      ang_margin = fabs(angle_diff(yaw, goal_yaw));
    } while (ang_margin >= res && n < 4096);
    // the last element is replaced by goal_yaw (vehicle.cpp:72-76)
    if (ok)
    {
      ok = check_pose_dev(v, x, y, goal_yaw, 0, &m);
      margin = fmin(margin, m);
      if (m < GA_EDGE_EPS) check_pose_robust(v, x, y, goal_yaw, 0, GA_EDGE_EPS, &amb);
    }
    const double fa = goal_yaw * e.yaw_res_inv;
    s.n_poses = n;
    s.collision_free = (uint8_t)ok;
    s.yaw = goal_yaw;
    s.xi = (int)(x * GA_CON2STAR);
    s.yi = (int)(y * GA_CON2STAR);
    s.yawi = (int)round(fa);
    margin = fmin(margin, fabs(fabs(fa - trunc(fa)) - 0.5));
    s.valid = (uint8_t)ok;
  }
  s.near_cell_edge = (uint8_t)(margin < GA_EDGE_EPS);
  s.collision_ambiguous = (uint8_t)amb;
  out[t] = s;
  if (ep.out)
  {
    // calcCost's heuristic lookup (hybrid_a_star_lib.cpp:242-250), getPathCosts' movement weight (:324), the
    // successor's own motion resolution (:1483), getDistance2goal's cost_dist_ (:781-798)
    ga_succ_cost c;
    c.heur_cost = GA_OUT_OF_HEURISTIC;
    c.heur_dist = GA_OUT_OF_HEURISTIC;
    c.movement_cost = 0.0;
    c.motion_res = 0.0;
    const bool on_grid = s.n_poses > 0 && (-1 < s.xi && s.xi < e.astar_dim) && (-1 < s.yi && s.yi < e.astar_dim);
    if (on_grid)
    {
      const size_t cell = (size_t)s.yi * e.astar_dim + s.xi;
      if (ep.hflags[cell] & GA_H_CLOSED)
      {
        c.heur_cost = ep.hcost[cell];
        c.heur_dist = ep.hdist[cell];
      }
      c.movement_cost = ep.mc[cell];
      c.motion_res = ep.mres[cell];
    }
    ep.out[t] = c;
    // getProxOfCorners(parent pose * con2star_) (:335-336): one value per node, the same for all its primitives
    if (k == 0) ep.parent_prox[node] = prox_of_corners(nd.x, nd.y, nd.yaw, ep.pa);
  }
}

static expand_view make_expand_view(const ga_ctx* c)
{
  expand_view e;
  memcpy(e.steer, c->steer, sizeof(e.steer));
  memcpy(e.ra_delta, c->ra_delta, sizeof(e.ra_delta));
  e.wheel_base = c->ep.wheel_base;
  e.arc_l = c->ep.arc_l;
  e.yaw_res_inv = 1.0 / (c->ep.yaw_res_deg * (M_PI / 180));
  e.yaw_res_coll_deg = (double)c->cp.yaw_res_deg;
  e.nb_steer = c->ep.nb_steer;
  e.n_ra = c->n_ra;
  e.astar_dim = c->astar_dim;
  return e;
}

static int expand_impl(ga_ctx* c, const ga_node* nodes, int n_nodes, ga_succ* out, int slot, ga_succ_cost* out_cost,
                       double* out_parent_prox)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp && c->have_ep, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, n_nodes >= 0 && (n_nodes == 0 || (nodes && out)), GA_ERR_ARG, "bad arguments");
  const bool epi = out_cost != nullptr;
  if (epi)
  {
    GA_REQUIRE(c, slot == 0 || slot == 1, GA_ERR_ARG, "slot must be 0 (guidance) or 1 (path)");
    GA_REQUIRE(c, out_parent_prox != nullptr, GA_ERR_ARG, "out_parent_prox is NULL");
    GA_REQUIRE(c, c->have_pp, GA_ERR_STATE, "ga_set_prox_params first (vehicle corners of getProxOfCorners)");
  }
  if (n_nodes == 0) return GA_OK;
  GA_CUDA(c, cudaSetDevice(c->device));
  const int stride = 2 * c->ep.nb_steer + c->n_ra;
  const size_t b_in = sizeof(ga_node) * (size_t)n_nodes;
  const size_t b_out = sizeof(ga_succ) * (size_t)n_nodes * stride;
  const size_t b_cost = epi ? sizeof(ga_succ_cost) * (size_t)n_nodes * stride : 0;
  const size_t b_prox = epi ? sizeof(double) * (size_t)n_nodes : 0;
  const size_t off_out = (b_in + 63) & ~(size_t)63;
  const size_t off_cost = (off_out + b_out + 63) & ~(size_t)63;
  const size_t off_prox = (off_cost + b_cost + 63) & ~(size_t)63;
  int rc = ga_stage_reserve(c, off_prox + b_prox);
  if (rc) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  char* h = (char*)c->h_stage;
  char* d = (char*)c->d_stage;
  memcpy(h, nodes, b_in);
  GA_CUDA(c, cudaMemcpyAsync(d, h, b_in, cudaMemcpyHostToDevice, c->stream));
  const expand_view e = make_expand_view(c);
  expand_epilogue ep{};
  if (epi)
  {
    ep.hcost = c->d_hcost[slot];
    ep.hdist = c->d_hdist[slot];
    ep.hflags = c->d_hflags[slot];
    ep.mc = c->d_mc;
    ep.mres = c->d_mres;
    ep.out = (ga_succ_cost*)(d + off_cost);
    ep.parent_prox = (double*)(d + off_prox);
    ep.pa.grid = c->d_prox;
    ep.pa.dim = c->astar_dim;
    ep.pa.p = c->pp;
  }
  {
    ga_launch_scope ls(c, GA_K_EXPAND);
    const long long threads = (long long)n_nodes * stride;
    expand_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, c->stream>>>(make_view(c), e, ep, (const ga_node*)d, n_nodes,
                                                                            (ga_succ*)(d + off_out), expand_done{});
  }
  GA_CUDA(c, cudaGetLastError());
  // results are contiguous in the staging area: one copy back
  GA_CUDA(c, cudaMemcpyAsync(h + off_out, d + off_out, (off_prox + b_prox) - off_out, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  memcpy(out, h + off_out, b_out);
  if (epi)
  {
    memcpy(out_cost, h + off_cost, b_cost);
    memcpy(out_parent_prox, h + off_prox, b_prox);
  }
  return GA_OK;
}

extern "C" int ga_expand(ga_ctx* c, const ga_node* nodes, int n_nodes, ga_succ* out)
{
  return expand_impl(c, nodes, n_nodes, out, 0, nullptr, nullptr);
}

extern "C" int ga_expand_ex(ga_ctx* c, const ga_node* nodes, int n_nodes, int slot, ga_succ* out, ga_succ_cost* out_cost,
                            double* out_parent_prox)
{
  if (c && !out_cost) return ga_fail(c, GA_ERR_ARG, "%s: %s", __func__, "out_cost is NULL (use ga_expand)");
  return expand_impl(c, nodes, n_nodes, out, slot, out_cost, out_parent_prox);
}

// ---- asynchronous expansion (submit / collect) ------------------------------------------------------
// For a caller that knows which nodes it will expand NEXT (the successors it has just created): the launch returns at
// once, the results arrive in mapped host memory while the caller goes on with its own work, and are picked up when
// the node is actually expanded.  Nodes travel in the kernel's parameter block, results are written straight to host
// memory, completion is a ticket word the host polls: no copy-engine operation and no driver call on the wait side.
#define GA_ASYNC_RING 4
struct expand_async_state
{
  ga_succ* h_out[GA_ASYNC_RING];  // mapped pinned, GA_EXPAND_ASYNC_MAX * stride entries each
  ga_succ* d_out[GA_ASYNC_RING];
  uint8_t* h_flags[GA_ASYNC_RING];  // mapped pinned, one byte per primitive
  uint8_t* d_flags[GA_ASYNC_RING];
  volatile unsigned* h_flag;      // GA_ASYNC_RING ticket words (mapped)
  unsigned* d_flag;
  unsigned* d_counter;            // GA_ASYNC_RING CTA counters (device)
  unsigned ticket[GA_ASYNC_RING];
  int n_nodes[GA_ASYNC_RING];
  int stride;
  unsigned next_ticket;
  int next_slot;
};

void ga_expand_async_free(ga_ctx* c)
{
  expand_async_state* a = static_cast<expand_async_state*>(c->expand_async);
  if (!a) return;
  for (int i = 0; i < GA_ASYNC_RING; ++i)
  {
    if (a->h_out[i]) cudaFreeHost(a->h_out[i]);
    if (a->h_flags[i]) cudaFreeHost(a->h_flags[i]);
  }
  if (a->h_flag) cudaFreeHost((void*)a->h_flag);
  if (a->d_counter) cudaFree(a->d_counter);
  delete a;
  c->expand_async = nullptr;
}

extern "C" int ga_expand_submit(ga_ctx* c, const ga_node* nodes, int n_nodes, int* ticket)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp && c->have_ep, GA_ERR_STATE, "no patch / params");
  GA_REQUIRE(c, nodes && ticket && n_nodes >= 1 && n_nodes <= GA_EXPAND_ASYNC_MAX, GA_ERR_ARG, "1 .. 32 nodes per submission");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int stride = 2 * c->ep.nb_steer + c->n_ra;
  expand_async_state* a = static_cast<expand_async_state*>(c->expand_async);
  if (a && a->stride != stride)
  {
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
    ga_expand_async_free(c);
    a = nullptr;
  }
  if (!a)
  {
    a = new expand_async_state();
    memset(a, 0, sizeof(*a));
    c->expand_async = a;
    a->stride = stride;
    for (int i = 0; i < GA_ASYNC_RING; ++i)
    {
      GA_CUDA(c, cudaHostAlloc((void**)&a->h_out[i], sizeof(ga_succ) * GA_EXPAND_ASYNC_MAX * (size_t)stride, cudaHostAllocMapped));
      GA_CUDA(c, cudaHostGetDevicePointer((void**)&a->d_out[i], a->h_out[i], 0));
      GA_CUDA(c, cudaHostAlloc((void**)&a->h_flags[i], GA_EXPAND_ASYNC_MAX * (size_t)stride, cudaHostAllocMapped));
      GA_CUDA(c, cudaHostGetDevicePointer((void**)&a->d_flags[i], a->h_flags[i], 0));
    }
    GA_CUDA(c, cudaHostAlloc((void**)&a->h_flag, sizeof(unsigned) * GA_ASYNC_RING, cudaHostAllocMapped));
    GA_CUDA(c, cudaHostGetDevicePointer((void**)&a->d_flag, (void*)a->h_flag, 0));
    for (int i = 0; i < GA_ASYNC_RING; ++i) a->h_flag[i] = 0;
    GA_CUDA(c, cudaMalloc(&a->d_counter, sizeof(unsigned) * GA_ASYNC_RING));
    GA_CUDA(c, cudaMemsetAsync(a->d_counter, 0, sizeof(unsigned) * GA_ASYNC_RING, c->stream));
  }
  const int slot = a->next_slot;
  // the slot's previous submission must have been collected or at least be complete before it is reused
  if (a->ticket[slot] && a->h_flag[slot] != a->ticket[slot]) GA_CUDA(c, cudaStreamSynchronize(c->stream));
  a->next_slot = (slot + 1) % GA_ASYNC_RING;
  unsigned tk = ++a->next_ticket;
  if (tk == 0) tk = ++a->next_ticket;
  a->ticket[slot] = tk;
  a->n_nodes[slot] = n_nodes;
  expand_nodes_param p;
  memcpy(p.n, nodes, sizeof(ga_node) * (size_t)n_nodes);
  expand_done done{ a->d_counter + slot, a->d_flag + slot, tk };
  {
    ga_launch_scope ls(c, GA_K_EXPAND);
    const int threads = n_nodes * stride;
    expand_param_kernel<<<(threads + 127) / 128, 128, 0, c->stream>>>(make_view(c), make_expand_view(c), p, n_nodes, a->d_out[slot],
                                                                        a->d_flags[slot], done);
  }
  GA_CUDA(c, cudaGetLastError());
  *ticket = (int)((tk << 2) | (unsigned)slot);
  return GA_OK;
}

static int expand_wait(ga_ctx* c, int ticket, int* slot_out)
{
  expand_async_state* a = static_cast<expand_async_state*>(c->expand_async);
  GA_REQUIRE(c, a != nullptr, GA_ERR_STATE, "nothing was submitted");
  const int slot = ticket & 3;
  const unsigned tk = (unsigned)ticket >> 2;
  GA_REQUIRE(c, (a->ticket[slot] & 0x3FFFFFFFu) == tk, GA_ERR_ARG, "stale ticket: the ring slot was reused (at most 4 submissions in flight)");
  const unsigned want = a->ticket[slot];
  for (unsigned long long spins = 0; a->h_flag[slot] != want; ++spins)
  {
    if ((spins & 0xFFFFF) == 0xFFFFF)
    {
      const cudaError_t q = cudaStreamQuery(c->stream);
      if (q != cudaErrorNotReady && a->h_flag[slot] != want)
      {
        if (q == cudaSuccess) break;
        return ga_fail(c, GA_ERR_CUDA, "%s: %s", __func__, cudaGetErrorString(q));
      }
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  *slot_out = slot;
  return GA_OK;
}

extern "C" int ga_expand_collect(ga_ctx* c, int ticket, ga_succ* out, int* n_nodes)
{
  if (!c || !out) return GA_ERR_ARG;
  int slot;
  const int rc = expand_wait(c, ticket, &slot);
  if (rc != GA_OK) return rc;
  expand_async_state* a = static_cast<expand_async_state*>(c->expand_async);
  memcpy(out, a->h_out[slot], sizeof(ga_succ) * (size_t)a->n_nodes[slot] * a->stride);
  if (n_nodes) *n_nodes = a->n_nodes[slot];
  return GA_OK;
}

extern "C" int ga_expand_collect_flags(ga_ctx* c, int ticket, uint8_t* flags, int* n_nodes)
{
  if (!c || !flags) return GA_ERR_ARG;
  int slot;
  const int rc = expand_wait(c, ticket, &slot);
  if (rc != GA_OK) return rc;
  expand_async_state* a = static_cast<expand_async_state*>(c->expand_async);
  memcpy(flags, a->h_flags[slot], (size_t)a->n_nodes[slot] * a->stride);
  if (n_nodes) *n_nodes = a->n_nodes[slot];
  return GA_OK;
}
