// ga_prox.cu -- proximity scoring of sampled poses: the last per-pose reads of h_prox_arr_ the
// search makes on the host.
//
// Replaces HybridAStar::getProxOfCorners (hybrid_a_star_lib.cpp:265-289) with util::getBilinInterp /
// util::bilinearInterpolation (util2.cpp:41-52, util1.cpp:152-200) and the prox-cost loops of
// getRAPathCosts / getRSPathCosts (hybrid_a_star_lib.cpp:588-594, 670-676).  fp64 with the reference's
// operand order and no FMA contraction; the one thing that is not the host's arithmetic is sin / cos
// (CUDA libm vs glibc, <= 2 ulp apart), so results agree to ~1e-15 relative, not bit for bit -- the
// interpolant is continuous across cell borders, a corner that lands on the other side of one moves
// the value by the same order.
#include "ga_internal.cuh"

struct prox_args
{
  const double* grid;  // AStar::h_prox_arr_, dim x dim row-major
  int dim;
  ga_prox_params p;
};

// libstdc++'s std::lerp (C++20, <cmath> __lerp): exact at t = 0 / 1, monotone
__device__ __forceinline__ double lerp_std(const double a, const double b, const double t)
{
  if ((a <= 0 && b >= 0) || (a >= 0 && b <= 0)) return t * b + (1 - t) * a;
  if (t == 1) return b;
  const double x = a + t * (b - a);
  return (t > 1) == (b > a) ? (b < x ? x : b) : (b > x ? x : b);
}

// util::bilinearInterpolation (util1.cpp:152-200)
__device__ __forceinline__ double bilinear(const double q11, const double q12, const double q21, const double q22, const int x1,
                                           const int x2, const int y1, const int y2, const double x, const double y)
{
  const int x2x1 = x2 - x1, y2y1 = y2 - y1;
  const double yy1 = y - y1, xx1 = x - x1;
  double result;
  if (x2x1 == 0 && y2y1 == 0) result = q11;
  else if (x2x1 == 0) result = lerp_std(q11, q12, yy1 / y2y1);
  else if (y2y1 == 0) result = lerp_std(q11, q21, xx1 / x2x1);
  else
  {
    const double x2x = static_cast<double>(x2) - x, y2y = y2 - y;
    result = 1.0 / (x2x1 * y2y1) * (q11 * x2x * y2y + q21 * xx1 * y2y + q12 * x2x * yy1 + q22 * xx1 * yy1);
  }
  return isnan(result) ? 0.0 : result;
}

// util::getBilinInterp (util2.cpp:41-52).  The reference clamps the cell indices at 0 only and reads
// past the array for a point beyond the last row / column; here they are clamped into the grid.
__device__ __forceinline__ double bilin_at(const double x, const double y, const double* __restrict__ grid, const int dim)
{
  const int hi = dim - 1;
  const int x1 = min(max(static_cast<int>(x), 0), hi), x2 = min(max(static_cast<int>(ceil(x)), 0), hi);
  const int y1 = min(max(static_cast<int>(y), 0), hi), y2 = min(max(static_cast<int>(ceil(y)), 0), hi);
  const double q11 = __ldg(grid + (size_t)y1 * dim + x1), q12 = __ldg(grid + (size_t)y2 * dim + x1);
  const double q21 = __ldg(grid + (size_t)y1 * dim + x2), q22 = __ldg(grid + (size_t)y2 * dim + x2);
  return bilinear(q11, q12, q21, q22, x1, x2, y1, y2, x, y);
}

// getProxOfCorners(pose * con2star_): the corner offsets stay in METRES while the pose is in astar
// cells -- that is what the reference computes (hybrid_a_star_lib.cpp:270-274), kept as it is
__device__ __forceinline__ double prox_of_corners(const double px, const double py, const double yaw, const prox_args& a)
{
  double s, c;
  sincos(yaw, &s, &c);
  const double ox = px * a.p.con2star, oy = py * a.p.con2star;
  const double hw = a.p.width / 2;
  const double cx[4] = { a.p.lf, a.p.lf, -a.p.lb, -a.p.lb }, cy[4] = { -hw, hw, -hw, hw };  // f_r, f_l, r_r, r_l (vehicle.cpp:24-27)
  double best = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
  {
    const double x = (cx[k] * c - cy[k] * s) + ox, y = (cx[k] * s + cy[k] * c) + oy;  // rotatePreCalc + offset
    const double v = bilin_at(x, y, a.grid, a.dim);
    best = k == 0 ? v : (best < v ? v : best);  // std::max({v_1 .. v_4}); bilinear never returns NaN
  }
  return best;
}

__global__ void __launch_bounds__(256) prox_poses_kernel(prox_args a, const double* __restrict__ poses, int n, double* __restrict__ out)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = prox_of_corners(poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], a);
}

// one warp per path: the lanes score 32 poses at a time, the sum runs in list order (the reference
// adds pose by pose, and fp64 addition is not associative)
__global__ void __launch_bounds__(256) prox_paths_kernel(prox_args a, const double* __restrict__ poses, const int* __restrict__ offsets,
                                                         int n_paths, double* __restrict__ out)
{
  const int lane = threadIdx.x & 31, path = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (path >= n_paths) return;
  const int b = offsets[path], e = offsets[path + 1];
  double acc = 0;
  for (int c0 = b; c0 < e; c0 += 32)
  {
    const int i = c0 + lane;
    double t = 0;
    if (i < e) t = prox_of_corners(poses[3 * i], poses[3 * i + 1], poses[3 * i + 2], a) * a.p.h_prox_cost * a.p.interp_res;
    const int m = min(32, e - c0);
    for (int k = 0; k < m; ++k) acc += __shfl_sync(0xFFFFFFFFu, t, k);
  }
  if (lane == 0) out[path] = acc;
}

extern "C" int ga_set_prox_params(ga_ctx* c, const ga_prox_params* p)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, p != nullptr, GA_ERR_ARG, "params are NULL");
  GA_REQUIRE(c, p->con2star > 0 && p->width >= 0, GA_ERR_ARG, "bad geometry");
  c->pp = *p;
  c->have_pp = true;
  return GA_OK;
}

static int prox_prepare(ga_ctx* c, const void* host_a, size_t bytes_a, const void* host_b, size_t bytes_b, size_t bytes_out,
                        unsigned char** d_a, unsigned char** d_b, unsigned char** d_out)
{
  GA_REQUIRE(c, c->have_pp, GA_ERR_STATE, "no prox params (call ga_set_prox_params)");
  GA_REQUIRE(c, c->astar_dim > 0 && c->d_prox != nullptr, GA_ERR_STATE, "no patch");
  GA_CUDA(c, cudaSetDevice(c->device));
  const size_t off_b = (bytes_a + 255) & ~(size_t)255, off_o = off_b + ((bytes_b + 255) & ~(size_t)255);
  const int rc = ga_stage_reserve(c, off_o + bytes_out);
  if (rc != GA_OK) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));  // the pinned staging buffer may still feed an earlier copy
  memcpy(c->h_stage, host_a, bytes_a);
  if (bytes_b) memcpy(static_cast<unsigned char*>(c->h_stage) + off_b, host_b, bytes_b);
  GA_CUDA(c, cudaMemcpyAsync(c->d_stage, c->h_stage, off_b + bytes_b, cudaMemcpyHostToDevice, c->stream));
  *d_a = static_cast<unsigned char*>(c->d_stage);
  *d_b = *d_a + off_b;
  *d_out = *d_a + off_o;
  return GA_OK;
}

static int prox_fetch(ga_ctx* c, const unsigned char* d_out, double* out, size_t bytes)
{
  unsigned char* h_out = static_cast<unsigned char*>(c->h_stage) + (d_out - static_cast<unsigned char*>(c->d_stage));
  GA_CUDA(c, cudaMemcpyAsync(h_out, d_out, bytes, cudaMemcpyDeviceToHost, c->stream));
  GA_CUDA(c, cudaStreamSynchronize(c->stream));
  memcpy(out, h_out, bytes);
  return GA_OK;
}

extern "C" int ga_prox_of_corners(ga_ctx* c, const double* poses, int n, double* out)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, n >= 0 && (n == 0 || (poses && out)), GA_ERR_ARG, "bad pose list");
  if (n == 0) return GA_OK;
  unsigned char *d_a, *d_b, *d_o;
  const int rc = prox_prepare(c, poses, sizeof(double) * 3 * n, nullptr, 0, sizeof(double) * n, &d_a, &d_b, &d_o);
  if (rc != GA_OK) return rc;
  const prox_args a{ c->d_prox, c->astar_dim, c->pp };
  {
    ga_launch_scope ls(c, GA_K_PROX);
    prox_poses_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(a, reinterpret_cast<const double*>(d_a), n, reinterpret_cast<double*>(d_o));
  }
  GA_CUDA(c, cudaGetLastError());
  return prox_fetch(c, d_o, out, sizeof(double) * n);
}

extern "C" int ga_path_prox_costs(ga_ctx* c, const double* poses, const int* path_offsets, int n_paths, double* out)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, n_paths >= 0 && (n_paths == 0 || (path_offsets && out)), GA_ERR_ARG, "bad path list");
  if (n_paths == 0) return GA_OK;
  GA_REQUIRE(c, path_offsets[0] == 0, GA_ERR_ARG, "path_offsets[0] must be 0");
  for (int i = 0; i < n_paths; ++i)
    GA_REQUIRE(c, path_offsets[i + 1] >= path_offsets[i], GA_ERR_ARG, "path_offsets must not decrease");
  const int n = path_offsets[n_paths];
  GA_REQUIRE(c, n == 0 || poses != nullptr, GA_ERR_ARG, "poses are NULL");
  unsigned char *d_a, *d_b, *d_o;
  const double none = 0;
  const int rc = prox_prepare(c, n ? static_cast<const void*>(poses) : &none, n ? sizeof(double) * 3 * n : sizeof none, path_offsets,
                              sizeof(int) * (n_paths + 1), sizeof(double) * n_paths, &d_a, &d_b, &d_o);
  if (rc != GA_OK) return rc;
  const prox_args a{ c->d_prox, c->astar_dim, c->pp };
  {
    ga_launch_scope ls(c, GA_K_PROX);
    const long long threads = (long long)n_paths * 32;
    prox_paths_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, c->stream>>>(a, reinterpret_cast<const double*>(d_a),
                                                                              reinterpret_cast<const int*>(d_b), n_paths,
                                                                              reinterpret_cast<double*>(d_o));
  }
  GA_CUDA(c, cudaGetLastError());
  return prox_fetch(c, d_o, out, sizeof(double) * n_paths);
}
