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
#include "ga_prox.cuh"

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
