// ga_raytrace.cu -- the simulated lidar ray-cast that produces each cycle's sensor window in the
// unknown-environment loop.
//
// Replaces GridMapSim::calculatePrecasts / raytracing (LIB/src/gridmap_sim_lib/gridmap_sim.cpp:14-80):
// 180 angular sectors of 2 degrees, each a list of the cells of the 801 x 801 window it covers sorted
// by distance from the centre; per list the first occupied cell of the ground truth is marked
// occupied, everything before it free, the rest stays unknown.  The reference walks the lists one
// after the other; an occupied mark only ever lands on a ground-truth-occupied cell and a free mark
// only on a non-occupied one, so the lists commute and run as 180 independent CTAs.
#include <algorithm>
#include <cmath>
#include <vector>

#include "ga_internal.cuh"

#define RT_DIM 801      // GM_DIM, gridmap_sim.hpp:17
#define RT_LISTS 180    // NB_PRECASTS, gridmap_sim.hpp:21
#define RT_THREADS 512

namespace
{
struct rt_cast  // PrecastDB (gridmap_sim.hpp:74-90): ordered by distance only
{
  double d;
  int ix, iy;
  bool operator<(const rt_cast& o) const { return d < o.d; }
};
}  // namespace

// gridmap_sim.cpp:14-40 on the host, once per context: fp64 libm hypot / atan2 and std::sort, exactly
// the reference's arithmetic and ordering algorithm (equal distances are ordered by std::sort's
// behaviour on the same input sequence).  entries = ix * 801 + iy, the flat index both
// grid_map_data(ix, iy) and pmap_(ix, iy) address.
static void build_precasts(std::vector<int>& entries, int* offsets)
{
  const double pi = 3.14159265358979323846, yaw_res = 2 * (pi / 180);
  const int middle = static_cast<int>(static_cast<double>(RT_DIM - 1) / 2);
  std::vector<std::vector<rt_cast>> lists(RT_LISTS);
  for (int ix = 0; ix < RT_DIM; ++ix)
    for (int iy = 0; iy < RT_DIM; ++iy)
    {
      const double dist = hypot(ix - middle, iy - middle);
      double angle = atan2(iy - middle, ix - middle);
      if (angle < 0) angle += 2 * pi;
      const int a1 = std::min(static_cast<int>(floor(angle / yaw_res)), RT_LISTS - 1);
      const int a2 = std::min(static_cast<int>(ceil(angle / yaw_res)), RT_LISTS - 1);
      lists[a1].push_back(rt_cast{ dist, ix, iy });
      lists[a2].push_back(rt_cast{ dist, ix, iy });
    }
  entries.clear();
  for (int k = 0; k < RT_LISTS; ++k)
  {
    std::sort(lists[k].begin(), lists[k].end());
    offsets[k] = (int)entries.size();
    for (const rt_cast& c : lists[k]) entries.push_back(c.ix * RT_DIM + c.iy);
  }
  offsets[RT_LISTS] = (int)entries.size();
}

// one CTA per precast list: chunk-wise search of the first occupied cell (the chunks are in list
// order, so the first chunk with a hit holds the first hit), then the marks
__global__ void __launch_bounds__(RT_THREADS)
raytrace_kernel(const uint8_t* __restrict__ gt, const int* __restrict__ entries, const int* __restrict__ offsets,
                uint8_t* __restrict__ pmap)
{
  __shared__ int s_hit;
  const int b = offsets[blockIdx.x], e = offsets[blockIdx.x + 1];
  if (threadIdx.x == 0) s_hit = e;
  __syncthreads();
  for (int c0 = b; c0 < e; c0 += RT_THREADS)
  {
    const int i = c0 + threadIdx.x;
    if (i < e && gt[__ldg(entries + i)] == 255) atomicMin(&s_hit, i);  // OCCUPIED, gridmap_sim.hpp:28
    __syncthreads();
    const int h = s_hit;
    __syncthreads();  // nobody may start the next chunk's atomicMin before everyone has read this one's result
    if (h < e) break;
  }
  const int hit = s_hit;
  for (int i = b + threadIdx.x; i < hit; i += RT_THREADS) pmap[__ldg(entries + i)] = 0;  // FREE
  if (threadIdx.x == 0 && hit < e) pmap[entries[hit]] = 255;
}

void ga_rt_free(ga_ctx* c)
{
  if (c->d_rt_entries) cudaFree(c->d_rt_entries);
  if (c->d_rt_offsets) cudaFree(c->d_rt_offsets);
  if (c->d_rt_gt) cudaFree(c->d_rt_gt);
  if (c->d_pmap) cudaFree(c->d_pmap);
  c->d_rt_entries = c->d_rt_offsets = nullptr;
  c->d_rt_gt = c->d_pmap = nullptr;
}

static int rt_init(ga_ctx* c)
{
  if (c->d_rt_entries) return GA_OK;
  std::vector<int> entries;
  int offsets[RT_LISTS + 1];
  build_precasts(entries, offsets);
  GA_CUDA(c, cudaMalloc(&c->d_rt_entries, sizeof(int) * entries.size()));
  GA_CUDA(c, cudaMalloc(&c->d_rt_offsets, sizeof offsets));
  GA_CUDA(c, cudaMalloc(&c->d_rt_gt, (size_t)RT_DIM * RT_DIM));
  GA_CUDA(c, cudaMalloc(&c->d_pmap, (size_t)RT_DIM * RT_DIM));
  GA_CUDA(c, cudaMemcpy(c->d_rt_entries, entries.data(), sizeof(int) * entries.size(), cudaMemcpyHostToDevice));
  GA_CUDA(c, cudaMemcpy(c->d_rt_offsets, offsets, sizeof offsets, cudaMemcpyHostToDevice));
  GA_CUDA(c, cudaMemsetAsync(c->d_pmap, 127, (size_t)RT_DIM * RT_DIM, c->stream));
  return GA_OK;
}

// GridMapSim::gmSimulation (gridmap_sim.cpp:92-103): grid_map_gt = 801 x 801 u8 ground-truth window
// in host memory; the result stays on the device (GA_GRID_PMAP) and is copied to pmap_out when that
// is not NULL
extern "C" int ga_raytrace(ga_ctx* c, const uint8_t* grid_map_gt, uint8_t* pmap_out)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, grid_map_gt != nullptr, GA_ERR_ARG, "grid_map_gt is NULL");
  GA_CUDA(c, cudaSetDevice(c->device));
  int rc = rt_init(c);
  if (rc != GA_OK) return rc;
  const size_t bytes = (size_t)RT_DIM * RT_DIM;
  rc = ga_stage_reserve(c, bytes);
  if (rc != GA_OK) return rc;
  GA_CUDA(c, cudaStreamSynchronize(c->stream));  // the pinned staging buffer may still feed an earlier copy
  memcpy(c->h_stage, grid_map_gt, bytes);
  GA_CUDA(c, cudaMemcpyAsync(c->d_rt_gt, c->h_stage, bytes, cudaMemcpyHostToDevice, c->stream));
  GA_CUDA(c, cudaMemsetAsync(c->d_pmap, 127, bytes, c->stream));  // pmap_.setTo(UNKNOWN), :46
  {
    ga_launch_scope ls(c, GA_K_RAYTRACE);
    raytrace_kernel<<<RT_LISTS, RT_THREADS, 0, c->stream>>>(c->d_rt_gt, c->d_rt_entries, c->d_rt_offsets, c->d_pmap);
  }
  GA_CUDA(c, cudaGetLastError());
  if (pmap_out)
  {
    GA_CUDA(c, cudaMemcpyAsync(pmap_out, c->d_pmap, bytes, cudaMemcpyDeviceToHost, c->stream));
    GA_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return GA_OK;
}
