// Floor of a launch + wait round trip on this box, and ga_check_poses on one expansion fan, timed in C++ (no Python).
// nvcc -O2 -gencode arch=compute_100a,code=sm_100a -I include tools/probe/latency_probe.cu -L guided-extended-hybrid-astar_b200/lib -l:libgridaccel.so -Xlinker -rpath=$PWD/guided-extended-hybrid-astar_b200/lib -o gpurun_out/latency_probe
#include <chrono>
#include <cstdio>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "gridaccel.h"

__global__ void empty_kernel() {}
__global__ void flag_kernel(volatile unsigned* f, unsigned v)
{
  __threadfence_system();
  *f = v;
}

static double now_us() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main()
{
  cudaStream_t s;
  cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  const int reps = 2000;
  for (int i = 0; i < 100; ++i) empty_kernel<<<1, 32, 0, s>>>();
  cudaStreamSynchronize(s);
  double t0 = now_us();
  for (int i = 0; i < reps; ++i)
  {
    empty_kernel<<<1, 32, 0, s>>>();
    cudaStreamSynchronize(s);
  }
  printf("empty kernel + cudaStreamSynchronize : %.2f us\n", (now_us() - t0) / reps);
  unsigned* hf;
  unsigned* df;
  cudaHostAlloc(&hf, 64, cudaHostAllocMapped);
  cudaHostGetDevicePointer(&df, hf, 0);
  *hf = 0;
  t0 = now_us();
  for (int i = 1; i <= reps; ++i)
  {
    flag_kernel<<<1, 32, 0, s>>>(df, (unsigned)i);
    while (*(volatile unsigned*)hf != (unsigned)i) __builtin_ia32_pause();
  }
  printf("flag kernel + spin on mapped word     : %.2f us\n", (now_us() - t0) / reps);
  t0 = now_us();
  for (int i = 0; i < reps; ++i) empty_kernel<<<1, 32, 0, s>>>();
  double t1 = now_us();
  cudaStreamSynchronize(s);
  printf("launch call alone (async, queue of %d): %.2f us each; drain %.2f us each\n", reps, (t1 - t0) / reps, (now_us() - t0) / reps);

  ga_ctx* c = nullptr;
  if (ga_create(0, &c) != 0) return 1;
  ga_astar_params ap{ 5.0, 5.0, 3.0, 0.0, 100.0, 7.0, 3.0, 0.1, 1.0, 3.0, 10.0, 200, GA_DT_IPPFLOAT };
  ga_set_astar_params(c, &ap);
  std::vector<int> cen(120 * 3 * 2, 0);
  ga_collision_params cp{ 100, 130, 10, 3, 3, 4 };
  ga_set_collision_params(c, &cp, cen.data());
  ga_patch_reset(c, 2048);
  ga_build_safety(c);
  ga_sync(c);
  for (int per : { 1, 3, 9 })
  {
    const int n_seg = 26, n = n_seg * per;
    std::vector<double> x(n, 100.0), y(n, 120.0), yaw(n, 0.3);
    std::vector<int> seg(n_seg + 1), hit(n_seg);
    for (int i = 0; i <= n_seg; ++i) seg[i] = i * per;
    for (int i = 0; i < 50; ++i) ga_check_poses(c, x.data(), y.data(), yaw.data(), n, seg.data(), n_seg, 0, hit.data());
    t0 = now_us();
    for (int i = 0; i < reps; ++i) ga_check_poses(c, x.data(), y.data(), yaw.data(), n, seg.data(), n_seg, 0, hit.data());
    printf("ga_check_poses 26 segments x %d poses   : %.2f us / call\n", per, (now_us() - t0) / reps);
  }
  ga_destroy(c);
  return 0;
}
