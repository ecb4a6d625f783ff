// Cost of a passing mbarrier wait executed by a whole warp vs one lane (plus the reconvergence), per iteration.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/kernels.cuh"
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb;
using namespace bsb::ptx;

__global__ void __launch_bounds__(128, 1) ops(int iters, long long* out) {
  __shared__ __align__(8) uint64_t bar[4];
  const int tid = threadIdx.x, lane = tid & 31;
  if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_mbar_init(); }
  __syncthreads();
  if (tid < 32) {
    const uint32_t b0 = smem_u32(&bar[0]);
    long long t0, t1; uint32_t acc = 0;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_wait(b0, 1); }
    t1 = clock64(); if (lane == 0) out[0] = (t1 - t0) / iters;
    __syncwarp();
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { if (lane == 0) mbar_wait(b0, 1); __syncwarp(); }
    t1 = clock64(); if (lane == 0) out[1] = (t1 - t0) / iters;
    __syncwarp();
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_spin(b0, 1); }
    t1 = clock64(); if (lane == 0) out[2] = (t1 - t0) / iters;
    __syncwarp();
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { if (lane == 0) mbar_spin(b0, 1); __syncwarp(); }
    t1 = clock64(); if (lane == 0) out[3] = (t1 - t0) / iters;
    __syncwarp();
    if (lane == 0) {
      t0 = clock64();
      for (int i = 0; i < iters; ++i) { mbar_wait(b0, 1); }
      t1 = clock64(); out[4] = (t1 - t0) / iters;
      t0 = clock64();
      for (int i = 0; i < iters; ++i) { acc += (uint32_t)clock64(); }
      t1 = clock64(); out[5] = (t1 - t0) / iters;
      t0 = clock64();
      for (int i = 0; i < iters; ++i) { acc += (uint32_t)clock(); }
      t1 = clock64(); out[6] = (t1 - t0) / iters;
      out[7] = acc;
    }
  }
}

int main() {
  long long* d; cudaMalloc(&d, 80);
  long long h[10];
  for (int rep = 0; rep < 2; ++rep) { ops<<<1, 128>>>(2000, d); cudaDeviceSynchronize(); }
  cudaMemcpy(h, d, 72, cudaMemcpyDeviceToHost);
  const char* names[] = {"try_wait, 32 lanes", "try_wait, lane 0 + syncwarp", "test_wait, 32 lanes", "test_wait, lane 0 + syncwarp",
                         "try_wait, lane 0 alone", "clock64", "clock"};
  for (int i = 0; i < 7; ++i) printf("%-30s %lld cycles\n", names[i], h[i]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
