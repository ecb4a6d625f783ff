// Consumer-loop decomposition: what a passing wait + release costs one thread, alone and with a partner thread.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/kernels.cuh"
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb;
using namespace bsb::ptx;

__global__ void __launch_bounds__(128, 1) ops(int iters, long long* out) {
  __shared__ __align__(8) uint64_t bar[16];
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { for (int i = 0; i < 16; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_mbar_init(); }
  if (tid < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t b0 = smem_u32(&bar[0]), b1 = smem_u32(&bar[1]);
  long long t0, t1;
  if (tid == 0) {
    t0 = clock64(); for (int i = 0; i < iters; ++i) { mbar_wait(b0, 1); } t1 = clock64(); out[0] = (t1 - t0) / iters;
    t0 = clock64(); for (int i = 0; i < iters; ++i) { mbar_wait(b0, 1); mma_commit(b1); } t1 = clock64(); out[1] = (t1 - t0) / iters;
    t0 = clock64(); for (int i = 0; i < iters; ++i) { mbar_wait(b0, 1); mbar_arrive(b1); } t1 = clock64(); out[2] = (t1 - t0) / iters;
    // wait on the phase the own previous arrive completed (same barrier)
    t0 = clock64(); for (int i = 0; i < iters; ++i) { mbar_arrive(smem_u32(&bar[2])); mbar_wait(smem_u32(&bar[2]), (uint32_t)(i & 1)); } t1 = clock64(); out[3] = (t1 - t0) / iters;
    t0 = clock64(); for (int i = 0; i < iters; ++i) { mma_commit(smem_u32(&bar[3])); mbar_wait(smem_u32(&bar[3]), (uint32_t)(i & 1)); } t1 = clock64(); out[4] = (t1 - t0) / iters;
  }
  __syncthreads();
  // two threads, ring of 4, producer never waits on the consumer beyond the ring: who pays what
  for (int variant = 0; variant < 3; ++variant) {
    uint64_t* full = &bar[4]; uint64_t* empty = &bar[8];
    __syncthreads();
    if (tid == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar[4 + i]), 1); fence_mbar_init(); }
    __syncthreads();
    if (warp == 0 && lane == 0) {
      int st = 0; uint32_t ph = 0; long long tw = 0, tr = 0;
      t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const long long a = clock64();
        mbar_wait(smem_u32(&full[st]), ph);
        const long long b = clock64();
        if (variant == 0) mma_commit(smem_u32(&empty[st])); else mbar_arrive(smem_u32(&empty[st]));
        const long long c = clock64();
        tw += b - a; tr += c - b;
        if (++st == 4) { st = 0; ph ^= 1; }
      }
      t1 = clock64();
      out[5 + 3 * variant] = (t1 - t0) / iters; out[6 + 3 * variant] = tw / iters; out[7 + 3 * variant] = tr / iters;
    } else if (warp == 1 && lane == 0) {
      int st = 0; uint32_t ph = 0;
      for (int i = 0; i < iters; ++i) {
        if (variant == 2) mbar_spin(smem_u32(&empty[st]), ph ^ 1); else mbar_wait(smem_u32(&empty[st]), ph ^ 1);
        mbar_arrive(smem_u32(&full[st]));
        if (++st == 4) { st = 0; ph ^= 1; }
      }
    }
  }
  tc_fence_before(); __syncthreads();
  if (tid < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8 * 16);
  long long h[16];
  for (int rep = 0; rep < 2; ++rep) { ops<<<1, 128>>>(2000, d); cudaDeviceSynchronize(); }
  cudaMemcpy(h, d, 8 * 16, cudaMemcpyDeviceToHost);
  const char* names[] = {"wait(passes)", "wait + commit(other bar)", "wait + arrive(other bar)", "arrive + wait(own phase)", "commit + wait(own phase)",
                         "ring4 commit/try_wait partner: iter", "   wait", "   release", "ring4 arrive/try_wait partner: iter", "   wait", "   release",
                         "ring4 arrive/test_wait partner: iter", "   wait", "   release"};
  for (int i = 0; i < 14; ++i) printf("%-40s %lld cycles\n", names[i], h[i]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
