// Per-operation cost (one thread, back to back, result consumed) of the synchronisation primitives of the GEMM loop.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/kernels.cuh"
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb;
using namespace bsb::ptx;

__global__ void __launch_bounds__(128, 1) ops(int iters, long long* out) {
  __shared__ __align__(8) uint64_t bar[4];
  __shared__ uint32_t slot;
  const int tid = threadIdx.x;
  if (tid == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_mbar_init(); }
  if (tid < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (tid == 0) {
    const uint32_t b0 = smem_u32(&bar[0]), b1 = smem_u32(&bar[1]);
    long long t0, t1; uint32_t acc = 0;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) acc += mbar_try_wait(b0, 1) ? 1 : 0;        // fresh barrier: parity 1 passes
    t1 = clock64(); out[0] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_spin(b0, 1); }
    t1 = clock64(); out[1] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_arrive(b1); }                        // count 1: every arrive completes a phase
    t1 = clock64(); out[2] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { tc_fence_after(); }
    t1 = clock64(); out[3] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mma_commit(b1); }
    t1 = clock64(); out[4] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_arrive(b1); mbar_spin(b1, (uint32_t)(i & 1)); }   // arrive then observe own phase
    t1 = clock64(); out[5] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mma_commit(smem_u32(&bar[2])); mbar_spin(smem_u32(&bar[2]), (uint32_t)(i & 1)); }
    t1 = clock64(); out[6] = (t1 - t0) / iters;
    t0 = clock64();
    for (int i = 0; i < iters; ++i) { mbar_try_wait(b0, 1); tc_fence_after(); mbar_arrive(b1); }
    t1 = clock64(); out[7] = (t1 - t0) / iters;
    out[8] = acc;
  }
  tc_fence_before(); __syncthreads();
  if (tid < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 80);
  long long h[10];
  for (int rep = 0; rep < 2; ++rep) { ops<<<1, 128>>>(2000, d); cudaDeviceSynchronize(); }
  cudaMemcpy(h, d, 72, cudaMemcpyDeviceToHost);
  const char* names[] = {"try_wait (passes)", "test_wait loop (passes)", "mbarrier.arrive", "tcgen05.fence::after", "tcgen05.commit (idle pipe)",
                         "arrive + observe", "commit + observe", "try_wait + fence + arrive"};
  for (int i = 0; i < 8; ++i) printf("%-28s %lld cycles\n", names[i], h[i]);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
