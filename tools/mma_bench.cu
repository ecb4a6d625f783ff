// Microbenchmark: cycles per tcgen05.mma (kind::f16, M=128, K=16) for several N, SS and TS operand modes.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

template <int MODE>   // 0: SS, 1: TS, 2: SS alternating two accumulators
__global__ void __launch_bounds__(128, 1) bench(int n, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, n);
    const uint64_t a = make_sw128_desc(base), b = make_sw128_desc(base + 16384);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t adv = (uint64_t)(kk * 2);
        if (MODE == 0) mma_bf16_ss(tm, a + adv, b + adv, idesc, 1);
        if (MODE == 1) mma_bf16_ts(tm, tm + 384 + kk * 8, b + adv, idesc, 1);
        if (MODE == 2) mma_bf16_ss(tm + (kk & 1) * 256, a + adv, b + adv, idesc, 1);
      }
    }
    mma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  const int smem = 16384 + 32768 + 2048;
  cudaFuncSetAttribute(bench<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(bench<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(bench<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 2000;
  for (int grid : {1, 148}) for (int mode = 0; mode < 3; ++mode) for (int n : {64, 128, 192, 256}) {
    if (mode == 2 && n > 256) continue;
    for (int rep = 0; rep < 2; ++rep) {
      if (mode == 0) bench<0><<<grid, 128, smem>>>(n, iters, d);
      if (mode == 1) bench<1><<<grid, 128, smem>>>(n, iters, d);
      if (mode == 2) bench<2><<<grid, 128, smem>>>(n, iters, d);
      cudaDeviceSynchronize();
    }
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("grid %3d mode %d N %3d : %.1f cycles/MMA  (%s)\n", grid, mode, n, (double)h / (iters * 4), cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
