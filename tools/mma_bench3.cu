// Microbenchmark 3: does tcgen05.mma throughput depend on operand reuse between consecutive MMAs?
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__global__ void __launch_bounds__(128, 1) bench(int mode, int n, int iters, long long* out) {
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
    const uint64_t a_hi = make_sw128_desc(base), a_lo = make_sw128_desc(base + 16384);
    const uint64_t w_hi = make_sw128_desc(base + 32768), w_lo = make_sw128_desc(base + 65536);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const uint64_t adv = (uint64_t)(kk * 2);
        if (mode == 0) {          // kernel order: (lo,hi) (hi,lo) (hi,hi)
          mma_bf16_ss(tm, a_lo + adv, w_hi + adv, idesc, 1);
          mma_bf16_ss(tm, a_hi + adv, w_lo + adv, idesc, 1);
          mma_bf16_ss(tm, a_hi + adv, w_hi + adv, idesc, 1);
        } else if (mode == 1) {   // B reused twice in a row: (lo,hi) (hi,hi) (hi,lo)
          mma_bf16_ss(tm, a_lo + adv, w_hi + adv, idesc, 1);
          mma_bf16_ss(tm, a_hi + adv, w_hi + adv, idesc, 1);
          mma_bf16_ss(tm, a_hi + adv, w_lo + adv, idesc, 1);
        } else if (mode == 2) {   // same operands every time
          mma_bf16_ss(tm, a_hi, w_hi, idesc, 1);
          mma_bf16_ss(tm, a_hi, w_hi, idesc, 1);
          mma_bf16_ss(tm, a_hi, w_hi, idesc, 1);
        } else if (mode == 3) {   // TS: A from TMEM, kernel order
          mma_bf16_ts(tm, tm + 384 + 32 + kk * 8, w_hi + adv, idesc, 1);
          mma_bf16_ts(tm, tm + 384 + kk * 8, w_lo + adv, idesc, 1);
          mma_bf16_ts(tm, tm + 384 + kk * 8, w_hi + adv, idesc, 1);
        }
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
  const int smem = 98304 + 2048;
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"SS (lo,hi)(hi,lo)(hi,hi)", "SS (lo,hi)(hi,hi)(hi,lo)", "SS same operands", "TS kernel order"};
  for (int mode = 0; mode < 4; ++mode) for (int n : {128, 192, 256}) {
    for (int rep = 0; rep < 2; ++rep) { bench<<<148, 128, smem>>>(mode, n, 1000, d); cudaDeviceSynchronize(); }
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-28s N %3d : %.1f cycles/MMA (%s)\n", names[mode], n, (double)h / 12000, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
