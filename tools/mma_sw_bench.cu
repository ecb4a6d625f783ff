// Microbenchmark: cycles per tcgen05.mma M128 x N x K16 (SS form, bf16) for SWIZZLE_128B vs SWIZZLE_64B operand
// images, in the issue pattern of tc_grouped_linear_kernel (6 MMAs per K chunk of a stage, commit per chunk,
// 4 stages in rotation). No loads: the smem holds zeros; only the tensor pipe and its smem operand reads are timed.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__device__ __forceinline__ uint64_t desc_sw64(uint32_t a) {
  uint64_t d = 0;
  d |= (uint64_t)((a & 0x3FFFF) >> 4); d |= (uint64_t)1 << 16; d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46; d |= (uint64_t)4 << 61;
  return d;
}

__global__ void __launch_bounds__(128, 1) bench(int mode, int N, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = make_idesc_bf16(128, N);
    uint32_t ph = 0;
    long long t0 = 0;
    for (int rep = 0; rep < 2; ++rep) {           // rep 0 = warm-up
      t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t sb = base + (i & 3) * 49152;
        if (elect_one()) {
          if (mode == 0) {                         // SW128, BK = 64: 4 k-steps, 3 MMAs each (A 2 x 16 KB, W 2 x 32 KB do not fit 48 KB: overlap, zeros anyway)
            const uint64_t ah = make_sw128_desc(sb), al = make_sw128_desc(sb + 16384), wh = make_sw128_desc(sb), wl = make_sw128_desc(sb + 16384);
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {       // two k-steps = the same 6 MMAs per "chunk" as the SW64 kernel
              const uint64_t adv = (uint64_t)(kk * 2);
              mma_bf16_ss(tm, al + adv, wh + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wl + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wh + adv, idesc, 1);
            }
          } else {                                 // SW64, BK = 32: the kernel's layout
            const uint64_t ah = desc_sw64(sb), al = desc_sw64(sb + 8192), wh = desc_sw64(sb + 16384), wl = desc_sw64(sb + 32768);
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const uint64_t adv = (uint64_t)(kk * 2);
              mma_bf16_ss(tm, al + adv, wh + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wl + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wh + adv, idesc, 1);
            }
          }
          if ((i & 7) == 7 || i == iters - 1) mma_commit(smem_u32(&bar));
        }
        __syncwarp();
        if ((i & 7) == 7 || i == iters - 1) { mbar_wait(smem_u32(&bar), ph); ph ^= 1; }
      }
    }
    if (threadIdx.x == 0) out[0] = clock64() - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

// mode 0: per chunk 6 MMAs + commit + wait (serial): cycles = T_mma + commit->arrive latency + wake
// mode 1: per chunk commit + wait with NO MMAs: the bare commit->arrive->wake round trip
// mode 2: 6 MMAs + commit per chunk, wait lagging `lag` chunks behind (a ring of `lag` barriers): throughput vs depth
__global__ void __launch_bounds__(128, 1) lat(int mode, int lag, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar[8];
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)))[i] = 0;
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar[i]), 1); fence_mbar_init(); }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x < 32) {
    const uint32_t idesc = make_idesc_bf16(128, 256);
    uint32_t phase_bits = 0;                  // per-barrier parity
    long long t0 = clock64();
    for (int i = 0; i < iters + lag; ++i) {
      if (i < iters) {
        const uint32_t sb = base + (i & 3) * 49152;
        if (elect_one()) {
          if (mode != 1) {
            const uint64_t ah = desc_sw64(sb), al = desc_sw64(sb + 8192), wh = desc_sw64(sb + 16384), wl = desc_sw64(sb + 32768);
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
              const uint64_t adv = (uint64_t)(kk * 2);
              mma_bf16_ss(tm, al + adv, wh + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wl + adv, idesc, 1);
              mma_bf16_ss(tm, ah + adv, wh + adv, idesc, 1);
            }
          }
          mma_commit(smem_u32(&bar[i & 7]));
        }
        __syncwarp();
      }
      const int j = i - lag;                  // wait for chunk j
      if (j >= 0) {
        const int b = j & 7;
        mbar_wait(smem_u32(&bar[b]), (phase_bits >> b) & 1);
        phase_bits ^= 1u << b;
        if (mode == 3) tc_fence_after();
        if (mode == 4) { tc_fence_before(); tc_fence_after(); }
      }
    }
    if (threadIdx.x == 0) out[0] = clock64() - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  {
    long long* d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(lat, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 1024;
    for (int mode : {0, 1, 3, 4})
      for (int lag : {0, 1, 3}) {
        if (mode == 1 && lag > 1) continue;
        lat<<<148, 128, 200 * 1024>>>(mode, lag, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("%s, wait lagging %d chunks: %.0f cycles per chunk %s\n", mode == 1 ? "commit only" : mode == 3 ? "6 MMAs + commit + fence::after_thread_sync after each wait" : mode == 4 ? "6 MMAs + commit + fence before+after" : "6 MMAs + commit", lag,
               (double)c / iters, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  }
  long long* d; cudaMalloc(&d, 8);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 2048;
  for (int N : {256, 128, 64})
    for (int mode = 0; mode < 2; ++mode)
      for (int grid : {1, 148}) {
        bench<<<grid, 128, 200 * 1024>>>(mode, N, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("N=%3d %s grid=%3d: %.1f cycles per MMA (%.0f per 6-MMA chunk) %s\n", N, mode ? "SW64 " : "SW128", grid,
               (double)c / iters / 6, (double)c / iters, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  return 0;
}
