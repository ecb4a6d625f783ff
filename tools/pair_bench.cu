// Throughput of back-to-back tcgen05.mma.cta_group::2 (M=256 over a CTA pair, N, K=16, SWIZZLE_64B operands)
// against cta_group::1 (M=128) on every SM, all 148 SMs busy.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/kernels.cuh"
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb;
using namespace bsb::ptx;

__device__ __forceinline__ uint64_t desc64(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;
  return d;
}

template <bool PAIR>
__global__ void __launch_bounds__(384, 1) bench(int N, int iters, long long* out, int mode) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_done, bar_e[8], bar_x;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int tid = threadIdx.x;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0;
  if (tid == 0) { mbar_init(smem_u32(&bar_done), 1); mbar_init(smem_u32(&bar_x), 1); for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar_e[i]), 1); fence_mbar_init(); }
  if (tid < 32) {
    if (PAIR) { tmem_alloc_pair(smem_u32(&slot), 512); tmem_relinquish_pair(); }
    else { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (rank == 0 && tid < 32 && mode >= 6) {
    const int lane = tid;
    const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, N);
    const uint64_t a_hi = desc64(base), a_lo = desc64(base + 8192), w_hi = desc64(base + 16384), w_lo = desc64(base + 32768);
    const long long t0 = clock64();
    int st = 0;
    for (int i = 0; i < iters; ++i) {
      if (i >= 4) mbar_spin(smem_u32(&bar_e[(st + 4) & 7]), (uint32_t)(((i - 4) >> 3) & 1));
      if (mode >= 7) tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          if (PAIR) {
            mma_bf16_ss_pair(tm, a_lo + 2 * kk, w_hi + 2 * kk, idesc, 1);
            mma_bf16_ss_pair(tm, a_hi + 2 * kk, w_lo + 2 * kk, idesc, 1);
            mma_bf16_ss_pair(tm, a_hi + 2 * kk, w_hi + 2 * kk, idesc, 1);
          } else {
            mma_bf16_ss(tm, a_lo + 2 * kk, w_hi + 2 * kk, idesc, 1);
            mma_bf16_ss(tm, a_hi + 2 * kk, w_lo + 2 * kk, idesc, 1);
            mma_bf16_ss(tm, a_hi + 2 * kk, w_hi + 2 * kk, idesc, 1);
          }
        }
        if (PAIR) mma_commit_pair(smem_u32(&bar_e[st]), 1); else mma_commit(smem_u32(&bar_e[st]));
      }
      __syncwarp();
      st = (st + 1) & 7;
    }
    if (lane == 0) { if (PAIR) mma_commit_pair(smem_u32(&bar_done), 3); else mma_commit(smem_u32(&bar_done)); }
    mbar_wait(smem_u32(&bar_done), 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
  } else if (rank == 0 && tid == 0) {
    const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, N);
    const uint64_t a_hi = desc64(base), a_lo = desc64(base + 8192), w_hi = desc64(base + 16384), w_lo = desc64(base + 32768);
    const long long t0 = clock64();
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      if (mode == 2) mbar_spin(smem_u32(&bar_x), 1);                 // passes at once (fresh barrier)
      if (mode >= 3 && i >= 4) {                                      // ring: the commit of 4 chunks ago
        const int ws = (st + 4) & 7;                                  // == (i-4) & 7
        mbar_spin(smem_u32(&bar_e[ws]), (uint32_t)(((i - 4) >> 3) & 1));
      }
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        if (PAIR) {
          mma_bf16_ss_pair(tm, a_lo + 2 * kk, w_hi + 2 * kk, idesc, 1);
          mma_bf16_ss_pair(tm, a_hi + 2 * kk, w_lo + 2 * kk, idesc, 1);
          mma_bf16_ss_pair(tm, a_hi + 2 * kk, w_hi + 2 * kk, idesc, 1);
        } else {
          mma_bf16_ss(tm, a_lo + 2 * kk, w_hi + 2 * kk, idesc, 1);
          mma_bf16_ss(tm, a_hi + 2 * kk, w_lo + 2 * kk, idesc, 1);
          mma_bf16_ss(tm, a_hi + 2 * kk, w_hi + 2 * kk, idesc, 1);
        }
      }
      if (mode >= 1) {
        if (PAIR) mma_commit_pair(smem_u32(&bar_e[st]), 1); else mma_commit(smem_u32(&bar_e[st]));
      }
      st = (st + 1) & 7; if (st == 0) ph ^= 1;
    }
    if (PAIR) mma_commit_pair(smem_u32(&bar_done), 3); else mma_commit(smem_u32(&bar_done));
    mbar_wait(smem_u32(&bar_done), 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  } else if (tid == 0 && rank == 1) {
    mbar_wait(smem_u32(&bar_done), 0);
  } else if (tid >= 128 && mode == 4) {
    mbar_spin(smem_u32(&bar_done), 0);        // 8 warps polling, like an idle epilogue
  } else if (tid >= 128 && mode == 5) {
    mbar_wait(smem_u32(&bar_done), 0);        // 8 warps in try_wait
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  if (tid < 32) { tc_fence_after(); if (PAIR) tmem_dealloc_pair(tm, 512); else tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  const int smem = 64 * 1024;
  cudaFuncSetAttribute(bench<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(bench<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 1000;
  for (int mode = 6; mode < 8; ++mode)
  for (int N : {256}) {
    printf("mode %d: ", mode);
    long long h;
    for (int rep = 0; rep < 2; ++rep) { bench<false><<<148, 384, smem>>>(N, iters, d, mode); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("cta_group::1  M=128 N=%3d K=16: %.1f cycles per MMA (%s)\n", N, (double)h / (iters * 6), cudaGetErrorString(cudaGetLastError()));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(148); cfg.blockDim = dim3(384); cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) { cudaLaunchKernelEx(&cfg, bench<true>, N, iters, d, mode); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("cta_group::2  M=256 N=%3d K=16: %.1f cycles per MMA (%s)\n", N, (double)h / (iters * 6), cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
