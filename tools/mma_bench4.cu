// Microbenchmark 4: how much LDG / STS progress do 8 warps make while the tensor core runs SS/TS MMAs?
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__global__ void __launch_bounds__(288, 1) bench(int mma_mode, int work, int iters, const float4* g, long long n4, long long* out, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  __shared__ unsigned long long count;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); done = 0; count = 0; }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, 256);
    const uint64_t a_hi = make_sw128_desc(base), a_lo = make_sw128_desc(base + 16384);
    const uint64_t w_hi = make_sw128_desc(base + 32768), w_lo = make_sw128_desc(base + 65536);
    long long t0 = clock64();
    if (mma_mode == 0) {
      while (clock64() - t0 < (long long)iters * 12 * 142) {}
    } else {
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adv = (uint64_t)(kk * 2);
          if (mma_mode == 1) {
            mma_bf16_ss(tm, a_lo + adv, w_hi + adv, idesc, 1);
            mma_bf16_ss(tm, a_hi + adv, w_lo + adv, idesc, 1);
            mma_bf16_ss(tm, a_hi + adv, w_hi + adv, idesc, 1);
          } else {
            mma_bf16_ts(tm, tm + 384 + 32 + kk * 8, w_hi + adv, idesc, 1);
            mma_bf16_ts(tm, tm + 384 + kk * 8, w_lo + adv, idesc, 1);
            mma_bf16_ts(tm, tm + 384 + kk * 8, w_hi + adv, idesc, 1);
          }
        }
      }
      mma_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), 0);
    }
    long long t1 = clock64();
    done = 1;
    if (blockIdx.x == 0) out[0] = t1 - t0;
  } else if (threadIdx.x >= 32) {
    float acc = 0.f;
    unsigned long long my = 0;
    uint8_t* scratch = gen + 98304 + 1024;
    long long base4 = (long long)blockIdx.x * 8 * 256;   // coalesced: 256 threads x 16 B consecutive, 8 rows
    while (!done) {
      if (work == 0) {            // LDG.128 stream, 8 independent loads per thread per iteration
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __ldg(g + ((base4 + j * 256 + (threadIdx.x - 32)) % n4));
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += v[j].x;
        base4 += 8 * 256 * 148;
      } else {                    // STS.128, 8 per thread per iteration
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(scratch + ((threadIdx.x - 32) * 16 + j * 4096)) = make_uint4(j, j, j, j);
      }
      ++my;
    }
    if (acc == 123.456f) sink[0] = acc;
    if ((threadIdx.x & 31) == 0) atomicAdd(&count, my);
  }
  __syncthreads();
  if (threadIdx.x == 0 && blockIdx.x == 0) out[1] = (long long)count;
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  float* sink; cudaMalloc(&sink, 4);
  const long long n4 = 64LL << 20;
  float4* g; cudaMalloc(&g, n4 * 16); cudaMemset(g, 0, n4 * 16);
  const int smem = 98304 + 1024 + 32768 + 1024;
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* mm[] = {"no MMA (spin)", "SS MMAs", "TS MMAs"};
  const char* wk[] = {"LDG.128 stream", "STS.128"};
  for (int work = 0; work < 2; ++work) for (int mode = 0; mode < 3; ++mode) {
    for (int rep = 0; rep < 2; ++rep) { bench<<<148, 288, smem>>>(mode, work, 400, g, n4, d, sink); cudaDeviceSynchronize(); }
    long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    const double bytes = (double)h[1] * 8 * 32 * 16;     // per warp-iteration: 8 x 32 lanes x 16 B
    printf("%-16s | %-14s : window %lld cycles, %.1f cycles/MMA, %.1f B/cycle/SM by the 8 warps (%s)\n", wk[work], mm[mode], h[0],
           (double)h[0] / 4800, bytes / h[0], cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
