// Microbenchmark 2: tcgen05.mma (SS, M=128,N=256,K=16) throughput under interference from other warps.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__global__ void __launch_bounds__(288, 1) bench(int mode, int iters, const float4* g, long long n4, long long* out, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar, bar2;
  __shared__ uint32_t slot;
  __shared__ volatile int done;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&bar2), 1); fence_mbar_init(); done = 0; }
  if (threadIdx.x < 32) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc_bf16(128, 256);
    const uint64_t a = make_sw128_desc(base), b = make_sw128_desc(base + 16384);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) mma_bf16_ss(tm, a + kk * 2, b + kk * 2, idesc, 1);
    }
    mma_commit(smem_u32(&bar));
    mbar_wait(smem_u32(&bar), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
    done = 1;
  } else if (warp >= 1) {
    float acc = 0.f;
    uint8_t* scratch = gen + 49152 + 1024;          // 64 KB region away from the operands
    long long idx = (long long)blockIdx.x * 100003 + threadIdx.x;
    uint32_t ph = 0;
    while (!done) {
      if (mode == 1) {                                // STS.128 traffic (8 warps)
#pragma unroll
        for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(scratch + ((threadIdx.x - 32) * 16 + j * 4096)) = make_uint4(j, j, j, j);
      } else if (mode == 2) {                         // bulk copies 32 KB global -> smem
        if (threadIdx.x == 32) {
          mbar_arrive_expect_tx(smem_u32(&bar2), 32768);
          bulk_g2s(smem_u32(scratch), g + (idx % (n4 - 4096)), 32768, smem_u32(&bar2));
          mbar_wait(smem_u32(&bar2), ph); ph ^= 1; idx += 7919;
        }
      } else if (mode == 3) {                         // LDG.128 streaming (8 warps)
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float4 v = __ldg(g + ((idx + j * 256) % n4)); acc += v.x + v.w; }
        idx += 8 * 256 * 37;
      } else if (mode == 4) {                         // tcgen05.ld of other TMEM columns (8 warps)
        uint32_t r[32];
        tmem_ld_32x32(tm + ((uint32_t)((warp & 3) * 32) << 16) + 256 + (warp >> 2) * 32, r);
        tmem_ld_wait();
        acc += __uint_as_float(r[lane & 31]);
      } else if (mode == 5) {                         // LDS.128 traffic
#pragma unroll
        for (int j = 0; j < 8; ++j) { const uint4 v = *reinterpret_cast<const uint4*>(scratch + ((threadIdx.x - 32) * 16 + j * 4096)); acc += __uint_as_float(v.x); }
      } else {
        __nanosleep(100);
      }
    }
    if (acc == 123.456f) sink[0] = acc;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  float* sink; cudaMalloc(&sink, 4);
  const long long n4 = 64LL << 20;   // 1 GiB of float4
  float4* g; cudaMalloc(&g, n4 * 16); cudaMemset(g, 0, n4 * 16);
  const int smem = 49152 + 1024 + 65536 + 1024;
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const char* names[] = {"idle", "STS.128 x8 warps", "bulk 32KB copies", "LDG.128 stream x8 warps", "tcgen05.ld x8 warps", "LDS.128 x8 warps"};
  for (int mode = 0; mode < 6; ++mode) {
    for (int rep = 0; rep < 2; ++rep) { bench<<<148, 288, smem>>>(mode, 2000, g, n4, d, sink); cudaDeviceSynchronize(); }
    long long h; cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-26s : %.1f cycles/MMA (%s)\n", names[mode], (double)h / 8000, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
