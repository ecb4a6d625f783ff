// Probe of the CTA-pair MMA (tcgen05 cta_group::2): D[256 x N] = A[256 x 32] . B[N x 32]^T with A rows and
// B rows split across the two CTAs of a cluster, for N in {64, 256}. Checks the operand/accumulator mapping
// against the CPU. SWIZZLE_64B K-major operands as in tc_linear.cu.
#include <cstdio>
#include <cstdlib>
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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int N) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_done, bar_peer;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const int tid = threadIdx.x;
  const uint32_t rank = cluster_ctarank();
  if (tid == 0) { mbar_init(smem_u32(&bar_done), 1); mbar_init(smem_u32(&bar_peer), 1); fence_mbar_init(); }
  if (tid < 32) { tmem_alloc_pair(smem_u32(&slot), 256); tmem_relinquish_pair(); }
  // own 128 rows of A, own half of the N rows of B
  for (int i = tid; i < 128 * 32; i += 128) {
    const int r = i / 32, k = i % 32;
    *reinterpret_cast<__nv_bfloat16*>(gen + sw64(r, k >> 3) + (k & 7) * 2) = A[(rank * 128 + r) * 32 + k];
  }
  const int nh = N / 2;
  for (int i = tid; i < nh * 32; i += 128) {
    const int r = i / 32, k = i % 32;
    *reinterpret_cast<__nv_bfloat16*>(gen + 8192 + sw64(r, k >> 3) + (k & 7) * 2) = B[(rank * nh + r) * 32 + k];
  }
  fence_proxy_async_smem();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tm = slot;
  if (rank == 1 && tid == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&bar_peer), 0));   // "my operands are in place"
  if (rank == 0 && tid == 0) {
    mbar_wait_cluster(smem_u32(&bar_peer), 0);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(256, N);
    const uint64_t da = desc64(base), db = desc64(base + 8192);
    mma_bf16_ss_pair(tm, da, db, idesc, 0);
    mma_bf16_ss_pair(tm, da + 2, db + 2, idesc, 1);
    mma_commit_pair(smem_u32(&bar_done), 3);
  }
  mbar_wait(smem_u32(&bar_done), 0);
  tc_fence_after();
  const int q = tid >> 5, lane = tid & 31;
  for (int cb = 0; cb < N / 32; ++cb) {
    uint32_t r[32];
    tmem_ld_32x32(tm + ((uint32_t)(q * 32) << 16) + cb * 32, r);
    tmem_ld_wait32(r);
    for (int j = 0; j < 32; ++j) D[(rank * 128 + q * 32 + lane) * N + cb * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  cluster_sync_all();
  if (tid < 32) { tc_fence_after(); tmem_dealloc_pair(tm, 256); }
}

int main() {
  static __nv_bfloat16 hA[256 * 32], hB[256 * 32];
  static float ref[256 * 256], hD[256 * 256];
  srand(2);
  for (int i = 0; i < 256 * 32; ++i) { hA[i] = __float2bfloat16((float)(rand() % 7 - 3)); hB[i] = __float2bfloat16((float)(rand() % 5 - 2)); }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, sizeof(hA)); cudaMalloc(&dB, sizeof(hB)); cudaMalloc(&dD, sizeof(hD));
  cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, sizeof(hB), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int N : {64, 128, 256}) {
    for (int m = 0; m < 256; ++m) for (int n = 0; n < N; ++n) {
      float s = 0; for (int k = 0; k < 32; ++k) s += __bfloat162float(hA[m * 32 + k]) * __bfloat162float(hB[n * 32 + k]);
      ref[m * N + n] = s;
    }
    cudaMemset(dD, 0, sizeof(hD));
    probe<<<2, 128, 64 * 1024>>>(dA, dB, dD, N);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N %d: %s\n", N, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(hD, dD, 256 * N * 4, cudaMemcpyDeviceToHost);
    float err = 0; int bad = 0;
    for (int i = 0; i < 256 * N; ++i) { const float d = fabsf(hD[i] - ref[i]); if (d > 0) ++bad; err = fmaxf(err, d); }
    printf("pair MMA M=256 N=%d: max |err| = %g, %d wrong of %d %s\n", N, err, bad, 256 * N, err == 0.f ? "  <== exact" : "");
    if (err != 0.f) {   // where do the wrong entries sit?
      int rows_bad[4] = {0, 0, 0, 0}, cols_bad[2] = {0, 0};
      for (int m = 0; m < 256; ++m) for (int n = 0; n < N; ++n) if (hD[m * N + n] != ref[m * N + n]) { rows_bad[m / 64]++; cols_bad[n >= N / 2]++; }
      printf("   wrong by row quarter: %d %d %d %d ; by column half: %d %d\n", rows_bad[0], rows_bad[1], rows_bad[2], rows_bad[3], cols_bad[0], cols_bad[1]);
    }
  }
  return 0;
}
