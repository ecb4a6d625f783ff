// Probe: which shared-memory image does a K-major SWIZZLE_64B tcgen05 operand descriptor read?
// One CTA computes D[128 x 64] = A[128 x 32] . B[64 x 32]^T (bf16, exact small integers) for a set of
// candidate (xor pattern, stride byte offset, layout code) and reports the max error against the CPU.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, int sbo, int layout) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(sbo >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}
__device__ __forceinline__ int unit_of(int r, int c, int mode) {     // 16-byte unit c (0..3) of row r -> physical unit
  if (mode == 0) return c ^ ((r >> 1) & 3);
  if (mode == 1) return c ^ (r & 3);
  return c;
}

__global__ void __launch_bounds__(128, 1) probe(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, int mode, int sbo, int layout) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));
  const int tid = threadIdx.x;
  if (tid == 0) { mbar_init(smem_u32(&bar), 1); fence_mbar_init(); }
  if (tid < 32) { tmem_alloc(smem_u32(&slot), 64); tmem_relinquish(); }
  // rows are 64 B; 8-row groups are `sbo` bytes apart
  for (int i = tid; i < 128 * 32; i += 128) {
    const int r = i / 32, k = i % 32;
    const int off = (r >> 3) * sbo + (r & 7) * 64 + unit_of(r & 7, k >> 3, mode) * 16 + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(gen + off) = A[i];
  }
  uint8_t* genb = gen + 16384;
  for (int i = tid; i < 64 * 32; i += 128) {
    const int r = i / 32, k = i % 32;
    const int off = (r >> 3) * sbo + (r & 7) * 64 + unit_of(r & 7, k >> 3, mode) * 16 + (k & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(genb + off) = B[i];
  }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tm = slot;
  if (tid == 0) {
    const uint32_t idesc = make_idesc_bf16(128, 64);
    const uint64_t da = make_desc(base, sbo, layout), db = make_desc(base + 16384, sbo, layout);
    mma_bf16_ss(tm, da, db, idesc, 0);
    mma_bf16_ss(tm, da + 2, db + 2, idesc, 1);      // second k16 step: +32 bytes
    mma_commit(smem_u32(&bar));
  }
  mbar_wait(smem_u32(&bar), 0);
  tc_fence_after();
  const int q = tid >> 5, lane = tid & 31;
  for (int cb = 0; cb < 2; ++cb) {
    uint32_t r[32];
    tmem_ld_32x32(tm + ((uint32_t)(q * 32) << 16) + cb * 32, r);
    tmem_ld_wait32(r);
    for (int j = 0; j < 32; ++j) D[(q * 32 + lane) * 64 + cb * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before(); __syncthreads();
  if (tid < 32) { tc_fence_after(); tmem_dealloc(tm, 64); }
}

int main() {
  __nv_bfloat16 hA[128 * 32], hB[64 * 32];
  float ref[128 * 64], hD[128 * 64];
  srand(1);
  for (int i = 0; i < 128 * 32; ++i) hA[i] = __float2bfloat16((float)(rand() % 7 - 3));
  for (int i = 0; i < 64 * 32; ++i) hB[i] = __float2bfloat16((float)(rand() % 5 - 2));
  for (int m = 0; m < 128; ++m) for (int n = 0; n < 64; ++n) {
    float s = 0; for (int k = 0; k < 32; ++k) s += __bfloat162float(hA[m * 32 + k]) * __bfloat162float(hB[n * 32 + k]);
    ref[m * 64 + n] = s;
  }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, sizeof(hA)); cudaMalloc(&dB, sizeof(hB)); cudaMalloc(&dD, sizeof(hD));
  cudaMemcpy(dA, hA, sizeof(hA), cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, sizeof(hB), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int layout : {4, 6, 2}) for (int sbo : {512, 1024}) for (int mode : {0, 1, 2}) {
    cudaMemset(dD, 0, sizeof(hD));
    probe<<<1, 128, 64 * 1024>>>(dA, dB, dD, mode, sbo, layout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("layout %d sbo %d mode %d: %s\n", layout, sbo, mode, cudaGetErrorString(e)); return 1; }
    cudaMemcpy(hD, dD, sizeof(hD), cudaMemcpyDeviceToHost);
    float err = 0; for (int i = 0; i < 128 * 64; ++i) err = fmaxf(err, fabsf(hD[i] - ref[i]));
    printf("layout code %d, SBO %4d, xor mode %d : max |err| = %g %s\n", layout, sbo, mode, err, err == 0.f ? "  <== exact" : "");
  }
  return 0;
}
