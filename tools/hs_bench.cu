// Handshake costs on one SM: tcgen05.commit -> mbarrier -> waiting thread -> mbarrier.arrive -> waiting thread,
// with a ring of D stages and no tensor work (what the GEMM main loop costs besides its MMAs).
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/kernels.cuh"
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb;
using namespace bsb::ptx;

// mode 0: consumer commits empty[s]; loader waits empty[s], arrives full[s]  (the kernel's loop without copies)
// mode 1: consumer does a plain mbarrier.arrive on empty[s] instead of tcgen05.commit
// mode 2: like 0 with three loader threads (full count 3)
__global__ void __launch_bounds__(384, 1) hs(int iters, int D, int mode, int spin, int one, long long* out) {
  __shared__ __align__(8) uint64_t bar_full[8], bar_empty[8];
  __shared__ uint32_t slot;
  __shared__ uint32_t ts[128], tl[128], tp[32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nload = mode == 2 ? 3 : 1;
  if (tid == 0) {
    for (int i = 0; i < 8; ++i) { mbar_init(smem_u32(&bar_full[i]), nload); mbar_init(smem_u32(&bar_empty[i]), 1); }
    fence_mbar_init();
  }
  if (warp == 0) { tmem_alloc(smem_u32(&slot), 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (warp == 0) {
    int st = 0; uint32_t ph = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      if (lane == 0 && i >= 1000 && i < 1128) ts[i - 1000] = (uint32_t)clock();
      if (!one || lane == 0) { if (spin) mbar_spin(smem_u32(&bar_full[st]), ph); else mbar_wait(smem_u32(&bar_full[st]), ph); }
      if (one) __syncwarp();
      if (lane == 0 && i >= 1000 && i < 1006) tp[4 * (i - 1000)] = (uint32_t)clock();
      tc_fence_after();
      if (lane == 0) {
        if (mode == 1) mbar_arrive(smem_u32(&bar_empty[st])); else mma_commit(smem_u32(&bar_empty[st]));
        if (i >= 1000 && i < 1006) tp[1 + 4 * (i - 1000)] = (uint32_t)clock();
      }
      __syncwarp();
      if (lane == 0 && i >= 1000 && i < 1006) tp[2 + 4 * (i - 1000)] = (uint32_t)clock();
      if (++st == D) { st = 0; ph ^= 1; }
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0 && lane == 0) out[0] = t1 - t0;
    __syncwarp();
    if (blockIdx.x == 0 && lane == 0) for (int i = 0; i < 24; ++i) { out[1 + i] = i < 6 ? ts[i] - ts[0] : tp[4 * ((i - 6) / 3) + (i - 6) % 3] - ts[0]; }
  } else if (warp >= 1 && warp <= nload && lane == 0) {
    int st = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      if (spin) mbar_spin(smem_u32(&bar_empty[st]), ph ^ 1); else mbar_wait(smem_u32(&bar_empty[st]), ph ^ 1);
      mbar_arrive(smem_u32(&bar_full[st]));
      if (warp == 1 && i >= 1000 && i < 1128) tl[i - 1000] = (uint32_t)clock();
      if (++st == D) { st = 0; ph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (blockIdx.x == 0 && tid == 0) for (int i = 0; i < 24; ++i) out[32 + i] = tl[i] - ts[0];
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 8 * 64);
  const int iters = 2000;
  for (int grid : {1, 148})
  for (int threads : {64, 384})
  for (int one = 0; one < 2; ++one)
  for (int mode = 0; mode < 2; ++mode)
    for (int spin = 0; spin < 1; ++spin)
      for (int D : {4}) {
        long long h; long long hh[64];
        for (int rep = 0; rep < 2; ++rep) { hs<<<grid, threads>>>(iters, D, mode, spin, one, d); cudaDeviceSynchronize(); }
        cudaMemcpy(hh, d, 8 * 64, cudaMemcpyDeviceToHost); h = hh[0];
        if (0) { printf("  consumer: 6 loop starts, then (waited, committed, synced) x6:"); for (int i = 0; i < 24; ++i) printf(" %lld", hh[1 + i]); printf("\n  loader arrives:      "); for (int i = 0; i < 24; ++i) printf(" %lld", hh[32 + i]); printf("\n"); }
        printf("grid %d threads %d %s mode %d %s depth %d: %.1f cycles per chunk (%s)\n", grid, threads, one ? "lane 0 waits" : "32 lanes wait", mode, spin ? "test_wait" : "try_wait ", D, (double)h / iters,
               cudaGetErrorString(cudaGetLastError()));
      }
  return 0;
}
