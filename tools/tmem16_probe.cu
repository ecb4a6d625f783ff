// Probe: register layout of tcgen05.ld.16x256b.x1 and which 16-lane windows a warp may address.
// Every TMEM cell (lane l, column c) is first filled with l * 1000 + c through 32x32b stores; each warp then
// reads 16 lanes x 8 columns at lane offsets 32*(warp%4) + {0, 16} and prints what every thread received.
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

__global__ void __launch_bounds__(128, 1) probe(uint32_t* out) {
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (w == 0) { tmem_alloc(smem_u32(&slot), 32); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = slot;
  uint32_t v[8];
  for (int c0 = 0; c0 < 32; c0 += 8) {
    for (int i = 0; i < 8; ++i) v[i] = (uint32_t)((w * 32 + lane) * 1000 + c0 + i);
    tmem_st_32x8(base + ((uint32_t)(w * 32) << 16) + c0, v);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int half = 0; half < 2; ++half) {
    uint32_t r0, r1, r2, r3;
    const uint32_t taddr = base + ((uint32_t)(w * 32 + half * 16) << 16) + 8;     // columns 8..15
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];\n\ttcgen05.wait::ld.sync.aligned;"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(taddr) : "memory");
    uint32_t* o = out + ((w * 2 + half) * 32 + lane) * 4;
    o[0] = r0; o[1] = r1; o[2] = r2; o[3] = r3;
  }
  tc_fence_before();
  __syncthreads();
  if (w == 0) { tc_fence_after(); tmem_dealloc(base, 32); }
}

int main() {
  uint32_t* d;
  cudaMalloc(&d, 4 * 2 * 32 * 4 * 4);
  cudaMemset(d, 0xff, 4 * 2 * 32 * 4 * 4);
  probe<<<1, 128>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  static uint32_t h[4 * 2 * 32 * 4];
  cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int w = 0; w < 4; ++w)
    for (int half = 0; half < 2; ++half)
      for (int t = 0; t < 32; ++t) {
        const uint32_t* o = h + ((w * 2 + half) * 32 + t) * 4;
        // expected (mma m16n8 C fragment): regs 0,1 = row t/4, cols 2(t%4), +1 ; regs 2,3 = row t/4 + 8
        const int l0 = w * 32 + half * 16 + t / 4, c0 = 8 + 2 * (t % 4);
        const uint32_t want[4] = {(uint32_t)(l0 * 1000 + c0), (uint32_t)(l0 * 1000 + c0 + 1), (uint32_t)((l0 + 8) * 1000 + c0),
                                  (uint32_t)((l0 + 8) * 1000 + c0 + 1)};
        const bool ok = o[0] == want[0] && o[1] == want[1] && o[2] == want[2] && o[3] == want[3];
        bad += !ok;
        if (w == 1 || !ok) printf("warp %d half %d thread %2d: %u %u %u %u %s\n", w, half, t, o[0], o[1], o[2], o[3], ok ? "" : "<-- not the C-fragment layout");
      }
  printf("%s\n", bad ? "MISMATCH" : "16x256b.x1 == mma.m16n8 C fragment layout at lane offsets 32*(warp%4) + {0,16}");
  return 0;
}
