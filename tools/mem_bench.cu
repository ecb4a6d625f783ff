// Microbenchmark 5: global->SM streaming rate per SM as a function of bytes in flight and of the
// access method (LDG.128 into registers vs cp.async.bulk into shared memory), all 148 SMs active.
//   src = 1 GB (DRAM resident) or 48 MB (L2 resident after the first pass)
#include <cstdio>
#include <cuda_runtime.h>
#include "../speechseparation_b200/csrc/ptx.cuh"
using namespace bsb::ptx;

template <int U>
__global__ void __launch_bounds__(512, 1) ldg_stream(const float4* g, long long n4, int iters, float* sink, long long* out) {
  float acc = 0.f;
  const int nthr = blockDim.x;
  long long base = (long long)blockIdx.x * U * nthr;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    float4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) v[j] = __ldg(g + ((base + j * nthr + threadIdx.x) % n4));
#pragma unroll
    for (int j = 0; j < U; ++j) acc += v[j].x;
    base += (long long)U * nthr * gridDim.x;
  }
  const long long t1 = clock64();
  if (acc == 123.456f) sink[0] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
}

// one thread keeps `stages` bulk copies of `bytes` each in flight
__global__ void __launch_bounds__(128, 1) bulk_stream(const uint8_t* g, long long nbytes, int stages, int bytes, int pieces, int iters, long long* out, int shared_src = 0) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar[8];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if (threadIdx.x == 0) { for (int s = 0; s < 8; ++s) mbar_init(smem_u32(&bar[s]), 1); fence_mbar_init(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long off = shared_src ? (long long)(blockIdx.x % shared_src) * bytes : (long long)blockIdx.x * bytes;   // shared_src: groups of SMs walk the same addresses
    const long long stride = shared_src ? (long long)shared_src * bytes : (long long)gridDim.x * bytes;
    const int pb = bytes / pieces;
    auto issue = [&](int s) {
      mbar_arrive_expect_tx(smem_u32(&bar[s]), bytes);
      for (int p = 0; p < pieces; ++p)
        bulk_g2s(base + s * bytes + p * pb, g + (off % nbytes) + p * pb, pb, smem_u32(&bar[s]));
      off += stride;
    };
    const long long t0 = clock64();
    for (int s = 0; s < stages; ++s) issue(s);
    int s = 0; uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(smem_u32(&bar[s]), ph);
      if (it + stages < iters) issue(s);
      if (++s == stages) { s = 0; ph ^= 1; }
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
}

// `nissue` threads (lane 0 of warps 0..nissue-1, or lanes 0..nissue-1 of warp 0 when same_warp) each keep their own
// ring of `stages` bulk copies of `bytes` in flight: are requests of different threads served concurrently?
__global__ void __launch_bounds__(256, 1) bulk_multi(const uint8_t* g, long long nbytes, int nissue, int same_warp, int stages, int bytes, int iters,
                                                     long long* out) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar[8][8];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if (threadIdx.x == 0) { for (int i = 0; i < 64; ++i) mbar_init(smem_u32(&bar[i / 8][i % 8]), 1); fence_mbar_init(); }
  __syncthreads();
  const int me = same_warp ? (int)threadIdx.x : (int)(threadIdx.x >> 5);
  const bool issuer = same_warp ? (threadIdx.x < nissue) : ((threadIdx.x & 31) == 0 && me < nissue);
  if (issuer) {
    long long off = ((long long)blockIdx.x * nissue + me) * bytes;
    const long long stride = (long long)gridDim.x * nissue * bytes;
    const uint32_t mybase = base + me * stages * bytes;
    auto issue = [&](int s) {
      mbar_arrive_expect_tx(smem_u32(&bar[me][s]), bytes);
      bulk_g2s(mybase + s * bytes, g + (off % nbytes), bytes, smem_u32(&bar[me][s]));
      off += stride;
    };
    const long long t0 = clock64();
    for (int s = 0; s < stages; ++s) issue(s);
    int s = 0; uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(smem_u32(&bar[me][s]), ph);
      if (it + stages < iters) issue(s);
      if (++s == stages) { s = 0; ph ^= 1; }
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0 && me == 0) out[0] = t1 - t0;
  }
}

int main() {
  long long* d; cudaMalloc(&d, 16);
  float* sink; cudaMalloc(&sink, 4);
  const long long big = 1LL << 30, small = 48LL << 20;
  uint8_t* g; cudaMalloc(&g, big); cudaMemset(g, 0, big);
  long long h;
  for (long long nb : {big, small}) {
    printf("---- source %lld MB (%s)\n", nb >> 20, nb == big ? "DRAM" : "L2 resident");
    const long long n4 = nb / 16;
#define RUN_LDG(U, T)                                                                                   \
    for (int rep = 0; rep < 2; ++rep) { ldg_stream<U><<<148, T>>>((const float4*)g, n4, 400, sink, d); cudaDeviceSynchronize(); } \
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);                                                       \
    printf("LDG.128  %2d warps x %2d per thread = %3d KB in flight : %6.1f B/clk/SM, period %lld cycles (%s)\n", T / 32, U, U * T * 16 / 1024, \
           400.0 * U * T * 16 / h, h / 400, cudaGetErrorString(cudaGetLastError()));
    RUN_LDG(4, 256) RUN_LDG(8, 256) RUN_LDG(16, 256) RUN_LDG(8, 512) RUN_LDG(16, 512) RUN_LDG(24, 512)
    cudaFuncSetAttribute(bulk_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int bytes : {16384, 32768}) for (int stages : {1, 2, 4, 6}) for (int pieces : {1, 64}) {
      if (stages * bytes > 196608) continue;
      for (int rep = 0; rep < 2; ++rep) { bulk_stream<<<148, 128, 200 * 1024>>>(g, nb, stages, bytes, pieces, 600, d); cudaDeviceSynchronize(); }
      cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
      printf("bulk     %d x %2d KB (%2d pieces of %5d B) = %3d KB in flight : %6.1f B/clk/SM (%s)\n", stages, bytes >> 10, pieces, bytes / pieces,
             stages * bytes >> 10, 600.0 * bytes / h, cudaGetErrorString(cudaGetLastError()));
    }
  }
  printf("---- request size (L2 resident source), pieces = 1\n");
  for (int bytes : {4096, 8192, 16384, 32768, 65536}) for (int stages : {2, 3, 6, 8}) {
    if (stages * bytes > 196608 || stages > 8) continue;
    for (int rep = 0; rep < 2; ++rep) { bulk_stream<<<148, 128, 200 * 1024>>>(g, small, stages, bytes, 1, 600, d); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("bulk %d x %2d KB in flight : %6.1f B/clk/SM, %6.0f cycles per request (%s)\n", stages, bytes >> 10, 600.0 * bytes / h, (double)h / 600,
           cudaGetErrorString(cudaGetLastError()));
  }
  printf("---- issuing threads (L2 resident source), 2 requests in flight per thread\n");
  cudaFuncSetAttribute(bulk_multi, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  for (int same_warp : {0, 1}) for (int bytes : {8192, 16384}) for (int nissue : {1, 2, 4}) {
    if (nissue * 2 * bytes > 196608) continue;
    for (int rep = 0; rep < 2; ++rep) { bulk_multi<<<148, 256, 200 * 1024>>>(g, small, nissue, same_warp, 2, bytes, 600, d); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%d issuing %s x 2 x %2d KB : %6.1f B/clk/SM, %6.0f cycles per request per thread (%s)\n", nissue, same_warp ? "lanes of one warp" : "warps (lane 0)",
           bytes >> 10, 600.0 * bytes * nissue / h, (double)h / 600, cudaGetErrorString(cudaGetLastError()));
  }
  printf("---- every SM streams the same addresses (weights): 4 MB region, 2 x 32 KB in flight\n");
  for (int groups : {1, 2, 4, 16, 148}) {
    for (int rep = 0; rep < 2; ++rep) { bulk_stream<<<148, 128, 200 * 1024>>>(g, 4LL << 20, 2, 32768, 1, 600, d, groups); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("bulk 2 x 32 KB, %3d distinct address streams over 148 SMs : %6.1f B/clk/SM (%s)\n", groups, 600.0 * 32768 / h, cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
