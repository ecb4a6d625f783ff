// Weight repacking: 288 state_dict tensors (bsrnn.py:328-376) -> one padded device buffer
// (layout.h). Runs once per load_state_dict / .to(); not on the hot path.
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

__global__ void pack_matrix_kernel(const float* __restrict__ src, int N, int K, float* __restrict__ dst, int ldd) {
  const long long total = (long long)N * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / K), k = (int)(i - (long long)n * K);
    dst[(long long)n * ldd + k] = src[i];
  }
}

__global__ void pack_bias_kernel(const float* __restrict__ b1, const float* __restrict__ b2, int N,
                                 float* __restrict__ dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < N) dst[i] = b2 ? b1[i] + b2[i] : b1[i];
}

__global__ void pack_split_kernel(const float* __restrict__ src, int N, int K, __nv_bfloat16* __restrict__ hi,
                                  __nv_bfloat16* __restrict__ lo, int ldd) {
  const long long total = (long long)N * K;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / K), k = (int)(i - (long long)n * K);
    const float w = src[i];
    const __nv_bfloat16 h = __float2bfloat16_rn(w);
    hi[(long long)n * ldd + k] = h;
    lo[(long long)n * ldd + k] = __float2bfloat16_rn(w - __bfloat162float(h));
  }
}

}  // namespace

cudaError_t launch_pack_split(const float* src, int N, int K, void* dst_hi, void* dst_lo, int ldd, cudaStream_t s) {
  const long long total = (long long)N * K;
  if (total == 0) return cudaSuccess;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  pack_split_kernel<<<blocks, 256, 0, s>>>(src, N, K, reinterpret_cast<__nv_bfloat16*>(dst_hi),
                                           reinterpret_cast<__nv_bfloat16*>(dst_lo), ldd);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_pack_matrix(const float* src, int N, int K, float* dst, int ldd, cudaStream_t s) {
  const long long total = (long long)N * K;
  if (total == 0) return cudaSuccess;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  pack_matrix_kernel<<<blocks, 256, 0, s>>>(src, N, K, dst, ldd);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_pack_bias(const float* b1, const float* b2, int N, float* dst, cudaStream_t s) {
  if (N == 0) return cudaSuccess;
  pack_bias_kernel<<<(N + 255) / 256, 256, 0, s>>>(b1, b2, N, dst);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
