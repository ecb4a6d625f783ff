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

// Tensor-core LSTM layer image (layout.h RnnBlock::tcw/tcb). Block n = 4*unit + gate, thread k of
// the concatenated K axis [x part (kin) | h part (64)]. fc_in (bsrnn.py:69,80) is folded into the
// layer-0 input weights in double precision: W_ih.(W_in.x + b_in) = (W_ih.W_in).x + W_ih.b_in.
__global__ void pack_lstm_tc_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                                    const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                    const float* __restrict__ w_in, const float* __restrict__ b_in, int kin,
                                    uint8_t* __restrict__ img, float* __restrict__ bias) {
  const int n = blockIdx.x, k = threadIdx.x;
  const int unit = n >> 2, gate = n & 3;
  const int row = gate * kFeat + unit;                 // nn.LSTM row: gates stacked i,f,g,o
  const double scale = (gate == 2 ? -2.0 : -1.0) * 1.4426950408889634;
  const int nch = (kin + kFeat) / 64;
  double w;
  if (k < kin) {
    if (w_in != nullptr) {
      w = 0.0;
      for (int j = 0; j < kFeat; ++j) w += (double)w_ih[row * kFeat + j] * (double)w_in[j * kFeat + k];
    } else {
      w = (double)w_ih[row * kin + k];
    }
  } else {
    w = (double)w_hh[row * kFeat + (k - kin)];
  }
  const float wf = (float)(w * scale);
  const __nv_bfloat16 h = __float2bfloat16_rn(wf);
  const __nv_bfloat16 l = __float2bfloat16_rn(wf - __bfloat162float(h));
  const int c = k >> 6, kk = k & 63;
  const size_t off = ((size_t)c * kGates + n) * 128 + ((((kk >> 3) ^ (n & 7))) << 4) + (kk & 7) * 2;
  const size_t plane = (size_t)nch * kGates * 128;
  *reinterpret_cast<__nv_bfloat16*>(img + off) = h;
  *reinterpret_cast<__nv_bfloat16*>(img + plane + off) = l;
  if (k == 0) {
    double b = (double)b_ih[row] + (double)b_hh[row];
    if (w_in != nullptr)
      for (int j = 0; j < kFeat; ++j) b += (double)w_ih[row * kFeat + j] * (double)b_in[j];
    bias[n] = (float)(b * scale);
  }
}

}  // namespace

cudaError_t launch_pack_lstm_tc(const float* w_ih, const float* w_hh, const float* b_ih, const float* b_hh,
                                const float* w_in, const float* b_in, int kin, void* img, float* bias,
                                cudaStream_t s) {
  pack_lstm_tc_kernel<<<kGates, kin + kFeat, 0, s>>>(w_ih, w_hh, b_ih, b_hh, w_in, b_in, kin,
                                                     reinterpret_cast<uint8_t*>(img), bias);
  count_launch();
  return cudaGetLastError();
}

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
