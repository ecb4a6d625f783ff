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

// W [N][K] fp32 -> bf16 hi/lo planes in the tensor-core kernel's shared-memory image: the plane is cut
// into blocks of (up to) 256 rows x 32 k, each stored contiguously as [rows][64 B] with the
// 16-byte units of a row XOR-swizzled by (row >> 1) & 3 (UMMA SWIZZLE_64B, K-major), blocks ordered
// n-tile major, k-chunk minor, so tc_linear.cu fetches a block with one cp.async.bulk.
// ldd = K padded to 64; rows are padded to 64 (zeros come from the memset of the packed buffer).
__global__ void pack_split_kernel(const float* __restrict__ src, int N, int K, uint8_t* __restrict__ hi,
                                  uint8_t* __restrict__ lo, int ldd) {
  const long long total = (long long)N * K;
  const int npad = (N + 63) / 64 * 64;
  const int nch = ldd >> 5;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / K), k = (int)(i - (long long)n * K);
    const float w = src[i];
    const __nv_bfloat16 h = __float2bfloat16_rn(w);
    const __nv_bfloat16 l = __float2bfloat16_rn(w - __bfloat162float(h));
    const int n0 = n & ~255, r = n & 255;
    const int bn = min(256, npad - n0);
    const int c = k >> 5, kk = k & 31;
    const size_t off = ((size_t)n0 * nch + (size_t)c * bn) * 64 + sw64(r, kk >> 3) + (kk & 7) * 2;
    *reinterpret_cast<__nv_bfloat16*>(hi + off) = h;
    *reinterpret_cast<__nv_bfloat16*>(lo + off) = l;
  }
}

// Element (row, k) of the concatenated LSTM matrix [W_ih' | W_hh] (row in nn.LSTM order, gates stacked
// i,f,g,o), scaled by -log2(e) (i,f,o) or -2.log2(e) (g) so that the gate pre-activation is directly the
// ex2 argument of the cell update. fc_in (bsrnn.py:69,80) is folded into the layer-0 input weights in
// double precision: W_ih.(W_in.x + b_in) = (W_ih.W_in).x + W_ih.b_in.
__device__ double lstm_gate_scale(int gate) { return (gate == 2 ? -2.0 : -1.0) * 1.4426950408889634; }
__device__ float lstm_cat_weight(int row, int k, int kin, const float* __restrict__ w_ih,
                                 const float* __restrict__ w_hh, const float* __restrict__ w_in) {
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
  return (float)(w * lstm_gate_scale(row / kFeat));
}
__device__ float lstm_cat_bias(int row, const float* __restrict__ w_ih, const float* __restrict__ b_ih,
                               const float* __restrict__ b_hh, const float* __restrict__ w_in,
                               const float* __restrict__ b_in) {
  double b = (double)b_ih[row] + (double)b_hh[row];
  if (w_in != nullptr)
    for (int j = 0; j < kFeat; ++j) b += (double)w_ih[row * kFeat + j] * (double)b_in[j];
  return (float)(b * lstm_gate_scale(row / kFeat));
}

// Tensor-core LSTM layer image (layout.h RnnBlock::tcw/tcb). Block n = 4*unit + gate, thread k of
// the concatenated K axis [x part (kin) | h part (64)].
__global__ void pack_lstm_tc_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                                    const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                    const float* __restrict__ w_in, const float* __restrict__ b_in, int kin,
                                    uint8_t* __restrict__ img, float* __restrict__ bias) {
  const int n = blockIdx.x, k = threadIdx.x;
  const int unit = n >> 2, gate = n & 3;
  const int row = gate * kFeat + unit;                 // nn.LSTM row: gates stacked i,f,g,o
  const int nch = (kin + kFeat) / 64;
  const float wf = lstm_cat_weight(row, k, kin, w_ih, w_hh, w_in);
  const __nv_bfloat16 h = __float2bfloat16_rn(wf);
  const __nv_bfloat16 l = __float2bfloat16_rn(wf - __bfloat162float(h));
  const int c = k >> 6, kk = k & 63;
  const size_t off = ((size_t)c * kGates + n) * 128 + ((((kk >> 3) ^ (n & 7))) << 4) + (kk & 7) * 2;
  const size_t plane = (size_t)nch * kGates * 128;
  *reinterpret_cast<__nv_bfloat16*>(img + off) = h;
  *reinterpret_cast<__nv_bfloat16*>(img + plane + off) = l;
  if (k == 0) bias[n] = lstm_cat_bias(row, w_ih, b_ih, b_hh, w_in, b_in);
}

// mma.sync fragment image of the same matrix for the time-axis kernel (lstm_seq2.cu, kin = 64):
// block (warp w, k-tile kt, plane p), thread (lane, reg): the m16n8k16 A fragment word `reg` of `lane`
// for the 16 gate rows of warp w (rows 0-7: i,f of units 4w..4w+3 interleaved; rows 8-15: g,o).
__device__ __forceinline__ int seq_row_of(int w, int rm) {
  const int unit = 4 * w + ((rm & 7) >> 1);
  const int gate = (rm < 8 ? 0 : 2) + (rm & 1);
  return gate * kFeat + unit;
}
__global__ void pack_lstm_seq_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                                     const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                     const float* __restrict__ w_in, const float* __restrict__ b_in,
                                     uint32_t* __restrict__ img, float* __restrict__ bias) {
  const int w = blockIdx.x, kt = blockIdx.y, p = blockIdx.z;
  const int lane = threadIdx.x >> 2, reg = threadIdx.x & 3;
  const int rm = (lane >> 2) + ((reg & 1) ? 8 : 0);
  const int k = kt * 16 + 2 * (lane & 3) + ((reg & 2) ? 8 : 0);
  const int row = seq_row_of(w, rm);
  float v[2];
  for (int e = 0; e < 2; ++e) {
    const float wf = lstm_cat_weight(row, k + e, kFeat, w_ih, w_hh, w_in);
    const float hi = __bfloat162float(__float2bfloat16_rn(wf));
    v[e] = p == 0 ? hi : wf - hi;
  }
  const __nv_bfloat162 pk = __floats2bfloat162_rn(v[0], v[1]);
  img[(((size_t)w * 8 + kt) * 2 + p) * 128 + lane * 4 + reg] = *reinterpret_cast<const uint32_t*>(&pk);
  if (kt == 0 && p == 0 && threadIdx.x < 16)
    bias[w * 16 + threadIdx.x] = lstm_cat_bias(seq_row_of(w, threadIdx.x), w_ih, b_ih, b_hh, w_in, b_in);
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

cudaError_t launch_pack_lstm_seq(const float* w_ih, const float* w_hh, const float* b_ih, const float* b_hh,
                                 const float* w_in, const float* b_in, void* img, float* bias, cudaStream_t s) {
  pack_lstm_seq_kernel<<<dim3(16, 8, 2), 128, 0, s>>>(w_ih, w_hh, b_ih, b_hh, w_in, b_in,
                                                      reinterpret_cast<uint32_t*>(img), bias);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_pack_split(const float* src, int N, int K, void* dst_hi, void* dst_lo, int ldd, cudaStream_t s) {
  const long long total = (long long)N * K;
  if (total == 0) return cudaSuccess;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 1184) blocks = 1184;
  pack_split_kernel<<<blocks, 256, 0, s>>>(src, N, K, reinterpret_cast<uint8_t*>(dst_hi),
                                           reinterpret_cast<uint8_t*>(dst_lo), ldd);
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
