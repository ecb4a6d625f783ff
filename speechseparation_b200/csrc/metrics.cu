// Evaluation sums of the forward-only consumer m_dataset.train_infer (m_dataset.py:202-226):
// everything that function reports -- L1 in time, L1 of Re and Im, SDR, SI-SDR -- is a function of five sums
// per row over an (estimate, reference) pair, gathered here in ONE pass over both arrays (HBM-bound: 8 B per
// element pair, 16-byte loads, fp32 partials per thread, fp64 from the warp level up, one fp64 atomicAdd per
// quantity per CTA).
#include "kernels.cuh"

namespace bsb {

constexpr int kStatThreads = 256;
constexpr int kStatChunk = 16384;        // elements of one row per CTA

__device__ __forceinline__ void stat_acc(float e, float r, float (&a)[5]) {
  const float d = e - r;
  a[0] += fabsf(d);          // sum |e - r|
  a[1] = fmaf(r, r, a[1]);   // sum r^2
  a[2] = fmaf(d, d, a[2]);   // sum (e - r)^2
  a[3] = fmaf(e, r, a[3]);   // sum e r
  a[4] = fmaf(e, e, a[4]);   // sum e^2
}

__global__ void __launch_bounds__(kStatThreads)
pair_stats_kernel(const float* __restrict__ est, long long est_stride, const float* __restrict__ ref,
                  long long ref_stride, long long n, double* __restrict__ stats) {
  const int row = blockIdx.y;
  const float* e = est + row * est_stride;
  const float* r = ref + row * ref_stride;
  const long long lo = (long long)blockIdx.x * kStatChunk;
  const long long hi = min(n, lo + kStatChunk);
  float a[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
  const bool vec = (((reinterpret_cast<uintptr_t>(e + lo) | reinterpret_cast<uintptr_t>(r + lo)) & 15) == 0);
  if (vec) {
    const long long nv = (hi - lo) >> 2;
    const float4* e4 = reinterpret_cast<const float4*>(e + lo);
    const float4* r4 = reinterpret_cast<const float4*>(r + lo);
    for (long long i = threadIdx.x; i < nv; i += kStatThreads) {
      const float4 x = __ldg(e4 + i), y = __ldg(r4 + i);
      stat_acc(x.x, y.x, a); stat_acc(x.y, y.y, a); stat_acc(x.z, y.z, a); stat_acc(x.w, y.w, a);
    }
    for (long long i = lo + (nv << 2) + threadIdx.x; i < hi; i += kStatThreads) stat_acc(e[i], r[i], a);
  } else {
    for (long long i = lo + threadIdx.x; i < hi; i += kStatThreads) stat_acc(e[i], r[i], a);
  }
  __shared__ double red[kStatThreads / 32][5];
  double d[5];
#pragma unroll
  for (int q = 0; q < 5; ++q) {
    d[q] = (double)a[q];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) d[q] += __shfl_xor_sync(0xffffffffu, d[q], o);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int q = 0; q < 5; ++q) red[threadIdx.x >> 5][q] = d[q];
  }
  __syncthreads();
  if (threadIdx.x < 5) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kStatThreads / 32; ++w) s += red[w][threadIdx.x];
    atomicAdd(stats + row * 5 + threadIdx.x, s);
  }
}

cudaError_t launch_pair_stats(const float* est, long long est_stride, const float* ref, long long ref_stride,
                              int rows, long long n, double* stats, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(stats, 0, sizeof(double) * 5 * rows, s);
  if (e != cudaSuccess) return e;
  const long long chunks = (n + kStatChunk - 1) / kStatChunk;
  dim3 grid((unsigned)chunks, (unsigned)rows);
  pair_stats_kernel<<<grid, kStatThreads, 0, s>>>(est, est_stride, ref, ref_stride, n, stats);
  count_launch();
  return cudaGetLastError();
}

// sdr2 of m_dataset.train_infer as the reference actually reduces it (m_dataset.py:219-222): `sample` keeps its
// leading batch-1 dimension there, so `dim=1` of [1][C][L] is the CHANNEL axis: per sample l,
//   q(l) = 10 log10( (sum_c s[c][l]^2 + 1e-9) / (sum_c (s[c][l] - m[c][l])^2 + 1e-9) ),   sdr2 = mean_l q(l).
// out[0] (double, zeroed by the launch) = sum_l q(l).
__global__ void __launch_bounds__(kStatThreads)
channel_sdr_kernel(const float* __restrict__ mix, long long mix_stride, const float* __restrict__ src,
                   long long src_stride, int rows, long long n, double* __restrict__ out) {
  double acc = 0.0;
  for (long long l = (long long)blockIdx.x * kStatThreads + threadIdx.x; l < n; l += (long long)gridDim.x * kStatThreads) {
    float ss = 0.f, dd = 0.f;
    for (int c = 0; c < rows; ++c) {
      const float sv = __ldg(src + c * src_stride + l), mv = __ldg(mix + c * mix_stride + l);
      ss = fmaf(sv, sv, ss);
      dd = fmaf(sv - mv, sv - mv, dd);
    }
    acc += 10.0 * (double)log10f((ss + 1e-9f) / (dd + 1e-9f));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ double red[kStatThreads / 32];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kStatThreads / 32; ++w) s += red[w];
    atomicAdd(out, s);
  }
}

cudaError_t launch_channel_sdr(const float* mix, long long mix_stride, const float* src, long long src_stride,
                               int rows, long long n, double* out, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(out, 0, sizeof(double), s);
  if (e != cudaSuccess) return e;
  long long blocks = (n + kStatThreads - 1) / kStatThreads;
  if (blocks > 148 * 8) blocks = 148 * 8;
  channel_sdr_kernel<<<(unsigned)blocks, kStatThreads, 0, s>>>(mix, mix_stride, src, src_stride, rows, n, out);
  count_launch();
  return cudaGetLastError();
}

// 16-bit PCM <-> float at the host-buffer boundary (what a wav file holds before torchaudio.load, infer.py:24-27):
// halves the host<->device bytes of the bulk path. in: x = s / 32768 (exactly torchaudio's normalisation);
// out: s = clamp(rint(x * 32768), -32768, 32767). HBM-bound elementwise passes, 8 samples per thread.
__global__ void __launch_bounds__(256)
pcm16_to_float_kernel(const short* __restrict__ pcm, long long n, float* __restrict__ wav) {
  const long long i8 = ((long long)blockIdx.x * 256 + threadIdx.x) * 8;
  if (i8 + 8 <= n && ((reinterpret_cast<uintptr_t>(pcm) | reinterpret_cast<uintptr_t>(wav)) & 15) == 0) {
    const uint4 p = __ldg(reinterpret_cast<const uint4*>(pcm + i8));
    const unsigned w[4] = {p.x, p.y, p.z, p.w};
    float o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[2 * k] = (float)(short)(w[k] & 0xffffu) * (1.f / 32768.f);
      o[2 * k + 1] = (float)(short)(w[k] >> 16) * (1.f / 32768.f);
    }
    *reinterpret_cast<float4*>(wav + i8) = make_float4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<float4*>(wav + i8 + 4) = make_float4(o[4], o[5], o[6], o[7]);
  } else {
    for (long long i = i8; i < n && i < i8 + 8; ++i) wav[i] = (float)pcm[i] * (1.f / 32768.f);
  }
}

__device__ __forceinline__ unsigned to_pcm16(float x) {
  const float y = fminf(fmaxf(rintf(x * 32768.f), -32768.f), 32767.f);
  return (unsigned)(int)y & 0xffffu;
}

__global__ void __launch_bounds__(256)
float_to_pcm16_kernel(const float* __restrict__ wav, long long n, short* __restrict__ pcm) {
  const long long i8 = ((long long)blockIdx.x * 256 + threadIdx.x) * 8;
  if (i8 + 8 <= n && ((reinterpret_cast<uintptr_t>(pcm) | reinterpret_cast<uintptr_t>(wav)) & 15) == 0) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(wav + i8)), b = __ldg(reinterpret_cast<const float4*>(wav + i8 + 4));
    uint4 o;
    o.x = to_pcm16(a.x) | (to_pcm16(a.y) << 16);
    o.y = to_pcm16(a.z) | (to_pcm16(a.w) << 16);
    o.z = to_pcm16(b.x) | (to_pcm16(b.y) << 16);
    o.w = to_pcm16(b.z) | (to_pcm16(b.w) << 16);
    *reinterpret_cast<uint4*>(pcm + i8) = o;
  } else {
    for (long long i = i8; i < n && i < i8 + 8; ++i) pcm[i] = (short)to_pcm16(wav[i]);
  }
}

cudaError_t launch_pcm16_to_float(const short* pcm, long long n, float* wav, cudaStream_t s) {
  const long long blocks = (n + 2047) / 2048;
  pcm16_to_float_kernel<<<(unsigned)blocks, 256, 0, s>>>(pcm, n, wav);
  count_launch();
  return cudaGetLastError();
}
cudaError_t launch_float_to_pcm16(const float* wav, long long n, short* pcm, cudaStream_t s) {
  const long long blocks = (n + 2047) / 2048;
  float_to_pcm16_kernel<<<(unsigned)blocks, 256, 0, s>>>(wav, n, pcm);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
