// STFT / iSTFT for the model's fixed framing: n_fft 2048, hop 1024, periodic Hann.
//
// A 2048-point real transform is done as one 1024-point complex Stockham radix-4 FFT in
// shared memory (5 passes, 256 threads = one butterfly each per pass) plus an O(N) split
// step, one frame per CTA.  Window multiply and reflect padding are fused into the load,
// the Re/Im interleave of m_dataset.py:189-190 into the store.  Twiddles come from tables
// computed in double precision on the host.
//
//   analysis  : torch.stft(center=True, pad_mode="reflect", onesided, unnormalised)
//               (m_dataset.py:188; infer.py:31)  and torch.fft.rfft(buf*win) (infer-streaming.py:119)
//   synthesis : torch.istft (m_dataset.py:195): irfft, * window, overlap-add, / sum(window^2), trim
//               and torch.fft.irfft + 2-frame OLA / (w[i]+w[i+1024]) (infer-streaming.py:127-145)
#include <cmath>
#include <mutex>
#include <vector>
#include "kernels.cuh"

namespace bsb {
namespace {

constexpr int kN = 2048;     // n_fft
constexpr int kM = 1024;     // complex FFT size
constexpr int kHop = 1024;
constexpr int kFftThreads = 256;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// In-place (ping-pong) 1024-point complex FFT on smem buffers. Input in `a`, returns the
// buffer holding the result (natural order). INV uses conjugate twiddles (unnormalised).
template <bool INV>
__device__ __forceinline__ float2* fft1024(float2* a, float2* b, const float2* __restrict__ tw) {
  const int j = threadIdx.x;           // butterfly index 0..255
#pragma unroll
  for (int ns = 1; ns < kM; ns *= 4) {
    const int k = j & (ns - 1);
    float2 v0 = a[j], v1 = a[j + 256], v2 = a[j + 512], v3 = a[j + 768];
    if (ns > 1) {
      const int tstep = k * (256 / ns);
      float2 t1 = __ldg(tw + tstep), t2 = __ldg(tw + 2 * tstep), t3 = __ldg(tw + 3 * tstep);
      if (INV) { t1.y = -t1.y; t2.y = -t2.y; t3.y = -t3.y; }
      v1 = cmul(v1, t1); v2 = cmul(v2, t2); v3 = cmul(v3, t3);
    }
    const float2 a0 = make_float2(v0.x + v2.x, v0.y + v2.y);
    const float2 a1 = make_float2(v0.x - v2.x, v0.y - v2.y);
    const float2 a2 = make_float2(v1.x + v3.x, v1.y + v3.y);
    const float2 d = make_float2(v1.x - v3.x, v1.y - v3.y);
    // forward: (v1-v3)*(-i) = (d.y, -d.x); inverse: (v1-v3)*(+i) = (-d.y, d.x)
    const float2 a3 = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);
    const int j0 = ((j - k) << 2) + k;      // (j/ns)*ns*4 + k
    b[j0] = make_float2(a0.x + a2.x, a0.y + a2.y);
    b[j0 + ns] = make_float2(a1.x + a3.x, a1.y + a3.y);
    b[j0 + 2 * ns] = make_float2(a0.x - a2.x, a0.y - a2.y);
    b[j0 + 3 * ns] = make_float2(a1.x - a3.x, a1.y - a3.y);
    __syncthreads();
    float2* t = a; a = b; b = t;
  }
  return a;
}

// Split step: Z = FFT_1024(x[2n] + i x[2n+1])  ->  X[k], k = 0..1024, written interleaved.
__device__ __forceinline__ void rfft_split_store(const float2* Z, const float2* __restrict__ tw2, float* out) {
  for (int k = threadIdx.x; k <= kM; k += kFftThreads) {
    const float2 zk = Z[k & (kM - 1)];
    const float2 zm = Z[(kM - k) & (kM - 1)];
    // E = (Zk + conj(Zm))/2 ; O = -i/2 (Zk - conj(Zm))
    const float2 e = make_float2(0.5f * (zk.x + zm.x), 0.5f * (zk.y - zm.y));
    const float2 o = make_float2(0.5f * (zk.y + zm.y), -0.5f * (zk.x - zm.x));
    const float2 w = __ldg(tw2 + k);
    const float2 wo = cmul(w, o);
    *reinterpret_cast<float2*>(out + 2 * k) = make_float2(e.x + wo.x, e.y + wo.y);
  }
}

// Inverse split: X[k] (interleaved, imag of DC/Nyquist ignored) -> Z[k], k = 0..1023, such
// that ifft_1024(Z)/1024 = x[2n] + i x[2n+1].
__device__ __forceinline__ void irfft_split_load(const float* in, const float2* __restrict__ tw2, float2* Z) {
  for (int k = threadIdx.x; k < kM; k += kFftThreads) {
    float2 xk = *reinterpret_cast<const float2*>(in + 2 * k);
    float2 xm = *reinterpret_cast<const float2*>(in + 2 * (kM - k));
    if (k == 0) { xk.y = 0.f; xm.y = 0.f; }
    // E = (Xk + conj(Xm))/2 ; O = (Xk - conj(Xm))/2 * conj(w^k) ; Z = E + i O
    const float2 e = make_float2(0.5f * (xk.x + xm.x), 0.5f * (xk.y - xm.y));
    const float2 dlt = make_float2(0.5f * (xk.x - xm.x), 0.5f * (xk.y + xm.y));
    float2 w = __ldg(tw2 + k);
    w.y = -w.y;
    const float2 o = cmul(dlt, w);
    Z[k] = make_float2(e.x - o.y, e.y + o.x);
  }
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFftThreads)
stft_frames_kernel(const float* __restrict__ wav, long long L, int T, const float* __restrict__ win,
                   const float2* __restrict__ tw, float* __restrict__ xf, int xs) {
  __shared__ __align__(16) float2 sa[kM];
  __shared__ __align__(16) float2 sb[kM];
  const int r = blockIdx.x;              // frame-row c*T + t
  const int c = r / T, t = r - c * T;
  const float* w0 = wav + (long long)c * L;
  const long long base = (long long)t * kHop - kHop;   // centre padding of n_fft/2
  for (int n = threadIdx.x; n < kM; n += kFftThreads) {
    float v[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      long long j = base + 2 * n + e;
      if (j < 0) j = -j;                                   // reflect (no edge repeat)
      if (j >= L) j = 2 * (L - 1) - j;
      v[e] = __ldg(w0 + j) * __ldg(win + 2 * n + e);
    }
    sa[n] = make_float2(v[0], v[1]);
  }
  __syncthreads();
  const float2* Z = fft1024<false>(sa, sb, tw);
  rfft_split_store(Z, tw + kM, xf + (long long)r * xs);
}

// One CTA produces `blocks_per_cta` consecutive 1024-sample output blocks of one row.
// Output block b (samples [1024b, 1024b+1024) of the trimmed signal) = second half of frame b
// + first half of frame b+1, each irfft'ed and multiplied by the window, divided by
// w^2[i] + w^2[i+1024] (torch.istft's window envelope; every kept sample has both frames).
__global__ void __launch_bounds__(kFftThreads)
istft_frames_kernel(const float* __restrict__ xf, int xs, int T, const float* __restrict__ win,
                    const float2* __restrict__ tw, float* __restrict__ wav, int blocks_per_cta) {
  __shared__ __align__(16) float2 sa[kM];
  __shared__ __align__(16) float2 sb[kM];
  __shared__ float prev[kHop];           // windowed second half of the previous frame
  const int c = blockIdx.y;
  const int b0 = blockIdx.x * blocks_per_cta;
  const int nb = min(blocks_per_cta, (T - 1) - b0);
  if (nb <= 0) return;
  const long long out_len = (long long)kHop * (T - 1);
  const float inv_m = 1.0f / kM;
  for (int f = 0; f <= nb; ++f) {
    const int t = b0 + f;
    irfft_split_load(xf + ((long long)c * T + t) * xs, tw + kM, sa);
    __syncthreads();
    const float2* z = fft1024<true>(sa, sb, tw);
    // z[n] * inv_m = (x[2n], x[2n+1]); first half = n < 512, second half = n >= 512
    if (f > 0) {
      float* o = wav + (long long)c * out_len + (long long)(t - 1) * kHop;
      for (int n = threadIdx.x; n < kM / 2; n += kFftThreads) {
        const float2 v = z[n];
        const int i = 2 * n;
        const float w0 = __ldg(win + i), w1 = __ldg(win + i + 1);
        const float u0 = __ldg(win + i + kHop), u1 = __ldg(win + i + 1 + kHop);
        const float y0 = (prev[i] + v.x * inv_m * w0) / (u0 * u0 + w0 * w0);
        const float y1 = (prev[i + 1] + v.y * inv_m * w1) / (u1 * u1 + w1 * w1);
        *reinterpret_cast<float2*>(o + i) = make_float2(y0, y1);
      }
      __syncthreads();                   // prev consumed before it is overwritten
    }
    if (f < nb) {
      for (int n = threadIdx.x; n < kM / 2; n += kFftThreads) {
        const float2 v = z[n + kM / 2];
        const int i = 2 * n;
        prev[i] = v.x * inv_m * __ldg(win + kHop + i);
        prev[i + 1] = v.y * inv_m * __ldg(win + kHop + i + 1);
      }
    }
    __syncthreads();                     // z (smem) is reused by the next frame's load
  }
}

// ------------------------------------------------------------------------------------------
// streaming (infer-streaming.py:116-121 and :124-145), one CTA per row
__global__ void __launch_bounds__(kFftThreads)
stream_analysis_kernel(const float* __restrict__ chunk, const float* __restrict__ win,
                       const float2* __restrict__ tw, float* __restrict__ ring, float* __restrict__ x) {
  __shared__ __align__(16) float2 sa[kM];
  __shared__ __align__(16) float2 sb[kM];
  const int c = blockIdx.x;
  float* rg = ring + (long long)c * kN;
  const float* ch = chunk + (long long)c * kHop;
  // new ring = [old ring second half, chunk]
  float2 nv[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int n = threadIdx.x + q * kFftThreads;         // 0..511 : pair index within a half
    nv[q] = *reinterpret_cast<const float2*>(rg + kHop + 2 * n);
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int n = threadIdx.x + q * kFftThreads;
    const float2 nw = *reinterpret_cast<const float2*>(ch + 2 * n);
    *reinterpret_cast<float2*>(rg + 2 * n) = nv[q];
    *reinterpret_cast<float2*>(rg + kHop + 2 * n) = nw;
    sa[n] = make_float2(nv[q].x * __ldg(win + 2 * n), nv[q].y * __ldg(win + 2 * n + 1));
    sa[n + kM / 2] = make_float2(nw.x * __ldg(win + kHop + 2 * n), nw.y * __ldg(win + kHop + 2 * n + 1));
  }
  __syncthreads();
  const float2* Z = fft1024<false>(sa, sb, tw);
  rfft_split_store(Z, tw + kM, x + (long long)c * kSpec);
}

__global__ void __launch_bounds__(kFftThreads)
stream_synthesis_kernel(const float* __restrict__ y, const float* __restrict__ win,
                        const float2* __restrict__ tw, float* __restrict__ tail, float* __restrict__ out) {
  __shared__ __align__(16) float2 sa[kM];
  __shared__ __align__(16) float2 sb[kM];
  const int c = blockIdx.x;
  irfft_split_load(y + (long long)c * kSpec, tw + kM, sa);
  __syncthreads();
  const float2* z = fft1024<true>(sa, sb, tw);
  const float inv_m = 1.0f / kM;
  float* tl = tail + (long long)c * kHop;
  float* o = out + (long long)c * kHop;
  for (int n = threadIdx.x; n < kM / 2; n += kFftThreads) {
    const int i = 2 * n;
    const float2 v = z[n];
    const float2 v2 = z[n + kM / 2];
    const float2 p = *reinterpret_cast<const float2*>(tl + i);
    // sum_of_window = 0 + w[i] + w[i+1024] ; new = 0 + cur[i] + prev[i+1024]
    const float s0 = __ldg(win + i) + __ldg(win + i + kHop);
    const float s1 = __ldg(win + i + 1) + __ldg(win + i + 1 + kHop);
    *reinterpret_cast<float2*>(o + i) = make_float2((v.x * inv_m + p.x) / s0, (v.y * inv_m + p.y) / s1);
    *reinterpret_cast<float2*>(tl + i) = make_float2(v2.x * inv_m, v2.y * inv_m);
  }
}

// ------------------------------------------------------------------------------------------
// boundary <-> frame-major transposes (32x32 tiles through padded smem)
__global__ void __launch_bounds__(256)
to_frames_kernel(const float* __restrict__ x, int T, float* __restrict__ xf, int xs) {
  __shared__ float tile[32][33];
  const int c = blockIdx.z;
  const int t0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* xc = x + (long long)c * kSpec * T;
  for (int i = ty; i < 32; i += 8) {
    const int f = f0 + i, t = t0 + tx;
    tile[i][tx] = (f < kSpec && t < T) ? xc[(long long)f * T + t] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, f = f0 + tx;
    if (t < T && f < kSpec) xf[((long long)c * T + t) * xs + f] = tile[tx][i];
  }
}

__global__ void __launch_bounds__(256)
from_frames_kernel(const float* __restrict__ xf, int xs, int T, float* __restrict__ x) {
  __shared__ float tile[32][33];
  const int c = blockIdx.z;
  const int t0 = blockIdx.x * 32, f0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int i = ty; i < 32; i += 8) {
    const int t = t0 + i, f = f0 + tx;
    tile[i][tx] = (t < T && f < kSpec) ? xf[((long long)c * T + t) * xs + f] : 0.f;
  }
  __syncthreads();
  float* xc = x + (long long)c * kSpec * T;
  for (int i = ty; i < 32; i += 8) {
    const int f = f0 + i, t = t0 + tx;
    if (f < kSpec && t < T) xc[(long long)f * T + t] = tile[tx][i];
  }
}

// ------------------------------------------------------------------------------------------
// constant tables (per device, created on first use; tiny, live for the process lifetime)
struct Tables {
  float* window = nullptr;   // [2048]
  float2* tw = nullptr;      // [1024] W_1024^k, then [1025] W_2048^k
};
std::mutex g_tab_mu;
Tables g_tab[64];

cudaError_t ensure_tables(Tables** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_tab_mu);
  Tables& t = g_tab[dev];
  if (t.tw == nullptr) {
    const double pi = 3.14159265358979323846;
    std::vector<float> w(kN);
    for (int n = 0; n < kN; ++n) w[n] = (float)(0.5 - 0.5 * std::cos(2.0 * pi * n / kN));
    std::vector<float2> tw(kM + kM + 1);
    for (int k = 0; k < kM; ++k) tw[k] = make_float2((float)std::cos(2.0 * pi * k / kM), (float)-std::sin(2.0 * pi * k / kM));
    for (int k = 0; k <= kM; ++k) tw[kM + k] = make_float2((float)std::cos(2.0 * pi * k / kN), (float)-std::sin(2.0 * pi * k / kN));
    float* dw = nullptr;
    float2* dt = nullptr;
    if ((e = cudaMalloc(&dw, w.size() * sizeof(float))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&dt, tw.size() * sizeof(float2))) != cudaSuccess) return e;
    // synchronous copies: table creation happens once, outside any stream capture
    if ((e = cudaMemcpy(dw, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    if ((e = cudaMemcpy(dt, tw.data(), tw.size() * sizeof(float2), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    t.window = dw;
    t.tw = dt;
  }
  *out = &t;
  return cudaSuccess;
}

}  // namespace

const float* builtin_window(cudaStream_t, cudaError_t* err) {
  Tables* t = nullptr;
  *err = ensure_tables(&t);
  return *err == cudaSuccess ? t->window : nullptr;
}
const float2* twiddle_tables(cudaStream_t, cudaError_t* err) {
  Tables* t = nullptr;
  *err = ensure_tables(&t);
  return *err == cudaSuccess ? t->tw : nullptr;
}

#define BSB_TABLES(win_var, tw_var)                                   \
  cudaError_t terr_;                                                  \
  const float2* tw_var = twiddle_tables(s, &terr_);                   \
  if (terr_ != cudaSuccess) return terr_;                             \
  if (win_var == nullptr) {                                           \
    win_var = builtin_window(s, &terr_);                              \
    if (terr_ != cudaSuccess) return terr_;                           \
  }

cudaError_t launch_stft_frames(const float* wav, int C, long long L, int T, const float* window,
                               float* xf, int xs, cudaStream_t s) {
  BSB_TABLES(window, tw);
  stft_frames_kernel<<<C * T, kFftThreads, 0, s>>>(wav, L, T, window, tw, xf, xs);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_istft_frames(const float* xf, int xs, int C, int T, const float* window,
                                float* wav, cudaStream_t s) {
  if (T < 2) return cudaSuccess;
  BSB_TABLES(window, tw);
  // enough CTAs for >= 2 waves of 148 SMs x 4 resident CTAs when the clip allows it
  int per = ((long long)(T - 1) * C + 1183) / 1184;
  if (per < 1) per = 1;
  if (per > 64) per = 64;
  dim3 grid((T - 1 + per - 1) / per, C);
  istft_frames_kernel<<<grid, kFftThreads, 0, s>>>(xf, xs, T, window, tw, wav, per);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_stream_analysis(const float* chunk, int C, const float* window, float* ring,
                                   float* x, cudaStream_t s) {
  BSB_TABLES(window, tw);
  stream_analysis_kernel<<<C, kFftThreads, 0, s>>>(chunk, window, tw, ring, x);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_stream_synthesis(const float* y, int C, const float* window, float* tail,
                                    float* out, cudaStream_t s) {
  BSB_TABLES(window, tw);
  stream_synthesis_kernel<<<C, kFftThreads, 0, s>>>(y, window, tw, tail, out);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_to_frames(const float* x, int C, int T, float* xf, int xs, cudaStream_t s) {
  dim3 grid((T + 31) / 32, (kSpec + 31) / 32, C);
  to_frames_kernel<<<grid, 256, 0, s>>>(x, T, xf, xs);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_from_frames(const float* xf, int xs, int C, int T, float* x, cudaStream_t s) {
  dim3 grid((T + 31) / 32, (kSpec + 31) / 32, C);
  from_frames_kernel<<<grid, 256, 0, s>>>(xf, xs, T, x);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
