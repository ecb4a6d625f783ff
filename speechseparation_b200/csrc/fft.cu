// STFT / iSTFT for the model's fixed framing: n_fft 2048, hop 1024, periodic Hann.
//
// A 2048-point real transform is done as one 1024-point complex FFT plus an O(N) split step.
// Offline (throughput) kernels: one warp per frame, two in-register 32-point FFTs around one
// shared-memory transpose. Streaming (latency) kernels: one CTA per row, Stockham radix-4 in
// shared memory (5 passes, 256 threads). Window multiply and reflect padding are fused into
// the load, the Re/Im interleave of m_dataset.py:189-190 into the store. Twiddles come from
// tables computed in double precision on the host.
//
//   analysis  : torch.stft(center=True, pad_mode="reflect", onesided, unnormalised)
//               (m_dataset.py:188; infer.py:31)  and torch.fft.rfft(buf*win) (infer-streaming.py:119)
//   synthesis : torch.istft (m_dataset.py:195): irfft, * window, overlap-add, / sum(window^2), trim
//               and torch.fft.irfft + 2-frame OLA / (w[i]+w[i+1024]) (infer-streaming.py:127-145)
#include <cmath>
#include <mutex>
#include <vector>
#include "kernels.cuh"
#include "ptx.cuh"

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

// Same with the plugin's dry/wet control applied to the spectrum first (speech-ladspa-onnx.cpp:215-226):
// mix >= 0: mix*y + (1-mix)*x ; mix < 0: x + mix*y   (x = the analysed spectrum of this hop).
__device__ __forceinline__ float2 mix_bin(float2 y, float2 x, float mix) {
  const float a = mix >= 0.f ? 1.f - mix : 1.f;
  return make_float2(fmaf(mix, y.x, a * x.x), fmaf(mix, y.y, a * x.y));
}
__device__ __forceinline__ void irfft_split_load_mix(const float* in, const float* dry, float mix,
                                                     const float2* __restrict__ tw2, float2* Z) {
  for (int k = threadIdx.x; k < kM; k += kFftThreads) {
    float2 xk = mix_bin(*reinterpret_cast<const float2*>(in + 2 * k),
                        *reinterpret_cast<const float2*>(dry + 2 * k), mix);
    float2 xm = mix_bin(*reinterpret_cast<const float2*>(in + 2 * (kM - k)),
                        *reinterpret_cast<const float2*>(dry + 2 * (kM - k)), mix);
    if (k == 0) { xk.y = 0.f; xm.y = 0.f; }
    const float2 e = make_float2(0.5f * (xk.x + xm.x), 0.5f * (xk.y - xm.y));
    const float2 dlt = make_float2(0.5f * (xk.x - xm.x), 0.5f * (xk.y + xm.y));
    float2 w = __ldg(tw2 + k);
    w.y = -w.y;
    const float2 o = cmul(dlt, w);
    Z[k] = make_float2(e.x - o.y, e.y + o.x);
  }
}

// ------------------------------------------------------------------------------------------
// Offline path: one WARP per frame. 1024 = 32 x 32: every lane runs a 32-point FFT in registers,
// the warp transposes once through (padded) shared memory, every lane runs a second 32-point
// FFT. No block barrier, 16 KB of smem traffic per frame instead of ~100 KB, and every global
// access is a 256-byte run of consecutive lanes: the kernels are bound by HBM.
//   z[n], n = l + 32 j ;  k = k1 + 32 k2 :
//   X[k1 + 32 k2] = sum_l W32^(l k2) * ( W1024^(l k1) * sum_j z[l + 32 j] W32^(j k1) )
constexpr int kFrameWarps = 4;                 // frames (warps) per CTA
constexpr int kTr = 33;                        // transpose row stride (float2)
constexpr int kTw32 = 2 * 1024 + 1;            // offset of the [32][32] inter-stage twiddle table in the twiddle array

__device__ __forceinline__ constexpr int br5(int i) {
  return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i & 8) >> 2) | ((i & 16) >> 4);
}
__device__ __forceinline__ float cos32(int m) {   // cos(2 pi m / 32), m = 1..15 (folds after unrolling)
  switch (m) {
    case 1: return 0.98078528040323043f;  case 2: return 0.92387953251128674f;
    case 3: return 0.83146961230254524f;  case 4: return 0.70710678118654752f;
    case 5: return 0.55557023301960223f;  case 6: return 0.38268343236508977f;
    case 7: return 0.19509032201612827f;  case 8: return 0.f;
    case 9: return -0.19509032201612827f; case 10: return -0.38268343236508977f;
    case 11: return -0.55557023301960223f; case 12: return -0.70710678118654752f;
    case 13: return -0.83146961230254524f; case 14: return -0.92387953251128674f;
    case 15: return -0.98078528040323043f; default: return 1.f;
  }
}
__device__ __forceinline__ float sin32(int m) { return m <= 8 ? cos32(8 - m) : cos32(m - 8); }   // sin x = cos(x - pi/2)

// W_2048^(lane + 32 j) = W_2048^lane * W_64^j: one table load per lane and frame plus compile-time constants instead of
// 32 table loads (the FFT kernels are limited by L1 bytes per frame, not by FP issue slots).
__device__ __forceinline__ float2 w64(int j) {   // exp(-2 pi i j / 64), j = 0..31 (folds after unrolling)
  constexpr float c[32] = {1.0f, 0.995184727f, 0.98078528f, 0.956940336f, 0.923879533f, 0.881921264f, 0.831469612f, 0.773010453f, 0.707106781f, 0.634393284f, 0.555570233f, 0.471396737f, 0.382683432f, 0.290284677f, 0.195090322f, 0.0980171403f, 6.123234e-17f, -0.0980171403f, -0.195090322f, -0.290284677f, -0.382683432f, -0.471396737f, -0.555570233f, -0.634393284f, -0.707106781f, -0.773010453f, -0.831469612f, -0.881921264f, -0.923879533f, -0.956940336f, -0.98078528f, -0.995184727f};
  constexpr float sn[32] = {0.0f, 0.0980171403f, 0.195090322f, 0.290284677f, 0.382683432f, 0.471396737f, 0.555570233f, 0.634393284f, 0.707106781f, 0.773010453f, 0.831469612f, 0.881921264f, 0.923879533f, 0.956940336f, 0.98078528f, 0.995184727f, 1.0f, 0.995184727f, 0.98078528f, 0.956940336f, 0.923879533f, 0.881921264f, 0.831469612f, 0.773010453f, 0.707106781f, 0.634393284f, 0.555570233f, 0.471396737f, 0.382683432f, 0.290284677f, 0.195090322f, 0.0980171403f};
  return make_float2(c[j], -sn[j]);
}

// in-register 32-point DIF FFT; result X[br5(i)] = v[i]. INV: conjugate twiddles, unnormalised.
template <bool INV>
__device__ __forceinline__ void fft32(float2 (&v)[32]) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
#pragma unroll
    for (int b = 0; b < 32; b += 2 * s) {
#pragma unroll
      for (int k = 0; k < s; ++k) {
        const float2 a = v[b + k], c = v[b + k + s];
        v[b + k] = make_float2(a.x + c.x, a.y + c.y);
        const float2 d = make_float2(a.x - c.x, a.y - c.y);
        const int m = k * (16 / s);
        if (m == 0) {
          v[b + k + s] = d;
        } else if (m == 8) {
          v[b + k + s] = INV ? make_float2(-d.y, d.x) : make_float2(d.y, -d.x);
        } else {
          const float cs = cos32(m), sn = sin32(m);
          v[b + k + s] = INV ? make_float2(d.x * cs - d.y * sn, d.y * cs + d.x * sn)
                             : make_float2(d.x * cs + d.y * sn, d.y * cs - d.x * sn);
        }
      }
    }
  }
}

// v[j] = z[lane + 32 j] on entry; v[br5(k2)] = Z[lane + 32 k2] on return
template <bool INV>
__device__ __forceinline__ void warp_fft1024(float2 (&v)[32], float2* tr, const float2* __restrict__ tw, int lane) {
  fft32<INV>(v);
#pragma unroll
  for (int k1 = 0; k1 < 32; ++k1) {
    float2 y = v[br5(k1)];
    if (k1 > 0) {
      float2 w = __ldg(tw + kTw32 + k1 * 32 + lane);   // W_1024^(lane k1), stored k1-major: one 256-byte run per load
      if (INV) w.y = -w.y;
      y = cmul(y, w);
    }
    tr[k1 * kTr + lane] = y;
  }
  __syncwarp();
#pragma unroll
  for (int l = 0; l < 32; ++l) v[l] = tr[lane * kTr + l];
  __syncwarp();
  fft32<INV>(v);
}

// 5 CTAs (20 warps) per SM: 96 registers with 52 bytes of spill beats 128 registers at 16 warps (88.6 -> 79.6 us at
// cfg 2: the kernel waits on its frame loads, more resident warps hide them; 6 CTAs = 80 registers spills 400 bytes: 131 us)
__global__ void __launch_bounds__(kFrameWarps * 32, 5)
stft_warp_kernel(const float* __restrict__ wav, long long L, int T, int nrows, const float* __restrict__ win,
                 const float2* __restrict__ tw, float* __restrict__ xf, int xs) {
  __shared__ __align__(16) float2 tr_all[kFrameWarps][32 * kTr];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * kFrameWarps + warp;          // frame-row c*T + t
  if (r >= nrows) return;                                  // warp-uniform; no block barrier below
  const int c = r / T, t = r - c * T;
  const float* w0 = wav + (long long)c * L;
  const long long base = (long long)t * kHop - kHop;       // centre padding of n_fft/2
  float2 v[32];
  const bool fast = base >= 0 && base + kN <= L && ((reinterpret_cast<uintptr_t>(w0 + base) & 7) == 0);
  if (fast) {
    const float2* p = reinterpret_cast<const float2*>(w0 + base);
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float2 x = __ldg(p + lane + 32 * j);
      const float2 w = __ldg(reinterpret_cast<const float2*>(win) + lane + 32 * j);
      v[j] = make_float2(x.x * w.x, x.y * w.y);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int n = lane + 32 * j;
      float e[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        long long i = base + 2 * n + q;
        if (i < 0) i = -i;                                 // reflect (no edge repeat)
        if (i >= L) i = 2 * (L - 1) - i;
        e[q] = __ldg(w0 + i) * __ldg(win + 2 * n + q);
      }
      v[j] = make_float2(e[0], e[1]);
    }
  }
  warp_fft1024<false>(v, tr_all[warp], tw, lane);
  // split step: X[k] = E + W2048^k O with Z[k], Z[1024-k]; lane holds k = lane + 32 k2
  float* out = xf + (long long)r * xs;
  const float2 wl = __ldg(tw + kM + lane);                 // W_2048^lane
  const int partner = (32 - lane) & 31;
#pragma unroll
  for (int k2 = 0; k2 < 32; ++k2) {
    const float2 zk = v[br5(k2)];
    const float2 pv = v[br5(31 - k2)];
    float2 zm = make_float2(__shfl_sync(0xffffffffu, pv.x, partner), __shfl_sync(0xffffffffu, pv.y, partner));
    if (lane == 0) zm = v[br5((32 - k2) & 31)];
    const int k = lane + 32 * k2;
    const float2 e = make_float2(0.5f * (zk.x + zm.x), 0.5f * (zk.y - zm.y));
    const float2 o = make_float2(0.5f * (zk.y + zm.y), -0.5f * (zk.x - zm.x));
    const float2 wo = cmul(k2 == 0 ? wl : cmul(wl, w64(k2)), o);
    *reinterpret_cast<float2*>(out + 2 * k) = make_float2(e.x + wo.x, e.y + wo.y);
  }
  if (lane == 0) *reinterpret_cast<float2*>(out + 2 * kM) = make_float2(v[0].x - v[0].y, 0.f);
}

// A warp produces `per` consecutive 1024-sample output blocks of one row; the windowed second
// half of the previous frame stays in registers (same lane, same slot as the next first half).
// STAGED: the frame's spectrum row (8,200 B, padded to 8,208) is brought into shared memory by one bulk copy
// (cp.async.bulk, completion on a per-warp mbarrier) that is issued for frame f+1 as soon as frame f has been
// read into registers, so the HBM latency of a frame hides under the ~3k instructions of the previous one;
// the direct variant (any row stride / alignment) issues its loads at the top of the frame and waits for them.
constexpr int kStageFloats = 2052;             // 2050 rounded up to a multiple of 16 bytes
constexpr int kIstftWarpSmem = 32 * kTr * 8 + kStageFloats * 4;   // transpose tile + staged row, per warp

template <bool STAGED>
__global__ void __launch_bounds__(kFrameWarps * 32, STAGED ? 3 : 4)
istft_warp_kernel(const float* __restrict__ xf, int xs, int T, int nseg, int nwork, const float* __restrict__ win,
                  const float2* __restrict__ tw, float* __restrict__ wav, int per) {
  extern __shared__ __align__(16) uint8_t istft_smem[];
  __shared__ __align__(8) uint64_t bar_all[kFrameWarps];
  // 1 / (w[i]^2 + w[i+1024]^2) of the overlap-add, once per CTA instead of two loads, an FMA pair and a
  // reciprocal per output sample (same values: __fdividef(a, b) is a * rcp(b))
  __shared__ float2 inv_env[kHop / 2];
  {
    const float2* wp = reinterpret_cast<const float2*>(win);
    for (int n = threadIdx.x; n < kHop / 2; n += kFrameWarps * 32) {
      const float2 w = __ldg(wp + n), u = __ldg(wp + n + kHop / 2);
      inv_env[n] = make_float2(__fdividef(1.f, fmaf(u.x, u.x, w.x * w.x)), __fdividef(1.f, fmaf(u.y, u.y, w.y * w.y)));
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* tr = reinterpret_cast<float2*>(istft_smem + (size_t)warp * (STAGED ? kIstftWarpSmem : 32 * kTr * 8));
  const float* stage = reinterpret_cast<const float*>(tr + 32 * kTr);
  const uint32_t bar = ptx::smem_u32(&bar_all[warp]);
  const int wk = blockIdx.x * kFrameWarps + warp;          // (row, segment)
  if (wk >= nwork) return;
  const int c = wk / nseg, b0 = (wk - c * nseg) * per;
  const int nb = min(per, (T - 1) - b0);
  if (nb <= 0) return;
  const long long out_len = (long long)kHop * (T - 1);
  const float inv_m = 0.5f / kM;                           // 1/1024 of the inverse FFT and the 1/2 of the split step
  const float2 wl = __ldg(tw + kM + lane);                 // W_2048^lane
  const float2* win2 = reinterpret_cast<const float2*>(win);
  if (STAGED) {
    if (lane == 0) {
      ptx::mbar_init(bar, 1);
      ptx::fence_mbar_init();
      ptx::mbar_arrive_expect_tx(bar, kStageFloats * 4);
      ptx::bulk_g2s(ptx::smem_u32(stage), xf + ((long long)c * T + b0) * xs, kStageFloats * 4, bar);
    }
    __syncwarp();
  }
  float2 prev[16];
  float2 v[32];
  for (int f = 0; f <= nb; ++f) {
    const int t = b0 + f;
    const float* in = STAGED ? stage : xf + ((long long)c * T + t) * xs;
    if (STAGED) ptx::mbar_wait(bar, (uint32_t)(f & 1));
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const int k = lane + 32 * j;
      float2 xk, xm;
      if (STAGED) {
        xk = *reinterpret_cast<const float2*>(in + 2 * k);
        xm = *reinterpret_cast<const float2*>(in + 2 * (kM - k));
      } else {
        xk = __ldg(reinterpret_cast<const float2*>(in + 2 * k));
        xm = __ldg(reinterpret_cast<const float2*>(in + 2 * (kM - k)));
      }
      if (k == 0) { xk.y = 0.f; xm.y = 0.f; }
      // twice the textbook E and O: the factor 1/2 is folded into `scale` below
      const float2 e = make_float2(xk.x + xm.x, xk.y - xm.y);
      const float2 dlt = make_float2(xk.x - xm.x, xk.y + xm.y);
      const float2 w = j == 0 ? wl : cmul(wl, w64(j));   // W_2048^k
      const float2 o = make_float2(dlt.x * w.x + dlt.y * w.y, dlt.y * w.x - dlt.x * w.y);   // dlt * conj(w)
      v[j] = make_float2(e.x - o.y, e.y + o.x);
    }
    if (STAGED) {
      __syncwarp();                                        // every lane has read the staged row
      if (f < nb && lane == 0) {
        ptx::mbar_arrive_expect_tx(bar, kStageFloats * 4);
        ptx::bulk_g2s(ptx::smem_u32(stage), xf + ((long long)c * T + t + 1) * xs, kStageFloats * 4, bar);
      }
    }
    warp_fft1024<true>(v, tr, tw, lane);
    // v[br5(k2)] / 1024 = (x[2n], x[2n+1]), n = lane + 32 k2; k2 < 16 is the first half of the frame
    if (f > 0) {
      float* o = wav + (long long)c * out_len + (long long)(t - 1) * kHop;
#pragma unroll
      for (int k2 = 0; k2 < 16; ++k2) {
        const int n = lane + 32 * k2;
        const float2 z = v[br5(k2)];
        const float2 w = __ldg(win2 + n), r = inv_env[n];
        const float y0 = fmaf(z.x * inv_m, w.x, prev[k2].x) * r.x;
        const float y1 = fmaf(z.y * inv_m, w.y, prev[k2].y) * r.y;
        *reinterpret_cast<float2*>(o + 2 * n) = make_float2(y0, y1);
      }
    }
    if (f < nb) {
#pragma unroll
      for (int k2 = 0; k2 < 16; ++k2) {
        const int n = lane + 32 * k2;
        const float2 z = v[br5(k2 + 16)];
        const float2 u = __ldg(win2 + n + kHop / 2);
        prev[k2] = make_float2(z.x * inv_m * u.x, z.y * inv_m * u.y);
      }
    }
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
stream_synthesis_kernel(const float* __restrict__ y, const float* __restrict__ dry, const float* __restrict__ mix,
                        const float* __restrict__ win, const float2* __restrict__ tw, float* __restrict__ tail,
                        float* __restrict__ out) {
  __shared__ __align__(16) float2 sa[kM];
  __shared__ __align__(16) float2 sb[kM];
  const int c = blockIdx.x;
  if (mix != nullptr) irfft_split_load_mix(y + (long long)c * kSpec, dry + (long long)c * kSpec, __ldg(mix), tw + kM, sa);
  else irfft_split_load(y + (long long)c * kSpec, tw + kM, sa);
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
  float2* tw = nullptr;      // [1024] W_1024^k, then [1025] W_2048^k, then [32][32] W_1024^(l k1) at [k1][l]
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
    std::vector<float2> tw(kM + kM + 1 + 1024);
    for (int k = 0; k < kM; ++k) tw[k] = make_float2((float)std::cos(2.0 * pi * k / kM), (float)-std::sin(2.0 * pi * k / kM));
    for (int k = 0; k <= kM; ++k) tw[kM + k] = make_float2((float)std::cos(2.0 * pi * k / kN), (float)-std::sin(2.0 * pi * k / kN));
    for (int k1 = 0; k1 < 32; ++k1)
      for (int l = 0; l < 32; ++l)
        tw[kTw32 + k1 * 32 + l] = make_float2((float)std::cos(2.0 * pi * (l * k1) / kM), (float)-std::sin(2.0 * pi * (l * k1) / kM));
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

cudaError_t init_signal_tables() {
  Tables* t = nullptr;
  return ensure_tables(&t);
}
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
  const int nrows = C * T;
  stft_warp_kernel<<<(nrows + kFrameWarps - 1) / kFrameWarps, kFrameWarps * 32, 0, s>>>(wav, L, T, nrows, window, tw, xf, xs);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_istft_frames(const float* xf, int xs, int C, int T, const float* window,
                                float* wav, cudaStream_t s) {
  if (T < 2) return cudaSuccess;
  BSB_TABLES(window, tw);
  // a warp walks `per` consecutive output blocks (one extra frame of overlap per warp). Sized so
  // that all warps are resident at once (148 SMs x 12 or 16 warps) when the input allows it: one
  // balanced wave and <= 1/per redundant transforms.
  const bool staged = (xs % 4) == 0 && xs >= kStageFloats && (reinterpret_cast<uintptr_t>(xf) & 15) == 0;
  const int resident = 148 * kFrameWarps * (staged ? 3 : 4);
  const long long blocks = (long long)C * (T - 1);
  int per = (int)((blocks + resident - 1) / resident);
  if (per < 1) per = 1;
  if (per > 32) per = 32;
  const int nseg = (T - 1 + per - 1) / per;
  const int nwork = C * nseg;
  const int grid = (nwork + kFrameWarps - 1) / kFrameWarps;
  if (staged) {
    constexpr int smem = kFrameWarps * kIstftWarpSmem;
    static SmemOptIn opt;
    const cudaError_t e = opt_in_smem(istft_warp_kernel<true>, smem, opt);
    if (e != cudaSuccess) return e;
    istft_warp_kernel<true><<<grid, kFrameWarps * 32, smem, s>>>(xf, xs, T, nseg, nwork, window, tw, wav, per);
  } else {
    istft_warp_kernel<false><<<grid, kFrameWarps * 32, kFrameWarps * 32 * kTr * 8, s>>>(xf, xs, T, nseg, nwork, window, tw, wav, per);
  }
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

cudaError_t launch_stream_synthesis(const float* y, const float* dry, const float* mix, int C, const float* window,
                                    float* tail, float* out, cudaStream_t s) {
  BSB_TABLES(window, tw);
  stream_synthesis_kernel<<<C, kFftThreads, 0, s>>>(y, dry, mix, window, tw, tail, out);
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

// ------------------------------------------------------------------------------------------
// bsb_stft_pack_static: the STFT + pack of bsb_stft_pack for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// bsb_stft_pack needs a caller workspace for the frame-major spectrum it transposes
// from, and its window / twiddle tables come from a cudaMalloc on first use; here BOTH live in file-scope __device__
// arrays: every GPU that loads the library holds them (4.03 MB) until the process exits.
//   * limit: C * T <= kStaticFrameRows frame-rows (512 = 11.9 s of audio over all rows); larger calls are rejected
//   * the scratch is shared by every call on a device and keeps the previous call's contents. Nothing in it has to
//     start at zero: the analysis kernel writes all 2050 floats of every frame-row the transpose reads. Two calls
//     whose kernels overlap on the GPU (different streams, no ordering between them) RACE on it.
// In host code the array names below are host-side stand-ins, not device addresses: the launches take the pointers
// that cudaGetSymbolAddress returned for the current device at set-up.
constexpr int kStaticFrameRows = 512;
constexpr int kStaticTw = kM + kM + 1 + 1024;
static __device__ __align__(256) float g_static_scratch[kStaticFrameRows * kXS];   // frame-major xf[C*T][2052]
static __device__ __align__(256) float g_static_window[kN];                        // periodic Hann
static __device__ __align__(256) float2 g_static_tw[kStaticTw];                    // the three twiddle tables (Tables::tw)

namespace {
struct StaticPtrs {
  float* scratch = nullptr;
  float* window = nullptr;
  float2* tw = nullptr;
  bool ready = false;
};
std::mutex g_static_mu;
StaticPtrs g_static_ptrs[64];

// once per device, with that device current: resolve the device addresses and upload the constant tables (plain
// cudaMemcpy to memory the module already owns: no allocation)
cudaError_t static_setup(StaticPtrs** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_static_mu);
  StaticPtrs& p = g_static_ptrs[dev];
  if (!p.ready) {
    if ((e = cudaGetSymbolAddress((void**)&p.scratch, g_static_scratch)) != cudaSuccess) return e;
    if ((e = cudaGetSymbolAddress((void**)&p.window, g_static_window)) != cudaSuccess) return e;
    if ((e = cudaGetSymbolAddress((void**)&p.tw, g_static_tw)) != cudaSuccess) return e;
    const double pi = 3.14159265358979323846;          // same host arithmetic as ensure_tables: identical tables
    static float w[kN];
    static float2 tw[kStaticTw];
    for (int n = 0; n < kN; ++n) w[n] = (float)(0.5 - 0.5 * std::cos(2.0 * pi * n / kN));
    for (int k = 0; k < kM; ++k) tw[k] = make_float2((float)std::cos(2.0 * pi * k / kM), (float)-std::sin(2.0 * pi * k / kM));
    for (int k = 0; k <= kM; ++k) tw[kM + k] = make_float2((float)std::cos(2.0 * pi * k / kN), (float)-std::sin(2.0 * pi * k / kN));
    for (int k1 = 0; k1 < 32; ++k1)
      for (int l = 0; l < 32; ++l)
        tw[kTw32 + k1 * 32 + l] = make_float2((float)std::cos(2.0 * pi * (l * k1) / kM), (float)-std::sin(2.0 * pi * (l * k1) / kM));
    if ((e = cudaMemcpy(p.window, w, sizeof(w), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    if ((e = cudaMemcpy(p.tw, tw, sizeof(tw), cudaMemcpyHostToDevice)) != cudaSuccess) return e;
    p.ready = true;
  }
  *out = &p;
  return cudaSuccess;
}
}  // namespace

int static_max_frame_rows() { return kStaticFrameRows; }

cudaError_t launch_stft_pack_static(const float* wav, int C, long long L, int T, const float* window, float* x,
                                    cudaStream_t s) {
  StaticPtrs* p = nullptr;
  cudaError_t e = static_setup(&p);
  if (e != cudaSuccess) return e;
  const int nrows = C * T;
  stft_warp_kernel<<<(nrows + kFrameWarps - 1) / kFrameWarps, kFrameWarps * 32, 0, s>>>(
      wav, L, T, nrows, window ? window : p->window, p->tw, p->scratch, kXS);
  count_launch();
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  dim3 grid((T + 31) / 32, (kSpec + 31) / 32, C);
  from_frames_kernel<<<grid, 256, 0, s>>>(p->scratch, kXS, T, x);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
