// LSTM recurrence (nn.LSTM semantics, bsrnn.py:70): gate order i,f,g,o,
// c' = sig(f)*c + sig(i)*tanh(g), h' = sig(o)*tanh(c').
//
// The input projections W_ih.x + b_ih + b_hh of all steps are hoisted into one GEMM
// (grouped_linear) and arrive here as `xproj`; this kernel runs only the serial part.
// One CTA owns S sequences for all their steps and never talks to another CTA (hidden size
// 64 and independent sequences: no grid-level sync needed).
//
// Thread map (256 threads): warp w, lane l -> hidden unit u = 8w + l/4, gate g = l%4.
// Each thread keeps its W_hh row (gate g, unit u: 64 floats) in registers for the whole
// kernel; h lives in shared memory (double buffered, broadcast reads), c in registers.
// After the mat-vec the 4 lanes of a unit hold the 4 gates of S sequences; a 4x4 transpose
// with warp shuffles gives lane g all four gates of sequence 4q+g, so the cell update is
// spread over all lanes with no shared-memory exchange and one __syncthreads per step.
#include "kernels.cuh"

namespace bsb {
namespace {

// same activation arithmetic as the tensor-core recurrences (lstm_tc.cu, lstm_seq2.cu): ex2.approx + rcp.approx, <= 2 ulp each
__device__ __forceinline__ float ex2_(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sigmoidf_(float x) { return rcp_(1.0f + ex2_(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanhf_(float x) { return 2.0f * rcp_(1.0f + ex2_(-2.8853900817779268f * x)) - 1.0f; }

__device__ __forceinline__ float sel4(float a, float b, float c, float d, int idx) {
  const float lo = (idx & 1) ? b : a;
  const float hi = (idx & 1) ? d : c;
  return (idx & 2) ? hi : lo;
}

// NV: sequences of the group that exist (compile time, so the mat-vec of the empty slots disappears): S, or 1 / 2
// for the single short group of a 1- or 2-row streaming hop (band axis: one sequence per row).
template <int S, int NV>
__global__ void __launch_bounds__(256)
lstm_rec_kernel(const LstmArgs a) {
  static_assert(S % 4 == 0, "S must be a multiple of 4");
  __shared__ __align__(16) float hs[2][S][kFeat];

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  const int u = warp * 8 + (lane >> 2);
  const int g = lane & 3;
  const int dir = blockIdx.y;
  const int seq0 = blockIdx.x * S;
  const bool reverse = (dir == 1);

  // W_hh row (gate g, unit u) -> registers
  float w[kFeat];
  {
    const float4* wp = reinterpret_cast<const float4*>(a.whh + ((long long)dir * kGates + g * kFeat + u) * kFeat);
#pragma unroll
    for (int k = 0; k < kFeat / 4; ++k) {
      const float4 v = __ldg(wp + k);
      w[4 * k + 0] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
    }
  }

  // per-sequence base offsets of xproj (this thread's gate column) and hout
  long long xoff[S];
#pragma unroll
  for (int s = 0; s < S; ++s) {
    const int sq = seq0 + s;
    xoff[s] = (sq < a.nseq)
                  ? (long long)(sq / a.n_inner) * a.xp_outer + (long long)(sq % a.n_inner) * a.xp_inner +
                        dir * kGates + g * kFeat + u
                  : -1;
  }

  // initial state: this lane owns (sequence 4q+g, unit u) for q = 0..S/4-1
  float c[S / 4];
#pragma unroll
  for (int q = 0; q < S / 4; ++q) {
    const int s = 4 * q + g;
    const int sq = seq0 + s;
    float h0 = 0.f, c0 = 0.f;
    if (sq < a.nseq && a.h0 != nullptr) {
      h0 = a.h0[(long long)sq * kFeat + u];
      c0 = a.c0[(long long)sq * kFeat + u];
    }
    c[q] = c0;
    hs[0][s][u] = h0;
  }

  float xnext[S];
  {
    const int st0 = reverse ? a.nsteps - 1 : 0;
#pragma unroll
    for (int s = 0; s < S; ++s)
      xnext[s] = (xoff[s] >= 0) ? __ldg(a.xproj + xoff[s] + (long long)st0 * a.xp_step) : 0.f;
  }
  __syncthreads();

  int cur = 0;
  for (int it = 0; it < a.nsteps; ++it) {
    const int step = reverse ? a.nsteps - 1 - it : it;
    float acc[S];
#pragma unroll
    for (int s = 0; s < S; ++s) acc[s] = xnext[s];
    if (it + 1 < a.nsteps) {
      const int stn = reverse ? step - 1 : step + 1;
#pragma unroll
      for (int s = 0; s < S; ++s)
        xnext[s] = (xoff[s] >= 0) ? __ldg(a.xproj + xoff[s] + (long long)stn * a.xp_step) : 0.f;
    }
    // gate pre-activation: acc[s] += W_hh[g,u,:] . h[s,:]
#pragma unroll
    for (int k4 = 0; k4 < kFeat / 4; ++k4) {
#pragma unroll
      for (int s = 0; s < NV; ++s) {
        const float4 hv = *reinterpret_cast<const float4*>(&hs[cur][s][4 * k4]);
        acc[s] = fmaf(w[4 * k4 + 0], hv.x, acc[s]);
        acc[s] = fmaf(w[4 * k4 + 1], hv.y, acc[s]);
        acc[s] = fmaf(w[4 * k4 + 2], hv.z, acc[s]);
        acc[s] = fmaf(w[4 * k4 + 3], hv.w, acc[s]);
      }
    }
    // 4x4 transpose across the unit's 4 lanes, then the cell update
#pragma unroll
    for (int q = 0; q < S / 4; ++q) {
      float recv[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        // lane with gate g' supplies its value for sequence 4q + (g' ^ r); the reader g = g'^r
        const float mine = sel4(acc[4 * q + 0], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3], g ^ r);
        recv[r] = (r == 0) ? mine : __shfl_xor_sync(0xffffffffu, mine, r);
      }
      // recv[r] is gate (g ^ r) of sequence 4q+g
      const float gi = sel4(recv[0], recv[1], recv[2], recv[3], g ^ 0);
      const float gf = sel4(recv[0], recv[1], recv[2], recv[3], g ^ 1);
      const float gg = sel4(recv[0], recv[1], recv[2], recv[3], g ^ 2);
      const float go = sel4(recv[0], recv[1], recv[2], recv[3], g ^ 3);
      const float cn = sigmoidf_(gf) * c[q] + sigmoidf_(gi) * tanhf_(gg);
      const float hn = sigmoidf_(go) * tanhf_(cn);
      c[q] = cn;
      const int s = 4 * q + g;
      const int sq = seq0 + s;
      hs[cur ^ 1][s][u] = hn;
      if (sq < a.nseq) {
        const long long o = (long long)(sq / a.n_inner) * a.ho_outer + (long long)(sq % a.n_inner) * a.ho_inner +
                            (long long)step * a.ho_step + dir * kFeat + u;
        a.hout[o] = hn;
        if (it == a.nsteps - 1 && a.hN != nullptr) {
          a.hN[(long long)sq * kFeat + u] = hn;
          a.cN[(long long)sq * kFeat + u] = cn;
        }
      }
    }
    __syncthreads();
    cur ^= 1;
  }
}

}  // namespace

cudaError_t launch_lstm(const LstmArgs& a, cudaStream_t s) {
  if (a.nseq <= 0 || a.nsteps <= 0) return cudaSuccess;
  // few long sequences (time axis): small S spreads them over the SMs;
  // many short sequences (band axis): larger S amortises the W_hh register load.
  if (a.nseq <= 2) {                             // 1- or 2-row hop, band axis
    dim3 grid(1, a.ndir);
    if (a.nseq == 1) lstm_rec_kernel<4, 1><<<grid, 256, 0, s>>>(a);
    else lstm_rec_kernel<4, 2><<<grid, 256, 0, s>>>(a);
  } else if (a.nseq <= 4 * 148 * 2) {
    constexpr int S = 4;
    dim3 grid((a.nseq + S - 1) / S, a.ndir);
    lstm_rec_kernel<S, S><<<grid, 256, 0, s>>>(a);
  } else {
    constexpr int S = 8;
    dim3 grid((a.nseq + S - 1) / S, a.ndir);
    lstm_rec_kernel<S, S><<<grid, 256, 0, s>>>(a);
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
