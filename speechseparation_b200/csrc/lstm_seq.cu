// Time-axis LSTM layer (TimewiseLSTM, bsrnn.py:101-128; nn.LSTM semantics bsrnn.py:70): few, long,
// independent sequences (C*12 of them, T steps). The step count is the serial depth, so the
// kernel is built for per-step latency, not throughput: a CTA owns 8 sequences for all their
// steps and never talks to another CTA.
//
//   gates^T[256 x 8] = [W_ih' | W_hh] . [x_t ; h_{t-1}]      (fc_in folded into W_ih', layout.h)
//
// 16 warps; warp w keeps the 16 gate rows of hidden units 4w..4w+3 (both K halves, bf16 hi/lo
// planes = 64 registers) as mma.sync m16n8k16 A fragments for the whole kernel; the 8 sequences
// are the n8 side. A tcgen05 tile needs >= 64 rows of sequences per CTA, which at 768 sequences
// would light 12 SMs; the warp-level MMA lets 96 CTAs each advance 8 sequences per ~0.5 us step.
// Split-bf16 arithmetic as everywhere else (hi.hi + hi.lo + lo.hi, fp32 accumulate).
//
// Per step only the W_hh half is on the critical path: the W_ih' half of step t+1 is computed
// while other warps finish step t (x is prefetched 3-4 steps ahead), h_t is handed to the next step
// through shared memory as bf16 hi/lo B fragments (double buffered, one __syncthreads per step)
// and flushed to HBM as fp32 rows with coalesced 16-byte stores.
// Gate rows are interleaved so that lane pairs (l, l^4) hold i,g / f,o of the same unit: two
// shuffles give every lane the four gates of one (unit, sequence) cell.
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int SQ = 8;            // sequences per CTA (the n8 of the MMA)
constexpr int NWARP = 16;
constexpr int NT = NWARP * 32;
constexpr int BS = 72;           // bf16 row stride of the B-fragment buffers (36 words: conflict free)
constexpr int FS = 68;           // fp32 row stride of the h staging buffer

__device__ __forceinline__ void hmma(float (&d)[4], const uint4& a, uint32_t b0, uint32_t b1) {
  asm(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b0), "r"(b1));
}

// acc += W(4 k-tiles, hi/lo) . B(hi/lo planes in smem); three independent accumulation chains
__device__ __forceinline__ void matpart(float (&acc)[4], const uint4 (&W)[4][2],
                                        const __nv_bfloat16 (&B)[2][SQ][BS], int gr, int tq) {
  float a1[4] = {0.f, 0.f, 0.f, 0.f}, a2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int kt = 0; kt < 4; ++kt) {
    const uint32_t* bh = reinterpret_cast<const uint32_t*>(&B[0][gr][kt * 16 + 2 * tq]);
    const uint32_t* bl = reinterpret_cast<const uint32_t*>(&B[1][gr][kt * 16 + 2 * tq]);
    const uint32_t bh0 = bh[0], bh1 = bh[4], bl0 = bl[0], bl1 = bl[4];
    hmma(acc, W[kt][0], bh0, bh1);
    hmma(a1, W[kt][0], bl0, bl1);
    hmma(a2, W[kt][1], bh0, bh1);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) acc[i] += a1[i] + a2[i];
}

__global__ void __launch_bounds__(NT, 1)
lstm_seq_kernel(const __grid_constant__ LstmSeqArgs a) {
  __shared__ __align__(16) __nv_bfloat16 xB[2][2][SQ][BS];   // [buffer][plane][sequence][k]
  __shared__ __align__(16) __nv_bfloat16 hB[2][2][SQ][BS];
  __shared__ __align__(16) float hF[2][SQ][FS];

  const int tid = threadIdx.x;
  const int w = tid >> 5, lane = tid & 31;
  const int gr = lane >> 2, tq = lane & 3;
  const int seq0 = blockIdx.x * SQ;

  // weights -> registers (A fragments), bias of this lane's two rows
  uint4 wx[4][2], wh[4][2];
  {
    const uint4* wf = reinterpret_cast<const uint4*>(a.wfrag) + (size_t)w * 8 * 2 * 32 + lane;
#pragma unroll
    for (int kt = 0; kt < 4; ++kt)
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        wx[kt][p] = __ldg(wf + (kt * 2 + p) * 32);
        wh[kt][p] = __ldg(wf + ((kt + 4) * 2 + p) * 32);
      }
  }
  const float b_lo = __ldg(a.bias + w * 16 + gr), b_hi = __ldg(a.bias + w * 16 + gr + 8);

  // the cell this lane updates: unit 4w + gr/2 of sequence 2tq + (gr&1)
  const bool odd = gr & 1;
  const int unit = 4 * w + (gr >> 1);
  const int sl = 2 * tq + (odd ? 1 : 0);
  const int gseq = seq0 + sl;
  const bool valid = gseq < a.nseq;
  float c = 0.f;
  {
    float h0 = 0.f;
    if (valid && a.h0 != nullptr) {
      h0 = __ldg(a.h0 + (long long)gseq * kFeat + unit);
      c = __ldg(a.c0 + (long long)gseq * kFeat + unit);
    }
    uint16_t hh, hl;
    split1(h0, hh, hl);
    reinterpret_cast<uint16_t&>(hB[0][0][sl][unit]) = hh;
    reinterpret_cast<uint16_t&>(hB[0][1][sl][unit]) = hl;
  }

  // x loader / h writer role of the first 128 threads: (sequence tid/16, 4 floats at 4*(tid%16))
  const bool io = tid < SQ * 16;
  const int lseq = tid >> 4, part = tid & 15;
  const int lg = seq0 + lseq;
  const bool lvalid = io && lg < a.nseq;
  const float* xptr = a.x;
  float* hptr = a.hout;
  if (lvalid) {
    xptr += (long long)(lg / a.n_inner) * a.x_outer + (long long)(lg % a.n_inner) * a.x_inner + part * 4;
    hptr += (long long)(lg / a.n_inner) * a.h_outer + (long long)(lg % a.n_inner) * a.h_inner + part * 4;
  }
  auto ldx = [&](int step) {
    return (lvalid && step < a.nsteps) ? __ldg(reinterpret_cast<const float4*>(xptr + (long long)step * a.x_step))
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
  };
  auto stx = [&](int buf, const float4& v) {
    uint32_t h0, l0, h1, l1;
    split2(v.x, v.y, h0, l0);
    split2(v.z, v.w, h1, l1);
    *reinterpret_cast<uint2*>(&xB[buf][0][lseq][part * 4]) = make_uint2(h0, h1);
    *reinterpret_cast<uint2*>(&xB[buf][1][lseq][part * 4]) = make_uint2(l0, l1);
  };
  float4 xr0 = make_float4(0.f, 0.f, 0.f, 0.f), xr1 = xr0;
  if (io) {
    stx(0, ldx(0));
    stx(1, ldx(1));
    xr0 = ldx(2);
    xr1 = ldx(3);
  }
  __syncthreads();
  float xacc[4] = {b_lo, b_lo, b_hi, b_hi};
  matpart(xacc, wx, xB[0], gr, tq);
  __syncthreads();

  for (int t = 0; t < a.nsteps; ++t) {
    const int cur = t & 1, nxt = cur ^ 1;
    // ---- critical path: W_hh . h_{t-1}, gate exchange, cell update --------------------------------
    float acc[4] = {xacc[0], xacc[1], xacc[2], xacc[3]};
    matpart(acc, wh, hB[cur], gr, tq);
    const float s0 = odd ? acc[0] : acc[1], s1 = odd ? acc[2] : acc[3];
    const float r0 = __shfl_xor_sync(0xffffffffu, s0, 4), r1 = __shfl_xor_sync(0xffffffffu, s1, 4);
    const float ti = fminf(odd ? r0 : acc[0], 40.f), tg = fminf(odd ? r1 : acc[2], 40.f);
    const float tf = fminf(odd ? acc[1] : r0, 40.f), to = fminf(odd ? acc[3] : r1, 40.f);
    const float ei = ex2_approx(ti), ef = ex2_approx(tf), eg = ex2_approx(tg), eo = ex2_approx(to);
    const float pi = 1.f + ei, pf = 1.f + ef, pg = 1.f + eg;
    const float pig = pi * pg;
    c = fmaf(c, pig, (1.f - eg) * pf) * rcp_approx(pf * pig);
    const float ec = ex2_approx(fminf(c * -2.885390081777927f, 40.f));
    const float h = (1.f - ec) * rcp_approx((1.f + eo) * (1.f + ec));
    {
      uint16_t hh, hl;
      split1(h, hh, hl);
      reinterpret_cast<uint16_t&>(hB[nxt][0][sl][unit]) = hh;
      reinterpret_cast<uint16_t&>(hB[nxt][1][sl][unit]) = hl;
      hF[nxt][sl][unit] = h;
    }
    if (t == a.nsteps - 1 && valid && a.hN != nullptr) {
      a.hN[(long long)gseq * kFeat + unit] = h;
      a.cN[(long long)gseq * kFeat + unit] = c;
    }
    // ---- off the critical path: input half of step t+1, x staging for step t+2 -------------------
    if (t + 1 < a.nsteps) {
      xacc[0] = xacc[1] = b_lo;
      xacc[2] = xacc[3] = b_hi;
      matpart(xacc, wx, xB[nxt], gr, tq);
    }
    if (io) {
      stx(cur, xr0);                 // x_{t+2} replaces x_t (consumed one barrier ago)
      xr0 = xr1;
      xr1 = ldx(t + 4);
    }
    __syncthreads();
    if (lvalid)
      *reinterpret_cast<float4*>(hptr + (long long)t * a.h_step) = *reinterpret_cast<const float4*>(&hF[nxt][lseq][part * 4]);
  }
}

}  // namespace

cudaError_t launch_lstm_seq(const LstmSeqArgs& a, cudaStream_t s) {
  if (a.nseq <= 0 || a.nsteps <= 0) return cudaSuccess;
  lstm_seq_kernel<<<(a.nseq + SQ - 1) / SQ, NT, 0, s>>>(a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
