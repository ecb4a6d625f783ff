// Time-axis LSTM layer (TimewiseLSTM, bsrnn.py:101-128; nn.LSTM semantics bsrnn.py:70): few, long,
// independent sequences (C*12 of them, T steps). The step count is the serial depth, so the
// kernel is built for per-step latency, not throughput: a CTA owns 8 sequences for all their
// steps and never talks to another CTA.
//
//   gates^T[256 x 8] = W_ih' . x_t  +  W_hh . h_{t-1}            (fc_in folded into W_ih', layout.h)
//
// Only the W_hh half depends on the previous step. It runs on mma.sync (m16n8k16, the 8 sequences are the
// n8 side): 16 warps, warp w keeps the 16 gate rows of hidden units 4w..4w+3 as bf16 hi/lo A fragments in
// registers for the whole kernel; h_t is handed to the next step through shared memory as B fragments
// (double buffered, one barrier per step) and flushed to HBM as fp32 rows with coalesced 16-byte stores.
// A tcgen05 tile for this half would need >= 64 sequence rows per CTA, which at 768 sequences would
// light 12 SMs; the warp-level MMA lets 96 CTAs each advance 8 sequences per step.
//
// The W_ih' half does not depend on the recurrence, and at 12 + 12 warp-MMAs per warp and step the legacy
// tensor path was the limiter (73 % busy). It therefore runs ahead of the recurrence on the 5th-gen tensor
// core of the same SM, 16 steps at a time: rows = (step, sequence) = 128, one tcgen05.mma tile
// [128 x 256] = x[128 x 64] . W_ih'^T per block, accumulators in TMEM (two blocks in flight), drained by four
// helper warps into a shared-memory ring of 2 x 4 steps from which the recurrence warps read their four
// pre-activations per step with two 8-byte loads:
//
//   warps 0-15  recurrence (mma.sync W_hh.h, gate exchange, cell update, h flush)
//   warps 16-19 helpers: x rows of the next block  global -> bf16 hi/lo, 128B-swizzled smem (A operand);
//               TMEM accumulator (lane quarter = warp % 4 = steps 4q..4q+3 of the block) -> ring
//   warp  20    one thread issues the 12 tcgen05.mma of a block (3 bf16 products per fp32 product)
//
// Split-bf16 arithmetic as everywhere else (hi.hi + hi.lo + lo.hi, fp32 accumulate). Gate rows are
// interleaved so that lane pairs (l, l^4) hold i,g / f,o of the same unit: two shuffles give every lane
// the four gates of one (unit, sequence) cell.
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int SQ = 8;            // sequences per CTA (the n8 of the MMA)
constexpr int NWARP = 16;        // recurrence warps
constexpr int NT = NWARP * 32;
constexpr int HELP_WARP0 = NWARP, N_HELP = 4, MMA_WARP = NWARP + N_HELP;
constexpr int THREADS = (MMA_WARP + 1) * 32;
constexpr int BS = 72;           // bf16 row stride of the B-fragment buffers (36 words: conflict free)
constexpr int FS = 68;           // fp32 row stride of the h staging buffer
constexpr int BLK = 16;          // steps per tcgen05 block (16 steps x 8 sequences = 128 rows)
constexpr int QSTEPS = 4;        // steps per ring half (= one TMEM lane quarter)
constexpr int W_PLANE = kGates * 128;                 // 32 KB: [256 rows][64 k] bf16
constexpr int A_PLANE = 128 * 128;                    // 16 KB: [128 rows][64 k] bf16
constexpr int STEP_STRIDE = kGates * SQ * 4 + 32;     // ring bytes per step: [256 cols][8 seq] fp32, +32 B: bank shift
constexpr int HALF_BYTES = QSTEPS * STEP_STRIDE;
constexpr int OFF_W = 0, OFF_A = 2 * W_PLANE, OFF_RING = OFF_A + 2 * 2 * A_PLANE;
constexpr int SMEM_BYTES = OFF_RING + 2 * HALF_BYTES + 1024;

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

__device__ __forceinline__ void bar_recurrence() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }

__global__ void __launch_bounds__(THREADS, 1)
lstm_seq_kernel(const __grid_constant__ LstmSeqArgs a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(16) __nv_bfloat16 hB[2][2][SQ][BS];   // [buffer][plane][sequence][k]
  __shared__ __align__(16) float hF[2][SQ][FS];
  __shared__ __align__(8) uint64_t bar_xfull[2], bar_xempty[2], bar_accfull[2], bar_accfree[2], bar_ringfull[2], bar_ringfree[4];
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x;
  const int w = tid >> 5, lane = tid & 31;
  const int seq0 = blockIdx.x * SQ;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int nblk = (a.nsteps + BLK - 1) / BLK;
  const int nq = (a.nsteps + QSTEPS - 1) / QSTEPS;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bar_xfull[i]), N_HELP * 32);
      mbar_init(smem_u32(&bar_xempty[i]), 1);
      mbar_init(smem_u32(&bar_accfull[i]), 1);
      mbar_init(smem_u32(&bar_accfree[i]), N_HELP * 32);
      mbar_init(smem_u32(&bar_ringfull[i]), 32);
    }
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar_ringfree[i]), NWARP);
    fence_mbar_init();
  }
  if (w == MMA_WARP) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  {  // W_ih' image (tcgen05 B operand): already the swizzled shared-memory layout
    const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(a.wfrag) + kSeqFragFloats * 4);
    uint4* dst = reinterpret_cast<uint4*>(smem_gen + OFF_W);
    for (int i = tid; i < 2 * W_PLANE / 16; i += THREADS) dst[i] = __ldg(src + i);
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (w < NWARP) {
    // ================= recurrence =================
    const int gr = lane >> 2, tq = lane & 3;
    uint4 wh[4][2];            // W_hh rows of this warp as A fragments
    {
      const uint4* wf = reinterpret_cast<const uint4*>(a.wfrag) + (size_t)w * 8 * 2 * 32 + lane;
#pragma unroll
      for (int kt = 0; kt < 4; ++kt)
#pragma unroll
        for (int p = 0; p < 2; ++p) wh[kt][p] = __ldg(wf + ((kt + 4) * 2 + p) * 32);
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
    // h writer role of the first 128 threads: (sequence tid/16, 4 floats at 4*(tid%16))
    const bool io = tid < SQ * 16;
    const int lseq = tid >> 4, part = tid & 15;
    const int lg = seq0 + lseq;
    const bool lvalid = io && lg < a.nseq;
    float* hptr = a.hout;
    long long hoff = 0;          // float offset of (sequence, step 0) in the output rows
    if (lvalid) hoff = (long long)(lg / a.n_inner) * a.h_outer + (long long)(lg % a.n_inner) * a.h_inner;
    hptr += hoff + part * 4;

    // W_ih'.x_t of this lane's four accumulator positions: ring[half][step % 4][column][sequence]
    const uint8_t* ring = smem_gen + OFF_RING + (w * 16 + gr) * (SQ * 4) + tq * 8;
    float2 xlo, xhi;
    auto fetch_x = [&](int t) {
      const uint8_t* p = ring + ((t >> 2) & 1) * HALF_BYTES + (t & 3) * STEP_STRIDE;
      xlo = *reinterpret_cast<const float2*>(p);
      xhi = *reinterpret_cast<const float2*>(p + 8 * SQ * 4);
    };
    mbar_wait(smem_u32(&bar_ringfull[0]), 0);
    fetch_x(0);
    bar_recurrence();          // h_{-1} is in place

    for (int t = 0; t < a.nsteps; ++t) {
      const int cur = t & 1, nxt = cur ^ 1;
      float acc[4] = {xlo.x + b_lo, xlo.y + b_lo, xhi.x + b_hi, xhi.y + b_hi};
      if ((t & 3) == 3 || t == a.nsteps - 1) {       // this warp has read its last value of the ring half
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_ringfree[(t >> 2) & 3]));
      }
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
      if (t + 1 < a.nsteps) {
        if (((t + 1) & 3) == 0) mbar_wait(smem_u32(&bar_ringfull[((t + 1) >> 2) & 1]), ((t + 1) >> 3) & 1);
        fetch_x(t + 1);
      }
      bar_recurrence();
      if (lvalid) {
        if (a.hplanes != nullptr) {
          // A operand of the Linear that follows (kernels.cuh planes: [m-tile][32-column chunk][128 x 64 B]): row =
          // output row index, columns 4 part .. +3; the bf16 hi/lo pairs are already in the B-fragment buffers
          const long long m = (hoff + (long long)t * a.h_step) >> 6;
          const int r = (int)(m & 127);
          uint8_t* p = reinterpret_cast<uint8_t*>(a.hplanes) + ((m >> 7) * 2 + (part >> 3)) * (2 * kPlaneBlock) + sw64(r, (part & 7) >> 1) +
                       (part & 1) * 8;
          *reinterpret_cast<uint2*>(p) = *reinterpret_cast<const uint2*>(&hB[nxt][0][lseq][part * 4]);
          *reinterpret_cast<uint2*>(p + a.hp_lo) = *reinterpret_cast<const uint2*>(&hB[nxt][1][lseq][part * 4]);
        } else {
          *reinterpret_cast<float4*>(hptr + (long long)t * a.h_step) = *reinterpret_cast<const float4*>(&hF[nxt][lseq][part * 4]);
        }
      }
    }
  } else if (w < MMA_WARP) {
    // ================= helpers: x rows -> A operand, accumulator -> ring =================
    const int dq = w - HELP_WARP0;              // TMEM lane quarter (HELP_WARP0 % 4 == 0)
    const int j = dq * 32 + lane;               // 0..127: (sequence j/16, 4 floats at 4*(j%16))
    const int lseq = j >> 4, part = j & 15;
    const int lg = seq0 + lseq;
    const bool lvalid = lg < a.nseq;
    const float* xptr = a.x;
    if (lvalid) xptr += (long long)(lg / a.n_inner) * a.x_outer + (long long)(lg % a.n_inner) * a.x_inner + part * 4;
    auto load_block = [&](int b) {              // 16 steps x 8 sequences x 64 floats -> bf16 hi/lo planes of buffer b&1
      uint8_t* a_hi = smem_gen + OFF_A + (b & 1) * 2 * A_PLANE;
      uint8_t* a_lo = a_hi + A_PLANE;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float4 v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int step = b * BLK + half * 8 + i;
          v[i] = (lvalid && step < a.nsteps) ? __ldg(reinterpret_cast<const float4*>(xptr + (long long)step * a.x_step))
                                             : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = (half * 8 + i) * SQ + lseq;           // row = (step in block, sequence)
          uint2 h, l;
          split2(v[i].x, v[i].y, h.x, l.x);
          split2(v[i].z, v[i].w, h.y, l.y);
          const uint32_t off = (uint32_t)(r * 128 + (((part >> 1) ^ (r & 7)) << 4) + (part & 1) * 8);
          *reinterpret_cast<uint2*>(a_hi + off) = h;
          *reinterpret_cast<uint2*>(a_lo + off) = l;
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(smem_u32(&bar_xfull[b & 1]));
    };
    const uint32_t tlane = tmem_base + ((uint32_t)(dq * 32) << 16);
    uint8_t* ring_w = smem_gen + OFF_RING + (lane >> 3) * STEP_STRIDE + (lane & 7) * 4;   // (step in quarter, sequence)
    load_block(0);
    for (int b = 0; b < nblk; ++b) {
      if (b + 1 < nblk) {
        if (b + 1 >= 2) mbar_wait(smem_u32(&bar_xempty[(b + 1) & 1]), ((b - 1) >> 1) & 1);   // MMAs of block b-1 retired
        load_block(b + 1);
      }
      const int g = b * (BLK / QSTEPS) + dq;    // global quarter index handled by this warp
      if (g < nq) {
        mbar_wait(smem_u32(&bar_accfull[b & 1]), (b >> 1) & 1);
        tc_fence_after();
        // quarter g-2 (the previous user of this ring half, written by warp dq ^ 2) must be consumed. One free
        // barrier per helper warp's quarters (index = quarter % 4), each waited on by exactly one warp (this one
        // waits on dq ^ 2): a completion can never run two phases ahead of its only observer, however long the
        // observer was held up elsewhere (e.g. waiting for the other layer's rows).
        if (g >= 2) mbar_wait(smem_u32(&bar_ringfree[dq ^ 2]), (uint32_t)(((g - 2) >> 2) & 1));
        uint8_t* dst = ring_w + (g & 1) * HALF_BYTES;
#pragma unroll 1
        for (int cb = 0; cb < kGates / 32; ++cb) {
          uint32_t r[32];
          tmem_ld_32x32(tlane + (uint32_t)((b & 1) * kGates + cb * 32), r);
          tmem_ld_wait32(r);
#pragma unroll
          for (int i = 0; i < 32; ++i) *reinterpret_cast<uint32_t*>(dst + (cb * 32 + i) * (SQ * 4)) = r[i];
        }
        mbar_arrive(smem_u32(&bar_ringfull[g & 1]));
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_accfree[b & 1]));
    }
  } else {
    // ================= tcgen05 issuer: x[128 x 64] . W_ih'^T -> accumulator b&1 =================
    const uint32_t idesc = make_idesc_bf16(128, kGates);
    const uint64_t w_hi = make_sw128_desc(smem_base + OFF_W);
    const uint64_t w_lo = make_sw128_desc(smem_base + OFF_W + W_PLANE);
    for (int b = 0; b < nblk; ++b) {
      mbar_wait(smem_u32(&bar_xfull[b & 1]), (b >> 1) & 1);
      if (b >= 2) mbar_wait(smem_u32(&bar_accfree[b & 1]), ((b >> 1) - 1) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint64_t a_hi = make_sw128_desc(smem_base + OFF_A + (b & 1) * 2 * A_PLANE);
        const uint64_t a_lo = make_sw128_desc(smem_base + OFF_A + (b & 1) * 2 * A_PLANE + A_PLANE);
        const uint32_t d_tmem = tmem_base + (uint32_t)((b & 1) * kGates);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adv = (uint64_t)(kk * 32 >> 4);       // 16 bf16 = 32 bytes along K
          mma_bf16_ss(d_tmem, a_lo + adv, w_hi + adv, idesc, kk != 0);   // small terms first
          mma_bf16_ss(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
          mma_bf16_ss(d_tmem, a_hi + adv, w_hi + adv, idesc, 1);
        }
        mma_commit(smem_u32(&bar_accfull[b & 1]));
        mma_commit(smem_u32(&bar_xempty[b & 1]));
      }
      __syncwarp();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (w == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

cudaError_t launch_lstm_seq(const LstmSeqArgs& a, cudaStream_t s) {
  if (a.nseq <= 0 || a.nsteps <= 0) return cudaSuccess;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(lstm_seq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  lstm_seq_kernel<<<(a.nseq + SQ - 1) / SQ, THREADS, SMEM_BYTES, s>>>(a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
