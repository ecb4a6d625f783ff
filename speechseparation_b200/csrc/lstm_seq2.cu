// Time-axis LSTM block, BOTH layers in one launch, every matrix product on tcgen05 (TimewiseLSTM,
// bsrnn.py:101-128; nn.LSTM semantics bsrnn.py:70).
//
// The step count T is the serial depth of this part of the model. Two things shorten it:
//
// 1. Wavefront over the layers: CTA 2g advances layer 0 of sequence group g (16 sequences), CTA 2g+1 advances
//    layer 1 of the same group a few steps behind it, reading layer 0's h rows from L2 as they are published
//    (one release flag per group, bumped every 4 steps). Depth T + lag instead of 2T.
//
// 2. The recurrent product itself runs TRANSPOSED on the 5th-gen tensor core:
//      gates^T[256 gate rows x 16 sequences] = W_ih' . x_t^T + W_hh . h_{t-1}^T
//    as two M = 128 tiles with N = 16, the weights (bf16 hi/lo) resident in TENSOR MEMORY as the A operand
//    (TS form: 256 of the 512 columns), x and h as small K-major B operands in shared memory. The warp-level
//    MMA this replaces costs ~20 cycles per m16n8k16 per SM sub-partition on this part (measured with a
//    clock trace): 1,900 cycles per step for 16 sequences, against a few hundred for the 24 tcgen05.mma here.
//    * W_ih'.x runs one block (4 steps x 16 sequences = 64 accumulator columns per tile) ahead, off the serial
//      path: issued right after the recurrent MMAs of the last step of the previous block, it executes while
//      the cell update of that step is computed.
//    * W_hh.h_{t-1} accumulates ONTO the 16 columns of step t, so the accumulator holds the complete
//      pre-activations. The gate rows are permuted so that the 16 rows cell-warp w needs sit in the TMEM lanes
//      it may read (32 (w % 4) + 16 ((w / 4) % 2), tile w / 8); tcgen05.ld.16x256b returns the m16n8 fragment
//      layout (tools/tmem16_probe.cu), i.e. lane pairs (l, l ^ 4) hold i,g / f,o of one (unit, sequence) cell
//      exactly as the packed gate order (pack.cu seq_row_of) was built for.
//    Per step: [MMA warp] wait h_{t-1} -> 24 MMAs -> commit | [16 cell warps] wait -> tcgen05.ld -> cell update
//    (ex2/rcp) -> h_t as bf16 hi/lo into the swizzled B operand -> fence.proxy.async -> arrive. No block-wide
//    barrier inside the loop.
//
//   warps 0-15  cell update (one (unit, sequence) cell per lane and n8 half)
//   warps 16-19 weights -> TMEM once; x rows (global/L2) -> swizzled bf16 hi/lo B operand, 4 blocks ahead
//   warps 20-21 tcgen05 issuers (one per gate-row tile); between steps they write the h rows of the previous step
//               (fp32 rows or bf16 planes, half of the sequences each), fence and publish them
#include <cstdlib>
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int SQ = 16;           // sequences per CTA (N of the recurrent MMA)
constexpr int TS = 8;            // sequences per accumulator fragment
constexpr int NWARP = 16;        // cell warps
constexpr int NT = NWARP * 32;
constexpr int HELP_WARP0 = NWARP, N_HELP = 4, MMA_WARP = NWARP + N_HELP;
constexpr int N_MMA = 2;         // issuer warps: one per M = 128 gate-row tile
constexpr int THREADS = (MMA_WARP + N_MMA) * 32;
constexpr int FS = 68;           // fp32 row stride of the h staging buffers
constexpr int BLK = 4;           // steps per projection block (4 steps x 16 sequences = 64 columns per tile)
constexpr int NXB = 4;           // x-block buffers in shared memory
constexpr int X_PLANE = BLK * SQ * 128;               // 8 KB: [64 (step, sequence) rows][64 k] bf16
constexpr int H_PLANE = SQ * 128;                     // 2 KB: [16 sequences][64 k] bf16
constexpr int H_BUF = 2 * H_PLANE;                    // hi + lo
constexpr int OFF_X = 0, OFF_H = NXB * 2 * X_PLANE;   // two h operand buffers: h_t lives in buffer (t + 1) % 2
constexpr int FC_SMEM = 2 * 2 * 128 * 128 + 2 * kFeat * 128;          // fc role (below): two A blocks + the weight operand
constexpr int SMEM_BYTES = (OFF_H + 2 * H_BUF > FC_SMEM ? OFF_H + 2 * H_BUF : FC_SMEM) + 1024;
// tensor memory columns
constexpr int COL_ACC = 0;       // [2 buffers][2 tiles][BLK steps][16 sequences] fp32
constexpr int COL_WIH = 256;     // [2 tiles][2 planes][32]: lane = gate row, column = two consecutive k (bf16)
constexpr int COL_WHH = 384;

// 16 TMEM lanes x 8 columns -> the m16n8 accumulator fragment (rows lane/4, lane/4 + 8; columns 2 (lane % 4), +1)
__device__ __forceinline__ void tmem_ld_frag(uint32_t taddr, float (&r)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.16x256b.x1.b32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait_frag2(float (&r)[4], float (&q)[4]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+f"(r[0]), "+f"(r[1]), "+f"(r[2]), "+f"(r[3]), "+f"(q[0]), "+f"(q[1]), "+f"(q[2]), "+f"(q[3]) : : "memory");
}

#ifdef BSB_SEQ2_TRACE
// timeline of cell warp 0 and the MMA warp of CTAs 0 and 1 over steps [kTr0, kTr0 + kTrN)
constexpr int kTr0 = 200, kTrN = 8, kTrP = 8;
__device__ long long g_seq2_trace[2][2][kTrN][kTrP];
#define TR(k) do { if (tr != nullptr && t >= kTr0 && t < kTr0 + kTrN) tr[(t - kTr0) * kTrP + (k)] = clock64(); } while (0)
#else
#define TR(k) do { } while (0)
#endif

__device__ __forceinline__ void bar_cells() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// ------------------------------------------------------------------------------------------------------------
// Third CTA of a group (optional): the block's output Linear + residual, out = fc(h_layer1) + x (bsrnn.py:85-86 via
// 119-120), riding the same wavefront: it consumes layer 1's fp32 rows a block of 128 (step, sequence) rows at a
// time as layer 1 publishes them, one M = 128, N = 64, K = 64 split-bf16 tcgen05 tile per block (SS form), and
// writes the fp32 feature rows of the next model block. Saves a separate GEMM launch over the whole batch and the
// bf16-plane round trip of h; the CTA is idle most of the time (12 MMAs + 128 rows in and out per ~7 us).
constexpr int FC_WARPS = 4;                              // loader / epilogue warps (TMEM lane quarters)
constexpr int FC_THREADS = (FC_WARPS + 1) * 32;          // + issuer warp
constexpr int FC_A_PLANE = 128 * 128;                    // [128 rows][64 k] bf16
constexpr int FC_W_PLANE = kFeat * 128;                  // [64 n][64 k] bf16
constexpr int FC_OFF_A = 0, FC_OFF_W = 2 * 2 * FC_A_PLANE;

__device__ __forceinline__ void bar_fc() { asm volatile("bar.sync 2, %0;" ::"n"(FC_THREADS) : "memory"); }

__device__ void fc_role(const LstmSeq2Args& a, int grp, uint8_t* smem_gen, uint32_t smem_base) {
  __shared__ __align__(8) uint64_t bar_afull[2], bar_aempty[2], bar_accfull[2], bar_accfree[2];
  __shared__ __align__(16) float fc_bias[kFeat];
  __shared__ uint32_t fc_tmem_slot;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (w > FC_WARPS) return;
  const int sq = a.group, seq0 = grp * sq, T = a.nsteps;
  const int spb = 128 / sq;                                // steps per block
  const int nblk = (T + spb - 1) / spb;
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&bar_afull[i]), FC_WARPS * 32);
      mbar_init(smem_u32(&bar_aempty[i]), 1);
      mbar_init(smem_u32(&bar_accfull[i]), 1);
      mbar_init(smem_u32(&bar_accfree[i]), FC_WARPS * 32);
    }
    fence_mbar_init();
  }
  if (w == FC_WARPS) {
    tmem_alloc(smem_u32(&fc_tmem_slot), 128);
    tmem_relinquish();
  } else {
    // W_fc [64 n][64 k] fp32 -> bf16 hi/lo, K-major 128B-swizzled B operand; bias
    for (int e = tid; e < kFeat * kFeat / 2; e += FC_WARPS * 32) {
      const int n = e >> 5, k = (e & 31) * 2;
      const float2 v = __ldg(reinterpret_cast<const float2*>(a.fcw + (long long)n * a.ldfcw + k));
      uint32_t h, l;
      split2(v.x, v.y, h, l);
      const uint32_t off = (uint32_t)(n * 128 + (((k >> 3) ^ (n & 7)) << 4) + (k & 7) * 2);
      *reinterpret_cast<uint32_t*>(smem_gen + FC_OFF_W + off) = h;
      *reinterpret_cast<uint32_t*>(smem_gen + FC_OFF_W + FC_W_PLANE + off) = l;
    }
    if (tid < kFeat) fc_bias[tid] = __ldg(a.fcb + tid);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  bar_fc();
  tc_fence_after();
  const uint32_t tmem_base = fc_tmem_slot;

  if (w < FC_WARPS) {
    const int lseq = tid >> 4, part = tid & 15;            // loader: (sequence, 4 floats at 4 part)
    const int* flag = a.flags + a.ngroups + grp;           // layer 1's publications
    // epilogue: this thread's TMEM lane = row r = (step in block r / sq, sequence r % sq)
    const int r = tid, ei = r / sq, es = r - ei * sq;
    const int eg = seq0 + es;
    const bool evalid = eg < a.nseq;
    const long long eoff = evalid ? (long long)(eg / a.n_inner) * a.outer + (long long)(eg % a.n_inner) * a.inner : 0;
    const uint32_t tlane = tmem_base + ((uint32_t)(w * 32) << 16);
    auto epilogue = [&](int b) {
      mbar_wait(smem_u32(&bar_accfull[b & 1]), (uint32_t)((b >> 1) & 1));
      tc_fence_after();
      const int step = b * spb + ei;
      const bool ok = evalid && step < T;
      const float* res = a.x + eoff + (long long)step * a.step;
      float* dst = a.fc_out + eoff + (long long)step * a.step;
      // plane output: sequence eg = (row c, band), frame-row rr = c * T + step -> band's block (rr / 128, chunk = half), row rr % 128
      const long long rr = (long long)(eg / a.n_inner) * T + step;
      const int prow = (int)(rr & 127);
      uint8_t* pbase = reinterpret_cast<uint8_t*>(a.fc_planes) + (long long)(eg % a.n_inner) * a.fcp_band + (rr >> 7) * (4LL * kPlaneBlock);
      uint4 ph, pl;
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        uint32_t v[32];
        tmem_ld_32x32(tlane + (uint32_t)((b & 1) * kFeat + half * 32), v);
        tmem_ld_wait32(v);
        if (ok) {
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 x4 = __ldg(reinterpret_cast<const float4*>(res + half * 32 + q * 4));
            const float4 b4 = *reinterpret_cast<const float4*>(&fc_bias[half * 32 + q * 4]);
            float4 o;
            o.x = __uint_as_float(v[4 * q + 0]) + b4.x + x4.x;
            o.y = __uint_as_float(v[4 * q + 1]) + b4.y + x4.y;
            o.z = __uint_as_float(v[4 * q + 2]) + b4.z + x4.z;
            o.w = __uint_as_float(v[4 * q + 3]) + b4.w + x4.w;
            if (!a.fc_skip_rows) *reinterpret_cast<float4*>(dst + half * 32 + q * 4) = o;
            if (a.fc_planes != nullptr) {       // two float4 = one 16-byte unit (8 bf16) of the hi and of the lo plane
              if ((q & 1) == 0) {
                split2(o.x, o.y, ph.x, pl.x);
                split2(o.z, o.w, ph.y, pl.y);
              } else {
                split2(o.x, o.y, ph.z, pl.z);
                split2(o.z, o.w, ph.w, pl.w);
                uint8_t* pp = pbase + (long long)half * (2 * kPlaneBlock) + sw64(prow, q >> 1);
                *reinterpret_cast<uint4*>(pp) = ph;
                *reinterpret_cast<uint4*>(pp + a.fcp_lo) = pl;
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_accfree[b & 1]));
    };
    for (int b = 0; b < nblk; ++b) {
      const int need = min((b + 1) * spb, T);
      if (lane == 0) {
        while (ld_acquire(flag) < need) __nanosleep(200);
      }
      __syncwarp();
      if (b >= 2) mbar_wait(smem_u32(&bar_aempty[b & 1]), (uint32_t)(((b >> 1) - 1) & 1));   // MMAs of block b-2 retired
      uint8_t* a_hi = smem_gen + FC_OFF_A + (b & 1) * 2 * FC_A_PLANE;
      uint8_t* a_lo = a_hi + FC_A_PLANE;
#pragma unroll
      for (int pass = 0; pass < 2; ++pass) {
        float4 v[8];
        int rows[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int qq = pass * 8 + q;
          const int s = sq == SQ ? (qq >> 3) * TS + lseq : lseq;
          const int i = sq == SQ ? (qq & 7) : qq;
          rows[q] = i * sq + s;
          const int g = seq0 + s, step = b * spb + i;
          v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (g < a.nseq && step < T) {
            const long long off = (long long)(g / a.n_inner) * a.outer + (long long)(g % a.n_inner) * a.inner + (long long)step * a.step;
            v[q] = __ldcg(reinterpret_cast<const float4*>(a.hout + off + part * 4));
          }
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          uint2 h, l;
          split2(v[q].x, v[q].y, h.x, l.x);
          split2(v[q].z, v[q].w, h.y, l.y);
          const uint32_t off = (uint32_t)(rows[q] * 128 + (((part >> 1) ^ (rows[q] & 7)) << 4) + (part & 1) * 8);
          *reinterpret_cast<uint2*>(a_hi + off) = h;
          *reinterpret_cast<uint2*>(a_lo + off) = l;
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(smem_u32(&bar_afull[b & 1]));
      if (b > 0) epilogue(b - 1);
    }
    epilogue(nblk - 1);
  } else {
    const uint32_t idesc = make_idesc_bf16(128, kFeat);
    const uint64_t w_hi = make_sw128_desc(smem_base + FC_OFF_W);
    const uint64_t w_lo = make_sw128_desc(smem_base + FC_OFF_W + FC_W_PLANE);
    for (int b = 0; b < nblk; ++b) {
      mbar_wait(smem_u32(&bar_afull[b & 1]), (uint32_t)((b >> 1) & 1));
      if (b >= 2) mbar_wait(smem_u32(&bar_accfree[b & 1]), (uint32_t)(((b >> 1) - 1) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint64_t a_hi = make_sw128_desc(smem_base + FC_OFF_A + (b & 1) * 2 * FC_A_PLANE);
        const uint64_t a_lo = make_sw128_desc(smem_base + FC_OFF_A + (b & 1) * 2 * FC_A_PLANE + FC_A_PLANE);
        const uint32_t d = tmem_base + (uint32_t)((b & 1) * kFeat);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adv = (uint64_t)(kk * 32 >> 4);
          mma_bf16_ss(d, a_lo + adv, w_hi + adv, idesc, kk != 0);   // small terms first
          mma_bf16_ss(d, a_hi + adv, w_lo + adv, idesc, 1);
          mma_bf16_ss(d, a_hi + adv, w_hi + adv, idesc, 1);
        }
        mma_commit(smem_u32(&bar_accfull[b & 1]));
        mma_commit(smem_u32(&bar_aempty[b & 1]));
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  bar_fc();
  if (w == FC_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 128);
  }
}

__global__ void __launch_bounds__(THREADS, 1)
lstm_seq2_kernel(const __grid_constant__ LstmSeq2Args a) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(16) float hF[3][SQ][FS];              // [step % 3][sequence][unit]
  __shared__ __align__(8) uint64_t bar_xfull[NXB], bar_xempty[NXB], bar_gates[2], bar_h;
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x;
  const int w = tid >> 5, lane = tid & 31;
  const int nrole = a.fuse_fc ? 3 : 2;
  // Logical CTA index = a ticket taken at start, not blockIdx.x: layer-1 and fc CTAs spin on flags published by the
  // CTA with the next-lower logical index of their group. With tickets that producer has always STARTED before its
  // consumer, whatever order the hardware dispatches CTAs in and even when the grid exceeds the SM count (the unfused
  // 16-sequence mode: 150 CTAs at 1,200 sequences), so a resident consumer can never wait on a CTA that has no SM.
  __shared__ int ticket_s;
  if (tid == 0) ticket_s = atomicAdd(a.flags + 2 * a.ngroups, 1);
  __syncthreads();
  const int cta = ticket_s;
  const int layer = cta % nrole, grp = cta / nrole;                    // role 0/1 = LSTM layer, 2 = output Linear
  const uint32_t smem_base0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
  if (layer == 2) {
    fc_role(a, grp, smem_raw + (smem_base0 - smem_u32(smem_raw)), smem_base0);
    return;
  }
  const int sq = a.group;          // sequences this CTA really owns: 16, or 8 (only fragment 0) when that still fits the SMs
  const int seq0 = grp * sq;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int T = a.nsteps;
  const int nblk = (T + BLK - 1) / BLK;
  const float* xin = layer ? a.hmid : a.x;
  float* out = layer ? a.hout : a.hmid;
  void* planes = (layer && !a.fuse_fc) ? a.hplanes : nullptr;
  const bool publish = layer == 0 || a.fuse_fc;          // somebody downstream in this launch reads this layer's rows
  int* my_flag = a.flags + layer * a.ngroups + grp;       // rows of steps < *my_flag of this group are in global memory

  if (tid == 0) {
    for (int i = 0; i < NXB; ++i) {
      mbar_init(smem_u32(&bar_xfull[i]), N_HELP * 32);
      mbar_init(smem_u32(&bar_xempty[i]), N_MMA);
    }
    mbar_init(smem_u32(&bar_gates[0]), 1);
    mbar_init(smem_u32(&bar_gates[1]), 1);
    mbar_init(smem_u32(&bar_h), NWARP);
    fence_mbar_init();
  }
  if (w == MMA_WARP) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  // ---------------- prologue: weights -> TMEM (helpers), h_{-1} -> B operand (cell warps) ----------------
  // cell of this lane (cell warps): unit 4w + gr/2 of sequence 2tq + (gr&1) of each n8 half
  const int gr = lane >> 2, tq = lane & 3;
  const bool odd = gr & 1;
  const int unit = 4 * w + (gr >> 1);
  const int sl = 2 * tq + (odd ? 1 : 0);
  float c[2] = {0.f, 0.f};
  bool valid[2] = {false, false};
  uint32_t hoffs[2] = {0, 0};    // byte offset of this lane's h element in a plane of the B operand
  if (w < NWARP) {
    const float* h0p = a.h0[layer];
    const float* c0p = a.c0[layer];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int s = j * TS + sl, gseq = seq0 + s;
      valid[j] = s < sq && gseq < a.nseq;
      hoffs[j] = (uint32_t)(s * 128 + (((unit >> 3) ^ (s & 7)) << 4) + (unit & 7) * 2);
      float h0 = 0.f;
      if (valid[j] && h0p != nullptr) {
        h0 = __ldg(h0p + (long long)gseq * kFeat + unit);
        c[j] = __ldg(c0p + (long long)gseq * kFeat + unit);
      }
      uint16_t hh, hl;
      split1(h0, hh, hl);
      *reinterpret_cast<uint16_t*>(smem_gen + OFF_H + hoffs[j]) = hh;
      *reinterpret_cast<uint16_t*>(smem_gen + OFF_H + H_PLANE + hoffs[j]) = hl;
      *reinterpret_cast<uint16_t*>(smem_gen + OFF_H + H_BUF + hoffs[j]) = 0;
      *reinterpret_cast<uint16_t*>(smem_gen + OFF_H + H_BUF + H_PLANE + hoffs[j]) = 0;
    }
    fence_proxy_async_smem();
  } else if (w < MMA_WARP) {
    // Weights -> TMEM: lane l of tile tl holds fragment row i = l % 16 of cell warp ww = tl*8 + ((l/16)%2)*4 + l/32 (so
    // that warp reads its own rows); 32-bit column cc of matrix m (0 = W_ih', 1 = W_hh) holds k = 2cc, 2cc+1. Source: the mma.sync A-fragment
    // image of [W_ih' | W_hh] (pack.cu pack_lstm_seq_kernel): [warp][k-tile 0..7][plane][lane][reg].
    const int l = (w & 3) * 32 + lane;
    const int i = l & 15;
    const uint32_t* img = reinterpret_cast<const uint32_t*>(a.wfrag[layer]);
    const uint32_t tl_addr = tmem_base + ((uint32_t)((w & 3) * 32) << 16);
#pragma unroll 1
    for (int q = 0; q < 8; ++q) {                 // (matrix, tile, plane)
      const int m = q >> 2, tile = (q >> 1) & 1, plane = q & 1;
      const int ww = (tile << 3) | (((l >> 4) & 1) << 2) | (l >> 5);
      uint32_t v[32];
#pragma unroll
      for (int cc = 0; cc < 32; ++cc) {
        const int kt = m * 4 + (cc >> 3);
        const int lf = (i & 7) * 4 + (cc & 3);
        const int reg = (i >= 8 ? 1 : 0) + (((cc & 7) >= 4) ? 2 : 0);
        v[cc] = __ldg(img + ((size_t)(ww * 8 + kt) * 2 + plane) * 128 + lf * 4 + reg);
      }
      const uint32_t col = (uint32_t)((m ? COL_WHH : COL_WIH) + (tile * 2 + plane) * 32);
      tmem_st_32x16(tl_addr + col, v);
      tmem_st_32x16(tl_addr + col + 16, v + 16);
    }
    tmem_st_wait();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (w < NWARP) {
    // ================= cell warps =================
    const float b_lo = __ldg(a.bias[layer] + w * 16 + gr), b_hi = __ldg(a.bias[layer] + w * 16 + gr + 8);
    float* hNp = a.hN[layer];
    float* cNp = a.cN[layer];
    // pre-activations of fragment j -> cell update, h to the B operand and the staging row
    auto cell = [&](const float (&g)[4], int j, int t) {
      const float acc0 = g[0] + b_lo, acc1 = g[1] + b_lo, acc2 = g[2] + b_hi, acc3 = g[3] + b_hi;
      const float s0 = odd ? acc0 : acc1, s1 = odd ? acc2 : acc3;
      const float r0 = __shfl_xor_sync(0xffffffffu, s0, 4), r1 = __shfl_xor_sync(0xffffffffu, s1, 4);
      const float ti = fminf(odd ? r0 : acc0, 40.f), tg = fminf(odd ? r1 : acc2, 40.f);
      const float tf = fminf(odd ? acc1 : r0, 40.f), to = fminf(odd ? acc3 : r1, 40.f);
      const float ei = ex2_approx(ti), ef = ex2_approx(tf), eg = ex2_approx(tg), eo = ex2_approx(to);
      const float pi = 1.f + ei, pf = 1.f + ef, pg = 1.f + eg;
      const float pig = pi * pg;
      const float cn = fmaf(c[j], pig, (1.f - eg) * pf) * rcp_approx(pf * pig);
      c[j] = cn;
      const float ec = ex2_approx(fminf(cn * -2.885390081777927f, 40.f));
      const float h = (1.f - ec) * rcp_approx((1.f + eo) * (1.f + ec));
      uint16_t hh, hl;
      split1(h, hh, hl);
      // the MMAs of step t (either tile's, possibly not even issued yet by the other issuer warp) read h_{t-1} from
      // the other buffer: no write-after-read hazard, whatever the skew between the two issuers
      uint8_t* hb = smem_gen + OFF_H + ((t + 1) & 1) * H_BUF;
      *reinterpret_cast<uint16_t*>(hb + hoffs[j]) = hh;
      *reinterpret_cast<uint16_t*>(hb + H_PLANE + hoffs[j]) = hl;
      hF[t % 3][j * TS + sl][unit] = h;
      if (t == T - 1 && valid[j] && hNp != nullptr) {
        const int gseq = seq0 + j * TS + sl;
        hNp[(long long)gseq * kFeat + unit] = h;
        cNp[(long long)gseq * kFeat + unit] = cn;
      }
    };
    // this warp's 16 gate rows: its own TMEM lanes, tile w / 8; columns of step t: buffer, step in block, sequence
    const uint32_t tw = tmem_base + ((uint32_t)(32 * (w & 3) + 16 * ((w >> 2) & 1)) << 16) + (uint32_t)(COL_ACC + (w >> 3) * (BLK * SQ));
#ifdef BSB_SEQ2_TRACE
    long long* tr = (blockIdx.x < 2 && w == 0 && lane == 0) ? &g_seq2_trace[blockIdx.x][0][0][0] : nullptr;
#endif
    // h rows leave one step late, after this warp's arrive: thread tid owns (sequence tid / 16, 16-byte part tid % 16)
    const int fseq = (tid >> 4) & 15, fpart = tid & 15;
    const int fg = seq0 + fseq;
    const bool fvalid = tid < sq * 16 && fg < a.nseq && !(a.exp & 16);
    const long long frow = fvalid ? (long long)(fg / a.n_inner) * a.outer + (long long)(fg % a.n_inner) * a.inner : 0;
    auto flush = [&](int t) {
      if (!fvalid) return;
      const long long hoff = frow + (long long)t * a.step;
      const float4 v = *reinterpret_cast<const float4*>(&hF[t % 3][fseq][fpart * 4]);
      if (planes != nullptr) {
        // A operand of the Linear that follows (kernels.cuh planes: [m-tile][32-column chunk][128 x 64 B]): row =
        // output row index, columns 4 part .. +3 as bf16 hi and lo
        const long long m = hoff >> 6;
        const int r = (int)(m & 127);
        uint8_t* p = reinterpret_cast<uint8_t*>(planes) + ((m >> 7) * 2 + (fpart >> 3)) * (2 * kPlaneBlock) +
                     sw64(r, (fpart & 7) >> 1) + (fpart & 1) * 8;
        uint2 h, l;
        split2(v.x, v.y, h.x, l.x);
        split2(v.z, v.w, h.y, l.y);
        *reinterpret_cast<uint2*>(p) = h;
        *reinterpret_cast<uint2*>(p + a.hp_lo) = l;
      } else {
        *reinterpret_cast<float4*>(out + hoff + fpart * 4) = v;
      }
    };
    for (int t = 0; t < T; ++t) {
      TR(0);
      mbar_wait(smem_u32(&bar_gates[w >> 3]), (uint32_t)(t & 1));    // (polling with test_wait measured slower)
      tc_fence_after();
      TR(1);
      const uint32_t col = (uint32_t)(((t / BLK) & 1) * (2 * BLK * SQ) + (t % BLK) * SQ);
      float ga[4], gb[4];
      tmem_ld_frag(tw + col, ga);
      tmem_ld_frag(tw + col + TS, gb);
      tmem_ld_wait_frag2(ga, gb);
      TR(2);
      if (!(a.exp & 1)) {
        cell(ga, 0, t);
        if (sq > TS) cell(gb, 1, t);     // (h of the unused half stays zero in the B operand)
      }
      TR(3);
      fence_proxy_async_smem();      // h_t (generic-proxy stores) -> visible to the tensor core's operand reads
      tc_fence_before();             // this warp's tcgen05.ld of the step are ordered before the next MMAs
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&bar_h));
      TR(4);
      // off the chain: rows of step t-1 (every cell warp wrote its units before its arrive on bar_h(t-1), which the
      // gates of step t waited for)
      if (t > 0) flush(t - 1);
      // Publication of the rows of steps < t-2, every BLK steps: every cell warp ran flush(t-3) before its arrive on
      // bar_h(t-1) in program order, and this thread has observed the gates of step t, i.e. all those arrives
      // (release/acquire at CTA scope through the mbarriers; the device-scope fence below is cumulative over them).
      if (publish && tid == 0 && t > 2 && ((t - 2) % BLK) == 0) {
        __threadfence();
        *reinterpret_cast<volatile int*>(my_flag) = t - 2;
      }
    }
    // last rows: cell warps only (named barrier 1), then one thread publishes everything
    bar_cells();
    flush(T - 1);
    if (publish) {
      bar_cells();
      if (tid == 0) {
        __threadfence();
        *reinterpret_cast<volatile int*>(my_flag) = T;
      }
    }
  } else if (w < MMA_WARP) {
    // ================= helpers: x rows -> bf16 hi/lo B operand of the projection =================
    const int j = (w - HELP_WARP0) * 32 + lane;     // 0..127: (sequence j/16 and j/16 + 8, 4 floats at 4*(j%16))
    const int lseq = j >> 4, part = j & 15;
    const float* xptr[2];
    bool lvalid[2];
#pragma unroll
    for (int s2 = 0; s2 < 2; ++s2) {
      const int lg = seq0 + s2 * TS + lseq;
      lvalid[s2] = s2 * TS < sq && lg < a.nseq;
      xptr[s2] = xin;
      if (lvalid[s2]) xptr[s2] += (long long)(lg / a.n_inner) * a.outer + (long long)(lg % a.n_inner) * a.inner + part * 4;
    }
    const int* flag = a.flags + grp;               // layer 0's publications
    for (int b = 0; b < nblk; ++b) {            // 4 steps x 16 sequences x 64 floats -> planes of buffer b % NXB
      if (b >= NXB) mbar_wait(smem_u32(&bar_xempty[b % NXB]), ((b / NXB) - 1) & 1);   // MMAs of block b - NXB retired
      uint8_t* x_hi = smem_gen + OFF_X + (b % NXB) * 2 * X_PLANE;
      uint8_t* x_lo = x_hi + X_PLANE;
      if (layer == 1 && !(a.exp & 8)) {         // layer 0 must have published the rows of this block
        const int need = min((b + 1) * BLK, T);
        if (lane == 0) {
          while (ld_acquire(flag) < need) __nanosleep(64);
        }
        __syncwarp();
      }
      float4 v[2][BLK];
#pragma unroll
      for (int s2 = 0; s2 < 2; ++s2)
#pragma unroll
        for (int i = 0; i < BLK; ++i) {
          const int step = b * BLK + i;
          const float4* p = reinterpret_cast<const float4*>(xptr[s2] + (long long)step * a.step);
          v[s2][i] = (lvalid[s2] && step < T && !(a.exp & 8)) ? (layer == 1 ? __ldcg(p) : __ldg(p)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
      for (int s2 = 0; s2 < 2; ++s2)
#pragma unroll
        for (int i = 0; i < BLK; ++i) {
          const int r = i * SQ + s2 * TS + lseq;              // row = accumulator column = (step in block, sequence)
          uint2 h, l;
          split2(v[s2][i].x, v[s2][i].y, h.x, l.x);
          split2(v[s2][i].z, v[s2][i].w, h.y, l.y);
          const uint32_t off = (uint32_t)(r * 128 + (((part >> 1) ^ (r & 7)) << 4) + (part & 1) * 8);
          *reinterpret_cast<uint2*>(x_hi + off) = h;
          *reinterpret_cast<uint2*>(x_lo + off) = l;
        }
      fence_proxy_async_smem();
      mbar_arrive(smem_u32(&bar_xfull[b % NXB]));
    }
  } else {
    // ================= tcgen05 issuers: warp MMA_WARP + tile owns the M = 128 gate-row tile `tile` =================
    // (the MMAs of the two tiles touch disjoint accumulator columns; two issuing threads halve the issue time
    // of a step, and the cell warps of a tile start as soon as their own 12 MMAs have retired)
    const int tile = w - MMA_WARP;
    const uint32_t idesc_x = make_idesc_bf16(128, BLK * SQ);
    const uint32_t idesc_h = make_idesc_bf16(128, SQ);
    const uint64_t h_hi0 = make_sw128_desc(smem_base + OFF_H);
    const uint64_t h_lo0 = make_sw128_desc(smem_base + OFF_H + H_PLANE);
    const uint32_t wih_hi = tmem_base + (uint32_t)(COL_WIH + (tile * 2 + 0) * 32), wih_lo = wih_hi + 32;
    const uint32_t whh_hi = tmem_base + (uint32_t)(COL_WHH + (tile * 2 + 0) * 32), whh_lo = whh_hi + 32;
    // W_ih'[tile] . x_block^T -> accumulator buffer b % 2 (overwrites: the block two back is consumed). ~50 cycles per
    // N = 64 MMA in TS form (~70 in SS form with the weights in shared memory; issuing it in four parts, one per
    // step, cost more than it hid).
    auto project = [&](int b) {
      mbar_wait(smem_u32(&bar_xfull[b % NXB]), (uint32_t)((b / NXB) & 1));
      tc_fence_after();
      if (elect_one()) {
        const uint64_t x_hi = make_sw128_desc(smem_base + OFF_X + (b % NXB) * 2 * X_PLANE);
        const uint64_t x_lo = make_sw128_desc(smem_base + OFF_X + (b % NXB) * 2 * X_PLANE + X_PLANE);
        const uint32_t d = tmem_base + (uint32_t)(COL_ACC + (b & 1) * (2 * BLK * SQ) + tile * (BLK * SQ));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint64_t adv = (uint64_t)(kk * 32 >> 4);       // 16 bf16 = 32 bytes along K
          mma_bf16_ts(d, wih_hi + kk * 8, x_lo + adv, idesc_x, kk != 0);   // small terms first
          mma_bf16_ts(d, wih_lo + kk * 8, x_hi + adv, idesc_x, 1);
          mma_bf16_ts(d, wih_hi + kk * 8, x_hi + adv, idesc_x, 1);
        }
        mma_commit(smem_u32(&bar_xempty[b % NXB]));
      }
      __syncwarp();
    };
    project(0);
#ifdef BSB_SEQ2_TRACE
    long long* tr = (blockIdx.x < 2 && tile == 0 && lane == 0) ? &g_seq2_trace[blockIdx.x][1][0][0] : nullptr;
#endif
    for (int t = 0; t < T; ++t) {
      TR(0);
      if (t > 0) {
        mbar_wait(smem_u32(&bar_h), (uint32_t)((t - 1) & 1));      // h_{t-1} is in the B operand
        tc_fence_after();
      }
      TR(1);
      if (elect_one()) {
        if (!(a.exp & 2)) {
          const uint32_t d = tmem_base + (uint32_t)(COL_ACC + ((t / BLK) & 1) * (2 * BLK * SQ) + tile * (BLK * SQ) + (t % BLK) * SQ);
          const uint64_t hsel = (uint64_t)((t & 1) * (H_BUF >> 4));      // h_{t-1} is in buffer t % 2
          const uint64_t h_hi = h_hi0 + hsel, h_lo = h_lo0 + hsel;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t adv = (uint64_t)(kk * 32 >> 4);
            mma_bf16_ts(d, whh_hi + kk * 8, h_lo + adv, idesc_h, 1);
            mma_bf16_ts(d, whh_lo + kk * 8, h_hi + adv, idesc_h, 1);
            mma_bf16_ts(d, whh_hi + kk * 8, h_hi + adv, idesc_h, 1);
          }
        }
        mma_commit(smem_u32(&bar_gates[tile]));
      }
      __syncwarp();
      TR(2);
      // the projection of the next block executes under the cell update of this step
      if ((t % BLK) == BLK - 1 && t / BLK + 1 < nblk) project(t / BLK + 1);
      TR(3);
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

#ifdef BSB_SEQ2_TRACE
extern "C" int bsb_debug_seq2_trace(long long* host, int n) {
  if (n != 2 * 2 * kTrN * kTrP) return 2 * 2 * kTrN * kTrP;
  return (int)cudaMemcpyFromSymbol(host, g_seq2_trace, sizeof(long long) * n);
}
#endif

cudaError_t launch_lstm_seq2(const LstmSeq2Args& a, cudaStream_t s, int* fused) {
  if (fused) *fused = 0;
  if (a.nseq <= 0 || a.nsteps <= 0) return cudaSuccess;
  static SmemOptIn opt;
  int dev = 0;
  cudaError_t e = opt_in_smem(lstm_seq2_kernel, SMEM_BYTES, opt, &dev);
  if (e != cudaSuccess) return e;
  const int sms = sm_count(dev);
  // 8 sequences per CTA halve the cell update of a step (one fragment per lane instead of two); taken when all
  // CTAs are still co-resident, i.e. for short batches and long-form clips (cfg 1, 3, 5)
  const int g8 = (a.nseq + TS - 1) / TS, g16 = (a.nseq + SQ - 1) / SQ;
  static const bool no_fc = std::getenv("BSB_SEQ2_NOFC") != nullptr;      // A/B switch, read once
  const bool can_fuse = a.fcw != nullptr && a.fc_out != nullptr && a.fc_out != a.x && !no_fc;
  // preference: fused output Linear (third CTA per group) as long as all CTAs are co-resident, small groups first
  int group = SQ, fuse = 0;
  if (can_fuse && 3 * g8 <= sms) { group = TS; fuse = 1; }
  else if (2 * g8 <= sms) { group = TS; }
  else if (can_fuse && 3 * g16 <= sms) { fuse = 1; }
  const int groups = (a.nseq + group - 1) / group;
  static const int exp_mode = std::getenv("BSB_SEQ2_EXP") ? std::atoi(std::getenv("BSB_SEQ2_EXP")) : 0;
  LstmSeq2Args b = a;
  b.exp = exp_mode;
  b.group = group;
  b.fuse_fc = fuse;
  b.ngroups = groups;
  e = cudaMemsetAsync(a.flags, 0, sizeof(int) * (2 * groups + 1), s);      // publication flags + the CTA ticket counter
  if (e != cudaSuccess) return e;
  lstm_seq2_kernel<<<(fuse ? 3 : 2) * groups, THREADS, SMEM_BYTES, s>>>(b);
  if (fused) *fused = fuse;
  count_launch();
  return cudaGetLastError();
}

}  // namespace bsb
