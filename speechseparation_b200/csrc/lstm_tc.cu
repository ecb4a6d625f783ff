// One LSTM layer (nn.LSTM semantics, bsrnn.py:70) of one direction on the 5th-gen tensor cores,
// input projection fused: for a tile of 128 sequences a CTA computes, step after step,
//
//   gates[128 x 256] = [x_t | h_{t-1}] . [W_ih' | W_hh]^T         (tcgen05.mma, fp32 in TMEM)
//   c_t = sig(f).c_{t-1} + sig(i).tanh(g),  h_t = sig(o).tanh(c_t)   (epilogue warps)
//
// so neither the gate pre-activations (2 KB per sequence-step) nor fc_in's output ever reach HBM
// (fc_in, bsrnn.py:80, is folded into W_ih' at pack time, layout.h). One CTA per SM:
//
//   shared memory : the whole [256 x (KIN+64)] weight matrix as bf16 hi/lo planes, 128B-swizzled
//                   K-major (128 KB for KIN=64, 192 KB for KIN=128), loaded once per CTA
//   tensor memory : cols [0,256) gate accumulators; [256,320) h_{t-1} as bf16 hi|lo (A operand);
//                   [320,320+KIN) x_t as bf16 hi|lo (A operand). A operands never touch smem:
//                   the thread that owns sequence row r writes them with tcgen05.st (lane = row).
//   warps 0-7     : epilogue; warp w owns TMEM lanes 32(w%4).. and, of every 16-unit slot, units 8(w/4)..+7
//   warp  8       : one thread issues the MMAs: 3 bf16 products per fp32 product (x = hi + lo,
//                   hi.hi + hi.lo + lo.hi), M=128, N=256, K=16
//
// Only W_hh.h_{t-1} is on the serial path. The W_ih'.x_t half of step t+1 is issued while the epilogue of
// step t is still running: the accumulator is handed back to the tensor core in four 64-column slots (16
// hidden units each) as soon as all epilogue warps have read a slot, and the x part refills it in two halves of
// two slots (N = 128 per MMA); when h_t is complete only the 12 MMAs of the h part (N = 256) remain before
// the next epilogue.
//
// Persistent grid (round 2): one CTA per SM, the CTAs split between the directions (even CTAs forward, odd reverse)
// and each walks the 128-sequence tiles of its direction, so the weight image (128-192 KB) is loaded ONCE per SM by
// four to six 32 KB bulk copies instead of once per tile with plain loads (432 CTAs at cfg 2: ~30-45 us per launch).
// Inside a step the epilogue keeps the tcgen05.ld of the NEXT 64-column slot and the tcgen05.st of x_{t+1} in
// flight under the cell arithmetic of the current slot (measured before: with no MMAs and no cell arithmetic the
// launch still took 57-63 % of its time -- four serialised TMEM loads, a TMEM store and their waits per step).
//
// Gate rows are permuted to n = 4*unit + gate so a thread reads the four gates of a unit from
// four adjacent TMEM columns, and pre-scaled by -log2(e) (i,f,o) / -2.log2(e) (g) so the
// accumulator is directly the ex2 argument. The cell update shares reciprocals
// (c' = [c.(1+ei)(1+eg) + (1-eg)(1+ef)] / [(1+ef)(1+ei)(1+eg)]): 5 ex2 + 2 rcp per cell.
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int TILE = 128;
constexpr int N_EPI_WARPS = 8;
constexpr int THREADS = (N_EPI_WARPS + 1) * 32;
constexpr int COL_ACC = 0, COL_AH = 256, COL_AX = 320;
constexpr int CHUNK_BYTES = kGates * 128;          // one 64-wide K chunk of one plane: 32 KB
constexpr float kClamp = 40.f;                     // ex2 argument cap: (1+2^40)^3 stays finite

template <int KIN>
struct Cfg {
  static constexpr int NCH = (KIN + kFeat) / 64;            // K chunks: x part then the h part
  static constexpr int PLANE = NCH * CHUNK_BYTES;
  static constexpr int W_BYTES = 2 * PLANE;
  static constexpr int SMEM = W_BYTES + kGates * 4 + 1024;  // + bias + alignment slack
  static constexpr int XH = KIN / 2;                        // x floats per epilogue thread
  // keep the tcgen05.ld of the next slot in flight under the cell arithmetic of the current one (second register set)
#ifndef BSB_LSTM_PIPE128
#define BSB_LSTM_PIPE128 0
#endif
  static constexpr bool PIPE = KIN == 64 || BSB_LSTM_PIPE128;
};

template <int KIN>
__global__ void __launch_bounds__(THREADS, 1)
lstm_tc_kernel(const __grid_constant__ LstmTcArgs a) {
  using C = Cfg<KIN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_ready, bar_full, bar_slot[4], bar_w;
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  // persistent: CTA b serves direction b % ndir and is the (b / ndir)-th of the gridDim.x / ndir CTAs of that direction
  const int dir = (int)blockIdx.x % a.ndir;
  const int cta_slot = (int)blockIdx.x / a.ndir;
  const int cta_slots = (int)gridDim.x / a.ndir;
  const int ntiles = (a.nseq + TILE - 1) / TILE;
  const bool reverse = dir == 1;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  float* bias_s = reinterpret_cast<float*>(smem_gen + C::W_BYTES);

  if (tid == 0) {
    mbar_init(smem_u32(&bar_ready), N_EPI_WARPS * 32);
    mbar_init(smem_u32(&bar_full), 1);
    mbar_init(smem_u32(&bar_w), 1);
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar_slot[i]), N_EPI_WARPS * 32);
    fence_mbar_init();
    // weights: the packed image is already the swizzled shared-memory layout; one bulk copy per 32 KB chunk
    mbar_arrive_expect_tx(smem_u32(&bar_w), C::W_BYTES);
    const uint8_t* src = reinterpret_cast<const uint8_t*>(a.wimg[dir]);
    for (int i = 0; i < C::W_BYTES / CHUNK_BYTES; ++i)
      bulk_g2s(smem_base + i * CHUNK_BYTES, src + (size_t)i * CHUNK_BYTES, CHUNK_BYTES, smem_u32(&bar_w));
  }
  if (warp == N_EPI_WARPS) {
    tmem_alloc(smem_u32(&tmem_slot), 512);
    tmem_relinquish();
  }
  for (int i = tid; i < kGates; i += THREADS) bias_s[i] = __ldg(a.bias[dir] + i);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp < N_EPI_WARPS) {
    // ================= epilogue / operand producer =================
    const int q = warp & 3, hf = warp >> 2;
    const int row = q * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(q * 32) << 16);
    uint32_t g = 0;                           // steps done by this CTA: the barrier phases run on across tiles
    for (int tile = cta_slot; tile < ntiles; tile += cta_slots) {
    const long long seq = (long long)tile * TILE + row;
    const bool valid = seq < a.nseq;
    const float* xrow = a.x + seq * a.x_seq + hf * C::XH;
    float* hrow = a.hout + seq * a.h_seq + dir * kFeat + hf * 8;

    float xr[C::XH];
    auto load_x = [&](int step) {
      if (valid) {
        const float4* p = reinterpret_cast<const float4*>(xrow + (long long)step * a.x_step);
#pragma unroll
        for (int i = 0; i < C::XH / 4; ++i) {
          const float4 v = __ldg(p + i);
          xr[4 * i + 0] = v.x; xr[4 * i + 1] = v.y; xr[4 * i + 2] = v.z; xr[4 * i + 3] = v.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < C::XH; ++i) xr[i] = 0.f;
      }
    };
    auto store_x = [&]() {     // bf16 hi/lo pairs -> TMEM A operand columns of this thread's k range
#pragma unroll
      for (int blk = 0; blk < C::XH / 32; ++blk) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) split2(xr[blk * 32 + 2 * i], xr[blk * 32 + 2 * i + 1], hi[i], lo[i]);
        const uint32_t col = COL_AX + hf * (KIN / 4) + blk * 16;
        tmem_st_32x16(tlane + col, hi);
        tmem_st_32x16(tlane + col + KIN / 2, lo);
      }
    };

    float c[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) c[i] = 0.f;

    // round 0 of a tile: x of the first step; h_{-1} = 0. (Every MMA of the previous tile has completed: its last
    // bar_full was waited for, so the x columns and the accumulator are free.)
    load_x(reverse ? a.nsteps - 1 : 0);
    store_x();
    tmem_st_wait();
    tc_fence_before();
#pragma unroll
    for (int sl = 0; sl < 4; ++sl) mbar_arrive(smem_u32(&bar_slot[sl]));
    mbar_arrive(smem_u32(&bar_ready));

    for (int it = 0; it < a.nsteps; ++it, ++g) {
      const int step = reverse ? a.nsteps - 1 - it : it;
      const bool more = it + 1 < a.nsteps;
      if (more) load_x(reverse ? step - 1 : step + 1);       // in flight while the MMAs run
      mbar_wait(smem_u32(&bar_full), g & 1);
      tc_fence_after();
      // slot = 16 units = 64 accumulator columns, 8 units (32 columns) per thread. The load of slot sl + 1 is issued
      // before the arithmetic of slot sl, and the store of x_{t+1} (all MMAs of this step are done: the x columns
      // are free) runs under the load of slot 0.
      uint32_t ra[32], rb[C::PIPE ? 32 : 1];
      if (!C::PIPE && more) {                                // (KIN = 128: no registers to spare for the overlap)
        store_x();
        tmem_st_wait();
      }
      tmem_ld_32x32(tlane + COL_ACC + hf * 32, ra);
      if (C::PIPE && more) store_x();
      float* hdst = hrow + (long long)step * a.h_step;
#pragma unroll
      for (int sl = 0; sl < 4; ++sl) {
        uint32_t (&r)[32] = (C::PIPE && (sl & 1)) ? reinterpret_cast<uint32_t (&)[32]>(rb) : ra;
        uint32_t (&rn)[32] = (C::PIPE && !(sl & 1)) ? reinterpret_cast<uint32_t (&)[32]>(rb) : ra;
        if (!C::PIPE && sl > 0) tmem_ld_32x32(tlane + COL_ACC + sl * 64 + hf * 32, r);
        tmem_ld_wait32(r);
        if (C::PIPE && sl == 0 && more) tmem_st_wait();      // x_{t+1} is in place
        if (C::PIPE && sl < 3) tmem_ld_32x32(tlane + COL_ACC + (sl + 1) * 64 + hf * 32, rn);
        if (more) {                                          // this thread is done with the slot
          tc_fence_before();
          mbar_arrive(smem_u32(&bar_slot[sl]));
        }
        float h[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b = *reinterpret_cast<const float4*>(bias_s + sl * 64 + hf * 32 + j * 4);
          const float ti = fminf(__uint_as_float(r[4 * j + 0]) + b.x, kClamp);
          const float tf = fminf(__uint_as_float(r[4 * j + 1]) + b.y, kClamp);
          const float tg = fminf(__uint_as_float(r[4 * j + 2]) + b.z, kClamp);
          const float to = fminf(__uint_as_float(r[4 * j + 3]) + b.w, kClamp);
          const float ei = ex2_approx(ti), ef = ex2_approx(tf), eg = ex2_approx(tg), eo = ex2_approx(to);
          const float pi = 1.f + ei, pf = 1.f + ef, pg = 1.f + eg;
          const float pig = pi * pg;
          const float num = fmaf(c[sl * 8 + j], pig, (1.f - eg) * pf);
          const float cn = num * rcp_approx(pf * pig);
          c[sl * 8 + j] = cn;
          const float ec = ex2_approx(fminf(cn * -2.885390081777927f, kClamp));
          h[j] = (1.f - ec) * rcp_approx((1.f + eo) * (1.f + ec));
        }
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split2(h[2 * i], h[2 * i + 1], hi[i], lo[i]);
        if (valid) {
          if (a.hplanes != nullptr) {
            // A operand of the Linear that follows (kernels.cuh planes: [m-tile][32-column chunk][128 x 64 B]): row
            // seq*nsteps + step; units 16 sl + 8 hf .. +7 of direction dir = chunk 2 dir + sl/2, 16-byte unit 2 (sl&1) + hf
            const long long m = seq * a.nsteps + step;
            const int r128 = (int)(m & 127);
            uint8_t* p = reinterpret_cast<uint8_t*>(a.hplanes) + ((m >> 7) * (2 * a.ndir) + 2 * dir + (sl >> 1)) * (2 * kPlaneBlock) +
                         sw64(r128, 2 * (sl & 1) + hf);
            *reinterpret_cast<uint4*>(p) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<uint4*>(p + a.hp_lo) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          } else {
            *reinterpret_cast<float4*>(hdst + sl * 16) = make_float4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<float4*>(hdst + sl * 16 + 4) = make_float4(h[4], h[5], h[6], h[7]);
          }
        }
        if (more) {
          const uint32_t col = COL_AH + sl * 8 + hf * 4;     // unit 16 sl + 8 hf + 2 i -> 32-bit column unit / 2
          tmem_st_32x4(tlane + col, hi);
          tmem_st_32x4(tlane + col + 32, lo);
        }
      }
      if (more) {
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(smem_u32(&bar_ready));
      }
    }
    }
  } else {
    // ================= MMA issuer =================
    const uint32_t idesc_half = make_idesc_bf16(TILE, 128);
    const uint32_t idesc_full = make_idesc_bf16(TILE, kGates);
    const uint32_t d_tmem = tmem_base + COL_ACC;
    constexpr int NCX = C::NCH - 1;                          // K chunks of the x part
    mbar_wait(smem_u32(&bar_w), 0);                          // the weight image has landed (once per CTA)
    uint32_t g = 0;
    for (int tile = cta_slot; tile < ntiles; tile += cta_slots)
    for (int it = 0; it < a.nsteps; ++it, ++g) {
      const uint32_t par = g & 1;
      // x part of this step, in two halves of 128 accumulator columns (32 units) as the epilogue of the previous step
      // releases them (slot barriers 0,1 then 2,3). N = 128 per MMA: a tcgen05.mma has a floor of ~82 cycles
      // whatever its N (tools/mma_sw_bench.cu: N = 64 -> 85 cycles, N = 128 -> 90, N = 256 -> 134), so the former
      // slot-wise N = 64 issue cost 8.2k cycles per step for KIN = 128 -- more than the epilogue it hides under.
      for (int hv = 0; hv < 2; ++hv) {
        mbar_wait(smem_u32(&bar_slot[2 * hv]), par);
        mbar_wait(smem_u32(&bar_slot[2 * hv + 1]), par);
        tc_fence_after();
        if (lane == 0) {
          uint32_t accumulate = 0;
#pragma unroll
          for (int ch = 0; ch < NCX; ++ch) {
            const uint32_t a_hi0 = tmem_base + COL_AX + ch * 32;
            const uint32_t a_lo0 = a_hi0 + KIN / 2;
            // rows 128 hv .. 128 hv + 127 of the weight chunk: 16 swizzle atoms of 8 rows x 128 B further
            const uint64_t w_hi = make_sw128_desc(smem_base + ch * CHUNK_BYTES + hv * 128 * 128);
            const uint64_t w_lo = make_sw128_desc(smem_base + C::PLANE + ch * CHUNK_BYTES + hv * 128 * 128);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t adv = (uint64_t)(kk * 32 >> 4);   // 16 bf16 = 32 bytes along K
              mma_bf16_ts(d_tmem + hv * 128, a_lo0 + kk * 8, w_hi + adv, idesc_half, accumulate);   // small terms first
              mma_bf16_ts(d_tmem + hv * 128, a_hi0 + kk * 8, w_lo + adv, idesc_half, 1);
              mma_bf16_ts(d_tmem + hv * 128, a_hi0 + kk * 8, w_hi + adv, idesc_half, 1);
              accumulate = 1;
            }
          }
        }
        __syncwarp();
      }
      // h part: needs every unit of h_{t-1}
      mbar_wait(smem_u32(&bar_ready), par);
      tc_fence_after();
      if (lane == 0) {
        if (it > 0) {                                        // h_{-1} = 0
          const uint32_t a_hi0 = tmem_base + COL_AH;
          const uint32_t a_lo0 = a_hi0 + 32;
          const uint64_t w_hi = make_sw128_desc(smem_base + NCX * CHUNK_BYTES);
          const uint64_t w_lo = make_sw128_desc(smem_base + C::PLANE + NCX * CHUNK_BYTES);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            const uint64_t adv = (uint64_t)(kk * 32 >> 4);
            mma_bf16_ts(d_tmem, a_lo0 + kk * 8, w_hi + adv, idesc_full, 1);
            mma_bf16_ts(d_tmem, a_hi0 + kk * 8, w_lo + adv, idesc_full, 1);
            mma_bf16_ts(d_tmem, a_hi0 + kk * 8, w_hi + adv, idesc_full, 1);
          }
        }
        mma_commit(smem_u32(&bar_full));
      }
      __syncwarp();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == N_EPI_WARPS) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

template <int KIN>
cudaError_t launch_kin(const LstmTcArgs& a, cudaStream_t s) {
  static SmemOptIn opt;
  cudaError_t e = opt_in_smem(lstm_tc_kernel<KIN>, Cfg<KIN>::SMEM, opt);
  if (e != cudaSuccess) return e;
  if (a.ndir < 1 || a.ndir > 2) return cudaErrorInvalidValue;
  // persistent: one CTA per SM, an equal number per direction, no more than there are tiles
  int dev = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  const int ntiles = (a.nseq + TILE - 1) / TILE;
  int per_dir = sm_count(dev) / a.ndir;
  if (per_dir > ntiles) per_dir = ntiles;
  lstm_tc_kernel<KIN><<<per_dir * a.ndir, THREADS, Cfg<KIN>::SMEM, s>>>(a);
  count_launch();
  return cudaGetLastError();
}

}  // namespace

cudaError_t launch_lstm_tc(const LstmTcArgs& a, cudaStream_t s) {
  if (a.nseq <= 0 || a.nsteps <= 0) return cudaSuccess;
  if (a.kin == 64) return launch_kin<64>(a, s);
  if (a.kin == 128) return launch_kin<128>(a, s);
  return cudaErrorInvalidValue;
}

}  // namespace bsb
