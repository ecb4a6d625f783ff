// Grouped Linear layers on the 5th-gen tensor cores (tcgen05 + TMEM), split-bf16 arithmetic.
//
//   O[m][n] = epi( sum_k A[m][k] * W[n][k] + bias[n] ),  A, O fp32 in HBM.
//
// Precision: the gate is 1e-3 against the fp32 reference, which plain TF32/bf16 misses
// (measured by emulation: TF32 1.7e-3). Each fp32 operand is split into two bf16 terms
// (x = hi + lo, 16 mantissa bits) and the product is formed from three MMAs
//   A_hi.W_hi + A_hi.W_lo + A_lo.W_hi      (fp32 accumulation in TMEM)
// which reproduces the fp32 path to ~1e-5 end to end at 3 bf16 MMAs per product.
//
// Persistent, warp-specialised CTA (one per SM), tiles of 128 rows x up to 256 columns:
//   8 warps    A producers : fp32 global (512-byte runs per warp request) -> registers, one K chunk
//                            ahead of the smem stage -> (hi, lo) bf16 (integer rounding on the ALU pipe,
//                            ptx.cuh split2) -> 128B-swizzled smem
//   1 warp     W loader    : the weights are packed as ready-made swizzled [bn x 64] bf16 blocks
//                            (pack.cu), so one thread fetches a hi and a lo block per K chunk with
//                            cp.async.bulk (TMA engine, completes on the stage's mbarrier)
//   1 warp     MMA issuer  : one thread issues tcgen05.mma (M=128, N<=256, K=16) x 12 per 64-wide
//                            K chunk; tcgen05.commit frees the smem stage / publishes the tile
//   4 warps    epilogue    : tcgen05.ld TMEM -> registers -> per-warp smem transpose, then bias +
//                            LeakyReLU (+residual, *x) on rows of 32 consecutive columns so that every
//                            global load/store of the epilogue is a full 128-byte line; the residual /
//                            mask operands of a chunk are all requested before the TMEM read
// Two smem stages (mbarrier full/empty ring) and two 256-column TMEM accumulators (full/empty
// barriers) let producer, tensor core and epilogue overlap across K chunks and across tiles.
//
// What bounds it (round-1 measurements, tools/mma_bench*.cu, tools/trace_tc.py, profiles/):
// tcgen05.mma itself runs at 128-142 cycles per M128xN256xK16 instruction in this operand pattern,
// also under concurrent LDG/STS/bulk-copy traffic, i.e. 1.7k cycles per chunk; the pipeline
// delivers a chunk every ~5k cycles. Each chunk makes the SM pull 32 KB of A and 64 KB of W
// through L2 (the weights are re-streamed for every 128-row tile); deeper A prefetch, fully
// coalesced loads, cheaper conversion and SMSP placement of the roles all left that period
// unchanged, a TMEM-resident A operand (lstm_tc.cu style) lost more to row-per-thread loads than
// it saved. The remaining lever is the CTA pair (cta_group::2: half of W per SM) - next round.
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int BM = 128;              // rows per tile (UMMA M)
constexpr int BK = 64;               // bf16 elements per K chunk = one 128-byte swizzle row
constexpr int BN_MAX = 256;          // columns per tile (UMMA N <= 256)
constexpr int STAGES = 2;
constexpr int A_PLANE = BM * 128;            // bytes of one A plane per stage (16 KB)
constexpr int W_PLANE = BN_MAX * 128;        // bytes of one W plane per stage (32 KB)
constexpr int STAGE_BYTES = 2 * A_PLANE + 2 * W_PLANE;   // 96 KB
constexpr int STG_LD = 36;                              // staging row stride (floats): conflict-free 16B access
constexpr int STG_WARP = 32 * STG_LD * 4;               // bytes of one epilogue warp's 32x32 staging tile
// Warp roles are placed by scheduler (SMSP = warp % 4). The thread that issues tcgen05.mma occupies its
// SMSP's dispatch port for the whole MMA batch (pipeline trace: warps sharing that SMSP made no
// progress while a batch was being issued), so no producer shares the MMA warp's SMSP:
//   warp 0            MMA issuer (SMSP 0)          warps 4, 8      idle
//   warps 1-3,5-7,9,10  A producers (SMSPs 1-3)    warp 11         W loader
//   warps 12-15       epilogue (TMEM lane quarter = warp % 4)
constexpr int N_A_WARPS = 8, N_EPI_WARPS = 4;
constexpr int W_WARP = 11, EPI_WARP0 = 12;
__device__ __forceinline__ int producer_rank(int warp) {      // 1,2,3,5,6,7,9,10 -> 0..7 ; others -1
  if ((warp & 3) == 0 || warp > 10) return -1;
  return (warp >> 2) * 3 + (warp & 3) - 1;
}
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + N_EPI_WARPS * STG_WARP + 1024;  // + alignment slack
constexpr int MMA_WARP = 0;
constexpr int THREADS = 16 * 32;
constexpr int N_PRODUCERS = N_A_WARPS * 32;                   // 256
constexpr int TMEM_COLS = 512;       // 2 accumulators x 256 fp32 columns

struct TcProb {
  const float* A;
  const __nv_bfloat16* Wh;   // [Npad][ldw] bf16, hi plane
  const __nv_bfloat16* Wl;   // lo plane
  const float* bias;
  float* O;
  const float* res;
  const float* mul;
  int lda, ldw, ldo, ldres, ldmul;
  int M, N, K, act;
  int tile_start, tiles_m;
};
struct TcArgs {
  TcProb p[kMaxProb];
  int nprob;
  int total_tiles;
};

struct Tile {
  int pi, m0, n0, bn;    // problem, first row, first column, columns in this tile (multiple of 64)
};

__device__ __forceinline__ Tile decode_tile(const TcArgs& args, int tile) {
  int pi = 0;
  while (pi + 1 < args.nprob && tile >= args.p[pi + 1].tile_start) ++pi;
  const TcProb& P = args.p[pi];
  const int local = tile - P.tile_start;
  Tile t;
  t.pi = pi;
  const int tiles_n = (P.N + BN_MAX - 1) / BN_MAX;
  t.m0 = (local / tiles_n) * BM;              // n fastest: the CTAs of a wave read different W blocks
  t.n0 = (local % tiles_n) * BN_MAX;
  const int npad = (P.N + 63) & ~63;
  t.bn = min(BN_MAX, npad - t.n0);
  return t;
}

// byte offset of 16-byte chunk `c` (0..7) of row `r` inside a 128B-swizzled K-major plane
__device__ __forceinline__ uint32_t sw128(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }

// ---- A producer: 128 rows x 64 k of fp32 -> hi/lo bf16 planes --------------------------------
// 256 threads; thread t owns the float4 units u = t + 256 i (row u/16, floats 4(u%16)..+3), i < 8, so
// every warp-level load covers 512 consecutive bytes (2 rows x 256 B): full sectors per request.
struct AChunk { float4 v[8]; };

__device__ __forceinline__ void load_a(const TcProb& P, int m0, int k0, int t, AChunk& a) {
  const bool vec = ((P.lda & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.A) & 15) == 0);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int u = t + 256 * i;
    const int r = u >> 4, c = u & 15;
    const int m = m0 + r, k = k0 + c * 4;
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m < P.M && k < P.K) {
      const float* p = P.A + (long long)m * P.lda + k;
      if (vec && k + 3 < P.K) {
        x = __ldg(reinterpret_cast<const float4*>(p));
      } else {
        x.x = __ldg(p);
        if (k + 1 < P.K) x.y = __ldg(p + 1);
        if (k + 2 < P.K) x.z = __ldg(p + 2);
        if (k + 3 < P.K) x.w = __ldg(p + 3);
      }
    }
    a.v[i] = x;
  }
}

__device__ __forceinline__ void store_a(const AChunk& a, uint8_t* a_hi, uint8_t* a_lo, int t) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int u = t + 256 * i;
    const int r = u >> 4, c = u & 15;
    uint2 h, l;
    split2(a.v[i].x, a.v[i].y, h.x, l.x);
    split2(a.v[i].z, a.v[i].w, h.y, l.y);
    const uint32_t off = sw128(r, c >> 1) + (c & 1) * 8;     // half of a 16-byte swizzle chunk
    *reinterpret_cast<uint2*>(a_hi + off) = h;
    *reinterpret_cast<uint2*>(a_lo + off) = l;
  }
}

__device__ __forceinline__ float lrelu(float v) { return v > 0.f ? v : v * 0.01f; }

#ifdef BSB_TRACE
// pipeline timeline of CTA 0 (debug builds only): three role regions of 3000 (code, clock) pairs,
// a private counter per role (no atomics: tracing must not stretch the pipeline it observes)
__device__ unsigned long long* g_trace = nullptr;
#define TRACE_DECL int trace_n = 0
#define TRACE(role, code)                                                            \
  do {                                                                               \
    if (blockIdx.x == 0 && g_trace != nullptr && trace_n < 3000) {                   \
      g_trace[(role) * 6000 + 2 * trace_n] = (unsigned long long)(code);             \
      g_trace[(role) * 6000 + 2 * trace_n + 1] = (unsigned long long)clock64();      \
      ++trace_n;                                                                     \
    }                                                                                \
  } while (0)
#else
#define TRACE_DECL
#define TRACE(role, code)
#endif

__global__ void __launch_bounds__(THREADS, 1)
tc_grouped_linear_kernel(const __grid_constant__ TcArgs args) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[STAGES], bar_empty[STAGES], bar_tfull[2], bar_tempty[2];
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  // 1024-byte aligned operand area (SWIZZLE_128B atoms)
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), N_PRODUCERS + 1);   // + the W issuer's expect_tx arrive
      mbar_init(smem_u32(&bar_empty[s]), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&bar_tfull[a]), 1);
      mbar_init(smem_u32(&bar_tempty[a]), N_EPI_WARPS * 32);
    }
    fence_mbar_init();
  }
  if (warp == MMA_WARP) {
    tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  TRACE_DECL;

  const int prank = producer_rank(warp);
  if (prank >= 0) {
    // ================= A producers =================
    const int pt = prank * 32 + lane;           // producer thread 0..255
    // One K chunk of A is prefetched into registers while the previous one is converted and stored.
    // (Two chunks in flight were measured and did not help: at this tile shape the kernel is bound by
    // the bytes each SM must pull from L2 - 32 KB of A plus 64 KB of W per 128x256x64 chunk - not by
    // the latency of any one of them; see DESIGN.md.)
    int stage = 0;
    uint32_t phase = 0;
    int tile = blockIdx.x, kc = 0, nk = 0;
    Tile T{};
    AChunk regs;
    if (tile < args.total_tiles) {
      T = decode_tile(args, tile);
      nk = (args.p[T.pi].K + BK - 1) / BK;
      load_a(args.p[T.pi], T.m0, 0, pt, regs);
    }
    while (tile < args.total_tiles) {
      mbar_wait(smem_u32(&bar_empty[stage]), phase ^ 1);
      if (pt == 0) TRACE(0, 1);
      uint8_t* sb = smem_gen + stage * STAGE_BYTES;
      store_a(regs, sb, sb + A_PLANE, pt);
      if (pt == 0) TRACE(0, 2);
      fence_proxy_async_smem();
      mbar_arrive(smem_u32(&bar_full[stage]));
      if (pt == 0) TRACE(0, 3);
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
      if (++kc == nk) {
        kc = 0;
        tile += gridDim.x;
        if (tile < args.total_tiles) {
          T = decode_tile(args, tile);
          nk = (args.p[T.pi].K + BK - 1) / BK;
        }
      }
      if (tile < args.total_tiles) load_a(args.p[T.pi], T.m0, kc * BK, pt, regs);
      if (pt == 0) TRACE(0, 4);
    }
  } else if (warp == W_WARP) {
    // ================= W loader (one thread) =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < args.total_tiles; tile += gridDim.x) {
        const Tile T = decode_tile(args, tile);
        const TcProb& P = args.p[T.pi];
        const int nk = (P.K + BK - 1) / BK;
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(smem_u32(&bar_empty[stage]), phase ^ 1);
          // W blocks [bn x 64] bf16, already swizzled: (n0 * nch + kc * bn) * 128 bytes into each plane
          const uint32_t bytes = (uint32_t)T.bn * 128u;
          const long long off = ((long long)T.n0 * (P.ldw >> 6) + (long long)kc * T.bn) * 128;
          const uint32_t bar = smem_u32(&bar_full[stage]);
          mbar_arrive_expect_tx(bar, 2 * bytes);
          bulk_g2s(smem_base + stage * STAGE_BYTES + 2 * A_PLANE, reinterpret_cast<const uint8_t*>(P.Wh) + off, bytes, bar);
          bulk_g2s(smem_base + stage * STAGE_BYTES + 2 * A_PLANE + W_PLANE, reinterpret_cast<const uint8_t*>(P.Wl) + off, bytes, bar);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ================= MMA issuer =================
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < args.total_tiles; tile += gridDim.x) {
      const Tile T = decode_tile(args, tile);
      const TcProb& P = args.p[T.pi];
      const int nk = (P.K + BK - 1) / BK;
      mbar_wait(smem_u32(&bar_tempty[acc]), acc_phase ^ 1);   // epilogue has drained this accumulator
      if (lane == 0) TRACE(1, 10);
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16(BM, T.bn);
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN_MAX);
      for (int kc = 0; kc < nk; ++kc) {
        mbar_wait(smem_u32(&bar_full[stage]), phase);
        if (lane == 0) TRACE(1, 11);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sb = smem_base + stage * STAGE_BYTES;
          const uint64_t a_hi = make_sw128_desc(sb);
          const uint64_t a_lo = make_sw128_desc(sb + A_PLANE);
          const uint64_t w_hi = make_sw128_desc(sb + 2 * A_PLANE);
          const uint64_t w_lo = make_sw128_desc(sb + 2 * A_PLANE + W_PLANE);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t adv = (uint64_t)(kk * 32 >> 4);    // 16 bf16 = 32 bytes along K
            mma_bf16_ss(d_tmem, a_lo + adv, w_hi + adv, idesc, (kc | kk) != 0);   // small terms first
            mma_bf16_ss(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
            mma_bf16_ss(d_tmem, a_hi + adv, w_hi + adv, idesc, 1);
          }
          mma_commit(smem_u32(&bar_empty[stage]));             // smem stage reusable once these MMAs retire
          if (kc == nk - 1) mma_commit(smem_u32(&bar_tfull[acc]));
          TRACE(1, 12);
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= EPI_WARP0) {
    // ================= epilogue =================
    const int q = warp & 3;                 // TMEM lane quarter this warp may access
    float* stg = reinterpret_cast<float*>(smem_gen + STAGES * STAGE_BYTES + (warp - EPI_WARP0) * STG_WARP);
    const int cg = lane & 7, sub = lane >> 3;   // transposed phase: 4 columns at 4*cg of rows sub + 4k
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < args.total_tiles; tile += gridDim.x) {
      const Tile T = decode_tile(args, tile);
      const TcProb& P = args.p[T.pi];
      mbar_wait(smem_u32(&bar_tfull[acc]), acc_phase);
      if (warp == EPI_WARP0 && lane == 0) TRACE(2, 20);
      tc_fence_after();
      const int mw = T.m0 + q * 32;           // first row of this warp
      const bool o_vec = ((P.ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.O) & 15) == 0);
      const bool r_vec = P.res && ((P.ldres & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.res) & 15) == 0);
      const bool m_vec = P.mul && ((P.ldmul & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.mul) & 15) == 0);
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN_MAX);
      for (int cc = 0; cc < T.bn / 32; ++cc) {
        const int nb = T.n0 + cc * 32;
        if (nb >= P.N) break;
        // operands of the fused "+ residual" / "* x": all loads of the chunk are issued up front
        const int n = nb + 4 * cg;
        const bool ncol = n < P.N;
        const bool full4 = n + 3 < P.N;
        float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 rr[8], mm[8];
        if (ncol) bb = __ldg(reinterpret_cast<const float4*>(P.bias + n));   // bias padded to Npad
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int m = mw + sub + 4 * k;
          rr[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          mm[k] = make_float4(1.f, 1.f, 1.f, 1.f);
          if (ncol && m < P.M) {
            if (P.res) {
              const float* rrow = P.res + (long long)m * P.ldres + n;
              if (r_vec && full4) {
                rr[k] = __ldg(reinterpret_cast<const float4*>(rrow));
              } else {
                rr[k].x = __ldg(rrow);
                if (n + 1 < P.N) rr[k].y = __ldg(rrow + 1);
                if (n + 2 < P.N) rr[k].z = __ldg(rrow + 2);
                if (n + 3 < P.N) rr[k].w = __ldg(rrow + 3);
              }
            }
            if (P.mul) {
              const float* mrow = P.mul + (long long)m * P.ldmul + n;
              if (m_vec && full4) {
                mm[k] = __ldg(reinterpret_cast<const float4*>(mrow));
              } else {
                mm[k].x = __ldg(mrow);
                if (n + 1 < P.N) mm[k].y = __ldg(mrow + 1);
                if (n + 2 < P.N) mm[k].z = __ldg(mrow + 2);
                if (n + 3 < P.N) mm[k].w = __ldg(mrow + 3);
              }
            }
          }
        }
        uint32_t r[32];
        tmem_ld_32x32(tbase + cc * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<uint4*>(stg + lane * STG_LD + 4 * j) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        __syncwarp();
        if (ncol) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int i = sub + 4 * k;
            const int m = mw + i;
            const float4 a4 = *reinterpret_cast<const float4*>(stg + i * STG_LD + 4 * cg);
            float v[4] = {a4.x + bb.x, a4.y + bb.y, a4.z + bb.z, a4.w + bb.w};
            if (P.act == ACT_LRELU) {
#pragma unroll
              for (int e = 0; e < 4; ++e) v[e] = lrelu(v[e]);
            }
            v[0] = (v[0] + rr[k].x) * mm[k].x;
            v[1] = (v[1] + rr[k].y) * mm[k].y;
            v[2] = (v[2] + rr[k].z) * mm[k].z;
            v[3] = (v[3] + rr[k].w) * mm[k].w;
            if (m < P.M) {
              float* orow = P.O + (long long)m * P.ldo + n;
              if (o_vec && full4) {
                *reinterpret_cast<float4*>(orow) = make_float4(v[0], v[1], v[2], v[3]);
              } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) if (n + e < P.N) orow[e] = v[e];
              }
            }
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      mbar_arrive(smem_u32(&bar_tempty[acc]));
      if (warp == EPI_WARP0 && lane == 0) TRACE(2, 21);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

int g_sm_count[64] = {};

}  // namespace

// Launch the problems of `args` that have K > 0 on the tensor-core kernel. Problems are taken
// in descending-K order so that the static round-robin schedule ends with the cheap tiles.
cudaError_t launch_tc_grouped_linear(const GroupedArgs& args, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc_grouped_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < 64 && g_sm_count[dev] == 0) {
    e = cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
  }
  const int sms = (dev >= 0 && dev < 64) ? g_sm_count[dev] : 148;

  int order[kMaxProb], n = 0;
  for (int i = 0; i < args.nprob; ++i)
    if (args.p[i].K > 0 && args.p[i].M > 0 && args.p[i].N > 0) order[n++] = i;
  for (int i = 1; i < n; ++i)               // insertion sort by K descending
    for (int j = i; j > 0 && args.p[order[j]].K > args.p[order[j - 1]].K; --j) {
      const int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t;
    }
  TcArgs ta{};
  int total = 0;
  for (int i = 0; i < n; ++i) {
    const LinProb& L = args.p[order[i]];
    TcProb& P = ta.p[i];
    P.A = L.A; P.Wh = reinterpret_cast<const __nv_bfloat16*>(L.Wh); P.Wl = reinterpret_cast<const __nv_bfloat16*>(L.Wl);
    P.bias = L.bias; P.O = L.O; P.res = L.res; P.mul = L.mul;
    P.lda = L.lda; P.ldw = L.ldw64; P.ldo = L.ldo; P.ldres = L.ldres; P.ldmul = L.ldmul;
    P.M = L.M; P.N = L.N; P.K = L.K; P.act = L.act;
    P.tile_start = total;
    P.tiles_m = (L.M + BM - 1) / BM;
    total += P.tiles_m * ((L.N + BN_MAX - 1) / BN_MAX);
  }
  ta.nprob = n;
  ta.total_tiles = total;
  if (total == 0) return cudaSuccess;
  const int grid = total < sms ? total : sms;
  tc_grouped_linear_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(ta);
  count_launch();
  return cudaGetLastError();
}

#ifdef BSB_TRACE
extern "C" int bsb_debug_set_trace(void* p) {
  return (int)cudaMemcpyToSymbol(g_trace, &p, sizeof(p));
}
#endif

}  // namespace bsb
