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
// Persistent, warp-specialised CTA (one per SM), tiles of 128 rows x up to 256 columns, K chunks of 32 (one 64-byte
// SWIZZLE_64B row = two k-steps of 16 = six MMAs), FOUR shared-memory stages of 48 KB (A hi|lo 16 KB + W hi 16 KB +
// W lo 16 KB), two 256-column fp32 accumulators in tensor memory:
//   A from planes (12 of the 14 launch kinds, 384 threads): activations between Linears live in HBM as bf16 hi/lo
//     "planes" in exactly the shared-memory image (kernels.cuh LinProb::Ap), so a stage is three bulk copies
//     (cp.async.bulk) issued by three loader warps; warp 0 issues the MMAs (elect.sync, tcgen05.commit frees the
//     stage); warps 4-11 are the epilogue (two per TMEM lane quarter, alternate 32-column steps).
//   A from fp32 rows (after the STFT and after the last LSTM block, 768 threads): two groups of 8 producer warps load
//     rows (16-byte __ldg), split them to bf16 hi/lo on the ALU pipe and store the swizzled stage; 4 epilogue warps
//     with registers moved to them by setmaxnreg.
//   Epilogue: tcgen05.ld gives a lane one accumulator row, 32 columns per step: bias, LeakyReLU, "+ residual",
//     "* multiplier" (operands requested one step ahead), then fp32 rows and/or planes. Plane output of launches
//     without residual/multiplier operands goes through a 4 KB staging buffer per warp and leaves as two 2 KB
//     cp.async.bulk shared -> global copies issued by one lane.
//
// What bounds it (round-2 measurements; tools/mma_sw_bench.cu, tools/exp_cost.py, tools/band_cost.py,
// profiles/r02_summary.md): a tcgen05.mma M128 x N256 x K16 takes 134 cycles with SWIZZLE_64B as with SWIZZLE_128B
// operands (807 per chunk), but has a floor of ~82 cycles whatever N (N = 64: 85, N = 128: 90), so narrow layers pay
// 2.6x per MAC. With the epilogue switched off the main loop runs close to that bound; with it, a d -> d layer takes
// 187 us against ~110 us of MMA time. The cost of the epilogue is proportional to the work it does (split arithmetic
// and stores ~45 us each of a 410 us pair of launches) and does NOT respond to how it is scheduled: coalescing its
// stores through shared memory, stores into an L2-resident window, pipelining its TMEM loads, bias from registers,
// sixteen epilogue warps instead of eight, a boustrophedon tile order and the CTA-pair kernel were all measured at
// +-1 % or worse; every variant that ADDS shared-memory traffic is slower (+5 .. +70 % on the launch). The shared
// memory port carries the MMA operand reads (12 KB per MMA = 92 B/clk at full rate) and the bulk-copy writes
// (48 KB per chunk = 60 B/clk): that sum, not the tensor pipe and not HBM, is what the kernel sits on. The step is
// also power-capped when sustained (sw_power_cap active, 1.6 GHz inside these launches under ncu).
#include <cstdlib>
#include <cuda.h>              // CUtensorMap types only: the encoder is fetched with cudaGetDriverEntryPoint (no libcuda link)
#include "kernels.cuh"
#include "ptx.cuh"

namespace bsb {
namespace {

using namespace ptx;

constexpr int BM = 128;              // rows per tile (UMMA M)
constexpr int BK = kPlaneK;          // bf16 elements per K chunk = one 64-byte swizzle row (two k16 MMA steps)
constexpr int BN_MAX = 256;          // columns per tile (UMMA N <= 256)
constexpr int A_PLANE = BM * BK * 2;         // bytes of one A plane per stage (8 KB) = kPlaneBlock
// PAIR: two CTAs of a cluster (one TPC) run one tcgen05.mma.cta_group::2 tile of 256 rows: each CTA holds its own
// 128 rows of A and HALF of the W rows, so a K chunk costs an SM 32 KB instead of 48 KB and six stages fit.
template <bool PAIR> struct Smem {
  static constexpr int W_PLANE = (PAIR ? BN_MAX / 2 : BN_MAX) * BK * 2;   // bytes of one W plane per stage (8 / 16 KB)
  static constexpr int STAGE_BYTES = 2 * A_PLANE + 2 * W_PLANE;           // 32 / 48 KB
  static constexpr int STAGES = PAIR ? 6 : 4;
};
constexpr int MAX_STAGES = 6;
// Warp roles (setmaxnreg works per warpgroup = 4 consecutive warps):
//   fp32-row A (A_PLANES = false), 24 warps at 80 registers:
//     warps 0-7   A producer group 0 (fills stage 0)      warps 8-15  A producer group 1 (fills stage 1)
//     warp 16     MMA issuer (+ TMEM alloc)               warps 17, 18  W loaders (hi, lo)   warp 19 idle
//     warps 20-23 epilogue, grown to 152 registers
//   plane A (A_PLANES = true), 12 warps, no register re-partitioning needed:
//     warp 0 MMA issuer, warps 1-3 loaders (A hi|lo, W hi, W lo: one bulk copy each per stage), warps 4-11 epilogue
// A thread's cp.async.bulk requests are served one at a time, ~540 cycles each whatever their size up to 32 KB
// (tools/mem_bench.cu), requests of different warps concurrently: one loader warp per piece of a stage.
// An epilogue warp may read the TMEM lane quarter warp % 4; with 8 epilogue warps the two warps of a quarter
// take alternate 32-column chunks.
// Two producer groups, each with its own register-resident chunk in flight: the global-load latency of one
// group's chunk overlaps the convert+store of the other's, and 16 warps of LDG.128 sustain ~2x the bytes per
// clock of 8 (tools/mem_bench.cu: 12.6 -> 22.3 B/clk/SM from DRAM, 17.3 -> 31.6 from L2).
constexpr int N_GROUP_WARPS = 8;
template <bool A_PLANES> struct Roles {
  static constexpr int PROD_WARPS = A_PLANES ? 0 : 2 * N_GROUP_WARPS;
  static constexpr int MMA_WARP = PROD_WARPS, LOAD_WARP = PROD_WARPS + 1;
  static constexpr int EPI_WARP0 = PROD_WARPS + 4;
  static constexpr int N_EPI_WARPS = A_PLANES ? 8 : 4;
  static constexpr int THREADS = (EPI_WARP0 + N_EPI_WARPS) * 32;
};
template <bool PAIR> constexpr int smem_bytes() { return Smem<PAIR>::STAGES * Smem<PAIR>::STAGE_BYTES + 1024; }   // + alignment slack
// Launches whose epilogue has no "+ residual" / "* multiplier" operand get, per epilogue warp, a 4 KB staging buffer
// behind the stages: plane output then leaves as cp.async.bulk shared -> global copies issued by one lane (measured:
// back -8 %, bandfc -6 %, pre -2 %). The two launches WITH such operands keep the smaller footprint: their operand
// loads depend on L1, which the extra 32 KB of shared memory would halve (measured +14 % on them).
constexpr int EPI_STAGE_BYTES = 4096;
constexpr int EPI_STAGE_TOTAL = 8 * EPI_STAGE_BYTES;
constexpr int REGS_PRODUCER = 64, REGS_CONTROL = 40, REGS_EPILOGUE = 184;   // released 512*16 + 128*40 = taken 128*104
static_assert(Smem<false>::STAGES == 4 && A_PLANE == kPlaneBlock, "two producer groups alternate over four stages");
constexpr int TMEM_COLS = 512;       // 2 accumulators x 256 fp32 columns

struct TcProb {
  const float* A;
  const __nv_bfloat16* Wh;   // [Npad][ldw] bf16, hi plane
  const __nv_bfloat16* Wl;   // lo plane
  const float* bias;
  float* O;
  const float* res;
  const float* mul;
  const uint8_t* Ah;         // A as blocked bf16 planes (kernels.cuh LinProb::Ap), or null: fp32 rows at A
  const uint8_t* Al;
  uint8_t* Oh;               // output planes, or null
  uint8_t* Ol;
  const uint8_t* Rh;         // residual operand as planes (geometry of the output), or null: fp32 rows at res
  const uint8_t* Rl;
  int lda, ldw, ldo, ldres, ldmul;
  int M, N, K, act;
  int tile_start, tiles_m;
  int o_tma;                 // maps.o[problem] is valid: fp32 output rows leave through staging + 2-D TMA stores
  int ops_tma;               // the residual / multiplier operands of this problem arrive through staging (TMA loads)
};
struct TcArgs {
  TcProb p[kMaxProb];
  int nprob;
  int total_tiles;
  int stage_out;      // plane output through the per-warp staging buffers + bulk stores (the launch reserved them)
  int trace;          // debug builds: this launch records the pipeline timeline
  int exp;            // timing-experiment switches (g_exp), 0 in production
};

// Tensor maps of a launch (second kernel parameter): per problem the fp32 output matrix as a 2-D tensor
// [M rows][N columns], row pitch ldo, boxes of 32 rows x 32 columns (= 128 B: one SWIZZLE_128B span).
struct TcMaps {
  alignas(64) CUtensorMap o[kMaxProb];
  alignas(64) CUtensorMap r[kMaxProb];     // "+ residual" operand as fp32 rows
  alignas(64) CUtensorMap m[kMaxProb];     // "* multiplier" operand (fp32 rows)
};

struct Tile {
  int pi, m0, n0, bn;    // problem, first row, first column, columns in this tile (multiple of 64)
};

// tile_rows = 128, rank 0 for single CTAs; 256 and the CTA's rank in the pair for cta_group::2
// `pi` carries the problem index from one tile of a role to its next (tiles of a CTA ascend), so the walk over
// the problem table is amortised: from scratch it costs ~10 dependent constant-bank loads for the widest bands.
__device__ __forceinline__ Tile decode_tile(const TcArgs& args, int tile, int& pi, int tile_rows = BM, int rank = 0) {
  while (pi + 1 < args.nprob && tile >= args.p[pi + 1].tile_start) ++pi;
  const TcProb& P = args.p[pi];
  const int local = tile - P.tile_start;
  Tile t;
  t.pi = pi;
  const int tiles_n = (P.N + BN_MAX - 1) / BN_MAX;
  t.m0 = (local / tiles_n) * tile_rows + rank * BM;   // n fastest: the CTAs of a wave read different W blocks
  t.n0 = (local % tiles_n) * BN_MAX;
  const int npad = (P.N + 63) & ~63;
  t.bn = min(BN_MAX, npad - t.n0);
  return t;
}

// K-major SWIZZLE_64B operand descriptor: rows of 64 B, 8-row groups 512 B apart (kernels.cuh sw64)
__device__ __forceinline__ uint64_t make_sw64_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(512 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)4 << 61;
  return d;
}

// ---- A producer: 128 rows x 32 k of fp32 -> hi/lo bf16 planes --------------------------------
// 256 threads; thread t owns the float4 units u = t + 256 i (row u/8, floats 4(u%8)..+3), i < 4, so
// every warp-level load covers 4 rows x 128 consecutive bytes.
struct AChunk { float4 v[4]; };

__device__ __forceinline__ void load_a(const TcProb& P, int m0, int k0, int t, AChunk& a) {
  const bool vec = ((P.lda & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.A) & 15) == 0);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int u = t + 256 * i;
    const int r = u >> 3, c = u & 7;
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
  for (int i = 0; i < 4; ++i) {
    const int u = t + 256 * i;
    const int r = u >> 3, c = u & 7;
    uint2 h, l;
    split2(a.v[i].x, a.v[i].y, h.x, l.x);
    split2(a.v[i].z, a.v[i].w, h.y, l.y);
    const uint32_t off = sw64(r, c >> 1) + (c & 1) * 8;      // half of a 16-byte swizzle unit
    *reinterpret_cast<uint2*>(a_hi + off) = h;
    *reinterpret_cast<uint2*>(a_lo + off) = l;
  }
}

__device__ __forceinline__ float lrelu(float v) { return fmaxf(v, v * 0.01f); }   // slope < 1: same value as v > 0 ? v : 0.01 v, two instructions

#ifndef BSB_TRACE_EPI
#define BSB_TRACE_EPI 0     // which epilogue warp the trace follows
#endif
#ifdef BSB_TRACE
// pipeline timeline of CTA 0 (debug builds only): five role regions of 3000 (code, clock) pairs,
// a private counter per role (no atomics: tracing must not stretch the pipeline it observes)
__device__ unsigned long long* g_trace = nullptr;
#define TRACE_DECL int trace_n = 0
#define TRACE(role, code)                                                            \
  do {                                                                               \
    if (blockIdx.x <= 1 && args.trace && g_trace != nullptr && trace_n < 3000) {                   \
      g_trace[((role) + 5 * blockIdx.x) * 6000 + 2 * trace_n] = (unsigned long long)(code);             \
      g_trace[((role) + 5 * blockIdx.x) * 6000 + 2 * trace_n + 1] = (unsigned long long)clock64();      \
      ++trace_n;                                                                     \
    }                                                                                \
  } while (0)
#else
#define TRACE_DECL
#define TRACE(role, code)
#endif


// A_PLANES: every problem of the launch takes its A operand from bf16 planes (two bulk copies per chunk by the
// loader thread); the 16 producer warps then have no work and only the loader arrives on the full barrier.
// PAIR (with A_PLANES only): launched as clusters of two CTAs; CTA rank 0 issues the MMAs for both, every CTA loads
// its own operands, drains its own 128 accumulator rows and reports to rank 0:
//   bar_full[s]   own bulk copies of stage s landed        bar_peer[s] (rank 0)  rank 1's stage s landed (relayed by its warp 0)
//   bar_empty[s]  MMAs reading stage s retired (commit multicast to both CTAs)
//   bar_tfull[a]  accumulator a complete (commit multicast) bar_tempty[a] (rank 0) drained by all 16 epilogue warps
// RMS ("residual / multiplier, staged"): the instantiation for plane-fed launches whose epilogue has a "+ residual" or
// "* multiplier" operand (post.2, the RNN blocks' fc): three main-loop stages, 8 KB of staging per epilogue warp, the
// operands brought in by TMA. Kept out of the other instantiations on purpose: compiled into the common kernel (runtime
// stage count, the extra epilogue code at the 168-register cap) it cost the launches that do not use it 4-5 %.
template <bool A_PLANES, bool PAIR, bool RMS = false>
__global__ void __launch_bounds__(Roles<A_PLANES>::THREADS, 1)
tc_grouped_linear_kernel(const __grid_constant__ TcArgs args, const __grid_constant__ TcMaps maps) {
  static_assert(A_PLANES || !PAIR, "the CTA-pair kernel takes its A operand from planes");
  static_assert(!RMS || (A_PLANES && !PAIR), "staged operands: plane-fed single-CTA launches only");
  using R = Roles<A_PLANES>;
  constexpr int MMA_WARP = R::MMA_WARP, W_WARP = R::LOAD_WARP, EPI_WARP0 = R::EPI_WARP0, N_EPI_WARPS = R::N_EPI_WARPS;
  constexpr int STAGES = Smem<PAIR>::STAGES, STAGE_BYTES = Smem<PAIR>::STAGE_BYTES, W_PLANE = Smem<PAIR>::W_PLANE;
  constexpr int nst = RMS ? 3 : STAGES;                           // stages in use
  constexpr int STG_BYTES = RMS ? 2 * EPI_STAGE_BYTES : EPI_STAGE_BYTES, Y_OFF = RMS ? EPI_STAGE_BYTES : 0;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bar_full[MAX_STAGES], bar_empty[MAX_STAGES], bar_peer[MAX_STAGES], bar_tfull[2], bar_tempty[2];
  __shared__ __align__(8) uint64_t bar_ld[8];              // per epilogue warp: its staged operand loads have landed
  __shared__ uint32_t tmem_slot;
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int tile0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;          // first tile and stride of this CTA / pair
  const int tile_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr int TILE_ROWS = PAIR ? 2 * BM : BM;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  // 1024-byte aligned operand area (SWIZZLE_128B atoms)
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(smem_u32(&bar_full[s]), A_PLANES ? 3 : N_GROUP_WARPS + 2);   // the loaders' expect_tx arrives (+ one arrive per producer warp)
      mbar_init(smem_u32(&bar_empty[s]), 1);
      mbar_init(smem_u32(&bar_peer[s]), 1);
    }
    for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bar_ld[i]), 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(smem_u32(&bar_tfull[a]), 1);
      mbar_init(smem_u32(&bar_tempty[a]), PAIR ? 2 * N_EPI_WARPS : N_EPI_WARPS * 32);   // pair: one lane per epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == MMA_WARP) {
    if (PAIR) {
      tmem_alloc_pair(smem_u32(&tmem_slot), TMEM_COLS);
      tmem_relinquish_pair();
    } else {
      tmem_alloc(smem_u32(&tmem_slot), TMEM_COLS);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  TRACE_DECL;

  if (warp < R::PROD_WARPS) {
    // ================= A producers: group g fills stages 2g and 2g+1 = chunks 4j+2g, 4j+2g+1 of the CTA's stream ====
    // (two chunks = eight 16-byte loads per thread are requested before the first is converted)
    setmaxnreg_dec<REGS_PRODUCER>();
    const int grp = warp / N_GROUP_WARPS;
    const int pt = (warp % N_GROUP_WARPS) * 32 + lane;      // thread 0..255 of the group
    int tile = blockIdx.x, kc = 0, nk = 0, pi = 0;
    Tile T{};
    auto open_tile = [&]() {
      if (tile < args.total_tiles) {
        T = decode_tile(args, tile, pi);
        nk = (args.p[T.pi].K + BK - 1) / BK;
      }
    };
    auto next_chunk = [&]() {
      if (tile < args.total_tiles && ++kc == nk) { kc = 0; tile += gridDim.x; open_tile(); }
    };
    open_tile();
    if (grp == 1) { next_chunk(); next_chunk(); }
    uint32_t phase = 0;
    while (tile < args.total_tiles) {
      AChunk r0, r1;
      load_a(args.p[T.pi], T.m0, kc * BK, pt, r0);
      next_chunk();
      const bool second = tile < args.total_tiles;
      if (second) load_a(args.p[T.pi], T.m0, kc * BK, pt, r1);
      next_chunk();
      next_chunk();
      next_chunk();
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (h == 1 && !second) break;
        const int stage = 2 * grp + h;
        uint8_t* sb = smem_gen + stage * STAGE_BYTES;
        // one lane per warp talks to the barriers: 32x fewer polls and arrives on the CTA's barrier unit
        if (lane == 0) mbar_wait(smem_u32(&bar_empty[stage]), phase ^ 1);
        __syncwarp();
        if (pt == 0) TRACE(grp ? 3 : 0, 1 + 100 * grp);
        store_a(h == 0 ? r0 : r1, sb, sb + A_PLANE, pt);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bar_full[stage]));
        if (pt == 0) TRACE(grp ? 3 : 0, 3 + 100 * grp);
      }
      phase ^= 1;
    }
  } else if (warp < EPI_WARP0) {
    if (!A_PLANES) setmaxnreg_dec<REGS_CONTROL>();
    constexpr int N_LOADERS = A_PLANES ? 3 : 2;
    if (warp >= W_WARP && warp < W_WARP + N_LOADERS) {
      // ================= loaders: one warp (one thread) per piece of a stage =================
      //   A_PLANES: piece 0 = A block (hi|lo, 16 KB), 1 = W hi, 2 = W lo;   fp32 A: piece 1 = W hi, 2 = W lo
      const int piece = warp - W_WARP + (A_PLANES ? 0 : 1);
      if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int pi = 0;
      for (int tile = tile0; tile < args.total_tiles; tile += tile_step) {
        const Tile T = decode_tile(args, tile, pi, TILE_ROWS, rank);
        const TcProb& P = args.p[T.pi];
        const int nk = (P.K + BK - 1) / BK;
        for (int kc = 0; kc < nk; ++kc) {
          mbar_wait(smem_u32(&bar_empty[stage]), phase ^ 1);
          if (piece == 0) TRACE(4, 30 + stage);
          const uint32_t bar = smem_u32(&bar_full[stage]);
          const uint32_t sb = smem_base + stage * STAGE_BYTES;
          if (piece == 0) {
            // plane block (m-tile, chunk): hi then lo, 16 KB
            const long long aoff = ((long long)(T.m0 / BM) * nk + kc) * (2 * A_PLANE);
            if (args.exp & 4) { mbar_arrive(bar); } else {
            mbar_arrive_expect_tx(bar, 2 * A_PLANE);
            bulk_g2s(sb, P.Ah + aoff, 2 * A_PLANE, bar); }
          } else {
            // W blocks [bn x 32] bf16, already swizzled: (n0 * nch + kc * bn) * 64 bytes into the plane; a CTA of a
            // pair takes the rows of its half of the tile's columns
            const uint32_t wrows = PAIR ? (uint32_t)T.bn >> 1 : (uint32_t)T.bn;
            const uint32_t bytes = wrows * (BK * 2);
            const long long off = ((long long)T.n0 * (P.ldw >> 5) + (long long)kc * T.bn + (long long)rank * wrows) * (BK * 2);
            const uint8_t* src = reinterpret_cast<const uint8_t*>(piece == 1 ? P.Wh : P.Wl) + off;
            if (args.exp & 4) { mbar_arrive(bar); } else {
            mbar_arrive_expect_tx(bar, bytes);
            bulk_g2s(sb + 2 * A_PLANE + (piece == 2 ? W_PLANE : 0), src, bytes, bar); }
          }
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
      }
      }
    } else if (PAIR && warp == MMA_WARP && rank == 1) {
      // ================= relay (rank 1): "my stage s has landed" -> rank 0 =================
      int stage = 0;
      uint32_t phase = 0;
      int pi = 0;
      for (int tile = tile0; tile < args.total_tiles; tile += tile_step) {
        const Tile T = decode_tile(args, tile, pi, TILE_ROWS, rank);
        const int nk = (args.p[T.pi].K + BK - 1) / BK;
        for (int kc = 0; kc < nk; ++kc) {
          if (lane == 0) {
            mbar_wait(smem_u32(&bar_full[stage]), phase);
            mbar_arrive_cluster(map_to_cta(smem_u32(&bar_peer[stage]), 0));
          }
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == MMA_WARP && rank == 0) {
    // ================= MMA issuer =================
    // The warp stays converged (every value below is warp-uniform) and one elected lane issues: see elect_one().
    {
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    int pi = 0;
    for (int tile = tile0; tile < args.total_tiles; tile += tile_step) {
      const Tile T = decode_tile(args, tile, pi, TILE_ROWS, rank);
      const TcProb& P = args.p[T.pi];
      const int nk = (P.K + BK - 1) / BK;
      mbar_wait(smem_u32(&bar_tempty[acc]), acc_phase ^ 1);   // epilogue has drained this accumulator
      tc_fence_after();
      if (lane == 0) TRACE(1, 10);
      const uint32_t idesc = make_idesc_bf16(TILE_ROWS, T.bn);
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN_MAX);
      for (int kc = 0; kc < nk; ++kc) {
        mbar_wait(smem_u32(&bar_full[stage]), phase);
        if (PAIR) mbar_wait(smem_u32(&bar_peer[stage]), phase);   // the peer's half of the stage (no data read by this thread)
        tc_fence_after();
        if (lane == 0) TRACE(1, 11);
        if (elect_one()) {
          const uint32_t sb = smem_base + stage * STAGE_BYTES;
          const uint64_t a_hi = make_sw64_desc(sb);
          const uint64_t a_lo = make_sw64_desc(sb + A_PLANE);
          const uint64_t w_hi = make_sw64_desc(sb + 2 * A_PLANE);
          const uint64_t w_lo = make_sw64_desc(sb + 2 * A_PLANE + W_PLANE);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t adv = (uint64_t)(kk * 32 >> 4);    // 16 bf16 = 32 bytes along K
            if (PAIR) {
              mma_bf16_ss_pair(d_tmem, a_lo + adv, w_hi + adv, idesc, (kc | kk) != 0);
              mma_bf16_ss_pair(d_tmem, a_hi + adv, w_lo + adv, idesc, 1);
              mma_bf16_ss_pair(d_tmem, a_hi + adv, w_hi + adv, idesc, 1);
            } else {
              if (args.exp & 16) continue;
              if (!(args.exp & 8)) {
              mma_bf16_ss(d_tmem, a_lo + adv, w_hi + adv, idesc, (kc | kk) != 0);   // small terms first
              mma_bf16_ss(d_tmem, a_hi + adv, w_lo + adv, idesc, 1); }
              mma_bf16_ss(d_tmem, a_hi + adv, w_hi + adv, idesc, 1);
            }
          }
          if (PAIR) {
            mma_commit_pair(smem_u32(&bar_empty[stage]), 3);   // both CTAs' stage s reusable once these MMAs retire
            if (kc == nk - 1) mma_commit_pair(smem_u32(&bar_tfull[acc]), 3);
          } else {
            mma_commit(smem_u32(&bar_empty[stage]));           // smem stage reusable once these MMAs retire
            if (kc == nk - 1) mma_commit(smem_u32(&bar_tfull[acc]));
          }
        }
        __syncwarp();
        if (++stage == nst) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    }
    }
  } else {
    // ================= epilogue =================
    // tcgen05.ld gives one thread one accumulator row (TMEM lane), 32 consecutive columns per step, and the
    // row is written from that layout: fp32 rows as eight 16-byte pieces of one 128-byte line, planes as four
    // 16-byte units per plane; L2 merges the pieces. (A transposition through shared memory so that a request
    // covers whole lines was measured slower: a single warp runs at ~4 cycles per dependent instruction, so
    // the instruction count and the exposed load latency per column bound the epilogue, not the store pattern.)
    // The "+ residual" / "* x" operands of a step are requested one step ahead - for the first step of a tile
    // before the accumulator is waited for - so their latency overlaps the main loop.
    if (!A_PLANES) setmaxnreg_inc<REGS_EPILOGUE>();
    const int ew = warp - EPI_WARP0;
    const int q = ew & 3;                   // TMEM lane quarter this warp may access (EPI_WARP0 % 4 == 0)
    uint8_t* const stg = smem_gen + nst * STAGE_BYTES + ew * STG_BYTES;            // this warp's staging (if the launch reserved it)
    const uint32_t stg_s = smem_base + nst * STAGE_BYTES + ew * STG_BYTES;
    const uint32_t bar_l = smem_u32(&bar_ld[ew & 7]);
    uint32_t ld_phase = 0;                  // phase of bar_l: runs on across the tiles of this warp
    constexpr int CC_STEP = N_EPI_WARPS / 4;
    const int rt = q * 32 + lane;           // row of the tile
    int acc = 0;
    uint32_t acc_phase = 0;
    int pi = 0;
    for (int tile = tile0; tile < args.total_tiles; tile += tile_step) {
      const Tile T = decode_tile(args, tile, pi, TILE_ROWS, rank);
      const TcProb& P = args.p[T.pi];
      const int m = T.m0 + rt;
      const bool row = m < P.M;
      const float* rrow = (P.res && row) ? P.res + (long long)m * P.ldres : nullptr;
      const float* mrow = (P.mul && row) ? P.mul + (long long)m * P.ldmul : nullptr;
      float* orow = (P.O && row) ? P.O + (long long)m * P.ldo : nullptr;
      // 16-byte access needs 16-byte aligned rows (band offsets 2 and 6 floats are not)
      const bool r_vec = ((P.ldres & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.res) & 15) == 0);
      const bool m_vec = ((P.ldmul & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.mul) & 15) == 0);
      const bool o_vec = ((P.ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(P.O) & 15) == 0);
      const int ncc = min(T.bn / 32, (P.N - T.n0 + 31) / 32);   // 32-column steps; a plane block is zero padded to 32
      const int nch_o = (P.N + BK - 1) / BK;   // 32-column blocks per 128-row tile of the output (and residual) planes
      const bool res_planes = P.Rh != nullptr && row;
      float4 pr[8], pm[8];
      auto fetch = [&](int nb) {             // residual / multiplier columns nb..nb+31 of this thread's row
        if (res_planes) {                    // raw bf16 units: pr[0..3] hi, pr[4..7] lo (decoded where they are used)
          const long long blk = ((long long)(T.m0 / BM) * nch_o + (nb >> 5)) * (2 * A_PLANE);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint32_t off = sw64(rt, u);
            pr[u] = __ldg(reinterpret_cast<const float4*>(P.Rh + blk + off));
            pr[4 + u] = __ldg(reinterpret_cast<const float4*>(P.Rl + blk + off));
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int n = nb + 4 * j;
          if (!res_planes) pr[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          pm[j] = make_float4(1.f, 1.f, 1.f, 1.f);
          if (rrow) {
            if (r_vec && n + 3 < P.N) {
              pr[j] = __ldg(reinterpret_cast<const float4*>(rrow + n));
            } else {
              if (n < P.N) pr[j].x = __ldg(rrow + n);
              if (n + 1 < P.N) pr[j].y = __ldg(rrow + n + 1);
              if (n + 2 < P.N) pr[j].z = __ldg(rrow + n + 2);
              if (n + 3 < P.N) pr[j].w = __ldg(rrow + n + 3);
            }
          }
          if (mrow) {
            if (m_vec && n + 3 < P.N) {
              pm[j] = __ldg(reinterpret_cast<const float4*>(mrow + n));
            } else {
              if (n < P.N) pm[j].x = __ldg(mrow + n);
              if (n + 1 < P.N) pm[j].y = __ldg(mrow + n + 1);
              if (n + 2 < P.N) pm[j].z = __ldg(mrow + n + 2);
              if (n + 3 < P.N) pm[j].w = __ldg(mrow + n + 3);
            }
          }
        }
      };
      // Staged operands (P.ops_tma; launches with 8 KB of staging per warp): region R (bytes 0..4095) receives the
      // residual -- two 2 KB bulk copies of this warp's rows of the hi and lo plane blocks, or one 32 x 32 fp32 box by
      // 2-D TMA --, region X (4096..8191) the multiplier box. One lane issues them one step ahead (as the per-lane
      // loads were), completion is counted on the warp's mbarrier; the rows are then read back conflict-free
      // (SWIZZLE_64B planes / SWIZZLE_128B boxes). No LSU global load, no dependence on L1.
      const bool ops_tma = RMS && P.ops_tma != 0;
      // the output box shares region X with the multiplier box: the next operands are requested only after the output
      // store has read the region (end of the step) instead of right after the operands were applied
      const bool y_in_x = RMS && P.ops_tma != 0 && P.o_tma != 0 && P.mul != nullptr;
      auto fetch_staged = [&](int nb) {
        if (lane == 0) {
          const uint32_t bytes = ((P.Rh != nullptr || P.res != nullptr) ? 4096u : 0u) + (P.mul != nullptr ? 4096u : 0u);
          mbar_arrive_expect_tx(bar_l, bytes);
          if (P.Rh != nullptr) {
            const long long blk = ((long long)(T.m0 / BM) * nch_o + (nb >> 5)) * (2 * A_PLANE) + q * 2048;
            bulk_g2s(stg_s, P.Rh + blk, 2048, bar_l);
            bulk_g2s(stg_s + 2048, P.Rl + blk, 2048, bar_l);
          } else if (P.res != nullptr) {
            tma_load_2d(stg_s, &maps.r[T.pi], nb, T.m0 + q * 32, bar_l);
          }
          if (P.mul != nullptr) tma_load_2d(stg_s + 4096, &maps.m[T.pi], nb, T.m0 + q * 32, bar_l);
        }
      };
      auto consume_staged = [&]() {          // staging -> this lane's own row in pr[] / pm[] (same layout as fetch())
        if (lane == 0) mbar_wait(bar_l, ld_phase);
        ld_phase ^= 1;
        __syncwarp();
        if (P.Rh != nullptr) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            pr[u] = *reinterpret_cast<const float4*>(stg + sw64(lane, u));
            pr[4 + u] = *reinterpret_cast<const float4*>(stg + 2048 + sw64(lane, u));
          }
        } else if (P.res != nullptr) {
#pragma unroll
          for (int j = 0; j < 8; ++j) pr[j] = *reinterpret_cast<const float4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4));
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) pr[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (P.mul != nullptr) {
#pragma unroll
          for (int j = 0; j < 8; ++j) pm[j] = *reinterpret_cast<const float4*>(stg + 4096 + lane * 128 + ((j ^ (lane & 7)) << 4));
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) pm[j] = make_float4(1.f, 1.f, 1.f, 1.f);
        }
        __syncwarp();                        // every lane has read: the regions may be refilled
      };
      const bool has_rm = P.res != nullptr || P.mul != nullptr || P.Rh != nullptr;      // warp-uniform
      int cc = (args.exp & 2) ? ncc : (ew >> 2);     // (timing experiment 2: no epilogue work at all, so no operand request either)
      if (has_rm && cc < ncc) { if (ops_tma) fetch_staged(T.n0 + cc * 32); else fetch(T.n0 + cc * 32); }
      if (lane == 0) mbar_wait(smem_u32(&bar_tfull[acc]), acc_phase);
      __syncwarp();
      if (ew == BSB_TRACE_EPI && lane == 0) TRACE(2, 20);
      tc_fence_after();
      const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN_MAX);
      for (; cc < ncc; cc += CC_STEP) {
        const int nb = T.n0 + cc * 32;
        uint32_t r[32];
        tmem_ld_32x32(tbase + cc * 32, r);
        float v[32];
        // bias (zero padded to Npad >= n0 + bn; same address in every lane: one L1 transaction per load)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(P.bias + nb) + j);
          v[4 * j] = b4.x; v[4 * j + 1] = b4.y; v[4 * j + 2] = b4.z; v[4 * j + 3] = b4.w;
        }
        tmem_ld_wait32(r);
        if (ew == BSB_TRACE_EPI && lane == 0) TRACE(2, 23);
        if (P.act == ACT_LRELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = lrelu(v[j] + __uint_as_float(r[j]));
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += __uint_as_float(r[j]);
        }
        if (ops_tma && has_rm) consume_staged();
        if (res_planes || (ops_tma && P.Rh != nullptr)) {
          auto bf2 = [](float hbits, float lbits, float& even, float& odd) {     // two bf16 pairs -> hi + lo
            const uint32_t h = __float_as_uint(hbits), l = __float_as_uint(lbits);
            even = __uint_as_float(h << 16) + __uint_as_float(l << 16);
            odd = __uint_as_float(h & 0xffff0000u) + __uint_as_float(l & 0xffff0000u);
          };
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            float e0, o0, e1, o1, e2, o2, e3, o3;
            bf2(pr[u].x, pr[4 + u].x, e0, o0);
            bf2(pr[u].y, pr[4 + u].y, e1, o1);
            bf2(pr[u].z, pr[4 + u].z, e2, o2);
            bf2(pr[u].w, pr[4 + u].w, e3, o3);
            v[8 * u] = (v[8 * u] + e0) * pm[2 * u].x;
            v[8 * u + 1] = (v[8 * u + 1] + o0) * pm[2 * u].y;
            v[8 * u + 2] = (v[8 * u + 2] + e1) * pm[2 * u].z;
            v[8 * u + 3] = (v[8 * u + 3] + o1) * pm[2 * u].w;
            v[8 * u + 4] = (v[8 * u + 4] + e2) * pm[2 * u + 1].x;
            v[8 * u + 5] = (v[8 * u + 5] + o2) * pm[2 * u + 1].y;
            v[8 * u + 6] = (v[8 * u + 6] + e3) * pm[2 * u + 1].z;
            v[8 * u + 7] = (v[8 * u + 7] + o3) * pm[2 * u + 1].w;
          }
          if (cc + CC_STEP < ncc && !y_in_x) { if (ops_tma) fetch_staged(nb + CC_STEP * 32); else fetch(nb + CC_STEP * 32); }
        } else if (has_rm) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            v[4 * j] = (v[4 * j] + pr[j].x) * pm[j].x;
            v[4 * j + 1] = (v[4 * j + 1] + pr[j].y) * pm[j].y;
            v[4 * j + 2] = (v[4 * j + 2] + pr[j].z) * pm[j].z;
            v[4 * j + 3] = (v[4 * j + 3] + pr[j].w) * pm[j].w;
          }
          if (cc + CC_STEP < ncc && !y_in_x) { if (ops_tma) fetch_staged(nb + CC_STEP * 32); else fetch(nb + CC_STEP * 32); }
        }
        if (args.exp & 1) {                 // timing experiment: no split, no stores (the deferred operand request still has to happen)
          if (v[0] == 123.456f && v[31] == 1.f) orow[0] = v[5];
          if (y_in_x && cc + CC_STEP < ncc) fetch_staged(nb + CC_STEP * 32);
          continue;
        }
        if (P.O != nullptr && P.o_tma) {
          // fp32 rows through the staging buffer ([32 rows][128 B], SWIZZLE_128B: unit j of row r at j ^ (r & 7)) and
          // ONE 2-D TMA store of the 32 x 32 box; rows >= M / columns >= N are clipped by the hardware
          if (lane == 0) bulk_wait_read();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(stg + Y_OFF + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&maps.o[T.pi], stg_s + Y_OFF, nb, T.m0 + q * 32);
            bulk_commit();
          }
          if (y_in_x && cc + CC_STEP < ncc) {
            if (lane == 0) bulk_wait_read();
            fetch_staged(nb + CC_STEP * 32);
          }
        } else if (orow) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int n = nb + 4 * j;
            if (o_vec && n + 3 < P.N) {
              *reinterpret_cast<float4*>(orow + n) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            } else {
#pragma unroll
              for (int e = 0; e < 4; ++e) if (n + e < P.N) orow[n + e] = v[4 * j + e];
            }
          }
        }
        if (P.Oh != nullptr && args.stage_out) {
          // block (m-tile, 32-column chunk), hi then lo; this warp's 32 rows are 2 KB contiguous in each plane (rows
          // of 64 B, SWIZZLE_64B): they go to the warp's staging buffer and leave as two 2 KB bulk copies issued by
          // one lane. Rows beyond M are written as zeros (the block exists: mtiles is rounded up).
          const long long blk = ((long long)((args.exp & 32) ? (T.m0 / BM) & 3 : (T.m0 / BM)) * nch_o + (nb >> 5)) * (2 * A_PLANE) + q * 2048;
          if (lane == 0) bulk_wait_read();          // the previous step's copies have read the staging buffer
          __syncwarp();
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            uint4 h, l;
            split2(v[8 * u], v[8 * u + 1], h.x, l.x);
            split2(v[8 * u + 2], v[8 * u + 3], h.y, l.y);
            split2(v[8 * u + 4], v[8 * u + 5], h.z, l.z);
            split2(v[8 * u + 6], v[8 * u + 7], h.w, l.w);
            if (!row) { h = make_uint4(0u, 0u, 0u, 0u); l = h; }
            const uint32_t off = sw64(lane, u);     // = sw64(rt, u) - q * 2048
            *reinterpret_cast<uint4*>(stg + off) = h;
            *reinterpret_cast<uint4*>(stg + 2048 + off) = l;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            bulk_s2g(P.Oh + blk, stg_s, 2048);
            bulk_s2g(P.Ol + blk, stg_s + 2048, 2048);
            bulk_commit();
          }
        } else if (P.Oh && row) {
          // block (m-tile, 32-column chunk), hi then lo: this step is the whole 64-byte row rt of both
          const long long blk = ((long long)((args.exp & 32) ? (T.m0 / BM) & 3 : (T.m0 / BM)) * nch_o + (nb >> 5)) * (2 * A_PLANE);
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            uint4 h, l;
            split2(v[8 * u], v[8 * u + 1], h.x, l.x);
            split2(v[8 * u + 2], v[8 * u + 3], h.y, l.y);
            split2(v[8 * u + 4], v[8 * u + 5], h.z, l.z);
            split2(v[8 * u + 6], v[8 * u + 7], h.w, l.w);
            const uint32_t off = sw64(rt, u);
            *reinterpret_cast<uint4*>(P.Oh + blk + off) = h;
            *reinterpret_cast<uint4*>(P.Ol + blk + off) = l;
          }
        }
        if (ew == BSB_TRACE_EPI && lane == 0) TRACE(2, 24);
      }
      tc_fence_before();
      if (PAIR) {
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&bar_tempty[acc]), 0));
      } else {
        mbar_arrive(smem_u32(&bar_tempty[acc]));
      }
      if (ew == BSB_TRACE_EPI && lane == 0) TRACE(2, 21);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  if (warp >= EPI_WARP0 && lane == 0) bulk_wait_read();   // outstanding plane stores still read this CTA's shared memory
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();   // the peer may still arrive on this CTA's barriers / read its smem
  if (warp == MMA_WARP) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// Plane-operand launches on CTA pairs (cta_group::2, 256-row tiles, each CTA loads half of the W rows) are verified
// bit-identical to the single-CTA path but measured slower on this workload (5.2 vs 4.5 ms/step, profiles/
// r01_summary.md): opt-in for experiments.
const bool g_pair = getenv("BSB_PAIR") != nullptr;
// timing experiments only (tools/band_cost.py, tools/exp_cost.py; results are then wrong on purpose):
// g_band_mask bit i = problem i of a grouped launch takes part;
// g_exp: 1 = epilogue skips split + stores, 2 = epilogue only releases the accumulator, 4 = loaders skip the bulk
// copies, 8 = one MMA of three, 16 = no MMAs (commit only), 32 = plane stores go to an L2-resident window
std::atomic<int> g_band_mask{0xfff};
std::atomic<int> g_exp{0};

}  // namespace

namespace {
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}
// fp32 matrix [rows][cols] with row pitch ld (floats) as boxes of 32 x 32; false when TMA's alignment rules do not hold
bool make_map_f32(CUtensorMap* m, const float* base, int rows, int cols, int ld) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc || base == nullptr || (reinterpret_cast<uintptr_t>(base) & 15) != 0 || (ld & 3) != 0 || cols < 1 || rows < 1) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
  const cuuint32_t box[2] = {32, 32};
  const cuuint32_t estr[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
}  // namespace

// Launch the problems of `args` that have K > 0 on the tensor-core kernel. Problems are taken
// in descending-K order so that the static round-robin schedule ends with the cheap tiles.
cudaError_t launch_tc_grouped_linear(const GroupedArgs& args, cudaStream_t s) {
  static SmemOptIn opt[3], opt_rms;
  int dev = 0;
  cudaError_t e = opt_in_smem(tc_grouped_linear_kernel<false, false>, smem_bytes<false>() + EPI_STAGE_TOTAL, opt[0], &dev);
  if (e != cudaSuccess) return e;
  if ((e = opt_in_smem(tc_grouped_linear_kernel<true, false>, smem_bytes<false>() + EPI_STAGE_TOTAL, opt[1])) != cudaSuccess) return e;
  if ((e = opt_in_smem(tc_grouped_linear_kernel<true, true>, smem_bytes<true>() + EPI_STAGE_TOTAL, opt[2])) != cudaSuccess) return e;
  if ((e = opt_in_smem(tc_grouped_linear_kernel<true, false, true>, smem_bytes<false>() + EPI_STAGE_TOTAL, opt_rms)) != cudaSuccess) return e;
  const int sms = sm_count(dev);

  int order[kMaxProb], n = 0;
  const int mask = g_band_mask.load(std::memory_order_relaxed);
  for (int i = 0; i < args.nprob; ++i)
    if (args.p[i].K > 0 && args.p[i].M > 0 && args.p[i].N > 0 && ((mask >> i) & 1)) order[n++] = i;
  for (int i = 1; i < n; ++i)               // insertion sort by K descending
    for (int j = i; j > 0 && args.p[order[j]].K > args.p[order[j - 1]].K; --j) {
      const int t = order[j]; order[j] = order[j - 1]; order[j - 1] = t;
    }
  TcArgs ta{};
  int total = 0, n_planes = 0;
  bool rm = false, planes_out = false, rows_out = false;
  for (int i = 0; i < n; ++i) {
    const LinProb& L = args.p[order[i]];
    n_planes += L.Ap != nullptr;
    rm = rm || L.res != nullptr || L.mul != nullptr || L.Rp != nullptr;
    planes_out = planes_out || L.Op != nullptr;
    rows_out = rows_out || L.O != nullptr;
  }
  // (the fp32-row-fed kernel keeps direct stores too: measured pre.0 236 -> 250 us, back.0 86 -> 99 us with staging)
  const bool pair_req = n_planes != 0 && g_pair;
  // launches with "+ residual" / "* multiplier" operands: 3 main-loop stages + 8 KB of staging per epilogue warp
  // ([R | X or Y]); the others: 4 stages + 4 KB ([planes or Y])
  const bool rm_stage = rm && n_planes != 0 && !pair_req;
  const bool reserve = rm_stage || ((planes_out || rows_out) && !rm && n_planes != 0);
  const int stages = rm_stage ? 3 : Smem<false>::STAGES;
  const int stg_bytes = rm_stage ? 2 * EPI_STAGE_BYTES : EPI_STAGE_BYTES;
  const int extra = (reserve ? 8 * stg_bytes : 0) - (Smem<false>::STAGES - stages) * Smem<false>::STAGE_BYTES;
  TcMaps maps;
  if (n_planes != 0 && n_planes != n) return cudaErrorInvalidValue;   // one A source per launch
  const bool pair = pair_req;                                     // plane operands: CTA pairs, 256-row tiles
  const int tile_rows = pair ? 2 * BM : BM;
  for (int i = 0; i < n; ++i) {
    const LinProb& L = args.p[order[i]];
    TcProb& P = ta.p[i];
    P.Ah = reinterpret_cast<const uint8_t*>(L.Ap);
    P.Al = L.Ap ? P.Ah + L.a_lo : nullptr;
    P.Oh = reinterpret_cast<uint8_t*>(L.Op);
    P.Ol = L.Op ? P.Oh + L.o_lo : nullptr;
    P.Rh = reinterpret_cast<const uint8_t*>(L.Rp);
    P.Rl = L.Rp ? P.Rh + L.r_lo : nullptr;
    P.A = L.A; P.Wh = reinterpret_cast<const __nv_bfloat16*>(L.Wh); P.Wl = reinterpret_cast<const __nv_bfloat16*>(L.Wl);
    P.bias = L.bias; P.O = L.O; P.res = L.res; P.mul = L.mul;
    P.lda = L.lda; P.ldw = L.ldw64; P.ldo = L.ldo; P.ldres = L.ldres; P.ldmul = L.ldmul;
    P.M = L.M; P.N = L.N; P.K = L.K; P.act = L.act;
    // (a TMA store moves whole 16-byte units: the extent must be a multiple of 4 floats, or run into padding the row owns;
    // in a residual/multiplier launch the output shares its staging region with the multiplier box)
    const int n4 = (L.N + 3) & ~3;
    const bool has_ops = L.res != nullptr || L.mul != nullptr || L.Rp != nullptr;
    P.ops_tma = (rm_stage && has_ops && (L.res == nullptr || make_map_f32(&maps.r[i], L.res, L.M, L.N, L.ldres)) &&
                 (L.mul == nullptr || make_map_f32(&maps.m[i], L.mul, L.M, L.N, L.ldmul))) ? 1 : 0;
    // (with a multiplier the output box reuses region X: only when that problem's operands are staged, see y_in_x)
    P.o_tma = (reserve && L.O != nullptr && n4 - L.N <= L.o_tail_pad && !(rm_stage && L.mul != nullptr && !P.ops_tma) &&
               make_map_f32(&maps.o[i], L.O, L.M, n4, L.ldo)) ? 1 : 0;
    P.tile_start = total;
    P.tiles_m = (L.M + tile_rows - 1) / tile_rows;
    total += P.tiles_m * ((L.N + BN_MAX - 1) / BN_MAX);
  }
  ta.nprob = n;
  ta.total_tiles = total;
  ta.exp = g_exp.load(std::memory_order_relaxed);
  ta.stage_out = (reserve && !rm_stage) ? 1 : 0;     // plane output through staging (never in a residual/multiplier launch)
#ifdef BSB_TRACE
  static int trace_launch = 0, trace_sel = -2;
  if (trace_sel == -2) trace_sel = getenv("BSB_TRACE_LAUNCH") ? atoi(getenv("BSB_TRACE_LAUNCH")) : -1;
  ta.trace = (trace_launch++ == trace_sel);
#endif

  if (total == 0) return cudaSuccess;
  if (pair) {
    const int pairs = total < sms / 2 ? total : sms / 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * pairs);
    cfg.blockDim = dim3(Roles<true>::THREADS);
    cfg.dynamicSmemBytes = smem_bytes<true>() + extra;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, tc_grouped_linear_kernel<true, true>, ta, maps);
    count_launch();
    return e != cudaSuccess ? e : cudaGetLastError();
  }
  const int grid = total < sms ? total : sms;
  if (rm_stage)
    tc_grouped_linear_kernel<true, false, true><<<grid, Roles<true>::THREADS, smem_bytes<false>() + extra, s>>>(ta, maps);
  else if (n_planes)
    tc_grouped_linear_kernel<true, false><<<grid, Roles<true>::THREADS, smem_bytes<false>() + extra, s>>>(ta, maps);
  else
    tc_grouped_linear_kernel<false, false><<<grid, Roles<false>::THREADS, smem_bytes<false>() + extra, s>>>(ta, maps);
  count_launch();
  return cudaGetLastError();
}

extern "C" int bsb_debug_set_band_mask(int mask) { g_band_mask.store(mask); return 0; }
extern "C" int bsb_debug_set_exp(int e) { g_exp.store(e); return 0; }

#ifdef BSB_TRACE
extern "C" int bsb_debug_set_trace(void* p) {
  return (int)cudaMemcpyToSymbol(g_trace, &p, sizeof(p));
}
#endif

}  // namespace bsb
