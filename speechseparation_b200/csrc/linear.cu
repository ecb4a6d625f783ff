// Grouped Linear layers (fp32 FFMA path).
//
// One launch evaluates the same MLP layer of all 11 bands (bsrnn.py:404-412, 420-430): each
// band is an independent problem O = act(A.W^T + b) with its own K/N (2..768); CTAs are
// mapped to (problem, m-tile, n-tile) through a prefix table in the kernel parameters.
// A is read K-major straight from the frame-major activations (band slice = column offset),
// W is the packed, zero padded nn.Linear weight (K contiguous), so neither operand is
// transposed in memory.  Bias, LeakyReLU(0.01), the decoder's "+ residual" (bsrnn.py:425)
// and the final "x * mask" (bsrnn.py:441) are fused into the epilogue.
#include "kernels.cuh"

#include <cstdlib>

namespace bsb {

std::atomic<long long> g_launches{0};

int sm_count(int dev) {
  static std::atomic<int> cache[64];
  if (dev < 0 || dev >= 64) return 148;
  int v = cache[dev].load(std::memory_order_relaxed);
  if (v <= 0) {
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) return 148;
    cache[dev].store(v, std::memory_order_relaxed);
  }
  return v;
}

namespace {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int TM = 8, TN = 4;
constexpr int NT = 256;
constexpr int AS = BM + 4;   // padded smem row strides (floats)
constexpr int BS = BN + 4;

__device__ __forceinline__ float lrelu(float v) { return v > 0.f ? v : v * 0.01f; }

__global__ void __launch_bounds__(NT, 2)
grouped_linear_kernel(const __grid_constant__ GroupedArgs args) {
  __shared__ __align__(16) float As[2][BK][AS];
  __shared__ __align__(16) float Bs[2][BK][BS];

  const int tile = blockIdx.x;
  int pi = 0;
  while (pi + 1 < args.nprob && tile >= args.p[pi + 1].tile_start) ++pi;
  const LinProb& P = args.p[pi];
  const int local = tile - P.tile_start;
  const int m0 = (local / P.tiles_n) * BM;
  const int n0 = (local % P.tiles_n) * BN;

  const int tid = threadIdx.x;
  const float* __restrict__ A = P.A;
  const float* __restrict__ W = P.W;
  const int M = P.M, K = P.K, lda = P.lda, ldw = P.ldw;
  const bool a_vec = ((lda & 3) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);

  // global->smem assignment: A tile 128x16 = 512 float4 (2 per thread), W tile 64x16 = 256 float4
  const int a_row = tid >> 2;          // 0..63 (+64)
  const int a_k = (tid & 3) * 4;
  const int w_row = tid >> 2;          // 0..63
  const int w_k = (tid & 3) * 4;

  float4 ra[2], rw;
  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int m = m0 + a_row + i * 64;
      const int k = k0 + a_k;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) {
        const float* p = A + (long long)m * lda + k;
        if (a_vec && k + 3 < K) {
          v = *reinterpret_cast<const float4*>(p);
        } else {
          if (k + 0 < K) v.x = p[0];
          if (k + 1 < K) v.y = p[1];
          if (k + 2 < K) v.z = p[2];
          if (k + 3 < K) v.w = p[3];
        }
      }
      ra[i] = v;
    }
    rw = *reinterpret_cast<const float4*>(W + (long long)(n0 + w_row) * ldw + k0 + w_k);
  };
  auto store_tiles = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int r = a_row + i * 64;
      As[buf][a_k + 0][r] = ra[i].x;
      As[buf][a_k + 1][r] = ra[i].y;
      As[buf][a_k + 2][r] = ra[i].z;
      As[buf][a_k + 3][r] = ra[i].w;
    }
    Bs[buf][w_k + 0][w_row] = rw.x;
    Bs[buf][w_k + 1][w_row] = rw.y;
    Bs[buf][w_k + 2][w_row] = rw.z;
    Bs[buf][w_k + 3][w_row] = rw.w;
  };

  const int tx = tid & 15;   // n direction
  const int ty = tid >> 4;   // m direction
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int nk = (K + BK - 1) / BK;
  if (nk > 0) {
    load_tiles(0);
    store_tiles(0);
  }
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) load_tiles((kt + 1) * BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * TM]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * TM + 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * TN]);
      const float av[TM] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[TN] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      store_tiles(buf ^ 1);
      __syncthreads();
    }
  }

  // epilogue
  const int N = P.N, ldo = P.ldo;
  float* __restrict__ O = P.O;
  const float* __restrict__ res = P.res;
  const float* __restrict__ mul = P.mul;
  const int n = n0 + tx * TN;
  if (n >= N) return;
  const float4 bb = *reinterpret_cast<const float4*>(P.bias + n);   // bias is padded to Npad
  const float bv[TN] = {bb.x, bb.y, bb.z, bb.w};
  const bool o_vec = ((ldo & 3) == 0) && ((reinterpret_cast<uintptr_t>(O) & 15) == 0) && (n + 3 < N);
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + ty * TM + i;
    if (m >= M) continue;
    float v[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      float t = acc[i][j] + bv[j];
      if (P.act == ACT_LRELU) t = lrelu(t);
      v[j] = t;
    }
    if (res != nullptr) {
      const float* r = res + (long long)m * P.ldres + n;
#pragma unroll
      for (int j = 0; j < TN; ++j)
        if (n + j < N) v[j] += r[j];
    }
    if (mul != nullptr) {
      const float* r = mul + (long long)m * P.ldmul + n;
#pragma unroll
      for (int j = 0; j < TN; ++j)
        if (n + j < N) v[j] *= r[j];
    }
    float* o = O + (long long)m * ldo + n;
    if (o_vec) {
      *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int j = 0; j < TN; ++j)
        if (n + j < N) o[j] = v[j];
    }
  }
}

// ---- small-M variant (streaming hops: M = rows of one frame, 1..127) ----------------------------------------
// With a handful of rows the layer is a weight read: 4.2 MB of fp32 weights for a d x d layer against 2 x 2050
// activations. The tiled kernel above walks K serially in 16-wide steps behind block barriers with 12 CTAs on its
// largest band (51 us per d x d layer at 2 rows); here a warp owns GV_COLS output columns, its lanes stride over K
// with 16-byte weight loads (one 512-byte run per row and iteration), the <= 8 activation rows of the CTA sit in
// shared memory, and the partial sums meet in a shuffle reduction. CTA = 8 warps = 16 columns of one band x 8 rows.
// (2-row hop with 2 / 4 / 1 columns per warp: 0.195 / 0.207 / 0.195 ms, 24 rows: 0.345 / 0.349 / 0.365; 4 or 16 warps per CTA: slower)
constexpr int GV_WARPS = 8, GV_COLS = 2, GV_MR = 8, GV_BN = GV_WARPS * GV_COLS;
constexpr int GV_KMAX = 768;                       // widest band (kMaxBandD); checked by the launcher

__global__ void __launch_bounds__(GV_WARPS * 32)
grouped_gemv_kernel(const __grid_constant__ GroupedArgs args) {
  __shared__ __align__(16) float As[GV_MR][GV_KMAX];
  const int tile = blockIdx.x;
  int pi = 0;
  while (pi + 1 < args.nprob && tile >= args.p[pi + 1].tile_start) ++pi;
  const LinProb& P = args.p[pi];
  const int local = tile - P.tile_start;
  const int m0 = (local / P.tiles_n) * GV_MR;
  const int n0 = (local % P.tiles_n) * GV_BN;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int M = P.M, K = P.K, N = P.N, ldw = P.ldw;
  const int Kp = (K + 127) & ~127;                 // a warp covers 128 k per iteration
  const float* __restrict__ A = P.A;
  for (int e = tid; e < GV_MR * Kp; e += GV_WARPS * 32) {
    const int r = e / Kp, k = e - r * Kp;
    As[r][k] = (m0 + r < M && k < K) ? __ldg(A + (long long)(m0 + r) * P.lda + k) : 0.f;
  }
  __syncthreads();
  const int nw = n0 + w * GV_COLS;
  float acc[GV_COLS][GV_MR];
#pragma unroll
  for (int c = 0; c < GV_COLS; ++c)
#pragma unroll
    for (int r = 0; r < GV_MR; ++r) acc[c][r] = 0.f;
  const float* __restrict__ W = P.W + (long long)nw * ldw;        // rows exist up to Npad (multiple of 64)
#pragma unroll 3
  for (int k = lane * 4; k < Kp; k += 128) {
    float4 wv[GV_COLS];
#pragma unroll
    for (int c = 0; c < GV_COLS; ++c)
      wv[c] = k < ldw ? __ldg(reinterpret_cast<const float4*>(W + (long long)c * ldw + k)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < GV_MR; ++r) {
      const float4 av = *reinterpret_cast<const float4*>(&As[r][k]);
#pragma unroll
      for (int c = 0; c < GV_COLS; ++c) {
        acc[c][r] = fmaf(av.x, wv[c].x, acc[c][r]);
        acc[c][r] = fmaf(av.y, wv[c].y, acc[c][r]);
        acc[c][r] = fmaf(av.z, wv[c].z, acc[c][r]);
        acc[c][r] = fmaf(av.w, wv[c].w, acc[c][r]);
      }
    }
  }
  float mine = 0.f;                                // lane c * 8 + r keeps the sum of (column c, row r)
#pragma unroll
  for (int c = 0; c < GV_COLS; ++c)
#pragma unroll
    for (int r = 0; r < GV_MR; ++r) {
      float v = acc[c][r];
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == c * GV_MR + r) mine = v;
    }
  if (lane >= GV_COLS * GV_MR) return;
  const int n = nw + lane / GV_MR, m = m0 + lane % GV_MR;
  if (n >= N || m >= M) return;
  float t = mine + __ldg(P.bias + n);
  if (P.act == ACT_LRELU) t = lrelu(t);
  if (P.res != nullptr) t += P.res[(long long)m * P.ldres + n];
  if (P.mul != nullptr) t *= P.mul[(long long)m * P.ldmul + n];
  P.O[(long long)m * P.ldo + n] = t;
}

}  // namespace

// every row count below the tensor-core threshold (hop of 40 / 64 / 100 rows: 0.65 / 0.66 / 0.69 -> 0.38 / 0.44 / 0.52 ms;
// the tiled kernel above then only serves BSB_FORCE_FP32 and the K = 0 bias rows)
constexpr int kGemvMaxRows = kTcMinRows - 1;

cudaError_t launch_gemv_grouped_linear(GroupedArgs& args, cudaStream_t s) {
  int total = 0;
  for (int i = 0; i < args.nprob; ++i) {
    LinProb& P = args.p[i];
    P.tile_start = total;
    P.tiles_n = (P.N + GV_BN - 1) / GV_BN;
    total += P.tiles_n * ((P.M + GV_MR - 1) / GV_MR);
  }
  args.total_tiles = total;
  if (total == 0) return cudaSuccess;
  grouped_gemv_kernel<<<total, GV_WARPS * 32, 0, s>>>(args);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_ffma_grouped_linear(GroupedArgs& args, cudaStream_t s) {
  int total = 0;
  for (int i = 0; i < args.nprob; ++i) {
    LinProb& P = args.p[i];
    P.tile_start = total;
    P.tiles_n = (P.N + BN - 1) / BN;
    const int tiles_m = (P.M + BM - 1) / BM;
    total += P.tiles_n * tiles_m;
  }
  args.total_tiles = total;
  if (total == 0) return cudaSuccess;
  grouped_linear_kernel<<<total, NT, 0, s>>>(args);
  count_launch();
  return cudaGetLastError();
}

static std::atomic<int> g_force_ffma{0};
void set_force_ffma(int on) { g_force_ffma.store(on); }
int force_ffma() { return g_force_ffma.load(std::memory_order_relaxed); }

cudaError_t launch_grouped_linear(GroupedArgs& args, cudaStream_t s) {
  bool tc = !force_ffma() && args.nprob > 0;
  for (int i = 0; i < args.nprob; ++i)
    if (args.p[i].K > 0 && (args.p[i].M < kTcMinRows || args.p[i].Wh == nullptr)) tc = false;
  if (!tc) {
    static const bool no_gemv = getenv("BSB_NO_GEMV") != nullptr;      // A/B switches
    static const int gemv_max = getenv("BSB_GEMV_MAX") ? atoi(getenv("BSB_GEMV_MAX")) : kGemvMaxRows;
    bool gemv = args.nprob > 0 && !no_gemv;
    for (int i = 0; i < args.nprob; ++i)
      if (args.p[i].M > gemv_max || args.p[i].K > GV_KMAX || args.p[i].O == nullptr || (args.p[i].K > 0 && args.p[i].A == nullptr)) gemv = false;
    return gemv ? launch_gemv_grouped_linear(args, s) : launch_ffma_grouped_linear(args, s);
  }
  cudaError_t e = launch_tc_grouped_linear(args, s);
  if (e != cudaSuccess) return e;
  GroupedArgs rest{};                       // K == 0 problems: out = bias broadcast
  for (int i = 0; i < args.nprob; ++i)
    if (args.p[i].K == 0) rest.p[rest.nprob++] = args.p[i];
  return rest.nprob ? launch_ffma_grouped_linear(rest, s) : cudaSuccess;
}

}  // namespace bsb
