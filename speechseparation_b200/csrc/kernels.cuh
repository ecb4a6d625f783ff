// Internal kernel-launch interface shared by the .cu files of libbsrnn_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include "layout.h"

namespace bsb {

// every kernel launch goes through this counter (bsb_launch_count); host threads may launch concurrently
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) applies to the CURRENT device only: one flag per (kernel, device).
// `*dev_out` receives the current device. Racing first calls both set the attribute, which is harmless.
struct SmemOptIn { std::atomic<bool> done[64]; };
template <class K>
inline cudaError_t opt_in_smem(K kernel, int bytes, SmemOptIn& st, int* dev_out = nullptr) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev_out) *dev_out = dev;
  if (dev < 0 || dev >= 64) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (!st.done[dev].load(std::memory_order_acquire)) {
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return e;
    st.done[dev].store(true, std::memory_order_release);
  }
  return cudaSuccess;
}
// SM count of a device, cached
int sm_count(int dev);

// ------------------------------------------------------------------------------------
// grouped Linear:  O[m][n] = epi( sum_k A[m][k] * W[n][k] + bias[n] )
// up to kMaxProb independent problems (the 11 bands of one MLP layer) in one launch.
// ------------------------------------------------------------------------------------
enum Act : int { ACT_NONE = 0, ACT_LRELU = 1 };

struct LinProb {
  const float* A;      // [M][lda], K valid columns
  const float* W;      // [Npad][ldw] zero padded (ldw = Kpad)
  const void* Wh;      // bf16 split planes [Npad][ldw64] (tensor-core path), hi and lo
  const void* Wl;
  const float* bias;   // [Npad]
  float* O;            // [M][ldo]
  const float* res;    // optional: v += res[m][n]   (after activation)
  const float* mul;    // optional: v *= mul[m][n]   (after res)
  int lda, ldw, ldw64, ldo, ldres, ldmul;
  int M, N, K;
  int act;
  int tile_start;      // first tile index of this problem in the launch
  int tiles_n;
  // Tensor-core path only: activations that go from one Linear straight into the next never exist as fp32
  // rows. The producing epilogue writes them as "planes": bf16 hi and lo (x = hi + lo), blocked as
  // [m-tile of 128 rows][k-chunk of 32 columns][hi | lo][128 rows x 64 B, SWIZZLE_64B] = the shared-memory image the
  // consuming layer's tcgen05.mma reads, so its A operand is one 16 KB bulk copy per chunk. The *_lo offsets are the
  // distance hi -> lo inside a block (kPlaneBlock).
  const void* Ap;      // optional input planes (hi at Ap, lo at Ap + a_lo); A may then be null
  void* Op;            // optional output planes (hi at Op, lo at Op + o_lo); O may then be null
  long long a_lo, o_lo;
  const void* Rp;      // optional: the residual operand as planes of the output's geometry (then res is null)
  long long r_lo;
  int o_tail_pad;      // floats after column N-1 of every O row that may be overwritten (row padding owned by nobody else):
                       // lets a 16-byte-granular TMA store serve an N that is not a multiple of 4 (band 10: 514 + 2 of kXS)
};
constexpr int kMaxProb = 12;
constexpr int kPlaneK = 32;                                              // columns per plane block = K chunk of the GEMM
constexpr int kPlaneBlock = 128 * kPlaneK * 2;                           // bytes of one [128 x 32] bf16 block
inline int plane_chunks(int n) { return (n + kPlaneK - 1) / kPlaneK; }
// byte offset of 16-byte unit `c` (0..3) of row `r` in a K-major SWIZZLE_64B image (rows of 64 B; verified on the
// device by tools/sw64_probe.cu: layout code 4, unit ^= (row >> 1) & 3, 8-row groups 512 B apart)
__host__ __device__ inline unsigned sw64(int r, int c) { return (unsigned)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4)); }
struct GroupedArgs {
  LinProb p[kMaxProb];
  int nprob;
  int total_tiles;
};

// fills tile_start/tiles_n/total_tiles and launches (no-op when there is no work).
// Dispatch by shape: M >= kTcMinRows -> tensor-core kernel (tc_linear.cu), else the FFMA kernel
// (streaming hops, M = rows of one frame); K == 0 problems (band 11 constant) always FFMA.
constexpr int kTcMinRows = 128;     // (measured: a hop of 128 / 160 / 256 rows takes 0.47 / 0.48 / 0.53 ms on the tensor-core
                                    // Linears, 0.54 / 0.61 / 0.81 ms on the small-M kernel: the threshold stays at one full tile)
cudaError_t launch_grouped_linear(GroupedArgs& args, cudaStream_t s);
cudaError_t launch_ffma_grouped_linear(GroupedArgs& args, cudaStream_t s);
cudaError_t launch_tc_grouped_linear(const GroupedArgs& args, cudaStream_t s);
// 0 = shape-based dispatch (default), 1 = force the fp32 FFMA kernel everywhere (A/B checks)
void set_force_ffma(int on);

// ------------------------------------------------------------------------------------
// LSTM recurrence over precomputed gate pre-activations.
//   gates(seq, step) = xproj(seq, step)[dir*256 .. +256] + W_hh[dir] . h(seq, step-1)
// element offset of (seq, step) = (seq / n_inner) * outer + (seq % n_inner) * inner + step * stepstride
// ------------------------------------------------------------------------------------
struct LstmArgs {
  const float* xproj;
  const float* whh;          // [ndir][256][64]
  float* hout;               // h of every step; column offset dir*64
  const float* h0;           // optional [nseq][64] (dir 0 only)
  const float* c0;
  float* hN;                 // optional final state [nseq][64]
  float* cN;
  long long xp_outer, xp_inner, xp_step;
  long long ho_outer, ho_inner, ho_step;
  int n_inner;
  int nseq, nsteps, ndir;
};
cudaError_t launch_lstm(const LstmArgs& a, cudaStream_t s);

// One LSTM layer with the input projection fused, on tcgen05 tensor cores (lstm_tc.cu): tiles of
// 128 sequences, zero initial state. element (seq, step, k) of the input is x[seq*x_seq + step*x_step + k],
// h(seq, step, dir*64 + u) is written to hout[seq*h_seq + step*h_step + dir*64 + u].
struct LstmTcArgs {
  const float* x;
  float* hout;
  const void* wimg[2];       // per direction: layout.h RnnBlock::tcw image
  const float* bias[2];      // per direction: RnnBlock::tcb
  long long x_seq, x_step, h_seq, h_step;
  int nseq, nsteps, ndir, kin;
  // optional: h goes out as bf16 hi/lo planes (LinProb::Ap geometry, row = seq * nsteps + step, 64-column chunk =
  // direction) for the Linear that follows, instead of fp32 rows at hout
  void* hplanes;
  long long hp_lo;
};
cudaError_t launch_lstm_tc(const LstmTcArgs& a, cudaStream_t s);
// Both layers of a time-axis block in ONE launch, 16 sequences per group, layer 1 a few steps behind layer 0
// (lstm_seq2.cu). x, hmid (layer-0 output = layer-1 input) and hout share one addressing:
// (seq / n_inner) * outer + (seq % n_inner) * inner + step * step.
struct LstmSeq2Args {
  const float* x;
  float* hmid;
  float* hout;
  const void* wfrag[2];      // per layer: layout.h RnnBlock::sqw
  const float* bias[2];      // RnnBlock::sqb
  const float* h0[2];        // optional initial state [nseq][64] per layer
  const float* c0[2];
  float* hN[2];              // optional final state
  float* cN[2];
  long long outer, inner, step;
  int n_inner, nseq, nsteps;
  void* hplanes;             // optional: layer-1 h as bf16 hi/lo planes instead of fp32 rows at hout
  long long hp_lo;
  int* flags;                // [2 * groups + 1] ints of scratch (zeroed by the launch): publication flags, CTA ticket counter
  int group;                 // sequences per CTA (8 or 16), chosen by the launch
  // optional: the block's output Linear + residual fused as a third CTA per group: fc_out = W_fc . h_layer1 + b + x
  const float* fcw;          // fp32 [64][ldfcw]
  const float* fcb;          // [64]
  float* fc_out;             // same addressing as x; must not alias x
  int ldfcw;
  // optional (fused fc only): the block's output ALSO / INSTEAD as bf16 hi/lo planes in the geometry of the grouped
  // Linear that consumes it next (LinProb::Ap with K = 64: per band [m-tile of 128 frame-rows][2 chunks][hi | lo]):
  // band b starts fcp_band bytes after band b-1, lo = hi + fcp_lo. With fc_skip_rows the fp32 rows are not written.
  void* fc_planes;
  long long fcp_band, fcp_lo;
  int fc_skip_rows;
  int fuse_fc, ngroups;      // set by the launch
  int exp;                   // timing-experiment switches (BSB_SEQ2_EXP: 1 no cell update, 2 no recurrent MMAs, 8 no x loads,
                             // 16 no h flush), 0 in production
};
// *fused (optional) = 1 when the launch carried the block's output Linear + residual (third CTA per group)
cudaError_t launch_lstm_seq2(const LstmSeq2Args& a, cudaStream_t s, int* fused = nullptr);
constexpr int kLstmSeqMinSteps = 2;  // single steps (streaming hops) stay on the FFMA recurrence
// Band-axis blocks below this many sequences (frame-rows) run as GEMM projections + the FFMA recurrence: the tcgen05
// layer kernel puts 128 sequences on one CTA for ~66-80 us whatever the count, so 128 rows = 2 CTAs on 148 SMs.
// Measured crossover (hop of R rows, device ms, tcgen05 vs split): 128: 0.61/0.50, 256: 0.65/0.54, 512: 0.74/0.67,
// 1024: 0.87/0.94; cfg 1 (346 frame-rows): 1.06/0.99 ms. Override for A/B runs: env BSB_LSTM_TC_MIN.
constexpr int kLstmTcMinSeq = 768;
int force_ffma();

// ------------------------------------------------------------------------------------
// signal kernels
// ------------------------------------------------------------------------------------
// wav[C][L] -> xf[C*T][xs] (frame-major, 2050 valid floats), torch.stft(center=True, reflect)
cudaError_t launch_stft_frames(const float* wav, int C, long long L, int T, const float* window,
                               float* xf, int xs, cudaStream_t s);
// xf[C*T][xs] -> wav[C][1024*(T-1)], torch.istft semantics
cudaError_t launch_istft_frames(const float* xf, int xs, int C, int T, const float* window,
                                float* wav, cudaStream_t s);
// streaming analysis: ring[C][2048] <- roll + chunk; x[C][2050] = pack(rfft(ring*window))
cudaError_t launch_stream_analysis(const float* chunk, int C, const float* window, float* ring,
                                   float* x, cudaStream_t s);
// streaming synthesis: cur = irfft(unpack(y)); out = (cur[:1024] + tail) / (w[i]+w[i+1024]); tail = cur[1024:]
cudaError_t launch_stream_synthesis(const float* y, const float* dry, const float* mix, int C, const float* window,
                                    float* tail, float* out, cudaStream_t s);
// x[C][2050][T] <-> xf[C*T][xs]
cudaError_t launch_to_frames(const float* x, int C, int T, float* xf, int xs, cudaStream_t s);
cudaError_t launch_from_frames(const float* xf, int xs, int C, int T, float* x, cudaStream_t s);
// wav[C][L] -> x[C][2050][T] through the file-scope __device__ scratch of fft.cu (no workspace, no allocation);
// C * T <= static_max_frame_rows()
cudaError_t launch_stft_pack_static(const float* wav, int C, long long L, int T, const float* window, float* x,
                                    cudaStream_t s);
int static_max_frame_rows();
const float* builtin_window(cudaStream_t s, cudaError_t* err);   // device float[2048], periodic Hann
const float2* twiddle_tables(cudaStream_t s, cudaError_t* err);  // device: [1024] W_1024^k then [1025] W_2048^k
// creates the window / twiddle tables of the CURRENT device (the library's only device allocation: 28 KB, once per
// device). bsb_init() calls it so that no launch call allocates; a launch on a device never initialised creates them lazily.
cudaError_t init_signal_tables();

// ------------------------------------------------------------------------------------
// evaluation sums (m_dataset.py:202-226): stats[rows][5] = sum|e-r|, sum r^2, sum (e-r)^2, sum e r, sum e^2
// ------------------------------------------------------------------------------------
cudaError_t launch_pair_stats(const float* est, long long est_stride, const float* ref, long long ref_stride,
                              int rows, long long n, double* stats, cudaStream_t s);

// 16-bit PCM <-> float at the host-buffer boundary (metrics.cu)
cudaError_t launch_pcm16_to_float(const short* pcm, long long n, float* wav, cudaStream_t s);
cudaError_t launch_float_to_pcm16(const float* wav, long long n, short* pcm, cudaStream_t s);
// sum over samples of the channel-summed log ratio of m_dataset.py:219-222 (see metrics.cu)
cudaError_t launch_channel_sdr(const float* mix, long long mix_stride, const float* src, long long src_stride,
                               int rows, long long n, double* out, cudaStream_t s);

// ------------------------------------------------------------------------------------
// weight packing helpers
// ------------------------------------------------------------------------------------
cudaError_t launch_pack_matrix(const float* src, int N, int K, float* dst, int ldd, cudaStream_t s);
// W -> bf16 hi/lo planes [N][ldd] (ldd in bf16 elements)
cudaError_t launch_pack_split(const float* src, int N, int K, void* dst_hi, void* dst_lo, int ldd, cudaStream_t s);
cudaError_t launch_pack_bias(const float* b1, const float* b2, int N, float* dst, cudaStream_t s);
// tensor-core LSTM layer image + bias (layout.h RnnBlock::tcw/tcb); w_in/b_in = fc_in to fold (layer 0) or null
cudaError_t launch_pack_lstm_tc(const float* w_ih, const float* w_hh, const float* b_ih, const float* b_hh,
                                const float* w_in, const float* b_in, int kin, void* img, float* bias,
                                cudaStream_t s);
// mma.sync fragment image + bias for the time-axis kernel (RnnBlock::sqw/sqb), kin = 64
cudaError_t launch_pack_lstm_seq(const float* w_ih, const float* w_hh, const float* b_ih, const float* b_hh,
                                 const float* w_in, const float* b_in, void* img, float* bias, cudaStream_t s);

}  // namespace bsb
