/*
 * bsrnn_b200.h -- C ABI of the B200-native BSRNN separation path (libbsrnn_b200.so).
 *
 * This is the drop-in boundary behind the reference's Python class bsrnn.BSRNN
 * (reference bsrnn.py:328-510) and the signal wrap its callers repeat around it
 * (m_dataset.py:187-195, infer.py:29-37, infer-streaming.py:84-147).
 *
 * Conventions
 *   - every function returns 0 on success, a negative BSB_E_* code for argument errors,
 *     or a positive cudaError_t; no C++ exception crosses this boundary;
 *   - all data pointers are DEVICE pointers owned by the caller (fp32 unless stated);
 *     nothing here allocates device memory, scratch space is a caller-provided workspace
 *     sized by bsb_workspace_bytes();
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream); work is only
 *     enqueued, never synchronised (except bsb_stage_times and bsb_init);
 *   - thread-compatible: host threads may call concurrently with different workspaces / streams (no per-call state
 *     lives in process globals; the stage-timing marks are per host thread); one workspace serves one stream at a time;
 *   - multi-device: every call acts on the CURRENT device (cudaSetDevice) of the calling thread; one process may
 *     use several devices in turn;
 *   - "rows" C = dim 0 of the reference tensors (channels x clips x sessions),
 *     T = frames, frame-row = one frame of one row.
 *
 * Layouts
 *   boundary spectrogram  x[C][2050][T]   row 2k = Re bin k, 2k+1 = Im bin k, T contiguous
 *                                         (what BSRNN.forward takes, bsrnn.py:385)
 *   frame-major (internal) xf[C*T][BSB_XS] frame-row r = c*T+t, 2050 valid floats per row
 *   features               f[C*T][12][64]  (bsrnn.py:415)
 *   streaming state        s[4][2][C*12][64] (bsrnn.py:125-128,475-481)
 */
#ifndef BSRNN_B200_H
#define BSRNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BSB_N_FFT      2048
#define BSB_HOP        1024
#define BSB_N_BINS     1025
#define BSB_SPEC       2050   /* floats per frame-row at the boundary */
#define BSB_XS         2052   /* padded frame-major row stride (16-byte aligned rows) */
#define BSB_N_BANDS    12
#define BSB_FEAT       64
#define BSB_N_TENSORS  288    /* entries of BSRNN.state_dict(), reference order */

#define BSB_E_ARG      (-1)   /* bad argument (null pointer, size <= 0, ...) */
#define BSB_E_SHAPE    (-2)   /* shape the path does not accept (e.g. L <= 1024 for reflect pad) */
#define BSB_E_WORKSPACE (-3)  /* workspace too small */
#define BSB_E_ALIGN    (-4)   /* pointer not 16-byte aligned where required */
#define BSB_E_TOO_LARGE (-5)  /* request exceeds a fixed limit (bsb_stft_pack_static: BSB_STATIC_MAX_FRAME_ROWS) */
#define BSB_STATIC_MAX_FRAME_ROWS 512   /* C*T limit of bsb_stft_pack_static */

/* --- introspection ------------------------------------------------------------------ */
int         bsb_version(void);
/* One-time set-up of `device` (-1 = the current one): uploads the periodic-Hann window and FFT twiddle tables
 * (28 KB, the library's only device allocation). Call it once per device before the first signal-wrap call so
 * that no launch call allocates or synchronises (the Python host does so in BSRNN.packed_weights and the plugin in
 * its constructor); a launch on a device that was never initialised creates the tables lazily. Thread-safe. */
int         bsb_init(int device);
const char* bsb_error_string(int code);
/* band widths in bins [1,2,3,6,12,24,48,96,192,384,257,0] -- replaces generate_bandsplits()
 * (bsrnn.py:247-326), which the reference re-evaluates on every forward. */
int         bsb_band_widths(int* out12);
/* element count of state_dict entry `idx` (0..287) in reference order; -1 if out of range */
long long   bsb_tensor_numel(int idx);

/* --- weights ---------------------------------------------------------------------------
 * Replaces the per-module parameter reads of nn.Linear / nn.LSTM (bsrnn.py:69-74,333-376):
 * the 288 state_dict tensors are repacked once into one padded device buffer. */
size_t bsb_packed_weights_bytes(void);
/* src: HOST array of BSB_N_TENSORS device pointers in state_dict order (entries of
 * zero-length tensors may be NULL). packed: device buffer of bsb_packed_weights_bytes(). */
int bsb_pack_weights(const void* const* src, int n_src, void* packed, void* stream);

/* --- workspace ------------------------------------------------------------------------- */
size_t bsb_workspace_bytes(int C, int T);

/* --- signal wrap -------------------------------------------------------------------------
 * window: device float[2048] (the caller's torch.hann_window(2048)), or NULL for the
 * built-in periodic Hann. */
/* frames of torch.stft(center=True): 1 + L/1024 */
int bsb_num_frames(long long L);
/* m_dataset.py:187-190: wav[C][L] -> x[C][2050][T] (boundary layout). ws: workspace. */
int bsb_stft_pack(const float* wav, int C, long long L, const float* window,
                  float* x, void* ws, size_t ws_bytes, void* stream);
/* bsb_stft_pack without any allocation: no workspace argument, no cudaMalloc / cudaFree / allocating library call,
 * ever -- for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). This entry point has NO counterpart in the reference.
 * The frame-major scratch that bsb_stft_pack takes from the caller, and the window / twiddle tables,
 * live in file-scope `static __device__` arrays of csrc/fft.cu (4.03 MB, held by every GPU that loads the library
 * until the process exits); their device addresses are resolved once per device with cudaGetSymbolAddress.
 *   limit:   C * T <= BSB_STATIC_MAX_FRAME_ROWS (512 frame-rows), else BSB_E_TOO_LARGE;
 *   sharing: every call on a device uses the same scratch, which keeps the previous call's contents (nothing in it
 *            must start at zero: each call writes every float it later reads). Calls whose kernels OVERLAP on the
 *            GPU (different streams without an ordering between them) race on it -- serialise them.
 * Same result as bsb_stft_pack, bit for bit. wav[C][L] -> x[C][2050][T]; window as above. */
int bsb_stft_pack_static(const float* wav, int C, long long L, const float* window, float* x, void* stream);
/* same analysis, frame-major output xf[C*T][BSB_XS] (internal layout, no transpose) */
int bsb_stft_frames(const float* wav, int C, long long L, const float* window,
                    float* xf, void* stream);
/* m_dataset.py:192-195: x[C][2050][T] -> wav[C][1024*(T-1)] (torch.istft semantics) */
int bsb_unpack_istft(const float* x, int C, int T, const float* window,
                     float* wav, void* ws, size_t ws_bytes, void* stream);
/* frame-major input variant */
int bsb_istft_frames(const float* xf, int C, int T, const float* window,
                     float* wav, void* stream);
/* layout converters between boundary and frame-major */
int bsb_to_frames(const float* x, int C, int T, float* xf, void* stream);
int bsb_from_frames(const float* xf, int C, int T, float* x, void* stream);

/* 16-bit PCM at the host-buffer boundary -- what the reference's wav files hold before torchaudio.load turns them into
 * floats (infer.py:24-27): wav = pcm / 32768 (torchaudio's normalisation, exact), and back
 * pcm = clamp(rint(wav * 32768), -32768, 32767). Halves the host<->device bytes of bulk separation; fp32 stays the
 * default contract of every other entry point. n = number of samples (all rows). */
int bsb_pcm16_to_float(const int16_t* pcm, long long n, float* wav, void* stream);
int bsb_float_to_pcm16(const float* wav, long long n, int16_t* pcm, void* stream);

/* --- model ------------------------------------------------------------------------------- */
/* BSRNN.forward (bsrnn.py:385-443): x[C][2050][T] -> y[C][2050][T], y = x * mask. */
int bsb_forward(const float* x, int C, int T, const void* packed,
                float* y, void* ws, size_t ws_bytes, void* stream);
/* same on frame-major buffers (xf, yf: [C*T][BSB_XS]); yf may not alias xf */
int bsb_forward_frames(const float* xf, int C, int T, const void* packed,
                       float* yf, void* ws, size_t ws_bytes, void* stream);
/* BSRNN.forward_recurrent (bsrnn.py:445-510): x[C][2050], state[4][2][C*12][64]
 * -> y[C][2050], new_state (distinct buffer). */
int bsb_forward_recurrent(const float* x, int C, const float* state, const void* packed,
                          float* y, float* new_state, void* ws, size_t ws_bytes, void* stream);
/* waveform in -> waveform out: stft -> forward -> istft (body of m_dataset.infer,
 * m_dataset.py:182-195) without materialising the boundary layout.
 * wav[C][L] -> out[C][1024*(T-1)]. */
int bsb_separate(const float* wav, int C, long long L, const float* window, const void* packed,
                 float* out, void* ws, size_t ws_bytes, void* stream);

/* per-stage entry points (used by the parity tests; same kernels bsb_forward_frames runs) */
/* bandFCs_pre + bandFCs (bsrnn.py:404-415): xf -> residual[C*T][BSB_XS], feat[C*T][12][64] */
int bsb_bandsplit(const float* xf, int rows, const void* packed,
                  float* residual, float* feat, void* ws, size_t ws_bytes, void* stream);
/* one block of self.lstms (bsrnn.py:417): block 0/2 = BandwiseLSTM, 1/3 = TimewiseLSTM.
 * feat_in/out [C*T][12][64]; state_in/out (block 1/3 only, may be NULL = zero state / discard)
 * are [2][2][C*12][64] = (h,c) x layers. */
int bsb_rnn_block(int block, const float* feat_in, int C, int T, const void* packed,
                  const float* state_in, float* state_out,
                  float* feat_out, void* ws, size_t ws_bytes, void* stream);
/* bandFCs_back + bandFCs_back_post + residual + x*mask (bsrnn.py:420-441) */
int bsb_decode_mask(const float* feat, const float* residual, const float* xf, int rows,
                    const void* packed, float* yf, void* ws, size_t ws_bytes, void* stream);

/* --- streaming (infer-streaming.py:104-147; speech-ladspa-onnx.cpp:171-267) -------------------
 * One hop for C rows: roll the 2048-sample ring, windowed rfft, forward_recurrent,
 * irfft, 2-frame overlap-add divided by w[i]+w[i+1024].
 *   chunk[C][1024] new samples;  ring[C][2048] (in/out, zero at stream start);
 *   state / new_state [4][2][C*12][64];  tail[C][1024] (in/out: second half of the previous
 *   synthesised frame, zero at stream start);  out[C][1024]. */
int bsb_stream_step(const float* chunk, int C, const float* window, const void* packed,
                    float* ring, const float* state, float* new_state, float* tail,
                    float* out, void* ws, size_t ws_bytes, void* stream);
/* same hop with the LADSPA plugin's dry/wet control applied to the spectrum before synthesis
 * (speech-ladspa-onnx.cpp:215-226): mix = *DEVICE* float (NULL = the plain step); with X the analysed
 * spectrum of the hop and Y the model output, mix >= 0 -> mix*Y + (1-mix)*X, mix < 0 -> X + mix*Y. */
int bsb_stream_step_mix(const float* chunk, int C, const float* window, const void* packed,
                        float* ring, const float* state, float* new_state, float* tail,
                        float* out, const float* mix, void* ws, size_t ws_bytes, void* stream);
size_t bsb_stream_workspace_bytes(int C);

/* --- evaluation sums (m_dataset.train_infer, m_dataset.py:202-226) ---------------------------------
 * One pass over an (estimate, reference) pair of [rows][n] fp32 arrays (row strides in floats, >= n):
 *   stats[row][0..4] (DEVICE double, zeroed by the call) =
 *     sum|e-r|, sum r^2, sum (e-r)^2, sum e*r, sum e^2.
 * L1 (time, Re, Im), SDR and SI-SDR of train_infer are closed forms of these sums
 * (speechseparation_b200/wrap.py eval_metrics). rows <= 65535. */
int bsb_pair_stats(const float* est, long long est_stride, const float* ref, long long ref_stride,
                   int rows, long long n, double* stats, void* stream);

/* sdr2 of m_dataset.train_infer exactly as the reference reduces it for its [1][C][L] samples (m_dataset.py:219-222,
 * `dim=1` = the channel axis): out[0] (DEVICE double, zeroed by the call) =
 *   sum_l 10 log10( (sum_c src[c][l]^2 + 1e-9) / (sum_c (src[c][l] - mix[c][l])^2 + 1e-9) ); sdr2 = out[0] / n. */
int bsb_channel_sdr(const float* mix, long long mix_stride, const float* src, long long src_stride,
                    int rows, long long n, double* out, void* stream);

/* --- measurement -----------------------------------------------------------------------------
 * When enabled, the orchestration functions record CUDA events on `stream` between their
 * stages. bsb_stage_times synchronises on the last event and returns per-stage milliseconds
 * of the most recent call (return value: number of stages written, or a negative code:
 * BSB_E_ARG or -(1000 + cudaError_t)). Stage names via bsb_stage_name. */
int         bsb_set_stage_timing(int enabled);
int         bsb_stage_count(void);
const char* bsb_stage_name(int i);
int         bsb_stage_times(float* ms, int n);
/* 0 (default): Linear layers with >= 128 rows run on the tensor cores (split-bf16, 3 MMAs per
 * product, ~1e-5 of fp32); 1: force the fp32 FFMA kernel everywhere (A/B numerics checks). */
int         bsb_set_force_fp32(int on);
/* number of kernels launched by this library since load (all streams) */
long long   bsb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* BSRNN_B200_H */
