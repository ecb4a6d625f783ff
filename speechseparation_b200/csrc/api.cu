// C ABI of libbsrnn_b200.so (include/bsrnn_b200.h): argument checks, workspace carving and the
// stage-by-stage orchestration of the kernels in linear.cu / lstm.cu / fft.cu.
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../include/bsrnn_b200.h"
#include "kernels.cuh"

using namespace bsb;

namespace {

#define CK(expr)                                  \
  do {                                            \
    cudaError_t e_ = (expr);                      \
    if (e_ != cudaSuccess) return (int)e_;        \
  } while (0)
#define CKI(expr)                                 \
  do {                                            \
    int r_ = (expr);                              \
    if (r_ != 0) return r_;                       \
  } while (0)

inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---- stage timing -----------------------------------------------------------------------
// (measurement aid: the switch is process-wide, the marks belong to the calling host thread, so two threads driving
// two streams do not overwrite each other's events; bsb_stage_times reads the calling thread's marks)
constexpr int kMaxMarks = 96;
std::atomic<bool> g_timing{false};
thread_local int g_nmarks = 0;
thread_local cudaEvent_t g_ev[kMaxMarks];
thread_local bool g_ev_made[kMaxMarks] = {};
thread_local char g_names[kMaxMarks][40];

void marks_reset() { if (g_timing) g_nmarks = 0; }
// record the END of stage `name` (the first mark of a call is named "begin")
void mark(const char* name, cudaStream_t s) {
  if (!g_timing || g_nmarks >= kMaxMarks) return;
  const int i = g_nmarks;
  if (!g_ev_made[i]) {
    if (cudaEventCreate(&g_ev[i]) != cudaSuccess) return;
    g_ev_made[i] = true;
  }
  if (cudaEventRecord(g_ev[i], s) != cudaSuccess) return;
  std::snprintf(g_names[i], sizeof(g_names[i]), "%s", name);
  ++g_nmarks;
}

// ---- workspace --------------------------------------------------------------------------
struct Ws {
  float* bufA;      // [R][kXS]  band layout
  float* bufB;      // [R][kXS]
  float* residual;  // [R][kXS]
  float* hid1;      // [R][11*128]
  float* hid2;      // [R][kHidStride]
  float* featA;     // [R*12][64]
  float* featB;     // [R*12][64]
  float* t64;       // [R*12][64]
  float* xproj;     // [R*12][512]
  float* hbuf;      // [R*12][128]
  float* xf;        // [R][kXS]   (only carved when with_io)
  float* yf;        // [R][kXS]
  char* planeA;     // buffers of bf16 hi|lo plane blocks for Linear -> Linear activations (kernels.cuh LinProb::Ap)
  char* planeB;
  char* planeR;     // the residual (bandFCs_pre output) when it lives only as planes (fused forward)
  long long mtiles;     // 128-row tiles
};
// widest layer output in 32-column chunks summed over the 11 bands (bandFCs_back.2: 7 x 4 + 6 + 12 + 24 + 17 = 87)
constexpr int kPlaneChunksMax = 88;

inline size_t rup256(size_t v) { return (v + 255) / 256 * 256; }

size_t carve(Ws* w, char* base, long long R, bool with_io) {
  size_t off = 0;
  auto take = [&](size_t floats) {
    float* p = base ? reinterpret_cast<float*>(base + off) : nullptr;
    off += rup256(floats * sizeof(float));
    return p;
  };
  float* p;
  p = take((size_t)R * kXS); if (w) w->bufA = p;
  p = take((size_t)R * kXS); if (w) w->bufB = p;
  p = take((size_t)R * kXS); if (w) w->residual = p;
  p = take((size_t)R * kRealBands * kMaskHidden); if (w) w->hid1 = p;
  p = take((size_t)R * kHidStride); if (w) w->hid2 = p;
  p = take((size_t)R * kBands * kFeat); if (w) w->featA = p;
  p = take((size_t)R * kBands * kFeat); if (w) w->featB = p;
  p = take((size_t)R * kBands * kFeat); if (w) w->t64 = p;
  p = take((size_t)R * kBands * 2 * kGates); if (w) w->xproj = p;
  p = take((size_t)R * kBands * 2 * kFeat); if (w) w->hbuf = p;
  {
    const long long mt = ((R + 127) / 128 + 1) & ~1LL;   // even: a CTA pair always loads two 128-row tiles
    const size_t plane_floats = (size_t)mt * kPlaneChunksMax * kPlaneBlock / sizeof(float);
    p = take(2 * plane_floats); if (w) w->planeA = reinterpret_cast<char*>(p);
    p = take(2 * plane_floats); if (w) w->planeB = reinterpret_cast<char*>(p);
    p = take(2 * plane_floats); if (w) w->planeR = reinterpret_cast<char*>(p);
    if (w) w->mtiles = mt;
  }
  if (with_io) {
    p = take((size_t)R * kXS); if (w) w->xf = p;
    p = take((size_t)R * kXS); if (w) w->yf = p;
  } else if (w) {
    w->xf = w->yf = nullptr;
  }
  return off;
}

int get_ws(Ws* w, void* ws, size_t ws_bytes, long long R, bool with_io) {
  if (ws == nullptr) return BSB_E_ARG;
  if (!aligned16(ws)) return BSB_E_ALIGN;
  const size_t need = carve(nullptr, nullptr, R, with_io);
  if (ws_bytes < need) return BSB_E_WORKSPACE;
  carve(w, reinterpret_cast<char*>(ws), R, with_io);
  return 0;
}

inline const float* PW(const void* packed, long long off) { return reinterpret_cast<const float*>(packed) + off; }

// fill one problem of a grouped launch
void set_prob(LinProb& P, const void* packed, const Lin& l, const float* A, int lda, float* O, int ldo, int M,
              int act, const float* res = nullptr, int ldres = 0, const float* mul = nullptr, int ldmul = 0) {
  P.A = A; P.lda = lda;
  P.W = PW(packed, l.w); P.ldw = l.Kpad;
  P.Wh = l.Kpad64 ? PW(packed, l.wh) : nullptr; P.Wl = l.Kpad64 ? PW(packed, l.wl) : nullptr; P.ldw64 = l.Kpad64;
  P.bias = PW(packed, l.b);
  P.O = O; P.ldo = ldo;
  P.res = res; P.ldres = ldres;
  P.mul = mul; P.ldmul = ldmul;
  P.M = M; P.N = l.N; P.K = l.K;
  P.act = act;
  P.tile_start = 0; P.tiles_n = 0;
  P.Ap = nullptr; P.Op = nullptr; P.Rp = nullptr; P.a_lo = P.o_lo = P.r_lo = 0;
  P.o_tail_pad = 0;
}

// Linear -> Linear activations as bf16 planes (tensor-core path): band i of a layer whose bands are n[0..10]
// wide starts mtiles * (chunks of the bands before it) blocks into the buffer.
struct BandPlanes {
  char* band[kRealBands];
  long long lo;
};
template <class WidthOf>
BandPlanes band_planes(const Ws& w, char* buf, WidthOf width_of) {
  BandPlanes b{};
  long long chunks = 0;
  for (int i = 0; i < kRealBands; ++i) {
    b.band[i] = buf + w.mtiles * chunks * (2 * kPlaneBlock);     // a block = hi then lo
    chunks += plane_chunks(width_of(i));
  }
  b.lo = kPlaneBlock;
  return b;
}
inline void from_planes(LinProb& P, const BandPlanes& b, int i) { P.Ap = b.band[i]; P.a_lo = b.lo; P.A = nullptr; }
inline void to_planes(LinProb& P, const BandPlanes& b, int i, bool keep_fp32 = false) {
  P.Op = b.band[i]; P.o_lo = b.lo;
  if (!keep_fp32) P.O = nullptr;
}
inline void res_planes(LinProb& P, const BandPlanes& b, int i) { P.Rp = b.band[i]; P.r_lo = b.lo; P.res = nullptr; }
inline bool use_planes(long long rows) { return rows >= kTcMinRows && !force_ffma(); }

int single_linear(const void* packed, const Lin& l, const float* A, int lda, float* O, int ldo, long long M, int act,
                  const float* res, int ldres, cudaStream_t s, const void* a_planes = nullptr, long long a_lo = 0) {
  GroupedArgs g{};
  g.nprob = 1;
  set_prob(g.p[0], packed, l, A, lda, O, ldo, (int)M, act, res, ldres);
  if (a_planes != nullptr) { g.p[0].Ap = a_planes; g.p[0].a_lo = a_lo; g.p[0].A = nullptr; }
  CK(launch_grouped_linear(g, s));
  return 0;
}

}  // namespace

// =========================================================================================
extern "C" {

int bsb_version(void) { return 200; }

int bsb_init(int device) {
  int prev = 0;
  CK(cudaGetDevice(&prev));
  if (device >= 0 && device != prev) CK(cudaSetDevice(device));
  const cudaError_t e = init_signal_tables();
  if (device >= 0 && device != prev) cudaSetDevice(prev);
  return (int)e;
}

const char* bsb_error_string(int code) {
  switch (code) {
    case 0: return "ok";
    case BSB_E_ARG: return "bsrnn_b200: bad argument";
    case BSB_E_SHAPE: return "bsrnn_b200: unsupported shape";
    case BSB_E_WORKSPACE: return "bsrnn_b200: workspace too small";
    case BSB_E_ALIGN: return "bsrnn_b200: pointer not 16-byte aligned";
    case BSB_E_TOO_LARGE: return "bsrnn_b200: C*T exceeds the static scratch of bsb_stft_pack_static (BSB_STATIC_MAX_FRAME_ROWS = 512 frame-rows); use bsb_stft_pack with a workspace";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "bsrnn_b200: unknown error";
  }
}

int bsb_band_widths(int* out12) {
  if (!out12) return BSB_E_ARG;
  for (int i = 0; i < kBands; ++i) out12[i] = kBandW[i];
  return 0;
}

long long bsb_tensor_numel(int idx) { return tensor_numel(idx); }

size_t bsb_packed_weights_bytes(void) { return (size_t)layout().total_floats * sizeof(float); }

int bsb_pack_weights(const void* const* src, int n_src, void* packed, void* stream) {
  const SrcIndex& I = src_index();
  const Layout& L = layout();
  if (!src || !packed || n_src != I.total || I.total != BSB_N_TENSORS) return BSB_E_ARG;
  if (!aligned16(packed)) return BSB_E_ALIGN;
  for (int i = 0; i < n_src; ++i)
    if (tensor_numel(i) > 0 && src[i] == nullptr) return BSB_E_ARG;
  cudaStream_t s = S(stream);
  float* dst = reinterpret_cast<float*>(packed);
  CK(cudaMemsetAsync(packed, 0, bsb_packed_weights_bytes(), s));
  auto F = [&](int i) { return reinterpret_cast<const float*>(src[i]); };
  auto lin = [&](const Lin& l, int iw) -> int {   // iw: index of .weight, bias follows
    CK(launch_pack_matrix(F(iw), l.N, l.K, dst + l.w, l.Kpad, s));
    CK(launch_pack_split(F(iw), l.N, l.K, dst + l.wh, dst + l.wl, l.Kpad64, s));
    CK(launch_pack_bias(F(iw + 1), nullptr, l.N, dst + l.b, s));
    return 0;
  };
  for (int i = 0; i < kRealBands; ++i) {
    CKI(lin(L.pre[i][0], I.pre[i]));
    CKI(lin(L.pre[i][1], I.pre[i] + 2));
    CKI(lin(L.fc[i][0], I.fc[i]));
    CKI(lin(L.fc[i][1], I.fc[i] + 2));
    CKI(lin(L.fc[i][2], I.fc[i] + 4));
    CKI(lin(L.back[i][0], I.back[i]));
    CKI(lin(L.back[i][1], I.back[i] + 2));
    CKI(lin(L.back[i][2], I.back[i] + 4));
    CKI(lin(L.post[i][0], I.post[i]));
    CKI(lin(L.post[i][1], I.post[i] + 2));
  }
  CK(launch_pack_bias(F(I.fc_const), nullptr, kFeat, dst + L.fc_const, s));
  for (int j = 0; j < 4; ++j) {
    const RnnBlock& R = L.rnn[j];
    int k = I.rnn[j];
    CKI(lin(R.fc_in, k));
    k += 2;
    for (int layer = 0; layer < 2; ++layer) {
      const Lin& ih = R.ih[layer];
      for (int d = 0; d < R.ndir; ++d) {
        // weight_ih, weight_hh, bias_ih, bias_hh
        CK(launch_pack_matrix(F(k), kGates, ih.K, dst + ih.w + (long long)d * kGates * ih.Kpad, ih.Kpad, s));
        CK(launch_pack_split(F(k), kGates, ih.K, dst + ih.wh + (long long)d * kGates * ih.Kpad64 / 2,
                             dst + ih.wl + (long long)d * kGates * ih.Kpad64 / 2, ih.Kpad64, s));
        CK(launch_pack_matrix(F(k + 1), kGates, kFeat, dst + R.hh[layer] + (long long)d * kGates * kFeat, kFeat, s));
        CK(launch_pack_bias(F(k + 2), F(k + 3), kGates, dst + ih.b + d * kGates, s));
        CK(launch_pack_lstm_tc(F(k), F(k + 1), F(k + 2), F(k + 3), layer == 0 ? F(I.rnn[j]) : nullptr,
                               layer == 0 ? F(I.rnn[j] + 1) : nullptr, ih.K, dst + R.tcw[layer][d],
                               dst + R.tcb[layer][d], s));
        if (R.ndir == 1)
          CK(launch_pack_lstm_seq(F(k), F(k + 1), F(k + 2), F(k + 3), layer == 0 ? F(I.rnn[j]) : nullptr,
                                  layer == 0 ? F(I.rnn[j] + 1) : nullptr, dst + R.sqw[layer], dst + R.sqb[layer], s));
        k += 4;
      }
    }
    CKI(lin(R.fc, k));
  }
  return 0;
}

size_t bsb_workspace_bytes(int C, int T) {
  if (C <= 0 || T <= 0) return 0;
  return carve(nullptr, nullptr, (long long)C * T, true);
}

size_t bsb_stream_workspace_bytes(int C) {
  if (C <= 0) return 0;
  // model workspace for C frame-rows + x[C][2050] + y[C][2050]
  return carve(nullptr, nullptr, C, false) + 2 * rup256((size_t)C * kSpec * sizeof(float));
}

int bsb_num_frames(long long L) { return L <= BSB_HOP ? -1 : (int)(1 + L / BSB_HOP); }

int bsb_set_stage_timing(int enabled) { g_timing.store(enabled != 0); g_nmarks = 0; return 0; }
int bsb_stage_count(void) { return g_nmarks > 0 ? g_nmarks - 1 : 0; }
const char* bsb_stage_name(int i) { return (i >= 0 && i + 1 < g_nmarks) ? g_names[i + 1] : ""; }
int bsb_stage_times(float* ms, int n) {
  // returns the number of stages written (>= 0), or a negative code: BSB_E_ARG / -(1000 + cudaError_t)
  if (!ms) return BSB_E_ARG;
  if (g_nmarks < 2) return 0;
  cudaError_t e = cudaEventSynchronize(g_ev[g_nmarks - 1]);
  if (e != cudaSuccess) return -(1000 + (int)e);
  int cnt = g_nmarks - 1;
  if (cnt > n) cnt = n;
  for (int i = 0; i < cnt; ++i) {
    e = cudaEventElapsedTime(&ms[i], g_ev[i], g_ev[i + 1]);
    if (e != cudaSuccess) return -(1000 + (int)e);
  }
  return cnt;
}
long long bsb_launch_count(void) { return g_launches.load(); }
int bsb_set_force_fp32(int on) { set_force_ffma(on); return 0; }

}  // extern "C"

// =========================================================================================
// model stages
// =========================================================================================
namespace {

// bandFCs_pre + bandFCs (bsrnn.py:404-415)
int run_bandsplit(const float* xf, int ldx, int rows, const void* packed, float* residual, float* feat,
                  const Ws& w, cudaStream_t s) {
  const Layout& L = layout();
  GroupedArgs g{};
  const bool pm = use_planes(rows);       // hidden activations as bf16 planes instead of fp32 rows
  const BandPlanes p_pre0 = band_planes(w, w.planeA, [&](int i) { return L.pre[i][0].N; });
  // residual == nullptr (fused forward, plane mode only): the residual exists as planes in planeR, which the
  // decoder reads back; otherwise fp32 rows for the caller plus planes for the next layer
  if (residual == nullptr && !pm) return BSB_E_ARG;
  float* rrows = residual ? residual : w.residual;     // only for pointer arithmetic when the rows are not used
  const BandPlanes p_res = band_planes(w, residual ? w.planeB : w.planeR, [&](int i) { return L.pre[i][1].N; });
  const BandPlanes p_fc0 = band_planes(w, w.planeA, [&](int i) { return L.fc[i][0].N; });
  const BandPlanes p_fc1 = band_planes(w, w.planeB, [&](int i) { return L.fc[i][1].N; });
  // pre.0 : x -> bufA (LReLU)
  g.nprob = kRealBands;
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.pre[i][0], xf + kBandOff[i], ldx, w.bufA + kBandOff[i], kXS, rows, ACT_LRELU);
    if (pm) to_planes(g.p[i], p_pre0, i);
  }
  CK(launch_grouped_linear(g, s));
  // pre.2 : bufA -> residual (LReLU); the residual is needed as fp32 rows (decoder) and as the next operand
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.pre[i][1], w.bufA + kBandOff[i], kXS, rrows + kBandOff[i], kXS, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_pre0, i); to_planes(g.p[i], p_res, i, residual != nullptr); }
    if (i == kRealBands - 1) g.p[i].o_tail_pad = kXS - kSpec;          // the row's own padding follows band 10
  }
  CK(launch_grouped_linear(g, s));
  mark("pre", s);
  // fc.0 : residual -> hid2 (LReLU)
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.fc[i][0], rrows + kBandOff[i], kXS, w.hid2 + hid_off(i), kHidStride, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_res, i); to_planes(g.p[i], p_fc0, i); }
  }
  CK(launch_grouped_linear(g, s));
  // fc.2 : hid2 -> featB[:, i, :] (LReLU)
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.fc[i][1], w.hid2 + hid_off(i), kHidStride, w.featB + i * kFeat, kBands * kFeat, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_fc0, i); to_planes(g.p[i], p_fc1, i); }
  }
  CK(launch_grouped_linear(g, s));
  // fc.4 : featB -> feat (no activation); band 11 = broadcast of the learned constant (K = 0 problem)
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.fc[i][2], w.featB + i * kFeat, kBands * kFeat, feat + i * kFeat, kBands * kFeat, rows, ACT_NONE);
    if (pm) from_planes(g.p[i], p_fc1, i);
  }
  {
    LinProb& P = g.p[kRealBands];
    Lin c{};
    c.w = L.fc_const; c.b = L.fc_const; c.N = kFeat; c.K = 0; c.Npad = kFeat; c.Kpad = 32;
    set_prob(P, packed, c, feat, kBands * kFeat, feat + kRealBands * kFeat, kBands * kFeat, rows, ACT_NONE);
  }
  g.nprob = kBands;
  CK(launch_grouped_linear(g, s));
  mark("bandfc", s);
  return 0;
}

// one NormRNNResidual block (bsrnn.py:78-98) with its layout adapter (bsrnn.py:106-162)
// feat_planes (time-axis blocks only, optional): the block's output as the A-operand planes of bandFCs_back.0 (K = 64,
// 12 bands, kernels.cuh LstmSeq2Args::fc_planes) instead of fp32 rows; *planes_written tells whether the launch did so
// (only the fused output Linear can).
int run_rnn_block(int j, const float* fin, int C, int T, const void* packed, const float* st_in, float* st_out,
                  float* fout, const Ws& w, cudaStream_t s, char* feat_planes = nullptr, bool* planes_written = nullptr) {
  if (planes_written) *planes_written = false;
  const Layout& L = layout();
  const RnnBlock& R = L.rnn[j];
  const bool band = (R.ndir == 2);
  const long long rows = (long long)C * T * kBands;
  const int ldx = kGates * R.ndir, ldh = kFeat * R.ndir;
  char nm[40];
  static const long long tc_min_seq = getenv("BSB_LSTM_TC_MIN") ? atoll(getenv("BSB_LSTM_TC_MIN")) : kLstmTcMinSeq;
  if (band && (long long)C * T >= tc_min_seq && !force_ffma()) {
    // tensor-core path: fc_in and the W_ih projections are fused into the layer kernels (lstm_tc.cu)
    float* hl[2] = {w.hbuf, w.xproj};                 // layer outputs [rows][128] (xproj is free here)
    for (int layer = 0; layer < 2; ++layer) {
      LstmTcArgs a{};
      a.kin = tc_lstm_kin(layer, R.ndir);
      a.x = layer == 0 ? fin : hl[0];
      a.x_step = a.kin; a.x_seq = (long long)kBands * a.kin;
      a.hout = hl[layer];
      a.h_step = ldh; a.h_seq = (long long)kBands * ldh;
      for (int d = 0; d < R.ndir; ++d) { a.wimg[d] = PW(packed, R.tcw[layer][d]); a.bias[d] = PW(packed, R.tcb[layer][d]); }
      a.nseq = C * T; a.nsteps = kBands; a.ndir = R.ndir;
      if (layer == 1) { a.hplanes = w.planeA; a.hp_lo = kPlaneBlock; }   // the fc below takes its A operand as planes
      CK(launch_lstm_tc(a, s));
      std::snprintf(nm, sizeof(nm), "rnn%d.rec%d", j, layer);
      mark(nm, s);
    }
    CKI(single_linear(packed, R.fc, hl[1], ldh, fout, kFeat, rows, ACT_NONE, fin, kFeat, s, w.planeA, kPlaneBlock));
    std::snprintf(nm, sizeof(nm), "rnn%d.fc", j);
    mark(nm, s);
    return 0;
  }
  if (!band && T >= kLstmSeqMinSteps && !force_ffma()) {
    // latency path: both layers in one launch, fc_in and W_ih fused into the recurrence (lstm_seq2.cu)
    float* hl[2] = {w.hbuf, w.xproj};
    int fused = 0;
    {
      LstmSeq2Args a{};
      a.x = fin; a.hmid = hl[0]; a.hout = hl[1];
      a.nseq = C * kBands; a.nsteps = T; a.n_inner = kBands;
      a.outer = (long long)T * kBands * kFeat; a.inner = kFeat; a.step = (long long)kBands * kFeat;
      const long long per = (long long)a.nseq * kFeat;       // state [2(h,c)][2 layers][nseq][64]
      for (int layer = 0; layer < 2; ++layer) {
        a.wfrag[layer] = PW(packed, R.sqw[layer]);
        a.bias[layer] = PW(packed, R.sqb[layer]);
        if (st_in) { a.h0[layer] = st_in + layer * per; a.c0[layer] = st_in + (2 + layer) * per; }
        if (st_out) { a.hN[layer] = st_out + layer * per; a.cN[layer] = st_out + (2 + layer) * per; }
      }
      if (rows >= kTcMinRows) { a.hplanes = w.planeA; a.hp_lo = kPlaneBlock; }
      a.flags = reinterpret_cast<int*>(w.t64);                // t64 is only used by the FFMA fallback path
      a.fcw = PW(packed, R.fc.w); a.ldfcw = R.fc.Kpad; a.fcb = PW(packed, R.fc.b); a.fc_out = fout;
      if (feat_planes != nullptr) {
        a.fc_planes = feat_planes;
        a.fcp_lo = kPlaneBlock;
        a.fcp_band = w.mtiles * plane_chunks(kFeat) * (2LL * kPlaneBlock);
        a.fc_skip_rows = 1;
      }
      CK(launch_lstm_seq2(a, s, &fused));
      if (planes_written) *planes_written = fused != 0 && feat_planes != nullptr;
      std::snprintf(nm, sizeof(nm), "rnn%d.rec", j);
      mark(nm, s);
    }
    if (fused) return 0;                        // the launch carried fc + residual (third CTA per group)
    if (rows >= kTcMinRows)
      CKI(single_linear(packed, R.fc, hl[1], ldh, fout, kFeat, rows, ACT_NONE, fin, kFeat, s, w.planeA, kPlaneBlock));
    else
      CKI(single_linear(packed, R.fc, hl[1], ldh, fout, kFeat, rows, ACT_NONE, fin, kFeat, s));
    std::snprintf(nm, sizeof(nm), "rnn%d.fc", j);
    mark(nm, s);
    return 0;
  }
  CKI(single_linear(packed, R.fc_in, fin, kFeat, w.t64, kFeat, rows, ACT_NONE, nullptr, 0, s));
  for (int layer = 0; layer < 2; ++layer) {
    const float* in = layer == 0 ? w.t64 : w.hbuf;
    const int ldin = layer == 0 ? kFeat : ldh;
    CKI(single_linear(packed, R.ih[layer], in, ldin, w.xproj, ldx, rows, ACT_NONE, nullptr, 0, s));
    std::snprintf(nm, sizeof(nm), "rnn%d.proj%d", j, layer);
    mark(nm, s);
    LstmArgs a{};
    a.xproj = w.xproj;
    a.whh = PW(packed, R.hh[layer]);
    a.hout = w.hbuf;
    a.ndir = R.ndir;
    if (band) {                      // sequences = frame-rows, steps = the 12 bands
      a.nseq = C * T; a.nsteps = kBands; a.n_inner = 1;
      a.xp_outer = (long long)kBands * ldx; a.xp_inner = 0; a.xp_step = ldx;
      a.ho_outer = (long long)kBands * ldh; a.ho_inner = 0; a.ho_step = ldh;
    } else {                         // sequences = (row, band), steps = frames
      a.nseq = C * kBands; a.nsteps = T; a.n_inner = kBands;
      a.xp_outer = (long long)T * kBands * ldx; a.xp_inner = ldx; a.xp_step = (long long)kBands * ldx;
      a.ho_outer = (long long)T * kBands * ldh; a.ho_inner = ldh; a.ho_step = (long long)kBands * ldh;
      const long long per = (long long)a.nseq * kFeat;       // state [2(h,c)][2 layers][nseq][64]
      if (st_in) { a.h0 = st_in + layer * per; a.c0 = st_in + (2 + layer) * per; }
      if (st_out) { a.hN = st_out + layer * per; a.cN = st_out + (2 + layer) * per; }
    }
    CK(launch_lstm(a, s));
    std::snprintf(nm, sizeof(nm), "rnn%d.rec%d", j, layer);
    mark(nm, s);
  }
  CKI(single_linear(packed, R.fc, w.hbuf, ldh, fout, kFeat, rows, ACT_NONE, fin, kFeat, s));
  std::snprintf(nm, sizeof(nm), "rnn%d.fc", j);
  mark(nm, s);
  return 0;
}

// bandFCs_back + bandFCs_back_post + residual + x*mask (bsrnn.py:420-441)
int run_decode(const float* feat, const float* residual, const float* xf, int ldx, int rows, const void* packed,
               float* yf, int ldy, const Ws& w, cudaStream_t s, const char* feat_planes = nullptr) {
  const Layout& L = layout();
  GroupedArgs g{};
  const bool pm = use_planes(rows);
  const BandPlanes p_b0 = band_planes(w, w.planeA, [&](int i) { return L.back[i][0].N; });
  const BandPlanes p_b1 = band_planes(w, w.planeB, [&](int i) { return L.back[i][1].N; });
  const BandPlanes p_b2 = band_planes(w, w.planeA, [&](int i) { return L.back[i][2].N; });
  const BandPlanes p_p0 = band_planes(w, w.planeB, [&](int i) { return L.post[i][0].N; });
  const BandPlanes p_rres = band_planes(w, w.planeR, [&](int i) { return L.pre[i][1].N; });
  if (residual == nullptr && !pm) return BSB_E_ARG;
  const float* rrows = residual ? residual : w.residual;
  g.nprob = kRealBands;
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.back[i][0], feat + i * kFeat, kBands * kFeat, w.hid1 + i * kMaskHidden,
             kRealBands * kMaskHidden, rows, ACT_LRELU);
    if (pm) to_planes(g.p[i], p_b0, i);
    if (pm && feat_planes != nullptr) {       // the last time-axis block wrote its output as this layer's A planes
      g.p[i].Ap = feat_planes + (long long)i * w.mtiles * plane_chunks(kFeat) * (2LL * kPlaneBlock);
      g.p[i].a_lo = kPlaneBlock;
      g.p[i].A = nullptr;
    }
  }
  CK(launch_grouped_linear(g, s));
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.back[i][1], w.hid1 + i * kMaskHidden, kRealBands * kMaskHidden, w.hid2 + hid_off(i),
             kHidStride, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_b0, i); to_planes(g.p[i], p_b1, i); }
  }
  CK(launch_grouped_linear(g, s));
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.back[i][2], w.hid2 + hid_off(i), kHidStride, w.bufA + kBandOff[i], kXS, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_b1, i); to_planes(g.p[i], p_b2, i); }
  }
  CK(launch_grouped_linear(g, s));
  mark("back", s);
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.post[i][0], w.bufA + kBandOff[i], kXS, w.bufB + kBandOff[i], kXS, rows, ACT_LRELU);
    if (pm) { from_planes(g.p[i], p_b2, i); to_planes(g.p[i], p_p0, i); }
  }
  CK(launch_grouped_linear(g, s));
  // post.2 : y = x * (residual + W.b + bias)
  for (int i = 0; i < kRealBands; ++i) {
    set_prob(g.p[i], packed, L.post[i][1], w.bufB + kBandOff[i], kXS, yf + kBandOff[i], ldy, rows, ACT_NONE,
             rrows + kBandOff[i], kXS, xf + kBandOff[i], ldx);
    if (pm) from_planes(g.p[i], p_p0, i);
    if (residual == nullptr) res_planes(g.p[i], p_rres, i);
    if (i == kRealBands - 1 && ldy == kXS) g.p[i].o_tail_pad = kXS - kSpec;
  }
  CK(launch_grouped_linear(g, s));
  mark("post_mask", s);
  return 0;
}

int run_model(const float* xf, int ldx, int C, int T, const void* packed, const float* state, float* new_state,
              float* yf, int ldy, const Ws& w, cudaStream_t s) {
  const int rows = C * T;
  float* residual = use_planes(rows) ? nullptr : w.residual;     // plane mode: the residual never exists as fp32 rows
  CKI(run_bandsplit(xf, ldx, rows, packed, residual, w.featA, w, s));
  float* fa = w.featA;
  float* fb = w.featB;
  const char* feat_planes = nullptr;
  const long long blk = (long long)4 * C * kBands * kFeat;   // one Timewise block's (h,c) x 2 layers
  for (int j = 0; j < 4; ++j) {
    const float* si = nullptr;
    float* so = nullptr;
    if (j % 2 == 1) {
      if (state) si = state + (j / 2) * blk;
      if (new_state) so = new_state + (j / 2) * blk;
    }
    // the last block's output goes straight to bandFCs_back.0 as planes (planeB is free from here until back.1 writes it)
    const bool to_planes_j = j == 3 && use_planes(rows) && T >= kLstmSeqMinSteps;
    bool wrote = false;
    CKI(run_rnn_block(j, fa, C, T, packed, si, so, fb, w, s, to_planes_j ? w.planeB : nullptr, &wrote));
    feat_planes = wrote ? w.planeB : nullptr;
    float* t = fa; fa = fb; fb = t;
  }
  CKI(run_decode(fa, residual, xf, ldx, rows, packed, yf, ldy, w, s, feat_planes));
  return 0;
}

}  // namespace

extern "C" {

int bsb_stft_frames(const float* wav, int C, long long L, const float* window, float* xf, void* stream) {
  if (!wav || !xf || C <= 0) return BSB_E_ARG;
  const int T = bsb_num_frames(L);
  if (T < 1) return BSB_E_SHAPE;
  CK(launch_stft_frames(wav, C, L, T, window, xf, kXS, S(stream)));
  return 0;
}

int bsb_istft_frames(const float* xf, int C, int T, const float* window, float* wav, void* stream) {
  if (!xf || !wav || C <= 0) return BSB_E_ARG;
  if (T < 2) return BSB_E_SHAPE;
  CK(launch_istft_frames(xf, kXS, C, T, window, wav, S(stream)));
  return 0;
}

int bsb_to_frames(const float* x, int C, int T, float* xf, void* stream) {
  if (!x || !xf || C <= 0 || T <= 0) return BSB_E_ARG;
  CK(launch_to_frames(x, C, T, xf, kXS, S(stream)));
  return 0;
}

int bsb_from_frames(const float* xf, int C, int T, float* x, void* stream) {
  if (!x || !xf || C <= 0 || T <= 0) return BSB_E_ARG;
  CK(launch_from_frames(xf, kXS, C, T, x, S(stream)));
  return 0;
}

int bsb_stft_pack(const float* wav, int C, long long L, const float* window, float* x, void* ws, size_t ws_bytes,
                  void* stream) {
  if (!wav || !x || C <= 0) return BSB_E_ARG;
  const int T = bsb_num_frames(L);
  if (T < 1) return BSB_E_SHAPE;
  if (!ws) return BSB_E_ARG;
  if (ws_bytes < (size_t)C * T * kXS * sizeof(float)) return BSB_E_WORKSPACE;
  float* xf = reinterpret_cast<float*>(ws);
  CK(launch_stft_frames(wav, C, L, T, window, xf, kXS, S(stream)));
  CK(launch_from_frames(xf, kXS, C, T, x, S(stream)));
  return 0;
}

int bsb_unpack_istft(const float* x, int C, int T, const float* window, float* wav, void* ws, size_t ws_bytes,
                     void* stream) {
  if (!x || !wav || C <= 0) return BSB_E_ARG;
  if (T < 2) return BSB_E_SHAPE;
  if (!ws) return BSB_E_ARG;
  if (ws_bytes < (size_t)C * T * kXS * sizeof(float)) return BSB_E_WORKSPACE;
  float* xf = reinterpret_cast<float*>(ws);
  CK(launch_to_frames(x, C, T, xf, kXS, S(stream)));
  CK(launch_istft_frames(xf, kXS, C, T, window, wav, S(stream)));
  return 0;
}

int bsb_bandsplit(const float* xf, int rows, const void* packed, float* residual, float* feat, void* ws,
                  size_t ws_bytes, void* stream) {
  if (!xf || !packed || !residual || !feat || rows <= 0) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, rows, false));
  marks_reset(); mark("begin", S(stream));
  return run_bandsplit(xf, kXS, rows, packed, residual, feat, w, S(stream));
}

int bsb_rnn_block(int block, const float* feat_in, int C, int T, const void* packed, const float* state_in,
                  float* state_out, float* feat_out, void* ws, size_t ws_bytes, void* stream) {
  if (block < 0 || block > 3 || !feat_in || !feat_out || !packed || C <= 0 || T <= 0) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, (long long)C * T, false));
  marks_reset(); mark("begin", S(stream));
  return run_rnn_block(block, feat_in, C, T, packed, state_in, state_out, feat_out, w, S(stream));
}

int bsb_decode_mask(const float* feat, const float* residual, const float* xf, int rows, const void* packed,
                    float* yf, void* ws, size_t ws_bytes, void* stream) {
  if (!feat || !residual || !xf || !yf || !packed || rows <= 0) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, rows, false));
  marks_reset(); mark("begin", S(stream));
  return run_decode(feat, residual, xf, kXS, rows, packed, yf, kXS, w, S(stream));
}

int bsb_forward_frames(const float* xf, int C, int T, const void* packed, float* yf, void* ws, size_t ws_bytes,
                       void* stream) {
  if (!xf || !yf || !packed || C <= 0 || T <= 0 || xf == yf) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, (long long)C * T, false));
  marks_reset(); mark("begin", S(stream));
  return run_model(xf, kXS, C, T, packed, nullptr, nullptr, yf, kXS, w, S(stream));
}

int bsb_forward(const float* x, int C, int T, const void* packed, float* y, void* ws, size_t ws_bytes,
                void* stream) {
  if (!x || !y || !packed || C <= 0 || T <= 0) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, (long long)C * T, true));
  cudaStream_t s = S(stream);
  marks_reset(); mark("begin", s);
  CK(launch_to_frames(x, C, T, w.xf, kXS, s));
  mark("to_frames", s);
  CKI(run_model(w.xf, kXS, C, T, packed, nullptr, nullptr, w.yf, kXS, w, s));
  CK(launch_from_frames(w.yf, kXS, C, T, y, s));
  mark("from_frames", s);
  return 0;
}

int bsb_forward_recurrent(const float* x, int C, const float* state, const void* packed, float* y,
                          float* new_state, void* ws, size_t ws_bytes, void* stream) {
  if (!x || !y || !state || !new_state || !packed || C <= 0 || state == new_state) return BSB_E_ARG;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, C, false));
  marks_reset(); mark("begin", S(stream));
  // one frame per row: the frame-major layout with row stride 2050 IS the boundary layout [C][2050]
  return run_model(x, kSpec, C, 1, packed, state, new_state, y, kSpec, w, S(stream));
}

int bsb_separate(const float* wav, int C, long long L, const float* window, const void* packed, float* out,
                 void* ws, size_t ws_bytes, void* stream) {
  if (!wav || !out || !packed || C <= 0) return BSB_E_ARG;
  const int T = bsb_num_frames(L);
  if (T < 2) return BSB_E_SHAPE;
  Ws w;
  CKI(get_ws(&w, ws, ws_bytes, (long long)C * T, true));
  cudaStream_t s = S(stream);
  marks_reset(); mark("begin", s);
  CK(launch_stft_frames(wav, C, L, T, window, w.xf, kXS, s));
  mark("stft", s);
  CKI(run_model(w.xf, kXS, C, T, packed, nullptr, nullptr, w.yf, kXS, w, s));
  CK(launch_istft_frames(w.yf, kXS, C, T, window, out, s));
  mark("istft", s);
  return 0;
}

int bsb_stream_step_mix(const float* chunk, int C, const float* window, const void* packed, float* ring,
                        const float* state, float* new_state, float* tail, float* out, const float* mix,
                        void* ws, size_t ws_bytes, void* stream) {
  if (!chunk || !packed || !ring || !state || !new_state || !tail || !out || C <= 0 || state == new_state)
    return BSB_E_ARG;
  if (!ws) return BSB_E_ARG;
  if (!aligned16(ws)) return BSB_E_ALIGN;
  if (ws_bytes < bsb_stream_workspace_bytes(C)) return BSB_E_WORKSPACE;
  const size_t io = rup256((size_t)C * kSpec * sizeof(float));
  float* x = reinterpret_cast<float*>(ws);
  float* y = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + io);
  Ws w;
  CKI(get_ws(&w, reinterpret_cast<char*>(ws) + 2 * io, ws_bytes - 2 * io, C, false));
  cudaStream_t s = S(stream);
  marks_reset(); mark("begin", s);
  CK(launch_stream_analysis(chunk, C, window, ring, x, s));
  mark("analysis", s);
  CKI(run_model(x, kSpec, C, 1, packed, state, new_state, y, kSpec, w, s));
  CK(launch_stream_synthesis(y, x, mix, C, window, tail, out, s));
  mark("synthesis", s);
  return 0;
}

int bsb_stream_step(const float* chunk, int C, const float* window, const void* packed, float* ring,
                    const float* state, float* new_state, float* tail, float* out, void* ws, size_t ws_bytes,
                    void* stream) {
  return bsb_stream_step_mix(chunk, C, window, packed, ring, state, new_state, tail, out, nullptr, ws, ws_bytes,
                             stream);
}

int bsb_pair_stats(const float* est, long long est_stride, const float* ref, long long ref_stride, int rows,
                   long long n, double* stats, void* stream) {
  if (!est || !ref || !stats || rows <= 0 || rows > 65535 || n <= 0 || est_stride < n || ref_stride < n)
    return BSB_E_ARG;
  if ((reinterpret_cast<uintptr_t>(stats) & 7) != 0) return BSB_E_ALIGN;
  CK(launch_pair_stats(est, est_stride, ref, ref_stride, rows, n, stats, S(stream)));
  return 0;
}

int bsb_channel_sdr(const float* mix, long long mix_stride, const float* src, long long src_stride, int rows,
                    long long n, double* out, void* stream) {
  if (!mix || !src || !out || rows <= 0 || n <= 0 || mix_stride < n || src_stride < n) return BSB_E_ARG;
  if ((reinterpret_cast<uintptr_t>(out) & 7) != 0) return BSB_E_ALIGN;
  CK(launch_channel_sdr(mix, mix_stride, src, src_stride, rows, n, out, S(stream)));
  return 0;
}

int bsb_stft_pack_static(const float* wav, int C, long long L, const float* window, float* x, void* stream) {
  if (!wav || !x || C <= 0) return BSB_E_ARG;
  const int T = bsb_num_frames(L);
  if (T < 1) return BSB_E_SHAPE;
  if ((long long)C * T > static_max_frame_rows()) return BSB_E_TOO_LARGE;
  CK(launch_stft_pack_static(wav, C, L, T, window, x, S(stream)));
  return 0;
}

int bsb_pcm16_to_float(const int16_t* pcm, long long n, float* wav, void* stream) {
  if (!pcm || !wav || n <= 0) return BSB_E_ARG;
  CK(launch_pcm16_to_float(pcm, n, wav, S(stream)));
  return 0;
}

int bsb_float_to_pcm16(const float* wav, long long n, int16_t* pcm, void* stream) {
  if (!pcm || !wav || n <= 0) return BSB_E_ARG;
  CK(launch_float_to_pcm16(wav, n, pcm, S(stream)));
  return 0;
}

}  // extern "C"
