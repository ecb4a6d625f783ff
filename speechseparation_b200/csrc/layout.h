// Packed-weight layout and model constants (host side, plain C++).
//
// The reference keeps 288 separate parameter tensors (bsrnn.py:328-376); the kernels read
// one device buffer in which every Linear is stored [Npad][Kpad] row-major (the nn.Linear
// orientation: output-major, K contiguous), zero padded to Kpad = roundup(K,32),
// Npad = roundup(N,64), followed by its bias [Npad].  Every Linear is stored a second time as two
// bf16 planes (hi = bf16(W), lo = bf16(W - hi)), K padded to 64, for the tensor-core kernel.  LSTM input weights of the two
// directions are stacked along N (fwd gates 0..255, reverse 256..511) and b_ih+b_hh are
// summed at pack time; W_hh is kept raw [256][64] per direction.
#pragma once
#include <cstddef>
#include <cstdint>

namespace bsb {

constexpr int kBands = 12;        // incl. the zero-width band 11 (bsrnn.py:320)
constexpr int kRealBands = 11;
constexpr int kFeat = 64;         // band_features (bsrnn.py:60)
constexpr int kMaskHidden = 128;  // bsrnn.py:360
constexpr int kSpec = 2050;
constexpr int kXS = 2052;         // frame-major row stride
constexpr int kGates = 256;       // 4*kFeat

// band widths in bins, dims (=2w) and float offsets inside a 2050-float frame-row
constexpr int kBandW[kBands] = {1, 2, 3, 6, 12, 24, 48, 96, 192, 384, 257, 0};
constexpr int kBandD[kBands] = {2, 4, 6, 12, 24, 48, 96, 192, 384, 768, 514, 0};
constexpr int kBandOff[kBands] = {0, 2, 6, 12, 24, 48, 96, 192, 384, 768, 1536, 2050};

inline constexpr int rup(int v, int m) { return (v + m - 1) / m * m; }
inline constexpr int imax(int a, int b) { return a > b ? a : b; }

// hidden-activation row layout used between MLP layers: band i owns kHidW(i) floats
inline constexpr int hid_w(int i) { return rup(imax(kBandD[i], kMaskHidden), 4); }
inline constexpr int hid_off(int i) { int o = 0; for (int b = 0; b < i; ++b) o += hid_w(b); return o; }
constexpr int kHidStride = hid_off(kRealBands);   // 2756

struct Lin {            // one packed Linear
  long long w;          // float offset of W[Npad][Kpad] (fp32)
  long long b;          // float offset of bias[Npad]
  long long wh, wl;     // float offsets of the bf16 split planes W ~= hi + lo, each [Npad][Kpad64] bf16
  int N, K, Npad, Kpad, Kpad64;
};

struct RnnBlock {
  Lin fc_in;            // 64 -> 64
  Lin ih[2];            // per layer: K = 64 | 128 (bi layer 1), N = 256*ndir
  long long hh[2];      // per layer: float offset of W_hh [ndir][256][64]
  Lin fc;               // 64*ndir -> 64
  int ndir;
  // tensor-core LSTM layer images (lstm_tc.cu), per [layer][dir]:
  //   tcw: bf16 hi/lo planes of the concatenated [W_ih' | W_hh] matrix, rows permuted to
  //        n = 4*unit + gate and pre-scaled by -log2(e) (i,f,o) / -2*log2(e) (g), stored as the
  //        128B-swizzled shared-memory image [plane][k-chunk of 64][256 rows][64] the kernel copies
  //        verbatim. Layer 0 has fc_in folded in: W_ih' = W_ih.W_in, b' = b + W_ih.b_in.
  //   tcb: fp32 [256] bias in the same row order and scaling.
  long long tcw[2][2];
  long long tcb[2][2];
  // time-axis blocks (ndir == 1) only: per layer, the same matrix as mma.sync A fragments
  // [warp 16][k-tile 8][plane 2][lane 32][4 words] and the bias [warp 16][16 rows] (lstm_seq2.cu gathers its TMEM weight image from it)
  long long sqw[2];
  long long sqb[2];
};
constexpr long long kSeqFragFloats = 16 * 8 * 2 * 32 * 4;         // mma.sync A fragments of [W_ih' | W_hh]
constexpr long long kSeqImgFloats = kSeqFragFloats;
inline constexpr int tc_lstm_kin(int layer, int ndir) { return layer == 0 ? kFeat : kFeat * ndir; }
inline constexpr int tc_lstm_chunks(int layer, int ndir) { return (tc_lstm_kin(layer, ndir) + kFeat) / 64; }
inline constexpr long long tc_lstm_img_floats(int layer, int ndir) {   // 2 planes x chunks x 256 x 64 bf16
  return (long long)tc_lstm_chunks(layer, ndir) * kGates * 64;
}

struct Layout {
  Lin pre[kRealBands][2];
  Lin fc[kRealBands][3];
  long long fc_const;   // band 11 learned [64] feature (bsrnn.py:348)
  RnnBlock rnn[4];
  Lin back[kRealBands][3];
  Lin post[kRealBands][2];
  long long total_floats;
};

const Layout& layout();

// index of the first state_dict entry of each family, reference order (see layout.cpp)
struct SrcIndex {
  int pre[kRealBands];      // .0.weight,.0.bias,.2.weight,.2.bias
  int fc[kRealBands];       // .0.weight,.0.bias,.2.*, .4.*
  int fc_const;
  int rnn[4];               // fc_in.weight, fc_in.bias, then per layer per dir: w_ih,w_hh,b_ih,b_hh; fc.weight, fc.bias
  int back[kRealBands];
  int post[kRealBands];
  int total;
};
const SrcIndex& src_index();
long long tensor_numel(int idx);

}  // namespace bsb
