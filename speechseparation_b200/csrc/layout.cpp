// Packed-weight layout tables. See layout.h.
#include "layout.h"

namespace bsb {

namespace {

struct Builder {
  long long cur = 0;
  long long take(long long n) {
    long long o = cur;
    cur += (n + 63) / 64 * 64;  // every block starts 256-byte aligned
    return o;
  }
  Lin lin(int N, int K) {
    Lin l;
    l.N = N;
    l.K = K;
    l.Npad = rup(N, 64);
    l.Kpad = rup(K, 32);
    l.Kpad64 = rup(K, 64);
    l.w = take((long long)l.Npad * l.Kpad);
    l.b = take(l.Npad);
    l.wh = take((long long)l.Npad * l.Kpad64 / 2);   // bf16 plane = half as many floats
    l.wl = take((long long)l.Npad * l.Kpad64 / 2);
    return l;
  }
};

Layout build_layout() {
  Layout L{};
  Builder B;
  for (int i = 0; i < kRealBands; ++i) {       // bandFCs_pre, bsrnn.py:333-340
    int d = kBandD[i];
    L.pre[i][0] = B.lin(d, d);
    L.pre[i][1] = B.lin(d, d);
  }
  for (int i = 0; i < kRealBands; ++i) {       // bandFCs, bsrnn.py:341-349
    int d = kBandD[i], h = imax(d, kFeat);
    L.fc[i][0] = B.lin(h, d);
    L.fc[i][1] = B.lin(kFeat, h);
    L.fc[i][2] = B.lin(kFeat, kFeat);
  }
  L.fc_const = B.take(kFeat);
  for (int j = 0; j < 4; ++j) {                // lstms, bsrnn.py:352-356 / 63-74
    RnnBlock& R = L.rnn[j];
    R.ndir = (j % 2 == 0) ? 2 : 1;
    R.fc_in = B.lin(kFeat, kFeat);
    for (int layer = 0; layer < 2; ++layer) {
      int K = layer == 0 ? kFeat : kFeat * R.ndir;
      R.ih[layer] = B.lin(kGates * R.ndir, K);
      R.hh[layer] = B.take((long long)R.ndir * kGates * kFeat);
    }
    R.fc = B.lin(kFeat, kFeat * R.ndir);
    for (int layer = 0; layer < 2; ++layer)
      for (int d = 0; d < R.ndir; ++d) {
        R.tcw[layer][d] = B.take(tc_lstm_img_floats(layer, R.ndir));
        R.tcb[layer][d] = B.take(kGates);
      }
    for (int layer = 0; layer < 2; ++layer) {
      R.sqw[layer] = R.ndir == 1 ? B.take(kSeqImgFloats) : -1;
      R.sqb[layer] = R.ndir == 1 ? B.take(kGates) : -1;
    }
  }
  for (int i = 0; i < kRealBands; ++i) {       // bandFCs_back, bsrnn.py:361-369
    int d = kBandD[i], h = imax(d, kMaskHidden);
    L.back[i][0] = B.lin(kMaskHidden, kFeat);
    L.back[i][1] = B.lin(h, kMaskHidden);
    L.back[i][2] = B.lin(d, h);
  }
  for (int i = 0; i < kRealBands; ++i) {       // bandFCs_back_post, bsrnn.py:370-376
    int d = kBandD[i];
    L.post[i][0] = B.lin(d, d);
    L.post[i][1] = B.lin(d, d);
  }
  L.total_floats = B.cur;
  return L;
}

// state_dict enumeration in the reference's registration order (bsrnn.py:333-376).
struct SrcTables {
  SrcIndex idx;
  long long numel[288];
};

SrcTables build_src() {
  SrcTables S{};
  int n = 0;
  auto lin = [&](int N, int K) {
    S.numel[n++] = (long long)N * K;
    S.numel[n++] = N;
  };
  for (int i = 0; i < kBands; ++i) {
    int d = kBandD[i];
    if (d > 0) { S.idx.pre[i] = n; lin(d, d); lin(d, d); }
    else S.numel[n++] = 0;                      // bandFCs_pre.11.0.trainable_constant [0]
  }
  for (int i = 0; i < kBands; ++i) {
    int d = kBandD[i], h = imax(d, kFeat);
    if (d > 0) { S.idx.fc[i] = n; lin(h, d); lin(kFeat, h); lin(kFeat, kFeat); }
    else { S.idx.fc_const = n; S.numel[n++] = kFeat; }
  }
  for (int j = 0; j < 4; ++j) {
    int ndir = (j % 2 == 0) ? 2 : 1;
    S.idx.rnn[j] = n;
    lin(kFeat, kFeat);
    for (int layer = 0; layer < 2; ++layer) {
      int K = layer == 0 ? kFeat : kFeat * ndir;
      for (int dir = 0; dir < ndir; ++dir) {
        S.numel[n++] = (long long)kGates * K;      // weight_ih
        S.numel[n++] = (long long)kGates * kFeat;  // weight_hh
        S.numel[n++] = kGates;                     // bias_ih
        S.numel[n++] = kGates;                     // bias_hh
      }
    }
    lin(kFeat, kFeat * ndir);
  }
  for (int i = 0; i < kBands; ++i) {
    int d = kBandD[i], h = imax(d, kMaskHidden);
    if (d > 0) { S.idx.back[i] = n; lin(kMaskHidden, kFeat); lin(h, kMaskHidden); lin(d, h); }
    else S.numel[n++] = 0;
  }
  for (int i = 0; i < kBands; ++i) {
    int d = kBandD[i];
    if (d > 0) { S.idx.post[i] = n; lin(d, d); lin(d, d); }
    else S.numel[n++] = 0;
  }
  S.idx.total = n;
  return S;
}

const SrcTables& src_tables() {
  static const SrcTables S = build_src();
  return S;
}

}  // namespace

const Layout& layout() {
  static const Layout L = build_layout();
  return L;
}

const SrcIndex& src_index() { return src_tables().idx; }

long long tensor_numel(int idx) {
  const SrcTables& S = src_tables();
  if (idx < 0 || idx >= S.idx.total) return -1;
  return S.numel[idx];
}

}  // namespace bsb
