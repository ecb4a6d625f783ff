"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the golden fixtures
generated from the reference.  All arithmetic is fp32; gate (north_star):
max|y - y_ref| / max|y_ref| <= 1e-3 per row on waveforms, |dSI-SDR| <= 0.01 dB.
Stage tests use a tighter bound (fp32 kernels): 2e-4 relative to the stage's max."""
import contextlib
import io
import os

import numpy as np
import pytest
import torch

from oracle import bsrnn_oracle as O

pytestmark = pytest.mark.gpu
torch.set_grad_enabled(False)
STAGE_TOL = 2e-4
GATE = 1e-3
XS = 2052


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def model(dev, oracle_sd):
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        m = BSRNN()
    m.load_state_dict(oracle_sd, strict=True)
    return m.to(dev).eval()


def rel(a, b):
    a = a.detach().double().cpu().reshape(-1)
    b = b.detach().double().cpu().reshape(-1)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def to_frames(x):            # [C,2050,T] -> [C*T, XS] padded (pad filled with NaN-free junk)
    C, F, T = x.shape
    xf = torch.full((C * T, XS), 7.0, dtype=torch.float32, device=x.device)
    xf[:, :F] = x.permute(0, 2, 1).reshape(C * T, F)
    return xf


def _ws(l, C, T, dev):
    return torch.empty(l.bsb_workspace_bytes(C, T), dtype=torch.uint8, device=dev)


def _s():
    return torch.cuda.current_stream().cuda_stream


# ---------------------------------------------------------------------------------------------
def test_stft_pack_and_istft_vs_golden(dev, golden_dir):
    from speechseparation_b200 import stft_pack, unpack_istft
    g = np.load(os.path.join(golden_dir, "wrap_small.npz"))
    wav = torch.from_numpy(g["wav"]).to(dev)
    win = torch.hann_window(2048, device=dev)
    for w in (win, None):                       # caller's window and the built-in table
        x = stft_pack(wav, w)
        assert x.shape == (2, 2050, 7)
        assert rel(x, torch.from_numpy(g["packed"])) < 2e-6
        assert rel(unpack_istft(x, w), torch.from_numpy(g["istft"])) < 2e-6
        # arbitrary (non-consistent) spectrogram, non-zero DC/Nyquist imaginary parts
        assert rel(unpack_istft(torch.from_numpy(g["xr"]).to(dev), w), torch.from_numpy(g["istft_r"])) < 2e-6


@pytest.mark.parametrize("C,L", [(1, 1025), (3, 5000), (2, 44100), (5, 1024 * 9)])
def test_stft_istft_vs_oracle_and_roundtrip(dev, C, L):
    from speechseparation_b200 import stft_pack, unpack_istft
    wav, _ = O.synth_audio(C, L, seed=L)
    x = stft_pack(wav.to(dev))
    assert rel(x, O.stft_pack(wav)) < 2e-6
    T = x.shape[2]
    if T >= 2:
        back = unpack_istft(x)
        assert back.shape == (C, 1024 * (T - 1))
        assert rel(back, O.unpack_istft(O.stft_pack(wav))) < 2e-6
        assert float((back.cpu() - wav[:, :back.shape[1]]).abs().max()) < 2e-6     # perfect reconstruction


def test_layout_transposes_are_exact(dev):
    from speechseparation_b200 import _lib
    l = _lib.lib()
    C, T = 3, 37
    x = torch.randn(C, 2050, T, device=dev)
    xf = torch.zeros(C * T, XS, device=dev)
    _lib.check(l.bsb_to_frames(x.data_ptr(), C, T, xf.data_ptr(), _s()))
    assert torch.equal(xf[:, :2050].reshape(C, T, 2050).permute(0, 2, 1), x)
    x2 = torch.zeros_like(x)
    _lib.check(l.bsb_from_frames(xf.data_ptr(), C, T, x2.data_ptr(), _s()))
    assert torch.equal(x2, x)


def test_stages_vs_oracle(dev, model, oracle_sd, golden_dir):
    """Each stage is fed the ORACLE's input for that stage, so one failure does not mask another."""
    from speechseparation_b200 import _lib
    l = _lib.lib()
    g = np.load(os.path.join(golden_dir, "offline_small.npz"))
    x = torch.from_numpy(g["x"])
    C, _, T = x.shape
    st = O.forward_stages(oracle_sd, x)
    packed = model.packed_weights(dev)
    ws = _ws(l, C, T, dev)
    xf = to_frames(x.to(dev))
    R = C * T
    # band split
    residual = torch.zeros(R, XS, device=dev)
    feat = torch.zeros(R, 12, 64, device=dev)
    _lib.check(l.bsb_bandsplit(xf.data_ptr(), R, packed.data_ptr(), residual.data_ptr(), feat.data_ptr(),
                               ws.data_ptr(), ws.numel(), _s()), "bandsplit")
    assert rel(residual[:, :2050].reshape(C, T, 2050), st["residual"]) < STAGE_TOL
    assert rel(feat.reshape(C, T, 12, 64), st["feat"]) < STAGE_TOL
    assert rel(feat.reshape(C, T, 12, 64), torch.from_numpy(g["feat"])) < STAGE_TOL      # reference golden
    # rnn blocks, each from the oracle's input
    prev = st["feat"]
    for j in range(4):
        fin = prev.contiguous().to(dev)
        fout = torch.zeros_like(fin)
        _lib.check(l.bsb_rnn_block(j, fin.data_ptr(), C, T, packed.data_ptr(), None, None, fout.data_ptr(),
                                   ws.data_ptr(), ws.numel(), _s()), f"rnn{j}")
        assert rel(fout, st[f"rnn{j}"]) < STAGE_TOL, j
        assert rel(fout, torch.from_numpy(g[f"rnn{j}"])) < STAGE_TOL, j
        prev = st[f"rnn{j}"]
    # decoder + mask
    yf = torch.zeros(R, XS, device=dev)
    res_f = to_frames(st["residual"].permute(0, 2, 1).contiguous().to(dev))
    _lib.check(l.bsb_decode_mask(st["rnn3"].contiguous().to(dev).data_ptr(), res_f.data_ptr(), xf.data_ptr(), R,
                                 packed.data_ptr(), yf.data_ptr(), ws.data_ptr(), ws.numel(), _s()), "decode")
    y = yf[:, :2050].reshape(C, T, 2050).permute(0, 2, 1)
    assert rel(y, st["y"]) < STAGE_TOL
    assert rel(y, torch.from_numpy(g["y"])) < STAGE_TOL


def test_forward_vs_golden_and_input_not_mutated(dev, model, golden_dir):
    g = np.load(os.path.join(golden_dir, "offline_small.npz"))
    x = torch.from_numpy(g["x"]).to(dev)
    keep = x.clone()
    y = model.forward(x)
    assert y.shape == x.shape and y.data_ptr() != x.data_ptr() and torch.equal(x, keep)
    assert rel(y, torch.from_numpy(g["y"])) < STAGE_TOL
    assert rel(model(x), torch.from_numpy(g["y"])) < STAGE_TOL      # via __call__ too


def test_forward_recurrent_vs_golden(dev, model, golden_dir):
    g = np.load(os.path.join(golden_dir, "recurrent_small.npz"))
    s = torch.from_numpy(g["state0"]).to(dev)
    for h in range(g["x"].shape[0]):
        y, s2 = model.forward_recurrent(torch.from_numpy(g["x"][h]).to(dev), s)
        assert s2.data_ptr() != s.data_ptr()
        assert rel(y, torch.from_numpy(g["y"][h])) < STAGE_TOL
        assert float((s2.cpu() - torch.from_numpy(g["state"][h])).abs().max()) < 1e-4
        s = s2


# (C*T < 128 frame-rows: the small-M Linear kernel, incl. its 8-row chunk edges 8|9, 16, 31|32|33 and the
# hand-over to the tensor-core Linears at 127|128|129)
@pytest.mark.parametrize("C,T", [(1, 1), (2, 3), (1, 8), (3, 3), (2, 8), (1, 31), (4, 8), (11, 3), (1, 127), (1, 128), (1, 129),
                                 (5, 130), (3, 257)])
def test_forward_ragged_shapes_vs_oracle(dev, model, oracle_sd, C, T):
    rs = np.random.RandomState(100 * C + T)
    x = torch.from_numpy((rs.standard_normal((C, 2050, T)) * 2.0).astype(np.float32))
    y = model.forward(x.to(dev))
    assert rel(y, O.forward(oracle_sd, x)) < STAGE_TOL


def test_recurrent_equals_offline_and_rows_independent(dev, model):
    rs = np.random.RandomState(2)
    C, T = 3, 40
    x = torch.from_numpy(rs.standard_normal((C, 2050, T)).astype(np.float32)).to(dev)
    y = model.forward(x)
    s = torch.zeros(4, 2, C * 12, 64, device=dev)
    ys = []
    for t in range(T):
        yy, s = model.forward_recurrent(x[:, :, t].contiguous(), s)
        ys.append(yy)
    # offline runs its >=128-row GEMMs on the tensor cores (split-bf16), the hop-by-hop path in fp32
    assert rel(torch.stack(ys, 2), y) < 1e-4
    # rows are independent, bit for bit, wherever the same kernels are selected: 40 and 120 frame-rows both take the
    # small-M fp32 Linear kernel (< 128 rows; their 480 / 1,440 LSTM rows the tensor-core one), and so do 1 and 2
    # frame-rows with their 12 / 24 LSTM rows
    assert torch.equal(model.forward(x[1:2]), y[1:2])
    xs = x[:2, :, :1].contiguous()
    assert torch.equal(model.forward(xs[1:2]), model.forward(xs)[1:2])
    assert float(model.forward(torch.zeros_like(x)).abs().max()) == 0.0


def test_separate_cfg1_vs_golden_gate(dev, model, golden_dir):
    """cfg 1: [2,176400] end to end against the reference's output (decimated golden + checksums)."""
    g = np.load(os.path.join(golden_dir, "cfg1.npz"))
    wav, src = O.synth_audio(1, 176400, seed=1234)
    wav = torch.cat((wav, wav), 0)
    out = model.separate(wav.to(dev)).cpu()
    assert list(out.shape) == list(g["shape_out"])
    ref_dec = torch.from_numpy(g["sep_dec"])
    scale = float(g["sep_absmax"][0])
    assert float((out[:, ::37] - ref_dec).abs().max()) / scale < GATE
    assert abs(float((out.double() ** 2).sum()) - float(g["sep_sumsq"][0])) / float(g["sep_sumsq"][0]) < 1e-3
    sis = O.si_sdr(out, torch.cat((src, src), 0)[:, :out.shape[1]]).numpy()
    assert np.abs(sis - g["si_sdr"]).max() < 0.01
    # the three-call path (stft_pack -> forward -> unpack_istft) agrees with the fused one
    from speechseparation_b200 import infer
    x, x_time = infer(model, wav.to(dev))
    assert rel(x_time, out) < 1e-5
    assert rel(x[:, ::7, ::5], torch.from_numpy(g["y_dec"])) < STAGE_TOL


def test_streaming_vs_golden_and_graph_replay(dev, model, golden_dir):
    from speechseparation_b200 import StreamingSeparator
    g = np.load(os.path.join(golden_dir, "stream_small.npz"))
    wav = torch.from_numpy(g["wav"])
    for use_graph in (False, True):
        sep = StreamingSeparator(model, rows=2, use_graph=use_graph)
        outs = [sep.step(wav[:, h * 1024:(h + 1) * 1024].to(dev)).clone() for h in range(6)]
        assert rel(torch.cat(outs, 1), torch.from_numpy(g["out"])) < STAGE_TOL
        assert float((sep.current_state.cpu() - torch.from_numpy(g["state"])).abs().max()) < 1e-4


def test_grad_request_raises_not_silent(dev, oracle_sd):
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        m = BSRNN().to(dev)
    with torch.enable_grad():
        x = torch.randn(1, 2050, 2, device=dev)
        y = m.forward(x)
        assert y.requires_grad
        with pytest.raises(NotImplementedError):
            y.sum().backward()


def test_opt_in_library_backward_matches_oracle_autograd(dev, oracle_sd):
    """SURVEY 8(f) rank 4 is not built as kernels; the opt-in `backward_mode = "library"` recomputes the graph with
    stock PyTorch CUDA ops so that train.py can step. Forward values still come from the B200 kernels; the gradients
    must equal autograd through the CPU oracle on the same weights and input."""
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        m = BSRNN()
    m.load_state_dict(oracle_sd, strict=True)
    m = m.to(dev).train()
    m.backward_mode = "library"
    rs = np.random.RandomState(77)
    x = torch.from_numpy(rs.standard_normal((2, 2050, 9)).astype(np.float32))
    wgt = torch.from_numpy(rs.standard_normal((2, 2050, 9)).astype(np.float32))
    with torch.enable_grad():
        xd = x.to(dev).requires_grad_(True)
        y = m.forward(xd)
        (y * wgt.to(dev)).sum().backward()
        sd = {k: v.clone().requires_grad_(True) for k, v in oracle_sd.items()}
        xc = x.clone().requires_grad_(True)
        yo = O.forward(sd, xc)
        (yo * wgt).sum().backward()
    assert rel(y, yo) < STAGE_TOL                                      # forward: the kernels
    assert rel(xd.grad, xc.grad) < 2e-4
    worst = 0.0
    for (k, p), (k2, q) in zip(m.state_dict(keep_vars=True).items(), sd.items()):
        assert k == k2
        if p.numel() == 0:
            continue
        assert p.grad is not None and q.grad is not None, k
        worst = max(worst, rel(p.grad, q.grad))
    assert worst < 5e-4, worst
    # an optimizer step through the kernels' weight cache: the packed weights are rebuilt, the output changes
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
    opt.step()
    with torch.no_grad():
        y2 = m.forward(x.to(dev))
    assert not torch.equal(y2, y.detach()) and bool(torch.isfinite(y2).all())


def test_repack_after_weight_update(dev, oracle_sd):
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        m = BSRNN().to(dev).eval()
    x = torch.randn(1, 2050, 3, device=dev)
    y0 = m.forward(x)
    m.load_state_dict(oracle_sd, strict=True)
    y1 = m.forward(x)
    assert rel(y1, O.forward(oracle_sd, x.cpu())) < STAGE_TOL and not torch.equal(y0, y1)


def test_kernels_really_launch(dev, model):
    from speechseparation_b200 import _lib
    n0 = _lib.lib().bsb_launch_count()
    model.forward(torch.randn(1, 2050, 2, device=dev))
    torch.cuda.synchronize()
    assert _lib.lib().bsb_launch_count() - n0 >= 25      # 2 layout transposes + the forward's kernels


def test_tensor_core_linears_match_fp32_kernels(dev, model, oracle_sd):
    """Rows >= 128 route every Linear through the tcgen05 split-bf16 kernel; compare it with the
    fp32 FFMA kernel (forced) and with the oracle, stage by stage, on ragged row counts."""
    from speechseparation_b200 import _lib
    l = _lib.lib()
    packed = model.packed_weights(dev)
    for C, T in [(1, 128), (3, 131), (2, 300), (1, 257)]:
        rs = np.random.RandomState(C * 1000 + T)
        x = torch.from_numpy((rs.standard_normal((C, 2050, T)) * 2.0).astype(np.float32))
        R = C * T
        ws = _ws(l, C, T, dev)
        xf = to_frames(x.to(dev))
        outs = {}
        for force in (1, 0):
            l.bsb_set_force_fp32(force)
            residual = torch.zeros(R, XS, device=dev)
            feat = torch.zeros(R, 12, 64, device=dev)
            _lib.check(l.bsb_bandsplit(xf.data_ptr(), R, packed.data_ptr(), residual.data_ptr(), feat.data_ptr(),
                                       ws.data_ptr(), ws.numel(), _s()), "bandsplit")
            f1 = torch.zeros_like(feat)
            _lib.check(l.bsb_rnn_block(0, feat.data_ptr(), C, T, packed.data_ptr(), None, None, f1.data_ptr(),
                                       ws.data_ptr(), ws.numel(), _s()), "rnn0")
            yf = torch.zeros(R, XS, device=dev)
            _lib.check(l.bsb_decode_mask(f1.data_ptr(), residual.data_ptr(), xf.data_ptr(), R, packed.data_ptr(),
                                         yf.data_ptr(), ws.data_ptr(), ws.numel(), _s()), "decode")
            torch.cuda.synchronize()
            outs[force] = (residual[:, :2050].clone(), feat.clone(), f1.clone(), yf[:, :2050].clone())
        l.bsb_set_force_fp32(0)
        for name, a, b in zip(("residual", "feat", "rnn0", "y"), outs[0], outs[1]):
            for i, off in enumerate(O.BAND_OFFS[:11]) if name in ("residual", "y") else []:
                d = O.BAND_DIMS[i]
                assert rel(a[:, off:off + d], b[:, off:off + d]) < 1e-4, (name, "band", i, C, T)
            assert rel(a, b) < 5e-5, (name, C, T)
        st = O.forward_stages(oracle_sd, x)
        assert rel(outs[0][0].reshape(C, T, 2050), st["residual"]) < 5e-5
        assert rel(outs[0][1].reshape(C, T, 12, 64), st["feat"]) < 5e-5


def _run_block(l, j, fin, C, T, packed, ws, state_in=None, state_out=None):
    from speechseparation_b200 import _lib
    fout = torch.zeros_like(fin)
    _lib.check(l.bsb_rnn_block(j, fin.data_ptr(), C, T, packed.data_ptr(),
                               None if state_in is None else state_in.data_ptr(),
                               None if state_out is None else state_out.data_ptr(),
                               fout.data_ptr(), ws.data_ptr(), ws.numel(), _s()), f"rnn{j}")
    torch.cuda.synchronize()
    return fout


@pytest.mark.parametrize("C,T", [(1, 768), (3, 257), (2, 450), (7, 128), (1, 128), (3, 131), (2, 300), (7, 64)])
def test_band_axis_tensor_core_lstm_vs_oracle(dev, model, oracle_sd, C, T):
    """BandwiseLSTM blocks (bsrnn.py:138-153) against the ATen LSTM of the oracle and the fp32 FFMA kernels: on the
    tcgen05 recurrence kernel (C*T >= 768 sequences; ragged last tile at 771, 900 and 896 rows), and below that on
    tensor-core projections + the FFMA recurrence (128..600 sequences: few 128-row tiles would leave the GPU to
    2-10 CTAs walking 12 steps)."""
    from speechseparation_b200 import _lib
    l = _lib.lib()
    packed = model.packed_weights(dev)
    ws = _ws(l, C, T, dev)
    rs = np.random.RandomState(7 * C + T)
    fin = torch.from_numpy(rs.standard_normal((C * T, 12, 64)).astype(np.float32))
    for j in (0, 2):
        ref, _ = O.rnn_block(oracle_sd, j, fin)
        got = _run_block(l, j, fin.to(dev), C, T, packed, ws)
        assert rel(got, ref) < 5e-5, (j, C, T)
        l.bsb_set_force_fp32(1)
        try:
            fp32 = _run_block(l, j, fin.to(dev), C, T, packed, ws)
        finally:
            l.bsb_set_force_fp32(0)
        assert rel(got, fp32) < 5e-5, (j, C, T)
        assert not torch.equal(got, fp32)          # really a different kernel


@pytest.mark.parametrize("C,T,with_state", [(1, 2, False), (1, 3, True), (2, 37, True), (3, 130, False), (5, 64, True),
                                             (1, 15, True), (1, 16, False), (2, 17, True), (1, 33, False), (2, 100, True),
                                             (1, 4, True), (1, 5, False), (1, 9, True), (4, 8, False), (100, 9, True), (50, 21, True), (51, 6, False),
                                             (40, 9, True), (49, 6, False), (33, 5, True),
                                             (64, 431, False), (64, 431, True), (55, 200, True), (55, 200, False)])
def test_time_axis_lstm_vs_oracle_with_state(dev, model, oracle_sd, C, T, with_state):
    """TimewiseLSTM blocks (bsrnn.py:101-128) on the two-layer wavefront kernel (lstm_seq2.cu) against the ATen LSTM,
    in all four modes of its dispatch (launch_lstm_seq2; nseq = 12 C sequences, 148 SMs):
      * 8 sequences per CTA + fused output Linear (3 CTAs per group): nseq <= 392  -> C <= 32
      * 8 per CTA, separate fc GEMM (2 CTAs per group):               nseq <= 592  -> C = 33..49: (33,5) (40,9) (49,6)
      * 16 per CTA + fused output Linear:                             nseq <= 784  -> C = 50..65: (50,21) (51,6)
        and the headline shape itself, cfg 2 = (64, 431): 768 sequences x 431 steps = 108 projection blocks / flag
        publications / staging-buffer cycles, and (55, 200) = 660 sequences with a ragged last group (4 of 16)
      * 16 per CTA, separate fc, more CTAs than SMs:                   (100, 9) = 1200 sequences -> 150 CTAs
    Ragged last groups (12, 24, 36, 60 sequences), step counts around the 4-step projection blocks, initial (h, c)
    in, final (h, c) out."""
    from speechseparation_b200 import _lib
    l = _lib.lib()
    packed = model.packed_weights(dev)
    ws = _ws(l, C, T, dev)
    rs = np.random.RandomState(11 * C + T)
    fin = torch.from_numpy(rs.standard_normal((C, T, 12, 64)).astype(np.float32))
    seqs = fin.permute(0, 2, 1, 3).reshape(C * 12, T, 64)               # the reference's view (bsrnn.py:112-114)
    for j in (1, 3):
        st = None
        if with_state:
            st = torch.from_numpy((0.5 * rs.standard_normal((2, 2, C * 12, 64))).astype(np.float32))
        ref, (hn, cn) = O.rnn_block(oracle_sd, j, seqs, None if st is None else (st[0], st[1]))
        ref = ref.reshape(C, 12, T, 64).permute(0, 2, 1, 3)
        so = torch.zeros(2, 2, C * 12, 64, device=dev)
        got = _run_block(l, j, fin.reshape(C * T, 12, 64).to(dev), C, T, packed, ws,
                         None if st is None else st.to(dev), so)
        tol = 5e-5 if T <= 130 else 1e-4            # 431 recurrent steps accumulate a little more rounding
        assert rel(got.reshape(C, T, 12, 64), ref) < tol, (j, C, T)
        assert float((so[0].cpu() - hn).abs().max()) < tol and float((so[1].cpu() - cn).abs().max()) < 2 * tol


def test_lstm_saturated_gates_stay_finite(dev, model):
    """Large activations drive every gate into saturation: the ex2/rcp cell must clamp, not overflow."""
    from speechseparation_b200 import _lib
    l = _lib.lib()
    packed = model.packed_weights(dev)
    C, T = 2, 130
    ws = _ws(l, C, T, dev)
    fin = (torch.randn(C * T, 12, 64, device=dev) * 1e4)
    for j in range(4):
        out = _run_block(l, j, fin, C, T, packed, ws)
        assert bool(torch.isfinite(out).all()), j
        assert float((out - fin).abs().max()) < 1e3          # |fc(h)| is bounded since |h| <= 1


def test_full_size_properties_cfg2_slice(dev, model):
    """Size-independent properties at a bench-like shape (16 rows x 10 s): row independence across CTA
    tiles, zero in -> zero out, and silence padding does not leak backwards past the LSTM's reach."""
    C, L = 16, 441000
    wav = torch.randn(C, L, device=dev) * 0.1
    out = model.separate(wav)
    assert out.shape == (C, 1024 * (L // 1024)) and bool(torch.isfinite(out).all())
    sub = model.separate(wav[5:7].contiguous())
    assert rel(sub, out[5:7]) < 1e-5                           # other rows share tiles but not results
    assert float(model.separate(torch.zeros(2, L, device=dev)).abs().max()) == 0.0


def test_cfg2_headline_shape_vs_oracle(dev, model, oracle_sd):
    """BASELINE.json configs[1] exactly -- 64 rows x 10 s (27,584 frame-rows; 768 time-axis sequences x 431 steps in
    the 16-sequences-per-CTA fused mode; 216 band-axis tiles) -- waveform against the CPU oracle, per row:
    max|y - y_ref| / max|y_ref| <= 1e-3 and |SI-SDR(y, s) - SI-SDR(y_ref, s)| <= 0.01 dB (north_star gate), plus
    SI-SDR(y, y_ref) as the figure of how far below the gate the path sits."""
    wav, src = O.synth_audio(64, 441000, seed=1234)
    ref = O.separate(oracle_sd, wav)
    out = model.separate(wav.to(dev)).cpu()
    assert out.shape == ref.shape == (64, 1024 * 430)
    y, r = out.double(), ref.double()
    per_row = (y - r).abs().amax(1) / r.abs().amax(1)
    assert float(per_row.max()) <= GATE, per_row
    s = src[:, :out.shape[1]]
    d_sisdr = (O.si_sdr(out, s) - O.si_sdr(ref, s)).abs()
    assert float(d_sisdr.max()) <= 0.01, d_sisdr
    fidelity = O.si_sdr(out, ref)
    print(f"cfg2 vs oracle: max rel err {float(per_row.max()):.3e}, max |dSI-SDR| {float(d_sisdr.max()):.2e} dB, "
          f"min SI-SDR(y, y_ref) {float(fidelity.min()):.1f} dB")
    assert float(fidelity.min()) > 60.0
    assert float(per_row.max()) <= GATE / 20            # measured ~1e-5: a regression to plain bf16/TF32 would show


def test_streaming_1000_hops_vs_oracle(dev, model, oracle_sd):
    """cfg 3 as SURVEY.md 8(d) specifies it: 2 rows, >= 1000 hops from a zero state. Through StreamingSeparator
    (two CUDA graphs, state ping-pong from hop 2 on) and through the LADSPA-shaped C++ host, against the oracle's
    restatement of infer-streaming.py:104-147 / speech-ladspa-onnx.cpp:171-267: waveform gate on the concatenated
    hops and the carried state after the last hop (drift over ~1000 graph replays)."""
    from speechseparation_b200 import StreamingSeparator, plugin
    hops = 1000
    wav, _ = O.synth_audio(2, hops * 1024, seed=77)
    ref, ref_state = O.stream(oracle_sd, wav)
    sep = StreamingSeparator(model, rows=2)
    dwav = wav.to(dev)
    got = torch.empty(2, hops * 1024, device=dev)
    for h in range(hops):
        got[:, h * 1024:(h + 1) * 1024] = sep.step(dwav[:, h * 1024:(h + 1) * 1024])
    got = got.cpu()
    assert O.max_rel_err(got, ref) < GATE / 5
    assert O.max_rel_err(got[:, -50 * 1024:], ref[:, -50 * 1024:]) < GATE / 5      # the last hops on their own scale
    assert float((sep.current_state.cpu() - ref_state).abs().max()) < 2e-4
    assert float((O.si_sdr(got, ref)).min()) > 60.0
    # the plugin host over the same number of hops (channel 0 only, output delayed by its 1024-sample FIFO)
    p = plugin.SpeechSeparatorPlugin(oracle_sd, device=dev.index or 0)
    left = wav[0].numpy().copy()
    out = p.process(left, mix=1.0, blocks=[1024])
    pref, pstate = O.plugin_run(oracle_sd, wav[0], 1.0)
    assert O.max_rel_err(torch.from_numpy(out)[None], pref[None]) < GATE / 5
    assert float((p.state() - pstate).abs().max()) < 2e-4 and p.hops == hops
    p.close()


def test_forward_unfused_8_sequence_mode_full_model(dev, model, oracle_sd):
    """C = 40 rows (480 time-axis sequences): the dispatch branch with 8 sequences per CTA whose layer-1 output
    leaves as bf16 planes for a separate tensor-core fc GEMM -- full forward against the oracle."""
    rs = np.random.RandomState(4040)
    x = torch.from_numpy((rs.standard_normal((40, 2050, 7)) * 2.0).astype(np.float32))
    assert rel(model.forward(x.to(dev)), O.forward(oracle_sd, x)) < STAGE_TOL


def test_two_devices_in_one_process(oracle_sd):
    """The dynamic-shared-memory opt-in and the signal tables are per device: a model on cuda:0 and then one on
    cuda:1 in the same process must both run (process-wide `static bool attr_set`)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from speechseparation_b200 import BSRNN
    wav, _ = O.synth_audio(2, 300 * 1024, seed=5)
    ref = O.separate(oracle_sd, wav)
    for d in (0, 1, 0):
        device = torch.device("cuda", d)
        with contextlib.redirect_stdout(io.StringIO()):
            m = BSRNN()
        m.load_state_dict(oracle_sd, strict=True)
        m = m.to(device).eval()
        out = m.separate(wav.to(device))
        assert O.max_rel_err(out.cpu(), ref) < GATE / 5, d
        y, _ = m.forward_recurrent(torch.randn(2, 2050, device=device), torch.zeros(4, 2, 24, 64, device=device))
        assert bool(torch.isfinite(y).all())


def test_streams_share_packed_weights_safely(dev, model, oracle_sd):
    """Weights packed on one stream and used at once on another: the consumer stream must wait for
    the packing kernels, and two streams must not share one workspace."""
    from speechseparation_b200 import BSRNN
    with contextlib.redirect_stdout(io.StringIO()):
        m = BSRNN()
    m.load_state_dict(oracle_sd, strict=True)
    m = m.to(dev).eval()
    wav, _ = O.synth_audio(3, 200 * 1024, seed=8)
    dwav = wav.to(dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    torch.cuda.synchronize()
    with torch.cuda.stream(s1):
        m.packed_weights(dev)                   # packing kernels queued on s1 ...
    with torch.cuda.stream(s2):
        a = m.separate(dwav)                    # ... consumed on s2 without a host sync in between
    with torch.cuda.stream(s1):
        b = m.separate(dwav)                    # concurrently on s1: its own workspace
    torch.cuda.synchronize()
    ref = O.separate(oracle_sd, wav)
    assert O.max_rel_err(a.cpu(), ref) < GATE / 5 and torch.equal(a, b)
    assert len(m._ws) == 2


def test_batch_separator_pipeline_equals_separate(dev, model):
    """Host-buffer pipeline (H2D | kernels | D2H on three streams): every batch equals model.separate."""
    from speechseparation_b200 import BatchSeparator
    C, L = 3, 20000
    pipe = BatchSeparator(model, C, L, device=dev)
    ins = [torch.randn(C, L).mul_(0.1).pin_memory() for _ in range(5)]
    outs = [torch.empty(C, pipe.out_samples).pin_memory() for _ in range(5)]
    tickets = [pipe.submit(a, b) for a, b in zip(ins, outs)]
    for a, t in zip(ins, tickets):
        got = t.wait()
        assert torch.equal(got, model.separate(a.to(dev)).cpu())
    with pytest.raises(RuntimeError):
        pipe.submit(torch.randn(C, L), outs[0])              # not pinned
    with pytest.raises(RuntimeError):
        pipe.submit(ins[0][:, :-1].contiguous().pin_memory(), outs[0])


def test_batch_separator_pcm16_mode(dev, model, oracle_sd):
    """int16 PCM in and out (half the host<->device bytes): the input is exactly what torchaudio.load would hand
    the reference for a 16-bit wav (s / 32768), the output is the float result rounded to 16 bits."""
    from speechseparation_b200 import BatchSeparator
    C, L = 3, 30000
    pipe = BatchSeparator(model, C, L, device=dev, pcm16=True)
    wav, _ = O.synth_audio(C, L, seed=12)
    pcm = torch.clamp(torch.round(wav * 32768.0), -32768, 32767).to(torch.int16)
    pcm[0, 0], pcm[0, 1] = -32768, 32767                                   # full-scale samples
    outs = [torch.empty(C, pipe.out_samples, dtype=torch.int16).pin_memory() for _ in range(3)]
    tickets = [pipe.submit(pcm.pin_memory(), o) for o in outs]
    got = [t.wait().clone() for t in tickets]
    assert torch.equal(got[0], got[1]) and torch.equal(got[1], got[2])
    x = pcm.float() / 32768.0                                              # torchaudio.load's normalisation
    ref = O.separate(oracle_sd, x)
    want = torch.clamp(torch.round(model.separate(x.to(dev)).cpu() * 32768.0), -32768, 32767).to(torch.int16)
    assert torch.equal(got[0], want)                                       # the conversion kernels, bit for bit
    # against the oracle: half an LSB of 16-bit rounding (2^-16 of full scale) plus the path's own error. (Relative
    # to max|y_ref| that can be large: a random-init model's output is far below full scale.)
    # Samples beyond 16-bit full scale clip, as they would in a PCM wav file.
    clipped = ref.clamp(-1.0, 32767.0 / 32768.0)
    err = (got[0].float() / 32768.0 - clipped).abs().max()
    assert float(err) <= 0.5 / 32768.0 + GATE * float(ref.abs().max())
    assert int((got[0] == 32767).sum()) > 0 or float(ref.max()) < 1.0      # this input does clip
    with pytest.raises(RuntimeError):
        pipe.submit(wav.pin_memory(), outs[0])                             # float tensor into a pcm16 pipe


def test_ladspa_plugin_host_vs_golden_and_oracle(dev, oracle_sd, golden_dir, tmp_path):
    """The LADSPA-shaped C++ host (csrc/plugin.cpp) against the reference plugin loop (golden) with ragged
    run() block sizes, three control-port values, graph replay after the second hop, reset, and the
    descriptor path a LADSPA host takes (instantiate from BSRNN_B200_WEIGHTS)."""
    from speechseparation_b200 import plugin
    g = np.load(os.path.join(golden_dir, "plugin_small.npz"))
    left = g["left"]
    blocks = [int(b) for b in g["blocks"]]
    p = plugin.SpeechSeparatorPlugin(oracle_sd, device=dev.index or 0)
    for tag in ("wet", "mix04", "neg05"):
        p.reset()
        out = p.process(left, mix=float(g["mix_" + tag][0]), blocks=blocks)
        assert rel(torch.from_numpy(out)[None], torch.from_numpy(g["out_" + tag])[None]) < STAGE_TOL, tag
        assert float((p.state() - torch.from_numpy(g["state_" + tag])).abs().max()) < 1e-4
    assert p.hops == 18 and 0.0 < p.last_hop_ms <= p.max_hop_ms
    # block-size independence: one sample at a time == one big block (bit-exact, same kernels)
    p.reset()
    a = p.process(left[:3 * 1024 + 5], mix=1.0, blocks=[1])
    p.reset()
    b = p.process(left[:3 * 1024 + 5], mix=1.0, blocks=[4096])
    assert np.array_equal(a, b)
    p.close()
    # through ladspa_descriptor(0), as an audio host would
    wpath = str(tmp_path / "weights.f32")
    plugin.export_weights(oracle_sd, wpath)
    os.environ["BSRNN_B200_WEIGHTS"] = wpath
    os.environ["BSRNN_B200_DEVICE"] = str(dev.index or 0)
    d = plugin.ladspa_descriptor(0)
    h = d.instantiate(None, 44100)
    assert h
    n = 4 * 1024
    ctl = np.array([1.0], dtype=np.float32)
    inp = np.ascontiguousarray(left[:n])
    o_l, o_r = np.zeros(n, np.float32), np.zeros(n, np.float32)
    for port, buf in enumerate((ctl, inp, inp, o_l, o_r)):
        d.connect_port(h, port, buf.ctypes.data)
    d.run(h, n)
    d.cleanup(h)
    ref, _ = O.plugin_run(oracle_sd, torch.from_numpy(inp), 1.0)
    assert rel(torch.from_numpy(o_l)[None], ref[None]) < STAGE_TOL and np.array_equal(o_l, o_r)
    del os.environ["BSRNN_B200_WEIGHTS"]
    assert not d.instantiate(None, 44100)          # no weights configured: NULL handle, message on stderr


@pytest.mark.parametrize("rows,n,off", [(1, 1, 0), (2, 4099, 0), (3, 50001, 1), (5, 16384 * 2, 3), (2, 441000, 0)])
def test_pair_stats_vs_float64(dev, rows, n, off):
    """bsb_pair_stats against float64 torch sums: ragged lengths, unaligned row starts (strided views), n = 1."""
    from speechseparation_b200 import pair_stats
    g = torch.Generator().manual_seed(rows * 131 + n)
    eb = torch.randn(rows, n + 7, generator=g)
    rb = torch.randn(rows, n + 7, generator=g) * 0.5 + 0.3 * eb
    e, r = eb[:, off:off + n], rb[:, off:off + n]
    st = pair_stats(e.to(dev)[:, :], r.to(dev)[:, :]).cpu()
    st2 = pair_stats(eb.to(dev)[:, off:off + n], rb.to(dev)[:, off:off + n]).cpu()      # non-contiguous rows, odd offset
    ed, rd = e.double(), r.double()
    want = torch.stack(((ed - rd).abs().sum(1), (rd * rd).sum(1), ((ed - rd) ** 2).sum(1), (ed * rd).sum(1),
                        (ed * ed).sum(1)), 1)
    scale = torch.stack(((ed - rd).abs().sum(1), (rd * rd).sum(1), ((ed - rd) ** 2).sum(1), (ed * rd).abs().sum(1),
                         (ed * ed).sum(1)), 1).clamp_min(1e-30)
    assert float(((st - want).abs() / scale).max()) < 2e-6
    assert float(((st2 - want).abs() / scale).max()) < 2e-6


def test_train_infer_metrics_vs_oracle(dev, model, oracle_sd):
    """Forward-only m_dataset.train_infer (loss, sdr, sdr2, si-sdr) with every reduction on the device."""
    from speechseparation_b200 import train_infer
    wav, src = O.synth_audio(2, 30000, seed=21)
    # [1, C, L] as the reference's DataLoader yields it (sdr2 reduces over channels, m_dataset.py:219-222) and [C, L]
    for batched in (True, False):
        sample = (wav[None], src[None]) if batched else (wav, src)
        want = O.train_infer(oracle_sd, *sample)
        got = [float(v) for v in train_infer(model, tuple(t.to(dev) for t in sample))]
        assert abs(got[0] - want[0]) < 1e-4 * abs(want[0])
        for a, b in zip(got[1:], want[1:]):
            assert abs(a - b) < 0.01, (batched, got, want)            # dB
    with pytest.raises(RuntimeError):
        train_infer(model, (wav, src))                   # host tensors: no CPU path


def test_many_session_streaming_cfg4_shape(dev, model, oracle_sd):
    """cfg 4 (hop-synchronous serving of many sessions): 65 sessions x 2 rows = 130 rows per hop, enough rows for
    the tensor-core Linear / band-LSTM kernels inside forward_recurrent; carried state, graph replay from hop 2."""
    from speechseparation_b200 import StreamingSeparator
    rows, hops = 130, 4
    wav, _ = O.synth_audio(rows, hops * 1024, seed=31)
    ref, ref_state = O.stream(oracle_sd, wav)
    sep = StreamingSeparator(model, rows=rows)
    outs = [sep.step(wav[:, h * 1024:(h + 1) * 1024].to(dev)).clone() for h in range(hops)]
    got = torch.cat(outs, 1).cpu()
    for r in range(0, rows, 13):
        assert rel(got[r], ref[r]) < STAGE_TOL, r
    assert rel(got, ref) < STAGE_TOL
    assert float((sep.current_state.cpu() - ref_state).abs().max()) < 1e-4


def test_long_form_cfg5_shape(dev, model, oracle_sd):
    """cfg 5 shape (one long clip per GPU): 1 row x 60 s = 2,584 frames; 12 time-axis sequences -> ONE sequence
    group walks 2,584 steps (646 projection blocks, flag publications and buffer cycles), the band-axis kernels see
    2,584 x 12-step sequences. Waveform gate against the oracle."""
    wav, _ = O.synth_audio(1, 60 * 44100, seed=41)
    ref = O.separate(oracle_sd, wav)
    out = model.separate(wav.to(dev))
    assert out.shape == ref.shape == (1, 1024 * 2583)
    assert O.max_rel_err(out.cpu(), ref) < GATE / 5


def test_traced_streaming_step_records_the_operator_and_replays(dev, model, golden_dir, tmp_path):
    """the reference exports by tracing ToTrace(model) (infer-streaming.py:20-28,57-74). The trace
    must hold the step as ONE recordable operator, survive torch.jit.save/load, and reproduce the eager C-ABI call
    bit for bit on the golden hops (state carried by the caller)."""
    from speechseparation_b200 import ToTrace, trace_streaming_step
    traced = trace_streaming_step(model, rows=2, device=dev)
    kinds = [n.kind() for n in traced.graph.nodes()]
    assert kinds.count("bsrnn_b200::forward_recurrent") == 1, kinds
    path = os.path.join(tmp_path, "model-script.pth")
    torch.jit.save(traced, path)
    loaded = torch.jit.load(path, map_location=dev)
    g = np.load(os.path.join(golden_dir, "recurrent_small.npz"))
    s_e = s_t = s_l = torch.from_numpy(g["state0"]).to(dev)
    for h in range(g["x"].shape[0]):
        x = torch.from_numpy(g["x"][h]).to(dev)
        y_e, s_e = model.forward_recurrent(x, s_e)
        y_t, s_t, empty = traced(x, s_t)
        y_l, s_l, _ = loaded(x, s_l)
        assert empty.numel() == 0                                     # the reference's third, empty output
        assert torch.equal(y_t, y_e) and torch.equal(s_t, s_e)
        assert torch.equal(y_l, y_e) and torch.equal(s_l, s_e)
        assert rel(y_t, torch.from_numpy(g["y"][h])) < STAGE_TOL
    # the eager wrapper behaves like the reference's ToTrace outside a trace too
    a, b, c = ToTrace(model)(torch.from_numpy(g["x"][0]).to(dev), torch.from_numpy(g["state0"]).to(dev))
    assert a.shape == (2, 2050) and b.shape == (4, 2, 24, 64) and c.numel() == 0


def test_traced_offline_forward_replays(dev, model, golden_dir):
    g = np.load(os.path.join(golden_dir, "offline_small.npz"))
    x = torch.from_numpy(g["x"]).to(dev)
    with torch.no_grad():
        traced = torch.jit.trace(model, (x,), check_trace=False)
    assert [n.kind() for n in traced.graph.nodes()].count("bsrnn_b200::forward") == 1
    assert torch.equal(traced(x), model.forward(x))
    assert rel(traced(x), torch.from_numpy(g["y"])) < STAGE_TOL
