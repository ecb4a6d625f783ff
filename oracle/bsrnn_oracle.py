"""CPU oracle for the BSRNN separation path.  TEST INFRASTRUCTURE ONLY.

This file is the checker, not the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it.  Nothing under ``speechseparation_b200/`` imports it, and the
product path raises if its CUDA library is missing.

It restates, as plain functions over a ``state_dict`` (fp32, CPU, stock PyTorch
ops -- the reference's arithmetic *is* stock PyTorch, torch 2.11.0 here):

  * ``band_table``            <- reference bsrnn.py:247-326  (generate_bandsplits)
  * ``forward``               <- reference bsrnn.py:385-443  (BSRNN.forward)
  * ``forward_recurrent``     <- reference bsrnn.py:445-510  (BSRNN.forward_recurrent)
  * ``rnn_block``             <- reference bsrnn.py:78-98    (NormRNNResidual), with the
                                 layout adapters of bsrnn.py:106-128 (Timewise) and
                                 bsrnn.py:138-162 (Bandwise)
  * ``stft_pack``             <- reference m_dataset.py:187-190 / infer.py:29-33
  * ``unpack_istft``          <- reference m_dataset.py:192-195 / infer.py:35-37
  * ``stream``                <- reference infer-streaming.py:84-147 (file IO removed)
  * ``si_sdr`` / ``sdr``      <- reference m_dataset.py:201,214-217,224

Parity pinning: the reference ships no tests or golden vectors, so
the pin is the reference itself: ``oracle/gen_golden.py`` imports the unmodified
``/root/reference/bsrnn.py`` in the build container, runs it on seeded weights and
inputs and commits the outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks this restatement against those files (and
directly against the imported reference when ``/root/reference`` is present).

A second, loop-level restatement of the LSTM cell (``lstm_numpy``) spells out the
gate order / bias handling that ``torch.nn.LSTM`` hides (bsrnn.py:70).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------------------
# constants of the live reference configuration
# ----------------------------------------------------------------------------------------
SAMPLE_RATE = 44100
N_FFT = 2048
HOP = 1024
N_BINS = 1025
FEAT = 64                  # band_features, bsrnn.py:60
MASK_HIDDEN = 2 * FEAT     # mask_estimation_mlp_hidden, bsrnn.py:360
LRELU_SLOPE = 0.01         # nn.LeakyReLU() default, bsrnn.py:336 etc.


def band_table() -> List[int]:
    """Band widths in bins. Restates generate_bandsplits() (bsrnn.py:289-320).

    Octave bands from bin 3 upward; the last octave that would overshoot 1025 is dropped
    and replaced by the remainder, then a zero-width band is appended.
    Returns [1,2,3,6,12,24,48,96,192,384,257,0].
    """
    widths = [1, 2]
    pos = 3
    while pos < N_BINS:
        nxt = pos * 2
        if nxt == pos:
            nxt += 1
        widths.append(nxt - pos)
        pos = nxt
    widths.pop()
    if sum(widths) != N_BINS:
        widths.append(N_BINS - sum(widths))
    widths.append(0)
    return widths


BAND_WIDTHS = band_table()
BAND_DIMS = [2 * w for w in BAND_WIDTHS]           # floats per band (Re/Im interleaved)
N_BANDS = len(BAND_WIDTHS)                          # 12 (band 11 is empty)
BAND_OFFS = [2 * sum(BAND_WIDTHS[:i]) for i in range(N_BANDS)]


# ----------------------------------------------------------------------------------------
# state_dict spec + deterministic weights
# ----------------------------------------------------------------------------------------
def state_dict_spec() -> List[Tuple[str, Tuple[int, ...], float]]:
    """(key, shape, init_bound) in the reference's state_dict order (bsrnn.py:328-376).

    init_bound is the U(-b, b) bound PyTorch's default init would use (1/sqrt(fan_in)
    for Linear, 1/sqrt(hidden) for LSTM, 0 for the trainable constants).
    """
    spec: List[Tuple[str, Tuple[int, ...], float]] = []

    def lin(prefix: str, n_out: int, n_in: int):
        b = 1.0 / math.sqrt(n_in)
        spec.append((prefix + ".weight", (n_out, n_in), b))
        spec.append((prefix + ".bias", (n_out,), b))

    for i, d in enumerate(BAND_DIMS):                       # bandFCs_pre, bsrnn.py:333-340
        if d > 0:
            lin(f"bandFCs_pre.{i}.0", d, d)
            lin(f"bandFCs_pre.{i}.2", d, d)
        else:
            spec.append((f"bandFCs_pre.{i}.0.trainable_constant", (0,), 0.0))
    for i, d in enumerate(BAND_DIMS):                       # bandFCs, bsrnn.py:341-349
        if d > 0:
            h = max(d, FEAT)
            lin(f"bandFCs.{i}.0", h, d)
            lin(f"bandFCs.{i}.2", FEAT, h)
            lin(f"bandFCs.{i}.4", FEAT, FEAT)
        else:
            spec.append((f"bandFCs.{i}.0.trainable_constant", (FEAT,), 0.0))
    for j in range(4):                                      # lstms, bsrnn.py:352-356
        bi = (j % 2 == 0)                                   # Bandwise blocks are bidirectional
        p = f"lstms.{j}.m"
        lin(p + ".fc_in", FEAT, FEAT)
        b = 1.0 / math.sqrt(FEAT)
        for layer in range(2):
            n_in = FEAT if layer == 0 else (2 * FEAT if bi else FEAT)
            for suffix in ([""] + (["_reverse"] if bi else [])):
                spec.append((f"{p}.rnn.weight_ih_l{layer}{suffix}", (4 * FEAT, n_in), b))
                spec.append((f"{p}.rnn.weight_hh_l{layer}{suffix}", (4 * FEAT, FEAT), b))
                spec.append((f"{p}.rnn.bias_ih_l{layer}{suffix}", (4 * FEAT,), b))
                spec.append((f"{p}.rnn.bias_hh_l{layer}{suffix}", (4 * FEAT,), b))
        lin(p + ".fc", FEAT, 2 * FEAT if bi else FEAT)
    for i, d in enumerate(BAND_DIMS):                       # bandFCs_back, bsrnn.py:361-369
        if d > 0:
            h = max(d, MASK_HIDDEN)
            lin(f"bandFCs_back.{i}.0", MASK_HIDDEN, FEAT)
            lin(f"bandFCs_back.{i}.2", h, MASK_HIDDEN)
            lin(f"bandFCs_back.{i}.4", d, h)
        else:
            spec.append((f"bandFCs_back.{i}.0.trainable_constant", (0,), 0.0))
    for i, d in enumerate(BAND_DIMS):                       # bandFCs_back_post, bsrnn.py:370-376
        if d > 0:
            lin(f"bandFCs_back_post.{i}.0", d, d)
            lin(f"bandFCs_back_post.{i}.2", d, d)
        else:
            spec.append((f"bandFCs_back_post.{i}.0.trainable_constant", (0,), 0.0))
    return spec


def make_state_dict(seed: int = 0, const_scale: float = 0.1) -> Dict[str, torch.Tensor]:
    """Deterministic weights that do not depend on torch's RNG stream.

    MT19937 (numpy RandomState: frozen stream) uniform in the default-init bounds, so
    the GPU box regenerates bit-identical weights without shipping 30 MB fixtures.
    ``const_scale`` makes band 11's learned constant non-zero so that it is exercised.
    """
    rs = np.random.RandomState(seed)
    sd: Dict[str, torch.Tensor] = {}
    for key, shape, bound in state_dict_spec():
        n = int(np.prod(shape)) if len(shape) else 1
        if key.endswith("trainable_constant"):
            v = (rs.standard_normal(n) * const_scale).astype(np.float32)
        else:
            v = rs.uniform(-bound, bound, size=n).astype(np.float32)
        sd[key] = torch.from_numpy(v.reshape(shape).copy())
    return sd


def synth_audio(rows: int, n_samples: int, seed: int = 1234) -> Tuple[torch.Tensor, torch.Tensor]:
    """Synthetic mixture + clean source, SURVEY.md 8(d): 150 Hz 5-harmonic tone with 4 Hz AM,
    peak 0.3, plus 0.1*randn noise. Different phase/noise per row. Returns (mix, source)."""
    rs = np.random.RandomState(seed)
    t = np.arange(n_samples, dtype=np.float64) / SAMPLE_RATE
    mix = np.empty((rows, n_samples), dtype=np.float32)
    src = np.empty((rows, n_samples), dtype=np.float32)
    for r in range(rows):
        ph = rs.uniform(0, 2 * np.pi, size=5)
        f0 = 150.0 * (1.0 + 0.02 * r)
        s = sum(np.sin(2 * np.pi * f0 * (h + 1) * t + ph[h]) / (h + 1) for h in range(5))
        s = s * (0.6 + 0.4 * np.sin(2 * np.pi * 4.0 * t + ph[0]))
        s = 0.3 * s / np.max(np.abs(s))
        n = 0.1 * rs.standard_normal(n_samples)
        src[r] = s.astype(np.float32)
        mix[r] = (s + n).astype(np.float32)
    return torch.from_numpy(mix), torch.from_numpy(src)


# ----------------------------------------------------------------------------------------
# model restatement
# ----------------------------------------------------------------------------------------
def _lrelu(x: torch.Tensor) -> torch.Tensor:
    return F.leaky_relu(x, LRELU_SLOPE)


def _lin(sd, prefix: str, x: torch.Tensor) -> torch.Tensor:
    return F.linear(x, sd[prefix + ".weight"], sd[prefix + ".bias"])


def band_pre(sd, i: int, band: torch.Tensor) -> torch.Tensor:
    """bandFCs_pre[i] (bsrnn.py:333-340): Linear,LReLU,Linear,LReLU on [..., d_i]."""
    if BAND_DIMS[i] == 0:                                   # TrainableConstantModule([0])
        return band.new_zeros(band.shape[:-1] + (0,))
    y = _lrelu(_lin(sd, f"bandFCs_pre.{i}.0", band))
    return _lrelu(_lin(sd, f"bandFCs_pre.{i}.2", y))


def band_fc(sd, i: int, y: torch.Tensor) -> torch.Tensor:
    """bandFCs[i] (bsrnn.py:341-349): d->max(d,64) LReLU ->64 LReLU ->64 (no final act).
    Band 11: the learned [64] constant broadcast over the leading dims (bsrnn.py:12-24)."""
    if BAND_DIMS[i] == 0:
        c = sd[f"bandFCs.{i}.0.trainable_constant"]
        return c.expand(y.shape[:-1] + (FEAT,)).clone()
    y = _lrelu(_lin(sd, f"bandFCs.{i}.0", y))
    y = _lrelu(_lin(sd, f"bandFCs.{i}.2", y))
    return _lin(sd, f"bandFCs.{i}.4", y)


def band_back(sd, i: int, z: torch.Tensor) -> torch.Tensor:
    """bandFCs_back[i] (bsrnn.py:361-369): 64->128 LReLU ->max(d,128) LReLU ->d LReLU."""
    if BAND_DIMS[i] == 0:
        return z.new_zeros(z.shape[:-1] + (0,))
    z = _lrelu(_lin(sd, f"bandFCs_back.{i}.0", z))
    z = _lrelu(_lin(sd, f"bandFCs_back.{i}.2", z))
    return _lrelu(_lin(sd, f"bandFCs_back.{i}.4", z))


def band_post(sd, i: int, z: torch.Tensor) -> torch.Tensor:
    """bandFCs_back_post[i] (bsrnn.py:370-376): d->d LReLU d->d (no final act)."""
    if BAND_DIMS[i] == 0:
        return z.new_zeros(z.shape[:-1] + (0,))
    z = _lrelu(_lin(sd, f"bandFCs_back_post.{i}.0", z))
    return _lin(sd, f"bandFCs_back_post.{i}.2", z)


def _lstm_params(sd, j: int, bi: bool) -> List[torch.Tensor]:
    p = f"lstms.{j}.m.rnn."
    out = []
    for layer in range(2):
        for suffix in ([""] + (["_reverse"] if bi else [])):
            out += [sd[f"{p}weight_ih_l{layer}{suffix}"], sd[f"{p}weight_hh_l{layer}{suffix}"],
                    sd[f"{p}bias_ih_l{layer}{suffix}"], sd[f"{p}bias_hh_l{layer}{suffix}"]]
    return out


def rnn_block(sd, j: int, x: torch.Tensor, state: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """NormRNNResidual (bsrnn.py:78-98) for block j on x [seqs, steps, 64].

    groupnorm is None in the live model (bsrnn.py:104,136). fc_in -> 2-layer LSTM
    (bidirectional for even j) -> fc -> + x. Returns (out, (h_n, c_n)).
    Uses the same ATen LSTM kernel nn.LSTM dispatches to (torch.lstm).
    """
    bi = (j % 2 == 0)
    p = f"lstms.{j}.m"
    y = _lin(sd, p + ".fc_in", x)
    n_dir = 2 if bi else 1
    if state is None:
        z = x.new_zeros((2 * n_dir, x.shape[0], FEAT))
        state = (z, z.clone())
    out, h_n, c_n = torch.lstm(y, (state[0].contiguous(), state[1].contiguous()),
                               _lstm_params(sd, j, bi), True, 2, 0.0, False, bi, True)
    out = _lin(sd, p + ".fc", out)
    return out + x, (h_n, c_n)


def split_bands(x: torch.Tensor, dim: int) -> Sequence[torch.Tensor]:
    return x.split(BAND_DIMS, dim=dim)


def forward_stages(sd, x: torch.Tensor) -> Dict[str, torch.Tensor]:
    """BSRNN.forward (bsrnn.py:385-443) with every stage boundary kept.

    x: [C, 2050, T] fp32. Returned dict:
      residual [C,T,2050]  (cat of bandFCs_pre outputs, bsrnn.py:408-409)
      feat     [C,T,12,64] (stack of bandFCs outputs, bsrnn.py:415)
      rnn0..3  [C,T,12,64] (after each block of self.lstms, bsrnn.py:417)
      mask     [C,2050,T]  (bsrnn.py:430-432)
      y        [C,2050,T]  (x * mask, bsrnn.py:441)
    """
    C, _, T = x.shape
    res, feats = [], []
    for i, band in enumerate(split_bands(x, 1)):
        band = band.permute(0, 2, 1)                       # [C,T,d_i], bsrnn.py:406
        r = band_pre(sd, i, band)
        res.append(r)
        feats.append(band_fc(sd, i, r))
    out: Dict[str, torch.Tensor] = {}
    out["residual"] = torch.cat(res, 2)
    z = torch.stack(feats, 2)                              # [C,T,12,64]
    out["feat"] = z
    for j in range(4):
        if j % 2 == 0:                                     # BandwiseLSTM.forward, bsrnn.py:138-153
            z = rnn_block(sd, j, z.reshape(C * T, N_BANDS, FEAT))[0].reshape(C, T, N_BANDS, FEAT)
        else:                                              # TimewiseLSTM.forward, bsrnn.py:106-120
            zz = z.permute(0, 2, 1, 3).reshape(C * N_BANDS, T, FEAT)
            zz = rnn_block(sd, j, zz)[0]
            z = zz.reshape(C, N_BANDS, T, FEAT).permute(0, 2, 1, 3)
        out[f"rnn{j}"] = z.contiguous()
    masks = []
    for i in range(N_BANDS):
        b = band_back(sd, i, z[:, :, i, :])
        masks.append(res[i] + band_post(sd, i, b))         # bsrnn.py:425
    mask = torch.cat(masks, 2).permute(0, 2, 1)
    out["mask"] = mask.contiguous()
    out["y"] = x * mask
    return out


def forward(sd, x: torch.Tensor) -> torch.Tensor:
    return forward_stages(sd, x)["y"]


def forward_recurrent(sd, x: torch.Tensor, state: torch.Tensor):
    """BSRNN.forward_recurrent (bsrnn.py:445-510). x [C,2050], state [4,2,C*12,64].

    Band blocks are stateless (bsrnn.py:155-162); Timewise block k (k=0,1) reads
    h=state[2k], c=state[2k+1] and advances one step (bsrnn.py:122-128,470-481).
    """
    C = x.shape[0]
    res, feats = [], []
    for i, band in enumerate(split_bands(x, 1)):
        r = band_pre(sd, i, band)
        res.append(r)
        feats.append(band_fc(sd, i, r))
    z = torch.stack(feats, 1)                              # [C,12,64]
    new_state = []
    k = 0
    for j in range(4):
        if j % 2 == 0:
            z = rnn_block(sd, j, z)[0]                     # 12-step bi-LSTM over bands
        else:
            zz, (h_n, c_n) = rnn_block(sd, j, z.reshape(C * N_BANDS, 1, FEAT),
                                       (state[2 * k], state[2 * k + 1]))
            z = zz.reshape(C, N_BANDS, FEAT)
            new_state.append(torch.stack((h_n, c_n), 0))
            k += 1
    masks = []
    for i in range(N_BANDS):
        b = band_back(sd, i, z[:, i, :])
        masks.append(res[i] + band_post(sd, i, b))
    mask = torch.cat(masks, 1)
    return x * mask, torch.cat(new_state, 0)


# ----------------------------------------------------------------------------------------
# signal wrap restatement
# ----------------------------------------------------------------------------------------
def hann(device=None) -> torch.Tensor:
    return torch.hann_window(N_FFT, device=device)         # periodic, m_dataset.py:187


def stft_pack(wav: torch.Tensor) -> torch.Tensor:
    """m_dataset.py:187-190: [C,L] -> [C,2050,T], row 2k = Re bin k, 2k+1 = Im bin k."""
    s = torch.stft(wav, n_fft=N_FFT, hop_length=HOP, return_complex=True, window=hann(wav.device))
    s = torch.stack((s.real, s.imag), dim=2)
    return s.reshape(s.shape[0], -1, s.shape[3]).contiguous()


def unpack_istft(x: torch.Tensor) -> torch.Tensor:
    """m_dataset.py:192-195 with x.shape[0] in place of the literal 2: [C,2050,T] -> [C,1024(T-1)]."""
    x = x.reshape(x.shape[0], -1, 2, x.shape[2])
    xc = torch.complex(x[:, :, 0, :].contiguous(), x[:, :, 1, :].contiguous())
    return torch.istft(xc, n_fft=N_FFT, hop_length=HOP, window=hann(x.device))


def separate(sd, wav: torch.Tensor) -> torch.Tensor:
    """stft -> forward -> istft, the body of m_dataset.infer (m_dataset.py:182-195)."""
    return unpack_istft(forward(sd, stft_pack(wav)))


def stream(sd, wav: torch.Tensor, state: Optional[torch.Tensor] = None):
    """infer-streaming.py:84-147 without file IO. wav [C, n*1024] -> (out [C, n*1024], state).

    Rolling 2048 buffer starting at zero, analysis window only, 2-frame overlap-add divided
    by hann[:1024]+hann[1024:] (infer-streaming.py:116-145).
    """
    C, L = wav.shape
    n_hops = L // HOP
    win = hann()
    buf = torch.zeros(C, N_FFT)
    prev = torch.zeros(C, N_FFT)
    if state is None:
        state = torch.zeros(4, 2, C * N_BANDS, FEAT)
    sum_w = torch.zeros(HOP) + win[:HOP] + win[HOP:]
    outs = []
    for h in range(n_hops):
        buf = torch.cat((buf[:, HOP:], wav[:, h * HOP:(h + 1) * HOP]), 1)
        s = torch.fft.rfft(buf * win)
        x = torch.stack((s.real, s.imag), dim=2).reshape(C, -1)
        y, state = forward_recurrent(sd, x, state)
        y = y.reshape(C, -1, 2)
        cur = torch.fft.irfft(torch.complex(y[:, :, 0].contiguous(), y[:, :, 1].contiguous()))
        new = torch.zeros(C, HOP) + cur[:, :HOP] + prev[:, HOP:]
        outs.append(new / sum_w)
        prev = cur
    return torch.cat(outs, 1), state


def plugin_run(sd, left: torch.Tensor, mix: float = 1.0):
    """The LADSPA plugin's signal path, speech-ladspa-onnx.cpp:174-267, minus FFTW/onnxruntime.

    left [n] = the samples of port "Input (Left)" (the right input is read into a buffer nobody uses,
    :178,199-203). Returns (out [n], state): what the plugin writes to BOTH output ports.
      * run() hands out the result of hop k-1 while it collects the samples of hop k (:176-183),
        so the output is zero for the first 1024 samples and a trailing partial chunk runs no hop;
      * every model row carries the spectrum of input L (:199-220);
      * control port: mix >= 0 -> mix*Y + (1-mix)*X, mix < 0 -> X + mix*Y on the spectrum (:215-226);
      * row 0 of the model output only (:217-218), Hermitian completion + backward FFT / 2048
        (:229-247), 2-frame overlap-add divided by the two window values (:250-262; no synthesis window).
    """
    n = left.shape[0]
    win = hann()
    buf = torch.zeros(2, N_FFT)
    prev = torch.zeros(N_FFT)
    state = torch.zeros(4, 2, 2 * N_BANDS, FEAT)
    sum_w = win[:HOP] + win[HOP:]
    out = torch.zeros(n)
    for h in range(n // HOP):
        chunk = left[h * HOP:(h + 1) * HOP]
        buf = torch.cat((buf[:, HOP:], chunk.expand(2, -1)), 1)
        spec = torch.fft.rfft(buf[0] * win)
        x = torch.stack((spec.real, spec.imag), dim=1).reshape(1, -1).repeat(2, 1)
        y, state = forward_recurrent(sd, x, state)
        y0 = y[0].reshape(-1, 2)
        wet = torch.complex(y0[:, 0].contiguous(), y0[:, 1].contiguous())
        mixed = mix * wet + (1.0 - mix) * spec if mix >= 0.0 else spec + mix * wet
        cur = torch.fft.irfft(mixed, n=N_FFT)
        res = (cur[:HOP] + prev[HOP:]) / sum_w
        prev = cur
        lo = (h + 1) * HOP
        hi = min(lo + HOP, n)
        if hi > lo:
            out[lo:hi] = res[:hi - lo]
    return out, state


# ----------------------------------------------------------------------------------------
# metrics (m_dataset.py:201-226)
# ----------------------------------------------------------------------------------------
def si_sdr(est: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """torchmetrics ScaleInvariantSignalDistortionRatio default (zero_mean=False), per row, dB."""
    eps = torch.finfo(est.dtype).eps
    est = est.double()
    ref = ref.double()
    alpha = ((est * ref).sum(-1, keepdim=True) + eps) / ((ref * ref).sum(-1, keepdim=True) + eps)
    tgt = alpha * ref
    noise = tgt - est
    val = ((tgt * tgt).sum(-1) + eps) / ((noise * noise).sum(-1) + eps)
    return 10.0 * torch.log10(val)


def sdr(est: torch.Tensor, ref: torch.Tensor) -> torch.Tensor:
    """m_dataset.py:214-217."""
    n2s = (ref.double() ** 2).sum(1) + 1e-9
    n2d = ((est.double() - ref.double()) ** 2).sum(1) + 1e-9
    return 10.0 * torch.log10(n2s / n2d)


def train_infer(sd, waveform: torch.Tensor, speech: torch.Tensor):
    """m_dataset.train_infer with discriminator=None and lossfn=L1Loss(mean) (m_dataset.py:202-226,
    train.py:54,140), `x.shape[0]` in place of the literal 2 of m_dataset.py:192.
    waveform / speech: [1,C,L] as the reference's DataLoader yields them, or [C,L]. sdr2 sums over `dim=1` of the
    UN-squeezed sample exactly as m_dataset.py:219-222 does: the channel axis for [1,C,L], time for [C,L].
    Returns (loss, sdr, sdr2, sdr3) as python floats (float64 arithmetic for the reductions).
    Pinned by tests/test_oracle_golden.py::test_train_infer_matches_literal_restatement."""
    sample0, sample1 = waveform, speech
    if waveform.dim() == 3:
        waveform, speech = waveform.squeeze(0), speech.squeeze(0)
    spec = stft_pack(waveform)
    x = forward(sd, spec)
    x_time = unpack_istft(x)
    sp_freq = stft_pack(speech)
    sp_time = speech[:, :x_time.shape[1]]
    C = x.shape[0]
    xr, xi = x.reshape(C, -1, 2, x.shape[2])[:, :, 0].double(), x.reshape(C, -1, 2, x.shape[2])[:, :, 1].double()
    sr, si = (sp_freq.reshape(C, -1, 2, x.shape[2])[:, :, 0].double(),
              sp_freq.reshape(C, -1, 2, x.shape[2])[:, :, 1].double())
    loss = (x_time.double() - sp_time.double()).abs().mean() + (xr - sr).abs().mean() + (xi - si).abs().mean()
    n2s = sp_time.double().square().sum(1) + 1e-9
    n2d = (x_time.double() - sp_time.double()).square().sum(1) + 1e-9
    sdr_ = (10 * torch.log10(n2s / n2d)).mean()
    n2s2 = sample1.double().square().sum(1) + 1e-9
    n2d2 = (sample1.double() - sample0.double()).square().sum(1) + 1e-9
    sdr2 = (10 * torch.log10(n2s2 / n2d2)).mean()
    sdr3 = si_sdr(x_time, sp_time).mean()
    return float(loss), float(sdr_), float(sdr2), float(sdr3)


def max_rel_err(y: torch.Tensor, y_ref: torch.Tensor) -> float:
    """max|y - y_ref| / max|y_ref| per row, worst row (the parity gate's definition)."""
    y = y.double().reshape(y.shape[0], -1)
    r = y_ref.double().reshape(y_ref.shape[0], -1)
    den = r.abs().amax(1).clamp_min(1e-30)
    return float(((y - r).abs().amax(1) / den).max())


# ----------------------------------------------------------------------------------------
# loop-level LSTM restatement (numpy), spelling out what nn.LSTM does (bsrnn.py:70)
# ----------------------------------------------------------------------------------------
def lstm_numpy(x: np.ndarray, w_ih, w_hh, b_ih, b_hh, h0=None, c0=None, reverse=False):
    """Single-direction single-layer LSTM. x [S, T, I]; gates stacked i,f,g,o;
    both biases add; c' = sig(f)c + sig(i)tanh(g); h' = sig(o)tanh(c')."""
    S, T, _ = x.shape
    Hn = w_hh.shape[1]
    h = np.zeros((S, Hn), np.float64) if h0 is None else h0.astype(np.float64)
    c = np.zeros((S, Hn), np.float64) if c0 is None else c0.astype(np.float64)
    out = np.zeros((S, T, Hn), np.float64)
    sig = lambda v: 1.0 / (1.0 + np.exp(-v))
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        g = x[:, t, :].astype(np.float64) @ w_ih.T.astype(np.float64) + b_ih + \
            h @ w_hh.T.astype(np.float64) + b_hh
        i_g, f_g, g_g, o_g = g[:, :Hn], g[:, Hn:2 * Hn], g[:, 2 * Hn:3 * Hn], g[:, 3 * Hn:]
        c = sig(f_g) * c + sig(i_g) * np.tanh(g_g)
        h = sig(o_g) * np.tanh(c)
        out[:, t, :] = h
    return out, h, c


def rnn_block_numpy(sd, j: int, x: np.ndarray, state=None):
    """rnn_block restated with explicit loops in float64 (small cases only)."""
    bi = (j % 2 == 0)
    g = lambda k: sd[f"lstms.{j}.m.{k}"].numpy().astype(np.float64)
    y = x.astype(np.float64) @ g("fc_in.weight").T + g("fc_in.bias")
    h_n, c_n = [], []
    inp = y
    for layer in range(2):
        outs = []
        for d, suffix in enumerate([""] + (["_reverse"] if bi else [])):
            idx = layer * (2 if bi else 1) + d
            h0 = None if state is None else state[0][idx].numpy()
            c0 = None if state is None else state[1][idx].numpy()
            o, h, c = lstm_numpy(inp, g(f"rnn.weight_ih_l{layer}{suffix}"), g(f"rnn.weight_hh_l{layer}{suffix}"),
                                 g(f"rnn.bias_ih_l{layer}{suffix}"), g(f"rnn.bias_hh_l{layer}{suffix}"),
                                 h0, c0, reverse=(d == 1))
            outs.append(o)
            h_n.append(h)
            c_n.append(c)
        inp = np.concatenate(outs, -1)
    out = inp @ g("fc.weight").T + g("fc.bias") + x
    return out, (np.stack(h_n), np.stack(c_n))
