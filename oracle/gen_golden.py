"""Generate tests/golden/* from the UNMODIFIED reference.  TEST INFRASTRUCTURE ONLY.

Run in the build container (where /root/reference exists):

    python oracle/gen_golden.py

It imports /root/reference/bsrnn.py by path (it needs only torch), loads the
deterministic weights of ``oracle.bsrnn_oracle.make_state_dict`` into the reference's
``BSRNN`` and records reference outputs.  The wrap code around the model cannot be
imported (m_dataset.py / infer*.py need torchmetrics, soundfile, torchaudio.io), so
the few wrap lines are restated here next to their citation and fed to the reference
model.  The GPU box has no /root/reference: tests there read only the committed files.
"""
from __future__ import annotations

import contextlib
import io
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bsrnn_oracle as O  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")


def load_reference():
    # by file path under a private name: the repo root ships its own `bsrnn` drop-in shim
    import importlib.util
    spec = importlib.util.spec_from_file_location("reference_bsrnn", os.path.join(REF, "bsrnn.py"))
    ref = importlib.util.module_from_spec(spec)
    with contextlib.redirect_stdout(io.StringIO()):      # ctor prints the band table, bsrnn.py:351
        spec.loader.exec_module(ref)
    return ref


def ref_model(ref, sd):
    with contextlib.redirect_stdout(io.StringIO()):
        m = ref.BSRNN()
    m.load_state_dict(sd, strict=True)
    return m.eval()


def plugin_loop(m, left, mix, blocks):
    """speech-ladspa-onnx.cpp:174-267 statement by statement around the reference's forward_recurrent:
    run() with the given block sizes (cycled), new_chunk() with double-precision FFTs like FFTW's
    fftw_complex, float tensors in and out of the model, float window."""
    win = (0.5 * (1 - torch.cos(2 * torch.pi * torch.arange(2048, dtype=torch.float64) / 2048))).float()   # :57-59
    current = torch.zeros(2, 2048)
    overlap = [torch.zeros(2048), torch.zeros(2048)]
    state = torch.zeros(4, 2, 24, 64)
    buf_in = torch.zeros(1024)
    buf_out = torch.zeros(1024)
    out = torch.zeros_like(left)
    pos = lp = i = 0
    while pos < left.numel():
        count = min(blocks[i % len(blocks)], left.numel() - pos)
        i += 1
        while count > 0:                                               # run(), :174-191
            n = min(1024 - lp, count)
            buf_in[lp:lp + n] = left[pos:pos + n]
            out[pos:pos + n] = buf_out[lp:lp + n]
            lp += n
            pos += n
            count -= n
            if lp == 1024:                                             # new_chunk(), :193-267
                current = torch.cat((current[:, 1024:], buf_in.expand(2, -1)), 1)        # :199-210
                freq = torch.fft.fft((current[0] * win).double())                         # :206-211
                tin = torch.stack((freq[:1025].real, freq[:1025].imag), 1).reshape(1, -1).float().repeat(2, 1)
                tout, state = m.forward_recurrent(tin, state)                             # :236 (+ :264 state copy)
                y = tout[0].reshape(-1, 2).double()
                wet = torch.complex(y[:, 0].contiguous(), y[:, 1].contiguous())
                half = freq[:1025]
                half = mix * wet + (1.0 - mix) * half if mix >= 0.0 else mix * wet + half    # :215-226
                full = torch.cat((half, half[1:1024].flip(0).conj()))                     # :229-233
                t = torch.fft.ifft(full).real.float()                                     # :236-247 (real part / fft_size)
                overlap = [overlap[1], t]
                s = overlap[0][1024:] + overlap[1][:1024]                                 # :250-258
                wsum = win[1024:] + win[:1024]
                buf_out = s / wsum
                lp = 0
    return out, state


def gen_plugin(m):
    left, _ = O.synth_audio(1, 6 * 1024 + 200, seed=11)
    left = left[0]
    blocks = [300, 1024, 7, 2048, 1500, 64]
    rec = {"left": left.numpy(), "blocks": np.array(blocks)}
    for tag, mix in (("wet", 1.0), ("mix04", 0.4), ("neg05", -0.5)):
        out, state = plugin_loop(m, left, mix, blocks)
        rec[f"out_{tag}"] = out.numpy()
        rec[f"state_{tag}"] = state.numpy()
        rec[f"mix_{tag}"] = np.array([mix], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "plugin_small.npz"), **rec)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "plugin":      # only (re)write plugin_small.npz
        torch.set_grad_enabled(False)
        torch.set_num_threads(1)
        gen_plugin(ref_model(load_reference(), O.make_state_dict(0)))
        return
    torch.set_grad_enabled(False)
    torch.set_num_threads(1)                              # fixed reduction order
    os.makedirs(OUT, exist_ok=True)
    ref = load_reference()

    # ---- contract: keys, shapes, default-init fingerprint ---------------------------------
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        m0 = ref.BSRNN()
    sd0 = m0.state_dict()
    contract = {
        "torch": torch.__version__,
        "bandsplits": ref.generate_bandsplits()[0],
        "keys": [[k, list(v.shape)] for k, v in sd0.items()],
        "n_params": int(sum(v.numel() for v in sd0.values())),
        # fingerprint of `torch.manual_seed(0); BSRNN()` default init
        "seed0_sum": {k: float(v.double().sum()) for k, v in sd0.items()},
        "seed0_abssum": {k: float(v.double().abs().sum()) for k, v in sd0.items()},
    }
    with open(os.path.join(OUT, "contract.json"), "w") as f:
        json.dump(contract, f, indent=0)

    # ---- offline forward, small --------------------------------------------------------------
    sd = O.make_state_dict(0)
    m = ref_model(ref, sd)
    rs = np.random.RandomState(7)
    x = torch.from_numpy((rs.standard_normal((2, 2050, 6)) * 3.0).astype(np.float32))
    taps = {}
    hooks = []
    for j in range(4):
        hooks.append(m.lstms[j].register_forward_hook(
            lambda mod, inp, out, j=j: taps.__setitem__(f"rnn{j}", out.detach().contiguous().clone())))
    hooks.append(m.lstms[0].register_forward_hook(
        lambda mod, inp, out: taps.__setitem__("feat", inp[0].detach().contiguous().clone())))
    res_taps = [None] * 12
    for i in range(12):
        hooks.append(m.bandFCs_pre[i].register_forward_hook(
            lambda mod, inp, out, i=i: res_taps.__setitem__(i, out.detach().clone())))
    y = m.forward(x)
    for h in hooks:
        h.remove()
    np.savez_compressed(
        os.path.join(OUT, "offline_small.npz"), x=x.numpy(), y=y.numpy(),
        residual=torch.cat(res_taps, 2).numpy(), feat=taps["feat"].numpy(),
        **{f"rnn{j}": taps[f"rnn{j}"].numpy() for j in range(4)})

    # ---- recurrent forward: 4 hops, C=2, from a non-zero state ---------------------------------
    xs = torch.from_numpy((rs.standard_normal((4, 2, 2050)) * 3.0).astype(np.float32))
    st = torch.from_numpy((rs.standard_normal((4, 2, 24, 64)) * 0.2).astype(np.float32))
    ys, sts = [], []
    s = st
    for h in range(4):
        yy, s = m.forward_recurrent(xs[h], s)
        ys.append(yy.numpy())
        sts.append(s.numpy())
    np.savez_compressed(os.path.join(OUT, "recurrent_small.npz"), x=xs.numpy(), state0=st.numpy(),
                        y=np.stack(ys), state=np.stack(sts))

    # ---- recurrent == offline on the reference (SURVEY.md 4 invariant 1) ---------------------------
    s = torch.zeros(4, 2, 24, 64)
    yr = []
    for t in range(x.shape[2]):
        yy, s = m.forward_recurrent(x[:, :, t], s)
        yr.append(yy)
    rec_vs_off = float((torch.stack(yr, 2) - y).abs().max())

    # ---- wrap: m_dataset.py:187-190 and :192-195 (infer.py:29-37) -------------------------------
    wav, src = O.synth_audio(2, 7000, seed=5)             # L not a multiple of hop on purpose
    win = torch.hann_window(2048)
    spec = torch.stft(wav, n_fft=2048, hop_length=1024, return_complex=True, window=win)
    spec = torch.stack((spec.real, spec.imag), dim=2)
    packed = spec.reshape((spec.shape[0], -1, spec.shape[3]))
    un = packed.reshape((packed.shape[0], -1, 2, packed.shape[2]))
    un = torch.complex(un[:, :, 0, :], un[:, :, 1, :])
    back = torch.istft(un, n_fft=2048, hop_length=1024, window=win)
    # istft of a non-STFT-consistent spectrogram (random), incl. non-zero DC/Nyquist imag parts
    xr = torch.from_numpy(rs.standard_normal((2, 2050, 5)).astype(np.float32))
    ur = xr.reshape((2, -1, 2, 5))
    ur = torch.complex(ur[:, :, 0, :].contiguous(), ur[:, :, 1, :].contiguous())
    back_r = torch.istft(ur, n_fft=2048, hop_length=1024, window=win)
    np.savez_compressed(os.path.join(OUT, "wrap_small.npz"), wav=wav.numpy(), packed=packed.numpy(),
                        istft=back.numpy(), xr=xr.numpy(), istft_r=back_r.numpy())

    # ---- cfg 1 end to end: [2,176400] (mono dup as infer.py:26-27) -----------------------------
    wav1, src1 = O.synth_audio(1, 176400, seed=1234)
    wav1 = torch.cat((wav1, wav1), 0)
    spec = torch.stft(wav1, n_fft=2048, hop_length=1024, return_complex=True, window=win)
    spec = torch.stack((spec.real, spec.imag), dim=2)
    xin = spec.reshape((spec.shape[0], -1, spec.shape[3]))
    yout = m.forward(xin)
    un = yout.reshape((2, -1, 2, yout.shape[2]))
    un = torch.complex(un[:, :, 0, :], un[:, :, 1, :])
    sep = torch.istft(un, n_fft=2048, hop_length=1024, window=win)
    np.savez_compressed(
        os.path.join(OUT, "cfg1.npz"),
        shape_x=np.array(xin.shape), shape_out=np.array(sep.shape),
        y_dec=yout[:, ::7, ::5].numpy(), sep_dec=sep[:, ::37].numpy(),
        sep_sum=np.array([float(sep.double().sum())]),
        sep_sumsq=np.array([float((sep.double() ** 2).sum())]),
        sep_absmax=np.array([float(sep.abs().max())]),
        si_sdr=O.si_sdr(sep, torch.cat((src1, src1), 0)[:, :sep.shape[1]]).numpy())

    # ---- streaming loop: infer-streaming.py:84-147 around the reference's forward_recurrent ----------
    wavs, _ = O.synth_audio(2, 6 * 1024, seed=9)
    cur = torch.zeros(2, 2048)
    state = torch.zeros(4, 2, 24, 64)
    prev = [torch.zeros(2, 2048)] * 8
    outs = []
    for h in range(6):
        chunk = wavs[:, h * 1024:(h + 1) * 1024]
        wv = torch.cat((cur[:, 1024:], chunk), 1)
        cur = wv
        xx = torch.fft.rfft(wv * win)
        xx = torch.stack((xx.real, xx.imag), dim=2).reshape((2, -1))
        xx, state = m.forward_recurrent(xx, state)
        xx = xx.reshape((2, -1, 2))
        xx = torch.complex(xx[:, :, 0], xx[:, :, 1])
        wv = torch.fft.irfft(xx)
        prev = [wv, prev[0]]
        sum_of_window = torch.zeros(1024)
        new_samples = torch.zeros((2, 1024))
        rng = (0, 1024)
        for wf in prev:
            sum_of_window = sum_of_window + win[rng[0]:rng[1]]
            new_samples = new_samples + wf[:, rng[0]:rng[1]]
            rng = (rng[0] + 1024, rng[1] + 1024)
        outs.append(new_samples / sum_of_window)
    np.savez_compressed(os.path.join(OUT, "stream_small.npz"), wav=wavs.numpy(),
                        out=torch.cat(outs, 1).numpy(), state=state.numpy())

    gen_plugin(m)

    meta = {"torch": torch.__version__, "numpy": np.__version__, "threads": 1,
            "weights": "oracle.bsrnn_oracle.make_state_dict(0)",
            "reference_recurrent_vs_offline_maxabs": rec_vs_off}
    with open(os.path.join(OUT, "meta.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("golden written to", OUT, meta)


if __name__ == "__main__":
    main()
