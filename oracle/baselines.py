"""Timed baselines that sit BESIDE the bench line.  BASELINE INFRASTRUCTURE ONLY (bench.py's cpu_baseline /
gpu_comparator / --impl reference legs import this; nothing under speechseparation_b200/ does).

What is timed is the reference itself where it can travel: the unmodified ``bsrnn.BSRNN`` staged by
``oracle/stage_ref.py`` (``kind: "reference"``) inside the ten wrap lines every reference caller repeats around it
(m_dataset.py:187-195 offline, infer-streaming.py:116-145 per hop; restated in oracle/bsrnn_oracle.py because
m_dataset.py / infer-streaming.py need packages that are not installed).  Without a staged file the oracle port of
the model is timed instead (``kind: "port"``: the same ATen kernels, called from a functional restatement).
"""
from __future__ import annotations

import os
import time
from typing import Callable, Dict

import torch

from . import bsrnn_oracle as O
from .stage_ref import reference_model

SR = O.SAMPLE_RATE


def _forward_fn(seed: int, device) -> "tuple[Callable, Callable, str]":
    """(forward(x), forward_recurrent(x, state), kind) of the reference model, or of the oracle port."""
    m, kind = reference_model(seed, device)
    if m is not None:
        return m.forward, m.forward_recurrent, kind
    sd = {k: v.to(device) for k, v in O.make_state_dict(seed).items()}
    return (lambda x: O.forward(sd, x)), (lambda x, s: O.forward_recurrent(sd, x, s)), "port"


def separate_with(forward: Callable, wav: torch.Tensor) -> torch.Tensor:
    """m_dataset.infer's forward-only body (m_dataset.py:187-195) around `forward`."""
    return O.unpack_istft(forward(O.stft_pack(wav)))


def _best(fn: Callable, repeats: int, sync: Callable = lambda: None) -> float:
    best = float("inf")
    for _ in range(repeats):
        sync()
        t0 = time.perf_counter()
        fn()
        sync()
        best = min(best, time.perf_counter() - t0)
    return best


def stream_hop_fn(forward_recurrent: Callable, device, rows: int = 2):
    """One hop of infer-streaming.py:116-145 (roll, rfft, pack, forward_recurrent, irfft, 2-frame OLA) as a closure
    over its rolling state."""
    win = O.hann(device)
    st = {"buf": torch.zeros(rows, O.N_FFT, device=device), "prev": torch.zeros(rows, O.N_FFT, device=device),
          "state": torch.zeros(4, 2, rows * O.N_BANDS, O.FEAT, device=device)}
    sum_w = win[:O.HOP] + win[O.HOP:]

    def hop(chunk: torch.Tensor) -> torch.Tensor:
        st["buf"] = torch.cat((st["buf"][:, O.HOP:], chunk), 1)
        s = torch.fft.rfft(st["buf"] * win)
        x = torch.stack((s.real, s.imag), dim=2).reshape(rows, -1)
        y, st["state"] = forward_recurrent(x, st["state"])
        y = y.reshape(rows, -1, 2)
        cur = torch.fft.irfft(torch.complex(y[:, :, 0].contiguous(), y[:, :, 1].contiguous()))
        out = (cur[:, :O.HOP] + st["prev"][:, O.HOP:]) / sum_w
        st["prev"] = cur
        return out
    return hop


def cpu_baseline(synth: Callable, rows: int, n: int, repeats: int = 3) -> Dict:
    """SURVEY.md 8(d) "CPU baseline beside it": the reference on the host cores, fp32, no grad, eval.
    * cfg 2 sample (rows x n) with all host threads (the headline `value`) and ONE row with one thread
      (torch.set_num_threads(1): the plugin's setting, speech-ladspa-onnx.cpp:70-72);
    * cfg 1 end to end (2 x 4 s, stft -> model -> istft), all threads and one thread;
    * the streaming hop (forward_recurrent + rfft/irfft/OLA, 2 rows), all threads and one thread."""
    cores = os.cpu_count() or 1
    fwd, fwd_rec, kind = _forward_fn(0, "cpu")
    wav = synth(rows, n, "cpu")
    wav1 = synth(2, 176400, "cpu")
    chunks = 0.1 * torch.randn(40, 2, 1024)
    res: Dict = {}
    with torch.no_grad():
        for tag, threads in (("all", cores), ("1thread", 1)):
            torch.set_num_threads(threads)
            separate_with(fwd, wav[:1, :44100])                                   # warm-up
            w = wav if threads > 1 else wav[:1]
            t = _best(lambda: separate_with(fwd, w), repeats if threads > 1 else 1)
            res[f"cfg2_{tag}"] = {"audio_s_per_s": w.shape[0] * n / SR / t, "threads": threads,
                                  "sample": f"{w.shape[0]} x {n / SR:.1f} s"}
            t1 = _best(lambda: separate_with(fwd, wav1), repeats)
            res[f"cfg1_{tag}"] = {"ms": 1e3 * t1, "audio_s_per_s": 2 * 176400 / SR / t1, "threads": threads,
                                  "sample": "2 x 4.0 s end to end (stft -> model -> istft)"}
            hop = stream_hop_fn(fwd_rec, "cpu")
            for i in range(5):
                hop(chunks[i])
            ts = []
            for i in range(5, 40):
                t0 = time.perf_counter()
                hop(chunks[i])
                ts.append(time.perf_counter() - t0)
            ts.sort()
            res[f"hop_{tag}"] = {"ms_p50": 1e3 * ts[len(ts) // 2], "threads": threads,
                                 "sample": "2 rows, forward_recurrent + rfft/irfft/OLA, 35 hops", "budget_ms": 1e3 * 1024 / SR}
        torch.set_num_threads(cores)
    return {"value": res["cfg2_all"]["audio_s_per_s"], "unit": "audio-sec/s", "cores": cores, "kind": kind,
            "sample": f"{rows} x {n / SR:.1f} s clips, best of {repeats}, torch {torch.__version__} CPU fp32, "
                      f"{'reference bsrnn.py (oracle/_ref) inside the restated wrap' if kind == 'reference' else 'oracle/bsrnn_oracle.py'}",
            "detail": res}


def gpu_comparator(synth: Callable, rows: int, n: int, device, repeats: int = 5) -> Dict:
    """The only "existing GPU kernel" there is: the reference module on STOCK PyTorch CUDA --
    nn.Linear -> cuBLAS, nn.LSTM -> cuDNN, torch.stft/istft -> cuFFT -- fp32 with TF32 off, same workload, inputs
    resident, timed with CUDA events. Reported beside the bench line; it is not the `--impl reference` arm."""
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    try:
        fwd, _, kind = _forward_fn(0, device)
        wav = synth(rows, n, device)
        with torch.no_grad():
            for _ in range(2):
                out = separate_with(fwd, wav)
            torch.cuda.synchronize(device)
            best = float("inf")
            for _ in range(repeats):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                out = separate_with(fwd, wav)
                e1.record()
                torch.cuda.synchronize(device)
                best = min(best, e0.elapsed_time(e1))
        return {"value": rows * n / SR / (best / 1e3), "unit": "audio-sec/s", "ms_per_step": best, "kind": kind,
                "what": "reference BSRNN on stock PyTorch CUDA (cuBLAS / cuDNN / cuFFT), fp32, TF32 off, inputs resident, "
                        f"best of {repeats}, {rows} x {n / SR:.1f} s", "out_shape": list(out.shape)}
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
