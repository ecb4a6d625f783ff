"""bsb_stft_pack_static: the allocation-free variant of bsb_stft_pack (include/bsrnn_b200.h).

No counterpart in the reference; the entry point was added for callers that
cannot allocate. Its frame-major scratch and its tables live in file-scope __device__ arrays of csrc/fft.cu, so the
call takes no workspace. This file is named to be collected last and passes both
alone in a fresh process and after the other GPU tests: the first call on a device does the once-per-device set-up
(cudaGetSymbolAddress + table upload), later calls reuse it, and the shared scratch keeps the previous call's bytes.
"""
import ctypes

import pytest
import torch

from oracle import bsrnn_oracle as O

pytestmark = pytest.mark.gpu
torch.set_grad_enabled(False)
TOL = 2e-6          # the tolerance of the STFT tests in test_gpu_parity.py


def _rel(a, b):
    a, b = a.double().cpu().reshape(-1), b.double().cpu().reshape(-1)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _static(l, wav, window=None):
    from speechseparation_b200 import _lib
    C, L = wav.shape
    T = l.bsb_num_frames(L)
    x = torch.full((C, 2050, T), float("nan"), dtype=torch.float32, device=wav.device)
    code = l.bsb_stft_pack_static(wav.data_ptr(), C, L, None if window is None else window.data_ptr(), x.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    _lib.check(code, "bsb_stft_pack_static")
    return x


def _existing(l, wav, window=None):
    from speechseparation_b200 import _lib
    C, L = wav.shape
    T = l.bsb_num_frames(L)
    x = torch.empty((C, 2050, T), dtype=torch.float32, device=wav.device)
    ws = torch.empty(C * T * 2052 * 4, dtype=torch.uint8, device=wav.device)
    _lib.check(l.bsb_stft_pack(wav.data_ptr(), C, L, None if window is None else window.data_ptr(), x.data_ptr(),
                               ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream), "bsb_stft_pack")
    return x


def test_static_scratch_stft_matches_the_workspace_entry_point_twice():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from speechseparation_b200 import _lib
    l = _lib.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # two calls in a row with DIFFERENT shapes and data: the second runs on a scratch that still holds the first
    # call's spectrum (more rows, more frames), so stale bytes would show
    shapes = [(3, 44100), (2, 9 * 1024 + 37)]
    for i, (C, L) in enumerate(shapes):
        wav, _ = O.synth_audio(C, L, seed=100 + i)
        d = wav.to(dev)
        got = _static(l, d)
        want = _existing(l, d)
        torch.cuda.synchronize()
        assert got.shape == want.shape
        assert bool(torch.isfinite(got).all())
        assert _rel(got, want) < TOL, (i, C, L)
        assert torch.equal(got, want), "same kernels, same tables: expected bit-identical output"
        assert _rel(got, O.stft_pack(wav)) < TOL                      # and the oracle (m_dataset.py:187-190)
    # caller-supplied window, and exactly the limit: C * T = 512 frame-rows
    win = torch.hann_window(2048, device=dev)
    wav, _ = O.synth_audio(8, 63 * 1024, seed=3)                      # T = 64 -> 8 * 64 = 512
    d = wav.to(dev)
    assert l.bsb_num_frames(d.shape[1]) * d.shape[0] == 512
    assert torch.equal(_static(l, d, win), _existing(l, d, win))


def test_static_scratch_stft_rejects_oversize_and_bad_arguments():
    from speechseparation_b200 import _lib
    l = _lib.lib()
    E_ARG, E_SHAPE, E_TOO_LARGE = -1, -2, -5
    dummy = ctypes.c_void_p(256)                                      # never dereferenced: rejected before any launch
    assert l.bsb_stft_pack_static(dummy, 8, 64 * 1024, None, dummy, None) == E_TOO_LARGE      # 8 x 65 = 520 > 512
    assert l.bsb_stft_pack_static(dummy, 1, 513 * 1024, None, dummy, None) == E_TOO_LARGE
    assert l.bsb_stft_pack_static(None, 1, 4096, None, dummy, None) == E_ARG
    assert l.bsb_stft_pack_static(dummy, 0, 4096, None, dummy, None) == E_ARG
    assert l.bsb_stft_pack_static(dummy, 1, 1024, None, dummy, None) == E_SHAPE               # reflect pad needs L > 1024
    msg = l.bsb_error_string(E_TOO_LARGE).decode()
    assert "512" in msg and "bsb_stft_pack_static" in msg
