"""Drop-in ``bsrnn.BSRNN`` backed by the hand-written sm_100a kernels of libbsrnn_b200.so.

Mirrors the reference module tree (reference bsrnn.py:12-24, 63-162, 328-376) so that
``state_dict()`` has the same 288 keys / shapes / order and ``load_state_dict(strict=True)``
of a reference checkpoint works, and keeps the reference signatures

    BSRNN()                                   (no arguments, bsrnn.py:329)
    BSRNN.forward(x[C,2050,T]) -> y[C,2050,T]                      (bsrnn.py:385)
    BSRNN.forward_recurrent(x[C,2050], state[4,2,C*12,64]) -> (y, new_state)  (bsrnn.py:445)

The ``nn.Linear`` / ``nn.LSTM`` sub-modules only hold parameters; the arithmetic runs in the
CUDA library through its C ABI (include/bsrnn_b200.h).  There is no CPU path: tensors that
are not on a CUDA device raise.  The path is forward-only (north_star): asking autograd for
a gradient through it raises NotImplementedError in backward.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch
from torch import nn

from . import _lib

band_features = 64          # reference bsrnn.py:60
merge_channels = False      # reference bsrnn.py:61
N_SPEC = 2050


def generate_bandsplits() -> Tuple[List[int], List[int]]:
    """Band widths (bins) and neighbour sums, as reference bsrnn.py:247-326 returns them.

    The widths come from the library's compile-time table (bsb_band_widths), the same table
    the kernels index with; ``w`` (unused by the live model) is derived as the reference does.
    """
    buf = (ctypes.c_int * 12)()
    _lib.check(_lib.lib().bsb_band_widths(buf), "bsb_band_widths")
    v = list(buf)
    w = [v[i] + v[i + 1] for i in range(len(v) - 1)] + [v[0] + v[-1]]
    return v, w


class TrainableConstantModule(nn.Module):
    """Parameter holder for the zero-width band 11 (reference bsrnn.py:12-24)."""

    def __init__(self, shape):
        super().__init__()
        self.trainable_constant = nn.Parameter(torch.zeros(shape, dtype=torch.float32))


class NormRNNResidual(nn.Module):
    """Parameter holder: fc_in -> 2-layer LSTM -> fc (+x) (reference bsrnn.py:63-98)."""

    def __init__(self, bidirectional: bool = False):
        super().__init__()
        self.fc_in = nn.Linear(band_features, band_features)
        self.rnn = nn.LSTM(band_features, band_features, batch_first=True, num_layers=2,
                           bidirectional=bidirectional)
        self.fc = nn.Linear(band_features * (2 if bidirectional else 1), band_features)


class TimewiseLSTM(nn.Module):      # reference bsrnn.py:101-128
    def __init__(self):
        super().__init__()
        self.m = NormRNNResidual()


class BandwiseLSTM(nn.Module):      # reference bsrnn.py:131-162
    def __init__(self):
        super().__init__()
        self.m = NormRNNResidual(bidirectional=True)


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


class _ForwardOnly(torch.autograd.Function):
    """Runs the CUDA forward. The hand-written path has no backward kernels (north_star is inference; SURVEY.md 8(f)
    rank 4 is not built): by default a gradient request raises. With ``model.backward_mode = "library"`` (opt-in) the
    backward pass RECOMPUTES the forward graph with stock PyTorch CUDA ops over the module's own parameters
    (cuBLAS / cuDNN -- library kernels, not this repo's) and differentiates that, so that the reference's train.py
    (train.py:98-115) can take optimizer steps; the forward values still come from the B200 kernels."""

    @staticmethod
    def forward(ctx, model, x, *params):
        ctx.model = model
        ctx.save_for_backward(x)
        return model._forward_impl(x)

    @staticmethod
    def backward(ctx, grad_y):
        model = ctx.model
        if model.backward_mode != "library":
            raise NotImplementedError(
                "speechseparation_b200.BSRNN is forward-only: the B200 kernels implement the separation "
                "forward pass (BASELINE.json north_star); no backward kernels are provided. Run under torch.no_grad(), "
                "or opt in to the stock-PyTorch recompute with `model.backward_mode = \"library\"`.")
        (x,) = ctx.saved_tensors
        params = list(model.parameters())
        with torch.enable_grad():
            xg = x.detach().requires_grad_(ctx.needs_input_grad[1])
            y = model._library_graph(xg)
            wanted = ([xg] if ctx.needs_input_grad[1] else []) + [p for p in params if p.requires_grad]
            got = torch.autograd.grad(y, wanted, grad_y, allow_unused=True) if wanted else ()
        it = iter(got)
        gx = next(it) if ctx.needs_input_grad[1] else None
        gp = [next(it) if p.requires_grad else None for p in params]
        return (None, gx, *gp)


class BSRNN(nn.Module):
    def __init__(self):
        super().__init__()
        band_split = generate_bandsplits()
        lrelu = nn.LeakyReLU
        # construction order = the reference's, so `torch.manual_seed(s); BSRNN()` draws the same
        # default init and state_dict() lists the same keys in the same order (bsrnn.py:333-376)
        self.bandFCs_pre = nn.ModuleList([
            (nn.Sequential(nn.Linear(x * 2, x * 2), lrelu(), nn.Linear(x * 2, x * 2), lrelu())
             if x > 0 else nn.Sequential(TrainableConstantModule([0]))) for x in band_split[0]])
        self.bandFCs = nn.ModuleList([
            (nn.Sequential(nn.Linear(x * 2, max(x * 2, band_features)), lrelu(),
                           nn.Linear(max(x * 2, band_features), band_features), lrelu(),
                           nn.Linear(band_features, band_features))
             if x > 0 else nn.Sequential(TrainableConstantModule([band_features]))) for x in band_split[0]])
        print(band_split)                                     # reference side effect, bsrnn.py:351
        self.lstms = nn.Sequential(BandwiseLSTM(), TimewiseLSTM(), BandwiseLSTM(), TimewiseLSTM())
        hidden = band_features * 2
        self.bandFCs_back = nn.ModuleList([
            (nn.Sequential(nn.Linear(band_features, hidden), lrelu(),
                           nn.Linear(hidden, max(x * 2, hidden)), lrelu(),
                           nn.Linear(max(x * 2, hidden), x * 2), lrelu())
             if x > 0 else nn.Sequential(TrainableConstantModule(0))) for x in band_split[0]])
        self.bandFCs_back_post = nn.ModuleList([
            (nn.Sequential(nn.Linear(x * 2, x * 2), lrelu(), nn.Linear(x * 2, x * 2))
             if x > 0 else nn.Sequential(TrainableConstantModule([0]))) for x in band_split[0]])
        # "raise" (default): the path is forward-only. "library": backward recomputes through stock PyTorch ops.
        self.backward_mode = "raise"
        self._reset_caches()

    def _library_graph(self, x: torch.Tensor) -> torch.Tensor:
        """BSRNN.forward as a differentiable graph of stock PyTorch ops over this module's parameters (used ONLY by
        the opt-in backward; values agree with the kernels to ~1e-5). Same arithmetic as reference bsrnn.py:385-443:
        per-band MLPs on [C,T,d], four residual LSTM blocks on [C,T,12,64], decoder MLPs, residual, real multiply."""
        C, _, T = x.shape
        dims = [2 * w for w in generate_bandsplits()[0]]
        xt = x.transpose(1, 2)                                           # [C, T, 2050]
        residual, feats, off = [], [], 0
        for i, d in enumerate(dims):
            if d == 0:                                                   # band 11: learned constant feature
                residual.append(None)
                feats.append(self.bandFCs[i][0].trainable_constant.expand(C, T, band_features))
                continue
            r = self.bandFCs_pre[i](xt[:, :, off:off + d])
            residual.append(r)
            feats.append(self.bandFCs[i](r))
            off += d
        z = torch.stack(feats, 2)                                        # [C, T, 12, 64]
        for j, block in enumerate(self.lstms):
            m = block.m
            seq = z.reshape(C * T, len(dims), band_features) if j % 2 == 0 else \
                z.transpose(1, 2).reshape(C * len(dims), T, band_features)
            h, _ = m.rnn(m.fc_in(seq))
            seq = m.fc(h) + seq
            z = seq.reshape(C, T, len(dims), band_features) if j % 2 == 0 else \
                seq.reshape(C, len(dims), T, band_features).transpose(1, 2)
        masks = [residual[i] + self.bandFCs_back_post[i](self.bandFCs_back[i](z[:, :, i]))
                 for i, d in enumerate(dims) if d > 0]
        return x * torch.cat(masks, 2).transpose(1, 2)

    def _reset_caches(self):
        self._packed: Optional[torch.Tensor] = None
        self._packed_sig = None
        self._packed_event = None         # recorded after bsb_pack_weights on the packing stream
        self._packed_stream = None
        self._plist = None                # cached list(self.parameters()): the module walk costs ~0.4 ms per call
        self._ws = {}                     # workspace per (device, stream): one workspace serves one stream at a time

    # ---- pickling: device caches are not part of the module's state (infer-streaming.py:49) ----
    def __getstate__(self):
        d = self.__dict__.copy()
        for k in ("_packed", "_packed_sig", "_packed_event", "_packed_stream", "_plist"):
            d[k] = None
        d["_ws"] = {}
        return d

    def __setstate__(self, d):
        super().__setstate__(d)
        self.__dict__.setdefault("backward_mode", "raise")
        self._reset_caches()

    def _apply(self, fn, *a, **kw):       # .to() / .cuda() / .float(): parameter storage is replaced
        self.invalidate_packed()
        return super()._apply(fn, *a, **kw)

    def invalidate_packed(self):
        """Drop the packed-weight cache. In-place updates (optimizer steps, load_state_dict, p.data.copy_) and
        .to() are tracked automatically; call this after REPLACING a Parameter object of a sub-module."""
        self._packed = self._packed_sig = self._packed_event = self._packed_stream = self._plist = None

    # ---- weight packing ---------------------------------------------------------------------
    def _signature(self):
        # tensor versions only increase, so their sum changes with every in-place update of any parameter;
        # the data pointers catch storage replacement (.data = ..., .to())
        pl = self._plist
        if pl is None:
            pl = self._plist = list(self.parameters())
        return (sum(p._version for p in pl), tuple(p.data_ptr() for p in pl))

    def packed_weights(self, device: torch.device) -> torch.Tensor:
        """Packed device copy of the 288 parameters; re-packed whenever a parameter's storage
        or version changes (load_state_dict, .to(), optimizer steps). The buffer is produced on the stream
        that is current at packing time; any other stream that asks for it is made to wait for that work."""
        sig = self._signature()
        if self._packed is not None and self._packed_sig == sig and self._packed.device == device:
            cur = torch.cuda.current_stream(device)
            if cur != self._packed_stream:
                cur.wait_event(self._packed_event)
            return self._packed
        params = list(self.state_dict(keep_vars=True).values())
        l = _lib.lib()
        if len(params) != 288:
            raise RuntimeError(f"expected 288 state_dict entries, found {len(params)}")
        for i, p in enumerate(params):
            if p.dtype != torch.float32:
                raise RuntimeError("BSRNN parameters must be float32 (the kernels compute in fp32)")
            if p.device != device:
                raise RuntimeError(f"parameter on {p.device}, input on {device}: move the model with .to()")
            if p.numel() != l.bsb_tensor_numel(i):
                raise RuntimeError(f"state_dict entry {i} has {p.numel()} elements, expected {l.bsb_tensor_numel(i)}")
        keep = [p.detach().contiguous() for p in params]
        table = (ctypes.c_void_p * 288)(*[(t.data_ptr() if t.numel() else None) for t in keep])
        with torch.cuda.device(device):
            _lib.check(l.bsb_init(device.index if device.index is not None else -1), "bsb_init")
            packed = torch.empty(l.bsb_packed_weights_bytes() // 4, dtype=torch.float32, device=device)
            _lib.check(l.bsb_pack_weights(table, 288, packed.data_ptr(), _stream_ptr()), "bsb_pack_weights")
            ev = torch.cuda.Event()
            ev.record()
            self._packed_stream = torch.cuda.current_stream(device)
        packed._bsb_keep = keep           # contiguous copies of non-contiguous parameters must outlive the packing kernels
        self._packed, self._packed_sig, self._packed_event = packed, sig, ev
        return packed

    def workspace(self, nbytes: int, device: torch.device) -> torch.Tensor:
        """Scratch buffer of the CURRENT stream on `device` (grown on demand). Two streams never share one: a
        workspace serves one stream at a time (include/bsrnn_b200.h)."""
        key = (device.index, torch.cuda.current_stream(device).cuda_stream)
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nbytes:
            if ws is not None:
                ws.record_stream(torch.cuda.current_stream(device))     # queued work may still use the old block
            self._ws[key] = None
            ws = self._ws[key] = torch.empty(nbytes, dtype=torch.uint8, device=device)
        return ws

    # ---- checks ------------------------------------------------------------------------------
    @staticmethod
    def _check_input(x: torch.Tensor, ndim: int, what: str):
        if not isinstance(x, torch.Tensor):
            raise TypeError(f"{what} must be a torch.Tensor")
        if not x.is_cuda:
            raise RuntimeError(f"{what} is on {x.device}: speechseparation_b200 has no CPU path; move it to a CUDA device")
        if x.dtype != torch.float32:
            raise RuntimeError(f"{what} must be float32, got {x.dtype}")
        if x.dim() != ndim:
            raise RuntimeError(f"{what} must have {ndim} dims, got shape {tuple(x.shape)}")
        if x.shape[1] != N_SPEC:
            raise RuntimeError(f"{what}: dim 1 must be {N_SPEC} (1025 bins x Re/Im), got shape {tuple(x.shape)}")
        if x.shape[0] < 1:
            raise RuntimeError(f"{what}: empty dim 0")

    def _needs_grad(self, *tensors) -> bool:
        if not torch.is_grad_enabled():
            return False
        return any(t.requires_grad for t in tensors) or any(p.requires_grad for p in self.parameters())

    # ---- reference API -------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x [C,2050,T] fp32 CUDA -> x * mask, same shape (reference bsrnn.py:385-443)."""
        self._check_input(x, 3, "x")
        if x.shape[2] < 1:
            raise RuntimeError("x: empty time axis")
        if torch.jit.is_tracing():             # export path: a recordable operator (export.py)
            from . import export
            with torch.no_grad():
                packed = self.packed_weights(x.device)
            return export.forward_op(x, packed)
        if self._needs_grad(x):
            return _ForwardOnly.apply(self, x, *self.parameters())
        return self._forward_impl(x)

    def _forward_impl(self, x: torch.Tensor) -> torch.Tensor:
        l = _lib.lib()
        x = x.detach().contiguous()
        C, _, T = x.shape
        dev = x.device
        packed = self.packed_weights(dev)
        with torch.cuda.device(dev):
            nbytes = l.bsb_workspace_bytes(C, T)
            ws = self.workspace(nbytes, dev)
            y = torch.empty_like(x)
            _lib.check(l.bsb_forward(x.data_ptr(), C, T, packed.data_ptr(), y.data_ptr(), ws.data_ptr(),
                                     ws.numel(), _stream_ptr()), "bsb_forward")
        return y

    def forward_recurrent(self, x: torch.Tensor, state: torch.Tensor):
        """x [C,2050], state [4,2,C*12,64] -> (y [C,2050], new_state) (reference bsrnn.py:445-510)."""
        self._check_input(x, 2, "x")
        C = x.shape[0]
        if not isinstance(state, torch.Tensor) or tuple(state.shape) != (4, 2, C * 12, band_features):
            shape = tuple(state.shape) if isinstance(state, torch.Tensor) else type(state)
            raise RuntimeError(f"state must have shape (4, 2, {C * 12}, {band_features}) for {C} rows, got {shape}")
        if state.device != x.device or state.dtype != torch.float32:
            raise RuntimeError("state must be float32 on the same CUDA device as x")
        if torch.jit.is_tracing():             # export path (infer-streaming.py:20-28,74): a recordable operator
            from . import export
            with torch.no_grad():
                packed = self.packed_weights(x.device)
            y, new_state = export.forward_recurrent_op(x, state, packed)
            return y, new_state
        if self._needs_grad(x, state):
            raise NotImplementedError("speechseparation_b200.BSRNN.forward_recurrent is forward-only; "
                                      "call it under torch.no_grad()")
        l = _lib.lib()
        x = x.detach().contiguous()
        state = state.detach().contiguous()
        dev = x.device
        packed = self.packed_weights(dev)
        with torch.cuda.device(dev):
            ws = self.workspace(l.bsb_workspace_bytes(C, 1), dev)
            y = torch.empty_like(x)
            new_state = torch.empty_like(state)
            _lib.check(l.bsb_forward_recurrent(x.data_ptr(), C, state.data_ptr(), packed.data_ptr(), y.data_ptr(),
                                               new_state.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr()),
                       "bsb_forward_recurrent")
        return y, new_state

    # ---- fused wrap: waveform in, waveform out (body of m_dataset.infer, m_dataset.py:182-195) ----
    def separate(self, wav: torch.Tensor, window: Optional[torch.Tensor] = None,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """wav [C,L] fp32 CUDA -> separated [C, 1024*(L//1024)]: stft -> forward -> istft in one
        C-ABI call, without materialising the [C,2050,T] boundary tensors."""
        if not wav.is_cuda or wav.dtype != torch.float32 or wav.dim() != 2:
            raise RuntimeError("wav must be a float32 CUDA tensor of shape [C, L]")
        if self._needs_grad(wav):
            raise NotImplementedError("speechseparation_b200.BSRNN.separate is forward-only")
        l = _lib.lib()
        wav = wav.detach().contiguous()
        C, L = wav.shape
        T = l.bsb_num_frames(L)
        if T < 2:
            raise RuntimeError(f"wav too short for reflect padding / overlap-add: L={L} (need > 1024)")
        dev = wav.device
        packed = self.packed_weights(dev)
        with torch.cuda.device(dev):
            ws = self.workspace(l.bsb_workspace_bytes(C, T), dev)
            if out is None:
                out = torch.empty((C, 1024 * (T - 1)), dtype=torch.float32, device=dev)
            _lib.check(l.bsb_separate(wav.data_ptr(), C, L, _p(window), packed.data_ptr(), out.data_ptr(),
                                      ws.data_ptr(), ws.numel(), _stream_ptr()), "bsb_separate")
        return out
