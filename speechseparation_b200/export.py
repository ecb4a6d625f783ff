"""Traceable operators for the export path.

The reference exports the streaming step by tracing ``ToTrace(model)`` (infer-streaming.py:20-28,
:62-74: ``forward(x, state) -> (y, new_state, tensor([]))``). A tracer records operators, and a
ctypes call into the C-ABI library is invisible to it, so the two model entry points are also
registered as ``torch.library`` operators:

    bsrnn_b200::forward(x [C,2050,T], packed) -> y
    bsrnn_b200::forward_recurrent(x [C,2050], state [4,2,12C,64], packed) -> (y, new_state)

``packed`` is the packed weight buffer of ``BSRNN.packed_weights`` (it becomes a constant of the
traced graph). ``BSRNN.forward`` / ``forward_recurrent`` route through these operators only while
``torch.jit.is_tracing()``; the eager product path keeps its direct C-ABI call and its cached
workspace. Both operators are CUDA-only (no CPU kernel is registered: calling them with CPU tensors
raises NotImplementedError from the dispatcher), with a meta kernel for shape propagation.

A traced module saved with ``torch.jit.save`` loads in any Python process that has imported
``speechseparation_b200`` (which registers the operators). ONNX export of these operators needs a
custom-domain symbolic and a runtime that implements it; ``onnx`` is not installed here, so that
leg is not provided (out of scope).
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from . import _lib

N_SPEC = 2050
_ws_cache: Dict[Tuple[int, int], torch.Tensor] = {}


def _workspace(C: int, T: int, dev: torch.device) -> torch.Tensor:
    """Scratch for one call from a traced graph: one buffer per (device, current stream), grown on demand."""
    nbytes = _lib.lib().bsb_workspace_bytes(C, T)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(dev).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        _ws_cache.pop(key, None)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _ws_cache[key] = ws
    return ws


def _check_packed(packed: torch.Tensor) -> None:
    want = _lib.lib().bsb_packed_weights_bytes() // 4
    if packed.dtype != torch.float32 or packed.numel() != want or not packed.is_contiguous():
        raise RuntimeError(f"packed must be the contiguous float32 buffer of BSRNN.packed_weights ({want} floats)")


@torch.library.custom_op("bsrnn_b200::forward", mutates_args=(), device_types="cuda")
def forward_op(x: torch.Tensor, packed: torch.Tensor) -> torch.Tensor:
    l = _lib.lib()
    if x.dtype != torch.float32 or x.dim() != 3 or x.shape[1] != N_SPEC or x.shape[0] < 1 or x.shape[2] < 1:
        raise RuntimeError(f"x must be float32 [C, {N_SPEC}, T], got {x.dtype} {tuple(x.shape)}")
    _check_packed(packed)
    x = x.contiguous()
    C, _, T = x.shape
    with torch.cuda.device(x.device):
        ws = _workspace(C, T, x.device)
        y = torch.empty_like(x)
        _lib.check(l.bsb_forward(x.data_ptr(), C, T, packed.data_ptr(), y.data_ptr(), ws.data_ptr(), ws.numel(),
                                 torch.cuda.current_stream().cuda_stream), "bsb_forward")
    return y


@forward_op.register_fake
def _(x, packed):
    return torch.empty_like(x)


@torch.library.custom_op("bsrnn_b200::forward_recurrent", mutates_args=(), device_types="cuda")
def forward_recurrent_op(x: torch.Tensor, state: torch.Tensor, packed: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    l = _lib.lib()
    if x.dtype != torch.float32 or x.dim() != 2 or x.shape[1] != N_SPEC or x.shape[0] < 1:
        raise RuntimeError(f"x must be float32 [C, {N_SPEC}], got {x.dtype} {tuple(x.shape)}")
    C = x.shape[0]
    if state.dtype != torch.float32 or tuple(state.shape) != (4, 2, C * 12, 64) or state.device != x.device:
        raise RuntimeError(f"state must be float32 (4, 2, {C * 12}, 64) on {x.device}, got {tuple(state.shape)}")
    _check_packed(packed)
    x, state = x.contiguous(), state.contiguous()
    with torch.cuda.device(x.device):
        ws = _workspace(C, 1, x.device)
        y = torch.empty_like(x)
        new_state = torch.empty_like(state)
        _lib.check(l.bsb_forward_recurrent(x.data_ptr(), C, state.data_ptr(), packed.data_ptr(), y.data_ptr(),
                                           new_state.data_ptr(), ws.data_ptr(), ws.numel(),
                                           torch.cuda.current_stream().cuda_stream), "bsb_forward_recurrent")
    return y, new_state


@forward_recurrent_op.register_fake
def _(x, state, packed):
    return torch.empty_like(x), torch.empty_like(state)


class ToTrace(torch.nn.Module):
    """The reference's tracing wrapper (infer-streaming.py:20-28): (x, state) -> (y, new_state, tensor([]))."""

    def __init__(self, model):
        super().__init__()
        self.model = model

    def forward(self, x, state):
        a, b = self.model.forward_recurrent(x, state)
        return a, b, torch.tensor([])


def trace_streaming_step(model, rows: int = 2, device: str = "cuda"):
    """torch.jit.trace of ToTrace(model) on the reference's example input shapes
    (infer-streaming.py:57: ones [2, 2050], zeros [4, 2, 24, 64]); the traced graph holds one
    bsrnn_b200::forward_recurrent node and the packed weights as a constant."""
    ex = (torch.ones(rows, N_SPEC, device=device), torch.zeros(4, 2, rows * 12, 64, device=device))
    with torch.no_grad():
        return torch.jit.trace(ToTrace(model), ex, check_trace=False)
