"""Python handle on the LADSPA-shaped streaming host (include/bsrnn_b200_plugin.h, csrc/plugin.cpp).

``SpeechSeparatorPlugin`` drives the same C object a LADSPA host would reach through
``ladspa_descriptor(0)`` (reference: class SpeechSeparator, speech-ladspa-onnx.cpp:17-267): host
float buffers connected to five ports, ``run(sample_count)`` with any block size. It exists for
tests and bench; an audio host dlopen()s ``csrc/libbsrnn_b200.so`` directly and sets
``BSRNN_B200_WEIGHTS`` to a file written by ``export_weights``.
"""
from __future__ import annotations

import ctypes as C
from typing import Mapping

import numpy as np
import torch

from . import _lib

PORT_CONTROL, PORT_IN_L, PORT_IN_R, PORT_OUT_L, PORT_OUT_R = range(5)
STATE_FLOATS = 4 * 2 * 24 * 64


def flat_weights(state_dict: Mapping[str, torch.Tensor]) -> np.ndarray:
    """The 288 state_dict tensors in reference order, C-contiguous fp32, back to back."""
    l = _lib.lib()
    vals = list(state_dict.values())
    if len(vals) != 288:
        raise RuntimeError(f"expected the 288-entry BSRNN state_dict, got {len(vals)} entries")
    for i, v in enumerate(vals):
        if v.numel() != l.bsb_tensor_numel(i):
            raise RuntimeError(f"state_dict entry {i} has {v.numel()} elements, the layout wants {l.bsb_tensor_numel(i)}")
    flat = torch.cat([v.detach().to("cpu", torch.float32).reshape(-1) for v in vals]).numpy()
    assert flat.size == l.bsb_plugin_weight_floats()
    return np.ascontiguousarray(flat)


def export_weights(state_dict: Mapping[str, torch.Tensor], path: str) -> None:
    """Write the raw fp32 weight file the plugin's instantiate() reads (env BSRNN_B200_WEIGHTS)."""
    flat_weights(state_dict).tofile(path)


class SpeechSeparatorPlugin:
    def __init__(self, state_dict: Mapping[str, torch.Tensor] = None, device: int = 0, path: str = None):
        l = _lib.lib()
        self._l = l
        h = C.c_void_p()
        if path is not None:
            _lib.check(l.bsb_plugin_create_from_file(C.byref(h), path.encode(), device), "bsb_plugin_create_from_file")
        else:
            w = flat_weights(state_dict)
            _lib.check(l.bsb_plugin_create(C.byref(h), w.ctypes.data, w.size, device), "bsb_plugin_create")
        self._h = h
        self._ports = [None] * 5

    def connect_port(self, port: int, buf: np.ndarray) -> None:
        """buf: float32 C-contiguous numpy array, kept alive by this object (host-owned, as in LADSPA)."""
        if buf.dtype != np.float32 or not buf.flags.c_contiguous:
            raise RuntimeError("port buffers must be C-contiguous float32 numpy arrays")
        if 0 <= port < 5:
            self._ports[port] = buf
        self._l.bsb_plugin_connect_port(self._h, port, buf.ctypes.data)

    def run(self, sample_count: int) -> None:
        _lib.check(self._l.bsb_plugin_run(self._h, sample_count), "bsb_plugin_run")

    def reset(self) -> None:
        _lib.check(self._l.bsb_plugin_reset(self._h), "bsb_plugin_reset")

    def process(self, left: np.ndarray, mix: float = 1.0, blocks=None) -> np.ndarray:
        """Convenience: feed ``left`` through run() in blocks of the given sizes (cycled; default 1024)
        and return the left output (the right output is the same signal)."""
        left = np.ascontiguousarray(left, dtype=np.float32)
        blocks = list(blocks) if blocks else [1024]
        ctl = np.array([mix], dtype=np.float32)
        out = np.zeros_like(left)
        self.connect_port(PORT_CONTROL, ctl)
        pos, i = 0, 0
        while pos < left.size:
            n = min(blocks[i % len(blocks)], left.size - pos)
            i += 1
            if n == 0:
                self.run(0)
                continue
            a, o = left[pos:pos + n], out[pos:pos + n]
            self.connect_port(PORT_IN_L, a)
            self.connect_port(PORT_IN_R, a)
            self.connect_port(PORT_OUT_L, o)
            self.connect_port(PORT_OUT_R, o)
            self.run(n)
            pos += n
        return out

    @property
    def hops(self) -> int:
        return self._l.bsb_plugin_hops(self._h)

    @property
    def last_hop_ms(self) -> float:
        return self._l.bsb_plugin_last_hop_ms(self._h)

    @property
    def max_hop_ms(self) -> float:
        return self._l.bsb_plugin_max_hop_ms(self._h)

    def state(self) -> torch.Tensor:
        s = np.empty(STATE_FLOATS, dtype=np.float32)
        _lib.check(self._l.bsb_plugin_get_state(self._h, s.ctypes.data, s.size), "bsb_plugin_get_state")
        return torch.from_numpy(s).reshape(4, 2, 24, 64)

    def close(self) -> None:
        if self._h:
            self._l.bsb_plugin_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LadspaDescriptor(C.Structure):
    """ctypes view of bsb_ladspa_descriptor (= LADSPA_Descriptor) for hosts/tests."""
    class Hint(C.Structure):
        _fields_ = [("HintDescriptor", C.c_int), ("LowerBound", C.c_float), ("UpperBound", C.c_float)]

    _fields_ = [("UniqueID", C.c_ulong), ("Label", C.c_char_p), ("Properties", C.c_int), ("Name", C.c_char_p),
                ("Maker", C.c_char_p), ("Copyright", C.c_char_p), ("PortCount", C.c_ulong),
                ("PortDescriptors", C.POINTER(C.c_int)), ("PortNames", C.POINTER(C.c_char_p)),
                ("PortRangeHints", C.POINTER(Hint)), ("ImplementationData", C.c_void_p),
                ("instantiate", C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_ulong)),
                ("connect_port", C.CFUNCTYPE(None, C.c_void_p, C.c_ulong, C.c_void_p)),
                ("activate", C.c_void_p),
                ("run", C.CFUNCTYPE(None, C.c_void_p, C.c_ulong)),
                ("run_adding", C.c_void_p), ("set_run_adding_gain", C.c_void_p), ("deactivate", C.c_void_p),
                ("cleanup", C.CFUNCTYPE(None, C.c_void_p))]


def ladspa_descriptor(idx: int = 0):
    p = _lib.lib().ladspa_descriptor(idx)
    return C.cast(p, C.POINTER(LadspaDescriptor)).contents if p else None
