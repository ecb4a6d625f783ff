"""ctypes binding of libbsrnn_b200.so (C ABI: include/bsrnn_b200.h).

There is no CPU or PyTorch fallback: if the shared library is missing, ``lib()`` raises.
Build it with ``python -c "import __graft_entry__ as g; g.build()"`` (or ``make -C
speechseparation_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libbsrnn_b200.so")

_lib: Optional[C.CDLL] = None

vp, ci, cll, csz, cfp = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t, C.c_void_p

# name -> (restype, argtypes); mirrors include/bsrnn_b200.h one to one
SIGNATURES = {
    "bsb_version": (ci, []),
    "bsb_init": (ci, [ci]),
    "bsb_error_string": (C.c_char_p, [ci]),
    "bsb_band_widths": (ci, [C.POINTER(ci)]),
    "bsb_tensor_numel": (cll, [ci]),
    "bsb_packed_weights_bytes": (csz, []),
    "bsb_pack_weights": (ci, [C.POINTER(vp), ci, vp, vp]),
    "bsb_workspace_bytes": (csz, [ci, ci]),
    "bsb_num_frames": (ci, [cll]),
    "bsb_stft_pack": (ci, [cfp, ci, cll, cfp, cfp, vp, csz, vp]),
    "bsb_stft_pack_static": (ci, [cfp, ci, cll, cfp, cfp, vp]),
    "bsb_pcm16_to_float": (ci, [vp, cll, cfp, vp]),
    "bsb_float_to_pcm16": (ci, [cfp, cll, vp, vp]),
    "bsb_stft_frames": (ci, [cfp, ci, cll, cfp, cfp, vp]),
    "bsb_unpack_istft": (ci, [cfp, ci, ci, cfp, cfp, vp, csz, vp]),
    "bsb_istft_frames": (ci, [cfp, ci, ci, cfp, cfp, vp]),
    "bsb_to_frames": (ci, [cfp, ci, ci, cfp, vp]),
    "bsb_from_frames": (ci, [cfp, ci, ci, cfp, vp]),
    "bsb_forward": (ci, [cfp, ci, ci, vp, cfp, vp, csz, vp]),
    "bsb_forward_frames": (ci, [cfp, ci, ci, vp, cfp, vp, csz, vp]),
    "bsb_forward_recurrent": (ci, [cfp, ci, cfp, vp, cfp, cfp, vp, csz, vp]),
    "bsb_separate": (ci, [cfp, ci, cll, cfp, vp, cfp, vp, csz, vp]),
    "bsb_bandsplit": (ci, [cfp, ci, vp, cfp, cfp, vp, csz, vp]),
    "bsb_rnn_block": (ci, [ci, cfp, ci, ci, vp, cfp, cfp, cfp, vp, csz, vp]),
    "bsb_decode_mask": (ci, [cfp, cfp, cfp, ci, vp, cfp, vp, csz, vp]),
    "bsb_stream_step": (ci, [cfp, ci, cfp, vp, cfp, cfp, cfp, cfp, cfp, vp, csz, vp]),
    "bsb_stream_step_mix": (ci, [cfp, ci, cfp, vp, cfp, cfp, cfp, cfp, cfp, cfp, vp, csz, vp]),
    "bsb_stream_workspace_bytes": (csz, [ci]),
    "bsb_pair_stats": (ci, [cfp, cll, cfp, cll, ci, cll, vp, vp]),
    "bsb_channel_sdr": (ci, [cfp, cll, cfp, cll, ci, cll, vp, vp]),
    "bsb_set_stage_timing": (ci, [ci]),
    "bsb_stage_count": (ci, []),
    "bsb_stage_name": (C.c_char_p, [ci]),
    "bsb_stage_times": (ci, [C.POINTER(C.c_float), ci]),
    "bsb_launch_count": (cll, []),
    "bsb_set_force_fp32": (ci, [ci]),
}

# include/bsrnn_b200_plugin.h (host pointers; the LADSPA-shaped streaming host)
cul, cdb = C.c_ulong, C.c_double
PLUGIN_SIGNATURES = {
    "ladspa_descriptor": (vp, [cul]),
    "bsb_plugin_weight_floats": (cll, []),
    "bsb_plugin_create": (ci, [C.POINTER(vp), vp, cll, ci]),
    "bsb_plugin_create_from_file": (ci, [C.POINTER(vp), C.c_char_p, ci]),
    "bsb_plugin_connect_port": (None, [vp, cul, vp]),
    "bsb_plugin_run": (ci, [vp, cul]),
    "bsb_plugin_reset": (ci, [vp]),
    "bsb_plugin_hops": (cll, [vp]),
    "bsb_plugin_last_hop_ms": (cdb, [vp]),
    "bsb_plugin_max_hop_ms": (cdb, [vp]),
    "bsb_plugin_get_state": (ci, [vp, vp, cll]),
    "bsb_plugin_destroy": (None, [vp]),
}


class BsrnnB200Error(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raise if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BsrnnB200Error(
                f"{LIB_PATH} not found: the CUDA library is not built and there is no fallback path. "
                "Run `python -c 'import __graft_entry__ as g; g.build()'` in the repo root.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in list(SIGNATURES.items()) + list(PLUGIN_SIGNATURES.items()):
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().bsb_error_string(code).decode()
        raise BsrnnB200Error(f"{what}: {msg} (code {code})")


def stage_times() -> "list[tuple[str, float]]":
    """Per-stage device milliseconds of the most recent orchestration call (timing enabled)."""
    l = lib()
    buf = (C.c_float * 96)()
    n = l.bsb_stage_times(buf, 96)
    if n < 0:
        raise BsrnnB200Error(f"bsb_stage_times failed: {n}")
    return [(l.bsb_stage_name(i).decode(), float(buf[i])) for i in range(n)]
