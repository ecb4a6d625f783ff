"""Per-band cost of the grouped Linear launches at cfg 2 (timing experiment: the band mask makes results wrong).
Prints device ms of the four MLP stages with subsets of the 11 bands taking part."""
import contextlib, io, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speechseparation_b200 import BSRNN, _lib
import ctypes
torch.set_grad_enabled(False)
l = _lib.lib()
l.bsb_debug_set_band_mask.argtypes = [ctypes.c_int]
dev = torch.device("cuda:0")
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().to(dev).eval()
wav = 0.1 * torch.randn(64, 441000, device=dev)
out = m.separate(wav)
masks = {"all": 0xfff, "0-5": 0x03f, "6": 0x040, "7": 0x080, "8": 0x100, "9": 0x200, "10": 0x400, "6-10": 0x7c0, "9+10": 0x600,
         "0-7": 0x0ff, "none": 0x800}
l.bsb_set_stage_timing(1)
print(f"{'bands':>6s} " + " ".join(f"{s:>10s}" for s in ("pre", "bandfc", "back", "post_mask", "rnn0.fc")))
for name, mk in masks.items():
    l.bsb_debug_set_band_mask(mk if name != "all" else 0xfff)
    best = {}
    for _ in range(4):
        m.separate(wav, out=out)
        for k, v in _lib.stage_times():
            best[k] = min(best.get(k, 1e9), v)
    print(f"{name:>6s} " + " ".join(f"{1e3 * best[s]:10.1f}" for s in ("pre", "bandfc", "back", "post_mask", "rnn0.fc")))
l.bsb_debug_set_band_mask(0xfff)
