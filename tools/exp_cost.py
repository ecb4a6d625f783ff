"""Timing experiments on the grouped Linear kernel (results are wrong on purpose): band 9 only and all bands,
with parts of the pipeline switched off (tc_linear.cu g_exp)."""
import contextlib, io, os, sys, ctypes
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speechseparation_b200 import BSRNN, _lib
torch.set_grad_enabled(False)
l = _lib.lib()
dev = torch.device("cuda:0")
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().to(dev).eval()
wav = 0.1 * torch.randn(64, 441000, device=dev)
out = m.separate(wav)
exps = {"baseline": 0, "stores to an L2-resident window": 32, "split but no plane stores": 64, "no stores": 1, "no epilogue": 2, "no copies": 4, "1 MMA of 3": 8, "no MMA": 16, "no copies, no epi": 6,
        "no copies, no stores": 5, "1 MMA, no epi": 10, "1 MMA, no copies": 12, "only MMA": 6, "nothing but sync": 22}
l.bsb_set_stage_timing(1)
for bname, mk in (("band 9", 0x200), ("bands 6-10", 0x7c0), ("all", 0xfff)):
    l.bsb_debug_set_band_mask(mk)
    print(f"--- {bname}: pre (2 launches, us) | post_mask")
    for name, e in exps.items():
        l.bsb_debug_set_exp(e)
        best = {}
        for _ in range(4):
            m.separate(wav, out=out)
            for k, v in _lib.stage_times():
                best[k] = min(best.get(k, 1e9), v)
        print(f"{name:>34s} {1e3 * best['pre']:8.1f} {1e3 * best['post_mask']:8.1f}")
l.bsb_debug_set_exp(0)
l.bsb_debug_set_band_mask(0xfff)
