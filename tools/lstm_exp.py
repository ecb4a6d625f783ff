"""Timing experiments on the band-axis LSTM layer kernel (lstm_tc.cu g_lstm_exp; results wrong on purpose)."""
import contextlib, io, os, sys
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
l.bsb_set_stage_timing(1)
for name, e in (("baseline", 0), ("no x-part MMAs", 1), ("no h-part MMAs", 2), ("no cell arithmetic", 4), ("no MMAs at all", 3),
                ("x-part only (no h, no cell)", 6), ("h-part only (no x, no cell)", 5), ("nothing", 7)):
    l.bsb_debug_set_lstm_exp(e)
    best = {}
    for _ in range(4):
        m.separate(wav, out=out)
        for k, v in _lib.stage_times():
            best[k] = min(best.get(k, 1e9), v)
    print(f"{name:>30s}  layer0 (KIN=64) {1e3 * best['rnn0.rec0']:7.1f} us   layer1 (KIN=128) {1e3 * best['rnn0.rec1']:7.1f} us")
l.bsb_debug_set_lstm_exp(0)
