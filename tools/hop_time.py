"""Device time per streaming hop (CUDA graph) for a few row counts: python tools/hop_time.py 2 32 64 100 128"""
import sys, contextlib, io, torch
sys.path.insert(0, ".")
from speechseparation_b200 import BSRNN, StreamingSeparator
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().to("cuda").eval()
for rows in [int(a) for a in sys.argv[1:]] or [2]:
    sep = StreamingSeparator(m, rows=rows)
    x = torch.randn(rows, 1024, device="cuda") * 0.1
    for _ in range(8):
        sep.step(x)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
    for a, b in ev:
        a.record(); sep.step(x); b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    print(f"rows {rows:5d}: p50 {t[100]:.3f} ms  p99 {t[197]:.3f} ms")
