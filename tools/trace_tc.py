"""Debug: pipeline timeline of CTA 0 of the tensor-core Linear kernel (needs a -DBSB_TRACE build)."""
import ctypes, sys, os, io, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speechseparation_b200 import BSRNN, _lib
torch.set_grad_enabled(False)
l = _lib.lib()
l.bsb_debug_set_trace.argtypes = [ctypes.c_void_p]
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().cuda().eval()
C, T = 64, 431
R = C * T
packed = m.packed_weights(torch.device("cuda", 0))
ws = torch.empty(l.bsb_workspace_bytes(C, T), dtype=torch.uint8, device="cuda")
xf = torch.randn(R, 2052, device="cuda")
residual = torch.zeros(R, 2052, device="cuda")
feat = torch.zeros(R, 12, 64, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    _lib.check(l.bsb_bandsplit(xf.data_ptr(), R, packed.data_ptr(), residual.data_ptr(), feat.data_ptr(), ws.data_ptr(), ws.numel(), s))
torch.cuda.synchronize()
l.bsb_debug_set_exp(int(os.environ.get("BSB_EXP", "0")))
l.bsb_debug_set_band_mask(int(os.environ.get("BSB_MASK", "0xfff"), 0))
tr = torch.zeros(10 * 6000, dtype=torch.int64, device="cuda")
l.bsb_debug_set_trace(tr.data_ptr())
_lib.check(l.bsb_bandsplit(xf.data_ptr(), R, packed.data_ptr(), residual.data_ptr(), feat.data_ptr(), ws.data_ptr(), ws.numel(), s))
torch.cuda.synchronize()
l.bsb_debug_set_trace(None)
t = tr.cpu().tolist()
ev = []
for role in range(10):
    for i in range(3000):
        code, clk = t[role * 6000 + 2 * i], t[role * 6000 + 2 * i + 1]
        if code: ev.append((clk, code + (1000 if role >= 5 else 0)))
ev.sort()
n = len(ev)
t0 = ev[0][0]
names = {1: "P0 empty-acq", 3: "P0 arrived", 4: "P0 loads-issued", 101: "P1 empty-acq", 103: "P1 arrived", 104: "P1 loads-issued", 40: "L issued", 30: "W issue s0", 31: "W issue s1", 10: "M tempty-acq", 13: "M full-own", 1014: "R1 relay full", 11: "M full-acq", 12: "M issued", 20: "E tfull-acq", 21: "E done", 22: "E loads-issued", 23: "E tmem-read", 24: "E chunk-done"}
with open("gpurun_out/trace_tc.txt", "w") as f:
    for c, code in [e for e in ev if e[1] < 1000][:400]:
        f.write(f"{c - t0:9d} {names.get(code, code)}\n")
print("events", n)
# summary: intervals between successive events of each kind
import collections
by = collections.defaultdict(list)
for c, code in ev:
    by[code].append(c - t0)
for code, v in sorted(by.items()):
    d = [b - a for a, b in zip(v, v[1:])]
    d.sort()
    if d:
        print(f"{str(names.get(code, code)):>16s}: n={len(v):5d} first={v[0]:8d} last={v[-1]:8d} gap p10={d[len(d)//10]:6d} p50={d[len(d)//2]:6d} p90={d[9*len(d)//10]:6d}")
