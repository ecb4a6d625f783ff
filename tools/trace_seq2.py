"""Timeline of lstm_seq2_kernel (library built with EXTRA=-DBSB_SEQ2_TRACE): cycles between the trace points of
cell warp 0 and the MMA warp of CTA 0 (layer 0) and CTA 1 (layer 1), steps 200..207 of cfg 2."""
import contextlib, ctypes, io, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speechseparation_b200 import BSRNN, _lib
torch.set_grad_enabled(False)
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().to("cuda:0").eval()
wav = torch.randn(64, 441000, device="cuda:0") * 0.1
for _ in range(3):
    m.separate(wav)
torch.cuda.synchronize()
l = _lib.lib()
n = 2 * 2 * 8 * 8
buf = (ctypes.c_longlong * n)()
l.bsb_debug_seq2_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert l.bsb_debug_seq2_trace(buf, n) == 0
v = torch.tensor(list(buf)).reshape(2, 2, 8, 8)
cell_names = ["wait gates", "ld+wait", "cell", "fence+arrive", "flush"]
mma_names = ["wait h", "issue+commit", "projection"]
for cta in range(2):
    for role in range(2):
        d = v[cta, role]
        k = 6 if role == 0 else 4
        seg = (d[:, 1:k] - d[:, :k - 1]).double().mean(0)
        step = (d[1:, 0] - d[:-1, 0]).double().mean()
        nm = cell_names if role == 0 else mma_names
        print(f"layer {cta} {'cell warp 0' if role == 0 else 'MMA warp'}: step {step:.0f} cycles; " +
              ", ".join(f"{n_} {x:.0f}" for n_, x in zip(nm, seg.tolist())))
