"""Eager (no graph) streaming hops for an ncu launch list: python tools/prof_hop.py [rows] [hops]."""
import sys, torch
sys.path.insert(0, ".")
from speechseparation_b200 import BSRNN, StreamingSeparator
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2
hops = int(sys.argv[2]) if len(sys.argv) > 2 else 6
torch.manual_seed(0)
m = BSRNN().to("cuda").eval()
sep = StreamingSeparator(m, rows=rows, use_graph=False)
x = torch.randn(rows, 1024, device="cuda") * 0.1
for _ in range(hops):
    sep.step(x)
torch.cuda.synchronize()
print("ok")
