"""Debug: one time-axis block on a tiny shape, compared with the fp32 kernels."""
import sys, os, io, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speechseparation_b200 import BSRNN, _lib
torch.set_grad_enabled(False)
l = _lib.lib()
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().cuda().eval()
C, T = int(sys.argv[1]), int(sys.argv[2])
packed = m.packed_weights(torch.device("cuda", 0))
ws = torch.empty(l.bsb_workspace_bytes(C, T), dtype=torch.uint8, device="cuda")
fin = torch.randn(C * T, 12, 64, device="cuda")
outs = []
for force in (1, 0):
    l.bsb_set_force_fp32(force)
    fout = torch.zeros_like(fin)
    _lib.check(l.bsb_rnn_block(1, fin.data_ptr(), C, T, packed.data_ptr(), None, None, fout.data_ptr(), ws.data_ptr(), ws.numel(),
                               torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    outs.append(fout.clone())
    print("force", force, "ok", float(fout.abs().max()))
print("max diff", float((outs[0] - outs[1]).abs().max()))
