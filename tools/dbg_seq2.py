"""Debug: one time-axis block through bsb_rnn_block vs the oracle, error per step / sequence."""
import contextlib, io, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bsrnn_oracle as O
from speechseparation_b200 import BSRNN, _lib
torch.set_grad_enabled(False)
C, T = int(sys.argv[1]), int(sys.argv[2])
sd = O.make_state_dict(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN()
m.load_state_dict(sd); m = m.to("cuda:0").eval()
l = _lib.lib(); dev = torch.device("cuda:0")
packed = m.packed_weights(dev)
ws = torch.empty(l.bsb_workspace_bytes(C, T), dtype=torch.uint8, device=dev)
rs = np.random.RandomState(11 * C + T)
fin = torch.from_numpy(rs.standard_normal((C, T, 12, 64)).astype(np.float32))
seqs = fin.permute(0, 2, 1, 3).reshape(C * 12, T, 64)
for j in (1, 3):
    ref, (hn, cn) = O.rnn_block(sd, j, seqs, None)
    ref = ref.reshape(C, 12, T, 64).permute(0, 2, 1, 3)
    for rep in range(2):
        fout = torch.zeros(C * T, 12, 64, device=dev)
        so = torch.zeros(2, 2, C * 12, 64, device=dev)
        _lib.check(l.bsb_rnn_block(j, fin.reshape(C * T, 12, 64).to(dev).data_ptr(), C, T, packed.data_ptr(), None, so.data_ptr(),
                                   fout.data_ptr(), ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream), "blk")
        torch.cuda.synchronize()
        d = (fout.cpu().reshape(C, T, 12, 64) - ref).abs()
        print("   hN err per layer", [f"{float((so[0, k].cpu() - hn[k]).abs().max()):.1e}" for k in range(2)],
              "cN err", [f"{float((so[1, k].cpu() - cn[k]).abs().max()):.1e}" for k in range(2)])
        print(f"block {j} rep {rep}: max err {float(d.max()):.3e}; per step", [f"{float(d[:, t].max()):.1e}" for t in range(min(T, 12))],
              "per band", [f"{float(d[:, :, b].max()):.1e}" for b in range(12)])
