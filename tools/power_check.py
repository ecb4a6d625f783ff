"""Is the step power-limited? ms/step for runs of different length, with nvidia-smi sampled every 20 ms."""
import contextlib, io, os, subprocess, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from speechseparation_b200 import BSRNN
torch.set_grad_enabled(False)
dev = torch.device("cuda:0")
torch.manual_seed(0)
with contextlib.redirect_stdout(io.StringIO()):
    m = BSRNN().to(dev).eval()
wav = 0.1 * torch.randn(64, 441000, device=dev)
out = m.separate(wav)
for _ in range(3):
    m.separate(wav, out=out)
torch.cuda.synchronize()
for steps in (5, 20, 100, 400):
    time.sleep(1.0)
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "20"],
                         stdout=subprocess.PIPE, text=True)
    time.sleep(0.2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        m.separate(wav, out=out)
    e1.record()
    torch.cuda.synchronize()
    time.sleep(0.05)
    p.terminate()
    lines = [l.strip().split(", ") for l in p.stdout.read().strip().splitlines()]
    busy = [l for l in lines if len(l) == 3 and float(l[1]) > 300]
    clk = sorted(float(l[0]) for l in busy)
    pw = sorted(float(l[1]) for l in busy)
    cap = sum(1 for l in busy if l[2].lower().startswith("active"))
    print(f"steps {steps:4d}: {e0.elapsed_time(e1) / steps:.3f} ms/step; samples under load {len(busy)}: "
          f"sm MHz min/med {clk[0] if clk else 0:.0f}/{clk[len(clk) // 2] if clk else 0:.0f}, power W med/max "
          f"{pw[len(pw) // 2] if pw else 0:.0f}/{pw[-1] if pw else 0:.0f}, sw_power_cap active in {cap}")
