# tests, then the time-axis stage timings of two bench runs (A/B against the numbers in profiles/)
python -m pytest tests -m gpu -q 2>&1 | tail -2
for i in 1 2 3; do python -m pytest tests -m gpu -q -k "time_axis or stream or forward or long_form" 2>&1 | tail -1; done
mkdir -p gpurun_out
for v in 0 1; do
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/chk$v.log 2> gpurun_out/chk$v.err || tail -5 gpurun_out/chk$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/chk$v.log").read().strip().splitlines()[-1])
print("run $v", d["ms_per_step"], " ".join(f"{k}={v['ms']:.3f}" for k,v in d["stages"].items() if "rnn1" in k or "rnn3" in k))
PY
done
