timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in "BSB_X=1"; do
  env $v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/bab.log 2> gpurun_out/bab.err || tail -5 gpurun_out/bab.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bab.log").read().strip().splitlines()[-1])
print("$v", round(d["ms_per_step"],3), " ".join(f"{k}={v['ms']:.3f}" for k,v in d["stages"].items() if not k.startswith("rnn") or k.endswith("fc")))
PY
done
