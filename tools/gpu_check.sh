timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bcur.log 2> gpurun_out/bcur.err || tail -20 gpurun_out/bcur.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bcur.log").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["e2e"])
print(d["roofline"])
print(" ".join(f"{k}={v['ms']:.3f}" for k,v in d["stages"].items()))
PY
