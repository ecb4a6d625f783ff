mkdir -p gpurun_out
for v in 0 1; do
  if [ $v = 1 ]; then export BSB_SEQ2_NOFC=1; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/ab$v.log 2> gpurun_out/ab$v.err || tail -5 gpurun_out/ab$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab$v.log").read().strip().splitlines()[-1])
print("nofc=$v", d["ms_per_step"], " ".join(f"{k}={v['ms']:.3f}" for k,v in d["stages"].items() if "rnn1" in k or "rnn3" in k))
PY
done
