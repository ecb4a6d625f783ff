python -m pytest tests -m gpu -q 2>&1 | tail -2
mkdir -p gpurun_out
for v in 0 1; do
  if [ $v = 1 ]; then export BSB_NO_GEMV=1; fi
  timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/gv$v.log 2> gpurun_out/gv$v.err || tail -5 gpurun_out/gv$v.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/gv$v.log").read().strip().splitlines()[-1])
print("no_gemv=$v", json.dumps(d.get("streaming"))[:400])
print({k:(v["device_ms_p50"], v["host_ms_p99"]) for k,v in d.get("streaming_sessions",{}).items() if isinstance(v,dict)})
PY
done
