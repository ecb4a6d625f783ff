# Multi-GPU lines of one build: cfg 2 (weak), cfg 4 sessions (strong), cfg 5 long-form (one clip per GPU).
# Usage on an N-GPU box: bash tools/gpu_multi.sh N
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_cfg2_${N}gpu.json 2> gpurun_out/r02_multi_${N}.err
$TR bench.py --gpus $N --steps 200 --warmup 5 --workload cfg4_sessions_512 > gpurun_out/r02_bench_cfg4_${N}gpu.json 2>> gpurun_out/r02_multi_${N}.err
$TR bench.py --gpus $N --steps 10 --warmup 3 --workload cfg5_longform_1x60s > gpurun_out/r02_bench_cfg5_${N}gpu.json 2>> gpurun_out/r02_multi_${N}.err
for f in cfg2 cfg4 cfg5; do python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/r02_bench_${f}_${N}gpu.json") if l.startswith("{")][-1])
    e = d["e2e"]
    print("${f}", "N=", d["n_gpus"], "value", round(d["value"]), "ms/step", round(d["ms_per_step"], 3), "e2e", round(e["value"]),
          "pcm16", round(e.get("pcm16_value", 0)), "copy ceiling GB/s", round(e.get("host_copy_ceiling_gbs", 0), 1), "parity", d.get("parity", {}).get("ok"))
except Exception as ex:
    print("${f}", "failed:", ex)
PY
done
tail -3 gpurun_out/r02_multi_${N}.err
