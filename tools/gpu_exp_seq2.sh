# timing experiments on lstm_seq2_kernel (BSB_SEQ2_EXP bits: 1 no cell, 2 no MMA, 4 all warps MMA-first); results are wrong by design
mkdir -p gpurun_out
for e in 0 32 64 96; do
  BSB_SEQ2_EXP=$e timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/exp$e.log 2> gpurun_out/exp$e.err || tail -5 gpurun_out/exp$e.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/exp$e.log").read().strip().splitlines()[-1])
print("exp $e", d["ms_per_step"], " ".join(f"{k}={v['ms']:.3f}" for k,v in d["stages"].items() if "rnn1" in k))
PY
done
