# Round-2 profile of ONE build: the bench line (no profiler), then the ncu launch list of one timed step of the same
# command. Usage on the GPU box: bash tools/prof_r02.sh <tag>     (outputs under gpurun_out/r02_<tag>_*)
set -e
TAG=${1:-cur}
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_${TAG}_bench.json 2> gpurun_out/r02_${TAG}_bench.err
cut -c1-700 gpurun_out/r02_${TAG}_bench.json
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-gpu-comparator > gpurun_out/r02_${TAG}_plain.log 2>&1
L=$(python -c "import json;print(json.load(open('gpurun_out/r02_${TAG}_plain.log'))['gpu_launches'])")
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k "regex:^(tc_|lstm_|stft|istft|grouped|mlp_)" -s $((3 * L)) -c $L --csv \
  --log-file gpurun_out/r02_${TAG}_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-gpu-comparator \
  > gpurun_out/r02_${TAG}_ncu.log 2>&1
python tools/ncu_launches.py gpurun_out/r02_${TAG}_launches.csv
