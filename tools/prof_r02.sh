# Round-2 profile of ONE build: the bench line (no profiler), the ncu launch list of one timed step of the same
# command, and one `--set full` capture of that step (every distinct kernel), all from the same binaries.
# Usage on the GPU box: bash tools/prof_r02.sh <tag> [full]     (outputs under gpurun_out/r02_<tag>_*)
set -e
TAG=${1:-cur}
python bench.py --steps 10 --warmup 3 > gpurun_out/r02_${TAG}_bench.json 2> gpurun_out/r02_${TAG}_bench.err
cut -c1-400 gpurun_out/r02_${TAG}_bench.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming --no-gpu-comparator"
$CMD > gpurun_out/r02_${TAG}_plain.log 2>&1
L=$(python -c "import json;print(json.load(open('gpurun_out/r02_${TAG}_plain.log'))['gpu_launches'])")
K="regex:^(tc_|lstm_|stft|istft|grouped)"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k "$K" -s $((3 * L)) -c $L --csv --log-file gpurun_out/r02_${TAG}_launches.csv $CMD > gpurun_out/r02_${TAG}_ncu.log 2>&1
python tools/ncu_launches.py gpurun_out/r02_${TAG}_launches.csv
if [ "$2" = "full" ]; then
  ncu --set full --clock-control none --import-source on -k "$K" -s $((3 * L)) -c $L -o gpurun_out/r02_${TAG}_full $CMD > gpurun_out/r02_${TAG}_ncufull.log 2>&1
  ls -la gpurun_out/r02_${TAG}_full.ncu-rep
fi
