# Round profile, part 1: the bench line (no profiler), then the ncu launch list of one step of the same command.
set -e
python bench.py --steps 5 --warmup 3 > gpurun_out/r01_bench_final.json 2> gpurun_out/bench_err.log
tail -c 600 gpurun_out/r01_bench_final.json
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:^(tc_|lstm_|stft|istft|grouped)" -s 63 -c 21 --csv --log-file gpurun_out/r01_launches_final.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu1.log 2>&1
