# Round profile: bench line, ncu launch list of one step, full capture of the dominant kernel.
set -e
python bench.py --steps 5 --warmup 3 > gpurun_out/r01_bench_final.json 2> gpurun_out/bench_err.log
tail -c 400 gpurun_out/r01_bench_final.json
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:^(tc_|lstm_|stft|istft|grouped)" -s 50 -c 25 --csv --log-file gpurun_out/r01_launches_final.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu1.log 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^tc_grouped" -s 14 -c 14 -o gpurun_out/prof_tc_final python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu2.log 2>&1
