set -e
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:^(tc_|lstm_|stft|istft|grouped)" -s 75 -c 25 --csv --log-file gpurun_out/launches_cur.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu1.log 2>&1
