set -e
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^lstm_tc" -s 12 -c 2 -o gpurun_out/prof_lstm_cur -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep
