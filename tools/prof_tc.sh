set -e
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/p_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^tc_grouped" -s 14 -c 3 -o gpurun_out/prof_tc_cur python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/p_ncu2.log 2>&1
