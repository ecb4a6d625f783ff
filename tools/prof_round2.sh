# Round profile, part 2 (separate call: one profiler run per call): full capture of the GEMM and FFT launches of one step.
set -e
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^(tc_grouped|stft|istft)" -s 48 -c 16 -o gpurun_out/prof_final python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu2.log 2>&1
