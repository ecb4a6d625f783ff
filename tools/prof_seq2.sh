# one `ncu --set full` capture of the time-axis kernel and the two FFT kernels (after the plain command exited 0)
set -e
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^(lstm_seq2|stft_warp|istft_warp)" -s 8 -c 4 -o gpurun_out/prof_seq2_fft -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/p_ncu4.log 2>&1
ls -la gpurun_out/*.ncu-rep
