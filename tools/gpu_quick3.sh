timeout -s KILL 120 python -m pytest tests/test_gpu_parity.py -x -q -k "time_axis or stages_vs_oracle" 2>&1 | tail -8 || exit 1
bash tools/gpu_check.sh
