# one tensor-core parity test first (short timeout: a pipeline deadlock must not hold the box), then the rest
timeout -s KILL 90 python -m pytest tests/test_gpu_parity.py -x -q -k "tensor_core_linears" 2>&1 | tail -5 || exit 1
bash tools/gpu_check.sh
