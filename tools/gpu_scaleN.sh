# N-GPU weak-scaling check of bench.py, launched the way the driver does it: bash tools/gpu_scaleN.sh N
N=${1:-8}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale$N.log 2> gpurun_out/scale$N.err || tail -30 gpurun_out/scale$N.err
tail -1 gpurun_out/scale$N.log | cut -c1-900
