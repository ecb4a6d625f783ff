# 2-GPU weak-scaling check of bench.py, launched the way the driver does it
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/scale2.log 2> gpurun_out/scale2.err || tail -30 gpurun_out/scale2.err
tail -1 gpurun_out/scale2.log | cut -c1-1500
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/scale1.log 2> gpurun_out/scale1.err || tail -20 gpurun_out/scale1.err
tail -1 gpurun_out/scale1.log | cut -c1-700
