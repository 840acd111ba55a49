#!/bin/bash
# r03d (8 GPUs): Euler step at N = 8 with the rebuilt default mode
o=gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29579 bench_euler.py --gpus 8 --no-cpu > $o/r03d_euler_8gpu.log 2>&1; echo "euler N=8 rc=$?"; tail -c 600 $o/r03d_euler_8gpu.log
