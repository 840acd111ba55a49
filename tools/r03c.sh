#!/bin/bash
# r03c (8 GPUs): the driver's scaling commands at N = 8 and N = 2 with the final kernels, plus the 2-rank NCCL parity test
o=gpurun_out
for N in 8 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 20 --warmup 5 > $o/r03c_bench_${N}gpu.log 2>&1; echo "bench N=$N rc=$?"; tail -c 400 $o/r03c_bench_${N}gpu.log
done
timeout 600 python -m pytest tests/test_gpu_slabs.py -x -q -m gpu > $o/r03c_pytest_slabs.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r03c_pytest_slabs.log
