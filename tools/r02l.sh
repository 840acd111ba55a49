#!/bin/bash
# r02l (8 GPUs): the driver's scaling commands, N = 8 then 4 (same box), plus the reference arm at N = 8
o=gpurun_out
for N in 8 4; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus $N --steps 20 --warmup 5 > $o/r02l_bench_${N}gpu.log 2>&1; echo "bench N=$N rc=$?"; tail -c 700 $o/r02l_bench_${N}gpu.log
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29578 bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > $o/r02l_ref_8gpu.log 2>&1; tail -c 500 $o/r02l_ref_8gpu.log
