#!/bin/bash
# r02m (2 GPUs): slab / Poisson parity after the keep-rhs change, then the driver's bench commands at N = 1 and 2
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_slabs.py tests/test_gpu_parity.py -x -q -m gpu -k "slab or poisson or nccl" > $o/r02m_tests.log 2>&1; echo "tests rc=$?"; tail -3 $o/r02m_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > $o/r02m_bench_1gpu.log 2>&1; echo "bench1 rc=$?"; tail -c 1500 $o/r02m_bench_1gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus 2 --steps 20 --warmup 5 > $o/r02m_bench_2gpu.log 2>&1; echo "bench2 rc=$?"; tail -c 1500 $o/r02m_bench_2gpu.log
