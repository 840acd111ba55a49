#!/bin/bash
# r02h (2 GPUs): real-NCCL parity test, then the slab bench with and without the overlapped halo exchange
o=gpurun_out
timeout 600 python -m pytest tests/test_gpu_slabs.py -m gpu -x -q > $o/r02h_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $o/r02h_pytest.log
timeout 900 python bench.py --gpus 2 --steps 20 --warmup 5 > $o/r02h_bench_2gpu.log 2>&1; echo "bench rc=$?"; tail -c 2500 $o/r02h_bench_2gpu.log
B200CFD_NO_OVERLAP=1 timeout 900 python bench.py --gpus 2 --steps 20 --warmup 5 --no-poisson > $o/r02h_bench_2gpu_nooverlap.log 2>&1; echo "bench(no overlap) rc=$?"; tail -c 1500 $o/r02h_bench_2gpu_nooverlap.log
timeout 600 python bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $o/r02h_ref_2gpu.log 2>&1; tail -c 600 $o/r02h_ref_2gpu.log
