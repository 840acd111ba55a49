#!/bin/bash
# r02x: Euler step: select-based min/max, march unrolled by two (A/B)
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_euler.py -x -q -m gpu > $o/r02x_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02x_pytest.log
for L in cfd_simulations_b200/libb200cfd.so cfd_simulations_b200/build/variants/eu_pf3.so cfd_simulations_b200/build/variants/eu_pf4.so cfd_simulations_b200/build/variants/eu_pf8.so; do
B200CFD_LIB=$L python bench_euler.py --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$L', d['value'], d['ms_per_step'])"
done
