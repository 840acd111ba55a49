#!/bin/bash
# r02e: streaming Poisson pipeline: parity tests, then the Poisson bench lines (stream vs tile kernel)
o=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_slabs.py -m gpu -x -q -k "poisson or jacobi or c3 or slab or sor" > $o/r02e_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $o/r02e_pytest.log
timeout 600 python bench_poisson.py --no-cpu --depths 1,0 > $o/r02e_poisson.log 2>&1; echo "poisson rc=$?"; cut -c1-330 $o/r02e_poisson.log
B200CFD_POISSON_STREAM=0 timeout 600 python bench_poisson.py --no-cpu --depths 0 > $o/r02e_poisson_tile.log 2>&1; cut -c1-330 $o/r02e_poisson_tile.log
for W in 4 5 6; do echo "== stream warps/SM $W"; B200CFD_STREAM_WARPS_PER_SM=$W timeout 300 python bench_poisson.py --no-cpu --depths 0 2>&1 | cut -c1-200; done > $o/r02e_knobs.txt 2>&1; cat $o/r02e_knobs.txt
