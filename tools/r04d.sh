#!/bin/bash
# r04d: streaming Poisson kernel with 32-byte-swizzled TMA rows (conflict-free 128-bit row loads): parity tests, A/B against
# the unswizzled build
o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests -x -q -m gpu -k "poisson or slabs or jacobi or sor_v2" > $o/r04d_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $o/r04d_pytest.txt
V=cfd_simulations_b200/build/variants
( echo "== default build: swizzled rows"; timeout 200 python tools/alpha_sweep.py --alphas 50,100
  for v in $VARIANTS; do echo "== $v"; B200CFD_LIB=$V/$v.so timeout 200 python tools/alpha_sweep.py --alphas 50,100; done ) > $o/r04d_alpha.txt 2>&1
cat $o/r04d_alpha.txt
