#!/bin/bash
# r04b: streaming Poisson kernel with the north-row shuffles issued one row step early and the block-parallel plan kernel:
# Poisson / slab parity tests, alpha sweep, A/B against the late-shuffle build and the red-black LDS.64 build
o=gpurun_out; mkdir -p $o
timeout 600 python -m pytest tests -x -q -m gpu -k "poisson or slabs or jacobi or sor_v2" > $o/r04b_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $o/r04b_pytest.txt
V=cfd_simulations_b200/build/variants
( echo "== default build"; timeout 200 python tools/alpha_sweep.py --alphas 0,35,50,70,100,150
  echo "== late shuffles (round-2 form)"; B200CFD_LIB=$V/late_shfl.so timeout 200 python tools/alpha_sweep.py --alphas 50,100
  echo "== red-black: two LDS.64 per level"; B200CFD_LIB=$V/rb_lds64.so timeout 200 python tools/alpha_sweep.py --alphas 0,50 ) > $o/r04b_alpha.txt 2>&1
cat $o/r04b_alpha.txt
