#!/bin/bash
# r04c: streaming Poisson kernel, 9 warps per SM (3 blocks of 3 warps, 168 registers) against 8 (2 blocks of 4, up to 255)
o=gpurun_out; mkdir -p $o
V=cfd_simulations_b200/build/variants
( echo "== default build: 8 warps per SM, Jacobi late shuffles, red-black early shuffles"; timeout 200 python tools/alpha_sweep.py --alphas 50,100
  for v in $VARIANTS; do echo "== $v"; B200CFD_LIB=$V/$v.so timeout 200 python tools/alpha_sweep.py --alphas 50,100; done ) > $o/r04c_alpha.txt 2>&1
cat $o/r04c_alpha.txt
