#!/bin/bash
# r04i: tile table in row-major order (default) against strip-major (the r04f order), and swizzled staging stores (variant)
o=gpurun_out; mkdir -p $o; V=cfd_simulations_b200/build/variants
K="poisson or slabs or jacobi or sor_v2"
( timeout 60 python -m pytest tests -x -q -m gpu -k "$K" 2>&1 | tail -1
  B200CFD_LIB=$V/stswz.so timeout 60 python -m pytest tests -x -q -m gpu -k "$K" 2>&1 | tail -1 ) > $o/r04i_pytest.txt 2>&1 &
wait
( echo "== default: row-major tile table"; python tools/alpha_sweep.py --alphas 100 --reps 2
  echo "== stripmajor (the r04f order)"; B200CFD_LIB=$V/stripmajor.so python tools/alpha_sweep.py --alphas 100 --reps 2
  echo "== stswz: row-major + swizzled staging stores"; B200CFD_LIB=$V/stswz.so python tools/alpha_sweep.py --alphas 100 --reps 2 ) > $o/r04i_alpha.txt 2>&1
cat $o/r04i_pytest.txt; grep -v identical $o/r04i_alpha.txt
