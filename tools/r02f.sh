#!/bin/bash
o=gpurun_out
Q="python bench_poisson.py --no-cpu --sweeps 24 --depths 0"
$Q > $o/r02f_plain.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:"k_poisson_stream" \
  --launch-skip 4 --launch-count 2 -f -o $o/prof_r02f_stream $Q > $o/ncu_stream.log 2>&1; echo "ncu rc=$?"
tail -3 $o/ncu_stream.log
