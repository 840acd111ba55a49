#!/bin/bash
# r02a: state of the code at the start of round 2 + the measurements the round-2 kernel work is planned on.
o=gpurun_out
python -m pytest tests -m gpu -x -q > $o/r02a_pytest.log 2>&1; echo "pytest rc=$?"
tools/ubench_chain > $o/r02a_ubench.txt 2>&1; echo "ubench rc=$?"
python bench.py > $o/r02a_bench.log 2>&1; echo "bench rc=$?"
P="python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base"
$P > $o/plain1.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_(clean_sweep_lean|finalize_fast)" \
  --launch-skip 9 --launch-count 3 -f -o $o/prof_r02a_step $P > $o/ncu_step.log 2>&1; echo "ncu step rc=$?"
Q="python bench_poisson.py --no-cpu --sweeps 24 --depths 1,0"
$Q > $o/r02a_poisson_short.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:"k_(poisson_tb|tb_make_rhs)" \
  --launch-count 10 -f -o $o/prof_r02a_tb $Q > $o/ncu_tb.log 2>&1; echo "ncu tb rc=$?"
timeout 500 ncu --set full --clock-control none -k regex:"k_(jacobi|rbgs_color)" --launch-skip 46 --launch-count 6 -f -o $o/prof_r02a_1sweep $Q > $o/ncu_1sweep.log 2>&1; echo "ncu 1sweep rc=$?"
python bench_poisson.py --no-cpu > $o/r02a_poisson.log 2>&1; echo "poisson rc=$?"
tail -3 $o/r02a_pytest.log; cat $o/r02a_ubench.txt; tail -c 600 $o/r02a_bench.log
