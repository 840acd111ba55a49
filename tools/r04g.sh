#!/bin/bash
# r04g: rhs rows loaded once per pair of levels (streaming Poisson) and swizzled TMA rows in the lexicographic sweep:
# full GPU suite, Poisson rates, step time against the unswizzled sweep build, bench line
o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests -x -q -m gpu > $o/r04g_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $o/r04g_pytest.txt
timeout 200 python tools/alpha_sweep.py --alphas 100 > $o/r04g_alpha.txt 2>&1; cat $o/r04g_alpha.txt
run() { python bench.py --steps 128 --warmup 32 --no-cpu --no-scaling-base --no-poisson "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('ms/step %.4f'%d['ms_per_step'], {k:round(v['ms']*1e3,1) for k,v in d['roofline']['phases'].items()})"; }
( echo "4096x1024 swizzled sweep rows"; run; echo "4096x1024 unswizzled (CLEAN_SWZ=0)"; B200CFD_LIB=cfd_simulations_b200/build/variants/noclean.so run
  echo "swizzled again"; run ) > $o/r04g_step.txt 2>&1; cat $o/r04g_step.txt
timeout 600 python bench.py > $o/r04g_bench.json 2> $o/r04g_bench.err; echo "bench rc=$?"; tail -c 2500 $o/r04g_bench.json
