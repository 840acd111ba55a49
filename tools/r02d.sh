#!/bin/bash
# r02d: ring slots x tiles per SM of the TMA sweep at 4096x1024 and 16384^2
o=gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "clean or steps_vs_oracle or full_c2" > $o/r02d_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02d_pytest.log
run() { python bench.py --steps 40 --warmup 5 --no-cpu --no-scaling-base --no-poisson "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f sweep1 %.4f sweep2 %.4f tile_h %d'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms'],d['clean_plan']['tile_h']))"; }
for S in 2 3 4; do for W in 6 7 8 10 12 14 16; do echo "== 4096x1024 nslot $S warps/SM $W"; B200CFD_CLEAN_NSLOT=$S B200CFD_CLEAN_WARPS_PER_SM=$W run --nx 4096 --ny 1024; done; done > $o/r02d_knobs.txt 2>&1
for S in 2 3 4; do for W in 8 12; do echo "== 16384x16384 nslot $S warps/SM $W"; B200CFD_CLEAN_NSLOT=$S B200CFD_CLEAN_WARPS_PER_SM=$W run --nx 16384 --ny 16384; done; done >> $o/r02d_knobs.txt 2>&1
cat $o/r02d_knobs.txt
