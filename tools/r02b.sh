#!/bin/bash
# r02b: first run of the TMA-staged sweep + fused corrections
o=gpurun_out
python -m pytest tests -m gpu -x -q > $o/r02b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $o/r02b_pytest.log
python bench.py > $o/r02b_bench.log 2>&1; echo "bench rc=$?"; tail -c 300 $o/r02b_bench.log
bash tools/phases.sh > $o/r02b_phases.txt 2>&1; cat $o/r02b_phases.txt
run() { python bench.py --steps 40 --warmup 5 --no-cpu --no-scaling-base --no-poisson "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f sweep1 %.4f sweep2 %.4f tile_h %d'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms'],d['clean_plan']['tile_h']))"; }
for sz in "4096 1024" "16384 16384"; do set -- $sz
for W in 4 6 8 12 16; do echo "== $1x$2 warps/SM $W"; B200CFD_CLEAN_WARPS_PER_SM=$W run --nx $1 --ny $2; done
echo "== $1x$2 TMA off"; B200CFD_CLEAN_TMA=0 B200CFD_FUSE_CORR=0 run --nx $1 --ny $2
echo "== $1x$2 TMA on, corrections unfused"; B200CFD_FUSE_CORR=0 run --nx $1 --ny $2
done > $o/r02b_knobs.txt 2>&1; cat $o/r02b_knobs.txt
