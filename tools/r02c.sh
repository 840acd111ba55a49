#!/bin/bash
# r02c: TMA sweep with multiplicative masks: parity, phases, and an ncu capture (source-level) of the sweeps at 4096x1024
o=gpurun_out
python -m pytest tests -m gpu -x -q > $o/r02c_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $o/r02c_pytest.log
bash tools/phases.sh > $o/r02c_phases.txt 2>&1; cat $o/r02c_phases.txt
run() { python bench.py --steps 40 --warmup 5 --no-cpu --no-scaling-base --no-poisson "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f sweep1 %.4f sweep2 %.4f tile_h %d'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms'],d['clean_plan']['tile_h']))"; }
for W in 3 4 5 6 7; do echo "== 4096x1024 warps/SM $W"; B200CFD_CLEAN_WARPS_PER_SM=$W run --nx 4096 --ny 1024; done > $o/r02c_knobs.txt 2>&1; cat $o/r02c_knobs.txt
P="python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base --no-poisson"
$P > $o/plain1.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_clean_sweep_tma" \
  --launch-skip 6 --launch-count 2 -f -o $o/prof_r02c_sweep $P > $o/ncu_sweep.log 2>&1; echo "ncu rc=$?"
