#!/bin/bash
# r02z: ncu --set full of the Euler step kernel (default mode) with source-level stalls
o=gpurun_out
E="python bench_euler.py --steps 3 --warmup 3 --no-cpu"
$E > $o/plain_euler.log 2>&1; echo "plain rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_euler_step" --launch-skip 4 --launch-count 1 -f -o $o/prof_r02z_euler $E > $o/ncu_e.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py full $o/prof_r02z_euler.ncu-rep $o/r02z_euler_full.md $o/r02z_traffic_euler.json
python tools/ncu_stalls.py $o/prof_r02z_euler.ncu-rep $o/r02z_euler_stalls.md
python tools/ncu_warpstate.py $o/prof_r02z_euler.ncu-rep $o/r02z_euler_warpstate.md
ncu -i $o/prof_r02z_euler.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
r=list(csv.reader(sys.stdin)); h=r[0]
for i,k in enumerate(h):
    if 'pipe' in k and 'pct' in k and ('inst_executed' in k or 'cycles_active' in k): print(k, r[2][i])
" > $o/r02z_euler_pipes.txt
rm -f $o/*.ncu-rep
cat $o/r02z_euler_full.md $o/r02z_euler_warpstate.md $o/r02z_euler_stalls.md
