#!/bin/bash
# Step-only part of capture_round.sh (after a change to a step kernel): bench line, phases, launch list, ncu --set full of the
# step kernels at both sizes.  Usage: tools/capture_step.sh r02z
tag=${1:-rXX}
o=gpurun_out
mkdir -p $o
python bench.py > $o/${tag}_bench.log 2>&1; echo "bench rc=$?"
bash tools/phases.sh > $o/${tag}_phases.txt 2>&1; cat $o/${tag}_phases.txt
P="python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base --no-poisson"
$P > $o/plain_step.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv $P > $o/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_(predictor|div_rb1|clean|finalize|ctl)" \
  --launch-skip 18 --launch-count 6 -f -o $o/prof_${tag}_step $P > $o/ncu_step.log 2>&1; echo "ncu step rc=$?"
PL="python bench.py --steps 2 --warmup 3 --no-cpu --no-scaling-base --no-poisson --nx 16384 --ny 16384"
$PL > $o/plain_step_large.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"k_(predictor|div_rb1|clean|finalize|ctl)" \
  --launch-skip 18 --launch-count 6 -f -o $o/prof_${tag}_step16384 $PL > $o/ncu_step_large.log 2>&1; echo "ncu step 16384 rc=$?"
for r in step step16384; do
  python tools/ncu_summary.py full $o/prof_${tag}_$r.ncu-rep $o/${tag}_${r}_full.md $o/${tag}_traffic_$r.json
  python tools/ncu_warpstate.py $o/prof_${tag}_$r.ncu-rep $o/${tag}_${r}_warpstate.md
done
python tools/ncu_stalls.py $o/prof_${tag}_step.ncu-rep $o/${tag}_step_stalls.md
rm -f $o/*.ncu-rep
tail -c 300 $o/${tag}_bench.log
