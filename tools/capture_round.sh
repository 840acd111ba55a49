#!/bin/bash
# One GPU-box session that produces everything profiles/ needs for a code state: plain runs first (never under ncu), then
# the ncu launch list and --set full captures of every kernel of the hot path.  Usage: tools/capture_round.sh r02j
tag=${1:-rXX}
o=gpurun_out
mkdir -p $o
timeout 900 python -m pytest tests -m gpu -x -q > $o/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/${tag}_pytest.log
python bench.py > $o/${tag}_bench.log 2>&1; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_ref.log 2>&1; echo "ref rc=$?"
python bench_poisson.py --depths 1,0 > $o/${tag}_poisson.log 2>&1; echo "poisson rc=$?"
python bench_euler.py > $o/${tag}_euler.log 2>&1; echo "euler rc=$?"
bash tools/phases.sh > $o/${tag}_phases.txt 2>&1; cat $o/${tag}_phases.txt
python tools/f32_speed.py 16384 8192 20 > $o/${tag}_f32_speed.txt 2>&1; python tools/f32_speed.py 4096 1024 50 >> $o/${tag}_f32_speed.txt 2>&1; python tools/f32_speed.py 600 180 200 >> $o/${tag}_f32_speed.txt 2>&1; grep ratio $o/${tag}_f32_speed.txt
# ---- ncu: launch list + full captures (each command ran plain right before)
P="python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base --no-poisson"
$P > $o/plain_step.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches.csv $P > $o/ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_(predictor|div_rb1|clean|finalize|ctl)" \
  --launch-skip 18 --launch-count 6 -f -o $o/prof_${tag}_step $P > $o/ncu_step.log 2>&1; echo "ncu step rc=$?"
PL="python bench.py --steps 2 --warmup 3 --no-cpu --no-scaling-base --no-poisson --nx 16384 --ny 16384"
$PL > $o/plain_step_large.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"k_(predictor|div_rb1|clean|finalize|ctl)" \
  --launch-skip 18 --launch-count 6 -f -o $o/prof_${tag}_step16384 $PL > $o/ncu_step_large.log 2>&1; echo "ncu step 16384 rc=$?"
Q="python bench_poisson.py --no-cpu --sweeps 24 --depths 1,0"
$Q > $o/plain_poisson.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_(poisson_stream|stream_make_rhs|stream_plan)" \
  --launch-skip 2 --launch-count 14 -f -o $o/prof_${tag}_poisson_stream $Q > $o/ncu_ps.log 2>&1; echo "ncu poisson stream rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:"k_(jacobi|rbgs_color)" --launch-skip 46 --launch-count 6 -f -o $o/prof_${tag}_poisson_1sweep $Q > $o/ncu_p1.log 2>&1; echo "ncu poisson one-sweep rc=$?"
QT="env B200CFD_POISSON_STREAM=0 python bench_poisson.py --no-cpu --sweeps 24 --depths 0"
$QT > $o/plain_poisson_tile.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"k_(poisson_tb|tb_make_rhs)" --launch-skip 2 --launch-count 6 -f -o $o/prof_${tag}_poisson_tile $QT > $o/ncu_pt.log 2>&1; echo "ncu poisson tile rc=$?"
E="python bench_euler.py --steps 3 --warmup 3 --no-cpu"
$E > $o/plain_euler.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_euler" --launch-skip 4 --launch-count 3 -f -o $o/prof_${tag}_euler $E > $o/ncu_e.log 2>&1; echo "ncu euler rc=$?"
H="python tools/slab_case.py"
$H > $o/plain_slab.log 2>&1 && timeout 600 ncu --set full --clock-control none -k regex:"k_(halo_rows|columns|sor_v2_bc)" --launch-count 6 -f -o $o/prof_${tag}_halo $H > $o/ncu_h.log 2>&1; echo "ncu halo rc=$?"
# ---- summarise on the box and drop the reports (gpurun_out/ comes back only if it stays under 64 MiB)
for r in step step16384 poisson_stream poisson_1sweep poisson_tile euler halo; do
  [ -f $o/prof_${tag}_$r.ncu-rep ] && python tools/ncu_summary.py full $o/prof_${tag}_$r.ncu-rep $o/${tag}_${r}_full.md $o/${tag}_traffic_$r.json
done
for r in step poisson_stream euler; do
  [ -f $o/prof_${tag}_$r.ncu-rep ] && python tools/ncu_stalls.py $o/prof_${tag}_$r.ncu-rep $o/${tag}_${r}_stalls.md
done
for r in step step16384 poisson_stream euler; do
  [ -f $o/prof_${tag}_$r.ncu-rep ] && python tools/ncu_warpstate.py $o/prof_${tag}_$r.ncu-rep $o/${tag}_${r}_warpstate.md
done
rm -f $o/*.ncu-rep
du -sh $o
tail -c 300 $o/${tag}_bench.log
