#!/bin/bash
# One GPU-box session that produces everything profiles/ needs for a code state: plain bench lines first (never under
# ncu), then the ncu launch list and one --set full capture of the step kernels.  Usage: tools/capture_round.sh r01e
tag=${1:-rXX}
mkdir -p gpurun_out
python bench.py > gpurun_out/bench_$tag.log 2>&1; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.log 2>&1; echo "ref rc=$?"
python bench_poisson.py > gpurun_out/poisson_$tag.log 2>&1; echo "poisson rc=$?"
python bench_euler.py > gpurun_out/euler_$tag.log 2>&1; echo "euler rc=$?"
bash tools/phases.sh > gpurun_out/phases_$tag.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base > gpurun_out/ncu1_$tag.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"k_(predictor|div_rb1|clean|finalize)" \
  --launch-skip 18 --launch-count 6 -f -o gpurun_out/prof_$tag python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base \
  > gpurun_out/ncu2_$tag.log 2>&1; echo "ncu full rc=$?"
tail -c 400 gpurun_out/bench_$tag.log
