#!/bin/bash
# r04a: streaming Poisson kernel with per-chunk class flags and the cost-balanced tile plan: GPU suite, alpha sweep, bench line,
# ncu --set full of the stream kernels
o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests -x -q -m gpu > $o/r04a_pytest.txt 2>&1; echo "pytest rc=$?"; tail -4 $o/r04a_pytest.txt
timeout 300 python tools/alpha_sweep.py > $o/r04a_alpha.txt 2>&1; echo "alpha rc=$?"; cat $o/r04a_alpha.txt
timeout 600 python bench.py > $o/r04a_bench.json 2> $o/r04a_bench.err; echo "bench rc=$?"; tail -c 3000 $o/r04a_bench.json
Q="python bench_poisson.py --no-cpu --sweeps 24 --depths 0"
$Q > $o/plain_poisson.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_(poisson_stream|stream_make_rhs|stream_plan)" \
  --launch-skip 2 --launch-count 16 -f -o $o/prof_r04a_poisson_stream $Q > $o/ncu_ps.log 2>&1; echo "ncu poisson stream rc=$?"
python tools/ncu_summary.py full $o/prof_r04a_poisson_stream.ncu-rep $o/r04a_poisson_stream_full.md $o/r04a_traffic_poisson_stream.json
python tools/ncu_warpstate.py $o/prof_r04a_poisson_stream.ncu-rep $o/r04a_poisson_stream_warpstate.md
python tools/ncu_stalls.py $o/prof_r04a_poisson_stream.ncu-rep $o/r04a_poisson_stream_stalls.md
rm -f $o/*.ncu-rep
cat $o/r04a_poisson_stream_full.md; head -30 $o/r04a_poisson_stream_warpstate.md
