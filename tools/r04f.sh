#!/bin/bash
# r04f: final capture of the round's streaming-Poisson work (per-chunk class selects, cost-balanced tile plan, swizzled TMA rows):
# full GPU suite, bench line, Poisson benchmark lines, ncu --set full of the stream kernels with the shared-memory conflict counters
o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests -x -q -m gpu > $o/r04f_pytest.txt 2>&1; echo "pytest rc=$?"; tail -3 $o/r04f_pytest.txt
timeout 600 python bench.py > $o/r04f_bench.json 2> $o/r04f_bench.err; echo "bench rc=$?"
timeout 300 python bench_poisson.py --no-cpu --depths 1,0 > $o/r04f_poisson_bench.jsonl 2> $o/r04f_poisson.err; echo "poisson rc=$?"
Q="python bench_poisson.py --no-cpu --sweeps 24 --depths 0"
$Q > $o/plain_poisson.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_(poisson_stream|stream_make_rhs|stream_plan)" \
  --launch-skip 2 --launch-count 16 -f -o $o/prof_r04f_poisson_stream $Q > $o/ncu_ps.log 2>&1; echo "ncu poisson stream rc=$?"
python tools/ncu_summary.py full $o/prof_r04f_poisson_stream.ncu-rep $o/r04f_poisson_stream_full.md $o/r04f_traffic_poisson_stream.json
python tools/ncu_warpstate.py $o/prof_r04f_poisson_stream.ncu-rep $o/r04f_poisson_stream_warpstate.md
python tools/ncu_stalls.py $o/prof_r04f_poisson_stream.ncu-rep $o/r04f_poisson_stream_stalls.md
ncu -i $o/prof_r04f_poisson_stream.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
r=list(csv.reader(sys.stdin)); h=r[0]
keys=[i for i,k in enumerate(h) if ('bank_conflict' in k and 'shared' in k) or 'wavefronts_mem_shared' in k or k in ('Kernel Name','gpu__time_duration.sum','sm__cycles_active.avg','sm__cycles_elapsed.max','smsp__inst_executed_op_shared_ld.sum','smsp__inst_executed_op_shared_st.sum')]
for row in r[2:]:
    print(' | '.join(h[i].replace('l1tex__','')+'='+row[i][:60] for i in keys))
" > $o/r04f_poisson_stream_smem.txt
rm -f $o/*.ncu-rep
cat $o/r04f_poisson_stream_full.md; head -12 $o/r04f_poisson_stream_warpstate.md; head -4 $o/r04f_poisson_stream_smem.txt; tail -c 1500 $o/r04f_bench.json
