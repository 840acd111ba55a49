#!/bin/bash
# r02g: TMA finalize kernel + streaming Poisson v3: full parity suite, phases, Poisson lines
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02g_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $o/r02g_pytest.log
bash tools/phases.sh > $o/r02g_phases.txt 2>&1; cat $o/r02g_phases.txt
timeout 600 python bench_poisson.py --no-cpu --depths 0 > $o/r02g_poisson.log 2>&1; echo "poisson rc=$?"; cut -c1-200 $o/r02g_poisson.log
for W in 6 7; do echo "== stream warps/SM $W"; B200CFD_STREAM_WARPS_PER_SM=$W timeout 300 python bench_poisson.py --no-cpu --depths 0 2>&1 | cut -c1-200; done > $o/r02g_knobs.txt 2>&1; cat $o/r02g_knobs.txt
echo "== finalize fast (TMA off)"; B200CFD_FIN_TMA=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize"
echo "== finalize nslot 3"; B200CFD_FIN_NSLOT=3 bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize"
