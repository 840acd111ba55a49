#!/bin/bash
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02k_pytest.log
bash tools/phases.sh > $o/r02k_phases.txt 2>&1; cat $o/r02k_phases.txt
echo "== old rpb rule"; B200CFD_RPB_MODEL=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|predictor|poisson_correct|finalize"
timeout 600 python bench_poisson.py --no-cpu --depths 0 > $o/r02k_poisson.log 2>&1; cut -c1-200 $o/r02k_poisson.log
