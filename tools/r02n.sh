#!/bin/bash
# r02n: lean interior rows of the finalize kernel — parity, then A/B phase times
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02n_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02n_pytest.log
bash tools/phases.sh > $o/r02n_phases.txt 2>&1; cat $o/r02n_phases.txt
echo "== general rows only"; B200CFD_FIN_LEAN=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize"
