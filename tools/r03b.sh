#!/bin/bash
# r03b: lean predictor rows: parity, A/B phase times
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r03b_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r03b_pytest.log
bash tools/phases.sh > $o/r03b_phases.txt 2>&1; cat $o/r03b_phases.txt
echo "== general predictor rows"; B200CFD_PRED_LEAN=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|predictor"
