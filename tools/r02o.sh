#!/bin/bash
# r02o: lean rows of K2+K3 — parity, then A/B phase times; finalize compiled for 4 blocks/SM
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02o_pytest.log
bash tools/phases.sh > $o/r02o_phases.txt 2>&1; cat $o/r02o_phases.txt
echo "== K2+K3 general rows only"; B200CFD_K23_LEAN=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|poisson_correct"
echo "== finalize 4 blocks/SM"; B200CFD_LIB=cfd_simulations_b200/build/variants/fin_minb4.so bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize"
