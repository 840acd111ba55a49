#!/bin/bash
# r02t: lean finalize rows with loads two rows ahead + single-barrier block reduce: parity, A/B phase times
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02t_pytest.log
bash tools/phases.sh > $o/r02t_phases.txt 2>&1; cat $o/r02t_phases.txt
echo "== loads one row ahead"; B200CFD_LIB=cfd_simulations_b200/build/variants/fin_shallow.so bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize"
