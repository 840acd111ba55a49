#!/bin/bash
# r02u: end-of-step control fused into the finalize kernel's last block: parity, A/B step time
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r02u_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02u_pytest.log
bash tools/phases.sh > $o/r02u_phases.txt 2>&1; cat $o/r02u_phases.txt
echo "== separate control kernel"; B200CFD_FUSE_CTL=0 bash tools/phases.sh 2>&1 | grep -E "ms/step|finalize|ctl"
