#!/bin/bash
# r03f: predictor prologue overlapped with the control kernel (programmatic dependent launch): parity, A/B
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r03f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r03f_pytest.log
for k in 1 0; do echo "== PDL_CTL=$k"; B200CFD_PDL_CTL=$k python bench.py --steps 128 --warmup 32 --no-cpu --no-scaling-base --no-poisson 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('4096x1024 ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'])"
B200CFD_PDL_CTL=$k python bench.py --steps 256 --warmup 32 --no-cpu --no-scaling-base --no-poisson --nx 600 --ny 180 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('600x180 ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'])"
B200CFD_PDL_CTL=$k python bench.py --steps 16 --warmup 8 --no-cpu --no-scaling-base --no-poisson --nx 16384 --ny 16384 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('16384^2 ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'])"
done
