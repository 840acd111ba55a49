#!/bin/bash
# r03e: several steps per CUDA-graph launch (B200CFD_GRAPH_STEPS): parity, then the step time at 4096x1024 and 600x180
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r03e_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r03e_pytest.log
for k in 1 4 8 16 32; do echo "== GRAPH_STEPS=$k"; B200CFD_GRAPH_STEPS=$k python bench.py --steps 128 --warmup 32 --no-cpu --no-scaling-base --no-poisson 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('4096x1024 ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'])"
B200CFD_GRAPH_STEPS=$k python bench.py --steps 256 --warmup 32 --no-cpu --no-scaling-base --no-poisson --nx 600 --ny 180 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('600x180 ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'])"
done
