#!/bin/bash
# r03g: programmatic dependent launch for every kernel of the step (B200CFD_PDL=1), with and without a release before the
# block epilogues of the finalize and K2+K3 kernels
run() { python bench.py --steps 128 --warmup 32 --no-cpu --no-scaling-base --no-poisson 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('4096x1024 ms/step %.4f'%d['ms_per_step'])"; }
echo "default"; run
echo "PDL=1"; B200CFD_PDL=1 run
echo "PDL=1 + late trigger"; B200CFD_PDL=1 B200CFD_LIB=cfd_simulations_b200/build/variants/late_trig.so run
