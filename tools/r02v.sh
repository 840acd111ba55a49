#!/bin/bash
# r02v: rows per block of the three marching kernels at 4096x1024 (developer knobs B200CFD_{PRED,K23,FIN}_RPB)
run() { python bench.py --steps 20 --warmup 3 --no-cpu --no-scaling-base --no-poisson --nx 4096 --ny 1024 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f'%d['ms_per_step'], ' '.join('%s %.1f'%(k[:9],v['ms']*1e3) for k,v in p.items()))"; }
echo "default"; run
for r in 8 12 20 24 28 32 40 44; do echo "FIN_RPB=$r"; B200CFD_FIN_RPB=$r run; done
for r in 8 12 20 24 28 32 40 44; do echo "K23_RPB=$r"; B200CFD_K23_RPB=$r run; done
for r in 4 6 10 12 16 20 24 32; do echo "PRED_RPB=$r"; B200CFD_PRED_RPB=$r run; done
