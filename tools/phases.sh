#!/bin/bash
# per-phase CUDA-event times of the step at C2 (4096x1024) and at 16384^2, printed compactly
for sz in "4096 1024" "16384 16384"; do
  set -- $sz
  python bench.py --steps 20 --warmup 3 --no-cpu --no-scaling-base --no-poisson --nx $1 --ny $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1])
print('$1x$2', 'ms/step %.4f'%d['ms_per_step'], 'MLUPS %.0f'%d['value'], 'step frac %.3f'%d['roofline']['whole_step']['frac'])
for k,v in d['roofline']['phases'].items(): print('   %-14s %8.4f ms %8s GB/s'%(k,v['ms'],v['algo_GBs']))
"
done
