#!/bin/bash
# developer sweep of the clean-divergence tuning knobs (environment variables read by cleandiv.cu)
for vec in 0 1 2; do for w in 8 12 16; do
  for sz in "4096 1024" "16384 16384"; do
    set -- $sz
    B200CFD_CLEAN_VEC=$vec B200CFD_CLEAN_WARPS_PER_SM=$w python bench.py --steps 20 --warmup 3 --no-cpu --no-scaling-base --nx $1 --ny $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('vec=$vec warps=$w $1x$2 ms/step %.4f sweep1 %.4f sweep2 %.4f'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms']), d['config']['clean_plan']['tile_h'])
"
  done
done; done
