#!/bin/bash
# developer sweep of the lean clean-divergence sweep's knobs (environment variables read by cleandiv.cu)
run() { python bench.py --steps 40 --warmup 5 --no-cpu --no-scaling-base "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f sweep1 %.4f sweep2 %.4f tile_h %d'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms'],d['config']['clean_plan']['tile_h']))"; }
for sz in "16384 16384" "4096 1024"; do set -- $sz
for B in 3 4; do for P in 6 12 24; do echo "== $1x$2 lean minb $B pfd $P"; B200CFD_CLEAN_MINB=$B B200CFD_CLEAN_PFD=$P run --nx $1 --ny $2; done; done
done
