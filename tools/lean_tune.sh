timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_lean.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lean.log; tail -3 gpurun_out/pytest_lean.log
run() { python bench.py --steps 50 --warmup 5 --no-cpu --no-scaling-base "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); p=d['roofline']['phases']
print('ms/step %.4f sweep1 %.4f sweep2 %.4f'%(d['ms_per_step'],p['clean_sweep_1']['ms'],p['clean_sweep_2']['ms']))"; }
for sz in "4096 1024" "16384 16384"; do set -- $sz
echo "== $1x$2 LEAN=0"; B200CFD_CLEAN_LEAN=0 run --nx $1 --ny $2
for U in 1 2 3 6; do echo "== $1x$2 lean unroll $U"; B200CFD_CLEAN_UNROLL=$U run --nx $1 --ny $2; done
done
