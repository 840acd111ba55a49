#!/bin/bash
# r02w: Euler step with fast reciprocal / square root in the default mode: parity, then the bench
o=gpurun_out
timeout 900 python -m pytest tests/test_gpu_euler.py -x -q -m gpu > $o/r02w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02w_pytest.log
python bench_euler.py --no-cpu > $o/r02w_euler.json 2> $o/r02w_euler.err; echo "euler rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02w_euler.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline'].get('frac'), {k:v for k,v in d['roofline'].items() if 'fp64' in k or 'flop' in k.lower()})
PY
