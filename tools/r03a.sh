#!/bin/bash
# r03a: float32 rounding by magic-number addition instead of two conversions: parity, f32 vs f64 step times
o=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $o/r03a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r03a_pytest.log
python tools/f32_speed.py 16384 8192 20; python tools/f32_speed.py 4096 1024 50; python tools/f32_speed.py 600 180 200
