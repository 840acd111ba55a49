#!/bin/bash
# compute-sanitizer over every kernel of the hot path (ONE tool per GPU-box call: tools/sanitize.sh memcheck | racecheck)
tool=${1:-memcheck}
o=gpurun_out
python tools/sanitize_case.py > $o/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 $o/sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py > $o/sanitize_$tool.log 2>&1
echo "compute-sanitizer $tool rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Invalid|Race reported|hazard" $o/sanitize_$tool.log | sort | uniq -c | head -20
tail -3 $o/sanitize_$tool.log
