#!/usr/bin/env python
"""Opcode histogram of every loop (backward branch) of one kernel in an object file / library.
python tools/sass_loops.py build/step_fast.o k_finalize_fastIdLi2E [min_instructions]"""
import collections
import re
import subprocess
import sys

obj, func = sys.argv[1], sys.argv[2]
minn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
ins, p = [], False
for l in out.splitlines():
    if "Function :" in l:
        p = func in l
        continue
    if not p:
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
print("total", len(ins))
for a, t in ins:
    if "BRA" not in t:
        continue
    mm = re.search(r"0x([0-9a-f]+)", t)
    if not mm or int(mm.group(1), 16) >= a:
        continue
    lo, n = int(mm.group(1), 16), (a - int(mm.group(1), 16)) // 16 + 1
    if n < minn:
        continue
    c = collections.Counter()
    for b, t2 in ins:
        if lo <= b <= a:
            c[re.sub(r"^@!?U?P\d\s+", "", t2).split()[0].split(".")[0]] += 1
    print(hex(lo), hex(a), n, c.most_common())
