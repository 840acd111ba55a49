#!/usr/bin/env python
"""Compact stall picture of every kernel in an ncu report captured with --import-source on: warp-stall samples by opcode
(what the warp was waiting to issue) and the ten hottest SASS lines.  python tools/ncu_stalls.py rep.ncu-rep out.md"""
import collections
import csv
import io
import subprocess
import sys

src, dst = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", src, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
kern, hdr, agg, hot, seen, seq = None, None, {}, {}, set(), {}
for r in rows:
    if r and r[0] == "Kernel Name":
        kern = r[1][:90]
        if kern in seen:
            kern = None          # one launch per kernel is enough
            continue
        seen.add(kern)
        agg[kern], hot[kern] = collections.Counter(), []
        continue
    if r and r[0] == "Address":
        hdr = r
        continue
    if kern is None or hdr is None or len(r) < 6:
        continue
    try:
        n = int(r[hdr.index("# Samples")])
    except (ValueError, IndexError):
        continue
    toks = r[1].split()
    op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
    agg[kern][op.split(".")[0]] += n
    hot[kern].append((n, r[1].strip()[:80]))
    seq.setdefault(kern, []).append(r[1].strip()[:80])
with open(dst, "w") as f:
    f.write(f"# warp-stall samples by opcode: {src}\n")
    for k, c in agg.items():
        tot = sum(c.values()) or 1
        f.write(f"\n## `{k}` ({tot} samples)\n\n")
        f.write(" ".join(f"{op} {100 * n / tot:.1f}%" for op, n in c.most_common(12)) + "\n\nhottest lines:\n\n")
        for n, line in sorted(hot[k], reverse=True)[:10]:
            f.write(f"    {100 * n / tot:5.1f}%  {line}\n")
        # the twelve instructions in front of the two hottest lines (what they wait for)
        for n, line in sorted(hot[k], reverse=True)[:2]:
            i = seq[k].index(line)
            f.write(f"\ncontext of `{line}`:\n\n")
            for l in seq[k][max(0, i - 12):i + 2]:
                f.write(f"    {l}\n")
