#!/usr/bin/env python
"""Warp-state picture of every kernel in an ncu --set full report: cycles a warp spends per issued instruction, by stall
reason (smsp__average_warps_issue_stalled_*_per_issue_active), plus issue-slot use and eligible warps.
python tools/ncu_warpstate.py rep.ncu-rep out.md"""
import csv
import io
import subprocess
import sys

src, dst = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
ki = hdr.index("Kernel Name")
stall = [(i, h) for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
extra = ["smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
         "smsp__warps_active.avg.per_cycle_active", "gpu__time_duration.sum", "smsp__inst_executed.sum", "sm__cycles_active.avg"]
seen = set()
with open(dst, "w") as f:
    f.write(f"# warp states per issued instruction: {src}\n")
    for r in rows[2:]:
        if len(r) <= ki or r[ki] in seen:
            continue
        seen.add(r[ki])
        f.write(f"\n## `{r[ki][:100]}`\n\n")
        for e in extra:
            if e in hdr:
                f.write(f"- {e}: {r[hdr.index(e)]}\n")
        vals = []
        for i, h in stall:
            try:
                vals.append((float(r[i].replace(",", "")), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
            except ValueError:
                pass
        vals.sort(reverse=True)
        f.write("\nstall cycles per issued instruction: " + ", ".join(f"{n} {v:.2f}" for v, n in vals[:9]) + "\n")
