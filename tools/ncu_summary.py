#!/usr/bin/env python
"""Summarise ncu output brought back in gpurun_out/ into small tracked files under profiles/.

  python tools/ncu_summary.py launches gpurun_out/launches_X.csv profiles/launches_X.md
  python tools/ncu_summary.py full gpurun_out/prof_X.ncu-rep profiles/full_X.md [profiles/traffic.json]
"""
import collections
import csv
import io
import json
import re
import subprocess
import sys


def short(name):
    name = name.replace("void ", "")
    name = re.sub(r"\(.*", "", name)
    name = name.replace("b200cfd::", "")
    return name


def launches(src, dst):
    txt = open(src).read()
    txt = txt[txt.index('"ID"'):]
    rows = list(csv.DictReader(io.StringIO(txt)))
    agg = collections.OrderedDict()
    for r in rows:
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1
        a[1] += float(r["Metric Value"])
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu launch list: {src}\n\n`gpu__time_duration.sum`, cold-cache and serialised: compare SHARES, not absolutes. "
                f"{len(rows)} launches, {tot/1e3:.1f} us total.\n\n| kernel | launches | avg us | share % | grid | block |\n|---|---|---|---|---|---|\n")
        for n, (c, t, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{n}` | {c} | {t/c/1e3:.2f} | {100*t/tot:.1f} | {g} | {b} |\n")


WANT = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "MB rd"), ("dram__bytes_write.sum", "MB wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
        ("launch__registers_per_thread", "regs"), ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("smsp__inst_executed.sum", "warp inst"),
        ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe %")]


def full(src, dst, traffic_json=None):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    cols = [(hdr.index(m), lab, m) for m, lab in WANT if m in hdr]
    traffic = {}
    with open(dst, "w") as f:
        f.write(f"# ncu --set full summary: {src}\n\n| kernel | " + " | ".join(l for _, l, _ in cols) + " |\n|---|" + "---|" * len(cols) + "\n")
        for r in rows[2:]:
            vals = []
            for i, lab, m in cols:
                v = r[i].replace(",", "")
                try:
                    x = float(v)
                    u = units[i]
                    if m.startswith("dram__bytes"):
                        x = x * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1.0)
                    if m == "gpu__time_duration.sum":
                        x = x * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}.get(u, 1.0)
                    vals.append(f"{x:.2f}" if abs(x) < 1e5 else f"{x:.3g}")
                except ValueError:
                    vals.append(v)
            f.write(f"| `{short(r[ki])}` | " + " | ".join(vals) + " |\n")
            try:
                rd = float(r[hdr.index('dram__bytes_read.sum')].replace(",", "")); wr = float(r[hdr.index('dram__bytes_write.sum')].replace(",", ""))
                sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
                b = rd * sc.get(units[hdr.index('dram__bytes_read.sum')], 1) + wr * sc.get(units[hdr.index('dram__bytes_write.sum')], 1)
                traffic.setdefault(short(r[ki]), []).append(b)
            except Exception:
                pass
    if traffic_json:
        try:
            old = json.load(open(traffic_json))
        except Exception:
            old = {}
        for k, v in traffic.items():
            old[k] = sum(v) / len(v)
        json.dump(old, open(traffic_json, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(*sys.argv[2:])
