"""Copy what tools/capture_step.sh TAG left in gpurun_out/ into profiles/ (summaries, bench line, launch list, traffic.json).
python tools/store_step_capture.py r02z"""
import collections
import csv
import json
import shutil
import subprocess
import sys

tag = sys.argv[1]
o = "gpurun_out"
for f in ("step16384_full", "step16384_warpstate", "step_full", "step_stalls", "step_warpstate"):
    shutil.copy(f"{o}/{tag}_{f}.md", "profiles/")
shutil.copy(f"{o}/{tag}_phases.txt", "profiles/")
line = [l for l in open(f"{o}/{tag}_bench.log") if l.startswith("{")][-1]
open(f"profiles/{tag}_bench.json", "w").write(line)
rows = [r for r in csv.reader(open(f"{o}/{tag}_launches.csv")) if len(r) > 5]
hdr, agg = None, collections.OrderedDict()
for r in rows:
    if r[0] == "ID":
        hdr = r
        continue
    if hdr is None:
        continue
    try:
        name, v = r[hdr.index("Kernel Name")], float(r[hdr.index("Metric Value")].replace(",", ""))
    except (ValueError, IndexError):
        continue
    unit = r[hdr.index("Metric Unit")]
    v = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
with open(f"profiles/{tag}_launches.md", "w") as f:
    f.write("# ncu launch list (gpu__time_duration.sum, --clock-control none): python bench.py --steps 3 --warmup 3 --no-cpu --no-scaling-base --no-poisson\n\n")
    f.write("Per-launch times under ncu are cold-cache and serialised: the SHARE of each kernel is what should agree with bench.py.\n\n"
            "| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{k[:80]}` | {n} | {t:.1f} | {t / n:.2f} | {100 * t / tot:.1f} % |\n")
t4, t16 = json.load(open(f"{o}/{tag}_traffic_step.json")), json.load(open(f"{o}/{tag}_traffic_step16384.json"))


def phase(d):
    out = {}
    for k, v in d.items():
        key = ("predictor" if "predictor" in k else "poisson_correct" if "div_rb1_correct" in k else
               "clean_sweep_1" if "clean_sweep_tma<double, 0" in k else "clean_sweep_2" if "clean_sweep_tma<double, 2" in k else
               "finalize2" if "finalize" in k else "ctl" if "ctl" in k else None)
        if key:
            out[key] = v
    return out


head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
tr = {"by_kernel_name": {"16384x16384": t16, "4096x1024": t4},
      "capture": f"{tag} @ {head} (tools/capture_step.sh {tag}: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch)",
      "grid": "4096x1024", "kernels": phase(t4), "kernels_16384": phase(t16)}
json.dump(tr, open("profiles/traffic.json", "w"), indent=1, sort_keys=True)
json.dump(tr, open(f"profiles/traffic_{tag}.json", "w"), indent=1, sort_keys=True)
d = json.loads(line)
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["kernel"], round(d["roofline"]["frac"], 3), d["roofline"]["whole_step"]["frac"])
print({k: v.get("GLUPS") for k, v in d["poisson"]["modes"].items()})
