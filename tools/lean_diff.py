"""Which lean path changes bits?  Runs the 1536x512 case with each of the three lean switches off in turn (subprocesses)
and prints the max |difference| of every field against the all-lean run."""
import os, subprocess, sys, tempfile
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
from test_gpu_lean_paths import CASE
def run(tag, env_over):
    out = Path(tempfile.gettempdir()) / f"lean_{tag}.npz"
    env = dict(os.environ); env.update(env_over)
    r = subprocess.run([sys.executable, "-c", CASE, str(out), "f64", os.environ.get("LEAN_STEPS", "12")], env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1500:]
    return np.load(out)
ref = run("all", {})
for knob in (() if os.environ.get("LEAN_ONLY_K23") else ("B200CFD_PRED_LEAN", "B200CFD_K23_LEAN", "B200CFD_FIN_LEAN")):
    b = run(knob, {knob: "0"})
    print(knob, "= 0:", {k: float(np.nanmax(np.abs(ref[k] - b[k]))) for k in ("u", "v", "phi", "us", "div")},
          "cells differing in v:", int((ref["v"] != b["v"]).sum()), "first:", np.argwhere(ref["v"] != b["v"])[:3].tolist())
b = run("k23", {"B200CFD_K23_LEAN": "0"})
for k in ("us", "vs", "div", "phi", "u", "v"):
    w = np.argwhere(ref[k] != b[k])
    print(k, len(w), "rows", sorted(set(w[:, 0].tolist()))[:20], "cols", sorted(set(w[:, 1].tolist()))[:20])
    for i, j in w[:4]:
        print("   ", i, j, repr(float(ref[k][i, j])), repr(float(b[k][i, j])))
d = ref["div"]
nz = np.argwhere(d != 0)
print("div != 0 cells:", len(nz), "of", d.size, "cols range", nz[:, 1].min(), nz[:, 1].max(), "rows range", nz[:, 0].min(), nz[:, 0].max())
print("max|div| per column band:", [float(np.abs(d[:, c:c + 128]).max()) for c in range(0, d.shape[1], 128)])
print("|div| around (16,1326):", np.abs(d[14:19, 1322:1330]))
