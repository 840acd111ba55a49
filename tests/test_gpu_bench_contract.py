"""GPU: the JSON line bench.py prints keeps the driver's contract (keys, units, roofline / e2e / cpu_baseline objects) —
run on a small grid so the whole check takes seconds."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def _run(*args):
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), *args], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    return json.loads(lines[0])


def test_bench_line_contract():
    d = _run("--steps", "5", "--warmup", "3", "--nx", "1024", "--ny", "256", "--cpu-budget", "1", "--no-scaling-base",
             "--poisson-n", "1024", "--poisson-sweeps", "24")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 5 and d["warmup"] == 3 and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["dtype"] == "f64" and d["data"] == "synthetic" and "workload" in d["config"]
    assert d["value"] > 0 and d["gpu_launches"] == 6 * 5                   # six kernels of ours per step
    po = d["poisson"]["modes"]
    for k in ("jacobi_fixed", "rbsor_fixed", "rbsor_tolerance"):
        assert po[k]["GLUPS"] > 0 and po[k]["frac"] > 0, k
    assert po["jacobi_fixed"]["depth"] >= 2 and po["rbsor_tolerance"]["iterations"] >= 1
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-12
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 2 * 1024 * 256 * 8 and e["d2h_bytes_per_step"] >= e["h2d_bytes_per_step"]
    assert e["value"] < d["value"]                                         # PCIe both ways every step
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] >= 1 and c["value"] > 0 and "sample" in c
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}


def test_reference_arm_contract():
    d = _run("--impl", "reference", "--steps", "2", "--warmup", "1")
    assert d["impl"] == "reference" and d["value"] > 0 and d["unit"] == "Mcell-updates/s"
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"]
    assert d["cpu_baseline"]["cores"] >= 1
    # both arms print the same workload description (the driver's same_config check)
    import argparse
    import bench
    ns = argparse.Namespace(nx=bench.NX, ny=bench.NY, slab_nx=16384, slab_ny=16384)
    assert d["config"] == bench.make_config(1, ns)
