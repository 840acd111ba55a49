"""TEST INFRASTRUCTURE — golden vectors of the v3 / v4 variants of the same step, from the UNMODIFIED reference scripts
(this container only):

    NUMBA_NUM_THREADS=1 python -m oracle.make_golden_variants

  steps_46x150_v4_from990_{f64,f32}.npz   v4.py (Python-float nu / dt, unrounded adaptive dt; v4.py:79-80,319,324):
                                          25 steps across step 1000
  steps_46x150_v3_from990_f64.npz         v3.py (inflow perturbation 0.005, v3.py:329; Python-float scalars; its default
                                          Re = 200): 25 steps across step 1000
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
from pathlib import Path

os.environ.setdefault("NUMBA_NUM_THREADS", "1")
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import refharness  # noqa: E402
from oracle.make_golden import OUT, meta  # noqa: E402


def run(m, f32, start, n):
    cfg = m.OptimizedTurbulentConfig(nx=150, ny=46, memory_efficient=f32)
    s = m.OptimizedTurbulentSolver(cfg)
    s.step = start
    out = dict(u0=s.u.copy(), v0=s.v.copy(), start_step=start)
    dts = []
    with contextlib.redirect_stdout(io.StringIO()):          # v3 prints per step
        for _ in range(n):
            dts.append(float(s.time_step()) if m.__name__ != "ref_v3" else float((s.time_step(), s.adaptive_time_step())[1]))
    out[f"u_{n}"], out[f"v_{n}"], out[f"phi_{n}"], out[f"ustar_{n}"] = s.u.copy(), s.v.copy(), s.phi.copy(), s.u_star.copy()
    out["dts"] = np.array(dts)
    return out


def main():
    v4 = refharness.load(refharness.CYL_DIR / "v4.py", "ref_v4")
    for f32, tag in ((False, "f64"), (True, "f32")):
        np.savez_compressed(OUT / f"steps_46x150_v4_from990_{tag}.npz",
                            meta=meta(dtype=tag, nx=150, ny=46, reference="python/flow_over_cylinder (Fischer)/v4.py"), **run(v4, f32, 990, 25))
    v3 = refharness.load(refharness.CYL_DIR / "v3.py", "ref_v3")
    np.savez_compressed(OUT / "steps_46x150_v3_from990_f64.npz",
                        meta=meta(dtype="f64", nx=150, ny=46, reference="python/flow_over_cylinder (Fischer)/v3.py"), **run(v3, False, 990, 25))
    print("variant goldens written")


if __name__ == "__main__":
    main()
