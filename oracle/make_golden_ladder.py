"""TEST INFRASTRUCTURE — golden vectors from the UNMODIFIED reference (this container only), second batch:

    NUMBA_NUM_THREADS=1 python -m oracle.make_golden_ladder

  steps_180x600_ladder_f64.npz   SURVEY.md section 4's step ladder: ONE run of the shipped 600x180 grid in float64 from
                                 step 0 through step 3000 without ever resetting `step` (it crosses the dt switch
                                 v5.py:320-326, the perturbation ramp :351 and the IBM ramp :406 at step 1000 the natural
                                 way); u, v, phi decimated [::4, ::4] after 999, 1001, 1100 and 3000 steps, every dt, every
                                 energy_history entry.
  steps_46x150_adaptive_{f64,f32}.npz   15 steps across step 1000 with cfl_target = 0.001, so that after the switch dt is
                                 the CFL value strictly between dt_min and dt_max (the unclipped branch of
                                 adaptive_time_step, v5.py:322-326, in both field dtypes).
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

os.environ.setdefault("NUMBA_NUM_THREADS", "1")
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import refharness  # noqa: E402
from oracle.make_golden import OUT, meta, step_run  # noqa: E402

LADDER = (999, 1001, 1100, 3000)
STRIDE = 4


def ladder(m):
    cfg = m.OptimizedTurbulentConfig(nx=600, ny=180, memory_efficient=False)
    s = m.OptimizedTurbulentSolver(cfg)
    out, dts = {}, []
    for k in range(1, LADDER[-1] + 1):
        dts.append(float(s.time_step()))
        if k in LADDER:
            for name in ("u", "v", "phi"):
                out[f"{name}_{k}"] = np.ascontiguousarray(getattr(s, name)[::STRIDE, ::STRIDE])
    out["dts"] = np.array(dts)
    out["energy"] = np.array([float(e) for _, e in s.energy_history])
    out["ladder"] = np.array(LADDER)
    out["stride"] = STRIDE
    return out


def main():
    m = refharness.load_v5()
    for f32, tag in ((False, "f64"), (True, "f32")):
        d = step_run(m, 150, 46, f32, 995, 15, {15}, cfl_target=0.001)
        assert 1e-6 < d["dts"][-1] < 1e-4, d["dts"]
        np.savez_compressed(OUT / f"steps_46x150_adaptive_{tag}.npz", meta=meta(dtype=tag, nx=150, ny=46, cfl_target=0.001), **d)
    d = ladder(m)
    np.savez_compressed(OUT / "steps_180x600_ladder_f64.npz", meta=meta(dtype="f64", nx=600, ny=180, steps=3000), **d)
    print("ladder goldens written to", OUT)


if __name__ == "__main__":
    main()
