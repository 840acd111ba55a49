"""Long run of the shipped cylinder case (BASELINE.json configs[0]/[1] physics check): wake drag/lift proxies and the
probe signal every `--every` steps, from the CUDA solver (--impl b200) or the CPU oracle (--impl oracle, test
infrastructure; used here only to mint tests/golden/longrun_*.npz).  Output: an .npz with t, cd, cl, vprobe, step.

  python tools/long_run.py --impl oracle --steps 150000 --every 25 --out tests/golden/longrun_c1_f32_oracle.npz
  python tools/long_run.py --impl b200   --steps 150000 --every 25 --out gpurun_out/longrun_c1_f32_b200.npz
"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from cfd_simulations_b200.config import OptimizedTurbulentConfig          # noqa: E402
from cfd_simulations_b200.diagnostics import drag_lift_proxies, probe_cell, strouhal, wake_column   # noqa: E402


def sample_run(solver, cfg, steps, every, on_device):
    jw = wake_column(cfg)
    ip, jp = probe_cell(cfg)
    T, CD, CL, VP, ST = [], [], [], [], []
    t = 0.0
    done = 0
    while done < steps:
        n = min(every, steps - done)
        if on_device:
            first = len(solver.scalars)
            solver.advance(n)
            solver._sync_scalars()
            t += sum(r["dt"] for r in solver.scalars[first:first + n])
            del solver.scalars[:]                      # keep the host history bounded
            solver._energy_history.clear(); solver._times.clear()
            ut, vt = solver.device_tensor("u"), solver.device_tensor("v")
            ucol, vcol = ut[:, jw].cpu().numpy(), vt[:, jw].cpu().numpy()
            vp = float(vt[ip, jp].item())
        else:
            for _ in range(n):
                t += float(solver.time_step())
            del solver.scalars[:]; del solver.energy_history[:]; del solver.times[:]
            ucol, vcol, vp = solver.u[:, jw].copy(), solver.v[:, jw].copy(), float(solver.v[ip, jp])
        done += n
        cd, cl = drag_lift_proxies(ucol, vcol, cfg)
        T.append(t); CD.append(cd); CL.append(cl); VP.append(vp); ST.append(done)
    return {k: np.asarray(v) for k, v in dict(t=T, cd=CD, cl=CL, vprobe=VP, step=ST).items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", default="b200", choices=["b200", "oracle"])
    ap.add_argument("--steps", type=int, default=150000)
    ap.add_argument("--every", type=int, default=25)
    ap.add_argument("--f64", action="store_true")
    ap.add_argument("--nx", type=int, default=0, help="grid override (0 = the shipped 600x180); with --ny")
    ap.add_argument("--ny", type=int, default=0)
    ap.add_argument("--field-stride", type=int, default=0, help="also store u, v, phi[::s, ::s] of the final state")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    kw = {"nx": a.nx, "ny": a.ny} if a.nx and a.ny else {}
    cfg = OptimizedTurbulentConfig(memory_efficient=not a.f64, **kw)    # default: the shipped case 600x180, Re 600
    t0 = time.time()
    if a.impl == "b200":
        from cfd_simulations_b200.solver import OptimizedTurbulentSolver
        s = OptimizedTurbulentSolver(cfg)
        out = sample_run(s, cfg, a.steps, a.every, True)
    else:
        from oracle.solver import OracleTurbulentSolver
        s = OracleTurbulentSolver(cfg)
        out = sample_run(s, cfg, a.steps, a.every, False)
    wall = time.time() - t0
    st = strouhal(out["t"], out["vprobe"], cfg, t_discard=min(10.0, 0.5 * out["t"][-1]))
    print(f"{a.impl}: {a.steps} steps, t_end {out['t'][-1]:.4f}, wall {wall:.1f} s, St {st}")
    if a.field_stride > 0:
        k = a.field_stride
        for name in ("u", "v", "phi"):
            out["final_" + name] = np.ascontiguousarray(np.asarray(getattr(s, name))[::k, ::k])
    np.savez_compressed(a.out, impl=a.impl, steps=a.steps, every=a.every, wall_s=wall, f64=a.f64, nx=cfg.nx, ny=cfg.ny, **out)


if __name__ == "__main__":
    main()
