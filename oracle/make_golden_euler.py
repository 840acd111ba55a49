"""TEST INFRASTRUCTURE — golden vectors of the compressible-Euler path,
minted by running the UNMODIFIED reference python/shockwave_analysis/v1_shock.py in this container.

    NUMBA_NUM_THREADS=1 python -m oracle.make_golden_euler

Writes tests/golden/euler_*.npz, each with a `meta` JSON string.  The reference has no tests of its own, so these
reference OUTPUTS pin the C oracle (oracle/cfd_oracle_euler.c) and the CUDA path (csrc/euler.cu).
"""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

os.environ.setdefault("NUMBA_NUM_THREADS", "1")
import numba
import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import refharness  # noqa: E402

OUT = ROOT / "tests" / "golden"


def meta(**kw):
    d = dict(numba=numba.__version__, numpy=np.__version__, numba_threads=1,
             reference="python/shockwave_analysis/v1_shock.py")
    d.update(kw)
    return np.array(json.dumps(d))


def random_states(rng, n, gamma=1.4):
    """Conservative states covering subsonic / supersonic both ways, strong jumps and a few vacuum cells."""
    rho = np.exp(rng.uniform(np.log(0.05), np.log(8.0), n))
    p = np.exp(rng.uniform(np.log(0.02), np.log(20.0), n))
    a = np.sqrt(gamma * p / rho)
    mach = rng.uniform(-3.5, 3.5, n)
    ang = rng.uniform(0, 2 * np.pi, n)
    u, v = mach * a * np.cos(ang), mach * a * np.sin(ang)
    E = p / (rho * (gamma - 1)) + 0.5 * (u * u + v * v)
    U = np.stack([rho, rho * u, rho * v, rho * E], axis=1).astype(np.float32)
    return U


def face_vectors(m):
    rng = np.random.default_rng(0)
    n = 3000
    UL, UR = random_states(rng, n), random_states(rng, n)
    # near-equal pairs (smooth flow), vacuum on either side, exact equality
    UR[:600] = (UL[:600] * (1 + 1e-3 * rng.standard_normal((600, 4)))).astype(np.float32)
    UR[600:700] = UL[600:700]
    UL[700:710, 0] = 0.0
    UR[710:720, 0] = 5e-9
    out = dict(UL=UL, UR=UR)
    for name, fn in (("hllc", m.hllc_solver), ("roe", m.roe_solver)):
        fx = np.zeros((n, 4), np.float32)
        fy = np.zeros((n, 4), np.float32)
        for k in range(n):
            fx[k] = fn(UL[k], UR[k], 1.4, 1.0, 0.0, 1e-8, 1e3)
            fy[k] = fn(UL[k], UR[k], 1.4, 0.0, 1.0, 1e-8, 1e3)
        out[name + "_x"], out[name + "_y"] = fx, fy
    F = np.zeros((n, 4), np.float32)
    G = np.zeros((n, 4), np.float32)
    for k in range(n):
        F[k], G[k] = m.compute_fluxes(UL[k], 1.4, 1e-8, 1e3)
    out["F_phys"], out["G_phys"] = F, G
    np.savez_compressed(OUT / "euler_faces.npz", meta=meta(kind="faces", seed=0, gamma=1.4, epsilon=1e-8, max_val=1e3), **out)


def update_vectors(m):
    rng = np.random.default_rng(1)
    nx, ny = 21, 17
    U = random_states(rng, nx * ny).reshape(nx, ny, 4)
    F = (0.5 * rng.standard_normal((nx, ny, 4))).astype(np.float32)
    G = (0.5 * rng.standard_normal((nx, ny, 4))).astype(np.float32)
    dx, dy, dt = 2.0 / nx, 1.0 / ny, 0.0123
    Un = m.update_state(U, F, G, dx, dy, dt, 1e-8, 1e3)
    # compute_dt through the reference method on a duck-typed object
    class S:
        pass
    s = S()
    s.config = m.ShockwaveConfig(nx=nx, ny=ny)
    s.U = U
    dtv = m.ShockwaveSolver.compute_dt(s)
    np.savez_compressed(OUT / "euler_update.npz", meta=meta(kind="update_state+compute_dt", seed=1),
                        U=U, F=F, G=G, dx=dx, dy=dy, dt=dt, U_new=Un, compute_dt=np.float64(dtv))


def run_steps(m, cfg_kw, snaps, perturb_seed=None, solver_type="hllc"):
    cfg = m.ShockwaveConfig(solver_type=solver_type, **cfg_kw)
    s = m.ShockwaveSolver(cfg)
    if perturb_seed is not None:
        rng = np.random.default_rng(perturb_seed)
        s.U = (s.U * (1 + 0.2 * rng.uniform(-1, 1, s.U.shape))).astype(np.float32)
    out = {"U0": s.U.copy(), "wedge_mask": s.wedge_mask.copy()}
    dts = []
    for n in range(1, max(snaps) + 1):
        dts.append(float(s.time_step()))
        if n in snaps:
            out[f"U{n}"] = s.U.copy()
    out["dts"] = np.array(dts, np.float64)
    out["times"] = np.array(s.times, np.float64)
    return cfg, out


def step_vectors(m):
    cases = [
        ("euler_steps_80x40", dict(nx=80, ny=40), (1, 2, 5, 10, 25, 50), None, "hllc"),
        ("euler_steps_80x40_perturbed", dict(nx=80, ny=40), (1, 2, 5, 10, 30), 7, "hllc"),
        ("euler_steps_61x33_reflect", dict(nx=61, ny=33, wedge_start_x=-0.1), (1, 3, 8), 3, "hllc"),
        ("euler_steps_48x24_roe", dict(nx=48, ny=24), (1, 4, 12), 5, "roe"),
        ("euler_steps_400x200", dict(), (1, 5, 20), None, "hllc"),
    ]
    for name, kw, snaps, seed, st in cases:
        cfg, out = run_steps(m, kw, snaps, seed, st)
        np.savez_compressed(OUT / f"{name}.npz",
                            meta=meta(kind="steps", config=kw, snaps=list(snaps), perturb_seed=seed, solver_type=st,
                                      dx=cfg.dx, dy=cfg.dy, u_inf=float(cfg.u_inf)), **out)
        print(name, "dts", out["dts"][:3], flush=True)


def main():
    assert refharness.SHOCK_DIR.exists(), "reference not mounted"
    m = refharness.load_shock()
    OUT.mkdir(parents=True, exist_ok=True)
    face_vectors(m)
    update_vectors(m)
    step_vectors(m)


if __name__ == "__main__":
    main()
