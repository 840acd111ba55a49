"""TEST INFRASTRUCTURE — mint golden vectors by running the UNMODIFIED reference in this container.

    NUMBA_NUM_THREADS=1 python -m oracle.make_golden

Writes tests/golden/*.npz.  Every file carries a `meta` JSON string (numba/numpy versions, threads=1,
config, seed).  The reference has no tests/fixtures of its own, so these reference
*outputs* are what pins the oracle and the CUDA path.
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
             reference="python/flow_over_cylinder (Fischer)/v5.py")
    d.update(kw)
    return np.array(json.dumps(d))


def kernel_vectors(m, ny, nx, dtype, seed):
    rng = np.random.default_rng(seed)
    dx, dy = 20.0 / (nx - 1), 4.0 / (ny - 1)
    f = lambda s=1.0: np.ascontiguousarray((s * rng.standard_normal((ny, nx))).astype(dtype))
    u, v, phi = f(), f(0.5), f()
    u[2, 3] = 0.0; v[2, 3] = 0.0          # exercises the |V| <= 1e-10 branch of tau
    dt = np.float32(2e-5)
    nu_eff_s = np.float32(1 / 600.0) + np.zeros((), dtype) + np.float32(0.001)
    nu_eff = np.full((ny, nx), nu_eff_s, dtype=dtype)
    ibm = np.clip(rng.standard_normal((ny, nx)), 0, 1)
    mask = rng.random((ny, nx)) < 0.1
    y = np.linspace(0.0, 4.0, ny)
    out = dict(u=u, v=v, phi=phi, ibm=ibm, mask=mask, y=y, dx=dx, dy=dy, dt=dt, nu_eff=nu_eff_s)
    tau = m.compute_supg_stabilization_fast(u, v, dx, dy, dt, nu_eff)
    out["tau"] = tau
    out["conv_supg"] = m.compute_convection_supg_fast(u, v, phi, dx, dy, tau)
    out["conv_upwind"] = m.compute_convection_fast(u, v, phi, dx, dy)
    out["lap"] = m.compute_laplacian_fast(phi, dx, dy, nu_eff)
    out["div"] = m.compute_divergence_fast(u, v, dx, dy)
    out["gx"], out["gy"] = m.compute_gradient_fast(phi, dx, dy)
    out["smag"] = m.compute_smagorinsky_viscosity_fast(u, v, dx, dy, 0.17)
    ui, vi = u.copy(), v.copy()
    m.apply_ibm_fast(ui, vi, ibm, 0.37)
    out["ibm_u"], out["ibm_v"] = ui, vi
    # boundary conditions via the reference method on a duck-typed object
    class S: pass
    s = S(); s.config = m.OptimizedTurbulentConfig(nx=nx, ny=ny); s.step = 437; s.y = y
    ub, vb = u.copy(), v.copy()
    m.OptimizedTurbulentSolver.apply_boundary_conditions(s, ub, vb)
    out["bc_u"], out["bc_v"], out["bc_step"] = ub, vb, 437
    # Poisson: as executed (parallel=True kernel, == 1 iteration) and as written (sequential re-jit)
    div = out["div"]
    p1 = m.solve_pressure_gauss_seidel_fast(np.zeros_like(u), div, dx, dy, np.float32(5e-5), mask, 1500, 1e-8)
    out["rbgs_as_executed"] = p1
    seq = numba.njit(fastmath=True)(m.solve_pressure_gauss_seidel_fast.py_func)
    out["rbgs_fixed40"] = seq(np.zeros_like(u), div, dx, dy, np.float32(5e-5), mask, 40, 0.0)
    tol = 1e-3 * float(np.abs(div).max()) / 5e-5 * dx * dx
    out["rbgs_tol"] = seq(np.zeros_like(u), div, dx, dy, np.float32(5e-5), mask, 100000, tol)
    out["rbgs_tol_value"] = tol
    # iteration count at which the tolerance break fired (bisect: output stops changing)
    lo, hi = 1, 100000
    ref = out["rbgs_tol"]
    while lo < hi:
        mid = (lo + hi) // 2
        r = seq(np.zeros_like(u), div, dx, dy, np.float32(5e-5), mask, mid, tol)
        if np.array_equal(r, ref): hi = mid
        else: lo = mid + 1
    out["rbgs_tol_iters"] = lo
    # clean divergence (lexicographic GS at 1 thread)
    uc, vc = u.copy(), v.copy()
    m.clean_divergence_fast(uc, vc, dx, dy, 2)
    out["clean_u"], out["clean_v"] = uc, vc
    return out


def jacobi_branch(m, ny, nx, dtype, sweeps):
    cfg = m.OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=(dtype == np.float32),
                                     use_fast_pressure=False, pressure_iterations=sweeps)
    s = m.OptimizedTurbulentSolver(cfg)
    rng = np.random.default_rng(5)
    div = (rng.standard_normal((ny, nx)) * 0.1).astype(dtype)
    phi = s.solve_pressure_fast(div).copy()
    return dict(div=div, phi=phi, mask=s.cylinder_mask, dx=cfg.dx, dt=cfg.dt, sweeps=sweeps)


def step_run(m, nx, ny, f32, start_step, nsteps, snaps, **cfgkw):
    cfg = m.OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=f32, **cfgkw)
    s = m.OptimizedTurbulentSolver(cfg)
    s.step = start_step
    out = dict(u0=s.u.copy(), v0=s.v.copy(), start_step=start_step)
    dts = []
    for k in range(1, nsteps + 1):
        dts.append(float(s.time_step()))
        if k in snaps:
            out[f"u_{k}"], out[f"v_{k}"], out[f"phi_{k}"] = s.u.copy(), s.v.copy(), s.phi.copy()
            out[f"ustar_{k}"] = s.u_star.copy()
    out["dts"] = np.array(dts)
    out["energy"] = np.array([float(e) for _, e in s.energy_history])
    return out


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    m = refharness.load_v5()
    for dtype, tag in ((np.float64, "f64"), (np.float32, "f32")):
        for (ny, nx) in ((12, 16), (37, 53)):
            d = kernel_vectors(m, ny, nx, dtype, seed=ny * 1000 + nx)
            np.savez_compressed(OUT / f"kernels_{ny}x{nx}_{tag}.npz", meta=meta(ny=ny, nx=nx, dtype=tag), **d)
        d = jacobi_branch(m, 24, 64, dtype, 50)
        np.savez_compressed(OUT / f"jacobi_24x64_{tag}.npz", meta=meta(dtype=tag), **d)
        # step runs: from step 0, and across the step-1000 switch of dt / IBM ramp / perturbation ramp
        d = step_run(m, 150, 46, dtype == np.float32, 0, 100, {1, 10, 100})
        np.savez_compressed(OUT / f"steps_46x150_from0_{tag}.npz", meta=meta(dtype=tag, nx=150, ny=46), **d)
        d = step_run(m, 150, 46, dtype == np.float32, 990, 25, {25})
        np.savez_compressed(OUT / f"steps_46x150_from990_{tag}.npz", meta=meta(dtype=tag, nx=150, ny=46), **d)
    # shipped grid (600x180), shipped dtype (f32) and f64, 10 steps
    d = step_run(m, 600, 180, True, 0, 10, {10})
    np.savez_compressed(OUT / "steps_180x600_from0_f32.npz", meta=meta(dtype="f32", nx=600, ny=180), **d)
    d = step_run(m, 600, 180, False, 0, 10, {10})
    np.savez_compressed(OUT / "steps_180x600_from0_f64.npz", meta=meta(dtype="f64", nx=600, ny=180), **d)
    # upwind branch (use_supg=False)
    d = step_run(m, 90, 30, False, 995, 10, {10}, use_supg=False)
    np.savez_compressed(OUT / "steps_30x90_upwind_f64.npz", meta=meta(dtype="f64", nx=90, ny=30, use_supg=False), **d)
    # v2 lexicographic SOR (pure Python in the reference -> tiny grid)
    m2 = refharness.load_v2()
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        cfg2 = m2.TurbulentConfig(nx=48, ny=24)
    if cfg2 is not None:
        with contextlib.redirect_stdout(io.StringIO()):
            s2 = m2.RobustTurbulentSolver(cfg2)
        rng = np.random.default_rng(9)
        div = rng.standard_normal((24, 48)) * 1e-3
        phi = s2.solve_pressure_iterative(div)
        np.savez_compressed(OUT / "sor_v2_24x48.npz", meta=meta(reference="v2.py:284-306"), div=div, phi=phi,
                            mask=s2.cylinder_mask, dx=cfg2.dx, dy=cfg2.dy, dt=cfg2.dt)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
