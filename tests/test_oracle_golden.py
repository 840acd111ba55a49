"""CPU: pin the oracle (oracle/liborcfd.so, a C restatement) against goldens minted from the reference
itself (oracle/make_golden.py; Numba 0.65.0, NUMBA_NUM_THREADS=1).  Tolerances: fp64 <= 1e-12 rel-L2
(north_star), fp32 <= 1e-5; the observed values are ~1e-16 / ~1e-7."""
import json

import numpy as np
import pytest

from conftest import GOLDEN, rel_l2
from oracle import orc
from oracle.solver import OracleTurbulentSolver
from cfd_simulations_b200.config import OptimizedTurbulentConfig

TOL = {"f64": 1e-12, "f32": 1e-5}


def load(name):
    d = np.load(GOLDEN / name)
    return d, json.loads(str(d["meta"]))


@pytest.mark.parametrize("tag", ["f64", "f32"])
@pytest.mark.parametrize("shape", ["12x16", "37x53"])
def test_kernels_match_reference(tag, shape):
    g, meta = load(f"kernels_{shape}_{tag}.npz")
    assert meta["numba_threads"] == 1
    tol = TOL[tag]
    u, v, phi = g["u"], g["v"], g["phi"]
    dx, dy, dt, nu_eff = float(g["dx"]), float(g["dy"]), float(g["dt"]), float(g["nu_eff"])
    tau = orc.compute_supg_stabilization_fast(u, v, dx, dy, dt, nu_eff)
    assert rel_l2(tau, g["tau"]) <= tol
    assert rel_l2(orc.compute_convection_supg_fast(u, v, phi, dx, dy, g["tau"]), g["conv_supg"]) <= tol
    assert rel_l2(orc.compute_convection_fast(u, v, phi, dx, dy), g["conv_upwind"]) <= tol
    assert rel_l2(orc.compute_laplacian_fast(phi, dx, dy, nu_eff), g["lap"]) <= tol
    assert rel_l2(orc.compute_divergence_fast(u, v, dx, dy), g["div"]) <= tol
    gx, gy = orc.compute_gradient_fast(phi, dx, dy)
    assert rel_l2(gx, g["gx"]) <= tol and rel_l2(gy, g["gy"]) <= tol
    assert rel_l2(orc.compute_smagorinsky_viscosity_fast(u, v, dx, dy, 0.17), g["smag"]) <= tol
    ui, vi = u.copy(), v.copy()
    orc.apply_ibm_fast(ui, vi, g["ibm"], 0.37)
    assert rel_l2(ui, g["ibm_u"]) <= tol and rel_l2(vi, g["ibm_v"]) <= tol
    ub, vb = u.copy(), v.copy()
    orc.apply_boundary_conditions(ub, vb, g["y"], 4.0, 1.0, int(g["bc_step"]))
    assert rel_l2(ub, g["bc_u"]) <= tol and rel_l2(vb, g["bc_v"]) <= tol
    # Poisson: as executed == exactly one red-black iteration (SURVEY finding 3)
    p, it, _ = orc.solve_pressure_gauss_seidel(np.zeros_like(u), g["div"], dx, dy, 5e-5 if False else float(np.float32(5e-5)),
                                               g["mask"], 1, 1e-8)
    assert rel_l2(p, g["rbgs_as_executed"]) <= tol
    p, it, _ = orc.solve_pressure_gauss_seidel(np.zeros_like(u), g["div"], dx, dy, float(np.float32(5e-5)),
                                               g["mask"], 40, 0.0)
    assert it == 40 and rel_l2(p, g["rbgs_fixed40"]) <= max(tol, 1e-12)
    p, it, last = orc.solve_pressure_gauss_seidel(np.zeros_like(u), g["div"], dx, dy, float(np.float32(5e-5)),
                                                  g["mask"], 100000, float(g["rbgs_tol_value"]))
    assert it == int(g["rbgs_tol_iters"]), (it, int(g["rbgs_tol_iters"]))
    assert rel_l2(p, g["rbgs_tol"]) <= max(tol, 1e-12)
    uc, vc = u.copy(), v.copy()
    orc.clean_divergence_fast(uc, vc, dx, dy, 2)
    assert rel_l2(uc, g["clean_u"]) <= tol and rel_l2(vc, g["clean_v"]) <= tol


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_jacobi_branch(tag):
    g, _ = load(f"jacobi_24x64_{tag}.npz")
    phi = orc.solve_pressure_jacobi(np.zeros_like(g["div"]), g["div"], float(g["dx"]), float(g["dt"]), g["mask"],
                                    int(g["sweeps"]))
    assert rel_l2(phi, g["phi"]) <= TOL[tag]


def test_sor_v2_bit_exact():
    g, _ = load("sor_v2_24x48.npz")
    phi, it, _ = orc.sor_v2(np.zeros_like(g["div"]), g["div"], g["mask"], float(g["dx"]), float(g["dy"]), float(g["dt"]))
    assert it == 500
    assert np.array_equal(phi, g["phi"])


def _run_steps(name, nx, ny, f32, nsteps, snaps, **kw):
    g, _ = load(name)
    cfg = OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=f32, **kw)
    s = OracleTurbulentSolver(cfg)
    assert np.array_equal(s.u, g["u0"]) or rel_l2(s.u, g["u0"]) < 1e-15
    s.step = int(g["start_step"])
    tol = TOL["f32" if f32 else "f64"]
    for k in range(1, nsteps + 1):
        dt = s.time_step()
        assert float(dt) == float(g["dts"][k - 1])
        if k in snaps:
            assert rel_l2(s.u, g[f"u_{k}"]) <= tol, k
            assert rel_l2(s.v, g[f"v_{k}"]) <= tol * (1 if not f32 else 3), k
            assert rel_l2(s.phi, g[f"phi_{k}"]) <= tol * (1 if not f32 else 10), k
    e = np.array([x for _, x in s.energy_history])
    assert np.allclose(e, g["energy"], rtol=1e-12 if not f32 else 1e-5)


@pytest.mark.parametrize("f32", [False, True])
def test_steps_from0(f32):
    _run_steps(f"steps_46x150_from0_{'f32' if f32 else 'f64'}.npz", 150, 46, f32, 100, {1, 10, 100})


@pytest.mark.parametrize("f32", [False, True])
def test_steps_across_step1000_switch(f32):
    _run_steps(f"steps_46x150_from990_{'f32' if f32 else 'f64'}.npz", 150, 46, f32, 25, {25})


@pytest.mark.parametrize("f32", [False, True])
def test_steps_shipped_grid(f32):
    _run_steps(f"steps_180x600_from0_{'f32' if f32 else 'f64'}.npz", 600, 180, f32, 10, {10})


def test_steps_upwind_branch():
    _run_steps("steps_30x90_upwind_f64.npz", 90, 30, False, 10, {10}, use_supg=False)


@pytest.mark.parametrize("f32", [False, True])
def test_steps_unclipped_adaptive_dt(f32):
    """cfl_target = 0.001: after step 1000 dt is the CFL value, strictly between dt_min and dt_max (the branch of
    adaptive_time_step, v5.py:322-326, that the clipped goldens never exercise); dt compared with == every step."""
    name = f"steps_46x150_adaptive_{'f32' if f32 else 'f64'}.npz"
    g, _ = load(name)
    assert 1e-6 < float(g["dts"][-1]) < 1e-4
    _run_steps(name, 150, 46, f32, 15, {15}, cfl_target=0.001)


def test_step_ladder_3000_fp64():
    """SURVEY section 4's ladder: one float64 run of the shipped 600x180 grid from step 0 through step 3000 (crossing the
    dt switch, the perturbation ramp and the IBM ramp at step 1000 without resetting `step`), u, v, phi against the
    reference's (decimated) fields after 999, 1001, 1100 and 3000 steps, every dt."""
    g, _ = load("steps_180x600_ladder_f64.npz")
    k = int(g["stride"])
    orc.set_threads(4)
    s = OracleTurbulentSolver(OptimizedTurbulentConfig(nx=600, ny=180, memory_efficient=False))
    ladder = [int(x) for x in g["ladder"]]
    for n in range(1, ladder[-1] + 1):
        dt = s.time_step()
        assert float(dt) == float(g["dts"][n - 1]), n
        if n in ladder:
            for name in ("u", "v", "phi"):
                assert rel_l2(getattr(s, name)[::k, ::k], g[f"{name}_{n}"]) <= 1e-12, (name, n)
    e = np.array([x for _, x in s.energy_history])
    assert np.allclose(e, g["energy"], rtol=1e-12)


def _variant(name, f32, check_dt, solver_kw, **cfg_kw):
    g, meta = load(name)
    cfg = OptimizedTurbulentConfig(nx=150, ny=46, memory_efficient=f32, **cfg_kw)
    s = OracleTurbulentSolver(cfg, **solver_kw)
    s.step = int(g["start_step"])
    tol = TOL["f32" if f32 else "f64"]
    for k in range(25):
        dt = s.time_step()
        if check_dt:
            assert float(dt) == float(g["dts"][k]), k
    assert rel_l2(s.u, g["u_25"]) <= tol and rel_l2(s.v, g["v_25"]) <= tol * (1 if not f32 else 3)
    assert rel_l2(s.phi, g["phi_25"]) <= tol * (1 if not f32 else 10) and rel_l2(s.u_star, g["ustar_25"]) <= tol
    return meta


@pytest.mark.parametrize("f32", [False, True])
def test_v4_variant_unrounded_scalars(f32):
    """v4.py = v5.py with Python-float nu / dt and an unrounded adaptive dt (v4.py:79-80,319,324): float32_scalars=False."""
    m = _variant(f"steps_46x150_v4_from990_{'f32' if f32 else 'f64'}.npz", f32, True, {}, float32_scalars=False)
    assert m["reference"].endswith("v4.py")


def test_v3_variant_half_perturbation():
    """v3.py: inflow perturbation 0.005 (v3.py:329), Python-float scalars, its own default Re = 200 (v3.py:41)."""
    m = _variant("steps_46x150_v3_from990_f64.npz", False, False, {}, float32_scalars=False,
                 perturbation_amplitude=0.005, Re=200.0)
    assert m["reference"].endswith("v3.py")
