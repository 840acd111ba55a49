"""GPU parity of the compressible-Euler path (csrc/euler.cu through the C ABI of include/b200cfd_euler.h) against
(1) outputs of the reference itself (tests/golden/euler_*.npz) and (2) the C oracle on larger seeded inputs, plus
size-independent properties at BASELINE scale.  Tolerance (north_star): fp32 fields within 1e-5 relative L2."""
import json
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = Path(__file__).parent / "golden"
TOL_F32 = 1e-5


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def load(name):
    z = np.load(GOLD / name)
    return z, json.loads(str(z["meta"]))


@pytest.mark.parametrize("kind", ["hllc", "roe"])
def test_riemann_solvers_vs_reference(kind):
    from cfd_simulations_b200 import shockwave as sw
    z, m = load("euler_faces.npz")
    fn = sw.hllc_solver if kind == "hllc" else sw.roe_solver
    for tag, nx_ in (("x", 1.0), ("y", 0.0)):
        got = fn(z["UL"], z["UR"], m["gamma"], nx_, 1.0 - nx_, m["epsilon"], m["max_val"])
        ref = z[f"{kind}_{tag}"]
        assert rel(got, ref) <= 1e-6, (kind, tag, rel(got, ref))
        assert np.mean(got == ref) > 0.99, (kind, tag, np.mean(got == ref))   # FMA contraction moves a few last bits
        assert np.all(got[700:720] == 0)
    # single-face call signature of the reference
    one = fn(z["UL"][5], z["UR"][5], m["gamma"], 1.0, 0.0, m["epsilon"], m["max_val"])
    assert one.shape == (4,) and rel(one, z[f"{kind}_x"][5]) <= 1e-6


def test_physical_fluxes_update_state_compute_dt_vs_reference():
    from cfd_simulations_b200 import shockwave as sw
    z, m = load("euler_faces.npz")
    F, G = sw.compute_fluxes(z["UL"], m["gamma"], m["epsilon"], m["max_val"])
    assert rel(F, z["F_phys"]) <= 1e-7 and rel(G, z["G_phys"]) <= 1e-7
    z, _ = load("euler_update.npz")
    nx, ny = z["U"].shape[:2]
    Un = sw.update_state(z["U"], z["F"], z["G"], float(z["dx"]), float(z["dy"]), float(z["dt"]), 1e-8, 1e3)
    assert rel(Un, z["U_new"]) <= 1e-7
    s = sw.ShockwaveSolver(sw.ShockwaveConfig(nx=nx, ny=ny), U0=z["U"])
    assert s.compute_dt() == pytest.approx(float(z["compute_dt"]), rel=1e-13)
    s.close()


@pytest.mark.parametrize("exact", [False, True])
@pytest.mark.parametrize("name", ["euler_steps_80x40", "euler_steps_80x40_perturbed", "euler_steps_61x33_reflect",
                                  "euler_steps_48x24_roe", "euler_steps_400x200"])
def test_time_steps_vs_reference(name, exact):
    from cfd_simulations_b200 import shockwave as sw
    z, m = load(name + ".npz")
    cfg = sw.ShockwaveConfig(solver_type=m["solver_type"], **m["config"])
    s = sw.ShockwaveSolver(cfg, U0=z["U0"], exact_division=exact)
    assert np.array_equal(s.wedge_mask, z["wedge_mask"])
    done = 0
    for n in m["snaps"]:
        if n - done == 1:
            dt = s.time_step()                      # the reference's entry point, one host round trip
            assert dt == pytest.approx(float(z["dts"][n - 1]), rel=1e-6)
        else:
            s.advance(n - done)                     # device-resident
        done = n
        r = rel(s.U, z[f"U{n}"])
        assert r <= TOL_F32, (name, n, r)
    np.testing.assert_allclose(np.array(s.dts), z["dts"], rtol=1e-6)
    assert s.step == done and len(s.times) == done
    np.testing.assert_allclose(np.array(s.times), z["times"], rtol=1e-6)
    assert s.check_health()
    s.close()


@pytest.mark.parametrize("shape", [(257, 131), (96, 1000), (1000, 35), (5, 3), (64, 62)])
def test_time_steps_vs_oracle_seeded(shape):
    """Larger / ragged grids (window and chunk remainders, ny not a multiple of 30) against the C oracle."""
    from cfd_simulations_b200 import shockwave as sw
    from oracle import orc_euler as oe
    nx, ny = shape
    cfg = sw.ShockwaveConfig(nx=nx, ny=ny)
    rng = np.random.default_rng(nx * 1000 + ny)
    U0 = (sw.initial_state(cfg) * (1 + 0.15 * rng.uniform(-1, 1, (nx, ny, 4)))).astype(np.float32)
    x, y = sw.grid(cfg)
    wm = sw.wedge_mask(cfg, x, y)
    c = oe.make_cfg(cfg, sw.n_reflect(cfg, x, y))
    Uo = U0.copy()
    dts_o = oe.step(c, Uo, wm, 12)
    s = sw.ShockwaveSolver(cfg, U0=U0)
    s.advance(12)
    assert rel(s.U, Uo) <= TOL_F32, rel(s.U, Uo)
    np.testing.assert_allclose(np.array(s.dts), dts_o, rtol=1e-6)
    # graph replay and plain launches are the same computation
    s2 = sw.ShockwaveSolver(cfg, U0=U0, use_graph=False)
    s2.advance(12)
    assert np.array_equal(s2.U, s.U)
    s.close(); s2.close()


def test_nonfinite_dt_skips_the_step():
    """v1_shock.py:294-296: a non-finite dt returns 0.0 and leaves U and step untouched."""
    from cfd_simulations_b200 import shockwave as sw
    cfg = sw.ShockwaveConfig(nx=40, ny=24)
    U0 = sw.initial_state(cfg)
    U0[7, 9, 1] = np.nan
    s = sw.ShockwaveSolver(cfg, U0=U0)
    assert s.time_step() == 0.0 and s.time_step() == 0.0
    assert s.step == 0 and s.times == []
    assert np.array_equal(s.U, U0, equal_nan=True)
    assert not s.check_health()
    s.close()


def test_host_array_entry_point_matches_resident_path():
    import torch
    from cfd_simulations_b200 import shockwave as sw
    cfg = sw.ShockwaveConfig(nx=120, ny=70)
    rng = np.random.default_rng(11)
    U0 = (sw.initial_state(cfg) * (1 + 0.1 * rng.uniform(-1, 1, (120, 70, 4)))).astype(np.float32)
    a = sw.ShockwaveSolver(cfg, U0=U0)
    a.advance(3)
    b = sw.ShockwaveSolver(cfg)
    Uh = torch.from_numpy(U0.copy()).pin_memory()
    dt = 0.0
    for _ in range(3):
        dt = b.time_step_host(Uh, 1)
    assert np.array_equal(Uh.numpy(), a.U)
    assert dt == a.dts[-1]
    a.close(); b.close()


def test_free_stream_is_preserved_and_large_grid_properties():
    """Size-independent properties at a BASELINE-scale grid (config 5 scaled): without a wedge the free stream is a
    fixed point of the scheme (every face flux equals the physical flux); with the wedge, the state stays finite,
    densities positive, rows upstream of the wedge's domain of influence stay exactly free stream."""
    from cfd_simulations_b200 import shockwave as sw
    cfg = sw.ShockwaveConfig(nx=4096, ny=2048, wedge_start_x=10.0)      # wedge outside the domain
    s = sw.ShockwaveSolver(cfg)
    U0 = s.U
    s.advance(5)
    assert np.array_equal(s.U, U0)
    s.close()
    cfg = sw.ShockwaveConfig(nx=4096, ny=2048)
    s = sw.ShockwaveSolver(cfg)
    s.advance(20)
    U = s.U
    assert s.check_health() and U[..., 0].min() > 0
    i_w = int(np.searchsorted(s.x, cfg.wedge_start_x))
    assert np.array_equal(U[: i_w - 25], U0[: i_w - 25])                # supersonic: no upstream influence beyond the stencil
    assert not np.array_equal(U[i_w:], U0[i_w:])
    assert all(d > 0 for d in s.dts)
    s.close()


@pytest.mark.parametrize("nslabs", [2, 3])
def test_x_slabs_match_single_slab(nslabs):
    """The slab decomposition (1 halo row, MAX-reduced wave speeds) reproduces the single-slab run bit for bit."""
    from cfd_simulations_b200 import shockwave as sw
    cfg = sw.ShockwaveConfig(nx=203, ny=95)
    rng = np.random.default_rng(5)
    U0 = (sw.initial_state(cfg) * (1 + 0.15 * rng.uniform(-1, 1, (203, 95, 4)))).astype(np.float32)
    one = sw.ShockwaveSolver(cfg, U0=U0)
    one.advance(9)
    many = sw.LocalShockSlabs(cfg, nslabs, U0=U0)
    many.advance(9)
    assert np.array_equal(many.gather(), one.U)
    assert many.dts == one.dts
    one.close(); many.close()
