"""CPU: the C oracle of the Euler step (oracle/cfd_oracle_euler.c) against outputs of the reference itself
(tests/golden/euler_*.npz, minted by oracle/make_golden_euler.py from python/shockwave_analysis/v1_shock.py)."""
import json
from pathlib import Path

import numpy as np
import pytest

from cfd_simulations_b200 import shockwave_setup as ss
from oracle import orc_euler as oe

GOLD = Path(__file__).parent / "golden"
TOL_F32 = 1e-5  # north_star: fp32 fields within 1e-5 relative L2


def rel(a, b):
    return float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / max(np.linalg.norm(b.astype(np.float64)), 1e-300))


def load(name):
    z = np.load(GOLD / name)
    return z, json.loads(str(z["meta"]))


@pytest.mark.parametrize("kind", ["hllc", "roe"])
def test_riemann_solvers(kind):
    z, m = load("euler_faces.npz")
    for tag, nx_ in (("x", 1.0), ("y", 0.0)):
        got = oe.riemann(kind, z["UL"], z["UR"], m["gamma"], nx_, 1.0 - nx_, m["epsilon"], m["max_val"])
        ref = z[f"{kind}_{tag}"]
        assert rel(got, ref) <= 1e-7, (kind, tag, rel(got, ref))
        # the restatement rounds where the reference rounds: essentially every component is bit-equal
        assert np.mean(got == ref) > 0.999, (kind, tag, np.mean(got == ref))
        assert np.all(got[700:720] == 0)          # vacuum early returns (v1_shock.py:152-153,160-161)


def test_physical_fluxes():
    z, m = load("euler_faces.npz")
    F, G = oe.compute_fluxes(z["UL"], m["gamma"], m["epsilon"], m["max_val"])
    assert np.array_equal(F, z["F_phys"]) and np.array_equal(G, z["G_phys"])


def test_update_state_and_compute_dt():
    z, _ = load("euler_update.npz")
    nx, ny = z["U"].shape[:2]
    cfg = ss.ShockwaveConfig(nx=nx, ny=ny)
    c = oe.make_cfg(cfg)
    Un = oe.update_state(c, z["U"], z["F"], z["G"], float(z["dt"]))
    assert rel(Un, z["U_new"]) <= 1e-7
    assert np.mean(Un == z["U_new"]) > 0.999
    assert oe.compute_dt(c, z["U"]) == pytest.approx(float(z["compute_dt"]), rel=1e-14)


@pytest.mark.parametrize("name", ["euler_steps_80x40", "euler_steps_80x40_perturbed", "euler_steps_61x33_reflect",
                                  "euler_steps_48x24_roe", "euler_steps_400x200"])
def test_time_steps(name):
    z, m = load(name + ".npz")
    cfg = ss.ShockwaveConfig(solver_type=m["solver_type"], **m["config"])
    x, y = ss.grid(cfg)
    wm = ss.wedge_mask(cfg, x, y)
    assert np.array_equal(wm, z["wedge_mask"])
    nr = ss.n_reflect(cfg, x, y)
    assert (nr > 0) == ("reflect" in name)
    if m["perturb_seed"] is None:
        assert np.array_equal(ss.initial_state(cfg), z["U0"])
    c = oe.make_cfg(cfg, nr)
    U = z["U0"].copy()
    done = 0
    dts = []
    for n in m["snaps"]:
        dts += list(oe.step(c, U, wm, n - done))
        done = n
        assert rel(U, z[f"U{n}"]) <= TOL_F32, (name, n, rel(U, z[f"U{n}"]))
    np.testing.assert_allclose(np.array(dts), z["dts"], rtol=1e-6)
