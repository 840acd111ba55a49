"""GPU: the chunked driver (device-resident advance + asynchronous snapshots) against the reference's step-by-step loop
(v5.py:642-670) on the same CUDA solver, and the device vorticity against compute_vorticity (v5.py:365-373)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cfg(tmp_path, f32=False, **kw):
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    return OptimizedTurbulentConfig(nx=150, ny=46, memory_efficient=f32, output_dir=str(tmp_path / "out"),
                                    hdf5_file="run.h5", save_interval=20, **kw)


class _StepwiseView:
    """Hides advance()/scalars so run_simulation falls back to one time_step() call per step, like main()."""

    def __init__(self, s):
        self._s = s
        self.config = s.config

    def time_step(self):
        return self._s.time_step()

    def health(self):
        return self._s.health()

    u = property(lambda self: self._s.u)
    v = property(lambda self: self._s.v)
    X = property(lambda self: self._s.X)
    Y = property(lambda self: self._s.Y)

    def compute_vorticity(self):
        return self._s.compute_vorticity()


@pytest.mark.parametrize("f32", [False, True])
def test_device_vorticity_matches_reference_expression(tmp_path, f32):
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    s = OptimizedTurbulentSolver(_cfg(tmp_path, f32))
    s.advance(7)
    w = s.compute_vorticity_device().cpu().numpy()
    ref = s.compute_vorticity()
    assert np.array_equal(np.isnan(w), np.isnan(ref)) and np.isnan(w).any()
    ok = ~np.isnan(ref)
    assert np.abs(w[ok] - ref[ok]).max() <= (1e-12 if not f32 else 1e-5) * np.abs(ref[ok]).max()
    s.close()


def test_chunked_driver_equals_stepwise_loop(tmp_path):
    from cfd_simulations_b200.driver import run_simulation
    from cfd_simulations_b200.snapshot import SnapshotWriter, list_steps, load_snapshot
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    a = OptimizedTurbulentSolver(_cfg(tmp_path / "a"))
    b = OptimizedTurbulentSolver(_cfg(tmp_path / "b"))
    with SnapshotWriter(a.config, backend="npz") as wa:
        ra = run_simulation(a, writer=wa, max_steps=45)
    with SnapshotWriter(b.config, backend="npz") as wb:
        rb = run_simulation(_StepwiseView(b), writer=wb, max_steps=45)
    assert ra.steps == rb.steps == 45 and ra.saved_steps == rb.saved_steps == [0, 20, 40]
    assert ra.dts == rb.dts and ra.current_time == rb.current_time
    pa, pb = tmp_path / "a" / "out" / "run.h5", tmp_path / "b" / "out" / "run.h5"
    assert list_steps(pa) == list_steps(pb) == [0, 20, 40]
    for st in (0, 20, 40):
        sa, sb = load_snapshot(pa, st), load_snapshot(pb, st)
        # the asynchronous copy captured the state of step `st`, not a later one
        assert np.array_equal(sa["u"], sb["u"]) and np.array_equal(sa["v"], sb["v"]) and sa["time"] == sb["time"]
        ok = ~np.isnan(sb["vorticity"])
        assert np.array_equal(np.isnan(sa["vorticity"]), ~ok)
        assert np.abs(sa["vorticity"][ok] - sb["vorticity"][ok]).max() <= 1e-12 * np.abs(sb["vorticity"][ok]).max()
    assert np.array_equal(a.u, b.u)
    a.close(); b.close()


def test_chunked_driver_stops_on_the_reference_step(tmp_path):
    from cfd_simulations_b200.driver import run_simulation
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    T = 100.5 * float(np.float32(2e-5))
    a = OptimizedTurbulentSolver(_cfg(tmp_path))
    b = OptimizedTurbulentSolver(_cfg(tmp_path))
    ra = run_simulation(a, T_total=T)
    rb = run_simulation(_StepwiseView(b), T_total=T)
    assert ra.steps == rb.steps and ra.current_time == rb.current_time >= T
    assert abs(ra.steps - 101) <= 1           # float32 clock (NumPy 2 promotion of `current_time += dt`)
    assert np.array_equal(a.u, b.u)
    a.close(); b.close()


def test_long_run_drag_lift_probe_match_oracle_series():
    """north_star acceptance: drag / lift (wake-column proxies, diagnostics.py) and the shedding probe must agree with the
    reference path within 0.5 %.  tests/golden/longrun_c1_f32_oracle.npz holds the oracle's series for the SHIPPED case
    (600x180, float32, Re 600) sampled every 25 steps over 150 000 steps (tools/long_run.py --impl oracle); the CUDA solver
    re-runs the first 100 000 steps here.  Measured on B200 over all 150 000 steps: max deviation 1e-4 of the series' scale."""
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root / "tools"))
    import long_run
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from cfd_simulations_b200.diagnostics import strouhal
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    g = np.load(root / "tests" / "golden" / "longrun_c1_f32_oracle.npz")
    steps, every = 100000, int(g["every"])
    n = steps // every
    cfg = OptimizedTurbulentConfig()
    s = OptimizedTurbulentSolver(cfg)
    out = long_run.sample_run(s, cfg, steps, every, True)
    s.close()
    assert np.array_equal(out["step"], g["step"][:n])
    assert np.allclose(out["t"], g["t"][:n], rtol=1e-12)                 # same dt sequence
    for k in ("cd", "cl", "vprobe"):
        scale = np.abs(g[k][:n]).max()
        assert np.abs(out[k] - g[k][:n]).max() <= 5e-3 * scale, k
    a = strouhal(out["t"], out["vprobe"], cfg, t_discard=5.0)
    b = strouhal(g["t"][:n], g["vprobe"][:n], cfg, t_discard=5.0)
    assert abs(a["St"] - b["St"]) <= 5e-3 * abs(b["St"])


def test_snapshot_of_device_setup_solver(tmp_path):
    """setup="device" keeps no host mask / grid: the snapshot must still hold the device vorticity (NaN only inside the
    cylinder) and the global X, Y of the reference layout (v5.py:454-470)."""
    from cfd_simulations_b200.snapshot import SnapshotWriter, load_snapshot
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    cfg = _cfg(tmp_path)
    a = OptimizedTurbulentSolver(cfg, setup="device")
    b = OptimizedTurbulentSolver(cfg)
    a.advance(7); b.advance(7)
    assert a.cylinder_mask is None
    wa = a.compute_vorticity()
    wb = b.compute_vorticity()
    assert np.array_equal(np.isnan(wa), np.isnan(wb)) and 0 < np.isnan(wa).sum() < wa.size // 4
    with SnapshotWriter(cfg, backend="npz") as w:
        w.save_data_to_hdf5(a, 7, 1.4e-4).result()
        s = load_snapshot(w.path, 7)
    assert s["X"].shape == (cfg.ny, cfg.nx) and s["Y"].shape == (cfg.ny, cfg.nx)
    assert np.allclose(s["X"][0], np.linspace(cfg.x_min, cfg.x_max, cfg.nx)) and np.array_equal(np.isnan(s["vorticity"]), np.isnan(wb))
    ok = ~np.isnan(wb)
    assert np.abs(s["vorticity"][ok] - wb[ok]).max() <= 1e-6 * np.abs(wb[ok]).max()   # device setup: transcendental ulps
    a.close(); b.close()


def test_without_the_by_product_attributes_the_state_is_bit_identical():
    """materialize_attributes=False: tau_supg / div_u_star are not written (NULL in b200cfd_bind_fields); everything else is
    the same computation."""
    from cfd_simulations_b200._lib import B200CFDError
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    cfg = OptimizedTurbulentConfig(nx=768, ny=256, memory_efficient=False)
    a = OptimizedTurbulentSolver(cfg, device=0)
    b = OptimizedTurbulentSolver(cfg, device=0, materialize_attributes=False)
    a.advance(15); b.advance(15)
    for k in ("u", "v", "phi", "u_star", "v_star"):
        assert np.array_equal(getattr(a, k), getattr(b, k)), k
    assert [e for _, e in a.energy_history] == [e for _, e in b.energy_history]
    with pytest.raises(B200CFDError):
        b.tau_supg
    with pytest.raises(B200CFDError):
        b.div_u_star
    a.close(); b.close()
