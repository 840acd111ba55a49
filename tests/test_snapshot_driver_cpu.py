"""CPU: the driver loop (main(), v5.py:636-672) and the snapshot layout (v5.py:454-470) on the oracle solver."""
import numpy as np
import pytest

from cfd_simulations_b200.config import OptimizedTurbulentConfig
from cfd_simulations_b200.driver import run_simulation
from cfd_simulations_b200.snapshot import SnapshotWriter, group_name, list_steps, load_snapshot
from oracle.solver import OracleTurbulentSolver


class _Oracle(OracleTurbulentSolver):
    def compute_vorticity(self):                       # v5.py:365-373
        cfg, u, v = self.config, self.u, self.v
        w = np.zeros_like(u)
        w[1:-1, 1:-1] = (v[1:-1, 2:] - v[1:-1, :-2]) / (2 * cfg.dx) - (u[2:, 1:-1] - u[:-2, 1:-1]) / (2 * cfg.dy)
        w[self.cylinder_mask] = np.nan
        return w


def _cfg(tmp_path, **kw):
    return OptimizedTurbulentConfig(nx=150, ny=46, memory_efficient=False, output_dir=str(tmp_path / "out"),
                                    hdf5_file="run.h5", save_interval=20, **kw)


def test_group_names_follow_the_reference():
    assert group_name(0) == "step_000000" and group_name(1234) == "step_001234"


def test_driver_saves_the_reference_layout(tmp_path):
    cfg = _cfg(tmp_path)
    s = _Oracle(cfg)
    with SnapshotWriter(cfg, backend="npz") as w:
        res = run_simulation(s, writer=w, max_steps=45)
    assert res.steps == 45 and res.healthy and res.saved_steps == [0, 20, 40]
    path = tmp_path / "out" / "run.h5"
    assert list_steps(path) == [0, 20, 40]
    # the snapshot named step_000020 is the state AFTER the time step with loop index 20, i.e. after 21 steps
    ref = _Oracle(cfg)
    t = np.float32(0.0)
    for _ in range(21):
        t = np.float32(t + ref.time_step())
    snap = load_snapshot(path, 20)
    assert set(snap) == {"u", "v", "vorticity", "X", "Y", "time"}
    assert np.array_equal(snap["u"], ref.u) and np.array_equal(snap["v"], ref.v)
    assert np.array_equal(snap["vorticity"], ref.compute_vorticity(), equal_nan=True)
    assert np.isnan(snap["vorticity"][ref.cylinder_mask]).all()
    assert np.array_equal(snap["X"], ref.X) and snap["time"] == float(t)
    s0 = load_snapshot(path, 0)
    assert s0["time"] == 0.0 and np.array_equal(s0["u"], _Oracle(cfg).u)     # initial state, not overwritten by step 0


def test_driver_stops_on_the_reference_step(tmp_path):
    cfg = _cfg(tmp_path)
    res = run_simulation(_Oracle(cfg), T_total=3.5 * 2e-5)
    assert res.steps == 4 and res.current_time >= 3.5 * 2e-5      # dt = f32(2e-5) during the first 1000 steps


def test_existing_group_is_not_rewritten(tmp_path):
    cfg = _cfg(tmp_path)
    s = _Oracle(cfg)
    with SnapshotWriter(cfg, backend="npz") as w:
        w.save_data_to_hdf5(s, 7, 1.0).result()
        s.time_step()
        w.save_data_to_hdf5(s, 7, 2.0).result()                     # v5.py:459: `if group_name not in f`
    assert load_snapshot(tmp_path / "out" / "run.h5", 7)["time"] == 1.0


def test_hdf5_backend_needs_h5py(tmp_path):
    try:
        import h5py  # noqa: F401
    except Exception:
        with pytest.raises(RuntimeError):
            SnapshotWriter(_cfg(tmp_path), backend="hdf5")
