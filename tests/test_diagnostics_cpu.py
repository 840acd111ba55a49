"""CPU: the build-defined physics diagnostics on synthetic inputs."""
import numpy as np

from cfd_simulations_b200.config import OptimizedTurbulentConfig
from cfd_simulations_b200.diagnostics import drag_lift_proxies, probe_cell, strouhal, wake_column


def test_probe_and_wake_locations_on_the_shipped_grid():
    cfg = OptimizedTurbulentConfig()
    x = np.linspace(cfg.x_min, cfg.x_max, cfg.nx)
    y = np.linspace(cfg.y_min, cfg.y_max, cfg.ny)
    jw = wake_column(cfg)
    assert abs(x[jw] - 7.0) <= cfg.dx / 2                     # x_c + 3 D = 4 + 3
    ip, jp = probe_cell(cfg)
    assert abs(x[jp] - 6.0) <= cfg.dx / 2 and abs(y[ip] - 2.0) <= cfg.dy / 2


def test_proxies_of_known_profiles():
    cfg = OptimizedTurbulentConfig()
    ny = cfg.ny
    free = np.ones(ny)
    assert drag_lift_proxies(free, np.zeros(ny), cfg) == (0.0, 0.0)           # no deficit, no cross flow
    u = np.where(np.abs(np.linspace(cfg.y_min, cfg.y_max, ny) - 2.0) < 0.5, 0.5, 1.0)   # top-hat wake, deficit 0.5 over D
    cd, cl = drag_lift_proxies(u, 0.1 * np.ones(ny), cfg)
    n_in = int((u == 0.5).sum())
    assert np.isclose(cd, 2.0 * n_in * 0.25 * cfg.dy) and np.isclose(cl, 2.0 * 0.1 * u.sum() * cfg.dy)
    # float32 inputs are promoted: same numbers
    assert drag_lift_proxies(u.astype(np.float32), (0.1 * np.ones(ny)).astype(np.float32), cfg)[0] == cd


def test_strouhal_recovers_a_known_frequency_below_the_bin_width():
    cfg = OptimizedTurbulentConfig()
    rng = np.random.default_rng(0)
    f0 = 0.2137
    t = np.cumsum(np.full(12000, 2.5e-3))                                       # 30 time units, window 20 after discard
    s = 0.3 * np.sin(2 * np.pi * f0 * t + 0.4) + 0.02 * rng.standard_normal(t.size) + 0.05
    r = strouhal(t, s, cfg, t_discard=10.0)
    assert abs(r["St"] - f0) / f0 < 5e-3 and r["n_periods"] > 4
    short = strouhal(t[:10], s[:10], cfg, t_discard=10.0)
    assert np.isnan(short["St"])
