"""Host-side mirror of the reference's ShockwaveConfig and solver setup (python/shockwave_analysis/v1_shock.py:35-67,
225-261): same field names and defaults, grid / wedge mask / free-stream initial state, vectorised so that large
grids (16384 x 8192) set up in seconds.  No compute of the hot path lives here."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class ShockwaveConfig:
    """v1_shock.py:35-67 (the plotting / HDF5 fields are kept for drop-in construction and otherwise unused)."""
    L: float = 1.0
    x_min: float = 0.0
    x_max: float = 2.0
    y_min: float = 0.0
    y_max: float = 1.0
    nx: int = 400
    ny: int = 200
    M_inf: float = 2.0
    p_inf: float = 1.0
    rho_inf: float = 1.0
    gamma: float = 1.4
    T_total: float = 5.0
    cfl: float = 0.99
    wedge_angle: float = float(np.deg2rad(10.0))
    wedge_start_x: float = 0.5
    save_interval: int = 200
    output_dir: str = "v1_ma_2"
    hdf5_file: str = "v1_ma_2.h5"
    dpi: int = 200
    memory_efficient: bool = True
    parallel_threads: int = 4
    epsilon: float = 1e-8
    max_val: float = 1e3
    solver_type: str = "hllc"  # "roe" | "hllc"

    def __post_init__(self):
        # v1_shock.py:61-66 — note dx = L/nx although the nodes are linspace(x_min, x_max, nx)
        self.dx = (self.x_max - self.x_min) / self.nx
        self.dy = (self.y_max - self.y_min) / self.ny
        self.a_inf = np.sqrt(self.gamma * self.p_inf / self.rho_inf)
        self.u_inf = self.M_inf * self.a_inf
        if self.solver_type not in ("hllc", "roe"):
            raise ValueError(f"solver_type {self.solver_type!r}")


def grid(cfg: ShockwaveConfig):
    """setup_grid, v1_shock.py:233-237 (1-D coordinates; X[i,j] = x[i], Y[i,j] = y[j])."""
    return np.linspace(cfg.x_min, cfg.x_max, cfg.nx), np.linspace(cfg.y_min, cfg.y_max, cfg.ny)


def wedge_mask(cfg: ShockwaveConfig, x, y, i0: int = 0, i1: int | None = None) -> np.ndarray:
    """setup_wedge, v1_shock.py:239-248, vectorised; rows [i0, i1) of the (nx, ny) mask."""
    xs = x[i0:i1]
    wy = np.tan(cfg.wedge_angle) * (xs - cfg.wedge_start_x)
    return (xs[:, None] >= cfg.wedge_start_x) & (y[None, :] <= wy[:, None])


def n_reflect(cfg: ShockwaveConfig, x, y) -> int:
    """The `if` of apply_boundary_conditions (v1_shock.py:284-285) is evaluated at X[0, j] = x[0]: it is true for the
    columns with y[j] <= tan(angle) (x[0] - wedge_start_x).  y is increasing, so those columns are a prefix; for the
    shipped geometry (x[0] = 0 < wedge_start_x, y >= 0) the prefix is empty."""
    rhs = np.tan(cfg.wedge_angle) * (x[0] - cfg.wedge_start_x)
    flags = y <= rhs
    n = int(flags.sum())
    assert not flags[n:].any()
    return n


def free_stream(cfg: ShockwaveConfig) -> np.ndarray:
    """The four conserved free-stream values of initialize_fields (v1_shock.py:250-261), rounded to f32 on store."""
    u, rho, p = cfg.u_inf, cfg.rho_inf, cfg.p_inf
    E = p / (rho * (cfg.gamma - 1)) + 0.5 * (u ** 2)
    return np.array([rho, rho * u, 0.0, rho * E], dtype=np.float32)


def initial_state(cfg: ShockwaveConfig, nx_rows: int | None = None) -> np.ndarray:
    U = np.empty((cfg.nx if nx_rows is None else nx_rows, cfg.ny, 4), dtype=np.float32)
    U[:] = free_stream(cfg)
    return U
