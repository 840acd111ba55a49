"""TEST INFRASTRUCTURE — OracleTurbulentSolver: the reference's OptimizedTurbulentSolver (v5.py:259-441)
driven by the C restatement.  Same attribute / method names as the reference class."""
from __future__ import annotations

import ctypes as C

import numpy as np

from cfd_simulations_b200 import hostsetup
from cfd_simulations_b200.config import OptimizedTurbulentConfig  # noqa: F401  (re-export)

from . import orc


class OracleTurbulentSolver:
    def __init__(self, config, poisson_mode="as_executed", omega=1.0, clean_iters=2):
        cfg = self.config = config
        self.x, self.y, self.X, self.Y = hostsetup.setup_grid(cfg)
        self.dist, self.cylinder_mask, self.ibm_mask = hostsetup.setup_boundary_masks(cfg, self.X, self.Y)
        dt = cfg.field_dtype
        z = lambda: np.zeros((cfg.ny, cfg.nx), dtype=dt, order="C")
        self.p, self.nu_t, self.tau_supg = z(), z(), z()
        self.u_star, self.v_star, self.div_u_star, self.phi = z(), z(), z(), z()
        self.u, self.v = hostsetup.initialize_potential_flow(cfg, self.X, self.Y, self.dist, self.ibm_mask, dt)
        self.step = 0
        self.energy_history = []
        self.times = []
        self.scalars = []
        if not cfg.use_fast_pressure:
            poisson_mode = "jacobi"
        self.poisson_mode = {"as_executed": orc.POISSON_AS_EXECUTED, "as_written": orc.POISSON_AS_WRITTEN,
                             "jacobi": orc.POISSON_JACOBI}[poisson_mode]
        self._c = orc.OrcConfig(
            ny=cfg.ny, nx=cfg.nx, dx=cfg.dx, dy=cfg.dy, nu=float(cfg.nu), av=float(cfg.artificial_viscosity),
            cfg_dt=float(cfg.dt), V_inf=cfg.V_inf, y_max=cfg.y_max, cfl_target=cfg.cfl_target,
            dt_min=cfg.dt_min, dt_max=cfg.dt_max, dt_base=cfg.dt_base, max_velocity=cfg.max_velocity,
            initial_steps=cfg.initial_steps, use_supg=int(cfg.use_supg), use_les=int(cfg.use_les),
            adaptive_dt=int(cfg.adaptive_dt), cs=cfg.smagorinsky_constant, poisson_mode=self.poisson_mode,
            pressure_iterations=cfg.pressure_iterations, pressure_tolerance=cfg.pressure_tolerance,
            omega=omega, clean_iters=clean_iters,
            pert_amplitude=float(getattr(cfg, "perturbation_amplitude", 0.01)),
            scalar_f32=int(getattr(cfg, "float32_scalars", True)))
        self._mask_u8 = np.ascontiguousarray(self.cylinder_mask, dtype=np.uint8)
        self._work = np.empty((8, cfg.ny, cfg.nx), dtype=dt)
        self._fn = getattr(orc.lib(), "orc_time_step" + ("_f32" if dt == np.float32 else "_f64"))

    def time_step(self):
        scal = (C.c_double * 8)()
        p = orc._p
        dt = self._fn(C.byref(self._c), p(self.u), p(self.v), p(self.phi), p(self.u_star), p(self.v_star),
                      p(self.div_u_star), p(self.tau_supg), p(self.nu_t), p(self.ibm_mask), p(self._mask_u8),
                      p(self.y), C.c_long(self.step), p(self._work), scal)
        s = list(scal)
        self.scalars.append(s)
        self.energy_history.append((self.step, s[5]))
        self.times.append(self.step * dt)
        self.step += 1
        return np.float32(dt) if (self.config.adaptive_dt and getattr(self.config, "float32_scalars", True)) else dt
