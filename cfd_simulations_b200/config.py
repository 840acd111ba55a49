"""Host-side configuration mirroring the reference's dataclass.

Reference: python/flow_over_cylinder (Fischer)/v5.py:41-94 (OptimizedTurbulentConfig).  Field names,
defaults and the derived quantities of __post_init__ (dx, dy, float32-rounded nu / dt /
artificial_viscosity) are the reference's; logging/psutil side effects are dropped.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np


@dataclass
class OptimizedTurbulentConfig:
    L: float = 1.0
    R_cylinder: float = 0.5
    V_inf: float = 1.0
    cylinder_center: tuple = (4.0, 2.0)
    x_min: float = 0.0
    x_max: float = 20.0
    y_min: float = 0.0
    y_max: float = 4.0
    nx: int = 600
    ny: int = 180
    T_total: float = 30.0
    dt_base: float = 0.00005
    cfl_target: float = 0.1
    adaptive_dt: bool = True
    dt_min: float = 1e-6
    dt_max: float = 0.0001
    Re: float = 600.0
    use_les: bool = False
    smagorinsky_constant: float = 0.0
    use_supg: bool = True
    artificial_viscosity: float = 0.001
    pressure_iterations: int = 1500
    pressure_tolerance: float = 1e-8
    max_velocity: float = 5.0
    initial_steps: int = 1000
    parallel_threads: int = 4
    use_fast_pressure: bool = True
    memory_efficient: bool = True
    vectorized_ops: bool = True
    save_interval: int = 200
    output_dir: str = "v5_re_600"
    hdf5_file: str = "v5_re_600.h5"
    dpi: int = 200
    # --- not in the reference's dataclass: select the v3 / v4 variants of the same step ---
    perturbation_amplitude: float = 0.01    # inflow perturbation: v5.py:351 and v4.py:349 use 0.01, v3.py:329 uses 0.005
    float32_scalars: bool = True            # v5.py:80-82,321,326 round nu, dt, artificial_viscosity and the step dt to
                                            # float32; v3 / v4 keep Python floats (v4.py:79-80,319,324)

    def __post_init__(self):
        # v5.py:78-83
        self.dx = (self.x_max - self.x_min) / (self.nx - 1)
        self.dy = (self.y_max - self.y_min) / (self.ny - 1)
        if self.float32_scalars:
            self.nu = np.float32(1.0 / self.Re)
            self.dt = np.float32(self.dt_base)
            self.artificial_viscosity = np.float32(self.artificial_viscosity)
        else:                                # v4.py:79-80
            self.nu = 1.0 / self.Re
            self.dt = self.dt_base
        self.parallel_threads = min(self.parallel_threads, os.cpu_count() or 1)

    @property
    def field_dtype(self):
        """v5.py:287 — float32 storage when memory_efficient, else float64."""
        return np.float32 if self.memory_efficient else np.float64
