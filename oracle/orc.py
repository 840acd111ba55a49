"""TEST INFRASTRUCTURE — ctypes wrapper around oracle/liborcfd.so (the CPU restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package (cfd_simulations_b200) never does.

Mirrors the reference's call signatures (v5.py:96-257, 316-441) on NumPy arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

POISSON_AS_EXECUTED, POISSON_AS_WRITTEN, POISSON_JACOBI = 0, 1, 2


def build(force: bool = False) -> Path:
    so = _HERE / "liborcfd.so"
    srcs = [_HERE / "cfd_oracle.c", _HERE / "cfd_oracle_impl.h", _HERE / "cfd_oracle_euler.c"]
    if force or not so.exists() or any(s.stat().st_mtime > so.stat().st_mtime for s in srcs):
        subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(str(build()))
        _LIB.orc_max_vorticity_f64.restype = C.c_double
        _LIB.orc_max_vorticity_f32.restype = C.c_double
        _LIB.orc_mean_energy_f64.restype = C.c_double
        _LIB.orc_mean_energy_f32.restype = C.c_double
        _LIB.orc_adaptive_dt_f64.restype = C.c_double
        _LIB.orc_adaptive_dt_f32.restype = C.c_double
        _LIB.orc_time_step_f64.restype = C.c_double
        _LIB.orc_time_step_f32.restype = C.c_double
    return _LIB


def set_threads(n: int) -> int:
    """Set the OpenMP team size of the oracle library and return the size a parallel region really gets."""
    lib().orc_set_threads(int(n))
    return int(lib().orc_get_threads())


class OrcConfig(C.Structure):
    _fields_ = [
        ("ny", C.c_int), ("nx", C.c_int),
        ("dx", C.c_double), ("dy", C.c_double),
        ("nu", C.c_double), ("av", C.c_double), ("cfg_dt", C.c_double),
        ("V_inf", C.c_double), ("y_max", C.c_double),
        ("cfl_target", C.c_double), ("dt_min", C.c_double), ("dt_max", C.c_double),
        ("dt_base", C.c_double), ("max_velocity", C.c_double),
        ("initial_steps", C.c_int), ("use_supg", C.c_int), ("use_les", C.c_int), ("adaptive_dt", C.c_int),
        ("cs", C.c_double),
        ("poisson_mode", C.c_int), ("pressure_iterations", C.c_int),
        ("pressure_tolerance", C.c_double), ("omega", C.c_double),
        ("clean_iters", C.c_int),
        ("pert_amplitude", C.c_double), ("scalar_f32", C.c_int),
    ]


def _sfx(a: np.ndarray) -> str:
    if a.dtype == np.float64:
        return "_f64"
    if a.dtype == np.float32:
        return "_f32"
    raise TypeError(f"unsupported dtype {a.dtype}")


def _p(a):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _d(x):
    return C.c_double(float(x))


def _fn(name, a):
    return getattr(lib(), name + _sfx(a))


def _mask_u8(mask):
    return None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)


# ---- kernel-level mirrors (same names as the reference functions) ---------------------------------

def compute_divergence_fast(u, v, dx, dy):
    ny, nx = u.shape
    out = np.empty_like(u)
    _fn("orc_divergence", u)(_p(u), _p(v), _p(out), ny, nx, _d(dx), _d(dy))
    return out


def compute_gradient_fast(phi, dx, dy):
    ny, nx = phi.shape
    gx, gy = np.empty_like(phi), np.empty_like(phi)
    _fn("orc_gradient", phi)(_p(phi), _p(gx), _p(gy), ny, nx, _d(dx), _d(dy))
    return gx, gy


def compute_supg_stabilization_fast(u, v, dx, dy, dt, nu_eff):
    """nu_eff: scalar or constant array (the reference passes a full array of one constant)."""
    ny, nx = u.shape
    out = np.empty_like(u)
    ne = float(np.asarray(nu_eff).flat[0])
    _fn("orc_supg_tau", u)(_p(u), _p(v), _p(out), ny, nx, _d(dx), _d(dy), _d(dt), _d(ne), None)
    return out


def compute_convection_supg_fast(u, v, phi, dx, dy, tau):
    ny, nx = u.shape
    out = np.empty_like(phi)
    _fn("orc_convection_supg", u)(_p(u), _p(v), _p(phi), _p(tau), _p(out), ny, nx, _d(dx), _d(dy))
    return out


def compute_convection_fast(u, v, phi, dx, dy):
    ny, nx = u.shape
    out = np.empty_like(phi)
    _fn("orc_convection_upwind", u)(_p(u), _p(v), _p(phi), _p(out), ny, nx, _d(dx), _d(dy))
    return out


def compute_laplacian_fast(phi, dx, dy, nu_eff):
    ny, nx = phi.shape
    out = np.empty_like(phi)
    ne = float(np.asarray(nu_eff).flat[0])
    _fn("orc_laplacian", phi)(_p(phi), _p(out), ny, nx, _d(dx), _d(dy), _d(ne), None)
    return out


def compute_smagorinsky_viscosity_fast(u, v, dx, dy, cs):
    ny, nx = u.shape
    out = np.empty_like(u)
    _fn("orc_smagorinsky", u)(_p(u), _p(v), _p(out), ny, nx, _d(dx), _d(dy), _d(cs))
    return out


def apply_boundary_conditions(u, v, y, y_max, V_inf, step):
    ny, nx = u.shape
    y = np.ascontiguousarray(y, dtype=np.float64)
    _fn("orc_apply_bc", u)(_p(u), _p(v), ny, nx, _p(y), _d(y_max), _d(V_inf), C.c_long(int(step)))


def apply_ibm_fast(u, v, ibm_mask, force_strength):
    ny, nx = u.shape
    ibm = np.ascontiguousarray(ibm_mask, dtype=np.float64)
    _fn("orc_apply_ibm", u)(_p(u), _p(v), _p(ibm), ny, nx, _d(force_strength))
    return u, v


def solve_pressure_gauss_seidel(phi, div, dx, dy, dt, mask, iterations, tolerance, omega=1.0):
    """AS WRITTEN semantics (true max reduction).  iterations=1 reproduces the reference as executed.
    Returns (phi, iterations_done, last_change)."""
    ny, nx = phi.shape
    m = _mask_u8(mask)
    last = C.c_double(0.0)
    fn = _fn("orc_rbgs", phi)
    fn.restype = C.c_int
    it = fn(_p(phi), _p(div), _p(m), ny, nx, _d(dx), _d(dy), _d(dt), int(iterations), _d(tolerance),
            _d(omega), C.byref(last))
    return phi, it, last.value


def solve_pressure_jacobi(phi, div, dx, dt, mask, sweeps):
    ny, nx = phi.shape
    m = _mask_u8(mask)
    tmp = np.empty_like(phi)
    _fn("orc_jacobi", phi)(_p(phi), _p(tmp), _p(div), _p(m), ny, nx, _d(dx), _d(dt), int(sweeps))
    return phi


def lexgs_sweep(psi, div, dx, dy):
    ny, nx = psi.shape
    _fn("orc_lexgs_sweep", psi)(_p(psi), _p(div), ny, nx, _d(dx), _d(dy))
    return psi


def clean_divergence_fast(u, v, dx, dy, iterations=2, return_psi=False):
    ny, nx = u.shape
    work = np.empty((4, ny, nx), dtype=u.dtype)
    psi = np.empty_like(u) if return_psi else None
    _fn("orc_clean_divergence", u)(_p(u), _p(v), ny, nx, _d(dx), _d(dy), int(iterations), _p(work), _p(psi))
    return (u, v, psi) if return_psi else (u, v)


def max_vorticity(u, v, mask, dx, dy):
    ny, nx = u.shape
    return _fn("orc_max_vorticity", u)(_p(u), _p(v), _p(_mask_u8(mask)), ny, nx, _d(dx), _d(dy))


def mean_energy(u, v):
    ny, nx = u.shape
    return _fn("orc_mean_energy", u)(_p(u), _p(v), ny, nx)


def sor_v2(phi, div, mask, dx, dy, dt, omega=1.5, max_iter=500, tol=1e-7):
    ny, nx = phi.shape
    last = C.c_double(0.0)
    fn = lib().orc_sor_v2
    fn.restype = C.c_int
    it = fn(_p(phi), _p(div), _p(_mask_u8(mask)), ny, nx, _d(dx), _d(dy), _d(dt), _d(omega), int(max_iter),
            _d(tol), C.byref(last))
    return phi, it, last.value
