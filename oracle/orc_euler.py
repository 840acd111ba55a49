"""TEST INFRASTRUCTURE — ctypes wrapper around the Euler part of oracle/liborcfd.so (oracle/cfd_oracle_euler.c),
mirroring the reference's python/shockwave_analysis/v1_shock.py call signatures on NumPy arrays.
Only tests/, __graft_entry__.smoke() and bench_euler.py's cpu_baseline / --impl reference legs may import it."""
from __future__ import annotations

import ctypes as C

import numpy as np

from oracle import orc


class EuCfg(C.Structure):
    _fields_ = [("gamma", C.c_double), ("epsilon", C.c_double), ("max_val", C.c_double), ("cfl", C.c_double),
                ("dx", C.c_double), ("dy", C.c_double), ("rho_inf", C.c_double), ("u_inf", C.c_double),
                ("p_inf", C.c_double), ("nx", C.c_int32), ("ny", C.c_int32), ("solver_type", C.c_int32),
                ("n_reflect", C.c_int32)]


_ready = False


def lib():
    global _ready
    L = orc.lib()
    if not _ready:
        L.eu_compute_dt.restype = C.c_double
        L.eu_compute_dt.argtypes = [C.POINTER(EuCfg), C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.eu_hllc.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_double, C.c_double, C.c_void_p]
        L.eu_roe.argtypes = L.eu_hllc.argtypes
        L.eu_phys_fluxes.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.eu_update_state.argtypes = [C.POINTER(EuCfg), C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
        L.eu_apply_bc.argtypes = [C.POINTER(EuCfg), C.c_void_p, C.c_void_p]
        L.eu_step.argtypes = [C.POINTER(EuCfg), C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        _ready = True
    return L


def make_cfg(cfg, n_reflect=0) -> EuCfg:
    """cfg: any object with ShockwaveConfig's attributes (v1_shock.py:35-67)."""
    return EuCfg(cfg.gamma, cfg.epsilon, cfg.max_val, cfg.cfl, cfg.dx, cfg.dy, cfg.rho_inf, float(cfg.u_inf), cfg.p_inf,
                 cfg.nx, cfg.ny, 0 if cfg.solver_type == "hllc" else 1, n_reflect)


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a


def riemann(kind, UL, UR, gamma, normal_x, normal_y, epsilon, max_val):
    """hllc_solver / roe_solver (v1_shock.py:147-209 / 97-145) on arrays of face states (n, 4)."""
    UL, UR = _f32(UL), _f32(UR)
    out = np.zeros_like(UL)
    fn = lib().eu_hllc if kind == "hllc" else lib().eu_roe
    for k in range(UL.shape[0]):
        fn(UL[k].ctypes.data, UR[k].ctypes.data, gamma, int(normal_x == 1.0), epsilon, max_val, out[k].ctypes.data)
    return out


def compute_fluxes(U, gamma, epsilon, max_val):
    U = _f32(U)
    F, G = np.zeros_like(U), np.zeros_like(U)
    for k in range(U.shape[0]):
        lib().eu_phys_fluxes(U[k].ctypes.data, gamma, epsilon, max_val, F[k].ctypes.data, G[k].ctypes.data)
    return F, G


def update_state(c: EuCfg, U, F, G, dt):
    U, F, G = _f32(U), _f32(F), _f32(G)
    Un = np.zeros_like(U)
    lib().eu_update_state(C.byref(c), U.ctypes.data, F.ctypes.data, G.ctypes.data, dt, Un.ctypes.data)
    return Un


def compute_dt(c: EuCfg, U):
    U = _f32(U)
    return lib().eu_compute_dt(C.byref(c), U.ctypes.data, None, None)


def step(c: EuCfg, U, wedge, n):
    """n x ShockwaveSolver.time_step (v1_shock.py:291-317) in place on U; returns the dt of each step."""
    assert U.dtype == np.float32 and U.flags.c_contiguous
    w = np.ascontiguousarray(wedge, dtype=np.uint8)
    dts = np.zeros(n, np.float64)
    rc = lib().eu_step(C.byref(c), U.ctypes.data, w.ctypes.data, n, dts.ctypes.data)
    assert rc == 0
    return dts
