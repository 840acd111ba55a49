"""Function-level mirrors of the reference's @njit kernels (v5.py:96-257), each one CUDA launch through the
C ABI.  Arguments are NumPy arrays (uploaded, result downloaded — convenient for parity tests) or torch CUDA
tensors (no copies).  Same names and argument order as the reference functions.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import Config, check

_handles = {}


def _handle(ny, nx, np_dtype, dx, dy, device=0, initial_steps=1000, V_inf=1.0, y_max=4.0):
    """A cached kernel-only handle for (shape, dtype, spacing)."""
    key = (ny, nx, np.dtype(np_dtype).str, float(dx), float(dy), device, initial_steps, float(V_inf), float(y_max))
    h = _handles.get(key)
    if h is None:
        if not torch.cuda.is_available():
            raise _lib.B200CFDError("no CUDA device: cfd_simulations_b200 has no CPU fallback")
        lib = _lib.load()
        cfg = Config(ny=ny, nx=nx, dtype=_lib.F32 if np.dtype(np_dtype) == np.float32 else _lib.F64,
                     dx=float(dx), dy=float(dy), nu=1e-3, av=0.0, cfg_dt=5e-5, V_inf=V_inf, y_max=y_max,
                     initial_steps=initial_steps, use_supg=1, clean_iters=2, nx_global=nx, nranks=1,
                     max_velocity=5.0, omega=1.0, cfl_target=0.1, dt_min=1e-6, dt_max=1e-4, dt_base=5e-5)
        h = C.c_void_p()
        check(lib.b200cfd_create(C.byref(cfg), device, C.byref(h)))
        _handles[key] = h
    return h


def _dev(a, dtype=None, device=0):
    if isinstance(a, torch.Tensor):
        return a
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(torch.device("cuda", device))


def _out(t, like):
    return t if isinstance(like, torch.Tensor) else t.cpu().numpy()


def _stream(device=0):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _npdt(a):
    if isinstance(a, torch.Tensor):
        return np.float32 if a.dtype == torch.float32 else np.float64
    return a.dtype


def compute_divergence_fast(u, v, dx, dy):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv = _dev(u), _dev(v)
    out = torch.empty_like(du)
    check(_lib.load().b200cfd_k_divergence(h, du.data_ptr(), dv.data_ptr(), out.data_ptr(), _stream()), h)
    return _out(out, u)


def compute_gradient_fast(phi, dx, dy):
    ny, nx = phi.shape
    h = _handle(ny, nx, _npdt(phi), dx, dy)
    dp = _dev(phi)
    gx, gy = torch.empty_like(dp), torch.empty_like(dp)
    check(_lib.load().b200cfd_k_gradient(h, dp.data_ptr(), gx.data_ptr(), gy.data_ptr(), _stream()), h)
    return _out(gx, phi), _out(gy, phi)


def compute_supg_stabilization_fast(u, v, dx, dy, dt, nu_eff):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv = _dev(u), _dev(v)
    out = torch.empty_like(du)
    ne = float(nu_eff.flatten()[0]) if hasattr(nu_eff, "flatten") else float(nu_eff)
    check(_lib.load().b200cfd_k_supg_tau(h, du.data_ptr(), dv.data_ptr(), out.data_ptr(), float(dt), ne, _stream()), h)
    return _out(out, u)


def compute_convection_supg_fast(u, v, phi, dx, dy, tau_supg):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv, dp, dtau = _dev(u), _dev(v), _dev(phi), _dev(tau_supg)
    out = torch.empty_like(du)
    check(_lib.load().b200cfd_k_convection(h, du.data_ptr(), dv.data_ptr(), dp.data_ptr(), dtau.data_ptr(),
                                           out.data_ptr(), 1, _stream()), h)
    return _out(out, u)


def compute_convection_fast(u, v, phi, dx, dy):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv, dp = _dev(u), _dev(v), _dev(phi)
    out = torch.empty_like(du)
    check(_lib.load().b200cfd_k_convection(h, du.data_ptr(), dv.data_ptr(), dp.data_ptr(), None, out.data_ptr(), 0,
                                           _stream()), h)
    return _out(out, u)


def compute_laplacian_fast(phi, dx, dy, nu_eff):
    ny, nx = phi.shape
    h = _handle(ny, nx, _npdt(phi), dx, dy)
    dp = _dev(phi)
    out = torch.empty_like(dp)
    ne = float(nu_eff.flatten()[0]) if hasattr(nu_eff, "flatten") else float(nu_eff)
    check(_lib.load().b200cfd_k_laplacian(h, dp.data_ptr(), out.data_ptr(), ne, _stream()), h)
    return _out(out, phi)


def compute_smagorinsky_viscosity_fast(u, v, dx, dy, cs):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv = _dev(u), _dev(v)
    out = torch.empty_like(du)
    check(_lib.load().b200cfd_k_smagorinsky(h, du.data_ptr(), dv.data_ptr(), out.data_ptr(), float(cs), _stream()), h)
    return _out(out, u)


def apply_boundary_conditions(u, v, y, y_max, V_inf, step, handle=None):
    """v5.py:349-360; in place (NumPy arrays are uploaded, updated, copied back)."""
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), 1.0, 1.0, V_inf=V_inf, y_max=y_max)
    du, dv = _dev(u), _dev(v)
    dy_ = _dev(np.ascontiguousarray(y, dtype=np.float64))
    check(_lib.load().b200cfd_k_apply_bc(h, du.data_ptr(), dv.data_ptr(), dy_.data_ptr(), int(step), _stream()), h)
    if not isinstance(u, torch.Tensor):
        u[...] = du.cpu().numpy()
        v[...] = dv.cpu().numpy()
    return u, v


def apply_ibm_fast(u, v, ibm_mask, force_strength):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), 1.0, 1.0)
    du, dv = _dev(u), _dev(v)
    dm = _dev(ibm_mask, torch.float64)
    check(_lib.load().b200cfd_k_apply_ibm(h, du.data_ptr(), dv.data_ptr(), dm.data_ptr(), float(force_strength),
                                          _stream()), h)
    if not isinstance(u, torch.Tensor):
        u[...] = du.cpu().numpy()
        v[...] = dv.cpu().numpy()
    return u, v


def clean_divergence_fast(u, v, dx, dy, iterations=2, mode="lexicographic", return_psi=False):
    ny, nx = u.shape
    h = _handle(ny, nx, _npdt(u), dx, dy)
    du, dv = _dev(u), _dev(v)
    if du is u:
        pass
    psi = torch.zeros_like(du) if return_psi else None
    check(_lib.load().b200cfd_k_clean_divergence(h, du.data_ptr(), dv.data_ptr(), int(iterations),
                                                 _lib.CLEANDIV_MODES[mode], psi.data_ptr() if return_psi else None,
                                                 _stream()), h)
    if not isinstance(u, torch.Tensor):
        u[...] = du.cpu().numpy()
        v[...] = dv.cpu().numpy()
    if return_psi:
        return u, v, _out(psi, u)
    return u, v


def solve_pressure_gauss_seidel_fast(phi, div_u_star, dx, dy, dt, mask, iterations, tolerance, omega=1.0,
                                     mode="as_written", temporal_depth=0, return_info=False):
    """v5.py:202-226.  mode="as_written" (true max reduction, tolerance break) or "as_executed" (1 iteration)."""
    ny, nx = phi.shape
    h = _handle(ny, nx, _npdt(phi), dx, dy)
    dp, dd = _dev(phi), _dev(div_u_star)
    dm = _dev(np.ascontiguousarray(mask, dtype=np.uint8)) if mask is not None and not isinstance(mask, torch.Tensor) else mask
    it, last = C.c_int(0), C.c_double(0.0)
    check(_lib.load().b200cfd_poisson(h, dp.data_ptr(), dd.data_ptr(), dm.data_ptr() if dm is not None else None,
                                      float(dx), float(dy), float(dt), _lib.POISSON_MODES[mode], int(iterations),
                                      float(tolerance), float(omega), int(temporal_depth), C.byref(it), C.byref(last),
                                      _stream()), h)
    if not isinstance(phi, torch.Tensor):
        phi[...] = dp.cpu().numpy()
    if return_info == "full":
        return phi, it.value, last.value, _poisson_info(h)
    return (phi, it.value, last.value) if return_info else phi


def _poisson_info(h):
    """{depth, passes, pass_bytes_per_cell} of the last b200cfd_poisson call on this handle."""
    out = (C.c_int * 4)()
    check(_lib.load().b200cfd_poisson_info(h, C.byref(out)), h)
    return {"depth": out[0], "passes": out[1], "pass_bytes_per_cell": out[2]} if out[1] else {"depth": 1}


def solve_pressure_jacobi(phi, div_u_star, dx, dt, mask, sweeps, temporal_depth=0, return_info=False):
    """The Jacobi branch of solve_pressure_fast (v5.py:336-346) on arbitrary arrays."""
    ny, nx = phi.shape
    h = _handle(ny, nx, _npdt(phi), dx, dx)
    dp, dd = _dev(phi), _dev(div_u_star)
    dm = _dev(np.ascontiguousarray(mask, dtype=np.uint8)) if mask is not None and not isinstance(mask, torch.Tensor) else mask
    check(_lib.load().b200cfd_poisson(h, dp.data_ptr(), dd.data_ptr(), dm.data_ptr() if dm is not None else None,
                                      float(dx), float(dx), float(dt), _lib.POISSON_JACOBI, int(sweeps), 0.0, 1.0,
                                      int(temporal_depth), None, None, _stream()), h)
    if not isinstance(phi, torch.Tensor):
        phi[...] = dp.cpu().numpy()
    if return_info == "full":
        return phi, int(sweeps), 0.0, _poisson_info(h)
    return (phi, int(sweeps), 0.0) if return_info else phi


def solve_pressure_iterative(div_u_star, dx, dy, dt, mask, max_iter=500, tol=1e-7, omega=1.5, return_info=False):
    """v2.py:284-306 (v1.py:275-297) solve_pressure_iterative: the SOR fallback of the v1/v2 solvers on
    lap(phi) = +div/dt, Neumann left/bottom/top, Dirichlet-0 right, masked cells 0, phi starts at 0, stop when
    max|dphi| < tol (<= max_iter sweeps).  Red-black ordering on the GPU: same converged field as the reference's
    lexicographic sweep."""
    return solve_pressure_gauss_seidel_fast(np.zeros_like(div_u_star) if not isinstance(div_u_star, torch.Tensor)
                                            else torch.zeros_like(div_u_star), div_u_star, dx, dy, dt, mask, max_iter, tol,
                                            omega=omega, mode="sor_v2", temporal_depth=1, return_info=return_info)
