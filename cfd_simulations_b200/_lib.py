"""ctypes binding of libb200cfd.so (include/b200cfd.h).  No CPU fallback: if the library or a CUDA device is
missing every compute entry point raises."""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

PKG = Path(__file__).resolve().parent
# B200CFD_LIB: developer knob, an A/B build of the same library (tools/build_variant.sh)
LIB_PATH = Path(os.environ["B200CFD_LIB"]).resolve() if os.environ.get("B200CFD_LIB") else PKG / "libb200cfd.so"

F64, F32 = 0, 1
POISSON_AS_EXECUTED, POISSON_AS_WRITTEN, POISSON_JACOBI, POISSON_SOR_V2 = 0, 1, 2, 3
CLEANDIV_LEXICOGRAPHIC, CLEANDIV_REDBLACK = 0, 1
POISSON_MODES = {"as_executed": POISSON_AS_EXECUTED, "as_written": POISSON_AS_WRITTEN, "jacobi": POISSON_JACOBI,
                 "sor_v2": POISSON_SOR_V2}
CLEANDIV_MODES = {"lexicographic": CLEANDIV_LEXICOGRAPHIC, "redblack": CLEANDIV_REDBLACK}


class B200CFDError(RuntimeError):
    pass


class Config(C.Structure):
    """b200cfd_config (include/b200cfd.h) — mirrors OptimizedTurbulentConfig, v5.py:41-94."""
    _fields_ = [
        ("ny", C.c_int32), ("nx", C.c_int32), ("dtype", C.c_int32),
        ("poisson_mode", C.c_int32), ("pressure_iterations", C.c_int32), ("cleandiv_mode", C.c_int32),
        ("clean_iters", C.c_int32),
        ("initial_steps", C.c_int32), ("use_supg", C.c_int32), ("use_les", C.c_int32), ("adaptive_dt", C.c_int32),
        ("dx", C.c_double), ("dy", C.c_double),
        ("nu", C.c_double), ("av", C.c_double), ("cfg_dt", C.c_double),
        ("V_inf", C.c_double), ("y_max", C.c_double),
        ("cfl_target", C.c_double), ("dt_min", C.c_double), ("dt_max", C.c_double), ("dt_base", C.c_double),
        ("max_velocity", C.c_double),
        ("cs", C.c_double),
        ("pressure_tolerance", C.c_double), ("omega", C.c_double),
        ("rank", C.c_int32), ("nranks", C.c_int32), ("nx_global", C.c_int32), ("x_offset", C.c_int32),
        ("halo_left", C.c_int32), ("halo_right", C.c_int32),
        ("pert_amplitude", C.c_double), ("scalar_f32", C.c_int32), ("variant_set", C.c_int32),
    ]


class StepScalars(C.Structure):
    _fields_ = [(n, C.c_double) for n in (
        "dt", "max_div_pre", "max_grad_phi", "max_div_post", "max_vorticity", "mean_energy", "max_abs_u",
        "max_abs_v", "poisson_iters", "poisson_change", "nonfinite", "step")]


class EulerConfig(C.Structure):
    """b200cfd_euler_config (include/b200cfd_euler.h) — mirrors ShockwaveConfig, v1_shock.py:35-67."""
    _fields_ = [
        ("nx", C.c_int32), ("ny", C.c_int32), ("solver_type", C.c_int32), ("n_reflect", C.c_int32),
        ("rank", C.c_int32), ("nranks", C.c_int32), ("halo_lo", C.c_int32), ("halo_hi", C.c_int32),
        ("exact_division", C.c_int32), ("reserved", C.c_int32),
        ("gamma", C.c_double), ("epsilon", C.c_double), ("max_val", C.c_double), ("cfl", C.c_double),
        ("dx", C.c_double), ("dy", C.c_double), ("rho_inf", C.c_double), ("u_inf", C.c_double), ("p_inf", C.c_double),
    ]


# every symbol include/b200cfd_euler.h declares
EULER_EXPORTS = [
    "b200cfd_euler_last_error", "b200cfd_euler_create", "b200cfd_euler_destroy", "b200cfd_euler_bind",
    "b200cfd_euler_step", "b200cfd_euler_step_host", "b200cfd_euler_step_begin", "b200cfd_euler_step_end",
    "b200cfd_euler_reduction_ptr", "b200cfd_euler_reduction_len", "b200cfd_euler_local_speeds",
    "b200cfd_euler_set_dt_from", "b200cfd_euler_state", "b200cfd_euler_fetch_dts", "b200cfd_euler_launches_per_step",
    "b200cfd_euler_k_riemann", "b200cfd_euler_k_fluxes", "b200cfd_euler_k_update_state", "b200cfd_euler_k_compute_dt",
    "b200cfd_euler_k_apply_bc",
]

# every symbol include/b200cfd.h declares (tests check the library exports all of them)
EXPORTS = [
    "b200cfd_abi_version", "b200cfd_last_error", "b200cfd_device_count", "b200cfd_create", "b200cfd_destroy",
    "b200cfd_bind_fields", "b200cfd_set_step", "b200cfd_get_step", "b200cfd_step", "b200cfd_step_host",
    "b200cfd_launches_per_step", "b200cfd_health", "b200cfd_profile_begin", "b200cfd_profile_end", "b200cfd_get_clean_plan", "b200cfd_fetch_scalars", "b200cfd_poisson", "b200cfd_poisson_info", "b200cfd_poisson_keep_rhs",
    "b200cfd_k_supg_tau", "b200cfd_k_convection", "b200cfd_k_laplacian", "b200cfd_k_divergence",
    "b200cfd_k_gradient", "b200cfd_k_smagorinsky", "b200cfd_k_vorticity", "b200cfd_k_apply_bc", "b200cfd_k_apply_ibm",
    "b200cfd_k_clean_divergence", "b200cfd_k_predictor",
    "b200cfd_required_halo", "b200cfd_step_begin", "b200cfd_step_end", "b200cfd_reduction_ptr", "b200cfd_reduction_len", "b200cfd_poisson_change_ptr",
    "b200cfd_set_step_global", "b200cfd_pack_columns", "b200cfd_unpack_columns",
    "b200cfd_step_splits", "b200cfd_step_part", "b200cfd_halo_rows", "b200cfd_jacobi_static", "b200cfd_stream_plan_host",
]

_lib = None


def load():
    """dlopen the in-tree library (built by cfd_simulations_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise B200CFDError(f"{LIB_PATH} is missing: run `python -m cfd_simulations_b200.build` (nvcc, sm_100a). "
                           "There is no CPU fallback.")
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, i64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_double
    lib.b200cfd_last_error.restype = C.c_char_p
    lib.b200cfd_last_error.argtypes = [vp]
    lib.b200cfd_create.argtypes = [C.POINTER(Config), i32, C.POINTER(vp)]
    lib.b200cfd_destroy.argtypes = [vp]
    lib.b200cfd_bind_fields.argtypes = [vp] * 12
    lib.b200cfd_set_step.argtypes = [vp, i64, vp]
    lib.b200cfd_get_step.argtypes = [vp, C.POINTER(i64)]
    lib.b200cfd_step.argtypes = [vp, i32, i32, vp]
    lib.b200cfd_step_host.argtypes = [vp, vp, vp, i32, i32, C.POINTER(StepScalars), vp]
    lib.b200cfd_launches_per_step.argtypes = [vp]
    lib.b200cfd_profile_begin.argtypes = [vp]
    lib.b200cfd_profile_end.argtypes = [vp, C.c_char_p, i32, C.POINTER(dbl), i32]
    lib.b200cfd_health.argtypes = [vp, C.POINTER(C.c_double * 4), vp]
    lib.b200cfd_get_clean_plan.argtypes = [vp, C.POINTER(C.c_int * 8)]
    lib.b200cfd_get_clean_plan.restype = None
    lib.b200cfd_fetch_scalars.argtypes = [vp, i64, i32, C.POINTER(StepScalars), vp]
    lib.b200cfd_poisson.argtypes = [vp, vp, vp, vp, dbl, dbl, dbl, i32, i32, dbl, dbl, i32, C.POINTER(i32),
                                    C.POINTER(dbl), vp]
    lib.b200cfd_poisson_info.argtypes = [vp, C.POINTER(C.c_int * 4)]
    lib.b200cfd_poisson_keep_rhs.argtypes = [vp, i32]
    lib.b200cfd_jacobi_static.argtypes = [i32, i32, i32, vp, vp, vp, dbl, dbl, i32, vp]
    lib.b200cfd_stream_plan_host.argtypes = [vp, i32, i32, i32, i32, i32, i32, i32, vp]
    lib.b200cfd_k_supg_tau.argtypes = [vp, vp, vp, vp, dbl, dbl, vp]
    lib.b200cfd_k_convection.argtypes = [vp, vp, vp, vp, vp, vp, i32, vp]
    lib.b200cfd_k_laplacian.argtypes = [vp, vp, vp, dbl, vp]
    lib.b200cfd_k_divergence.argtypes = [vp, vp, vp, vp, vp]
    lib.b200cfd_k_gradient.argtypes = [vp, vp, vp, vp, vp]
    lib.b200cfd_k_smagorinsky.argtypes = [vp, vp, vp, vp, dbl, vp]
    lib.b200cfd_k_vorticity.argtypes = [vp, vp, vp, vp, vp, vp]
    lib.b200cfd_k_apply_bc.argtypes = [vp, vp, vp, vp, i64, vp]
    lib.b200cfd_k_apply_ibm.argtypes = [vp, vp, vp, vp, dbl, vp]
    lib.b200cfd_k_clean_divergence.argtypes = [vp, vp, vp, i32, i32, vp, vp]
    lib.b200cfd_k_predictor.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i64, dbl, vp]
    lib.b200cfd_required_halo.argtypes = [C.POINTER(Config), C.POINTER(i32), C.POINTER(i32)]
    lib.b200cfd_step_begin.argtypes = [vp, i32, vp]
    lib.b200cfd_step_end.argtypes = [vp, vp, i32, vp]
    lib.b200cfd_reduction_ptr.argtypes = [vp]
    lib.b200cfd_reduction_ptr.restype = vp
    lib.b200cfd_poisson_change_ptr.argtypes = [vp]
    lib.b200cfd_poisson_change_ptr.restype = vp
    lib.b200cfd_set_step_global.argtypes = [vp, i64, dbl, dbl, vp]
    lib.b200cfd_pack_columns.argtypes = [vp, vp, i32, i32, vp, vp]
    lib.b200cfd_unpack_columns.argtypes = [vp, vp, i32, i32, vp, vp]
    lib.b200cfd_step_splits.argtypes = [vp]
    lib.b200cfd_step_part.argtypes = [vp, i32, i32, vp]
    lib.b200cfd_halo_rows.argtypes = [vp, i32, C.POINTER(vp), i32, i32, i32, vp, i32, i32, vp, vp]
    # ---- include/b200cfd_euler.h ----
    lib.b200cfd_euler_last_error.restype = C.c_char_p
    lib.b200cfd_euler_last_error.argtypes = [vp]
    lib.b200cfd_euler_create.argtypes = [C.POINTER(EulerConfig), i32, C.POINTER(vp)]
    lib.b200cfd_euler_destroy.argtypes = [vp]
    lib.b200cfd_euler_bind.argtypes = [vp, vp, vp, vp, i64, vp]
    lib.b200cfd_euler_step.argtypes = [vp, i32, i32, vp]
    lib.b200cfd_euler_step_host.argtypes = [vp, vp, i32, i32, C.POINTER(dbl), vp]
    lib.b200cfd_euler_step_begin.argtypes = [vp, vp]
    lib.b200cfd_euler_step_end.argtypes = [vp, vp, vp]
    lib.b200cfd_euler_reduction_ptr.argtypes = [vp]
    lib.b200cfd_euler_reduction_ptr.restype = vp
    lib.b200cfd_euler_local_speeds.argtypes = [vp, vp]
    lib.b200cfd_euler_set_dt_from.argtypes = [vp, vp, i64, vp]
    lib.b200cfd_euler_state.argtypes = [vp, C.POINTER(i64), C.POINTER(C.c_int32), C.POINTER(dbl), C.POINTER(i64), vp]
    lib.b200cfd_euler_fetch_dts.argtypes = [vp, i64, i32, C.POINTER(dbl), vp]
    lib.b200cfd_euler_launches_per_step.argtypes = [vp]
    lib.b200cfd_euler_k_riemann.argtypes = [vp, vp, i32, i32, i32, dbl, dbl, dbl, vp, vp]
    lib.b200cfd_euler_k_fluxes.argtypes = [vp, i32, dbl, dbl, dbl, vp, vp, vp]
    lib.b200cfd_euler_k_update_state.argtypes = [vp, vp, vp, vp, dbl, vp, vp]
    lib.b200cfd_euler_k_compute_dt.argtypes = [vp, vp, C.POINTER(dbl), vp]
    lib.b200cfd_euler_k_apply_bc.argtypes = [vp, vp, vp, vp]
    _lib = lib
    return lib


def check(rc: int, handle=None):
    if rc != 0:
        msg = load().b200cfd_last_error(handle)
        raise B200CFDError(f"libb200cfd error {rc}: {msg.decode() if msg else ''}")


def check_euler(rc: int, handle=None):
    if rc != 0:
        msg = load().b200cfd_euler_last_error(handle)
        raise B200CFDError(f"libb200cfd (euler) error {rc}: {msg.decode() if msg else ''}")
