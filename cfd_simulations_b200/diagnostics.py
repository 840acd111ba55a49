"""Physics diagnostics shared by every implementation.

The reference computes no drag, lift or Strouhal number (SURVEY finding 7), while the acceptance criterion asks for their
agreement within 0.5 %.  They are therefore DEFINED here as functionals of the velocity field only (phi is not a physical
pressure in this scheme) and the identical code is applied to reference / oracle fields and to CUDA fields:

  Cd_w(t) = 2/(U^2 D) * sum_i u[i,jw] (U - u[i,jw]) dy      wake momentum deficit through the grid column jw nearest
                                                             x_c + 3 D
  Cl_w(t) = 2/(U^2 D) * sum_i u[i,jw] v[i,jw] dy             net cross-stream momentum flux through the same column
  St      = f D / U, f = dominant frequency of v at the probe cell nearest (x_c + 2 D, y_c), transient discarded

Everything here is plain NumPy on host arrays (a column and a cell per sample: a few KB cross PCIe, not the fields).
"""
from __future__ import annotations

import numpy as np


def wake_column(cfg, distance_in_diameters: float = 3.0) -> int:
    """Index of the grid column nearest x_c + distance * D."""
    x = np.linspace(cfg.x_min, cfg.x_max, cfg.nx)
    xw = cfg.cylinder_center[0] + distance_in_diameters * 2.0 * cfg.R_cylinder
    return int(np.argmin(np.abs(x - xw)))


def probe_cell(cfg, distance_in_diameters: float = 2.0) -> tuple[int, int]:
    """(row, column) of the cell nearest (x_c + distance * D, y_c)."""
    x = np.linspace(cfg.x_min, cfg.x_max, cfg.nx)
    y = np.linspace(cfg.y_min, cfg.y_max, cfg.ny)
    xp = cfg.cylinder_center[0] + distance_in_diameters * 2.0 * cfg.R_cylinder
    return int(np.argmin(np.abs(y - cfg.cylinder_center[1]))), int(np.argmin(np.abs(x - xp)))


def drag_lift_proxies(u_col, v_col, cfg) -> tuple[float, float]:
    """(Cd_w, Cl_w) from the u and v values of the wake column (1-D arrays of length ny), float64 arithmetic."""
    u = np.asarray(u_col, dtype=np.float64)
    v = np.asarray(v_col, dtype=np.float64)
    U, D = float(cfg.V_inf), 2.0 * float(cfg.R_cylinder)
    k = 2.0 / (U * U * D) * float(cfg.dy)
    return float(k * np.sum(u * (U - u))), float(k * np.sum(u * v))


def strouhal(t, v_probe, cfg, t_discard: float = 10.0) -> dict:
    """Dominant shedding frequency of the probe signal after t_discard, on a uniform resampling of (t, v_probe).

    Returns {'St', 'f', 'amplitude', 'n_periods', 'window'}.  The FFT peak is refined by parabolic interpolation of the
    log-magnitude (Hann window), so two runs of the same limit cycle agree far below the bin width 1/window."""
    t = np.asarray(t, dtype=np.float64)
    s = np.asarray(v_probe, dtype=np.float64)
    keep = t >= t_discard
    t, s = t[keep], s[keep]
    if t.size < 16:
        return {"St": float("nan"), "f": float("nan"), "amplitude": 0.0, "n_periods": 0.0, "window": 0.0}
    n = t.size
    tu = np.linspace(t[0], t[-1], n)
    su = np.interp(tu, t, s)
    su = su - su.mean()
    w = np.hanning(n)
    spec = np.abs(np.fft.rfft(su * w))
    freqs = np.fft.rfftfreq(n, d=(tu[-1] - tu[0]) / (n - 1))
    k = int(np.argmax(spec[1:]) + 1)
    f = freqs[k]
    if 1 <= k < spec.size - 1 and spec[k - 1] > 0 and spec[k + 1] > 0:
        a, b, c = np.log(spec[k - 1]), np.log(spec[k]), np.log(spec[k + 1])
        den = a - 2 * b + c
        if den != 0:
            f = freqs[k] + 0.5 * (a - c) / den * (freqs[1] - freqs[0])
    D, U = 2.0 * float(cfg.R_cylinder), float(cfg.V_inf)
    window = float(tu[-1] - tu[0])
    return {"St": float(f * D / U), "f": float(f), "amplitude": float(np.abs(su).max()),
            "n_periods": float(f * window), "window": window}
