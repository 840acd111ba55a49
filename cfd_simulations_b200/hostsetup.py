"""One-time host-side setup of grid, masks and the initial field (stays on the CPU, as in the reference).

Reference: v5.py:269-314 (setup_grid, setup_boundary_masks, initialize_fields, initialize_potential_flow).
initialize_potential_flow is an O(nx*ny) pure-Python loop in the reference (34 s at 4096x1024); here it
is the same arithmetic vectorised with NumPy (tests check it equals the loop on small grids).
"""
from __future__ import annotations

import numpy as np


def setup_grid(cfg, cols=None):
    """v5.py:269-273.  cols = (first global column, number of columns) restricts x to one x-slab."""
    x = np.linspace(cfg.x_min, cfg.x_max, cfg.nx)
    if cols is not None:
        x = x[cols[0]:cols[0] + cols[1]]
    y = np.linspace(cfg.y_min, cfg.y_max, cfg.ny)
    X, Y = np.meshgrid(x, y, indexing="xy")
    return x, y, X, Y


def setup_boundary_masks(cfg, X, Y):
    """v5.py:275-283: bool cylinder mask, float64 Gaussian-band IBM mask."""
    x_c, y_c = cfg.cylinder_center
    dist = np.sqrt((X - x_c) ** 2 + (Y - y_c) ** 2)
    cylinder_mask = dist <= cfg.R_cylinder
    sigma = 2 * cfg.dx
    ibm = np.exp(-((dist - cfg.R_cylinder) / sigma) ** 2)
    ibm = np.where(dist < cfg.R_cylinder, 1.0, np.where(dist < cfg.R_cylinder + 5 * cfg.dx, ibm, 0.0))
    return dist, cylinder_mask, ibm


def initialize_potential_flow(cfg, X, Y, dist, ibm_mask, dtype):
    """v5.py:299-314, vectorised: potential flow away from the body, quadratic blend near it."""
    x_c, y_c = cfg.cylinder_center
    R = cfg.R_cylinder
    far = dist > R + 4 * cfg.dx
    with np.errstate(divide="ignore", invalid="ignore"):
        theta = np.arctan2(Y - y_c, X - x_c)
        factor = (R / dist) ** 2
        u_far = cfg.V_inf * (1 - factor * np.cos(2 * theta)) * (1 - ibm_mask)
        v_far = -cfg.V_inf * factor * np.sin(2 * theta) * (1 - ibm_mask)
    blend = np.minimum(1.0, ((dist - R) / (4 * cfg.dx)) ** 2)
    u_near = cfg.V_inf * blend * (1 - ibm_mask)
    u = np.where(far, u_far, u_near).astype(dtype)
    v = np.where(far, v_far, 0.0).astype(dtype)
    return np.ascontiguousarray(u), np.ascontiguousarray(v)


def setup_on_device(cfg, device, cols=None):
    """The same three setup steps (v5.py:269-314) evaluated with torch on the GPU, for grids where the host arrays
    would not fit comfortably (16384^2 per slab).  One-time setup, not part of the time step; transcendental
    functions may differ from NumPy's in the last ulp, so parity tests against the oracle use the host path.
    Returns x, y (host float64), cylinder_mask (uint8), ibm_mask (float64), u0, v0 (field dtype) as device tensors."""
    import torch
    x = np.linspace(cfg.x_min, cfg.x_max, cfg.nx)
    if cols is not None:
        x = x[cols[0]:cols[0] + cols[1]]
    y = np.linspace(cfg.y_min, cfg.y_max, cfg.ny)
    X = torch.from_numpy(np.ascontiguousarray(x)).to(device)[None, :]
    Y = torch.from_numpy(y).to(device)[:, None]
    x_c, y_c = cfg.cylinder_center
    R = cfg.R_cylinder
    dist = torch.sqrt((X - x_c) ** 2 + (Y - y_c) ** 2)
    mask = (dist <= R).to(torch.uint8)
    sigma = 2 * cfg.dx
    ibm = torch.exp(-((dist - R) / sigma) ** 2)
    zero = torch.zeros((), dtype=torch.float64, device=device)
    one = torch.ones((), dtype=torch.float64, device=device)
    ibm = torch.where(dist < R, one, torch.where(dist < R + 5 * cfg.dx, ibm, zero))
    far = dist > R + 4 * cfg.dx
    theta2 = 2 * torch.atan2(Y - y_c, X - x_c)
    factor = (R / dist) ** 2
    om = 1 - ibm
    tdt = torch.float32 if cfg.field_dtype == np.float32 else torch.float64
    u = torch.where(far, cfg.V_inf * (1 - factor * torch.cos(theta2)) * om,
                    cfg.V_inf * torch.clamp(((dist - R) / (4 * cfg.dx)) ** 2, max=1.0) * om).to(tdt).contiguous()
    v = torch.where(far, -cfg.V_inf * factor * torch.sin(theta2) * om, zero).to(tdt).contiguous()
    del dist, far, theta2, factor, om
    return x, y, mask.contiguous(), ibm.contiguous(), u, v
