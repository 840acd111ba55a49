"""Workload for compute-sanitizer (tools/sanitize.sh): every hand-written kernel of the hot path once, on small grids.
smoke() (projection step 300x90 + Euler 96x50), a 1024x256 fp64 and fp32 step (TMA sweeps, both fused corrections),
the general (non-fast) path, Poisson solves through the streaming TMA pipeline (Jacobi, red-black fixed and tolerance mode) and the tile
kernel, and an emulated 2-slab step (halo pack / unpack, split finalize)."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import __graft_entry__ as g                                      # noqa: E402
from cfd_simulations_b200 import kernels as K                    # noqa: E402
from cfd_simulations_b200.config import OptimizedTurbulentConfig   # noqa: E402
from cfd_simulations_b200.multigpu import LocalSlabs             # noqa: E402
from cfd_simulations_b200.solver import OptimizedTurbulentSolver   # noqa: E402

g.smoke()
for f32 in (False, True):
    s = OptimizedTurbulentSolver(OptimizedTurbulentConfig(nx=1024, ny=256, memory_efficient=f32))
    s.advance(3)
    assert np.isfinite(s.u).all()
    s.close()
s = OptimizedTurbulentSolver(OptimizedTurbulentConfig(nx=150, ny=46, memory_efficient=False, use_supg=False))   # general kernels
s.advance(3)
s.close()
rng = np.random.default_rng(0)
N = 384
div = 0.1 * rng.standard_normal((N, N))
yy, xx = np.mgrid[0:N, 0:N]
mask = (xx - N // 5) ** 2 + (yy - N // 2) ** 2 <= (N // 40) ** 2
h, cdt = 1.0 / (N - 1), float(np.float32(5e-5))
z = lambda: np.zeros((N, N))
K.solve_pressure_jacobi(z(), div, h, cdt, mask, 13, temporal_depth=0)
K.solve_pressure_gauss_seidel_fast(z(), div, h, h, cdt, mask, 8, 0.0, omega=1.5, temporal_depth=0)
tol = 2e-2 * float(np.abs(div).max()) * h * h / cdt
_, it, _ = K.solve_pressure_gauss_seidel_fast(z(), div, h, h, cdt, mask, 500, tol, return_info=True)
print("tolerance mode iterations", it)
os.environ["B200CFD_POISSON_STREAM"] = "0"                       # (read once per process: the tile kernel needs its own run)
ls = LocalSlabs(OptimizedTurbulentConfig(nx=1600, ny=64, memory_efficient=False), 2)
ls.advance(3)
ls.close()
print("sanitize case done")
