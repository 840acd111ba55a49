"""Small workload for the ncu capture of the slab plumbing kernels: two emulated x-slabs of a 4096x1024 grid on one GPU
(k_halo_rows pack / unpack, split finalize) and one v1/v2 SOR solve (k_sor_v2_bc)."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cfd_simulations_b200 import kernels as K                    # noqa: E402
from cfd_simulations_b200.config import OptimizedTurbulentConfig   # noqa: E402
from cfd_simulations_b200.multigpu import LocalSlabs             # noqa: E402

ls = LocalSlabs(OptimizedTurbulentConfig(nx=4096, ny=1024, memory_efficient=False), 2)
ls.advance(3)
ls.close()
rng = np.random.default_rng(0)
div = 1e-3 * rng.standard_normal((256, 512))
K.solve_pressure_iterative(div, 0.01, 0.01, 5e-5, np.zeros((256, 512), bool), max_iter=6, tol=0.0)
print("slab case done")
