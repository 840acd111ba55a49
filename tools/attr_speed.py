"""Step time with and without the two by-product attributes (tau_supg, div_u_star).  python tools/attr_speed.py nx ny steps"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cfd_simulations_b200.config import OptimizedTurbulentConfig   # noqa: E402
from cfd_simulations_b200.solver import OptimizedTurbulentSolver   # noqa: E402
nx, ny, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
for attrs in (True, False):
    s = OptimizedTurbulentSolver(OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=False), device=0,
                                 setup="device" if nx * ny > (1 << 25) else "host", materialize_attributes=attrs)
    s.advance(5); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s.advance(steps); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    prof = s.profile_steps(min(steps, 20))
    print("attributes" if attrs else "no attributes", f"{nx}x{ny} {ms:.4f} ms/step {nx * ny / ms / 1e3:.0f} Mcell/s", " ".join(f"{k} {v:.4f}" for k, v in prof.items()), flush=True)
    s.close(); del s; torch.cuda.empty_cache()
