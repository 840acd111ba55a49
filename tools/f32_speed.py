"""Step time of the float32-storage mode against fp64 on the same grid (CUDA-graph replay, CUDA events).
python tools/f32_speed.py [nx ny steps]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cfd_simulations_b200.config import OptimizedTurbulentConfig   # noqa: E402
from cfd_simulations_b200.solver import OptimizedTurbulentSolver   # noqa: E402

nx, ny, steps = (int(a) for a in (sys.argv[1:4] + ["16384", "8192", "20"][len(sys.argv) - 1:]))
out = {}
for f32 in (False, True):
    cfg = OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=f32)
    s = OptimizedTurbulentSolver(cfg, device=0, setup="device" if nx * ny > (1 << 25) else "host")
    s.advance(5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); s.advance(steps); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    prof = s.profile_steps(min(steps, 20))
    out["f32" if f32 else "f64"] = ms
    print(("f32" if f32 else "f64"), f"{nx}x{ny} {ms:.4f} ms/step  {nx * ny / ms / 1e3:.0f} Mcell/s ", " ".join(f"{k} {v:.4f}" for k, v in prof.items()), flush=True)
    s.close()
    del s
    torch.cuda.empty_cache()
print(f"f64/f32 time ratio {out['f64'] / out['f32']:.3f}")
