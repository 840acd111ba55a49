#!/usr/bin/env python
"""tools/alpha_sweep.py — streaming Poisson kernel at 8192^2 fp64 (BASELINE.json configs[2] inputs, as bench_poisson.py builds
them) for several values of B200CFD_STREAM_ALPHA, the cost the tile plan (csrc/poisson_stream_plan.h) assigns to a chunk that
runs the class selects (percent of a plain chunk; 0 = equal row ranges).  The library reads the variable on every call, so one
process does the whole sweep.  Prints GLUPS per mode and alpha, and the checksum of the result (must not depend on alpha:
the plan moves tile boundaries, never values)."""
import argparse
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from cfd_simulations_b200 import kernels as K

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=8192)
ap.add_argument("--sweeps", type=int, default=240)
ap.add_argument("--alphas", default="0,20,35,50")
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
N = args.n
dev = torch.device("cuda", 0)
g = torch.Generator(device="cpu").manual_seed(0)
rhs = (0.1 * torch.randn((N, N), generator=g, dtype=torch.float64)).to(dev)
yy, xx = torch.meshgrid(torch.arange(N, device=dev), torch.arange(N, device=dev), indexing="ij")
mask = (((xx - N // 5) ** 2 + (yy - N // 2) ** 2) <= (N // 40) ** 2).to(torch.uint8).contiguous()
del yy, xx
dx = dy = 1.0 / (N - 1)
cdt = float(np.float32(5e-5))
cells = (N - 2) * (N - 2)
phi = torch.zeros((N, N), dtype=torch.float64, device=dev)
modes = (
    ("jacobi", lambda: K.solve_pressure_jacobi(phi.zero_(), rhs, dx, cdt, mask, args.sweeps, temporal_depth=0)),
    ("rbgs_omega1", lambda: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, args.sweeps, 0.0, omega=1.0, temporal_depth=0)),
    ("rbsor_omega1.5", lambda: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, args.sweeps, 0.0, omega=1.5, temporal_depth=0)),
)
for name, fn in modes:
    sums = set()
    for a in args.alphas.split(","):
        os.environ["B200CFD_STREAM_ALPHA"] = a
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.reps
        chk = float(out.abs().sum().item())
        sums.add(chk)
        print(f"{name:16s} alpha {a:>3s}: {ms:8.3f} ms for {args.sweeps} sweeps  {cells * args.sweeps / ms / 1e6:8.1f} GLUPS  checksum {chk!r}", flush=True)
    print(f"{name}: checksums {'identical' if len(sums) == 1 else 'DIFFER: ' + repr(sums)}", flush=True)
