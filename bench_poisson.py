#!/usr/bin/env python
"""bench_poisson.py — BASELINE.json configs[2]: pressure-Poisson-only benchmark on one B200.

  python bench_poisson.py [--n 8192] [--sweeps 200] [--dtype f64|f32] [--no-cpu]

Jacobi (v5.py:336-346) and red-black Gauss-Seidel / SOR (v5.py:202-226; omega = 1.5 = the v2.py:284-306 relaxation) on an
N x N grid, phi0 = 0, Dirichlet-0 border, disc mask of radius N/40 at (N/2, N/5), rhs = 0.1 * standard_normal (seed 0),
dt = float32(5e-5), dx = dy = 1/(N-1)  (SURVEY.md 8(d), C3).  Modes: fixed sweep count with the one-sweep-per-pass kernels
(depth 1) and with the TMA temporally blocked kernels (depth T), and tolerance-terminated red-black.
One JSON line per mode: MLUPS = interior cells * sweeps / t; algo_GBs = 25 B/cell/sweep * that rate (what an un-blocked
sweep must move: phi in, rhs, mask, phi out); pass_GBs = the bytes the blocked kernel really moves per pass (24 B/cell);
frac = algo_GBs / measured HBM peak (can exceed 1 with temporal blocking: that is the point).  The CPU oracle port is
timed beside it on a bounded sample (a few sweeps of the same grid).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def run_slabs(args):
    """--gpus N > 1: the same red-black solve on x-slabs of n x n cells per GPU (global grid n x (n * N)), one process per
    GPU: phi halos refreshed over NCCL every 7 iterations (temporally blocked passes in between), the last change
    all-reduced.  One JSON line per mode, value = global interior cells * sweeps / max-over-ranks device time."""
    import subprocess
    if "RANK" not in os.environ:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29541", __file__] + sys.argv[1:]
        return subprocess.call(cmd)
    import numpy as np
    import torch
    import torch.distributed as dist
    from bench import measured_peak
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    N = args.n
    h = 1.0 / (N - 1)
    cfg = OptimizedTurbulentConfig(nx=N * world, ny=N, memory_efficient=(args.dtype == "f32"), x_min=0.0, x_max=(N * world - 1) * h,
                                   y_min=0.0, y_max=(N - 1) * h, cylinder_center=((N // 5) * h, (N // 2) * h),
                                   R_cylinder=(N // 40) * h)
    s = OptimizedTurbulentSolver(cfg, device=local, nranks=world, rank=rank, setup="device")
    t = s.device_tensor("phi")
    g = torch.Generator(device=s.device).manual_seed(1234 + rank)
    rhs = 0.1 * torch.randn(t.shape, generator=g, dtype=t.dtype, device=s.device)
    s._phi_exchanger().exchange([rhs])                     # halo columns of the right-hand side = the neighbours' owned values
    es = 8 if args.dtype == "f64" else 4
    algo_b = 2 * es + es + 1
    peak, peak_src = measured_peak()
    cells = (N - 2) * (N * world - 2)
    for name, omega in (("rbgs_omega1", 1.0), ("rbsor_omega1.5", 1.5)):
        s.solve_pressure_slabs(rhs, max_iter=15, tol=0.0, omega=omega)          # warm-up (tensor maps, buffers)
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, it, last = s.solve_pressure_slabs(rhs, max_iter=args.sweeps, tol=0.0, omega=omega, return_info=True)
        e1.record()
        torch.cuda.synchronize(); dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=s.device)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = float(ms.item())
        if rank == 0:
            mlups = cells * it / (ms * 1e-3) / 1e6
            print(json.dumps({"bench": "poisson", "mode": name, "n_gpus": world, "grid": [N, N * world], "per_gpu": [N, N],
                              "dtype": args.dtype, "sweeps": it, "ms": ms, "MLUPS": mlups, "scaling": "weak",
                              "algo_GBs_per_gpu": mlups * 1e6 * algo_b / 1e9 / world,
                              "frac_of_measured_hbm_peak_per_gpu": mlups * 1e6 * algo_b / 1e9 / world / peak,
                              "hbm_peak_GBs": peak, "peak_source": peak_src, "last_change": last,
                              "kernel": "TMA temporally blocked, 7 iterations per NCCL halo refresh",
                              "halo_bytes_per_refresh_per_gpu": s._phi_exchanger().bytes_per_exchange}), flush=True)
    s.close()
    dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--n", type=int, default=8192)
    ap.add_argument("--sweeps", type=int, default=240)
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--depths", default="1,0,2,3,4,6")
    args = ap.parse_args()
    if args.gpus > 1 or int(os.environ.get("WORLD_SIZE", "1")) > 1:
        sys.exit(run_slabs(args))
    import numpy as np
    import torch
    from cfd_simulations_b200 import kernels as K
    from bench import measured_peak
    N = args.n
    npdt = np.float64 if args.dtype == "f64" else np.float32
    tdt = torch.float64 if args.dtype == "f64" else torch.float32
    es = 8 if args.dtype == "f64" else 4
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(0)
    rhs = (0.1 * torch.randn((N, N), generator=g, dtype=torch.float64)).to(tdt).to(dev)
    yy, xx = torch.meshgrid(torch.arange(N, device=dev), torch.arange(N, device=dev), indexing="ij")
    mask = (((xx - N // 5) ** 2 + (yy - N // 2) ** 2) <= (N // 40) ** 2).to(torch.uint8).contiguous()
    del yy, xx
    dx = dy = 1.0 / (N - 1)
    cdt = float(np.float32(5e-5))
    peak, peak_src = measured_peak()
    cells = (N - 2) * (N - 2)
    algo_b = 2 * es + es + 1        # phi in + rhs + mask + phi out per cell per sweep (25 B fp64, 13 B fp32)

    def timed(fn, reps=1):
        fn()                         # warm-up (allocations, tensor maps, smem attribute)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            out = fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps, out

    def emit(name, ms, sweeps, extra):
        mlups = cells * sweeps / (ms * 1e-3) / 1e6
        line = {"bench": "poisson", "mode": name, "grid": [N, N], "dtype": args.dtype, "sweeps": sweeps, "ms": ms,
                "MLUPS": mlups, "algo_GBs": mlups * 1e6 * algo_b / 1e9, "frac_of_measured_hbm_peak": mlups * 1e6 * algo_b / 1e9 / peak,
                "hbm_peak_GBs": peak, "peak_source": peak_src}
        line.update(extra)
        print(json.dumps(line), flush=True)

    phi = torch.zeros((N, N), dtype=tdt, device=dev)
    ref = {}
    for depth in [int(d) for d in args.depths.split(",")]:
        for name, fn in (
            ("jacobi", lambda d=depth: K.solve_pressure_jacobi(phi.zero_(), rhs, dx, cdt, mask, args.sweeps, temporal_depth=d)),
            ("rbgs_omega1", lambda d=depth: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, args.sweeps, 0.0, omega=1.0, temporal_depth=d)),
            ("rbsor_omega1.5", lambda d=depth: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, args.sweeps, 0.0, omega=1.5, temporal_depth=d)),
        ):
            ms, out = timed(fn)
            chk = float(out.double().abs().sum().item())
            if depth == 1:
                ref[name] = out.clone()
                dev_rel = 0.0
            else:
                dev_rel = float(((out - ref[name]).double().norm() / ref[name].double().norm()).item()) if name in ref else None
            pass_b = 3 * es
            eff_depth = depth if depth > 0 else (4 if name == "jacobi" else 3)
            emit(name, ms, args.sweeps, {"temporal_depth": "auto" if depth == 0 else depth,
                                         "kernel": "one sweep per pass" if depth == 1 else "TMA temporally blocked",
                                         "pass_GBs": None if depth == 1 else cells * pass_b * (args.sweeps / eff_depth) / (ms * 1e-3) / 1e9,
                                         "rel_l2_vs_depth1": dev_rel, "checksum": chk})
    # tolerance-terminated red-black (exact per-iteration max|dphi|, same stopping iteration as the sequential algorithm)
    tol = 1e-3 * float(rhs.abs().max().item()) * dx * dx / cdt
    t0 = time.perf_counter()
    out, iters, last = K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, 20000, tol, return_info=True)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    emit("rbgs_tolerance", ms, iters, {"tol": tol, "iterations": iters, "last_change": last, "timing": "wall clock incl. host checks"})
    if not args.no_cpu:
        from oracle import orc
        cores = len(os.sched_getaffinity(0))
        os.environ.setdefault("OMP_NUM_THREADS", str(cores))
        h_rhs, h_mask = rhs.cpu().numpy(), mask.cpu().numpy().astype(bool)
        for name, call in (("jacobi", lambda n: orc.solve_pressure_jacobi(np.zeros((N, N), npdt), h_rhs, dx, cdt, h_mask, n)),
                           ("rbgs_omega1", lambda n: orc.solve_pressure_gauss_seidel(np.zeros((N, N), npdt), h_rhs, dx, dy, cdt, h_mask, n, 0.0, 1.0))):
            call(1)
            t0 = time.perf_counter(); call(3); t = time.perf_counter() - t0
            print(json.dumps({"bench": "poisson", "mode": name, "impl": "cpu oracle port", "cores": cores, "grid": [N, N],
                              "sample": "3 sweeps of the same grid", "MLUPS": cells * 3 / t / 1e6}), flush=True)


if __name__ == "__main__":
    main()
