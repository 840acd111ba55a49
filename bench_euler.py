#!/usr/bin/env python
"""bench_euler.py — BASELINE.json configs[4]: the explicit compressible-Euler finite-volume update of
python/shockwave_analysis/v1_shock.py (HLLC, Mach 2 wedge) scaled to a large structured grid, 1 GPU or N GPUs
(x-slabs, weak scaling: --nx rows per GPU).

  python bench_euler.py [--gpus N] [--nx 16384] [--ny 8192] [--steps K] [--warmup W] [--impl b200|reference]

One JSON line (rank 0) in bench.py's format.  A "step" is one ShockwaveSolver.time_step().  The kernel computes in
float64 on float32 storage exactly like the reference, so it is bound by the FP64 pipe: the HBM roofline fraction is
reported (33 B/cell algorithmic) next to the achieved FP64 instruction rate.
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
from bench import ClockSampler, measured_peak  # noqa: E402

METRIC = "cell-updates/s per Euler finite-volume time step (nx*ny*steps/t)"
UNIT = "Mcell-updates/s"
ALGO_BYTES = 16 + 16 + 1       # R U (float4), W U', R wedge mask


def workload(nx, ny, n):
    return (f"compressible Euler HLLC step, Mach 2 over a 10 deg wedge, {nx}x{ny} cells per GPU x {n} GPU(s), f32 storage / "
            "f64 arithmetic (BASELINE.json configs[4], v1_shock.py scaled)")


def cpu_rate(budget_s, nx=1024, ny=512):
    """CPU arm: the C/OpenMP oracle port of v1_shock.py's time step on a bounded sample grid."""
    import numpy as np
    from cfd_simulations_b200 import shockwave_setup as ss
    from oracle import orc_euler as oe
    cores = len(os.sched_getaffinity(0))
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cfg = ss.ShockwaveConfig(nx=nx, ny=ny)
    x, y = ss.grid(cfg)
    wm = ss.wedge_mask(cfg, x, y)
    c = oe.make_cfg(cfg, ss.n_reflect(cfg, x, y))
    U = ss.initial_state(cfg)
    oe.step(c, U, wm, 2)
    t0 = time.perf_counter(); oe.step(c, U, wm, 1); one = time.perf_counter() - t0
    n = max(2, min(200, int(budget_s / max(one, 1e-4))))
    t0 = time.perf_counter(); oe.step(c, U, wm, n); t = time.perf_counter() - t0
    assert np.isfinite(U).all()
    return nx * ny * n / t / 1e6, cores, f"{n} time steps of a {nx}x{ny} grid of the same case (oracle port, OpenMP {cores} threads)", t / n * 1e3


def reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import numpy as np
    from cfd_simulations_b200 import shockwave_setup as ss
    from oracle import orc_euler as oe
    cores = len(os.sched_getaffinity(0))
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    nx, ny = 2048, 1024
    cfg = ss.ShockwaveConfig(nx=nx, ny=ny)
    x, y = ss.grid(cfg)
    wm = ss.wedge_mask(cfg, x, y)
    c = oe.make_cfg(cfg, ss.n_reflect(cfg, x, y))
    U = ss.initial_state(cfg)
    oe.step(c, U, wm, args.warmup)
    t0 = time.perf_counter(); oe.step(c, U, wm, args.steps); t = time.perf_counter() - t0
    v = nx * ny * args.steps / t / 1e6
    sample = f"each step = one time step on a {nx}x{ny} sub-grid of the workload (oracle port of v1_shock.py, OpenMP {cores} threads)"
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "f32 storage / f64 arithmetic", "data": "synthetic",
                      "config": {"workload": workload(args.nx, args.ny, args.gpus)},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


FP64_FLOP_PER_CELL = 300       # from the ncu instruction mix of k_euler_step (see the roofline note)
FP64_PEAK_TFLOPS = 36.8        # measured: profiles/r02a_ubench_chain.txt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nx", type=int, default=16384, help="rows (x) per GPU")
    ap.add_argument("--ny", type=int, default=8192)
    ap.add_argument("--cpu-budget", type=float, default=10.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--exact-division", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist
    from cfd_simulations_b200 import shockwave as sw
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nx_g = args.nx * world
    # weak scaling: rows per GPU fixed; the wedge geometry is kept in index-proportional coordinates
    cfg = sw.ShockwaveConfig(nx=nx_g, ny=args.ny)
    s = sw.ShockwaveSolver(cfg, device=local, nranks=world, rank=rank, exact_division=args.exact_division)
    cells = nx_g * args.ny

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    s.advance(args.warmup)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    s.advance(args.steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device=s.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = cells * args.steps / (ms * 1e-3) / 1e6
    healthy = s.check_health()
    dts = s.dts[-args.steps:]

    # dominant kernel alone: the fused step kernel timed with CUDA events around single launches (no graph)
    roof = e2e = cpu = None
    if world == 1:
        peak, peak_src = measured_peak()
        s2 = sw.ShockwaveSolver(cfg, device=local, use_graph=False, exact_division=args.exact_division)
        s2.advance(3)
        torch.cuda.synchronize()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nk = min(args.steps, 20)
        k0.record(); s2.advance(nk); k1.record(); torch.cuda.synchronize()
        kms = k0.elapsed_time(k1) / nk
        ach = ALGO_BYTES * cells / (kms * 1e-3) / 1e9
        traffic = None
        tf = ROOT / "profiles" / "traffic_euler.json"
        if tf.exists():
            try:
                traffic = json.loads(tf.read_text()).get(f"{args.nx}x{args.ny}")
            except Exception:
                traffic = None
        roof = {"bound": "hbm", "kernel": "k_euler_step (fused HLLC fluxes + update + BC + wedge + dt reduction) + k_euler_ctl_end",
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                "algo_bytes_per_cell": ALGO_BYTES, "ms_per_launch": kms,
                "note": "arithmetic is float64 as in the reference: the kernel is bound by instruction issue, not HBM (DESIGN.md 9)",
                # second roof: fp64.  ~210 fp64 math instructions per cell (ncu, profiles/r01d_euler_full.md: ~1000 thread
                # instructions per cell, 21 % of them DFMA/DADD/DMUL; 45 % of those are FMAs) ~ 300 flop per cell; the peak is
                # the dependent-free DFMA rate measured on this pool's B200 with tools/ubench_chain.cu
                "fp64": {"flop_per_cell": FP64_FLOP_PER_CELL, "achieved": FP64_FLOP_PER_CELL * cells / (kms * 1e-3) / 1e12,
                         "peak": FP64_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": FP64_FLOP_PER_CELL * cells / (kms * 1e-3) / 1e12 / FP64_PEAK_TFLOPS,
                         "peak_source": "tools/ubench_chain.cu on B200: 36.8 TFLOP/s fp64 FMA (profiles/r02a_ubench_chain.txt)"}}
        s2.close()
        # end to end through the host-array entry point: pinned HOST U up, time_step(), U down, every step
        Uh = torch.empty((cfg.nx, cfg.ny, 4), dtype=torch.float32).pin_memory()
        Uh.copy_(s.device_tensor())
        ne = max(3, min(args.steps, 10))
        s.time_step_host(Uh, 1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(ne):
            s.time_step_host(Uh, 1)
        te = time.perf_counter() - t0
        nb = Uh.numel() * 4
        e2e = {"value": cells * ne / te / 1e6, "unit": UNIT, "h2d_bytes_per_step": nb, "d2h_bytes_per_step": nb + 8, "steps": ne,
               "ms_per_step": te / ne * 1e3, "note": "every step: pinned-host U -> HBM, time_step(), HBM -> host U + dt"}
        if not args.no_cpu:
            v, cores, sample, cms = cpu_rate(args.cpu_budget)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "ms_per_step": cms}
    if world > 1 and rank == 0:
        # per GPU: algorithmic bytes of one slab / the step time (max over ranks); e2e is an N = 1 measurement
        peak, peak_src = measured_peak()
        ach = ALGO_BYTES * (cells // world) / (ms / args.steps * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": "k_euler_step + k_euler_ctl_end + halo rows + all_reduce(MAX) per step", "achieved": ach,
                "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                "algo_bytes_per_cell": ALGO_BYTES, "note": "per GPU; float64 arithmetic: bound by fp64 issue, not HBM (DESIGN.md 9)"}
    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 storage / f64 arithmetic", "data": "synthetic",
            "config": {"workload": workload(args.nx, args.ny, world), "solver_type": "hllc", "cfl": cfg.cfl,
                       "exact_division": bool(args.exact_division),
                       "l2": f"state {cells // world * 16 / 1e6:.0f} MB per GPU, read once and written once per step"
                             + (" (> 126 MB L2)" if cells // world * 16 > 126e6 else " (fits L2: not an HBM measurement)"),
                       "healthy": healthy, "dt_last": dts[-1] if dts else None},
            "clocks": clocks, "e2e": e2e, "gpu_launches": s.launches_per_step() * args.steps, "roofline": roof, "cpu_baseline": cpu}))
    s.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
