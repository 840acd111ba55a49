#!/usr/bin/env python
"""bench.py — the reference's headline metric on B200: cell-updates/s of the projection time step
(python/flow_over_cylinder (Fischer)/v5.py:375-441) on BASELINE.json configs[1] (4096x1024, fp64), with the
roofline of the dominant kernel and the CPU baseline beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One JSON line on stdout (rank 0).  A "step" is one OptimizedTurbulentSolver.time_step().
  value    cell-updates/s with the state resident in HBM (CUDA-graph replay of the step), CUDA-event timed
  e2e      the same through the reference-facing host API with HOST arrays: every step uploads u, v from pinned
           memory, runs time_step(), downloads u, v and the step's scalars
  roofline dominant kernel: algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle port (oracle/, OpenMP) timed on this box's host cores on a bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

NX, NY = 4096, 1024            # BASELINE.json configs[1]
METRIC = "cell-updates/s per projection time step (MLUPS_step = nx*ny*steps/t)"
UNIT = "Mcell-updates/s"
FALLBACK_HBM_GBS = 6650.0      # B200_PROFILING.md fallback


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# algorithmic HBM bytes per cell of each phase, fp64 (DESIGN.md "Kernels")
ALGO_BYTES = {
    "predictor": 16 + 16 + 8,            # R u,v  W u*,v* (+ W tau_supg, a reference attribute); ibm only in flagged tiles
    "poisson": 16 + 8 + 8,               # as executed: R u*,v*  W phi, div_u_star; mask only in flagged tiles
    "correct": 24 + 16,                  # R u*,v*,phi  W u,v
    "poisson_correct": 16 + 8 + 8 + 16,  # K2 + K3 fused (default): R u*,v*  W phi, div_u_star, u, v
    "clean_sweep_1": 16 + 8,             # R u,v  W psi (psi_old = 0)
    "clean_apply_1": 24 + 16,            # R psi,u,v  W u,v
    "clean_sweep_apply_1": 16 + 24,      # sweep + correction fused (default): R u,v  W psi,u,v (psi_old = 0)
    "clean_sweep_apply_2": 24 + 24,      # (clean_iters > 2 only)
    "clean_sweep_2": 16 + 8 + 8,         # R u,v,psi_old  W psi
    "clean_apply_2": 24 + 16,            # (general path only; the fast path applies it inside finalize)
    "finalize": 24 + 16,                 # R psi,u,v  W u,v (last correction fused); ibm, mask only in flagged tiles
    "ctl": 0,
}

# the kernels of the default path (as executed Poisson, lexicographic clean-divergence, fast kernels), one launch each
DEFAULT_PHASES = ("predictor", "poisson_correct", "clean_sweep_1", "clean_apply_1", "clean_sweep_2", "finalize")


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.p = None
        self.idx = gpu_index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_rate(nx, ny, budget_s, f32=False):
    """The CPU arm: oracle port (C + OpenMP restatement of v5.py's time step) on the host cores.
    Returns (Mcell-updates/s, cores, sample description, ms/step)."""
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from oracle.solver import OracleTurbulentSolver
    cores = len(os.sched_getaffinity(0))
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    cfg = OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=f32)
    s = OracleTurbulentSolver(cfg)
    s.time_step()                          # warm-up (page faults, OpenMP pool)
    t0 = time.perf_counter(); s.time_step(); one = time.perf_counter() - t0
    n = max(2, min(50, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        s.time_step()
    t = time.perf_counter() - t0
    rate = nx * ny * n / t / 1e6
    return rate, cores, f"{n} time steps of the {nx}x{ny} {'fp32' if f32 else 'fp64'} workload (oracle port, OpenMP {cores} threads)", t / n * 1e3


def run_reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores, same config/metric/unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from oracle.solver import OracleTurbulentSolver
    cores = len(os.sched_getaffinity(0))
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    # bounded sample: each "step" is one time step on an (nx_s x ny_s) sub-grid of the workload, sized so that
    # steps+warmup finish in ~150 s
    cfg = OptimizedTurbulentConfig(nx=NX, ny=NY, memory_efficient=False)
    s = OracleTurbulentSolver(cfg)
    s.time_step()
    t0 = time.perf_counter(); s.time_step(); one = time.perf_counter() - t0
    total = args.steps + args.warmup
    nx_s, ny_s = NX, NY
    while one * total * (nx_s * ny_s) / (NX * NY) > 150.0 and nx_s > 512:
        nx_s //= 2; ny_s //= 2
    if (nx_s, ny_s) != (NX, NY):
        cfg = OptimizedTurbulentConfig(nx=nx_s, ny=ny_s, memory_efficient=False)
        s = OracleTurbulentSolver(cfg)
    for _ in range(args.warmup):
        s.time_step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s.time_step()
    t = time.perf_counter() - t0
    rate = nx_s * ny_s * args.steps / t / 1e6
    sample = (f"{args.steps} time steps on a {nx_s}x{ny_s} fp64 grid (same solver config; the full workload is "
              f"{NX}x{NY}); oracle port = C/OpenMP restatement of v5.py, {cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": (f"cylinder case {NX}x{NY} fp64 (BASELINE.json configs[1])" if args.gpus == 1 else
                                    f"cylinder/channel case weak-scaled along x: {args.slab_nx}x{args.slab_ny} cells per GPU slab, fp64 "
                                    f"(BASELINE.json configs[3]) - timed on the bounded sample below: the rate per cell of this "
                                    f"stencil code does not depend on the grid size once it exceeds the caches"),
                       "sample_grid": [nx_s, ny_s]},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def slab_config(nx_global, ny, h=1.0 / 1024):
    """BASELINE.json configs[3]: the cylinder/channel case weak-scaled along x, ny x (nx per GPU * N) cells with the
    spacing fixed at h (the reference's warm-up dt = 2e-5 needs h >= 4.6e-4), cylinder at x = 4 in the
    first slab, mid-height."""
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    y_max = (ny - 1) * h
    return OptimizedTurbulentConfig(nx=nx_global, ny=ny, memory_efficient=False, x_min=0.0, x_max=(nx_global - 1) * h,
                                    y_min=0.0, y_max=y_max, cylinder_center=(4.0, y_max / 2))


def time_slab_steps(s, steps, warmup, world):
    """warmup + timed steps of a (possibly decomposed) solver: barrier + synchronize on both sides, CUDA events,
    max over ranks.  Returns milliseconds for `steps` steps."""
    import torch
    import torch.distributed as dist
    s.advance(warmup)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s.advance(steps)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=s.device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def run_slabs(args):
    """N > 1: one process per GPU (torchrun), x-slabs of slab_nx x slab_ny cells per GPU (weak scaling)."""
    import torch
    import torch.distributed as dist
    if "RANK" not in os.environ:
        # launched as plain `python bench.py --gpus N`: re-launch under torchrun on this node
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", __file__] + sys.argv[1:]
        return subprocess.call(cmd)
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    nxs, ny = args.slab_nx, args.slab_ny
    cfg = slab_config(nxs * world, ny)
    s = OptimizedTurbulentSolver(cfg, device=local, nranks=world, rank=rank, setup="device")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = time_slab_steps(s, args.steps, args.warmup, world)
    clocks = sampler.stop() if rank == 0 else None
    cells = nxs * world * ny
    value = cells * args.steps / (ms * 1e-3) / 1e6
    # end to end: every step uploads this slab's u, v from pinned host memory, steps, downloads them and the scalars
    n_e2e = max(1, min(args.steps, 3))
    uh = torch.empty_like(s.device_tensor("u"), device="cpu").pin_memory()
    vh = torch.empty_like(uh).pin_memory()
    uh.copy_(s.device_tensor("u")); vh.copy_(s.device_tensor("v"))
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        s.device_tensor("u").copy_(uh, non_blocking=True); s.device_tensor("v").copy_(vh, non_blocking=True)
        s.advance(1)
        uh.copy_(s.device_tensor("u"), non_blocking=True); vh.copy_(s.device_tensor("v"), non_blocking=True)
        s._sync_scalars()
        torch.cuda.synchronize()
    dist.barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=s.device)
    dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    t_e2e = float(t_e2e.item())
    slab_bytes = uh.numel() * 8 * 2
    if rank == 0:
        peak, peak_src = measured_peak()
        step_bytes = sum(ALGO_BYTES[k] for k in DEFAULT_PHASES) * cells
        ach = step_bytes / (ms / args.steps * 1e-3) / 1e9 / world
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"cylinder/channel case weak-scaled along x: {nxs}x{ny} cells per GPU slab, fp64, "
                                       f"h = 1/1024 (BASELINE.json configs[3]); global grid {nxs * world}x{ny}",
                           "poisson_mode": "as_executed", "cleandiv_mode": "lexicographic",
                           "decomposition": {"halo_left": s.plan.halo_left if rank else s.plan.send_right,
                                             "halo_right": s.plan.halo_right, "exchanges_per_step": 1,
                                             "halo_bytes_per_step_per_gpu": s._halo.bytes_per_exchange,
                                             "collectives_per_step": "1 all_gather (128 B/rank) + 1 batched send/recv"},
                           "l2": f"working set {13 * nxs * ny * 8 / 1e9:.1f} GB per GPU >> 126 MB L2"},
                "clocks": clocks,
                "e2e": {"value": cells * n_e2e / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": slab_bytes * world,
                        "d2h_bytes_per_step": slab_bytes * world + 96 * world, "steps": n_e2e,
                        "ms_per_step": t_e2e / n_e2e * 1e3,
                        "note": "every step, every rank: pinned-host slab u,v -> HBM, time step, HBM -> host u,v + scalars"},
                "gpu_launches": (s.launches_per_step() + 8) * args.steps * world,
                "roofline": {"bound": "hbm", "kernel": "whole step (sum of phases)", "achieved": ach, "peak": peak,
                             "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                             "note": "per GPU: algorithmic bytes of all phases / step time"},
                "cpu_baseline": None}
        print(json.dumps(line), flush=True)
    s.close()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nx", type=int, default=NX)
    ap.add_argument("--ny", type=int, default=NY)
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for cpu_baseline")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--slab-nx", type=int, default=16384, help="columns per GPU slab (N > 1, weak scaling)")
    ap.add_argument("--slab-ny", type=int, default=16384)
    ap.add_argument("--no-scaling-base", action="store_true", help="skip the 1-GPU run of one weak-scaling slab")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 or args.gpus > 1:
        return run_slabs(args)

    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    torch.cuda.set_device(local)
    nx, ny = args.nx, args.ny
    cfg = OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=False)
    s = OptimizedTurbulentSolver(cfg, device=local, setup="device" if nx * ny > (1 << 25) else "host")
    cells = nx * ny

    # ---- device-resident throughput: CUDA-graph replay, CUDA events, working set (10 fp64 fields = 336 MB) > L2
    s.advance(args.warmup)
    torch.cuda.synchronize()
    sampler = ClockSampler(local); sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    s.advance(args.steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    value = cells * args.steps / (ms * 1e-3) / 1e6
    launches = s.launches_per_step() * args.steps

    # ---- per-phase CUDA-event profile of the same step (stream launches, events between kernels)
    prof = s.profile_steps(min(args.steps, 50))
    peak, peak_src = measured_peak()
    dom = max((k for k in prof if k != "ctl"), key=lambda k: prof[k])
    phases = {}
    for k, t_ms in prof.items():
        b = ALGO_BYTES.get(k, 0) * cells
        phases[k] = {"ms": round(t_ms, 5), "algo_GBs": round(b / (t_ms * 1e-3) / 1e9, 1) if t_ms > 0 else None}
    ach = ALGO_BYTES[dom] * cells / (prof[dom] * 1e-3) / 1e9
    step_bytes = sum(ALGO_BYTES.get(k, 0) for k in prof) * cells
    traffic = None
    tf = ROOT / "profiles" / "traffic.json"
    if tf.exists():
        try:
            traffic = json.loads(tf.read_text()).get(dom)
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src,
                "whole_step": {"algo_bytes_per_cell": step_bytes // cells, "achieved": step_bytes / (ms / args.steps * 1e-3) / 1e9,
                               "frac": step_bytes / (ms / args.steps * 1e-3) / 1e9 / peak},
                "phases": phases}

    # ---- end to end through the host API: H2D u,v (pinned) -> time_step -> D2H u,v + scalars, every step
    tdt = torch.float64
    uh = torch.empty((ny, nx), dtype=tdt).pin_memory(); vh = torch.empty((ny, nx), dtype=tdt).pin_memory()
    uh.copy_(s.device_tensor("u")); vh.copy_(s.device_tensor("v"))
    un, vn = uh.numpy(), vh.numpy()
    for _ in range(3):
        s.time_step_host(un, vn, 1)
    torch.cuda.synchronize()
    n_e2e = min(args.steps, 50)
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        s.time_step_host(un, vn, 1)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    e2e = {"value": cells * n_e2e / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": 2 * cells * 8,
           "d2h_bytes_per_step": 2 * cells * 8 + 96, "steps": n_e2e, "ms_per_step": t_e2e / n_e2e * 1e3,
           "note": "every step: pinned-host u,v -> HBM, time_step(), HBM -> host u,v + 96 B of scalars"}

    # ---- CPU baseline on this box's host cores (bounded sample)
    cpu = None
    if not args.no_cpu:
        r, cores, sample, cpu_ms = cpu_reference_rate(nx, ny, args.cpu_budget)
        cpu = {"value": r, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "ms_per_step": cpu_ms}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"flow-over-cylinder projection step, {nx}x{ny} grid, fp64, Re 600 (BASELINE.json configs[1])",
                       "poisson_mode": "as_executed (1 red-black iteration: what the reference runs under Numba)",
                       "cleandiv_mode": "lexicographic (reference at NUMBA_NUM_THREADS=1)",
                       "l2": "working set 336 MB > 126 MB L2; no flush between steps (each step streams all fields)",
                       "clean_plan": s.clean_plan()},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu}
    s.close()
    if not args.no_scaling_base:
        # the 1-GPU point of the weak-scaling series (same slab as `--gpus N` uses per GPU), so that scaling efficiency
        # can be formed against the SAME per-GPU workload: value(N) / (N * scaling_base.value)
        del s
        torch.cuda.empty_cache()
        cfgs = slab_config(args.slab_nx, args.slab_ny)
        sb = OptimizedTurbulentSolver(cfgs, device=local, setup="device")
        nb = max(5, min(args.steps, 30))
        msb = time_slab_steps(sb, nb, 3, 1)
        line["scaling_base"] = {"workload": f"{args.slab_nx}x{args.slab_ny} fp64 slab on 1 GPU (BASELINE.json configs[3] per-GPU share)",
                                "value": args.slab_nx * args.slab_ny * nb / (msb * 1e-3) / 1e6, "unit": UNIT,
                                "ms_per_step": msb / nb, "steps": nb}
        sb.close()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
