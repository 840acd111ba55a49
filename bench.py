#!/usr/bin/env python
"""bench.py — the reference's headline metric on B200: cell-updates/s of the projection time step
(python/flow_over_cylinder (Fischer)/v5.py:375-441) and the Poisson solve's HBM rate (v5.py:202-226, 336-346), with
the roofline of the dominant kernel and the CPU baseline beside them.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One JSON line on stdout (rank 0).  A "step" is one OptimizedTurbulentSolver.time_step().
  N = 1   BASELINE.json configs[1]: 4096x1024 fp64
  N > 1   BASELINE.json configs[3]: x-slabs of 16384^2 cells per GPU (weak scaling), one process per GPU
  value    cell-updates/s with the state resident in HBM (CUDA-graph replay of the step), CUDA-event timed
  e2e      the same through the reference-facing host API with HOST arrays: every step uploads u, v from pinned
           memory, runs time_step(), downloads u, v and the step's scalars
  roofline dominant kernel: algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json
  poisson  BASELINE.json configs[2] (8192^2 fp64 Jacobi / red-black SOR, fixed count and tolerance mode): GLUPS, GB/s, frac
  parity   (N > 1) the N-rank slab run against one GPU on a small grid, over real NCCL, before anything is timed
  weak_scaling (N > 1) the same slab stepped by rank 0 alone in the same run, and the efficiency against it
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
PARITY_GRID = (1600, 128)      # (nx, ny) of the N-rank-vs-one-GPU check run inside `--gpus N`
PARITY_STEPS = 15
PARITY_TOL = 1e-13


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


# algorithmic HBM bytes per cell of each phase, fp64 (DESIGN.md "Kernels"): the compulsory traffic of
# the kernel AS FUSED (a fused kernel is charged its own reads and writes, never the sum of the kernels it replaced)
ALGO_BYTES = {
    "predictor": 16 + 16 + 8,            # R u,v  W u*,v* (+ W tau_supg, a reference attribute); ibm only in flagged tiles
    "poisson": 16 + 8 + 8,               # as executed: R u*,v*  W phi, div_u_star; mask only in flagged tiles
    "correct": 24 + 16,                  # R u*,v*,phi  W u,v
    "poisson_correct": 16 + 8 + 8 + 16,  # K2 + K3 fused (default): R u*,v*  W phi, div_u_star, u, v
    "clean_sweep_1": 16 + 8,             # R u,v  W psi (psi_old = 0)
    "clean_apply_1": 24 + 16,            # R psi,u,v  W u,v   (B200CFD_FUSE_CORR=0 only)
    "clean_sweep_apply_1": 16 + 24,      # sweep + correction in one kernel (B200CFD_FUSE_APPLY only)
    "clean_sweep_apply_2": 24 + 24,
    "clean_sweep_2": 16 + 8 + 8,         # R u,v,psi_1  W psi_2 (the first correction is applied on the load side)
    "clean_apply_2": 24 + 16,            # (general path only; the fast path applies it inside finalize)
    "finalize": 24 + 16,                 # R psi,u,v  W u,v (last correction fused); ibm, mask only in flagged tiles
    "finalize2": 32 + 16,                # R psi_1,psi_2,u,v  W u,v (both corrections fused: no clean_apply pass)
    "ctl": 0,
}


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


# ---------------------------------------------------------------------------------------------------------------
# the workload description: ONE function, used verbatim by both arms, so that the driver's same_config check holds
def make_config(n_gpus, args):
    if n_gpus <= 1:
        return {"workload": f"flow-over-cylinder projection step, {args.nx}x{args.ny} grid, fp64, Re 600 (BASELINE.json configs[1])",
                "poisson_mode": "as_executed (1 red-black iteration: what the reference runs under Numba)",
                "cleandiv_mode": "lexicographic (reference at NUMBA_NUM_THREADS=1)",
                "l2": "working set > 126 MB L2; no flush between steps (each step streams all fields)"}
    return {"workload": f"cylinder/channel case weak-scaled along x: {args.slab_nx}x{args.slab_ny} cells per GPU slab, fp64, "
                        f"h = 1/1024 (BASELINE.json configs[3]); global grid {args.slab_nx * n_gpus}x{args.slab_ny}",
            "poisson_mode": "as_executed (1 red-black iteration: what the reference runs under Numba)",
            "cleandiv_mode": "lexicographic (reference at NUMBA_NUM_THREADS=1)",
            "l2": "working set >> 126 MB L2; no flush between steps (each step streams all fields)"}


def slab_config(nx_global, ny, h=1.0 / 1024):
    """BASELINE.json configs[3]: the cylinder/channel case weak-scaled along x, ny x (nx per GPU * N) cells with the
    spacing fixed at h (the reference's warm-up dt = 2e-5 needs h >= 4.6e-4), cylinder at x = 4 in the
    first slab, mid-height."""
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    y_max = (ny - 1) * h
    return OptimizedTurbulentConfig(nx=nx_global, ny=ny, memory_efficient=False, x_min=0.0, x_max=(nx_global - 1) * h,
                                    y_min=0.0, y_max=y_max, cylinder_center=(4.0, y_max / 2))


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port (C + OpenMP restatement of v5.py's time step) on the host cores
def omp_team():
    """Set the oracle's OpenMP team to all the cores this process may run on and return (requested, actual).
    torchrun exports OMP_NUM_THREADS=1 to its workers, so the environment value is overwritten, not defaulted."""
    cores = len(os.sched_getaffinity(0))
    os.environ["OMP_NUM_THREADS"] = str(cores)
    from oracle import orc
    return cores, orc.set_threads(cores)


def cpu_reference_rate(cfg, budget_s):
    """Returns (Mcell-updates/s, threads really used, sample description, ms/step)."""
    from oracle.solver import OracleTurbulentSolver
    _, threads = omp_team()
    s = OracleTurbulentSolver(cfg)
    s.time_step()                          # warm-up (page faults, OpenMP pool)
    t0 = time.perf_counter(); s.time_step(); one = time.perf_counter() - t0
    n = max(2, min(50, int(budget_s / max(one, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        s.time_step()
    t = time.perf_counter() - t0
    rate = cfg.nx * cfg.ny * n / t / 1e6
    return rate, threads, (f"{n} time steps of the {cfg.nx}x{cfg.ny} fp64 workload (oracle port = C/OpenMP restatement of "
                           f"v5.py, {threads} threads)"), t / n * 1e3


def run_reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores, same config / metric / unit as the
    b200 arm at this N.  Each step is a bounded sample of the workload (a sub-grid of the same solver configuration)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from oracle.solver import OracleTurbulentSolver
    n_gpus = max(args.gpus, int(os.environ.get("WORLD_SIZE", "1")))
    _, threads = omp_team()
    if n_gpus <= 1:
        full = (args.nx, args.ny)
        mk = lambda nx, ny: OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=False)
    else:
        full = (args.slab_nx * n_gpus, args.slab_ny)
        mk = lambda nx, ny: slab_config(nx, ny)
    # bounded sample: a sub-grid of the workload sized so that steps + warmup finish in about 150 s.  The rate per cell of
    # this stencil code does not depend on the grid size once it exceeds the caches.
    nx_s, ny_s = min(full[0], 4096), min(full[1], 1024)
    s = OracleTurbulentSolver(mk(nx_s, ny_s))
    s.time_step()
    t0 = time.perf_counter(); s.time_step(); one = time.perf_counter() - t0
    total = args.steps + args.warmup
    while one * total > 150.0 and nx_s > 512:
        nx_s //= 2; ny_s = max(ny_s // 2, 128); one /= 4
        s = None
    if s is None:
        s = OracleTurbulentSolver(mk(nx_s, ny_s))
    for _ in range(args.warmup):
        s.time_step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        s.time_step()
    t = time.perf_counter() - t0
    rate = nx_s * ny_s * args.steps / t / 1e6
    sample = (f"{args.steps} time steps on a {nx_s}x{ny_s} fp64 sub-grid of the workload ({full[0]}x{full[1]}, same solver "
              f"configuration); oracle port = C/OpenMP restatement of v5.py, {threads} OpenMP threads (team size read back)")
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": make_config(n_gpus, args),
            "sample_grid": [nx_s, ny_s],
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
def load_traffic():
    """DRAM bytes per launch from the committed ncu capture (profiles/traffic.json), with the capture it came from:
    {"capture": "r02x @ <commit>", "grid": "4096x1024", "kernels": {phase: bytes}}.  Never produced by this run."""
    tf = ROOT / "profiles" / "traffic.json"
    try:
        d = json.loads(tf.read_text())
    except Exception:
        return {}
    return d if "kernels" in d else {"capture": "r01g (untagged)", "grid": "4096x1024", "kernels": d}


def phase_table(prof, cells, peak):
    phases = {}
    for k, t_ms in prof.items():
        b = ALGO_BYTES.get(k, 0) * cells
        gbs = b / (t_ms * 1e-3) / 1e9 if t_ms > 0 else None
        phases[k] = {"ms": round(t_ms, 5), "algo_GBs": round(gbs, 1) if gbs else None, "frac": round(gbs / peak, 3) if gbs else None}
    return phases


def time_steps(s, steps, warmup, world=1):
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


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json configs[2]: the pressure-Poisson benchmark (SURVEY 8(d) C3 inputs)
def poisson_inputs(N, tdt, dev, seed=0):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    rhs = (0.1 * torch.randn((N, N), generator=g, dtype=torch.float64)).to(tdt).to(dev)
    yy, xx = torch.meshgrid(torch.arange(N, device=dev), torch.arange(N, device=dev), indexing="ij")
    mask = (((xx - N // 5) ** 2 + (yy - N // 2) ** 2) <= (N // 40) ** 2).to(torch.uint8).contiguous()
    return rhs, mask


def poisson_object(N=8192, sweeps=240, device=0):
    """Jacobi (v5.py:336-346) and red-black SOR (v5.py:202-226 with omega = 1.5) on an N x N fp64 grid: fixed sweep count
    and tolerance-terminated.  algo bytes = 25 B per cell per sweep (phi in, rhs, mask, phi out: what an un-blocked sweep
    must move); pass bytes = what the temporally blocked kernel moves per HBM pass."""
    import numpy as np
    import torch
    from cfd_simulations_b200 import kernels as K
    dev = torch.device("cuda", device)
    rhs, mask = poisson_inputs(N, torch.float64, dev)
    dx = dy = 1.0 / (N - 1)
    cdt = float(np.float32(5e-5))
    peak, peak_src = measured_peak()
    cells = (N - 2) * (N - 2)
    phi = torch.zeros((N, N), dtype=torch.float64, device=dev)
    out = {"workload": f"pressure-Poisson {N}x{N} fp64, disc mask radius N/40, rhs = 0.1*randn (BASELINE.json configs[2])",
           "algo_bytes_per_cell_sweep": 25, "peak": peak, "peak_source": peak_src, "modes": {}}

    def timed(fn):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); r = fn(); e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1), r

    def put(name, ms, n, info):
        glups = cells * n / (ms * 1e-3) / 1e9
        d = {"sweeps": n, "ms": round(ms, 4), "GLUPS": round(glups, 2), "algo_GBs": round(glups * 25, 1), "frac": round(glups * 25 / peak, 3)}
        d.update(info)
        out["modes"][name] = d

    for name, fn in (
        ("jacobi_fixed", lambda: K.solve_pressure_jacobi(phi.zero_(), rhs, dx, cdt, mask, sweeps, temporal_depth=0, return_info="full")),
        ("rbsor_fixed", lambda: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, sweeps, 0.0, omega=1.5,
                                                                   temporal_depth=0, return_info="full")),
        ("jacobi_fixed_depth1", lambda: K.solve_pressure_jacobi(phi.zero_(), rhs, dx, cdt, mask, sweeps, temporal_depth=1, return_info="full")),
        ("rbsor_fixed_depth1", lambda: K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, sweeps, 0.0, omega=1.5,
                                                                         temporal_depth=1, return_info="full")),
    ):
        ms, r = timed(fn)
        info = dict(r[3]) if isinstance(r, tuple) and len(r) > 3 and isinstance(r[3], dict) else {}
        depth = info.get("depth", 1)
        passes = info.get("passes", sweeps)
        pass_b = info.get("pass_bytes_per_cell", 25)
        put(name, ms, sweeps, {"depth": depth, "pass_GBs": round(cells * pass_b * passes / (ms * 1e-3) / 1e9, 1),
                               "pass_frac": round(cells * pass_b * passes / (ms * 1e-3) / 1e9 / peak, 3),
                               "pass_bytes_per_cell": pass_b})
    # tolerance-terminated red-black SOR: device-side stop, iteration count = the sequential algorithm's
    tol = 1e-3 * float(rhs.abs().max().item()) * dx * dx / cdt
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = K.solve_pressure_gauss_seidel_fast(phi.zero_(), rhs, dx, dy, cdt, mask, 20000, tol, omega=1.5, return_info=True)
    e1.record(); torch.cuda.synchronize()
    put("rbsor_tolerance", e0.elapsed_time(e1), int(r[1]), {"tol": tol, "iterations": int(r[1]), "last_change": float(r[2])})
    del phi, rhs, mask
    torch.cuda.empty_cache()
    return out


def poisson_object_slabs(s, N, sweeps, world, rank):
    """The red-black solve on x-slabs of N x N cells per GPU (what `bench_poisson.py --gpus N` prints)."""
    import torch
    import torch.distributed as dist
    peak, _ = measured_peak()
    t = s.device_tensor("phi")
    g = torch.Generator(device=s.device).manual_seed(1234 + rank)
    rhs = 0.1 * torch.randn(t.shape, generator=g, dtype=t.dtype, device=s.device)
    s._phi_exchanger().exchange([rhs])
    cells = (N - 2) * (N * world - 2)
    out = {"workload": f"red-black SOR (omega 1.5) on x-slabs of {N}x{N} fp64 cells per GPU, {sweeps} sweeps, NCCL halo refresh",
           "modes": {}}
    s.solve_pressure_slabs(rhs, max_iter=15, tol=0.0, omega=1.5)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _, it, last = s.solve_pressure_slabs(rhs, max_iter=sweeps, tol=0.0, omega=1.5, return_info=True)
    e1.record()
    torch.cuda.synchronize(); dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=s.device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    glups = cells * it / (ms * 1e-3) / 1e9
    out["modes"]["rbsor_fixed"] = {"sweeps": it, "ms": round(ms, 3), "GLUPS": round(glups, 1), "algo_GBs_per_gpu": round(glups * 25 / world, 1),
                                   "frac_per_gpu": round(glups * 25 / world / peak, 3), "last_change": last}
    return out


# ---------------------------------------------------------------------------------------------------------------
def nccl_parity(world, rank, local):
    """The N-rank slab run against ONE GPU on a small grid, over the real NCCL path, before anything is timed:
    rel-L2 of u, v, phi over the owned columns after PARITY_STEPS steps.  Every rank returns the same dict."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    nx, ny = PARITY_GRID
    cfg = OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=False)
    s = OptimizedTurbulentSolver(cfg, device=local, nranks=world, rank=rank)
    s.advance(PARITY_STEPS)
    got = {k: s.gather(k) for k in ("u", "v", "phi")}
    dts = [r["dt"] for r in (s._sync_scalars() or s.scalars)]
    s.close()
    res = torch.zeros(4, dtype=torch.float64, device=torch.device("cuda", local))
    if rank == 0:
        one = OptimizedTurbulentSolver(cfg, device=local)
        one.advance(PARITY_STEPS)
        errs = []
        for k in ("u", "v", "phi"):
            a, b = got[k], getattr(one, k)
            errs.append(float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)))
        one._sync_scalars()
        same_dt = float(all(x == y["dt"] for x, y in zip(dts, one.scalars)))
        one.close()
        res = torch.tensor(errs + [same_dt], dtype=torch.float64, device=torch.device("cuda", local))
    dist.broadcast(res, 0)
    e = res.tolist()
    worst = max(e[:3])
    return {"nranks": world, "grid": [nx, ny], "steps": PARITY_STEPS, "rel_l2": worst, "rel_l2_u_v_phi": e[:3],
            "dt_sequence_equal": bool(e[3]), "tol": PARITY_TOL, "ok": bool(worst <= PARITY_TOL and e[3]),
            "what": "N-rank x-slabs over NCCL vs the same grid on one GPU (owned columns, fp64)"}


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
    # ---- numerical parity over real NCCL first: a wrong exchange must not produce a number
    parity = nccl_parity(world, rank, local)
    if not parity["ok"]:
        if rank == 0:
            print(json.dumps({"error": "N-rank slab run does not match one GPU", "parity": parity}), flush=True)
        dist.destroy_process_group()
        return 3
    nxs, ny = args.slab_nx, args.slab_ny
    # ---- weak-scaling base: rank 0 alone steps ONE slab of the same configuration (the others wait)
    base_ms = None
    nb = max(5, min(args.steps, 30))
    if not args.no_scaling_base:
        if rank == 0:
            sb = OptimizedTurbulentSolver(slab_config(nxs, ny), device=local, setup="device")
            base_ms = time_steps(sb, nb, 3, 1) / nb
            sb.close(); del sb
            torch.cuda.empty_cache()
        dist.barrier()
    cfg = slab_config(nxs * world, ny)
    s = OptimizedTurbulentSolver(cfg, device=local, nranks=world, rank=rank, setup="device")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = time_steps(s, args.steps, args.warmup, world)
    clocks = sampler.stop() if rank == 0 else None
    cells = nxs * world * ny
    value = cells * args.steps / (ms * 1e-3) / 1e6
    # ---- where the slab step's time goes: CUDA events around the phases of a step (stream launches)
    slab_phases = s.profile_slab_steps(min(args.steps, 10))
    # ---- end to end: every step uploads this slab's u, v from pinned host memory, steps, downloads them and the scalars
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
    phase_names = [k for k in slab_phases["kernels"] if k != "ctl"]
    launches = s.launches_per_step_slab() * args.steps * world
    halo = {"halo_left": s._halo_req[0], "halo_right": s._halo_req[1], "exchanges_per_step": 1,
            "halo_bytes_per_step_per_gpu": s._halo.bytes_per_exchange, "collectives_per_step": s.collectives_per_step()}
    s.close(); del s, uh, vh
    torch.cuda.empty_cache()
    # ---- the Poisson half of the metric on slabs (8192^2 per GPU)
    pois = None
    if not args.no_poisson:
        from cfd_simulations_b200.config import OptimizedTurbulentConfig
        N = args.poisson_n
        h = 1.0 / (N - 1)
        pcfg = OptimizedTurbulentConfig(nx=N * world, ny=N, memory_efficient=False, x_min=0.0, x_max=(N * world - 1) * h, y_min=0.0,
                                        y_max=(N - 1) * h, cylinder_center=((N // 5) * h, (N // 2) * h), R_cylinder=(N // 40) * h)
        ps = OptimizedTurbulentSolver(pcfg, device=local, nranks=world, rank=rank, setup="device")
        pois = poisson_object_slabs(ps, N, args.poisson_sweeps, world, rank)
        ps.close()
    if rank == 0:
        peak, peak_src = measured_peak()
        step_bytes = sum(ALGO_BYTES[k] for k in phase_names) * cells
        ach = step_bytes / (ms / args.steps * 1e-3) / 1e9 / world
        tr = load_traffic()
        tk = tr.get("kernels_16384", None)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "vs_baseline_note": "the reference publishes no number for this metric (BASELINE.md section 1)",
                "dtype": "f64", "data": "synthetic", "config": make_config(world, args),
                "decomposition": halo, "clocks": clocks, "parity": parity,
                "weak_scaling": None if base_ms is None else {
                    "base_ms": base_ms, "base_steps": nb, "efficiency": base_ms / (ms / args.steps),
                    "base": f"the same {nxs}x{ny} slab configuration stepped by rank 0 alone in this run"},
                "slab_phases_ms": slab_phases,
                "e2e": {"value": cells * n_e2e / t_e2e / 1e6, "unit": UNIT, "h2d_bytes_per_step": slab_bytes * world,
                        "d2h_bytes_per_step": slab_bytes * world + 96 * world, "steps": n_e2e,
                        "ms_per_step": t_e2e / n_e2e * 1e3,
                        "note": "every step, every rank: pinned-host slab u,v -> HBM, time step, HBM -> host u,v + scalars"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "kernel": "whole step (sum of phases)", "achieved": ach, "peak": peak,
                             "unit": "GB/s", "frac": ach / peak,
                             "traffic": sum(tk.values()) if tk else None,
                             "traffic_source": (tr.get("capture") if tk else None),
                             "algo_bytes_per_cell": step_bytes // cells, "peak_source": peak_src,
                             "note": "per GPU: algorithmic bytes of all phases / step time; traffic = DRAM bytes of the step's "
                                     "kernels at 16384^2 on one GPU from the committed ncu capture (not this run)"},
                "poisson": pois, "cpu_baseline": None}
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()
    return 0


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
    ap.add_argument("--no-poisson", action="store_true", help="skip the configs[2] Poisson object")
    ap.add_argument("--poisson-n", type=int, default=8192)
    ap.add_argument("--poisson-sweeps", type=int, default=240)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.warmup < 3:
        args.warmup = 3

    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
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

    # ---- device-resident throughput: CUDA-graph replay, CUDA events, working set > L2
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
    ach = ALGO_BYTES[dom] * cells / (prof[dom] * 1e-3) / 1e9
    step_bytes = sum(ALGO_BYTES.get(k, 0) for k in prof) * cells
    tr = load_traffic()
    same_grid = tr.get("grid") == f"{nx}x{ny}"
    roofline = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": tr.get("kernels", {}).get(dom) if same_grid else None,
                "traffic_source": (f"ncu --set full capture {tr.get('capture')} of this command at {tr.get('grid')}: committed "
                                   f"under profiles/, not produced by this run") if same_grid else None,
                "peak_source": peak_src,
                "whole_step": {"algo_bytes_per_cell": step_bytes // cells, "achieved": step_bytes / (ms / args.steps * 1e-3) / 1e9,
                               "frac": step_bytes / (ms / args.steps * 1e-3) / 1e9 / peak},
                "phases": phase_table(prof, cells, peak),
                "phases_note": "phase times: CUDA events between the kernels of a step launched on the stream; `value` is "
                               "measured under CUDA-graph replay"}

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
    plan = s.clean_plan()
    s.close(); del s, uh, vh
    torch.cuda.empty_cache()

    # ---- CPU baseline on this box's host cores (bounded sample)
    cpu = None
    if not args.no_cpu:
        r, threads, sample, cpu_ms = cpu_reference_rate(cfg, args.cpu_budget)
        cpu = {"value": r, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample, "ms_per_step": cpu_ms}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "vs_baseline_note": "the reference publishes no number for this metric (BASELINE.md section 1); cpu_baseline is the same-run CPU figure",
            "dtype": "f64", "data": "synthetic", "config": make_config(1, args), "clean_plan": plan,
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu}
    # ---- the Poisson half of the metric: BASELINE.json configs[2]
    if not args.no_poisson:
        line["poisson"] = poisson_object(args.poisson_n, args.poisson_sweeps, local)
    if not args.no_scaling_base:
        # the 1-GPU point of the weak-scaling series (same slab as `--gpus N` uses per GPU)
        cfgs = slab_config(args.slab_nx, args.slab_ny)
        sb = OptimizedTurbulentSolver(cfgs, device=local, setup="device")
        nb = max(5, min(args.steps, 30))
        msb = time_steps(sb, nb, 3, 1)
        cb = args.slab_nx * args.slab_ny
        pb = sb.profile_steps(5)
        sbytes = sum(ALGO_BYTES.get(k, 0) for k in pb) * cb
        line["scaling_base"] = {"workload": f"{args.slab_nx}x{args.slab_ny} fp64 slab on 1 GPU (BASELINE.json configs[3] per-GPU share)",
                                "value": cb * nb / (msb * 1e-3) / 1e6, "unit": UNIT,
                                "ms_per_step": msb / nb, "steps": nb,
                                "whole_step_frac": sbytes / (msb / nb * 1e-3) / 1e9 / peak,
                                "phases": phase_table(pb, cb, peak)}
        sb.close()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    sys.exit(main() or 0)
