"""The reference's time loop (main(), v5.py:636-672) around the B200 solver — SURVEY §8(f) rank 2.

Reference loop, per step: dt = solver.time_step(); current_time += dt; every 20 steps monitor_simulation_health (break
when it fails, v5.py:656-660); every save_interval steps save_data_to_hdf5 (v5.py:662-664); initial state saved as step 0
(v5.py:648); stop when current_time >= T_total.

Here the same loop is driven in CHUNKS: between two events (health check, snapshot, possible end of the run) the solver
advances with `advance(n)` = n CUDA-graph replays and no host round trip; dt of every step is read afterwards from the
device ring (`solver.scalars`).  The run still ends on exactly the step the reference ends on: a chunk is never longer
than the number of steps that are certainly needed, (T_total - t) / dt_max (dt <= dt_max, v5.py:326), and once that
bound drops below one the loop single-steps.

`solver` may also be any object with the reference's surface (time_step(), u, v, compute_vorticity()) — the oracle or the
reference itself; it is then stepped one call at a time.
"""
from __future__ import annotations

import time
from dataclasses import dataclass, field

import numpy as np


@dataclass
class RunResult:
    steps: int = 0                 # time steps taken
    current_time: float = 0.0      # sum of dt
    healthy: bool = True           # False: the health monitor stopped the run (v5.py:658-660)
    saved_steps: list = field(default_factory=list)
    wall_s: float = 0.0
    dts: list = field(default_factory=list)


def _health_ok(solver, step) -> bool:
    if hasattr(solver, "health"):
        from .solver import monitor_simulation_health
        return bool(monitor_simulation_health(solver, step))
    u, v = solver.u, solver.v                                   # v5.py:599-613 on host arrays
    if not (np.all(np.isfinite(u)) and np.all(np.isfinite(v))):
        return False
    cfg = solver.config
    if max(np.max(np.abs(u)), np.max(np.abs(v))) > cfg.max_velocity:         # v5.py:604
        return False
    div = np.zeros_like(u)
    div[1:-1, 1:-1] = (u[1:-1, 2:] - u[1:-1, :-2]) / (2 * cfg.dx) + (v[2:, 1:-1] - v[:-2, 1:-1]) / (2 * cfg.dy)
    return float(np.max(np.abs(div))) <= (20.0 if step <= 1000 else 2.0)


def run_simulation(solver, writer=None, T_total=None, max_steps=None, health_interval=20, save_interval=None,
                   on_chunk=None) -> RunResult:
    """Run `solver` until current_time >= T_total (config.T_total by default) or `max_steps` steps, like main().

    writer: a snapshot.SnapshotWriter (or anything with save_data_to_hdf5(solver, step, t)); None = no snapshots.
    on_chunk(result): optional progress callback after every chunk (the reference's tqdm bar)."""
    cfg = solver.config
    T_total = float(cfg.T_total if T_total is None else T_total)
    save_interval = int(cfg.save_interval if save_interval is None else save_interval)
    dt_max = float(cfg.dt_max) if getattr(cfg, "adaptive_dt", True) else float(cfg.dt_base)
    chunked = hasattr(solver, "advance") and hasattr(solver, "scalars")
    acc32 = bool(getattr(cfg, "adaptive_dt", True))
    res = RunResult()
    t0 = time.perf_counter()
    step = 0                                   # the loop's own counter (v5.py:643), not solver.step
    if writer is not None:
        writer.save_data_to_hdf5(solver, 0, 0.0)                # v5.py:646-648
        res.saved_steps.append(0)
    while res.current_time < T_total and (max_steps is None or step < max_steps):
        # steps until the next event that needs the host: `step % interval == 0` is tested AFTER the time step with
        # the pre-increment counter, so the event fires once the step with index k*interval has been taken
        n = 1
        if chunked:
            to_health = (-step) % health_interval + 1
            to_save = (-step) % save_interval + 1 if writer is not None else 1 << 30
            # steps that cannot overshoot T_total.  The clock accumulates in float32 (see below): near t ~ 30 one add can
            # round up by ulp32(t) ~ 2e-6 against dt ~ 1e-4, so the bound counts that rounding in and keeps one step in hand
            ulp32 = float(np.spacing(np.float32(max(T_total, 1e-30)))) if acc32 else 0.0
            certain = int((T_total - res.current_time) / (dt_max + ulp32)) - 1
            n = max(1, min(to_health, to_save, certain, 4096))
            if max_steps is not None:
                n = max(1, min(n, max_steps - step))
            first = len(solver.scalars)
            solver.advance(n)
            solver._sync_scalars()
            dts = [r["dt"] for r in solver.scalars[first:first + n]]
        else:
            dts = [float(solver.time_step())]
        for dt in dts:
            # `current_time += dt` with dt an np.float32 (adaptive_time_step) accumulates in float32 under NumPy 2's
            # promotion rules (0.0 is a weak Python scalar) — the reference's clock, kept so the run ends on the same step
            res.current_time = float(np.float32(res.current_time) + np.float32(dt)) if acc32 else res.current_time + float(dt)
        res.dts.extend(float(d) for d in dts)
        last = step + n - 1                    # index of the last step taken (the reference's `step` at its checks)
        step += n
        res.steps = step
        if last % health_interval == 0 and not _health_ok(solver, last):
            res.healthy = False
            break
        if writer is not None and last % save_interval == 0:
            writer.save_data_to_hdf5(solver, last, res.current_time)
            if last not in res.saved_steps:
                res.saved_steps.append(last)
        if on_chunk is not None:
            on_chunk(res)
    if writer is not None:
        writer.flush()
    res.wall_s = time.perf_counter() - t0
    return res
