"""Snapshot writer — the on-disk side of the reference's driver.

Mirrors OptimizedVisualizer.save_data_to_hdf5 (v5.py:454-470): one group per saved step, named `step_%06d`, holding the
datasets `u`, `v`, `vorticity`, `X`, `Y` and the attribute `time`; a step that is already present is not rewritten
(v5.py:459).  That layout is what the reference's frame generator (v5.py:472-490) and animate_cylinder.py:20-35 read.

Backends
  * "hdf5": h5py, gzip level 4, exactly the reference's calls — used when h5py is importable;
  * "npz" : h5py is not in this image, so the default here is a directory `<hdf5_file>.d/` with one compressed
            `step_%06d.npz` per group (arrays u, v, vorticity, X, Y, 0-d `time`).  `load_snapshot` / `list_steps`
            read either backend, so post-processing code does not care.

B200 side: the time loop must not wait for PCIe or the disk.  For a GPU solver `save_data_to_hdf5`
  1. computes the vorticity ON THE DEVICE (b200cfd_k_vorticity) and copies u, v, vorticity into staging buffers with
     device-to-device copies on the compute stream (microseconds; the solver may overwrite u, v right after),
  2. copies the staging buffers to pinned host memory on a separate copy stream, ordered after (1) by an event,
  3. hands compression + file I/O to a worker thread (the reference's ThreadPoolExecutor(max_workers=2), v5.py:452)
     that first waits for the copy event.
Two staging sets are cycled, so at most two snapshots are in flight; a third `save` waits for the oldest.
A solver without `device_tensor` (the oracle, the reference itself) is written synchronously from its host arrays.
"""
from __future__ import annotations

from concurrent.futures import Future, ThreadPoolExecutor
from pathlib import Path

import numpy as np

DATASETS = ("u", "v", "vorticity", "X", "Y")


def group_name(step: int) -> str:
    return f"step_{int(step):06d}"          # v5.py:458


def _have_h5py() -> bool:
    try:
        import h5py  # noqa: F401
        return True
    except Exception:
        return False


class _Staging:
    """Device staging + pinned host buffers of one in-flight snapshot."""

    def __init__(self, like):
        import torch
        self.dev = {k: torch.empty_like(like) for k in ("u", "v", "vorticity")}
        self.host = {k: torch.empty(like.shape, dtype=like.dtype, pin_memory=True) for k in ("u", "v", "vorticity")}
        self.done = torch.cuda.Event()
        self.future: Future | None = None


class SnapshotWriter:
    """Drop-in for the data-saving half of OptimizedVisualizer (v5.py:443-470)."""

    def __init__(self, config, backend: str = "auto", max_workers: int = 2, compress: bool = True):
        self.config = config
        self.output_path = Path(config.output_dir)
        self.output_path.mkdir(parents=True, exist_ok=True)
        if backend == "auto":
            backend = "hdf5" if _have_h5py() else "npz"
        if backend not in ("hdf5", "npz"):
            raise ValueError(f"unknown snapshot backend {backend!r}")
        if backend == "hdf5" and not _have_h5py():
            raise RuntimeError("backend 'hdf5' needs h5py, which is not installed")
        self.backend = backend
        self.compress = compress
        self.path = self.output_path / config.hdf5_file if backend == "hdf5" else self.output_path / (config.hdf5_file + ".d")
        if backend == "npz":
            self.path.mkdir(exist_ok=True)
        # one writer thread for HDF5 (a single file is not safe to append from two threads), two for npz files
        self.executor = ThreadPoolExecutor(max_workers=1 if backend == "hdf5" else max_workers)
        self._staging: list[_Staging] = []
        self._next = 0
        self._copy_stream = None
        self._grid_saved: dict[str, np.ndarray] = {}
        self.saved_steps: list[int] = []

    # ---- the reference's entry point -------------------------------------------------------------------------
    def save_data_to_hdf5(self, solver, step: int, current_time: float) -> Future:
        """Queue the snapshot of `solver` at `step`; returns a Future that resolves to the group name once it is on
        disk.  Errors are kept in the Future (the reference logs and carries on, v5.py:469-470)."""
        step, current_time = int(step), float(current_time)
        if self._exists(step):
            f: Future = Future()
            f.set_result(group_name(step))
            return f
        nranks = getattr(solver, "_nranks", 1)
        X, Y = getattr(solver, "X", None), getattr(solver, "Y", None)
        if X is None or Y is None or nranks > 1:
            # solvers set up on the device keep no host grid, and a slab only knows its own columns: the file holds the
            # GLOBAL grid of the reference layout (v5.py:269-272)
            cfg = solver.config
            X, Y = np.meshgrid(np.linspace(cfg.x_min, cfg.x_max, cfg.nx), np.linspace(cfg.y_min, cfg.y_max, cfg.ny))
        grids = {"X": np.asarray(X), "Y": np.asarray(Y)}
        if nranks > 1:
            # x-slabs: every rank calls this; the owned columns of u, v and of the (device-computed) vorticity are gathered
            # and rank 0 alone writes the global field.  The other ranks get an already resolved Future.
            w = solver.compute_vorticity_device()
            arrays = {"u": solver.gather_to_root("u"), "v": solver.gather_to_root("v"), "vorticity": solver.gather_to_root(w)}
            if getattr(solver, "_rank", 0) == 0:
                fut = self.executor.submit(self._write, step, current_time, arrays, grids)
            else:
                fut = Future()
                fut.set_result(group_name(step))
        elif hasattr(solver, "device_tensor"):
            fut = self._save_device(solver, step, current_time, grids)
        else:
            arrays = {"u": np.array(solver.u), "v": np.array(solver.v), "vorticity": np.array(solver.compute_vorticity())}
            fut = self.executor.submit(self._write, step, current_time, arrays, grids)
        self.saved_steps.append(step)
        return fut

    def flush(self):
        """Wait until every queued snapshot is on disk; re-raises the first writer error."""
        for st in self._staging:
            if st.future is not None:
                st.future.result()
        self.executor.submit(lambda: None).result()

    def close(self):
        self.flush()
        self.executor.shutdown(wait=True)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- device path --------------------------------------------------------------------------------------------
    def _save_device(self, solver, step, current_time, grids) -> Future:
        import torch
        u, v = solver.device_tensor("u"), solver.device_tensor("v")
        dev = u.device
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        if len(self._staging) < 2:
            self._staging.append(_Staging(u))
            st = self._staging[-1]
        else:
            st = self._staging[self._next % 2]          # the older of the two sets
        self._next += 1
        if st.future is not None:
            st.future.result()                      # the set is still being written: wait for it (back-pressure)
        compute = torch.cuda.current_stream(dev)
        with torch.cuda.device(dev):
            solver.compute_vorticity_device(out=st.dev["vorticity"])     # device kernel, compute stream
            st.dev["u"].copy_(u, non_blocking=True)
            st.dev["v"].copy_(v, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(compute)
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(ready)
                for k in ("u", "v", "vorticity"):
                    st.host[k].copy_(st.dev[k], non_blocking=True)
                st.done.record(self._copy_stream)

        def job():
            st.done.synchronize()
            return self._write(step, current_time, {k: st.host[k].numpy() for k in ("u", "v", "vorticity")}, grids)

        st.future = self.executor.submit(job)
        return st.future

    # ---- file formats ---------------------------------------------------------------------------------------------
    def _exists(self, step) -> bool:
        if step in self.saved_steps:
            return True
        if self.backend == "npz":
            return (self.path / (group_name(step) + ".npz")).exists()
        if not self.path.exists():
            return False
        import h5py
        with h5py.File(self.path, "r") as f:
            return group_name(step) in f

    def _write(self, step, current_time, arrays, grids) -> str:
        name = group_name(step)
        data = dict(arrays)
        for k in ("X", "Y"):
            if grids.get(k) is not None:
                data[k] = grids[k]
        if self.backend == "hdf5":
            import h5py
            with h5py.File(self.path, "a") as f:                       # v5.py:457-467
                if name not in f:
                    g = f.create_group(name)
                    g.attrs["time"] = current_time
                    for k, a in data.items():
                        g.create_dataset(k, data=a, compression="gzip", compression_opts=4)
        else:
            tmp = self.path / (name + ".tmp.npz")
            (np.savez_compressed if self.compress else np.savez)(tmp, time=np.float64(current_time), **data)
            tmp.replace(self.path / (name + ".npz"))                    # readers never see a partial file
        return name


# ---- readers (what generate_frames_from_hdf5, v5.py:472-490, does with the file) ------------------------------------
def _resolve(path) -> tuple[Path, str]:
    p = Path(path)
    if p.is_dir():
        return p, "npz"
    if p.exists():
        return p, "hdf5"
    d = p.with_name(p.name + ".d")
    if d.is_dir():
        return d, "npz"
    raise FileNotFoundError(path)


def list_steps(path) -> list[int]:
    """Saved step numbers, ascending (v5.py:480-481)."""
    p, kind = _resolve(path)
    if kind == "npz":
        return sorted(int(f.stem.split("_")[1]) for f in p.glob("step_*.npz") if ".tmp" not in f.name)
    import h5py
    with h5py.File(p, "r") as f:
        return sorted(int(k.split("_")[1]) for k in f.keys() if k.startswith("step_"))


def load_snapshot(path, step) -> dict:
    """{'u','v','vorticity','X','Y','time'} of one saved step."""
    p, kind = _resolve(path)
    if kind == "npz":
        with np.load(p / (group_name(step) + ".npz")) as z:
            out = {k: z[k] for k in z.files}
        out["time"] = float(out["time"])
        return out
    import h5py
    with h5py.File(p, "r") as f:
        g = f[group_name(step)]
        out = {k: g[k][:] for k in g.keys()}
        out["time"] = float(g.attrs["time"])
        return out
