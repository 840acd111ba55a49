"""Host-side mirror of the reference's solver class, backed by libb200cfd.so (CUDA, sm_100a).

Reference: python/flow_over_cylinder (Fischer)/v5.py:259-441 (OptimizedTurbulentSolver) and :599-613
(monitor_simulation_health).  Method and attribute names are the reference's; the field state lives in HBM
(torch CUDA tensors own the buffers, nothing else of torch is used) and only scalars come back per step.
There is no CPU fallback: constructing a solver without the CUDA library / a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib, hostsetup
from ._lib import B200CFDError, Config, StepScalars, check
from .config import OptimizedTurbulentConfig  # noqa: F401


def _torch_dtype(np_dtype):
    return torch.float32 if np_dtype == np.float32 else torch.float64


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def make_c_config(cfg, ny, nx, poisson_mode, cleandiv_mode, omega, clean_iters, rank=0, nranks=1, nx_global=None,
                  x_offset=0, halo_left=0, halo_right=0):
    if not cfg.use_fast_pressure:
        poisson_mode = "jacobi"          # v5.py:336: the NumPy Jacobi branch
    return Config(
        ny=ny, nx=nx, dtype=_lib.F32 if cfg.field_dtype == np.float32 else _lib.F64,
        poisson_mode=_lib.POISSON_MODES[poisson_mode], pressure_iterations=int(cfg.pressure_iterations),
        cleandiv_mode=_lib.CLEANDIV_MODES[cleandiv_mode], clean_iters=int(clean_iters),
        initial_steps=int(cfg.initial_steps), use_supg=int(cfg.use_supg), use_les=int(cfg.use_les),
        adaptive_dt=int(cfg.adaptive_dt), dx=float(cfg.dx), dy=float(cfg.dy), nu=float(cfg.nu),
        av=float(cfg.artificial_viscosity), cfg_dt=float(cfg.dt), V_inf=float(cfg.V_inf), y_max=float(cfg.y_max),
        cfl_target=float(cfg.cfl_target), dt_min=float(cfg.dt_min), dt_max=float(cfg.dt_max),
        dt_base=float(cfg.dt_base), max_velocity=float(cfg.max_velocity), cs=float(cfg.smagorinsky_constant),
        pressure_tolerance=float(cfg.pressure_tolerance), omega=float(omega),
        rank=rank, nranks=nranks, nx_global=nx if nx_global is None else nx_global, x_offset=x_offset,
        halo_left=halo_left, halo_right=halo_right,
        pert_amplitude=float(getattr(cfg, "perturbation_amplitude", 0.01)),
        scalar_f32=int(getattr(cfg, "float32_scalars", True)), variant_set=1)


class OptimizedTurbulentSolver:
    """Drop-in for the reference class (v5.py:259).  Extra keyword arguments select how the reference's
    ambiguous pieces are interpreted (SURVEY findings 3, 4):
      poisson_mode  "as_executed" (1 red-black iteration: what Numba parallel=True really runs, default),
                    "as_written" (tolerance-terminated, <= pressure_iterations), "jacobi" (use_fast_pressure=False)
      cleandiv_mode "lexicographic" (the 1-thread reference order, default) | "redblack" (not reference parity)
    """

    def __init__(self, config, device=0, poisson_mode="as_executed", cleandiv_mode="lexicographic", omega=1.0,
                 clean_iters=2, use_graph=True, nranks=1, rank=0, group=None, setup="host", materialize_attributes=True):
        """materialize_attributes=False: the two arrays the reference keeps only as by-products of a step, `tau_supg`
        (v5.py:392) and `div_u_star` (v5.py:409), are not written (b200cfd_bind_fields takes NULL for them): 176 instead of
        192 B/cell per step; u, v, phi, u_star, v_star and every logged scalar are unchanged, reading the two attributes raises.
        nranks > 1: this process holds x-slab `rank` of `nranks` (multigpu.py; torch.distributed must be
        initialised, NCCL).  setup = "host" (NumPy, bit-identical to the reference's arrays) | "device" (torch on the
        GPU, for slabs too large for comfortable host setup; p is not allocated)."""
        if not torch.cuda.is_available():
            raise B200CFDError("no CUDA device: cfd_simulations_b200 has no CPU fallback")
        self.config = cfg = config
        self.device = torch.device("cuda", device)
        self._lib = _lib.load()
        self._use_graph = bool(use_graph)
        self._nranks, self._rank, self._group = int(nranks), int(rank), group
        from . import multigpu
        hl = hr = C.c_int(0)
        if nranks > 1:
            hl, hr = C.c_int(0), C.c_int(0)
            probe = make_c_config(cfg, cfg.ny, cfg.nx, poisson_mode, cleandiv_mode, omega, clean_iters)
            check(self._lib.b200cfd_required_halo(C.byref(probe), C.byref(hl), C.byref(hr)))
        self._halo_req = (hl.value, hr.value)       # the same numbers on every rank
        self.plan = plan = multigpu.make_plan(cfg.nx, self._nranks, self._rank, hl.value, hr.value)
        cols = None if nranks == 1 else (plan.x_offset, plan.nx_local)
        npdt = cfg.field_dtype
        tdt = _torch_dtype(npdt)
        dev = self.device
        nxl = plan.nx_local
        z = lambda: torch.zeros((cfg.ny, nxl), dtype=tdt, device=dev)
        if setup == "device":
            self.x, self.y, mask_d, ibm_d, u_d, v_d = hostsetup.setup_on_device(cfg, dev, cols)
            self.X = self.Y = self.dist = self.cylinder_mask = self.ibm_mask = None
        else:
            # ---- host setup exactly as the reference (v5.py:269-314) ----
            self.x, self.y, self.X, self.Y = hostsetup.setup_grid(cfg, cols)
            self.dist, self.cylinder_mask, self.ibm_mask = hostsetup.setup_boundary_masks(cfg, self.X, self.Y)
            u0, v0 = hostsetup.initialize_potential_flow(cfg, self.X, self.Y, self.dist, self.ibm_mask, npdt)
            u_d, v_d = torch.from_numpy(u0).to(dev), torch.from_numpy(v0).to(dev)
            ibm_d = torch.from_numpy(np.ascontiguousarray(self.ibm_mask, dtype=np.float64)).to(dev)
            mask_d = torch.from_numpy(np.ascontiguousarray(self.cylinder_mask, dtype=np.uint8)).to(dev)
        self._d = {
            "u": u_d, "v": v_d,
            "phi": z(), "u_star": z(), "v_star": z(), "div_u_star": z(), "nu_t": z(), "tau_supg": z(),
            "ibm": ibm_d, "mask": mask_d,
            "y": torch.from_numpy(np.ascontiguousarray(self.y, dtype=np.float64)).to(dev),
        }
        if setup != "device":
            self._d["p"] = z()           # v5.py:290: allocated, never updated in v3-v5
        self._ccfg = make_c_config(cfg, cfg.ny, nxl, poisson_mode, cleandiv_mode, omega, clean_iters, rank=self._rank,
                                   nranks=self._nranks, nx_global=cfg.nx, x_offset=plan.x_offset,
                                   halo_left=plan.halo_left, halo_right=plan.halo_right)
        self._h = C.c_void_p()
        check(self._lib.b200cfd_create(C.byref(self._ccfg), device, C.byref(self._h)))
        d = self._d
        self._attributes = bool(materialize_attributes)
        if not self._attributes:
            del d["div_u_star"], d["tau_supg"]
        check(self._lib.b200cfd_bind_fields(self._h, d["u"].data_ptr(), d["v"].data_ptr(), d["phi"].data_ptr(),
                                            d["u_star"].data_ptr(), d["v_star"].data_ptr(),
                                            d["div_u_star"].data_ptr() if self._attributes else None,
                                            d["tau_supg"].data_ptr() if self._attributes else None, d["nu_t"].data_ptr(), d["ibm"].data_ptr(),
                                            d["mask"].data_ptr(), d["y"].data_ptr()), self._h)
        self._step = 0
        self._fetched = 0            # scalars of steps [first_step, _fetched) are in self.scalars
        self._first_step = 0
        self.scalars = []            # list of dicts, one per step (the five logged values + dt ...)
        self._energy_history = []
        self._times = []
        self._halo = self._phi_halo = None
        if self._nranks > 1:
            n_red = int(self._lib.b200cfd_reduction_len())
            self._red = multigpu.device_view(self._lib.b200cfd_reduction_ptr(self._h), n_red, dev)
            self._gathered = torch.zeros((self._nranks, n_red), dtype=torch.float64, device=dev)
            lib, h = self._lib, self._h
            sp = lambda: _stream_ptr(dev)
            self._halo = multigpu.HaloExchanger(
                plan, cfg.ny, 2, alloc=lambda shape: torch.empty(shape, dtype=tdt, device=dev),
                pack=lambda f, j0, n, out: check(lib.b200cfd_pack_columns(h, f.data_ptr(), j0, n, out.data_ptr(), sp()), h),
                unpack=lambda f, j0, n, src: check(lib.b200cfd_unpack_columns(h, f.data_ptr(), j0, n, src.data_ptr(), sp()), h),
                group=group, halo_rows=self._halo_rows)
            # halo exchange off the compute stream: its own stream, ordered against the step by events
            self._comm_stream = torch.cuda.Stream(device=dev)
            self._ev_edges = torch.cuda.Event()
            self._ev_halo = torch.cuda.Event()
            self._overlap = bool(lib.b200cfd_step_splits(self._h)) and not bool(os.environ.get("B200CFD_NO_OVERLAP"))
        torch.cuda.synchronize(dev)

    # ---- lifetime ----
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b200cfd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference attributes: host views of device state (lazy D2H) ----
    def _get(self, name):
        if name not in self._d:
            raise B200CFDError(f"{name} is not materialised (materialize_attributes=False)")
        return self._d[name].cpu().numpy()

    def _set(self, name, value):
        t = self._d[name]
        t.copy_(torch.as_tensor(np.ascontiguousarray(value), dtype=t.dtype))
        if name in ("u", "v"):
            self._set_step(self._step)

    u = property(lambda s: s._get("u"), lambda s, a: s._set("u", a))
    v = property(lambda s: s._get("v"), lambda s, a: s._set("v", a))
    p = property(lambda s: s._get("p"), lambda s, a: s._set("p", a))
    phi = property(lambda s: s._get("phi"), lambda s, a: s._set("phi", a))
    u_star = property(lambda s: s._get("u_star"), lambda s, a: s._set("u_star", a))
    v_star = property(lambda s: s._get("v_star"), lambda s, a: s._set("v_star", a))
    div_u_star = property(lambda s: s._get("div_u_star"), lambda s, a: s._set("div_u_star", a))
    nu_t = property(lambda s: s._get("nu_t"), lambda s, a: s._set("nu_t", a))
    tau_supg = property(lambda s: s._get("tau_supg"), lambda s, a: s._set("tau_supg", a))

    def device_tensor(self, name):
        """The torch CUDA tensor that owns a field (no copy)."""
        return self._d[name]

    @property
    def step(self):
        return self._step

    @step.setter
    def step(self, value):
        self._sync_scalars()
        self._step = int(value)
        self._fetched = self._first_step = self._step
        self._set_step(self._step)

    def _set_step(self, step):
        """self.step = step: re-derive what adaptive_time_step reads (max|u|, max|v| of the current state)."""
        if self._nranks == 1:
            check(self._lib.b200cfd_set_step(self._h, step, _stream_ptr(self.device)), self._h)
        else:
            hv = self.health()       # global maxima (all_reduce inside)
            check(self._lib.b200cfd_set_step_global(self._h, step, hv["max_abs_u"], hv["max_abs_v"],
                                                    _stream_ptr(self.device)), self._h)

    @property
    def energy_history(self):
        self._sync_scalars()
        return self._energy_history

    @property
    def times(self):
        self._sync_scalars()
        return self._times

    # ---- stepping ----
    def advance(self, n_steps):
        """n x time_step() with no host round trip (state and per-step scalars stay on the device)."""
        cap = 8192
        done = 0
        if self._nranks > 1:
            return self._advance_slabs(n_steps, cap)
        while done < n_steps:
            room = cap - (self._step - self._fetched)
            if room <= 0:
                self._sync_scalars()
                continue
            m = min(n_steps - done, room)
            check(self._lib.b200cfd_step(self._h, m, int(self._use_graph), _stream_ptr(self.device)), self._h)
            self._step += m
            done += m

    def _halo_rows(self, pack, fields, jl, nl, bl, jr, nr, br):
        """All fields and both sides of a halo refresh in one launch (b200cfd_halo_rows) on the current stream."""
        ptrs = (C.c_void_p * len(fields))(*[f.data_ptr() for f in fields])
        check(self._lib.b200cfd_halo_rows(self._h, int(bool(pack)), ptrs, len(fields), int(jl), int(nl if bl is not None else 0),
                                          bl.data_ptr() if bl is not None else None, int(jr), int(nr if br is not None else 0),
                                          br.data_ptr() if br is not None else None, _stream_ptr(self.device)), self._h)

    def _advance_slabs(self, n_steps, cap):
        """Slab stepping.  Per step: one all_gather of the 128-byte reduction vector and one halo refresh of u, v.  When the
        step splits (fast kernels), the refresh runs on its own stream and overlaps with the bulk of the finalize kernel:
          compute stream: part 1 (... finalize of the edge columns) | part 2 (finalize of the rest, local reduction) |
                          all_gather | closing kernel | wait for the halos
          comm stream:                 wait part 1 | pack (1 launch) | NCCL send/recv | unpack (1 launch)"""
        import torch.distributed as dist
        lib, h, sp = self._lib, self._h, lambda: _stream_ptr(self.device)
        u, v = self._d["u"], self._d["v"]
        main = torch.cuda.current_stream(self.device)
        for _ in range(n_steps):
            if self._step - self._fetched >= cap:
                self._sync_scalars()
            if self._overlap:
                check(lib.b200cfd_step_part(h, 1, int(self._use_graph), sp()), h)
                self._ev_edges.record(main)
                with torch.cuda.stream(self._comm_stream):
                    self._comm_stream.wait_event(self._ev_edges)
                    self._halo.exchange([u, v])
                    self._ev_halo.record(self._comm_stream)
                check(lib.b200cfd_step_part(h, 2, int(self._use_graph), sp()), h)
                dist.all_gather_into_tensor(self._gathered.view(-1), self._red, group=self._group)
                check(lib.b200cfd_step_end(h, self._gathered.data_ptr(), self._nranks, sp()), h)
                main.wait_event(self._ev_halo)
            else:
                check(lib.b200cfd_step_begin(h, int(self._use_graph), sp()), h)
                dist.all_gather_into_tensor(self._gathered.view(-1), self._red, group=self._group)
                check(lib.b200cfd_step_end(h, self._gathered.data_ptr(), self._nranks, sp()), h)
                self._halo.exchange([u, v])
            self._step += 1

    def gather(self, name):
        """Global (ny, nx) array of a field assembled from the owned columns of every slab (all ranks get it)."""
        loc = self._get(name)[:, self.plan.own0:self.plan.own1]
        if self._nranks == 1:
            return loc
        import torch.distributed as dist
        parts = [None] * self._nranks
        dist.all_gather_object(parts, loc, group=self._group)
        return np.concatenate(parts, axis=1)

    def set_global(self, name, array):
        """Assign a field from a GLOBAL (ny, nx) host array (each slab takes its columns, halos included)."""
        p = self.plan
        self._set(name, np.ascontiguousarray(array[:, p.x_offset:p.x_offset + p.nx_local]))

    def time_step(self):
        """v5.py:375-441.  Returns this step's dt (np.float32 when adaptive, as the reference)."""
        self.advance(1)
        self._sync_scalars()
        dt = self.scalars[-1]["dt"]
        return np.float32(dt) if (self.config.adaptive_dt and getattr(self.config, "float32_scalars", True)) else dt

    def time_step_host(self, u_host, v_host, n_steps=1):
        """time_step() on HOST arrays, the way the reference owns its state: uploads u, v, runs n_steps,
        downloads u, v and the last step's scalars (used for the end-to-end measurement)."""
        if self._nranks > 1:
            raise B200CFDError("time_step_host is a single-GPU entry point")
        last = StepScalars()
        check(self._lib.b200cfd_step_host(self._h, u_host.ctypes.data, v_host.ctypes.data, n_steps,
                                          int(self._use_graph), C.byref(last), _stream_ptr(self.device)), self._h)
        self._step += n_steps
        self._fetched = self._first_step = self._step     # skipped records are not replayed into the histories
        return last

    def _sync_scalars(self):
        n = self._step - self._fetched
        if n <= 0:
            return
        buf = (StepScalars * n)()
        check(self._lib.b200cfd_fetch_scalars(self._h, self._fetched, n, buf, _stream_ptr(self.device)), self._h)
        for r in buf:
            rec = {k: getattr(r, k) for k, _ in StepScalars._fields_}
            self.scalars.append(rec)
            st = int(rec["step"])
            self._energy_history.append((st, rec["mean_energy"]))     # v5.py:433
            # v5.py:434 `self.step * dt`: with dt an np.float32 (v5) the product is a float32 (NumPy 2 promotion)
            self._times.append(float(np.float32(st) * np.float32(rec["dt"])) if getattr(self.config, "float32_scalars", True)
                               and self.config.adaptive_dt else st * rec["dt"])
        self._fetched = self._step

    def adaptive_time_step(self):
        """v5.py:316-326 — evaluated on the host from the device-reduced post-clip maxima."""
        cfg = self.config
        if not cfg.adaptive_dt:
            return cfg.dt_base
        if self._step < 1000:
            return np.float32(0.00002)
        hv = self.health()
        vel_max = max(hv["max_abs_u"], hv["max_abs_v"], 1e-10)
        dt_cfl = cfg.cfl_target * min(cfg.dx, cfg.dy) / vel_max
        nu_total = cfg.nu + np.mean(self.nu_t) + cfg.artificial_viscosity
        dt_visc = 0.4 * min(cfg.dx, cfg.dy) ** 2 / nu_total
        return np.float32(np.clip(min(dt_cfl, dt_visc), cfg.dt_min, cfg.dt_max))

    # ---- the reference's "solve" entry point ----
    def solve_pressure_fast(self, div_u_star, mode=None, max_iter=None, tol=None, omega=None, temporal_depth=0,
                            return_info=False):
        """v5.py:328-347: zero phi, then the configured Poisson solve on `div_u_star` (numpy or CUDA tensor)."""
        cfg = self.config
        t = self._d["phi"]
        if self._nranks > 1:
            raise B200CFDError("solve_pressure_fast is the single-slab entry point: on x-slabs use solve_pressure_slabs "
                               "(halo refresh + all-reduced max|dphi|)")
        if isinstance(div_u_star, torch.Tensor):
            rhs = div_u_star.to(device=self.device, dtype=t.dtype).contiguous()
        else:
            rhs = torch.as_tensor(np.ascontiguousarray(div_u_star), dtype=t.dtype).to(self.device)
        t.zero_()
        mode_i = self._ccfg.poisson_mode if mode is None else _lib.POISSON_MODES[mode]
        it, last = C.c_int(0), C.c_double(0.0)
        check(self._lib.b200cfd_poisson(
            self._h, t.data_ptr(), rhs.data_ptr(), self._d["mask"].data_ptr(), cfg.dx, cfg.dy, float(cfg.dt), mode_i,
            int(cfg.pressure_iterations if max_iter is None else max_iter),
            float(cfg.pressure_tolerance if tol is None else tol), float(self._ccfg.omega if omega is None else omega),
            int(temporal_depth), C.byref(it), C.byref(last), _stream_ptr(self.device)), self._h)
        out = t if isinstance(div_u_star, torch.Tensor) else t.cpu().numpy()
        return (out, it.value, last.value) if return_info else out

    # ---- iterative Poisson across x-slabs (halo refresh per pass + all-reduced residual) ----
    def _poisson_iterations(self, rhs, n, omega, temporal_depth=1):
        """n red-black iterations of solve_pressure_gauss_seidel_fast on the LOCAL array (owned columns + halos), no
        tolerance stop, no host synchronisation; max|dphi| of the last one stays on the device.  temporal_depth = 1: one
        launch pair per iteration (the per-iteration change is observable); 0: the TMA temporally blocked kernel."""
        cfg = self.config
        check(self._lib.b200cfd_poisson(
            self._h, self._d["phi"].data_ptr(), rhs.data_ptr(), self._d["mask"].data_ptr(), cfg.dx, cfg.dy, float(cfg.dt),
            _lib.POISSON_AS_WRITTEN, int(n), 0.0, float(omega), int(temporal_depth), None, None,
            _stream_ptr(self.device)), self._h)

    def _jacobi_sweeps(self, rhs, n):
        """n sweeps of the Jacobi branch (v5.py:336-346: dx only, 0.25, cfg.dt) on the LOCAL array, no host sync."""
        cfg = self.config
        check(self._lib.b200cfd_poisson(
            self._h, self._d["phi"].data_ptr(), rhs.data_ptr(), self._d["mask"].data_ptr(), cfg.dx, cfg.dy, float(cfg.dt),
            _lib.POISSON_JACOBI, int(n), 0.0, 1.0, 0, None, None, _stream_ptr(self.device)), self._h)

    def _phi_exchanger(self):
        if self._phi_halo is None and self._nranks > 1:
            from . import multigpu
            lib, h, dev = self._lib, self._h, self.device
            sp = lambda: _stream_ptr(dev)
            tdt = self._d["phi"].dtype
            self._phi_halo = multigpu.HaloExchanger(
                self.plan, self.config.ny, 1, alloc=lambda shape: torch.empty(shape, dtype=tdt, device=dev),
                pack=lambda f, j0, n, out: check(lib.b200cfd_pack_columns(h, f.data_ptr(), j0, n, out.data_ptr(), sp()), h),
                unpack=lambda f, j0, n, src: check(lib.b200cfd_unpack_columns(h, f.data_ptr(), j0, n, src.data_ptr(), sp()), h),
                group=self._group)
        return self._phi_halo

    def solve_pressure_slabs(self, div_u_star, max_iter=None, tol=None, omega=None, return_info=False, mode="as_written"):
        """solve_pressure_fast / solve_pressure_gauss_seidel_fast "as written" (v5.py:202-226, 328-347) on a grid that is
        slab-decomposed over the ranks.  `div_u_star`: this rank's LOCAL (ny, nx_local) right-hand side, halo columns
        included (`set_global`-style slice of the global array), numpy or CUDA tensor.  Communication-avoiding: the halos
        are refreshed (NCCL send/recv) once per T = min(halo)/2 iterations, the region of validity shrinking by one column
        per colour sweep in between; in tolerance mode the per-iteration max|dphi| of a chunk is all-reduced (MAX) in ONE
        collective and, if the stopping iteration lies inside the chunk, the chunk is replayed up to exactly that
        iteration — so iterates, iteration count and last change are those of the single-GPU solve."""
        from . import multigpu
        import torch.distributed as dist
        cfg, t = self.config, self._d["phi"]
        if isinstance(div_u_star, torch.Tensor):
            rhs = div_u_star.to(device=self.device, dtype=t.dtype).contiguous()
        else:
            rhs = torch.as_tensor(np.ascontiguousarray(div_u_star), dtype=t.dtype).to(self.device)
        if tuple(rhs.shape) != tuple(t.shape):
            raise ValueError(f"local right-hand side must have the slab shape {tuple(t.shape)}, got {tuple(rhs.shape)}")
        ex = self._phi_exchanger()

        def reduce_max(parts):
            g = parts[0].clone()
            if self._nranks > 1:
                dist.all_reduce(g, op=dist.ReduceOp.MAX, group=self._group)
            return g

        if mode == "jacobi":       # v5.py:336-346: fixed sweep count, halos refreshed once per min(halo) sweeps
            it = multigpu.slab_jacobi([self], [rhs], int(cfg.pressure_iterations if max_iter is None else max_iter),
                                      exchange=(lambda: ex.exchange([t])) if ex is not None else (lambda: None))
            out = t if isinstance(div_u_star, torch.Tensor) else t.cpu().numpy()
            return (out, it, 0.0) if return_info else out
        if mode != "as_written":
            raise ValueError(f"solve_pressure_slabs: mode {mode!r} (as_written | jacobi)")
        it, last = multigpu.slab_poisson(
            [self], [rhs], int(cfg.pressure_iterations if max_iter is None else max_iter),
            float(cfg.pressure_tolerance if tol is None else tol), float(self._ccfg.omega if omega is None else omega),
            exchange=(lambda: ex.exchange([t])) if ex is not None else (lambda: None), reduce_max=reduce_max)
        out = t if isinstance(div_u_star, torch.Tensor) else t.cpu().numpy()
        return (out, it, last) if return_info else out

    def solve_pressure_iterative(self, div_u_star, max_iter=500, tol=1e-7, omega=1.5, return_info=False):
        """v2.py:284-306 (RobustTurbulentSolver.solve_pressure_iterative; v1.py:275-297 with 200 / 1e-6): the SOR fallback
        of the v1/v2 solvers on lap(phi) = +div/dt with Neumann left/bottom/top and Dirichlet-0 right walls, masked cells 0.
        Red-black ordering on the GPU: converges to the same field as the reference's lexicographic sweep."""
        return self.solve_pressure_fast(div_u_star, mode="sor_v2", max_iter=max_iter, tol=tol, omega=omega, temporal_depth=1,
                                        return_info=return_info)

    def apply_boundary_conditions(self, u, v):
        """v5.py:349-360, in place on numpy arrays or CUDA tensors, at the solver's current step."""
        from . import kernels
        kernels.apply_boundary_conditions(u, v, self.y, self.config.y_max, self.config.V_inf, self._step, handle=self)

    # ---- host-side diagnostics the reference's visualizer calls (not on the hot path) ----
    def compute_energy(self):
        u, v = self.u, self.v
        return 0.5 * (u ** 2 + v ** 2)

    def compute_vorticity(self):
        """v5.py:365-373 on the LOCAL array (host copy).  Solvers set up on the device (setup="device") hold no host
        mask: the device kernel (same formula, NaN inside the cylinder) is used instead."""
        if self.cylinder_mask is None:
            return self.compute_vorticity_device().cpu().numpy()
        cfg = self.config
        u, v = self.u, self.v
        w = np.zeros_like(u)
        w[1:-1, 1:-1] = (v[1:-1, 2:] - v[1:-1, :-2]) / (2 * cfg.dx) - (u[2:, 1:-1] - u[:-2, 1:-1]) / (2 * cfg.dy)
        w[self.cylinder_mask] = np.nan
        return w

    def gather_to_root(self, array_or_name, root=0):
        """Global (ny, nx) host array assembled on rank `root` from the OWNED columns of every slab (None on the other
        ranks).  `array_or_name`: a field name or a LOCAL (ny, nx_local) array / CUDA tensor (e.g. the device vorticity)."""
        a = self._get(array_or_name) if isinstance(array_or_name, str) else (
            array_or_name.cpu().numpy() if isinstance(array_or_name, torch.Tensor) else np.asarray(array_or_name))
        loc = np.ascontiguousarray(a[:, self.plan.own0:self.plan.own1])
        if self._nranks == 1:
            return loc
        import torch.distributed as dist
        parts = [None] * self._nranks if self._rank == root else None
        dist.gather_object(loc, parts, dst=root, group=self._group)
        return np.concatenate(parts, axis=1) if self._rank == root else None

    def compute_vorticity_device(self, out=None):
        """compute_vorticity (v5.py:365-373) on the device: a CUDA tensor (NaN inside the cylinder), nothing crosses PCIe."""
        u, v = self._d["u"], self._d["v"]
        w = torch.empty_like(u) if out is None else out
        check(self._lib.b200cfd_k_vorticity(self._h, u.data_ptr(), v.data_ptr(), self._d["mask"].data_ptr(), w.data_ptr(),
                                            _stream_ptr(self.device)), self._h)
        return w

    def _local_health(self):
        out = (C.c_double * 4)()
        check(self._lib.b200cfd_health(self._h, C.byref(out), _stream_ptr(self.device)), self._h)
        return list(out)

    def health(self):
        """Device reductions monitor_simulation_health needs (v5.py:599-613): no field leaves the GPU."""
        out = self._local_health()
        if self._nranks > 1:
            import torch.distributed as dist
            t = torch.tensor(list(out), dtype=torch.float64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self._group)
            out = t.tolist()
        return {"nonfinite": bool(out[0]), "max_abs_u": out[1], "max_abs_v": out[2], "max_div": out[3]}

    def profile_steps(self, n_steps):
        """Per-phase mean milliseconds per step over n_steps (CUDA events between the kernels; stream launches)."""
        check(self._lib.b200cfd_profile_begin(self._h), self._h)
        graph, self._use_graph = self._use_graph, False
        try:
            self.advance(n_steps)
        finally:
            self._use_graph = graph
        names = C.create_string_buffer(1024)
        ms = (C.c_double * 32)()
        n = self._lib.b200cfd_profile_end(self._h, names, 1024, ms, 32)
        if n < 0:
            check(n, self._h)
        return dict(zip(names.value.decode().split(";"), list(ms)[:n]))

    def profile_slab_steps(self, n_steps):
        """Where a slab-decomposed step's time goes (mean ms per step over n_steps, this rank, CUDA events on the compute
        stream, stream launches instead of the graph): the kernels of step_begin one by one, the all_gather of the
        reduction vector, the closing kernel, and pack / NCCL send-recv / unpack of the halo refresh."""
        import torch.distributed as dist
        if self._nranks == 1:
            return {"kernels": self.profile_steps(n_steps)}
        lib, h, sp = self._lib, self._h, lambda: _stream_ptr(self.device)
        u, v = self._d["u"], self._d["v"]
        ev = lambda: torch.cuda.Event(enable_timing=True)
        if self._overlap:
            return self._profile_overlapped(n_steps)
        check(lib.b200cfd_profile_begin(h), h)
        marks, halo_marks = [], []
        for _ in range(n_steps):
            e = [ev() for _ in range(5)]
            e[0].record()
            check(lib.b200cfd_step_begin(h, 0, sp()), h)
            e[1].record()
            dist.all_gather_into_tensor(self._gathered.view(-1), self._red, group=self._group)
            e[2].record()
            check(lib.b200cfd_step_end(h, self._gathered.data_ptr(), self._nranks, sp()), h)
            e[3].record()
            hm = []
            self._halo.exchange([u, v], marks=hm)
            e[4].record()
            marks.append(e); halo_marks.append(hm)
            self._step += 1
        names = C.create_string_buffer(1024)
        ms = (C.c_double * 32)()
        n = lib.b200cfd_profile_end(h, names, 1024, ms, 32)
        if n < 0:
            check(n, h)
        torch.cuda.synchronize(self.device)
        mean = lambda a, b: sum(m[a].elapsed_time(m[b]) for m in marks) / len(marks)
        out = {"kernels": dict(zip(names.value.decode().split(";"), [round(x, 5) for x in list(ms)[:n]])),
               "step_begin_total": round(mean(0, 1), 5), "all_gather": round(mean(1, 2), 5),
               "step_end": round(mean(2, 3), 5), "halo_exchange_total": round(mean(3, 4), 5)}
        if halo_marks and len(halo_marks[0]) == 4:
            hmean = lambda a, b: sum(m[a].elapsed_time(m[b]) for m in halo_marks) / len(halo_marks)
            out.update({"halo_pack": round(hmean(0, 1), 5), "halo_send_recv": round(hmean(1, 2), 5),
                        "halo_unpack": round(hmean(2, 3), 5)})
        return out

    def _profile_overlapped(self, n_steps):
        """The overlapped slab step (graph replay, as timed): CUDA events on both streams.  `exposed_*` are the parts of the
        step that are NOT hidden behind compute: what the compute stream spends in the all_gather + closing kernel, and how
        long it then still waits for the halos."""
        import torch.distributed as dist
        lib, h, sp = self._lib, self._h, lambda: _stream_ptr(self.device)
        u, v = self._d["u"], self._d["v"]
        main = torch.cuda.current_stream(self.device)
        ev = lambda: torch.cuda.Event(enable_timing=True)
        recs = []
        for _ in range(n_steps):
            e = [ev() for _ in range(6)]
            hm = []
            e[0].record(main)
            check(lib.b200cfd_step_part(h, 1, int(self._use_graph), sp()), h)
            e[1].record(main)
            with torch.cuda.stream(self._comm_stream):
                self._comm_stream.wait_event(e[1])
                self._halo.exchange([u, v], marks=hm)
                self._ev_halo.record(self._comm_stream)
            check(lib.b200cfd_step_part(h, 2, int(self._use_graph), sp()), h)
            e[2].record(main)
            dist.all_gather_into_tensor(self._gathered.view(-1), self._red, group=self._group)
            e[3].record(main)
            check(lib.b200cfd_step_end(h, self._gathered.data_ptr(), self._nranks, sp()), h)
            e[4].record(main)
            main.wait_event(self._ev_halo)
            e[5].record(main)
            recs.append((e, hm))
            self._step += 1
        torch.cuda.synchronize(self.device)
        mean = lambda f: round(sum(f(e, hm) for e, hm in recs) / len(recs), 5)
        out = {"overlap": True,
               "part1_to_edge_finalize": mean(lambda e, hm: e[0].elapsed_time(e[1])),
               "part2_rest_of_finalize": mean(lambda e, hm: e[1].elapsed_time(e[2])),
               "exposed_all_gather": mean(lambda e, hm: e[2].elapsed_time(e[3])),
               "exposed_step_end": mean(lambda e, hm: e[3].elapsed_time(e[4])),
               "exposed_wait_for_halos": mean(lambda e, hm: e[4].elapsed_time(e[5])),
               "step_total": mean(lambda e, hm: e[0].elapsed_time(e[5]))}
        if recs and len(recs[0][1]) == 4:
            out.update({"comm_stream_halo_pack": mean(lambda e, hm: hm[0].elapsed_time(hm[1])),
                        "comm_stream_halo_send_recv": mean(lambda e, hm: hm[1].elapsed_time(hm[2])),
                        "comm_stream_halo_unpack": mean(lambda e, hm: hm[2].elapsed_time(hm[3]))})
        # per-kernel times of the same step come from the unsplit stream-launch profile (b200cfd_profile_*)
        check(lib.b200cfd_profile_begin(h), h)
        for _ in range(min(n_steps, 3)):
            check(lib.b200cfd_step_begin(h, 0, sp()), h)
            dist.all_gather_into_tensor(self._gathered.view(-1), self._red, group=self._group)
            check(lib.b200cfd_step_end(h, self._gathered.data_ptr(), self._nranks, sp()), h)
            self._halo.exchange([u, v])
            self._step += 1
        names = C.create_string_buffer(1024)
        ms = (C.c_double * 32)()
        n = lib.b200cfd_profile_end(h, names, 1024, ms, 32)
        if n < 0:
            check(n, h)
        out["kernels"] = dict(zip(names.value.decode().split(";"), [round(x, 5) for x in list(ms)[:n]]))
        return out

    # ---- introspection ----
    def launches_per_step(self):
        return int(self._lib.b200cfd_launches_per_step(self._h))

    def launches_per_step_slab(self):
        """Kernels of ours per slab step: the step's own plus the halo pack / unpack launches (NCCL's kernels not counted)."""
        n = self.launches_per_step()
        return n + (self._halo.launches_per_exchange if self._halo is not None else 0)

    def collectives_per_step(self):
        return ("1 all_gather (128 B per rank) + 1 batched NCCL send/recv of the u, v halo columns" +
                (" on a communication stream, overlapped with the finalize kernel" if getattr(self, "_overlap", False) else ""))

    def clean_plan(self):
        out = (C.c_int * 8)()
        self._lib.b200cfd_get_clean_plan(self._h, C.byref(out))
        keys = ("strip_w", "tile_h", "apron_h", "apron_v", "n_strips", "n_tiles", "n_sm", "n_energy_partials")
        return dict(zip(keys, list(out)))


B200TurbulentSolver = OptimizedTurbulentSolver


def monitor_simulation_health(solver, step):
    """v5.py:599-613 on device reductions of the current u, v (finite check, max|u|,|v|, max|div|)."""
    cfg = solver.config
    hv = solver.health()
    if hv["nonfinite"]:
        return False
    if max(hv["max_abs_u"], hv["max_abs_v"]) > cfg.max_velocity:
        return False
    if hv["max_div"] > (20.0 if step <= 1000 else 2.0):
        return False
    return True
