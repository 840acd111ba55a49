"""Host-side mirror of the reference's ShockwaveSolver and its module-level functions, backed by libb200cfd.so
(CUDA, sm_100a; include/b200cfd_euler.h).

Reference: python/shockwave_analysis/v1_shock.py:69-328.  Names and argument meaning are the reference's.  The state
lives in HBM as (nx, ny, 4) float32 (two torch CUDA tensors own the ping-pong buffers; torch does nothing else) and
`time_step()` returns the reference's dt.  `advance(n)` runs n steps without any host round trip.
nranks > 1: x-slabs, one process per GPU (rows are contiguous, so halos are plain row copies; NCCL send/recv +
one all_reduce(MAX) of 4 doubles per step).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import B200CFDError, check_euler
from .shockwave_setup import ShockwaveConfig, free_stream, grid, initial_state, n_reflect, wedge_mask  # noqa: F401

HLLC, ROE = 0, 1


def _sp(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def row_split(nx, nranks, rank):
    """Owned global rows [i0, i1) of slab `rank` (even split; the first nx % nranks slabs get one more)."""
    base, rem = divmod(nx, nranks)
    i0 = rank * base + min(rank, rem)
    return i0, i0 + base + (1 if rank < rem else 0)


class ShockwaveSolver:
    """Drop-in for the reference class (v1_shock.py:225).  exact_division=True puts an IEEE division wherever the
    reference divides; the default uses one reciprocal per cell (results within 1 ulp of f64 before the f32 store)."""

    def __init__(self, config: ShockwaveConfig, device=0, use_graph=True, exact_division=False, nranks=1, rank=0,
                 group=None, U0=None):
        if not torch.cuda.is_available():
            raise B200CFDError("no CUDA device: cfd_simulations_b200 has no CPU fallback")
        self.config = cfg = config
        self.device = dev = torch.device("cuda", device)
        self._lib = _lib.load()
        self._use_graph = bool(use_graph)
        self._nranks, self._rank, self._group = int(nranks), int(rank), group
        # setup_grid / setup_wedge / initialize_fields (v1_shock.py:233-261), vectorised
        self.x, self.y = grid(cfg)
        self._i0, self._i1 = row_split(cfg.nx, self._nranks, self._rank)
        self._hlo = 1 if self._rank > 0 else 0
        self._hhi = 1 if self._rank < self._nranks - 1 else 0
        lo, hi = self._i0 - self._hlo, self._i1 + self._hhi
        self._lo, self._hi = lo, hi
        self.wedge_mask = wedge_mask(cfg, self.x, self.y, lo, hi)
        if U0 is None:
            U_host = initial_state(cfg, hi - lo)
        else:
            U0 = np.asarray(U0)
            if U0.shape != (cfg.nx, cfg.ny, 4):
                raise ValueError(f"U0 must have shape {(cfg.nx, cfg.ny, 4)}")
            U_host = np.ascontiguousarray(U0[lo:hi], dtype=np.float32)
        self._buf = [torch.from_numpy(U_host).to(dev), torch.zeros((hi - lo, cfg.ny, 4), dtype=torch.float32, device=dev)]
        self._wedge = torch.from_numpy(np.ascontiguousarray(self.wedge_mask, dtype=np.uint8)).to(dev)
        self._ccfg = _lib.EulerConfig(
            nx=hi - lo, ny=cfg.ny, solver_type=HLLC if cfg.solver_type == "hllc" else ROE,
            n_reflect=n_reflect(cfg, self.x, self.y), rank=self._rank, nranks=self._nranks, halo_lo=self._hlo,
            halo_hi=self._hhi, exact_division=int(bool(exact_division)), reserved=0, gamma=cfg.gamma,
            epsilon=cfg.epsilon, max_val=cfg.max_val, cfl=cfg.cfl, dx=cfg.dx, dy=cfg.dy, rho_inf=cfg.rho_inf,
            u_inf=float(cfg.u_inf), p_inf=cfg.p_inf)
        self._h = C.c_void_p()
        check_euler(self._lib.b200cfd_euler_create(C.byref(self._ccfg), device, C.byref(self._h)))
        self._cur = 0
        self._calls = 0
        self._fetched = 0
        self._dts = []
        self.step = 0
        self.times = []
        self._red = None
        if self._nranks > 1:
            from .multigpu import device_view
            self._red = device_view(self._lib.b200cfd_euler_reduction_ptr(self._h), 4, dev)
        self._bind(0)

    def _init_local(self, config, device, nranks, rank, U0, device_view, **kw):
        """LocalShockSlabs: the constructor without its collective (the caller reduces over the slabs itself)."""
        self._collectives = False
        ShockwaveSolver.__init__(self, config, device=device, nranks=nranks, rank=rank, U0=U0, **kw)

    # ---- plumbing -------------------------------------------------------------------------------------------------
    def _bind(self, step):
        a, b = self._buf[self._cur], self._buf[self._cur ^ 1]
        check_euler(self._lib.b200cfd_euler_bind(self._h, a.data_ptr(), b.data_ptr(), self._wedge.data_ptr(), int(step),
                                                 _sp(self.device)), self._h)
        self._buf = [a, b]
        self._cur = 0
        if self._nranks > 1 and getattr(self, "_collectives", True):
            import torch.distributed as dist
            dist.all_reduce(self._red, op=dist.ReduceOp.MAX, group=self._group)
            check_euler(self._lib.b200cfd_euler_set_dt_from(self._h, self._red.data_ptr(), int(step), _sp(self.device)), self._h)
        self._calls = self._fetched = int(step)

    def _exchange_halos(self, buf):
        from .multigpu import exchange_halo_rows
        exchange_halo_rows(buf, self._rank, self._nranks, self._hlo, self._hhi, self._group)

    def advance(self, n_steps: int):
        """n x time_step() with the state and dt resident in HBM (no host round trip)."""
        n_steps = int(n_steps)
        if self._nranks == 1:
            check_euler(self._lib.b200cfd_euler_step(self._h, n_steps, int(self._use_graph), _sp(self.device)), self._h)
            self._cur ^= n_steps & 1
        else:
            import torch.distributed as dist
            for _ in range(n_steps):
                check_euler(self._lib.b200cfd_euler_step_begin(self._h, _sp(self.device)), self._h)
                dist.all_reduce(self._red, op=dist.ReduceOp.MAX, group=self._group)
                self._cur ^= 1
                self._exchange_halos(self._buf[self._cur])
                check_euler(self._lib.b200cfd_euler_step_end(self._h, self._red.data_ptr(), _sp(self.device)), self._h)
        self._calls += n_steps

    def _sync_dts(self):
        n = self._calls - self._fetched
        if n <= 0:
            return
        out = (C.c_double * n)()
        check_euler(self._lib.b200cfd_euler_fetch_dts(self._h, self._fetched, n, out, _sp(self.device)), self._h)
        for dt in out:
            self._dts.append(float(dt))
            if dt != 0.0:                       # v1_shock.py:315-316 (a skipped step appends nothing)
                self.times.append(self.step * float(dt))
                self.step += 1
        self._fetched = self._calls

    # ---- the reference's method surface ---------------------------------------------------------------------------
    def time_step(self) -> float:
        """v1_shock.py:291-317.  Returns dt (0.0 if dt was not finite and the step was skipped)."""
        self.advance(1)
        self._sync_dts()
        return self._dts[-1]

    def compute_dt(self) -> float:
        """v1_shock.py:263-275 of the current state (local slab only when nranks > 1)."""
        out = (C.c_double * 3)()
        check_euler(self._lib.b200cfd_euler_k_compute_dt(self._h, self._buf[self._cur].data_ptr(), out, _sp(self.device)), self._h)
        return float(out[0])

    def apply_boundary_conditions(self):
        """v1_shock.py:277-289 (without the wedge zeroing) on the current state, in place."""
        check_euler(self._lib.b200cfd_euler_k_apply_bc(self._h, self._buf[self._cur].data_ptr(), None, _sp(self.device)), self._h)

    def check_health(self) -> bool:
        """v1_shock.py:319-328: no non-finite value, no negative density."""
        U = self._buf[self._cur]
        own = U[self._hlo:U.shape[0] - self._hhi]
        return bool(torch.isfinite(own).all().item()) and not bool((own[..., 0] < 0).any().item())

    @property
    def dts(self):
        self._sync_dts()
        return list(self._dts)

    @property
    def U(self) -> np.ndarray:
        """Host copy of this process's OWNED rows of the state, (rows, ny, 4) float32."""
        U = self._buf[self._cur]
        return U[self._hlo:U.shape[0] - self._hhi].cpu().numpy()

    @U.setter
    def U(self, value):
        value = np.ascontiguousarray(value, dtype=np.float32)
        if value.shape != (self.config.nx, self.config.ny, 4):
            raise ValueError(f"U must have shape {(self.config.nx, self.config.ny, 4)}")
        self._sync_dts()
        self._buf[self._cur].copy_(torch.from_numpy(value[self._lo:self._hi]))
        self._bind(self._calls)

    def device_tensor(self) -> torch.Tensor:
        return self._buf[self._cur]

    def time_step_host(self, U_host: torch.Tensor, n_steps=1) -> float:
        """The same step through a HOST array (pinned or pageable), as the reference's time_step() sees self.U:
        upload, n steps, download.  Single GPU.  Returns the last dt."""
        dt = C.c_double(0.0)
        a = self._buf[self._cur]
        assert U_host.dtype == torch.float32 and U_host.numel() == a.numel()
        check_euler(self._lib.b200cfd_euler_step_host(self._h, C.c_void_p(U_host.data_ptr()), int(n_steps), int(self._use_graph),
                                                      C.byref(dt), _sp(self.device)), self._h)
        self._cur ^= int(n_steps) & 1
        self._calls += int(n_steps)
        return dt.value

    def launches_per_step(self) -> int:
        return int(self._lib.b200cfd_euler_launches_per_step(self._h))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.b200cfd_euler_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class LocalShockSlabs:
    """All x-slabs of a decomposition stepped in lockstep inside ONE process on ONE GPU (no NCCL): the all_reduce is a
    torch.maximum over the slabs' reduction vectors and the halo refresh is a row copy.  Same kernels, same halo rows,
    same reduction path as the multi-process run: it makes the decomposition testable on a single-GPU box."""

    def __init__(self, config, nslabs, device=0, U0=None, **kw):
        from .multigpu import device_view
        self.config = config
        self.slabs = []
        for r in range(nslabs):
            s = ShockwaveSolver.__new__(ShockwaveSolver)
            s._local_emulation = True
            self.slabs.append(s)
        # construct each slab without collectives: nranks > 1 but the reductions are done here
        for r, s in enumerate(self.slabs):
            s._init_local(config, device, nslabs, r, U0, device_view, **kw)
        self._reduce_and(lambda s, red: check_euler(
            s._lib.b200cfd_euler_set_dt_from(s._h, red.data_ptr(), 0, _sp(s.device)), s._h))

    def _reduce_and(self, fn):
        red = torch.stack([s._red for s in self.slabs]).max(dim=0).values
        for s in self.slabs:
            fn(s, red)

    def advance(self, n_steps):
        S = self.slabs
        for _ in range(int(n_steps)):
            for s in S:
                check_euler(s._lib.b200cfd_euler_step_begin(s._h, _sp(s.device)), s._h)
                s._cur ^= 1
            for a, b in zip(S[:-1], S[1:]):          # a holds the rows just below b's
                ua, ub = a._buf[a._cur], b._buf[b._cur]
                ub[0].copy_(ua[ua.shape[0] - 2])
                ua[ua.shape[0] - 1].copy_(ub[1])
            self._reduce_and(lambda s, red: check_euler(
                s._lib.b200cfd_euler_step_end(s._h, red.data_ptr(), _sp(s.device)), s._h))
            for s in S:
                s._calls += 1

    def gather(self):
        return np.concatenate([s.U for s in self.slabs], axis=0)

    @property
    def dts(self):
        return self.slabs[0].dts

    def close(self):
        for s in self.slabs:
            s.close()


# ---- module-level functions of the reference (v1_shock.py:84-223) on NumPy arrays, computed on the GPU ---------------
def _dev(a, device):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(torch.device("cuda", device))


def _riemann(kind, U_L, U_R, gamma, normal_x, normal_y, epsilon, max_val, device=0):
    lib = _lib.load()
    UL, UR = _dev(np.atleast_2d(U_L), device), _dev(np.atleast_2d(U_R), device)
    out = torch.empty_like(UL)
    check_euler(lib.b200cfd_euler_k_riemann(UL.data_ptr(), UR.data_ptr(), UL.shape[0], kind, int(normal_x == 1.0), gamma,
                                            epsilon, max_val, out.data_ptr(), _sp(UL.device)))
    res = out.cpu().numpy()
    return res[0] if np.ndim(U_L) == 1 else res


def hllc_solver(U_L, U_R, gamma, normal_x, normal_y, epsilon, max_val, device=0):
    """v1_shock.py:147-209; U_L, U_R may be (4,) or (n, 4)."""
    return _riemann(HLLC, U_L, U_R, gamma, normal_x, normal_y, epsilon, max_val, device)


def roe_solver(U_L, U_R, gamma, normal_x, normal_y, epsilon, max_val, device=0):
    """v1_shock.py:97-145."""
    return _riemann(ROE, U_L, U_R, gamma, normal_x, normal_y, epsilon, max_val, device)


def compute_fluxes(U, gamma, epsilon, max_val, device=0):
    """v1_shock.py:84-95; U may be (4,) or (n, 4)."""
    lib = _lib.load()
    Ud = _dev(np.atleast_2d(U), device)
    F, G = torch.empty_like(Ud), torch.empty_like(Ud)
    check_euler(lib.b200cfd_euler_k_fluxes(Ud.data_ptr(), Ud.shape[0], gamma, epsilon, max_val, F.data_ptr(), G.data_ptr(), _sp(Ud.device)))
    F, G = F.cpu().numpy(), G.cpu().numpy()
    return (F[0], G[0]) if np.ndim(U) == 1 else (F, G)


def update_state(U, F, G, dx, dy, dt, epsilon, max_val, gamma=1.4, device=0):
    """v1_shock.py:211-223."""
    lib = _lib.load()
    nx, ny, _ = U.shape
    cc = _lib.EulerConfig(nx=nx, ny=ny, solver_type=HLLC, n_reflect=0, rank=0, nranks=1, halo_lo=0, halo_hi=0, exact_division=1,
                          reserved=0, gamma=gamma, epsilon=epsilon, max_val=max_val, cfl=0.99, dx=dx, dy=dy, rho_inf=1.0,
                          u_inf=1.0, p_inf=1.0)
    h = C.c_void_p()
    check_euler(lib.b200cfd_euler_create(C.byref(cc), device, C.byref(h)))
    try:
        Ud, Fd, Gd = _dev(U, device), _dev(F, device), _dev(G, device)
        out = torch.empty_like(Ud)
        check_euler(lib.b200cfd_euler_k_update_state(h, Ud.data_ptr(), Fd.data_ptr(), Gd.data_ptr(), float(dt), out.data_ptr(),
                                                     _sp(Ud.device)), h)
        return out.cpu().numpy()
    finally:
        lib.b200cfd_euler_destroy(h)
