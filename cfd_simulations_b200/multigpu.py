"""x-slab decomposition of the projection time step over the GPUs of one node (one process per GPU).

The reference is single-process; this is the B200-native scaling of its time step.

Design: communication-avoiding slabs.  Rank r owns global columns [x0, x1) and stores [x0 - HL, x1 + HR): HL / HR halo
columns on interior sides.  Every kernel of the step runs on the whole local array, so halo cells are recomputed
redundantly and their region of validity shrinks by the stencil radius of each phase; HL, HR (b200cfd_required_halo) are
chosen so that after a full step all OWNED cells are exact.  Per time step there is therefore exactly
  * one all_gather of a 16-double reduction vector (max|div|, max|grad phi|, max|vorticity|, energy sum, max|u|,|v| for the
    adaptive dt, ...) between step_begin and step_end, and
  * one halo refresh of u and v (pack columns -> NCCL send/recv over NVLink -> unpack),
and no mid-step exchange.  The left halo is wide (2 x the upstream apron of the lexicographic clean-divergence sweep,
about 80 columns) because that sweep's dependence runs left-to-right; the right halo is 14 columns.  At 16384 columns
per slab the redundant work is 0.6 %.

SlabPlan / HaloExchanger are pure host logic (tested on CPU with gloo, world_size 2); the pack/unpack callables the
product passes in are CUDA kernels behind the C ABI.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist


@dataclass(frozen=True)
class SlabPlan:
    """Column ranges of one slab.  All numbers are columns; g = global index, l = local index."""
    rank: int
    nranks: int
    nx_global: int
    x0: int            # first owned global column
    x1: int            # one past the last owned global column
    halo_left: int     # halo columns stored left of x0 (0 on rank 0)
    halo_right: int    # halo columns stored right of x1 (0 on the last rank)
    send_left: int     # columns the LEFT neighbour needs from me  (= its halo_right)
    send_right: int    # columns the RIGHT neighbour needs from me (= its halo_left)

    @property
    def x_offset(self):
        return self.x0 - self.halo_left

    @property
    def nx_local(self):
        return (self.x1 - self.x0) + self.halo_left + self.halo_right

    @property
    def own0(self):
        return self.halo_left

    @property
    def own1(self):
        return self.halo_left + (self.x1 - self.x0)


def make_plan(nx_global, nranks, rank, halo_left, halo_right):
    """Even split of nx_global columns (the first nx_global % nranks slabs get one more)."""
    if not (0 <= rank < nranks):
        raise ValueError(f"rank {rank} outside [0, {nranks})")
    base, rem = divmod(nx_global, nranks)
    x0 = rank * base + min(rank, rem)
    x1 = x0 + base + (1 if rank < rem else 0)
    if nranks > 1 and base < max(halo_left, halo_right):
        raise ValueError(f"slabs of {base} columns are narrower than the halo ({halo_left}/{halo_right}): use fewer ranks")
    hl = halo_left if rank > 0 else 0
    hr = halo_right if rank < nranks - 1 else 0
    return SlabPlan(rank, nranks, nx_global, x0, x1, hl, hr,
                    send_left=halo_right if rank > 0 else 0, send_right=halo_left if rank < nranks - 1 else 0)


class HaloExchanger:
    """Refreshes the halo columns of a list of fields with one batched send/recv per neighbour.

    pack(field, j0, ncols, out2d) / unpack(field, j0, ncols, in2d) move field[:, j0:j0+ncols] to / from a contiguous
    (ny, ncols) buffer; buffers are allocated once with `alloc(shape)` on the device the backend communicates from."""

    def __init__(self, plan: SlabPlan, ny, nfields, alloc, pack, unpack, group=None, halo_rows=None):
        """halo_rows(pack: bool, fields, j_left, n_left, buf_left, j_right, n_right, buf_right): optional fused kernel that
        moves all fields and both sides in one launch (the product passes b200cfd_halo_rows); without it pack / unpack are
        called once per field and side."""
        self.plan, self.ny, self.nf, self.pack, self.unpack, self.group = plan, ny, nfields, pack, unpack, group
        self.halo_rows = halo_rows
        p = plan
        mk = lambda n: alloc((nfields, ny, n)) if n > 0 else None
        self.sbuf_l, self.sbuf_r = mk(p.send_left), mk(p.send_right)
        self.rbuf_l, self.rbuf_r = mk(p.halo_left), mk(p.halo_right)
        self.bytes_per_exchange = sum(b.numel() * b.element_size() for b in (self.sbuf_l, self.sbuf_r) if b is not None)
        # pack launches for the sides that send + unpack launches for the sides that receive, per field
        self.launches_per_exchange = (2 if halo_rows is not None else
                                      nfields * sum(b is not None for b in (self.sbuf_l, self.sbuf_r, self.rbuf_l, self.rbuf_r)))

    def exchange(self, fields, marks=None):
        """marks: optional list that receives 4 CUDA events (before pack, after pack, after send/recv, after unpack)."""
        p = self.plan
        assert len(fields) == self.nf
        ops = []
        def mark():
            if marks is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                marks.append(e)
        mark()
        # post receives first, then pack and send
        if self.rbuf_l is not None:
            ops.append(dist.P2POp(dist.irecv, self.rbuf_l, p.rank - 1, self.group))
        if self.rbuf_r is not None:
            ops.append(dist.P2POp(dist.irecv, self.rbuf_r, p.rank + 1, self.group))
        if self.halo_rows is not None:
            self.halo_rows(True, fields, p.own0, p.send_left, self.sbuf_l, p.own1 - p.send_right, p.send_right, self.sbuf_r)
        else:
            for k, f in enumerate(fields):
                if self.sbuf_l is not None:
                    self.pack(f, p.own0, p.send_left, self.sbuf_l[k])
                if self.sbuf_r is not None:
                    self.pack(f, p.own1 - p.send_right, p.send_right, self.sbuf_r[k])
        if self.sbuf_l is not None:
            ops.append(dist.P2POp(dist.isend, self.sbuf_l, p.rank - 1, self.group))
        if self.sbuf_r is not None:
            ops.append(dist.P2POp(dist.isend, self.sbuf_r, p.rank + 1, self.group))
        if not ops:
            return
        mark()
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        mark()
        if self.halo_rows is not None:
            self.halo_rows(False, fields, 0, p.halo_left, self.rbuf_l, p.own1, p.halo_right, self.rbuf_r)
        else:
            for k, f in enumerate(fields):
                if self.rbuf_l is not None:
                    self.unpack(f, 0, p.halo_left, self.rbuf_l[k])
                if self.rbuf_r is not None:
                    self.unpack(f, p.own1, p.halo_right, self.rbuf_r[k])
        mark()


def exchange_halo_rows(buf, rank, nranks, halo_lo, halo_hi, group=None):
    """Euler x-slabs (shockwave.py): x is the SLOW axis of the (rows, ny, 4) state, so a halo is one contiguous row.
    Row 0 / row -1 of `buf` are the halo rows (when present); they are refreshed from the neighbour's outermost owned
    row (its row 1 / row -2 ... i.e. index halo_lo / -1 - halo_hi) with one batched send/recv per neighbour."""
    ops = []
    n = buf.shape[0]
    if halo_lo:
        ops.append(dist.P2POp(dist.irecv, buf[0], rank - 1, group))
        ops.append(dist.P2POp(dist.isend, buf[halo_lo], rank - 1, group))
    if halo_hi:
        ops.append(dist.P2POp(dist.irecv, buf[n - 1], rank + 1, group))
        ops.append(dist.P2POp(dist.isend, buf[n - 1 - halo_hi], rank + 1, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return 2 * buf[0].numel() * buf.element_size() * (int(bool(halo_lo)) + int(bool(halo_hi)))


class _DevPtr:
    """Expose a raw device pointer to torch (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr, n, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def device_view(ptr, n, device):
    return torch.as_tensor(_DevPtr(ptr, n), device=device)


class LocalSlabs:
    """All slabs of a decomposition stepped in lockstep inside ONE process on ONE GPU (no NCCL): the all_gather is a
    device copy and the halo refresh packs from one slab and unpacks into its neighbour.  Same kernels, same halo
    widths, same reduction path as the multi-process run — it makes the decomposition numerics testable on a
    single-GPU box (SURVEY §4 "virtual slabs") and is what tests/test_gpu_parity.py uses."""

    def __init__(self, config, nslabs, device=0, **kw):
        from .solver import OptimizedTurbulentSolver
        self.slabs = [OptimizedTurbulentSolver(config, device=device, nranks=nslabs, rank=r, **kw) for r in range(nslabs)]
        self.config = config

    def advance(self, n_steps):
        from .solver import _stream_ptr
        from ._lib import check
        S = self.slabs
        for _ in range(n_steps):
            for s in S:
                check(s._lib.b200cfd_step_begin(s._h, int(s._use_graph), _stream_ptr(s.device)), s._h)
            g = torch.stack([s._red for s in S])
            for s in S:
                s._gathered.copy_(g)
                check(s._lib.b200cfd_step_end(s._h, s._gathered.data_ptr(), len(S), _stream_ptr(s.device)), s._h)
                s._step += 1
            self.refresh_halos()

    def refresh_halos(self, names=("u", "v")):
        S = self.slabs
        if tuple(names) == ("u", "v") and all(s._halo is not None and s._halo.halo_rows is not None for s in S):
            # the product's path: all fields and both sides in one launch per slab (b200cfd_halo_rows), the "message" being
            # the neighbour's send buffer
            for s in S:
                p, hx = s.plan, s._halo
                hx.halo_rows(True, [s._d[n] for n in names], p.own0, p.send_left, hx.sbuf_l, p.own1 - p.send_right, p.send_right, hx.sbuf_r)
            for i, s in enumerate(S):
                p, hx = s.plan, s._halo
                left = S[i - 1]._halo.sbuf_r if i > 0 else None
                right = S[i + 1]._halo.sbuf_l if i + 1 < len(S) else None
                hx.halo_rows(False, [s._d[n] for n in names], 0, p.halo_left, left, p.own1, p.halo_right, right)
            return
        for a, b in zip(S[:-1], S[1:]):          # a is the left neighbour of b
            pa, pb = a.plan, b.plan
            for k, name in enumerate(names):
                fa, fb = a._d[name], b._d[name]
                a._halo.pack(fa, pa.own1 - pa.send_right, pa.send_right, a._halo.sbuf_r[k])
                b._halo.unpack(fb, 0, pb.halo_left, a._halo.sbuf_r[k])
                b._halo.pack(fb, pb.own0, pb.send_left, b._halo.sbuf_l[k])
                a._halo.unpack(fa, pa.own1, pa.halo_right, b._halo.sbuf_l[k])

    def solve_pressure(self, div_global, max_iter, tol, omega=1.0, mode="as_written"):
        """The slab-decomposed Poisson solve (solver.solve_pressure_slabs) with the collectives emulated:
        returns (global phi, iterations, last change)."""
        S = self.slabs
        rhs = []
        for s in S:
            p = s.plan
            rhs.append(torch.as_tensor(np.ascontiguousarray(div_global[:, p.x_offset:p.x_offset + p.nx_local]),
                                       dtype=s._d["phi"].dtype).to(s.device))
        if mode == "jacobi":
            it, last = slab_jacobi(S, rhs, int(max_iter), exchange=lambda: self.refresh_halos(names=("phi",))), 0.0
        else:
            it, last = slab_poisson(S, rhs, int(max_iter), float(tol), float(omega),
                                    exchange=lambda: self.refresh_halos(names=("phi",)),
                                    reduce_max=lambda parts: torch.stack(parts).max(dim=0).values)
        return self.gather("phi"), it, last

    def set_step(self, step):
        S = self.slabs
        hv = [s._local_health() for s in S]
        mu, mv = max(h[1] for h in hv), max(h[2] for h in hv)
        from .solver import _stream_ptr
        from ._lib import check
        for s in S:
            s._sync_scalars()
            check(s._lib.b200cfd_set_step_global(s._h, int(step), mu, mv, _stream_ptr(s.device)), s._h)
            s._step = s._fetched = s._first_step = int(step)

    def set_global(self, name, array):
        for s in self.slabs:
            p = s.plan
            s._d[name].copy_(torch.as_tensor(np.ascontiguousarray(array[:, p.x_offset:p.x_offset + p.nx_local]),
                                             dtype=s._d[name].dtype))

    def gather(self, name):
        return np.concatenate([s._get(name)[:, s.plan.own0:s.plan.own1] for s in self.slabs], axis=1)

    @property
    def scalars(self):
        self.slabs[0]._sync_scalars()
        return self.slabs[0].scalars

    def close(self):
        for s in self.slabs:
            s.close()


def slab_jacobi(slabs, rhs, sweeps, exchange):
    """The Jacobi branch of solve_pressure_fast (v5.py:336-346: a fixed number of sweeps, no tolerance) on slabs: the
    region of validity of the halos shrinks by ONE column per sweep, so they are refreshed once per min(halo) sweeps;
    the sweeps in between run through the (temporally blocked) Jacobi kernel on the local array."""
    hl, hr = slabs[0]._halo_req
    T = max(1, min(hl, hr)) if slabs[0]._nranks > 1 else 64
    for s in slabs:
        s._d["phi"].zero_()
    it = 0
    while it < sweeps:
        m = min(T, sweeps - it)
        for i, s in enumerate(slabs):
            s._jacobi_sweeps(rhs[i], m)
        it += m
        exchange()
    return it


def slab_poisson(slabs, rhs, max_iter, tol, omega, exchange, reduce_max):
    """Control flow of the slab-decomposed red-black Poisson solve, shared by a torchrun rank (`slabs` = [its solver])
    and LocalSlabs (`slabs` = all of them).  exchange(): refresh the phi halos of every slab.  reduce_max(list of one
    1-D device tensor per local slab) -> tensor of the GLOBAL per-iteration maxima.  Returns (iterations done, last
    max|dphi|) with the reference's meaning (v5.py:212-225): the iteration whose change falls below `tol` is the last."""
    hl, hr = slabs[0]._halo_req
    nr = slabs[0]._nranks
    T = max(1, min(hl, hr) // 2) if nr > 1 else 64          # iterations per halo refresh (validity shrinks 2 columns each)
    if not (tol > 0.0) and T >= 3:
        T -= T % 3                                           # whole passes of the streaming kernel (3 iterations each)
    views = [device_view(s._lib.b200cfd_poisson_change_ptr(s._h), 1, s.device) for s in slabs]
    for s in slabs:
        s._d["phi"].zero_()                                   # v5.py:331
    check_tol = tol > 0.0
    backups = [torch.empty_like(s._d["phi"]) for s in slabs] if check_tol else None
    chg = [torch.zeros(T, dtype=torch.float64, device=s.device) for s in slabs]

    def run(n, record, final=False):
        if not record and not check_tol:
            # fixed count: the chunk in one temporally blocked pass; the very last iteration of the solve is run by the
            # plain kernels so that its max|dphi| is on the device for the report
            nb = n - 1 if final else n
            for i, s in enumerate(slabs):
                if nb >= 2:
                    s._poisson_iterations(rhs[i], nb, omega, temporal_depth=0)
                elif nb == 1:
                    s._poisson_iterations(rhs[i], 1, omega)
                if final:
                    s._poisson_iterations(rhs[i], 1, omega)
            return
        for k in range(n):
            for i, s in enumerate(slabs):
                s._poisson_iterations(rhs[i], 1, omega)
                if record:
                    chg[i][k:k + 1].copy_(views[i])
    it, last = 0, 0.0
    for s in slabs:                                           # the right-hand side does not change between the chunks
        s._lib.b200cfd_poisson_keep_rhs(s._h, 0)
    first = True
    try:
        while it < max_iter:
            m = min(T, max_iter - it)
            if not first:
                for s in slabs:
                    s._lib.b200cfd_poisson_keep_rhs(s._h, 1)
            first = False
            if check_tol:
                for s, b in zip(slabs, backups):
                    b.copy_(s._d["phi"])
            run(m, check_tol, final=(it + m >= max_iter))
            g = None
            stop = None
            if check_tol:
                g = reduce_max([c[:m] for c in chg]).tolist()     # one collective + one host read per chunk
                last = g[m - 1]
            elif it + m >= max_iter:                              # fixed count: only the final change is reported
                last = float(reduce_max([v.clone() for v in views]).tolist()[0])
            if check_tol:
                for k, c in enumerate(g):
                    if c < tol:
                        stop = k
                        break
            if stop is not None:
                if stop + 1 < m:                                  # overshot: replay the chunk up to the stopping iteration
                    for s, b in zip(slabs, backups):
                        s._d["phi"].copy_(b)
                    run(stop + 1, False)
                it += stop + 1
                last = g[stop]
                exchange()
                break
            it += m
            exchange()
    finally:
        for s in slabs:
            s._lib.b200cfd_poisson_keep_rhs(s._h, 0)
    return it, last
