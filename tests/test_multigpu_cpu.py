"""Host-side logic of the x-slab decomposition (cfd_simulations_b200/multigpu.py) on CPU: the slab plan, and the halo
exchange plumbing with the gloo backend at world_size 2 and 3 (pack/unpack are NumPy slices here; the product passes
CUDA kernels).  No compute of the library is called."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cfd_simulations_b200 import multigpu


def test_plan_covers_grid_without_overlap():
    for nxg, n in ((1000, 1), (1000, 3), (4096, 8), (1031, 4)):
        plans = [multigpu.make_plan(nxg, n, r, 80, 14) for r in range(n)]
        assert plans[0].x0 == 0 and plans[-1].x1 == nxg
        for a, b in zip(plans[:-1], plans[1:]):
            assert a.x1 == b.x0
            assert a.send_right == b.halo_left and b.send_left == a.halo_right
        assert plans[0].halo_left == 0 and plans[-1].halo_right == 0
        for p in plans:
            assert p.x_offset >= 0 and p.x_offset + p.nx_local <= nxg
            assert p.own1 - p.own0 == p.x1 - p.x0


def test_plan_rejects_slabs_narrower_than_halo():
    with pytest.raises(ValueError):
        multigpu.make_plan(300, 4, 1, 80, 14)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, ny, nxg, hl, hr, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = multigpu.make_plan(nxg, world, rank, hl, hr)
        glob = [np.arange(ny * nxg, dtype=np.float64).reshape(ny, nxg) * (k + 1) for k in range(2)]
        # local arrays: owned columns correct, halo columns poisoned
        loc = []
        for g in glob:
            a = torch.from_numpy(g[:, plan.x_offset:plan.x_offset + plan.nx_local].copy())
            a[:, :plan.own0] = -1.0
            a[:, plan.own1:] = -1.0
            loc.append(a)

        def pack(f, j0, n, o):
            o.copy_(f[:, j0:j0 + n])

        def unpack(f, j0, n, src):
            f[:, j0:j0 + n] = src

        ex = multigpu.HaloExchanger(plan, ny, 2, alloc=lambda shape: torch.empty(shape, dtype=torch.float64), pack=pack, unpack=unpack)
        ex.exchange(loc)
        ok = all(np.array_equal(a.numpy(), g[:, plan.x_offset:plan.x_offset + plan.nx_local]) for a, g in zip(loc, glob))
        # the reduction path: all_gather of per-rank vectors
        red = torch.full((16,), float(rank))
        gathered = torch.zeros(world * 16)
        dist.all_gather_into_tensor(gathered, red)
        ok = ok and gathered.view(world, 16)[:, 0].tolist() == [float(r) for r in range(world)]
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_halo_exchange_gloo(world):
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 7, 64, 9, 4, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(out[r] for r in range(world))


# ---- Euler x-slabs (shockwave.py): rows are the slow axis, a halo is one contiguous row ----
def test_euler_row_split_covers_rows():
    from cfd_simulations_b200.shockwave import row_split
    for nx, n in ((400, 1), (400, 3), (16384 * 8, 8), (1031, 4)):
        r = [row_split(nx, n, k) for k in range(n)]
        assert r[0][0] == 0 and r[-1][1] == nx
        assert all(a[1] == b[0] for a, b in zip(r[:-1], r[1:]))
        assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def _row_worker(rank, world, port, nx, ny, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cfd_simulations_b200.shockwave import row_split
        i0, i1 = row_split(nx, world, rank)
        hlo, hhi = int(rank > 0), int(rank < world - 1)
        glob = np.arange(nx * ny * 4, dtype=np.float32).reshape(nx, ny, 4)
        loc = torch.from_numpy(glob[i0 - hlo:i1 + hhi].copy())
        if hlo:
            loc[0] = -1.0
        if hhi:
            loc[-1] = -1.0
        nbytes = multigpu.exchange_halo_rows(loc, rank, world, hlo, hhi)
        ok = np.array_equal(loc.numpy(), glob[i0 - hlo:i1 + hhi]) and nbytes == 2 * ny * 16 * (hlo + hhi)
        # the per-step reduction: all_reduce(MAX) of the 4-double vector
        red = torch.tensor([1.0 + rank, 5.0 - rank, float(rank == 1), 0.0], dtype=torch.float64)
        dist.all_reduce(red, op=dist.ReduceOp.MAX)
        ok = ok and red.tolist() == [float(world), 5.0, 1.0, 0.0]
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_euler_halo_rows_gloo(world):
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_row_worker, args=(r, world, port, 23, 11, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert all(out[r] for r in range(world))
