"""x-slab decomposition on the GPU: the decomposed step must reproduce the single-slab step.
 * LocalSlabs: all slabs in one process on one GPU (device copies instead of NCCL) — runs on a 1-GPU box;
 * real NCCL run with one process per GPU when >= 2 GPUs are visible."""
import os
import socket

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu


def _cfg(nx, ny, f32=False, **kw):
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    return OptimizedTurbulentConfig(nx=nx, ny=ny, memory_efficient=f32, **kw)


@pytest.mark.parametrize("nslabs,f32", [(2, False), (3, False), (2, True)])
def test_local_slabs_match_single_slab(nslabs, f32):
    from cfd_simulations_b200.multigpu import LocalSlabs
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    cfg = _cfg(1200, 160, f32)
    one = OptimizedTurbulentSolver(cfg)
    many = LocalSlabs(cfg, nslabs)
    tol = 1e-13 if not f32 else 1e-6
    for start, n in ((0, 12), (995, 10)):
        if start:
            u, v = one.u, one.v
            one.step = start
            many.set_global("u", u); many.set_global("v", v); many.set_step(start)
        one.advance(n); many.advance(n)
        for name in ("u", "v", "phi", "u_star", "v_star"):
            assert rel_l2(many.gather(name), getattr(one, name)) <= tol * (10 if name == "phi" and f32 else 1), (name, start)
        one._sync_scalars()
        a, b = many.scalars, one.scalars
        for k in range(1, n + 1):
            for key in ("dt", "max_div_pre", "max_grad_phi", "max_div_post", "max_vorticity", "mean_energy", "max_abs_u",
                        "max_abs_v", "poisson_change", "step"):
                assert abs(a[-k][key] - b[-k][key]) <= (1e-12 if not f32 else 1e-5) * max(abs(b[-k][key]), 1e-30), (key, k)
    one.close(); many.close()


def test_pack_unpack_columns_roundtrip():
    import ctypes as C
    import torch
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver, _stream_ptr
    from cfd_simulations_b200._lib import check
    s = OptimizedTurbulentSolver(_cfg(300, 70))
    f = torch.arange(70 * 300, dtype=torch.float64, device="cuda").reshape(70, 300)
    buf = torch.empty((70, 17), dtype=torch.float64, device="cuda")
    check(s._lib.b200cfd_pack_columns(s._h, f.data_ptr(), 5, 17, buf.data_ptr(), _stream_ptr(s.device)), s._h)
    assert torch.equal(buf, f[:, 5:22])
    g = torch.zeros_like(f)
    check(s._lib.b200cfd_unpack_columns(s._h, g.data_ptr(), 280, 17, buf.data_ptr(), _stream_ptr(s.device)), s._h)
    assert torch.equal(g[:, 280:297], f[:, 5:22]) and float(g[:, :280].abs().sum()) == 0.0
    assert s._lib.b200cfd_pack_columns(s._h, f.data_ptr(), 290, 17, buf.data_ptr(), _stream_ptr(s.device)) != 0
    s.close()


def _nccl_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from cfd_simulations_b200.solver import OptimizedTurbulentSolver
        cfg = _cfg(1600, 128)
        s = OptimizedTurbulentSolver(cfg, device=rank, nranks=world, rank=rank)
        s.advance(15)
        res = {k: s.gather(k) for k in ("u", "v", "phi")}
        s._sync_scalars()
        if rank == 0:
            out["fields"] = res
            out["scalars"] = s.scalars[-1]
        s.close()
    finally:
        dist.destroy_process_group()


def test_nccl_slabs_match_single_gpu():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    world = min(n, 4)
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    sk = socket.socket(); sk.bind(("127.0.0.1", 0)); port = sk.getsockname()[1]; sk.close()
    procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    one = OptimizedTurbulentSolver(_cfg(1600, 128))
    one.advance(15)
    for k, a in out["fields"].items():
        assert rel_l2(a, getattr(one, k)) <= 1e-13, k
    one._sync_scalars()
    for key in ("dt", "max_div_post", "mean_energy", "max_vorticity"):
        assert abs(out["scalars"][key] - one.scalars[-1][key]) <= 1e-12 * abs(one.scalars[-1][key])
    one.close()
