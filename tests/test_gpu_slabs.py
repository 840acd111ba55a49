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


def _poisson_case(nx=1200, ny=160):
    cfg = _cfg(nx, ny)
    div = 0.1 * np.random.default_rng(11).standard_normal((ny, nx))
    return cfg, div


@pytest.mark.parametrize("nslabs", [2, 3])
def test_local_slabs_poisson_matches_single_slab(nslabs):
    """the iterative ("as written") red-black Poisson solve on slabs — halo refresh every T iterations,
    per-iteration max|dphi| reduced over the slabs, exact stop — must give the single-slab iterates, iteration count and
    last change, for fixed counts that are not multiples of the refresh period and in tolerance mode."""
    from cfd_simulations_b200.multigpu import LocalSlabs
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    cfg, div = _poisson_case()
    one = OptimizedTurbulentSolver(cfg)
    many = LocalSlabs(cfg, nslabs)
    single = lambda iters, tol, omega: one.solve_pressure_fast(div, mode="as_written", max_iter=iters, tol=tol, omega=omega,
                                                               temporal_depth=1, return_info=True)
    for omega, iters in ((1.0, 23), (1.5, 16), (1.0, 1)):
        p1, it1, l1 = single(iters, 0.0, omega)
        pm, itm, lm = many.solve_pressure(div, iters, 0.0, omega)
        # fixed-count chunks run through the temporally blocked kernel: same iterates up to FMA contraction (1e-15)
        assert itm == iters and rel_l2(pm, p1) <= 1e-14, (omega, iters, rel_l2(pm, p1))
        assert abs(lm - l1) <= 1e-12 * l1
    # tolerance mode: a tolerance between the changes of iterations 37 and 38 stops both after exactly 38 iterations
    l37, l38 = single(37, 0.0, 1.0)[2], single(38, 0.0, 1.0)[2]
    assert l38 < l37
    tol = 0.5 * (l37 + l38)
    p1, it1, l1 = single(5000, tol, 1.0)
    pm, itm, lm = many.solve_pressure(div, 5000, tol, 1.0)
    assert it1 == 38 and itm == 38 and lm == l1 and np.array_equal(pm, p1)
    one.close(); many.close()


@pytest.mark.parametrize("nslabs", [2, 3])
def test_local_slabs_jacobi_matches_single_slab(nslabs):
    """The Jacobi branch (v5.py:336-346) on slabs: halos refreshed once per min(halo) sweeps, same field as one slab."""
    from cfd_simulations_b200.multigpu import LocalSlabs
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    cfg, div = _poisson_case()
    one = OptimizedTurbulentSolver(cfg)
    many = LocalSlabs(cfg, nslabs)
    for sweeps in (37, 16, 1):
        p1 = one.solve_pressure_fast(div, mode="jacobi", max_iter=sweeps, temporal_depth=1)
        pm, itm, _ = many.solve_pressure(div, sweeps, 0.0, mode="jacobi")
        assert itm == sweeps and rel_l2(pm, p1) <= 1e-14, (sweeps, rel_l2(pm, p1))
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
        # snapshot of a slab-decomposed solver: every rank calls the writer, rank 0 alone writes the GLOBAL field
        import tempfile
        from cfd_simulations_b200.snapshot import SnapshotWriter, load_snapshot
        snap_dir = out["snap_dir"]
        cfg_s = _cfg(1600, 128)
        cfg_s.output_dir, cfg_s.hdf5_file = snap_dir, "slabs.h5"
        with SnapshotWriter(cfg_s, backend="npz") as w:
            w.save_data_to_hdf5(s, 15, 3e-4).result()
            dist.barrier()
            if rank == 0:
                out["snapshot"] = load_snapshot(w.path, 15)
        # iterative Poisson on the slabs: NCCL halo refresh + all_reduce(MAX) of the per-iteration change
        _, div = _poisson_case(1600, 128)
        p = s.plan
        loc = div[:, p.x_offset:p.x_offset + p.nx_local]
        _, itf, lf = s.solve_pressure_slabs(loc, max_iter=23, tol=0.0, omega=1.0, return_info=True)
        phi_fixed = s.gather("phi")
        _, itt, lt = s.solve_pressure_slabs(loc, max_iter=5000, tol=float(out["tol"]), omega=1.0, return_info=True)
        phi_tol = s.gather("phi")
        _, itj, _ = s.solve_pressure_slabs(loc, max_iter=37, return_info=True, mode="jacobi")
        phi_jac = s.gather("phi")
        if rank == 0:
            out["fields"] = res
            out["scalars"] = s.scalars[-1]
            out["poisson"] = {"fixed": (phi_fixed, itf, lf), "tol": (phi_tol, itt, lt)}
            out["jacobi"] = (phi_jac, itj)
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
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    ref = OptimizedTurbulentSolver(_cfg(1600, 128))
    _, div = _poisson_case(1600, 128)
    single = lambda iters, tol: ref.solve_pressure_fast(div, mode="as_written", max_iter=iters, tol=tol, omega=1.0,
                                                        temporal_depth=1, return_info=True)
    import tempfile
    out["snap_dir"] = tempfile.mkdtemp(prefix="b200cfd_slab_snap_")
    out["tol"] = 0.5 * (single(37, 0.0)[2] + single(38, 0.0)[2])
    pf = single(23, 0.0)
    pt = single(5000, float(out["tol"]))
    pf, pt = (pf[0].copy(), pf[1], pf[2]), (pt[0].copy(), pt[1], pt[2])
    pj = ref.solve_pressure_fast(div, mode="jacobi", max_iter=37, temporal_depth=1).copy()
    ref.close()
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
    snap = out["snapshot"]
    assert snap["u"].shape == (128, 1600) and snap["X"].shape == (128, 1600)
    assert rel_l2(snap["u"], one.u) <= 1e-13 and rel_l2(snap["v"], one.v) <= 1e-13
    w1 = one.compute_vorticity()
    ok = ~np.isnan(w1)
    assert np.array_equal(np.isnan(snap["vorticity"]), ~ok)
    assert np.abs(snap["vorticity"][ok] - w1[ok]).max() <= 1e-10 * np.abs(w1[ok]).max()
    one._sync_scalars()
    for key in ("dt", "max_div_post", "mean_energy", "max_vorticity"):
        assert abs(out["scalars"][key] - one.scalars[-1][key]) <= 1e-12 * abs(one.scalars[-1][key])
    for key, exp in (("fixed", pf), ("tol", pt)):
        got = out["poisson"][key]
        assert got[1] == exp[1] and abs(got[2] - exp[2]) <= 1e-12 * exp[2] and rel_l2(got[0], exp[0]) <= 1e-14, key
    assert pt[1] == 38
    assert out["jacobi"][1] == 37 and rel_l2(out["jacobi"][0], pj) <= 1e-14
    one.close()
