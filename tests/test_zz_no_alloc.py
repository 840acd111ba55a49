"""b200cfd_jacobi_static (include/b200cfd.h): the allocation-free Jacobi entry point, through the C ABI, against the
existing entry point b200cfd_poisson(mode = JACOBI) on the same inputs, twice in a row (the static scratch array keeps the
first call's contents), plus the size limit.  Collected last on purpose (file name).
Tolerances as the other GPU parity tests: rel-L2 <= 1e-12 in fp64, <= 1e-5 in fp32."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_l2

pytestmark = pytest.mark.gpu

TOL = {"f64": 1e-12, "f32": 1e-5}
SCRATCH_BYTES = 8 << 20          # B200CFD_STATIC_SCRATCH_BYTES


def _inputs(ny, nx, dtype, seed):
    rng = np.random.default_rng(seed)
    div = (0.1 * rng.standard_normal((ny, nx))).astype(dtype)
    phi0 = (0.01 * rng.standard_normal((ny, nx))).astype(dtype)
    yy, xx = np.mgrid[0:ny, 0:nx]
    mask = ((xx - nx // 3) ** 2 + (yy - ny // 2) ** 2 <= (min(ny, nx) // 6) ** 2)
    return phi0, div, mask


def _static(lib, torch, phi0, div, mask, dx, dt, sweeps):
    dp = torch.from_numpy(phi0.copy()).cuda()
    dd = torch.from_numpy(div).cuda()
    dm = torch.from_numpy(mask.astype(np.uint8)).cuda() if mask is not None else None
    ny, nx = phi0.shape
    rc = lib.b200cfd_jacobi_static(0 if phi0.dtype == np.float64 else 1, ny, nx, dp.data_ptr(), dd.data_ptr(),
                                   dm.data_ptr() if dm is not None else None, float(dx), float(dt), int(sweeps),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.b200cfd_last_error(None)
    torch.cuda.synchronize()
    return dp.cpu().numpy()


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_static_jacobi_matches_poisson_entry_point_twice(tag):
    import torch
    from cfd_simulations_b200 import _lib, kernels as K
    lib = _lib.load()
    dtype = np.float64 if tag == "f64" else np.float32
    dx, dt = 1.0 / 64, 5e-4
    # two calls in a row on different shapes and sweep parities: the second finds the first's values in the shared array
    for ny, nx, sweeps, seed in [(96, 160, 7, 1), (130, 258, 12, 2)]:
        phi0, div, mask = _inputs(ny, nx, dtype, seed)
        got = _static(lib, torch, phi0, div, mask, dx, dt, sweeps)
        want = K.solve_pressure_jacobi(phi0.copy(), div, dx, dt, mask, sweeps, temporal_depth=1)
        assert np.isfinite(got).all()
        assert rel_l2(got, want) <= TOL[tag]
        want_auto = K.solve_pressure_jacobi(phi0.copy(), div, dx, dt, mask, sweeps)        # temporally blocked default
        assert rel_l2(got, want_auto) <= TOL[tag]
    # and the same call repeated gives the same bits (nothing carried over through the array)
    again = _static(lib, torch, phi0, div, mask, dx, dt, sweeps)
    assert np.array_equal(again, got)
    # no mask
    got = _static(lib, torch, phi0, div, None, dx, dt, 5)
    want = K.solve_pressure_jacobi(phi0.copy(), div, dx, dt, None, 5, temporal_depth=1)
    assert rel_l2(got, want) <= TOL[tag]


def test_static_jacobi_rejects_sizes_beyond_the_array():
    import torch
    from cfd_simulations_b200 import _lib
    lib = _lib.load()
    ny, nx = 1024, 1025                                   # 8 B cells: 8 200 KiB + 8 KiB > 8 MiB
    assert ny * nx * 8 > SCRATCH_BYTES
    d = torch.zeros((ny, nx), dtype=torch.float64, device="cuda")
    rc = lib.b200cfd_jacobi_static(0, ny, nx, d.data_ptr(), d.data_ptr(), None, 0.01, 1e-3, 2, None)
    assert rc == -1                                       # B200CFD_EINVAL
    assert b"B200CFD_STATIC_SCRATCH_BYTES" in lib.b200cfd_last_error(None)
    # the largest size that fits is accepted (fp32: twice the cells)
    f = torch.zeros((1024, 2048), dtype=torch.float32, device="cuda")
    g = torch.zeros_like(f)
    assert lib.b200cfd_jacobi_static(1, 1024, 2048, f.data_ptr(), g.data_ptr(), None, 0.01, 1e-3, 3, None) == 0
    torch.cuda.synchronize()
    assert float(f.abs().max()) == 0.0
