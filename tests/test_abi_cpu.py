"""CPU: the C-ABI library loads and exports every symbol include/b200cfd.h declares; without a CUDA device
the product path fails loudly (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from cfd_simulations_b200 import _lib

ROOT = Path(__file__).resolve().parent.parent


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    hdr = "".join((ROOT / "include").glob("*.h").__next__().read_text())
    declared = set()
    for f in (ROOT / "include").glob("*.h"):
        declared |= set(re.findall(r"\b(b200cfd_[a-z0-9_]+)\s*\(", f.read_text()))
    assert declared, hdr[:100]
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert set(_lib.EXPORTS) <= declared
    assert lib.b200cfd_abi_version() == 3


def test_config_struct_layout_matches_header():
    # field order / types of b200cfd_config as declared in the header
    txt = (ROOT / "include" / "b200cfd.h").read_text()
    body = txt[txt.index("typedef struct b200cfd_config {"):txt.index("} b200cfd_config;")]
    names = []
    for line in body.splitlines()[1:]:
        line = line.split("/*")[0].strip()
        m = re.match(r"(int32_t|double)\s+([^;]+);", line)
        if m:
            names += [(n.strip(), m.group(1)) for n in m.group(2).split(",")]
    got = [(n, "int32_t" if t is C.c_int32 else "double") for n, t in _lib.Config._fields_]
    assert got == names
    body = txt[txt.index("typedef struct b200cfd_step_scalars {"):txt.index("} b200cfd_step_scalars;")]
    assert C.sizeof(_lib.StepScalars) == 12 * 8


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    cfg = _lib.Config(ny=16, nx=16, dx=0.1, dy=0.1)
    assert lib.b200cfd_create(C.byref(cfg), 0, C.byref(h)) == -2     # B200CFD_ECUDA
    assert b"no CPU fallback" in lib.b200cfd_last_error(None)
    from cfd_simulations_b200.solver import OptimizedTurbulentSolver
    from cfd_simulations_b200.config import OptimizedTurbulentConfig
    with pytest.raises(_lib.B200CFDError):
        OptimizedTurbulentSolver(OptimizedTurbulentConfig(nx=32, ny=16))


def test_product_package_never_imports_oracle():
    for f in (ROOT / "cfd_simulations_b200").rglob("*.py"):
        src = f.read_text()
        assert "import oracle" not in src and "from oracle" not in src, f


def test_euler_exports_and_struct_layout():
    lib = _lib.load()
    txt = (ROOT / "include" / "b200cfd_euler.h").read_text()
    declared = set(re.findall(r"\b(b200cfd_euler_[a-z0-9_]+)\s*\(", txt))
    assert declared == set(_lib.EULER_EXPORTS)
    assert not [s for s in declared if not hasattr(lib, s)]
    body = txt[txt.index("typedef struct b200cfd_euler_config {"):txt.index("} b200cfd_euler_config;")]
    names = []
    for line in body.splitlines()[1:]:
        line = line.split("/*")[0].strip()
        m = re.match(r"(int32_t|double)\s+([^;]+);", line)
        if m:
            names += [(n.strip(), m.group(1)) for n in m.group(2).split(",")]
    got = [(n, "int32_t" if t is C.c_int32 else "double") for n, t in _lib.EulerConfig._fields_]
    assert got == names


def test_euler_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    h = C.c_void_p()
    cfg = _lib.EulerConfig(nx=16, ny=16, nranks=1, dx=0.1, dy=0.1, gamma=1.4)
    assert lib.b200cfd_euler_create(C.byref(cfg), 0, C.byref(h)) == -2
    assert b"no CPU fallback" in lib.b200cfd_euler_last_error(None)
    from cfd_simulations_b200.shockwave import ShockwaveConfig, ShockwaveSolver
    with pytest.raises(_lib.B200CFDError):
        ShockwaveSolver(ShockwaveConfig(nx=32, ny=16))
