"""TEST INFRASTRUCTURE — import the UNMODIFIED reference scripts from /root/reference in this container.

Recipe of SURVEY.md Appendix C: stub matplotlib/h5py (absent from the image), register the module in
sys.modules before exec (Numba cache=True needs it), writable NUMBA_CACHE_DIR, scratch cwd (the scripts
create logs/ at import), logger silenced.  Must run with NUMBA_NUM_THREADS=1: with more threads the
reference's clean_divergence_fast is a data race (v5.py:250-253) and is not reproducible.
Only oracle/make_golden.py and opt-in tests use this; /root/reference does not exist on the GPU box.
"""
from __future__ import annotations

import importlib.util
import logging
import os
import sys
import tempfile
import types
from pathlib import Path

REF_ROOT = Path("/root/reference")
CYL_DIR = REF_ROOT / "python" / "flow_over_cylinder (Fischer)"
SHOCK_DIR = REF_ROOT / "python" / "shockwave_analysis"


def available() -> bool:
    return (CYL_DIR / "v5.py").exists()


def _stub_modules():
    noop = lambda *a, **k: None
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.animation",
                 "matplotlib.colors", "matplotlib.cm", "h5py"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    plt = sys.modules["matplotlib.pyplot"]
    plt.switch_backend = noop
    plt.style = types.SimpleNamespace(use=noop)
    plt.rcParams = {}
    sys.modules["matplotlib"].pyplot = plt
    sys.modules["matplotlib"].patches = sys.modules["matplotlib.patches"]
    sys.modules["matplotlib"].use = noop


def load(path: Path, modname: str):
    assert os.environ.get("NUMBA_NUM_THREADS") == "1", "run the reference with NUMBA_NUM_THREADS=1"
    if modname in sys.modules:
        return sys.modules[modname]
    _stub_modules()
    os.environ.setdefault("NUMBA_CACHE_DIR", os.path.join(tempfile.gettempdir(), "numba_cache_ref"))
    scratch = tempfile.mkdtemp(prefix="refcwd_")
    cwd = os.getcwd()
    os.chdir(scratch)
    os.makedirs("logs", exist_ok=True)          # v3 / v4 open logs/<name>.log at import without creating the directory
    try:
        spec = importlib.util.spec_from_file_location(modname, str(path))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
    finally:
        os.chdir(cwd)
    if hasattr(mod, "logger"):
        mod.logger.setLevel(logging.CRITICAL)
        for h in list(mod.logger.handlers):
            mod.logger.removeHandler(h)
    logging.getLogger().setLevel(logging.CRITICAL)
    return mod


def load_v5():
    return load(CYL_DIR / "v5.py", "ref_v5")


def load_v2():
    return load(CYL_DIR / "v2.py", "ref_v2")


def load_shock():
    return load(SHOCK_DIR / "v1_shock.py", "ref_v1_shock")
