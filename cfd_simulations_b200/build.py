"""Build libb200cfd.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo snapshot).
One object per translation unit (compiled in parallel, recompiled only when its source or a header changed), then one
link."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = PKG / "build"
LIB = PKG / "libb200cfd.so"
SOURCES = ["step_kernels.cu", "step_fast.cu", "poisson.cu", "poisson_tb.cu", "poisson_stream.cu", "cleandiv.cu", "cleandiv_tma.cu", "finalize_tma.cu", "euler.cu",
           "halo.cu", "static_scratch.cu", "c_api.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC"]


def sources():
    return [CSRC / s for s in SOURCES if (CSRC / s).exists()]


def _headers():
    return list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((PKG.parent / "include").glob("*.h"))


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    return any(d.stat().st_mtime > t for d in sources() + _headers())


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    OBJ.mkdir(exist_ok=True)
    hdr_t = max(h.stat().st_mtime for h in _headers())
    extra = os.environ.get("B200CFD_NVCC_EXTRA", "").split()

    def compile_one(src: Path):
        obj = OBJ / (src.stem + ".o")
        if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, hdr_t):
            return obj, None
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", "-o", str(obj), str(src)]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        r = subprocess.run(cmd, capture_output=True, text=True)
        return obj, (None if r.returncode == 0 else r.stdout + r.stderr)

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 2)) as ex:
        res = list(ex.map(compile_one, sources()))
    errs = [e for _, e in res if e]
    if errs:
        raise RuntimeError("nvcc failed:\n" + "\n".join(errs))
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", str(LIB), *[str(o) for o, _ in res]]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
