"""In-tree build of ``libflashgmm_b200.so`` for sm_100a (nvcc cross-compiles without a GPU).

    python -m flashgmm_b200.build [--force]

The shared library lands next to this file so that it travels with the repo
snapshot to the GPU box; it is git-ignored (``*.so``).
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
INCLUDE = PKG.parent / "include"
LIB = PKG / "libflashgmm_b200.so"

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-O3,-fno-strict-aliasing,-g1", "-I", str(INCLUDE)]

# (source, extra flags). Kernel TUs of the bit-exact path are built with FMA
# contraction off on top of the explicit _rn intrinsics; host-heavy TUs get AVX2.
UNITS = [
    ("gmm_kernels.cu", ["-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false"]),
    ("gmm_ileave.cu", ["-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false"]),
    ("gmm_pop.cu", ["-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false", "-Xptxas", "-v"]),
    ("gmm_push.cu", ["-Xptxas", "-v"]),
    ("gmm_prologue.cu", []),
    ("gmm_likelihood.cu", ["-Xptxas", "-v"]),
    ("gmm_table.cu", []),
    ("eb_likelihood.cu", []),
    ("flashgmm_abi.cu", ["-Xcompiler", "-march=x86-64-v3"]),
    ("abi_glue.cu", ["-Xcompiler", "-march=x86-64-v3"]),
    ("abi_table.cu", ["-Xcompiler", "-march=x86-64-v3"]),
]
# plain C++ units built by g++ with the reference's own flags (setup.py:72-76; FMA available)
CXX_UNITS = [
    ("scalar_gaussian.cpp", ["-O3", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-ffp-contract=fast"]),
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found; flashgmm_b200 needs the CUDA toolkit to build")


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def build(force: bool = False, verbose: bool = True) -> Path:
    hdrs = list(CSRC.glob("*.h")) + list(CSRC.glob("*.hpp")) + list(CSRC.glob("*.cuh")) + list(INCLUDE.glob("*.h"))
    objdir = PKG / "build"
    objdir.mkdir(exist_ok=True)
    objs = []
    cc = nvcc()
    for src, extra in UNITS:
        s = CSRC / src
        o = objdir / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s, Path(__file__)] + hdrs):
            cmd = [cc, *ARCH, *COMMON, *extra, "-c", str(s), "-o", str(o)]
            if verbose:
                print("+", " ".join(cmd), flush=True)
            subprocess.run(cmd, check=True)
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    for src, flags in CXX_UNITS:
        s = CSRC / src
        o = objdir / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s, Path(__file__)]):
            cmd = [cxx, *flags, "-c", str(s), "-o", str(o)]
            if verbose:
                print("+", " ".join(cmd), flush=True)
            subprocess.run(cmd, check=True)
    if force or _stale(LIB, objs):
        cmd = [cc, *ARCH, "-shared", "-o", str(LIB), *map(str, objs), "-Xcompiler", "-pthread"]
        if verbose:
            print("+", " ".join(cmd), flush=True)
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    a = ap.parse_args()
    build(a.force)
    print(LIB)
