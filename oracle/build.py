"""Build recipe for the parity checkers (TEST INFRASTRUCTURE, never the product path).

Two artefacts:

* ``oracle/_build/libgmm_oracle.so`` -- the plain-C CPU restatement in
  ``oracle/gmm_oracle.c`` (compiled with ``-ffp-contract=off`` so that only the
  explicit ``fmaf`` calls fuse).
* ``oracle/_ref/ans_<variant>.so`` -- the UNMODIFIED reference rANS extension,
  compiled from the sources where they lie under ``/root/reference``
  (``compressai/cpp_exts/rans/rans_interface.cpp`` + ``third_party/ryg_rans``)
  with the flags of the reference's ``setup.py:72-76`` (``-std=c++17 -O3``):
    - ``native``: ``-march=native`` exactly like ``setup.py``;
    - ``v3``: ``-march=x86-64-v3`` (AVX2+FMA, no AVX-512) so the prebuilt file
      still runs if the GPU box's host CPU is older than the build container's.
  Both contract the same ``mul+add`` pairs into FMAs (tests assert identical
  bytes). Only built when ``/root/reference`` exists; no reference source is
  copied into this repo. ``oracle/_ref/`` is git-ignored but travels with gpurun.

Usage: ``python oracle/build.py [--force] [--no-ref]``
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import sysconfig
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_ROOT = Path(os.environ.get("FLASHGMM_REFERENCE", "/root/reference"))
REF_OUT = HERE / "_ref"
C_OUT = HERE / "_build"


def _run(cmd):
    print("+", " ".join(str(c) for c in cmd), flush=True)
    subprocess.run([str(c) for c in cmd], check=True)


def _newer(target: Path, *sources: Path) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(s.exists() and s.stat().st_mtime <= t for s in sources)


def cpu_flags() -> set:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def build_c_oracle(force: bool = False) -> Path:
    src = HERE / "gmm_oracle.c"
    out = C_OUT / "libgmm_oracle.so"
    if not force and _newer(out, src):
        return out
    C_OUT.mkdir(exist_ok=True)
    # -ffp-contract=off: the restatement spells every fused op as fmaf();
    # -mfma so fmaf() is one instruction (the value is the same either way).
    _run(["gcc", "-O2", "-std=c11", "-fPIC", "-shared", "-ffp-contract=off",
          "-fno-fast-math", "-mfma", "-mavx2", "-o", out, src, "-lm"])
    return out


def build_ref(force: bool = False):
    """Compile the reference extension in place (sources stay in /root/reference)."""
    src = REF_ROOT / "compressai/cpp_exts/rans/rans_interface.cpp"
    if not src.exists():
        print(f"[oracle] reference sources not found under {REF_ROOT}; keeping prebuilt oracle/_ref")
        return []
    import pybind11
    import torch
    from torch.utils import cpp_extension

    REF_OUT.mkdir(exist_ok=True)
    incs = [sysconfig.get_paths()["include"], pybind11.get_include(),
            REF_ROOT / "third_party/ryg_rans", REF_ROOT / "compressai/cpp_exts/rans"]
    incs += cpp_extension.include_paths()
    torch_lib = Path(torch.__file__).parent / "lib"
    outs = []
    for variant, march in (("native", "-march=native"), ("v3", "-march=x86-64-v3")):
        out = REF_OUT / f"ans_{variant}.so"
        outs.append(out)
        if not force and _newer(out, src):
            continue
        cmd = ["g++", "-std=c++17", march, "-O3", "-fPIC", "-shared", "-w",
               "-DTORCH_EXTENSION_NAME=ans", "-DTORCH_API_INCLUDE_EXTENSION_H",
               f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
        for i in incs:
            cmd += ["-I", i]
        cmd += [src, "-o", out, f"-L{torch_lib}", f"-Wl,-rpath,{torch_lib}",
                "-ltorch", "-ltorch_cpu", "-lc10", "-ltorch_python"]
        _run(cmd)
    # compressai._CXX (pmf_to_quantized_cdf, used by EntropyBottleneck.update): pybind11 only, setup.py:86-97
    ops_src = REF_ROOT / "compressai/cpp_exts/ops/ops.cpp"
    ops_out = REF_OUT / "_CXX.so"
    if ops_src.exists() and (force or not _newer(ops_out, ops_src)):
        _run(["g++", "-std=c++17", "-O3", "-fPIC", "-shared", "-w", "-I", sysconfig.get_paths()["include"],
              "-I", pybind11.get_include(), ops_src, "-o", ops_out])
    outs.append(ops_out)
    meta = {
        "compiler": subprocess.run(["g++", "--version"], capture_output=True, text=True).stdout.splitlines()[0],
        "flags": "-std=c++17 -march={native|x86-64-v3} -O3 (setup.py:72-76)",
        "build_cpu_flags": sorted(f for f in cpu_flags() if f.startswith(("avx", "fma", "bmi", "sse", "f16c"))),
        "torch": torch.__version__,
    }
    (REF_OUT / "build_meta.json").write_text(json.dumps(meta, indent=1))
    return outs


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--no-ref", action="store_true")
    a = ap.parse_args(argv)
    build_c_oracle(a.force)
    if not a.no_ref:
        build_ref(a.force)


if __name__ == "__main__":
    main()
