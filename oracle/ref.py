"""Client for the reference's own compiled extension (``oracle/_ref/ans_*.so``).

TEST INFRASTRUCTURE ONLY. The reference caches ``APPROX_MODE`` / ``USE_SIMD`` in
function statics on first use (rans_interface.cpp:99-130), so every mode gets its
own worker process (``python -m oracle.ref_worker``) with the environment set
before the module is imported. Requests/replies are pickles over pipes.
"""
from __future__ import annotations

import importlib.util
import json
import os
import pickle
import struct
import subprocess
import sys
from pathlib import Path

_HERE = Path(__file__).resolve().parent
REF_DIR = _HERE / "_ref"


def _cpu_flags():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def pick_variant() -> str | None:
    """'native' if this host has every ISA flag the build host had, else 'v3'."""
    if not (REF_DIR / "ans_v3.so").exists() and not (REF_DIR / "ans_native.so").exists():
        return None
    v = os.environ.get("FLASHGMM_REF_VARIANT")
    if v:
        return v
    meta = REF_DIR / "build_meta.json"
    if (REF_DIR / "ans_native.so").exists() and meta.exists():
        need = set(json.loads(meta.read_text()).get("build_cpu_flags", []))
        if need and need <= _cpu_flags():
            return "native"
    return "v3" if (REF_DIR / "ans_v3.so").exists() else "native"


def available() -> bool:
    return pick_variant() is not None


def load_ans(variant: str | None = None):
    """Import the reference pybind module in THIS process (mode fixed by os.environ)."""
    import torch  # noqa: F401  (libtorch must be loaded first)
    variant = variant or pick_variant()
    path = REF_DIR / f"ans_{variant}.so"
    spec = importlib.util.spec_from_file_location("ans", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class RefWorker:
    """One subprocess running the reference extension with a fixed APPROX_MODE."""

    def __init__(self, mode: int = 0, use_simd: int = 1, variant: str | None = None):
        env = dict(os.environ)
        env["APPROX_MODE"] = str(mode)
        env["USE_SIMD"] = str(use_simd)
        env["OMP_NUM_THREADS"] = "1"
        env["CUDA_VISIBLE_DEVICES"] = ""
        if variant:
            env["FLASHGMM_REF_VARIANT"] = variant
        self.mode = mode
        self.p = subprocess.Popen([sys.executable, "-m", "oracle.ref_worker"], cwd=str(_HERE.parent),
                                  env=env, stdin=subprocess.PIPE, stdout=subprocess.PIPE)

    def call(self, op: str, **kw):
        blob = pickle.dumps((op, kw), protocol=4)
        self.p.stdin.write(struct.pack("<Q", len(blob)))
        self.p.stdin.write(blob)
        self.p.stdin.flush()
        hdr = self.p.stdout.read(8)
        if len(hdr) != 8:
            raise RuntimeError("reference worker died")
        (n,) = struct.unpack("<Q", hdr)
        ok, val = pickle.loads(self.p.stdout.read(n))
        if not ok:
            raise RuntimeError(f"reference worker error: {val}")
        return val

    # numpy in / numpy out; params are (N,4) f32 arrays handed over as K-major
    # (N,K) views with strides (1,N) when kmajor=True, as the reference's caller does.
    def gmm_encode(self, symbols, scales, means, weights, max_value=0, kmajor=True, repeat=1):
        return self.call("gmm_encode", symbols=symbols, scales=scales, means=means, weights=weights,
                         max_value=max_value, kmajor=kmajor, repeat=repeat)

    def gmm_decode(self, encoded, scales, means, weights, max_bs_value, kmajor=True, repeat=1):
        return self.call("gmm_decode", encoded=encoded, scales=scales, means=means, weights=weights,
                         max_bs_value=max_bs_value, kmajor=kmajor, repeat=repeat)

    def table_encode(self, symbols, indexes, cdfs, cdf_sizes, offsets):
        return self.call("table_encode", symbols=symbols, indexes=indexes, cdfs=cdfs,
                         cdf_sizes=cdf_sizes, offsets=offsets)

    def table_decode(self, encoded, indexes, cdfs, cdf_sizes, offsets):
        return self.call("table_decode", encoded=encoded, indexes=indexes, cdfs=cdfs,
                         cdf_sizes=cdf_sizes, offsets=offsets)

    def scalar_encode(self, symbols, scales, max_value=0):
        return self.call("scalar_encode", symbols=symbols, scales=scales, max_value=max_value)

    def scalar_decode(self, encoded, scales, max_bs_value):
        return self.call("scalar_decode", encoded=encoded, scales=scales, max_bs_value=max_bs_value)

    def close(self):
        try:
            self.p.stdin.close()
            self.p.wait(timeout=10)
        except Exception:
            self.p.kill()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
