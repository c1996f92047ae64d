"""Import the reference's own PYTHON modules for this path (TEST INFRASTRUCTURE).

Only usable where ``/root/reference`` exists (the build container): golden-vector
generators under ``tests/golden/`` call it; nothing on the GPU box does.

``import compressai`` as a whole drags in torch_geometric, pytorch_msssim and
matplotlib (absent here), so a synthetic ``compressai`` package namespace is
registered whose ``__path__`` points into the reference tree: sub-modules are then
imported from the reference's files unmodified, without running the package
``__init__``s. ``compressai.ans`` is the reference's compiled extension
(``oracle/_ref``), ``compressai._CXX`` a stub (only ``EntropyBottleneck.update`` uses it).
"""
from __future__ import annotations

import importlib
import os
import sys
import types
from pathlib import Path

REF_ROOT = Path(os.environ.get("FLASHGMM_REFERENCE", "/root/reference"))


def available() -> bool:
    return (REF_ROOT / "compressai/entropy_models/entropy_models.py").exists()


def _pkg(name: str, path: Path):
    m = types.ModuleType(name)
    m.__path__ = [str(path)]
    m.__package__ = name
    sys.modules[name] = m
    return m


def load():
    """-> (entropy_models module, gaussian_mixture_conditional latent-codec module) of the reference."""
    if "compressai" in sys.modules and not getattr(sys.modules["compressai"], "_flashgmm_ref_stub", False):
        raise RuntimeError("a real `compressai` is already imported")
    if "compressai" not in sys.modules:
        from . import ref

        root = _pkg("compressai", REF_ROOT / "compressai")
        root._flashgmm_ref_stub = True
        root.ans = ref.load_ans()
        sys.modules["compressai.ans"] = root.ans
        cxx = types.ModuleType("compressai._CXX")
        cxx.pmf_to_quantized_cdf = None
        sys.modules["compressai._CXX"] = cxx
        root._CXX = cxx
        root.available_entropy_coders = lambda: ["ans"]
        root.get_entropy_coder = lambda: "ans"
        _pkg("compressai.entropy_models", REF_ROOT / "compressai/entropy_models")
        _pkg("compressai.latent_codecs", REF_ROOT / "compressai/latent_codecs")
        reg = types.ModuleType("compressai.registry")      # the real one imports torch_geometric
        reg.register_module = lambda name: (lambda cls: cls)
        sys.modules["compressai.registry"] = reg
    importlib.import_module("compressai.ops")
    em = importlib.import_module("compressai.entropy_models.entropy_models")
    sys.modules["compressai.entropy_models"].GaussianMixtureConditional = em.GaussianMixtureConditional
    lc = importlib.import_module("compressai.latent_codecs.gaussian_mixture_conditional")
    return em, lc
