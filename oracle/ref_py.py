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


def _load_cxx():
    """The reference's ``compressai._CXX`` (oracle/_ref/_CXX.so, built by oracle/build.py) or a stub."""
    so = Path(__file__).resolve().parent / "_ref" / "_CXX.so"
    if so.exists():
        import importlib.machinery
        import importlib.util

        loader = importlib.machinery.ExtensionFileLoader("_CXX", str(so))
        spec = importlib.util.spec_from_loader("_CXX", loader)
        mod = importlib.util.module_from_spec(spec)
        loader.exec_module(mod)
        return mod
    cxx = types.ModuleType("compressai._CXX")
    cxx.pmf_to_quantized_cdf = None
    return cxx


def load_models():
    """-> namespace with the reference's own model classes (``Cheng2020AnchorCheckerboardGMMv2``,
    ``Elic2022GMM``) and latent codecs, imported from the files under /root/reference unmodified
    (models/ckbd_gmm.py:61-141, models/elic_gmm.py:64-235)."""
    em, lc = load()
    reg = sys.modules["compressai.registry"]
    reg.register_model = lambda name: (lambda cls: cls)
    for sub in ("layers", "models"):
        if f"compressai.{sub}" not in sys.modules:
            _pkg(f"compressai.{sub}", REF_ROOT / "compressai" / sub)
    layers = importlib.import_module("compressai.layers.layers")
    gdn = importlib.import_module("compressai.layers.gdn")
    lpk = sys.modules["compressai.layers"]
    for k in dir(layers):
        if not k.startswith("_"):
            setattr(lpk, k, getattr(layers, k))
    lpk.GDN = gdn.GDN
    ems = sys.modules["compressai.entropy_models"]
    for k in ("EntropyBottleneck", "EntropyModel", "GaussianConditional", "GaussianMixtureConditional"):
        setattr(ems, k, getattr(em, k))
    lcs = sys.modules["compressai.latent_codecs"]
    for modname, names in (("base", ["LatentCodec"]), ("gaussian_conditional", ["GaussianConditionalLatentCodec"]),
                           ("gaussian_mixture_conditional", ["GaussianMixtureConditionalLatentCodec"]),
                           ("hyper", ["HyperLatentCodec"]), ("hyperprior", ["HyperpriorLatentCodec"]),
                           ("checkerboard", ["CheckerboardLatentCodec"]), ("channel_groups", ["ChannelGroupsLatentCodec"])):
        m = importlib.import_module(f"compressai.latent_codecs.{modname}")
        for n in names:
            setattr(lcs, n, getattr(m, n))
    ns = types.SimpleNamespace(entropy_models=em, latent_codecs=lcs, layers=lpk)
    ns.ckbd_gmm = importlib.import_module("compressai.models.ckbd_gmm")
    ns.elic_gmm = importlib.import_module("compressai.models.elic_gmm")
    ns.Cheng2020AnchorCheckerboardGMMv2 = ns.ckbd_gmm.Cheng2020AnchorCheckerboardGMMv2
    ns.Elic2022GMM = ns.elic_gmm.Elic2022GMM
    return ns


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
        cxx = _load_cxx()
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
