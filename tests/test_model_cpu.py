"""CPU-side checks of the entropy-model layer (no GPU): the C oracle, driven through a numpy
restatement of GaussianMixtureConditional.compress's glue (entropy_models.py:833-867,
810-828), reproduces the bytes the reference's own Python + C++ stack produced
(tests/golden/model_small.npz); module surface and error behaviour."""
from pathlib import Path

import numpy as np
import pytest
import torch

import flashgmm_b200 as fg
from golden_inputs import CODEC_CASES
from oracle import port

GOLD = np.load(Path(__file__).resolve().parent / "golden" / "model_small.npz")


def numpy_compress_glue(y, scales, means, weights, K=4):
    """-> symbols int32[N], three (N,K) arrays with strides (1,N), abs_max, zero_bitmap."""
    abs_max = max(int(abs(y.max())), int(abs(y.min()))) + 1
    abs_max = max(abs_max, 1)
    yq = np.round(y)                                   # numpy rounds half to even like torch
    bitmap = (np.abs(yq).sum((3, 2))[0] != 0).astype(np.int64)
    nz = np.nonzero(bitmap)[0]
    sym = yq[:, nz].reshape(-1).astype(np.int32)
    B, KC, H, W = scales.shape

    def planes(t):
        return np.ascontiguousarray(t.reshape(B, K, KC // K, -1)[:, :, nz].transpose(1, 0, 2, 3).reshape(K, -1)).T

    return sym, np.clip(planes(scales), 0.11, 256).astype(np.float32), planes(means), planes(weights), abs_max, bitmap


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("name", sorted(CODEC_CASES))
def test_oracle_reproduces_reference_model_streams(name, mode):
    y, sc, mu, w = (GOLD[f"codec/{name}/em_{k}"] for k in ("y", "scales", "means", "weights"))
    sym, s2, m2, w2, abs_max, bitmap = numpy_compress_glue(y, sc, mu, w)
    key = f"codec/{name}/mode{mode}"
    assert abs_max == int(GOLD[f"{key}/abs_max"])
    assert np.array_equal(bitmap, GOLD[f"{key}/zero_bitmap"])
    if sym.size == 0:
        s2 = m2 = w2 = np.zeros((0, 4), np.float32)
    b = port.gmm_encode(sym, s2, m2, w2, mode)
    assert b == GOLD[f"{key}/bytes"].tobytes()
    d = port.gmm_decode(b, s2, m2, w2, abs_max + 1, mode)
    assert np.array_equal(d, sym)


def test_module_surface():
    m = fg.GaussianMixtureConditional(K=4)
    assert m.K == 4 and m.use_likelihood_bound
    assert float(m.lower_bound_scale.bound) == pytest.approx(0.11)
    assert float(m.likelihood_lower_bound.bound) == pytest.approx(1e-9)
    assert hasattr(m.entropy_coder, "_encoder") and hasattr(m.entropy_coder, "_decoder")
    for meth in ("forward", "compress", "decompress", "reshape_entropy_parameters", "quantize", "_likelihood"):
        assert callable(getattr(m, meth))
    with pytest.raises(ValueError):
        fg.GaussianMixtureConditional(K=4, scale_table=[0.5, 0.1])
    with pytest.raises(ValueError):
        m.quantize(torch.zeros(2), "bogus")
    assert m.quantize(torch.tensor([0.5, 1.5, 2.5, -0.5]), "symbols").tolist() == [0, 2, 2, 0]


def test_reshape_entropy_parameters_is_a_kmajor_view():
    m = fg.GaussianMixtureConditional(K=4)
    B, C, H, W = 1, 6, 3, 5
    sc = torch.arange(B * 4 * C * H * W, dtype=torch.float32).reshape(B, 4 * C, H, W) / 100
    nz = [0, 2, 5]
    s, mu, w = m.reshape_entropy_parameters(sc, sc + 1, sc + 2, nz)
    N = len(nz) * H * W
    assert s.shape == (N, 4) and mu.stride() == (1, N) and w.stride() == (1, N)
    # symbol i = (kept channel c', spatial p); component k reads input channel k*C + nz[c']
    i = 1 * H * W + 7
    for k in range(4):
        assert float(mu[i, k]) == float(sc[0, k * C + nz[1]].reshape(-1)[7] + 1)
    assert float(s.min()) >= 0.11 - 1e-7


def test_no_cpu_fallback():
    m = fg.GaussianMixtureConditional(K=4)
    y = torch.zeros(1, 2, 2, 2)
    p = torch.zeros(1, 8, 2, 2)
    with pytest.raises(RuntimeError):
        m.compress(y, p, p, p)
    with pytest.raises(RuntimeError):
        m.decompress(b"\0" * 8, 1, torch.ones(2, dtype=torch.int64), p, p, p)
    with pytest.raises(RuntimeError):
        m(y, p, p, p)


def test_latent_codec_chunks_and_softmax():
    c = fg.GaussianMixtureConditionalLatentCodec(K=4)
    p = torch.randn(1, 3 * 4 * 5, 2, 3)
    s, m_, w = c._chunk(p)
    assert s.shape == m_.shape == w.shape == (1, 20, 2, 3)
    ww = c._reshape_gmm_weight(w)
    assert ww.shape == w.shape
    torch.testing.assert_close(ww.reshape(1, 4, 5, 2, 3).sum(1), torch.ones(1, 5, 2, 3))
    with pytest.raises(AssertionError):
        fg.GaussianMixtureConditionalLatentCodec(K=4, quantizer="ste")


# ---- the model replicas (tests/gmm_models.py) against the reference's own classes (build container only) ----
def _ref_models():
    from oracle import ref_py

    if not ref_py.available():
        pytest.skip("/root/reference is not present on this machine")
    return ref_py.load_models()


SMALL = {"Cheng2020AnchorCheckerboardGMMv2": dict(N=32, K=4), "Elic2022GMM": dict(N=32, M=144, K=4)}


@pytest.mark.parametrize("name", sorted(SMALL))
def test_replicas_have_the_reference_state_dict_and_seeded_weights(name):
    import gmm_models as gm

    ns = _ref_models()
    torch.manual_seed(0)
    ref = getattr(ns, name)(**SMALL[name]).eval()
    torch.manual_seed(0)
    mine = getattr(gm, name)(**SMALL[name]).eval()
    for stage in ("init", "updated"):
        a, b = ref.state_dict(), mine.state_dict()
        assert list(a.keys()) == list(b.keys())
        for k in a:
            assert a[k].shape == b[k].shape and torch.equal(a[k], b[k]), (stage, k)
        ref.update()
        mine.update()
    x = torch.rand(1, 3, 64, 64)
    with torch.no_grad():
        y = ref.g_a(x)
        assert torch.equal(y, mine.g_a(x)) and torch.equal(ref.g_s(y), mine.g_s(y))


@pytest.mark.parametrize("name", sorted(SMALL))
@pytest.mark.parametrize("updated", [False, True])
def test_reference_checkpoints_load_strict(name, updated):
    """state dicts produced by the REFERENCE's module, before and after net.update(), load with
    strict=True (the 64-entry scale table and the CDF buffers are resized on load)."""
    import gmm_models as gm

    ns = _ref_models()
    torch.manual_seed(3)
    ref = getattr(ns, name)(**SMALL[name]).eval()
    if updated:
        ref.update()
    sd = ref.state_dict()
    torch.manual_seed(4)
    mine = getattr(gm, name)(**SMALL[name]).eval()
    mine.load_state_dict(sd, strict=True)
    for k, v in mine.state_dict().items():
        assert torch.equal(v, sd[k]), k
    gmc = next(m for m in mine.modules() if isinstance(m, fg.GaussianMixtureConditional))
    assert gmc._offset.numel() == (64 if updated else 0)
    # and the other way round: what this package saves loads into the reference's model
    torch.manual_seed(5)
    back = getattr(ns, name)(**SMALL[name]).eval()
    back.load_state_dict(mine.state_dict(), strict=True)


@pytest.mark.parametrize("name", sorted(SMALL))
def test_mount_on_the_reference_model_tree(name):
    """mount() swaps the entropy models and rebinds the codecs' coding methods on the reference's OWN instances;
    state_dict keys and values, and forward's module tree, are untouched."""
    ns = _ref_models()
    torch.manual_seed(0)
    net = getattr(ns, name)(**SMALL[name]).eval()
    net.update()
    before = {k: v.clone() for k, v in net.state_dict().items()}
    fg.mount(net)
    after = net.state_dict()
    assert list(before.keys()) == list(after.keys())
    for k in before:
        assert torch.equal(before[k], after[k]), k
    kinds = {type(m).__name__: type(m).__module__ for m in net.modules()}
    assert kinds["GaussianMixtureConditional"].startswith("flashgmm_b200")
    assert kinds["EntropyBottleneck"].startswith("flashgmm_b200")
    assert kinds["CheckerboardLatentCodec"].startswith("compressai")          # identity kept, methods rebound
    ck = next(m for m in net.modules() if type(m).__name__ == "CheckerboardLatentCodec")
    assert ck.compress.__func__ is fg.CheckerboardLatentCodec.compress
    assert ck.latent_codec["y"].gaussian_mixture_conditional is ck.y.gaussian_mixture_conditional
    hp = net.latent_codec
    assert hp.latent_codec["hyper"].entropy_bottleneck is hp.hyper.entropy_bottleneck
    # checkerboard (un)embedding of this package == the reference's (data movement only)
    t = torch.arange(2 * 3 * 5 * 8, dtype=torch.float32).reshape(2, 3, 5, 8)
    ref_ck = ns.latent_codecs.CheckerboardLatentCodec(forward_method="onepass")
    for parity in ("even", "odd"):
        ref_ck.anchor_parity = parity
        mine = fg.CheckerboardLatentCodec(anchor_parity=parity, forward_method="onepass")
        assert torch.equal(mine.unembed(t), ref_ck.unembed(t))
        assert torch.equal(mine.embed(ref_ck.unembed(t)), t)
    # a mounted model still refuses CPU tensors on the coding path (no CPU fallback)
    with pytest.raises(RuntimeError):
        net.compress(torch.rand(1, 3, 64, 64))


def test_pmf_to_quantized_cdf_matches_the_reference_extension():
    from oracle import ref_py
    from flashgmm_b200.entropy_models import pmf_to_quantized_cdf

    if not ref_py.available():
        pytest.skip("/root/reference is not present on this machine")
    cxx = ref_py._load_cxx()
    if cxx.pmf_to_quantized_cdf is None:
        pytest.skip("oracle/_ref/_CXX.so not built")
    rng = np.random.default_rng(0)
    for n in (2, 3, 17, 64, 300):
        for sharp in (0.5, 3.0, 12.0):
            p = rng.random(n).astype(np.float32) ** sharp
            p[rng.integers(0, n)] += 1.0
            p = (p / p.sum()).astype(np.float32)
            assert pmf_to_quantized_cdf(p.tolist(), 16) == list(cxx.pmf_to_quantized_cdf(p.tolist(), 16))


def test_net_runner_eager_modes_on_cpu_tensors():
    """graphs.NetRunner without a GPU: both batching modes run the step eagerly and agree (a 1x1 conv is batch-invariant
    on the CPU); per-image mode assembles image b's rows from the step run on image b alone."""
    import torch.nn as nn
    from flashgmm_b200.graphs import NetRunner

    torch.manual_seed(0)
    conv = nn.Conv2d(6, 4, 1)
    x = torch.randn(5, 6, 3, 7)
    side = torch.randn(5, 2, 3, 7)

    def step(a, b):
        return (torch.cat((conv(a), b), dim=1), a.sum(dim=1, keepdim=True))
    outs = {}
    for mode in ("per_image", "batched"):
        r = NetRunner(mode=mode)
        outs[mode] = r(("k", 1), step, x, side)
        assert r.captures == 0 and r.replays == 0               # no CUDA: nothing is captured
        assert outs[mode][0].shape == (5, 6, 3, 7) and outs[mode][1].shape == (5, 1, 3, 7)
    for a, b in zip(outs["per_image"], outs["batched"]):
        torch.testing.assert_close(a, b)
    one = NetRunner(mode="per_image")(("k", 1), step, x[2:3], side[2:3])
    assert torch.equal(one[0], outs["per_image"][0][2:3])
    with pytest.raises(ValueError):
        NetRunner(mode="bogus")


def test_mount_is_idempotent_and_keeps_own_classes():
    import gmm_models as gm

    torch.manual_seed(0)
    net = gm.Cheng2020AnchorCheckerboardGMMv2(N=16, K=4).eval()
    before = {id(m) for m in net.modules()}
    assert fg.mount(net) is net and fg.mount(net) is net
    assert {id(m) for m in net.modules()} == before          # nothing to swap in a tree built from this package's classes


def test_thread_bindings_of_ctx_and_runner_are_per_thread():
    """flashgmm_b200.Lane's plumbing: a ctx / runner bound on one thread is invisible on another."""
    import threading
    from flashgmm_b200 import _lib
    from flashgmm_b200.graphs import NetRunner, bind_thread_runner, default_runner, thread_runner
    a, b = object(), object()
    seen = {}
    assert _lib.bind_thread_ctx(3, a) is None
    r_main = NetRunner(graphs=False)
    assert bind_thread_runner(r_main) is None

    def other():
        seen["ctx_before"] = getattr(_lib._tls, "ctxs", None)
        _lib.bind_thread_ctx(3, b)
        seen["ctx"] = _lib.ctx(3)
        seen["runner"] = thread_runner()

    t = threading.Thread(target=other)
    t.start()
    t.join()
    try:
        assert not seen["ctx_before"] and seen["ctx"] is b and seen["runner"] is None
        assert _lib.ctx(3) is a and default_runner() is r_main
    finally:
        assert _lib.bind_thread_ctx(3, None) is a
        assert bind_thread_runner(None) is r_main
    assert thread_runner() is None
