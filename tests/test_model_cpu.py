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
