"""GaussianMixtureConditional / latent codec on the GPU against golden vectors produced
by the reference's own Python modules (tests/golden/make_model_golden.py) and against a
plain PyTorch fp32 restatement of the likelihood.

Tolerances. The coder is integer work: bit-exact. The likelihood is fp32: the contract is 1e-5
relative with the likelihood bound 1e-9 as the absolute floor, and that is what the
forward tests use. Gradients are differences of pdf values (more cancellation): see the
per-test tolerances; profiles/r02_likelihood_error_hist.txt has the measured error histograms
(kernel vs reference goldens, and each of them vs float64).
"""
from pathlib import Path

import numpy as np
import pytest
import torch

import flashgmm_b200 as fg
from flashgmm_b200 import ans
from flashgmm_b200.entropy_models import _GMMLikelihoodFn
from golden_inputs import CODEC_CASES, LIK_CASES, codec_inputs, lik_inputs

pytestmark = pytest.mark.gpu
GOLD = np.load(Path(__file__).resolve().parent / "golden" / "model_small.npz")
RTOL, ATOL = 1e-5, 1e-9

QUANT = {k: v[3] for k, v in CODEC_CASES.items()}


def torch_likelihood(y, sc, mu, w, K, scale_bound=0.11):
    """Plain PyTorch restatement of entropy_models.py:774-790 + 720-737 (any device)."""
    M = y.size(1)
    out = torch.zeros_like(y)
    c = -(2 ** -0.5)
    for k in range(K):
        s = torch.clamp_min(sc[:, M * k: M * (k + 1)], scale_bound)
        v = torch.abs(y - mu[:, M * k: M * (k + 1)])
        up = 0.5 * torch.erfc(c * ((0.5 - v) / s))
        lo = 0.5 * torch.erfc(c * ((-0.5 - v) / s))
        out = out + (up - lo) * w[:, M * k: M * (k + 1)]
    return out


def close(a, b, rtol=RTOL, atol=ATOL):
    a = a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a
    err = np.abs(a.astype(np.float64) - b.astype(np.float64))
    lim = rtol * np.abs(b.astype(np.float64)) + atol
    assert (err <= lim).all(), f"max excess {np.max(err - lim):.3e} at {np.unravel_index(np.argmax(err - lim), err.shape)}"


@pytest.mark.parametrize("name", sorted(LIK_CASES))
def test_forward_matches_reference_golden(name):
    B, M, H, W, K, seed = LIK_CASES[name]
    x, sc, mu, w = (t.cuda() for t in lik_inputs(B, M, H, W, K, seed))
    m = fg.GaussianMixtureConditional(K=K).cuda().eval()
    y, lik = m(x, sc, mu, w)
    assert np.array_equal(y.cpu().numpy(), GOLD[f"lik/{name}/y"])         # round-half-even, exact
    close(lik, GOLD[f"lik/{name}/likelihood"])
    assert float(lik.min()) >= float(np.float32(1e-9))


@pytest.mark.parametrize("name", sorted(LIK_CASES))
def test_backward_matches_reference_golden(name):
    B, M, H, W, K, seed = LIK_CASES[name]
    x, sc, mu, w = lik_inputs(B, M, H, W, K, seed)
    y = torch.from_numpy(GOLD[f"lik/{name}/y"])
    xs = [t.cuda().requires_grad_(True) for t in (y, sc, mu, w)]
    _, lik = _GMMLikelihoodFn.apply(xs[0], None, 0, xs[1], xs[2], xs[3], K, 0.11, 1e-9)
    torch.log(lik).sum().backward()
    for t, nm in zip(xs, ("gy", "gscales", "gmeans", "gweights")):
        want = GOLD[f"lik/{name}/{nm}"]
        # gradients of log-likelihood: quotient of fp32 quantities -> looser absolute floor, scaled
        close(t.grad, want, rtol=2e-4, atol=2e-4 * max(1.0, float(np.abs(want).max())) * 1e-2)


@pytest.mark.parametrize("shape", [(4, 192, 16, 16), (1, 320, 32, 48), (3, 7, 5, 3)])
def test_forward_backward_match_torch_fp32(shape):
    B, M, H, W = shape
    K = 4
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(shape, device="cuda", generator=g) * 4
    sc = torch.rand(B, K * M, H, W, device="cuda", generator=g) * 2 + 0.02
    mu = torch.randn(B, K * M, H, W, device="cuda", generator=g) * 2
    w = torch.softmax(torch.randn(B, K, M, H, W, device="cuda", generator=g), 1).reshape(B, K * M, H, W)
    y = torch.round(x)
    a = [t.clone().requires_grad_(True) for t in (y, sc, mu, w)]
    b = [t.clone().requires_grad_(True) for t in (y, sc, mu, w)]
    _, lik = _GMMLikelihoodFn.apply(a[0], None, 0, a[1], a[2], a[3], K, 0.11, 1e-9)
    lb, sb = fg.LowerBound(1e-9).cuda(), fg.LowerBound(0.11).cuda()
    ref = lb(torch_likelihood(b[0], sb(b[1]), b[2], b[3], K, scale_bound=0.0))
    close(lik, ref.detach().cpu().numpy())
    gout = torch.randn(shape, device="cuda", generator=g)
    lik.backward(gout)
    ref.backward(gout)
    for ta, tb in zip(a, b):
        scale = float(tb.grad.abs().max())
        close(ta.grad, tb.grad.cpu().numpy(), rtol=1e-4, atol=1e-5 * max(scale, 1e-3))


def test_training_mode_noise_and_grad_flow():
    torch.manual_seed(0)
    m = fg.GaussianMixtureConditional(K=2).cuda().train()
    x = torch.rand(1, 128, 32, 32, device="cuda", requires_grad=True)
    sc = torch.rand(1, 256, 32, 32, device="cuda", requires_grad=True)
    mu = torch.rand(1, 256, 32, 32, device="cuda", requires_grad=True)
    w = torch.softmax(torch.rand(1, 2, 128, 32, 32, device="cuda"), 1).reshape(1, 256, 32, 32).requires_grad_(True)
    y, lik = m(x, sc, mu, w)                                    # reference test: test_entropy_models.py:436-454
    assert y.shape == x.shape and lik.shape == x.shape
    d = (y - x).detach()
    assert float(d.min()) >= -0.5 and float(d.max()) <= 0.5 and float(d.abs().mean()) > 0.1
    (y.sum() + torch.log(lik).sum()).backward()
    for t in (x, sc, mu, w):
        assert t.grad is not None and torch.isfinite(t.grad).all()
    assert float((x.grad - 1).abs().max()) > 0                  # likelihood path reaches the input


def test_eval_mode_rounds_and_blocks_input_grad():
    m = fg.GaussianMixtureConditional(K=2).cuda().eval()         # reference: test_entropy_models.py:477-491
    x = (torch.rand(1, 16, 8, 8, device="cuda") * 10).requires_grad_(True)
    p = torch.rand(1, 32, 8, 8, device="cuda")
    y, lik = m(x, p, p, torch.full_like(p, 0.5))
    assert torch.equal(y, torch.round(x))
    lik.sum().backward()
    assert float(x.grad.abs().max()) == 0.0


def test_k1_equals_single_gaussian():
    """K=1 with weight 1 is GaussianConditional._likelihood (reference: test_entropy_models.py:456-475)."""
    m = fg.GaussianMixtureConditional(K=1).cuda().eval()
    x = torch.randn(2, 8, 6, 6, device="cuda") * 3
    sc = torch.rand(2, 8, 6, 6, device="cuda") + 0.2
    mu = torch.randn(2, 8, 6, 6, device="cuda")
    y, lik = m(x, sc, mu, torch.ones_like(sc))
    ref = torch_likelihood(torch.round(x), sc, mu, torch.ones_like(sc), 1)
    close(lik, torch.clamp_min(ref, 1e-9).cpu().numpy())


def test_cpu_tensors_are_refused():
    m = fg.GaussianMixtureConditional(K=4)
    t = torch.zeros(1, 4, 2, 2)
    with pytest.raises(RuntimeError):
        m(t, t.repeat(1, 4, 1, 1), t.repeat(1, 4, 1, 1), t.repeat(1, 4, 1, 1))


# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("name", sorted(CODEC_CASES))
def test_entropy_model_compress_decompress_golden(name, mode):
    """Same tensors the reference's entropy model received -> identical bytes / abs_max / bitmap / y_hat."""
    ans.set_approx_mode(mode)
    try:
        y, sc, mu, w = (torch.from_numpy(GOLD[f"codec/{name}/em_{k}"]).cuda() for k in ("y", "scales", "means", "weights"))
        m = fg.GaussianMixtureConditional(K=4).cuda()
        (bts, abs_max, bitmap), y_q = m.compress(y, sc, mu, w)
        key = f"codec/{name}/mode{mode}"
        assert isinstance(bts, bytes) and isinstance(abs_max, int) and bitmap.dtype == torch.int64
        assert abs_max == int(GOLD[f"{key}/abs_max"])
        assert np.array_equal(bitmap.cpu().numpy(), GOLD[f"{key}/zero_bitmap"])
        assert bts == GOLD[f"{key}/bytes"].tobytes()
        y_hat = m.decompress(bts, abs_max, bitmap, sc, mu, w)
        assert y_hat.shape == y.shape
        assert torch.equal(y_hat, torch.round(y))
        assert torch.equal(y_q, torch.round(y))
        # and the reference's own stream decodes to the reference's own result
        y_ref = m.decompress(GOLD[f"{key}/bytes"].tobytes(), abs_max, bitmap, sc, mu, w)
        want = GOLD[f"{key}/y_hat_dec"]
        if QUANT[name] == "noise":
            assert np.array_equal(y_ref.cpu().numpy(), want)
        else:   # the codec adds the weighted mean back after the entropy model
            assert np.array_equal(y_ref.cpu().numpy(), np.round(GOLD[f"codec/{name}/em_y"]))
    finally:
        ans.set_approx_mode(0)


@pytest.mark.parametrize("name", sorted(CODEC_CASES))
def test_latent_codec_roundtrip_and_golden(name):
    """Codec level: softmax / recentring run on the GPU here and on the CPU in the golden, so the
    parameters differ in the last ulp; symbols, side info and the round trip must still agree."""
    C, h, w_, _q, seed, gain = CODEC_CASES[name]
    y, params = (t.cuda() for t in codec_inputs(C, h, w_, seed, gain))
    codec = fg.GaussianMixtureConditionalLatentCodec(K=4, quantizer=QUANT[name]).cuda()
    enc = codec.compress(y, params)
    (bts, abs_max, bitmap) = enc["strings"][0]
    key = f"codec/{name}/mode0"
    assert abs_max == int(GOLD[f"{key}/abs_max"])
    assert np.array_equal(bitmap.cpu().numpy(), GOLD[f"{key}/zero_bitmap"])
    assert abs(len(bts) - GOLD[f"{key}/bytes"].size) <= 8
    assert tuple(enc["shape"]) == (h, w_)
    dec = codec.decompress(enc["strings"], enc["shape"], params)
    if QUANT[name] == "noise":
        assert np.array_equal(enc["y_hat"].cpu().numpy(), GOLD[f"{key}/y_hat_enc"])
        assert np.array_equal(dec["y_hat"].cpu().numpy(), GOLD[f"{key}/y_hat_dec"])
        assert torch.equal(dec["y_hat"], enc["y_hat"])
    else:
        np.testing.assert_allclose(dec["y_hat"].cpu().numpy(), GOLD[f"{key}/y_hat_dec"], rtol=0, atol=1e-4)
    out = codec.eval()(y, params)
    assert out["y_hat"].shape == y.shape and out["likelihoods"]["y"].shape == y.shape


def test_state_dict_keys_match_reference():
    want = ["_offset", "_quantized_cdf", "_cdf_length", "scale_table", "scale_bound",
            "likelihood_lower_bound.bound", "lower_bound_scale.bound"]     # reference module, printed in this container
    assert sorted(fg.GaussianMixtureConditional(K=4).state_dict().keys()) == sorted(want)
    codec = fg.GaussianMixtureConditionalLatentCodec(K=4)
    assert sorted(codec.state_dict().keys()) == sorted("gaussian_mixture_conditional." + k for k in want)


# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(1, 7, 5, 3), (1, 64, 16, 24), (1, 320, 32, 48), (1, 1, 1, 1)])
def test_compress_glue_kernels_match_the_tensor_ops(shape):
    """csrc/gmm_prologue.cu against the reference's own tensor ops (entropy_models.py:810-851, 902-909):
    abs_max, y_quantized, zero bitmap, compacted symbols, K-major parameter views, scatter back."""
    _B, Cc, H, W = shape
    g = torch.Generator().manual_seed(sum(shape))
    y = (torch.randn(shape, generator=g) * 3).cuda()
    y[:, ::3] *= 0.05                                   # channels that round to all zeros
    y[0, -1, 0, 0] = -17.6                              # the extremum is negative
    sc = (torch.rand((1, 4 * Cc, H, W), generator=g) * 3 + 0.05).cuda()      # some below the 0.11 clamp
    mu = torch.randn((1, 4 * Cc, H, W), generator=g).cuda()
    w = torch.softmax(torch.randn((1, 4, Cc, H, W), generator=g), 1).reshape(1, 4 * Cc, H, W).cuda()
    m = fg.GaussianMixtureConditional(K=4).cuda()

    # the reference's ops
    hi, lo = torch.stack((y.max(), y.min())).abs().int().tolist()
    want_abs_max = max(max(hi, lo) + 1, 1)
    yq = torch.round(y)
    want_bitmap = torch.where(torch.sum(torch.abs(yq), (3, 2)).squeeze(0) == 0, 0, 1)
    nz = torch.nonzero(want_bitmap).flatten().tolist()
    want_sym = yq[:, nz].reshape(-1).int()
    want_sc, want_mu, want_w = m.reshape_entropy_parameters(sc, mu, w, nz)

    (bts, abs_max, bitmap), y_q = m.compress(y, sc, mu, w)
    assert abs_max == want_abs_max
    assert bitmap.dtype == torch.int64 and torch.equal(bitmap, want_bitmap)
    assert torch.equal(y_q, yq)
    assert torch.equal(m.y_q, want_sym)
    ((sym, a, b, c),) = m._gather(yq, sc, mu, w, want_bitmap.reshape(1, -1).contiguous(), [len(nz)])
    assert torch.equal(sym, want_sym)
    for got, want in ((a, want_sc), (b, want_mu), (c, want_w)):
        assert got.shape == want.shape and got.stride() == (1, got.shape[0]) and torch.equal(got, want)
    # same bytes as the coder fed with the reference's tensors, and the round trip
    assert bts == ans.RansEncoder().encode_with_indexes_gmm(want_sym, want_sc, want_mu, want_w, abs_max + 1)
    m2 = fg.GaussianMixtureConditional(K=4).cuda()
    y_hat = m2.decompress(bts, abs_max, bitmap.cpu(), sc, mu, w)      # bitmap may arrive on the CPU
    assert y_hat.shape == y.shape and torch.equal(y_hat, yq)


def test_batched_glue_equals_per_image_glue():
    """B images per launch (scan / gather / scatter): image b's symbols, parameter views, abs_max, bitmap and
    bytes are those of coding image b alone; ragged kept-channel counts, one image with nothing kept."""
    B, Cc, H, W = 5, 24, 8, 6
    g = torch.Generator().manual_seed(77)
    y = (torch.randn((B, Cc, H, W), generator=g) * 4).cuda()
    y[0, ::2] *= 0.01
    y[2] *= 0.001                                       # image 2: every channel rounds to zero
    y[3, 5:] *= 0.01
    y[4, 0, 0, 0] = 300.7
    sc = (torch.rand((B, 4 * Cc, H, W), generator=g) * 3 + 0.05).cuda()
    mu = torch.randn((B, 4 * Cc, H, W), generator=g).cuda()
    w = torch.softmax(torch.randn((B, 4, Cc, H, W), generator=g), 1).reshape(B, 4 * Cc, H, W).cuda()
    m = fg.GaussianMixtureConditional(K=4).cuda()
    items, y_q = m.compress_batch(y, sc, mu, w)
    assert len(items) == B and torch.equal(y_q, torch.round(y))
    for b in range(B):
        one = fg.GaussianMixtureConditional(K=4).cuda()
        (bts, abs_max, bitmap), yq1 = one.compress(y[b:b + 1].contiguous(), sc[b:b + 1], mu[b:b + 1], w[b:b + 1])
        assert items[b][0] == bts and items[b][1] == abs_max and torch.equal(items[b][2], bitmap)
        assert torch.equal(yq1, y_q[b:b + 1])
    assert items[2][0] == bytes.fromhex("0000008000000000") and int(items[2][2].sum()) == 0
    y_hat = m.decompress_batch(items, sc, mu, w)
    assert torch.equal(y_hat, torch.round(y))
    # bitmaps arriving on the CPU (from a container), tuples or lists
    cpu_items = [(it[0], it[1], it[2].cpu()) for it in items]
    assert torch.equal(m.decompress_batch(cpu_items, sc, mu, w), torch.round(y))


def test_bounds_are_read_from_the_loaded_buffers():
    """a checkpoint whose bounds differ from the constructor's must be honoured."""
    m = fg.GaussianMixtureConditional(K=2).cuda().eval()
    sd = m.state_dict()
    sd["lower_bound_scale.bound"] = torch.tensor([0.5])
    sd["likelihood_lower_bound.bound"] = torch.tensor([1e-3])
    m.load_state_dict(sd)
    x = torch.tensor([[[[0.0, 9.0]]]]).cuda()
    sc = torch.full((1, 2, 1, 2), 0.2).cuda()
    mu = torch.zeros(1, 2, 1, 2).cuda()
    w = torch.full((1, 2, 1, 2), 0.5).cuda()
    _, lik = m(x, sc, mu, w)
    want = torch_likelihood(x, sc, mu, w, 2, scale_bound=0.5).clamp_min(1e-3)
    close(lik, want.cpu().numpy())
    assert float(lik[0, 0, 0, 1]) == pytest.approx(1e-3)


def test_gradcheck_finite_differences():
    """gradcheck for the fused fwd / bwd pair, in float32 (the kernel has no float64 form): central
    differences of sum(likelihood * r) on a 3 x 7 x 5 x 3 case, away from the bounds' kinks."""
    g = torch.Generator().manual_seed(5)
    B, M, H, W, K = 3, 7, 5, 3, 4
    x = (torch.randn(B, M, H, W, generator=g) * 2).cuda()
    sc = (torch.rand(B, K * M, H, W, generator=g) * 2 + 0.5).cuda()
    mu = torch.randn(B, K * M, H, W, generator=g).cuda()
    w = torch.softmax(torch.randn(B, K, M, H, W, generator=g), 1).reshape(B, K * M, H, W).cuda()
    r = torch.randn(B, M, H, W, generator=g).cuda()
    m = fg.GaussianMixtureConditional(K=K, likelihood_bound=0.0).cuda()

    def f(x_, sc_, mu_, w_):
        _, lik = _GMMLikelihoodFn.apply(x_, None, 0, sc_, mu_, w_, K, 0.11, 0.0)
        return (lik.double() * r.double()).sum()

    leaves = [t.clone().requires_grad_(True) for t in (x, sc, mu, w)]
    f(*leaves).backward()
    rng = np.random.default_rng(0)
    for which, leaf in enumerate(leaves):
        flat = leaf.detach().reshape(-1)
        for idx in rng.choice(flat.numel(), 12, replace=False):
            eps = 2e-2
            args_p = [t.detach().clone() for t in leaves]
            args_m = [t.detach().clone() for t in leaves]
            args_p[which].reshape(-1)[idx] += eps
            args_m[which].reshape(-1)[idx] -= eps
            fd = float(f(*args_p) - f(*args_m)) / (2 * eps)
            an = float(leaf.grad.reshape(-1)[idx])
            assert abs(fd - an) <= 2e-3 * max(1.0, abs(an)) + 2e-4, (which, int(idx), fd, an)


def test_entropy_bottleneck_fused_likelihood_matches_tensor_ops():
    """csrc/eb_likelihood.cu against the module's own tensor-op formula (entropy_models.py:421-463): forward within 1e-5
    relative + 2e-7 absolute, gradients equal to autograd through the tensor ops (the backward recomputes them).
    The absolute term is the formula's own fp32 noise: sigmoid(upper) - sigmoid(lower) in the upper tail subtracts two
    numbers next to 1.0, whose last bit is 6e-8 -- two correct implementations differ by that much."""
    torch.manual_seed(4)
    eb = fg.EntropyBottleneck(24).cuda()
    with torch.no_grad():
        for p in eb.parameters():
            p.add_(0.3 * torch.randn_like(p))
    z = (torch.randn(3, 24, 8, 12, device="cuda") * 4).requires_grad_(True)
    for training in (False, True):
        torch.manual_seed(9)
        y1, lik1 = eb(z, training=training)
        perm = z.transpose(0, 1).contiguous().reshape(24, 1, -1)
        torch.manual_seed(9)
        vals = eb.quantize(perm, "noise" if training else "dequantize", eb._get_medians())
        want, _, _ = eb._likelihood(vals)
        want = eb.likelihood_lower_bound(want).reshape(24, 3, 8, 12).transpose(0, 1)
        close(lik1, want.detach().cpu().numpy(), rtol=1e-5, atol=2e-7)
        assert torch.equal(y1, vals.reshape(24, 3, 8, 12).transpose(0, 1))
    # gradients: the backward recomputes the tensor-op formula, so d(sum lik) agrees tightly; d(sum log lik) also carries
    # 1 / lik of the fused forward, which amplifies the formula's absolute noise on tail values (lik ~ 1e-6)
    for reduce, rtol, atol in ((lambda t: t.sum(), 1e-5, 1e-6), (lambda t: t.log().sum(), 5e-3, 1e-3)):
        torch.manual_seed(9)
        _, lik = eb(z, training=True)
        g1 = torch.autograd.grad(reduce(lik), [z] + list(eb.parameters()), allow_unused=True)
        torch.manual_seed(9)
        vals = eb.quantize(z.transpose(0, 1).contiguous().reshape(24, 1, -1), "noise", eb._get_medians())
        want = eb.likelihood_lower_bound(eb._likelihood(vals)[0]).reshape(24, 3, 8, 12).transpose(0, 1)
        g2 = torch.autograd.grad(reduce(want), [z] + list(eb.parameters()), allow_unused=True)
        for a, b in zip(g1, g2):
            if a is None or b is None:
                assert a is None and b is None
            else:
                torch.testing.assert_close(a, b, rtol=rtol, atol=atol)
