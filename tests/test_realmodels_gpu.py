"""BASELINE.json configs 1-3 through the MODELS: Cheng2020AnchorCheckerboardGMMv2 (2 passes) and
Elic2022GMM (10 passes), compress + pass-sequential decompress on the GPU, every stream compared with
the compiled reference (``oracle/_ref``) run on the very tensors the device coder saw.

The architectures are built by ``tests/gmm_models.py`` (the reference's Python package cannot travel to the
GPU box); ``tests/test_model_cpu.py`` proves, where the reference is present, that they have the reference's
state_dict keys and bit-identical seeded weights. Protocol of SURVEY.md 8(d) configs 1/2: seed 0, random-init
weights, ``x = rand(1, 3, 512, 768)``, latent gain x20 (without it every latent rounds to 0 and nothing is coded),
``net.update()``; plus a "plausible parameters" variant in which every pass's entropy parameters are replaced by
the synthetic mixture of config 4 and y is drawn from it (random-init networks clamp sigma to 0.11: ~100 % bypass).
"""
import numpy as np
import pytest
import torch

import flashgmm_b200 as fg
from flashgmm_b200 import ans
from flashgmm_b200.graphs import NetRunner

import gmm_models as gm

pytestmark = pytest.mark.gpu

GAIN = 20.0
MODELS = {
    "ckbd": (gm.Cheng2020AnchorCheckerboardGMMv2, dict(N=192, K=4), 2),
    "elic": (gm.Elic2022GMM, dict(N=192, M=320, K=4), 10),
}
_cache = {}


def build(name):
    if name not in _cache:
        cls, kw, _ = MODELS[name]
        torch.manual_seed(0)
        net = cls(**kw).eval()
        net.update()
        _cache[name] = net.cuda()
    return _cache[name]


def gmcs(net):
    return [m for m in net.modules() if isinstance(m, fg.GaussianMixtureConditional)]


def latent(net, B=1, seed=0, hw=(512, 768)):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 3, *hw, generator=g).cuda()
    with torch.no_grad():
        return GAIN * net.g_a(x)


def to_np(stream):
    sym, sc, mu, w = stream
    f = lambda t: np.ascontiguousarray(t.detach().cpu().numpy())  # noqa: E731
    return (f(sym) if sym is not None else None), f(sc), f(mu), f(w)


def start_trace(net):
    trace = []
    for m in gmcs(net):
        m.trace = trace
    return trace


def stop_trace(net):
    for m in gmcs(net):
        m.trace = None


def flat(strings):
    """per-pass entries of `strings` (without the z entry) as a list of lists of tuples [pass][image]."""
    return [[s] if isinstance(s, tuple) else list(s) for s in strings[:-1]]


@pytest.fixture(scope="module")
def refs():
    from oracle import ref

    if not ref.available():
        pytest.skip("compiled reference (oracle/_ref) not present")
    workers = {}

    def get(mode):
        if mode not in workers:
            workers[mode] = ref.RefWorker(mode=mode)
        return workers[mode]
    yield get
    for w in workers.values():
        w.close()


def check_against_reference(trace, strings_by_pass, worker, kind):
    """every traced stream, replayed through the compiled reference on the same tensors."""
    n_checked = 0
    calls = [t for t in trace if t[0] == kind]
    assert len(calls) == len(strings_by_pass), (len(calls), len(strings_by_pass))
    for (_, per_image, abs_max), items in zip(calls, strings_by_pass):
        assert len(per_image) == len(items)
        for b, (stream, item) in enumerate(zip(per_image, items)):
            sym, sc, mu, w = to_np(stream)
            assert int(item[1]) == abs_max[b]
            if sc.shape[0] == 0:
                assert item[0] == bytes.fromhex("0000008000000000")
                continue
            if kind == "enc":
                want, _dt = worker.gmm_encode(sym, sc, mu, w)
                assert item[0] == want, f"stream of image {b}: bytes differ from the compiled reference"
            else:
                want, _dt = worker.gmm_decode(item[0], sc, mu, w, abs_max[b] + 1)
                assert np.array_equal(sym, want), f"image {b}: decoded symbols differ from the compiled reference"
            n_checked += 1
    return n_checked


# ---- configs 1, 2, 3: the model flow, all three approximation modes ------------------------------------
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("name", ["ckbd", "elic"])
def test_model_streams_byte_identical_and_latents_bit_exact(name, mode, refs):
    net = build(name)
    n_pass = MODELS[name][2]
    ans.set_approx_mode(mode)
    try:
        y = latent(net)
        trace = start_trace(net)
        enc = net.latent_codec.compress(y)
        dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
    finally:
        stop_trace(net)
        ans.set_approx_mode(0)
    strings = enc["strings"]
    assert len(strings) == n_pass + 1                                   # + the z stream
    by_pass = flat(strings)
    assert all(isinstance(s, tuple) and isinstance(s[0], bytes) for s in strings[:-1])   # B = 1: the reference's structure
    assert check_against_reference(trace, by_pass, refs(mode), "enc") == n_pass
    assert check_against_reference(trace, by_pass, refs(mode), "dec") == n_pass
    assert torch.equal(dec["y_hat"], enc["y_hat"])                      # decoder reproduces the encoder's latent
    assert torch.equal(enc["y_hat"], torch.round(y))                    # ... which is round(y) for the "noise" quantiser
    coded = sum(int(st[0].numel()) for t in trace if t[0] == "enc" for st in t[1])
    assert coded == y.numel(), "the x20 latent gain keeps every channel"


@pytest.mark.parametrize("name", ["ckbd", "elic"])
def test_hyper_stream_matches_reference_table_coder(name, refs):
    net = build(name)
    y = latent(net)
    hyper = net.latent_codec.latent_codec["hyper"]
    eb = hyper.entropy_bottleneck
    with torch.no_grad():
        z = hyper.h_a(y)
        out = hyper.compress(y)
    (z_strings,) = out["strings"]
    med = eb._get_medians().detach().reshape(1, -1, 1, 1)
    sym = torch.round(z - med).int()
    idx = eb._build_indexes(z.size())
    want = refs(0).table_encode(sym.reshape(-1).cpu().numpy(), idx.reshape(-1).numpy(),
                                eb._quantized_cdf.cpu().numpy(), eb._cdf_length.cpu().numpy(), eb._offset.cpu().numpy())
    assert z_strings[0] == want
    z_hat = eb.decompress(z_strings, z.shape[-2:])
    assert torch.equal(z_hat, sym.float() + med)
    got = refs(0).table_decode(want, idx.reshape(-1).numpy(), eb._quantized_cdf.cpu().numpy(),
                               eb._cdf_length.cpu().numpy(), eb._offset.cpu().numpy())
    assert np.array_equal(np.asarray(got, np.int32), sym.reshape(-1).cpu().numpy())


def test_hyper_bypass_symbols_roundtrip(refs):
    """values outside the table's support take the sentinel + bypass nibbles (rans_interface.cpp:356-397)."""
    eb = build("ckbd").latent_codec.latent_codec["hyper"].entropy_bottleneck
    g = torch.Generator().manual_seed(5)
    z = (torch.randn(3, eb.channels, 8, 12, generator=g) * 40).cuda()
    z[0, 0, 0, 0], z[1, 3, 2, 2], z[2, 5, 7, 11] = 70000.0, -123456.0, 0.0
    strings = eb.compress(z)
    med = eb._get_medians().detach().reshape(1, -1, 1, 1)
    sym = torch.round(z - med).int()
    idx = eb._build_indexes(z.size())
    for b in range(3):
        want = refs(0).table_encode(sym[b].reshape(-1).cpu().numpy(), idx[b].reshape(-1).numpy(),
                                    eb._quantized_cdf.cpu().numpy(), eb._cdf_length.cpu().numpy(), eb._offset.cpu().numpy())
        assert strings[b] == want
    assert torch.equal(eb.decompress(strings, (8, 12)), sym.float() + med)


# ---- batches: image b of a batch codes exactly like image b alone --------------------------------------
@pytest.mark.parametrize("name,B", [("ckbd", 3), ("elic", 2)])
def test_batch_streams_equal_single_image_streams(name, B):
    net = build(name)
    n_pass = MODELS[name][2]
    y = latent(net, B=B, seed=11)
    enc = net.latent_codec.compress(y)
    assert len(enc["strings"]) == n_pass + 1 and all(len(s) == B for s in enc["strings"])
    dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
    assert torch.equal(dec["y_hat"], enc["y_hat"]) and torch.equal(enc["y_hat"], torch.round(y))
    # through the file container (flashgmm_b200/container.py): bytes -> the same latents
    from flashgmm_b200 import container
    blob = container.pack(enc["strings"], enc["shape"])
    strings2, shape2, info = container.unpack(blob)
    assert info["images"] == B and info["passes"] == n_pass
    assert torch.equal(net.latent_codec.decompress(strings2, shape2)["y_hat"], enc["y_hat"])
    for b in range(B):
        one = net.latent_codec.compress(y[b:b + 1].contiguous())
        for p in range(n_pass):
            got, want = enc["strings"][p][b], one["strings"][p]
            assert got[0] == want[0] and got[1] == want[1] and torch.equal(got[2], want[2]), (b, p)
        assert enc["strings"][-1][b] == one["strings"][-1][0]
        assert torch.equal(one["y_hat"], enc["y_hat"][b:b + 1])


def test_graph_replay_equals_eager():
    net = build("elic")
    y = latent(net, B=2, seed=3, hw=(256, 256))
    outs = []
    for graphs in (True, False):
        runner = NetRunner(mode="per_image", graphs=graphs)
        fg.mount(net, runner=runner)
        enc = net.latent_codec.compress(y)
        dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
        assert torch.equal(dec["y_hat"], enc["y_hat"])
        outs.append(enc)
        if graphs:
            assert runner.captures > 0 and runner.replays > runner.captures
    fg.mount(net, runner=None)
    for m in net.modules():
        if hasattr(m, "_net_runner"):
            del m._net_runner
    for a, b in zip(outs[0]["strings"][:-1], outs[1]["strings"][:-1]):
        assert [x[0] for x in a] == [x[0] for x in b]


def test_lanes_two_threads_on_one_model_give_the_single_thread_bytes():
    """flashgmm_b200.Lane: two host threads code different batches through the SAME model concurrently (own ctx, own
    graphs, own stream each); every stream equals what the plain single-thread call produces."""
    import threading
    net = build("elic")
    ys = [latent(net, B=2, seed=21 + k, hw=(256, 384)) for k in range(2)]
    want = [net.latent_codec.compress(y) for y in ys]
    torch.cuda.synchronize()
    lanes = [fg.Lane(device=0) for _ in range(2)]
    for lane, y in zip(lanes, ys):                            # graphs are captured with the lane running alone
        with lane:
            net.latent_codec.compress(y)
    got, errors = [None, None], []
    gate = threading.Barrier(2)

    def work(k):
        try:
            with lanes[k]:
                gate.wait()
                for _ in range(3):
                    enc = net.latent_codec.compress(ys[k])
                    dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
                    if not torch.equal(dec["y_hat"], enc["y_hat"]):
                        errors.append((k, "round trip"))
                got[k] = enc
        except Exception as e:                                # noqa: BLE001
            errors.append((k, repr(e)))
            gate.abort()

    threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    [t.start() for t in threads]
    [t.join() for t in threads]
    assert not errors, errors
    for k in range(2):
        assert lanes[k].runner.replays > 0
        assert torch.equal(got[k]["y_hat"], want[k]["y_hat"])
        for a, b in zip(got[k]["strings"][:-1], want[k]["strings"][:-1]):
            assert [x[0] for x in a] == [x[0] for x in b] and [x[1] for x in a] == [x[1] for x in b]
        assert got[k]["strings"][-1] == want[k]["strings"][-1]


# ---- plausible parameters: the synthetic mixture of config 4 in place of the random-init networks' output ----
from gmm_models import SyntheticMixtures  # noqa: E402


@pytest.mark.parametrize("name", ["ckbd", "elic"])
def test_plausible_parameters_through_the_model_flow(name, refs):
    net = build(name)
    n_pass = MODELS[name][2]
    ycodec = net.latent_codec.latent_codec["y"]
    groups = [ycodec] if name == "ckbd" else [ycodec.latent_codec[f"y{k}"] for k in range(len(ycodec.groups))]
    widths = [192] if name == "ckbd" else list(ycodec.groups)
    H, W = 32, 48
    syn = SyntheticMixtures(net)
    syn.channels = widths
    # y: every (group, pass) half drawn from its own mixture, halves joined on the checkerboard
    parts = []
    for i, (codec, C) in enumerate(zip(groups, widths)):
        halves = [syn.params(i, p, 1, C, H, W // 2, 4, "cuda")[3] for p in range(2)]
        parts.append(codec._join(halves[0], halves[1]))
    y = torch.cat(parts, dim=1)
    with syn:
        trace = start_trace(net)
        try:
            enc = net.latent_codec.compress(y)
            syn.calls.clear()
            dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
        finally:
            stop_trace(net)
    by_pass = flat(enc["strings"])
    assert check_against_reference(trace, by_pass, refs(0), "enc") == n_pass
    assert check_against_reference(trace, by_pass, refs(0), "dec") == n_pass
    assert torch.equal(dec["y_hat"], torch.round(y))
    n_sym = y.numel()
    bits = 8 * sum(len(s[0]) for s in enc["strings"][:-1])
    assert 2.5 < bits / n_sym < 4.5, f"{bits / n_sym:.2f} bit/symbol: the synthetic mixture codes at ~3.5"


# ---- checkpoints ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["ckbd", "elic"])
@pytest.mark.parametrize("updated", [False, True])
def test_state_dicts_load_strict_and_roundtrip(name, updated):
    cls, kw, _ = MODELS[name]
    torch.manual_seed(0)
    src = cls(**kw).eval()
    if updated:
        src.update()
    sd = {k: v.clone() for k, v in src.state_dict().items()}
    torch.manual_seed(123)                                              # different init: the load must overwrite it
    dst = cls(**kw).eval()
    dst.load_state_dict(sd, strict=True)
    if not updated:
        dst.update()
    dst = dst.cuda()
    y = latent(dst, hw=(256, 256))
    enc = dst.latent_codec.compress(y)
    dec = dst.latent_codec.decompress(enc["strings"], enc["shape"])
    assert torch.equal(dec["y_hat"], torch.round(y))
    ref_net = build(name)                                               # same seed-0 weights: same streams
    enc2 = ref_net.latent_codec.compress(y)
    assert [s[0] for s in enc["strings"][:-1]] == [s[0] for s in enc2["strings"][:-1]]
    assert enc["strings"][-1] == enc2["strings"][-1]


def test_model_flow_on_the_device_pop_path_too():
    """Small batches decode on the host workers by the cost model (csrc/flashgmm_abi.cu::want_device_pop); the same model
    tests must hold with every batch forced onto the rows + pop kernels (FLASHGMM_POP_POLICY is read at ctx creation:
    child process)."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    if os.environ.get("FLASHGMM_POP_POLICY"):
        pytest.skip("already forced")
    env = dict(os.environ, FLASHGMM_POP_POLICY="device")
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-q", "-x", "-m", "gpu", "-k",
                        "batch_streams_equal or plausible_parameters"], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and " passed" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
