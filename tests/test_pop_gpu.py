"""Device-resident decode (batched rows kernel + one-warp-per-stream pop kernel,
flashgmm_b200/csrc/gmm_pop.cu) against the oracle's restatement of
RansDecoder::decode_with_indexes_gmm<4> (rans_interface.cpp:766-883): bit-exact symbols on
valid streams, on garbage streams (every fall-off-the-end case of the reference's search),
on bypass-only streams, on rows that are not monotone, for every interval count the kernel
is instantiated for, across chunk boundaries, and the truncation report."""
import ctypes as C

import numpy as np
import pytest
import torch

from flashgmm_b200 import _lib, synth
from oracle import port

pytestmark = pytest.mark.gpu


@pytest.fixture()
def pop_ctx(monkeypatch):
    """A private ctx whose decode calls always take the device path."""
    made = []

    def make(rows_bytes=None):
        monkeypatch.setenv("FLASHGMM_POP_MIN_STREAMS", "1")
        monkeypatch.setenv("FLASHGMM_POP_POLICY", "device")
        if rows_bytes is not None:
            monkeypatch.setenv("FLASHGMM_POP_ROWS_BYTES", str(rows_bytes))
        h = C.c_void_p()
        _lib.check(_lib.lib().flashgmm_ctx_create(0, C.byref(h)))
        made.append(h)
        return h

    yield make
    for h in made:
        _lib.lib().flashgmm_ctx_destroy(h)


def kmajor(a, device):
    """(N,4) numpy view -> torch (N,4) view with strides (1,N) on `device`."""
    t = torch.from_numpy(np.ascontiguousarray(a.T))
    if device == "cuda":
        t = t.cuda()
    elif torch.cuda.is_available():
        t = t.pin_memory()
    return t.t()


def decode_batch(h, items, mode, device="cuda"):
    """items: (bytes, sg, mu, w, R) -> (list of int32 arrays, per-stream status, rc)."""
    n = len(items)
    arr = (_lib.Stream * max(1, n))()
    keep, outs = [], []
    for i, (b, sg, mu, w, R) in enumerate(items):
        a, m, ww = (kmajor(x, device) for x in (sg, mu, w))
        buf = np.frombuffer(b, np.uint8)
        N = sg.shape[0]
        o = torch.full((max(N, 1),), -77777, dtype=torch.int32, device=device)
        keep += [a, m, ww, buf]
        outs.append(o)
        arr[i].scales, arr[i].means, arr[i].weights = a.data_ptr(), m.data_ptr(), ww.data_ptr()
        arr[i].n, arr[i].stride_n, arr[i].stride_k = N, a.stride(0), a.stride(1)
        arr[i].max_bs_value = R
        arr[i].bytes, arr[i].bytes_cap = buf.ctypes.data, buf.size
        arr[i].symbols_out = o.data_ptr()
    torch.cuda.synchronize()
    rc = _lib.lib().flashgmm_gmm_decode_batch(h, arr, n, mode, _lib.MEM_DEVICE if device == "cuda" else _lib.MEM_HOST)
    res = [o[: it[1].shape[0]].cpu().numpy() for o, it in zip(outs, items)]
    return res, [arr[i].status for i in range(n)], rc


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("device", ["cuda", "cpu"])
def test_elic_shaped_batch_matches_oracle(pop_ctx, mode, device):
    sizes = [12288, 12288, 12288, 12288, 24576, 24576, 49152, 49152, 147456, 147456, 0, 1, 5]
    items, want = [], []
    for i, n in enumerate(sizes):
        sym, sg, mu, w = synth.gmm_numpy(n, seed=300 + i)
        b = port.gmm_encode(sym, sg, mu, w, mode)
        R = synth.search_radius(sym)
        items.append((b, sg, mu, w, R))
        want.append(sym)
    got, status, rc = decode_batch(pop_ctx(), items, mode, device)
    assert rc == 0 and not any(status)
    for g, wsym in zip(got, want):
        assert np.array_equal(g, wsym)


@pytest.mark.parametrize("R", [2, 15, 16, 31, 32, 47, 48, 63, 64, 200])
def test_every_interval_count_and_the_generic_kernel(pop_ctx, R):
    """NI = 1..4 intervals per lane and the any-R kernel; half the streams are garbage so
    that cum falls outside every interval of many rows."""
    rng = np.random.default_rng(R)
    items, want = [], []
    for i in range(4):
        n = 3000 + 17 * i
        sym, sg, mu, w = synth.gmm_numpy(n, seed=400 + 10 * R + i, mean_scale=min(R / 4.0, 12.0), sigma_span=min(R / 6.0, 4.0))
        sym = np.clip(sym, -(R - 1), R - 1).astype(np.int32)
        if i % 2 == 0:
            b = port.gmm_encode(sym, sg, mu, w, 0)
        else:
            b = rng.integers(0, 256, 4 * (12 * n + 16), dtype=np.uint8).tobytes()
        items.append((b, sg, mu, w, R))
        want.append(port.gmm_decode(b, sg, mu, w, R, 0))
        if i % 2 == 0:
            assert np.array_equal(want[-1], sym)
    got, status, rc = decode_batch(pop_ctx(), items, 0)
    assert rc == 0 and not any(status)
    for g, wd in zip(got, want):
        assert np.array_equal(g, wd)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_bypass_only_and_extreme_parameters(pop_ctx, mode):
    r = np.random.default_rng(500 + mode)
    items, want = [], []
    # (a) sigma at the clamp, symbols far from every mean: every symbol is bypassed, incl. int32 extremes
    sym, sg, mu, w = synth.gmm_numpy(6000, seed=501, sigma_span=0.0, offcentre=300.0)
    sym[:8] = [2**31 - 1, -2**31, 0, 1, -1, 65535, -65536, 123456789]
    b = port.gmm_encode(sym, sg, mu, w, mode)
    items.append((b, sg, mu, w, 40))
    want.append(port.gmm_decode(b, sg, mu, w, 40, mode))
    assert np.array_equal(want[-1], sym)
    # (b) parameters over many orders of magnitude (rows that are not monotone show up here), garbage stream
    n = 20000
    mu2 = (r.standard_normal((4, n)) * 10 ** r.uniform(-3, 2, (4, n))).astype(np.float32)
    sg2 = np.clip(10 ** r.uniform(-1, 2.5, (4, n)), 0.11, 256).astype(np.float32)
    w2 = np.ascontiguousarray(r.dirichlet(np.ones(4), n).T, dtype=np.float32)
    w2[:, ::7] *= 1.5                                      # weights that do not sum to one
    b2 = r.integers(0, 256, 4 * (12 * n + 16), dtype=np.uint8).tobytes()
    items.append((b2, sg2.T, mu2.T, w2.T, 30))
    want.append(port.gmm_decode(b2, sg2.T, mu2.T, w2.T, 30, mode))
    got, status, rc = decode_batch(pop_ctx(), items, mode)
    assert rc == 0 and not any(status)
    for g, wd in zip(got, want):
        assert np.array_equal(g, wd)


def test_rows_that_are_not_monotone_replay_the_reference_search(pop_ctx):
    """NaN / negative weights make q16(cdf) go up and down along a row: the flag slot is set and
    the reference's binary search is replayed probe by probe."""
    r = np.random.default_rng(77)
    n = 8000
    sym, sg, mu, w = synth.gmm_numpy(n, seed=78)
    w = np.array(w.T, copy=True)
    w[1, ::3] = -w[1, ::3]                                  # a component with negative weight
    w[2, ::11] = 3.0
    w = w.T
    R = 24
    tab = port.gmm_cdf_table(sg, mu, w, R, 0)
    assert (np.diff(tab.astype(np.int32), axis=1) < 0).any(1).sum() > 100
    b = r.integers(0, 256, 4 * (12 * n + 16), dtype=np.uint8).tobytes()
    want = port.gmm_decode(b, sg, mu, w, R, 0)
    got, status, rc = decode_batch(pop_ctx(), [(b, sg, mu, w, R)], 0)
    assert rc == 0 and not any(status)
    assert np.array_equal(got[0], want)


def test_many_chunks_and_double_buffering(pop_ctx):
    """A 1 MiB rows budget cuts the batch into dozens of chunks: the state carried between
    pop launches and the two alternating rows buffers must not show."""
    items, want = [], []
    for i, n in enumerate([40000, 9000, 25000, 1, 300]):
        sym, sg, mu, w = synth.gmm_numpy(n, seed=600 + i)
        items.append((port.gmm_encode(sym, sg, mu, w, 0), sg, mu, w, synth.search_radius(sym)))
        want.append(sym)
    for device in ("cuda", "cpu"):
        got, status, rc = decode_batch(pop_ctx(rows_bytes=1 << 20), items, 0, device)
        assert rc == 0 and not any(status)
        for g, wd in zip(got, want):
            assert np.array_equal(g, wd)


def test_truncated_stream_reported_per_stream(pop_ctx):
    items = []
    for i in range(3):
        sym, sg, mu, w = synth.gmm_numpy(20000, seed=700 + i)
        b = port.gmm_encode(sym, sg, mu, w, 0)
        if i == 1:
            b = b[: (len(b) // 8) * 4]
        items.append((b, sg, mu, w, synth.search_radius(sym)))
    got, status, rc = decode_batch(pop_ctx(), items, 0)
    assert rc == _lib.ETRUNC
    assert status == [0, _lib.ETRUNC, 0]
    want1, ov = port.gmm_decode(items[1][0], *items[1][1:4], items[1][4], 0, return_overrun=True)
    assert ov and np.array_equal(got[1], want1)          # zeros are fed past the end, like the oracle does


def test_default_ctx_routes_batches_by_the_cost_model():
    """The public batch API on the default ctx: a batch of many long streams decodes on the device (the pop kernel's
    launches show up in the profile), a handful of short streams on the host workers (no pop kernel launch) -- same
    symbols either way."""
    from flashgmm_b200 import batch
    L = _lib.lib()

    def run(sizes, seed):
        data = [synth.gmm_numpy(n, seed=seed + i) for i, n in enumerate(sizes)]
        streams = []
        for sym, sg, mu, w in data:
            b = port.gmm_encode(sym, sg, mu, w, 0)
            streams.append((b, kmajor(sg, "cuda"), kmajor(mu, "cuda"), kmajor(w, "cuda"), synth.search_radius(sym)))
        L.flashgmm_profile_reset()
        L.flashgmm_profile_enable(1)
        outs = batch.decode_batch(streams, 0)
        L.flashgmm_profile_enable(0)
        ms, nl, by = C.c_double(), C.c_uint64(), C.c_double()
        _lib.check(L.flashgmm_profile_read(4, C.byref(ms), C.byref(nl), C.byref(by)))
        for o, (sym, *_r) in zip(outs, data):
            assert np.array_equal(o.cpu().numpy(), sym)
        return nl.value

    assert run([60000] * 32, 800) >= 1          # rows would cost more on the link than the warps' walk: device
    assert run([5000 + i for i in range(6)], 850) == 0   # six short streams: host workers


@pytest.mark.parametrize("n_streams", [1, 6])
def test_pinned_host_tensors_are_read_in_place(n_streams):
    """Page-locked host inputs take the zero-copy path (kernels read them over PCIe): same bytes,
    same symbols as the oracle, for the host-pop path (1 stream) and the device path (6 streams)."""
    from flashgmm_b200 import batch
    data = [synth.gmm_numpy(30_000 + 1000 * i + 3, seed=900 + i) for i in range(n_streams)]
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    enc_in = [(pin(sym), kmajor(sg, "cpu"), kmajor(mu, "cpu"), kmajor(w, "cpu")) for sym, sg, mu, w in data]
    assert all(t.is_pinned() for st in enc_in for t in (st[0], st[1].t()))
    got = batch.encode_batch(enc_in, 0)
    want = [port.gmm_encode(sym, sg, mu, w, 0) for sym, sg, mu, w in data]
    assert got == want
    outs = batch.decode_batch([(b, st[1], st[2], st[3], synth.search_radius(d[0])) for b, st, d in zip(want, enc_in, data)], 0)
    for o, d in zip(outs, data):
        assert o.device.type == "cpu" and np.array_equal(o.numpy(), d[0])


def test_more_length_classes_than_pop_groups(pop_ctx):
    """Stream lengths over four orders of magnitude: more length classes than the decoder has pop
    streams (the shortest classes share the last group), ragged lengths inside a group."""
    sizes = [3, 9, 40, 130, 500, 1100, 2300, 5000, 5100, 9000, 20000, 41000, 90000, 7, 2400]
    items, want = [], []
    for i, n in enumerate(sizes):
        sym, sg, mu, w = synth.gmm_numpy(n, seed=2000 + i)
        items.append((port.gmm_encode(sym, sg, mu, w, 0), sg, mu, w, synth.search_radius(sym)))
        want.append(sym)
    for device in ("cuda", "cpu"):
        got, status, rc = decode_batch(pop_ctx(), items, 0, device)
        assert rc == 0 and not any(status)
        for g, wd in zip(got, want):
            assert np.array_equal(g, wd)


def test_safe_and_fast_pop_variants_agree():
    """the shipped pop loop relies on the LSU keeping one warp's shared-memory post and read-back
    in program order. The __syncwarp() variant (FLASHGMM_POP_SAFE=1, read once per process) must give the same
    symbols on the whole pop test set -- run it in a child process and let the same oracle comparisons judge it."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    if os.environ.get("FLASHGMM_POP_SAFE"):
        pytest.skip("already the safe variant")
    env = dict(os.environ, FLASHGMM_POP_SAFE="1")
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-q", "-x", "-m", "gpu", "-k",
                        "elic_shaped or every_interval or bypass_only or not_monotone or many_chunks or truncated"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " passed" in r.stdout


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_peaky_mixtures_with_saturated_boundaries(pop_ctx, mode):
    """Small sigmas (what trained latents look like): most of the 33 window boundaries are saturated and the rows kernel
    takes them from the closed form (0 below, q16(sum of weights) above) instead of evaluating Phi. Valid AND garbage
    streams (garbage makes cum land in every interval, saturated ones included), large search radius, components both
    clustered and far apart, and scales right at the 0.11 clamp."""
    rng = np.random.default_rng(50 + mode)
    items, want = [], []
    for i, (span, mscale, R) in enumerate([(0.4, 0.7, 40), (1.2, 2.0, 200), (0.0, 6.0, 90), (2.5, 0.2, 1000)]):
        n = 6000 + 11 * i
        sym, sg, mu, w = synth.gmm_numpy(n, seed=900 + 10 * mode + i, sigma_span=span, mean_scale=mscale)
        sym = np.clip(sym, -(R - 1), R - 1).astype(np.int32)
        for garbage in (False, True):
            b = (rng.integers(0, 256, 4 * (12 * n + 16), dtype=np.uint8).tobytes() if garbage
                 else port.gmm_encode(sym, sg, mu, w, mode))
            items.append((b, sg, mu, w, R))
            want.append(port.gmm_decode(b, sg, mu, w, R, mode))
            if not garbage:
                assert np.array_equal(want[-1], sym)
    got, status, rc = decode_batch(pop_ctx(), items, mode)
    assert rc == 0 and not any(status)
    for g, wd in zip(got, want):
        assert np.array_equal(g, wd)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_parameter_corners_encode_and_decode_like_the_oracle(pop_ctx, mode):
    """tests/helpers.extreme_params: scales far outside the model's clamp, one-hot / vanishing / unnormalised weights,
    means on interval boundaries, off-model symbols, a search radius of ~20 000 -- the corners on which the oracle is
    pinned to the compiled reference in tests/test_oracle.py. Bytes and symbols must be the oracle's."""
    from flashgmm_b200 import batch
    from helpers import extreme_params
    sets = [extreme_params(300 + 10 * mode + k, n=12000) for k in range(2)]
    want_bytes = [port.gmm_encode(sym, sg, mu, w, mode) for sym, sg, mu, w in sets]
    streams = [(torch.from_numpy(sym).cuda(), kmajor(sg, "cuda"), kmajor(mu, "cuda"), kmajor(w, "cuda")) for sym, sg, mu, w in sets]
    got_bytes = batch.encode_batch(streams, mode)
    assert got_bytes == want_bytes
    items = [(b, sg, mu, w, synth.search_radius(sym)) for b, (sym, sg, mu, w) in zip(want_bytes, sets)]
    want = [port.gmm_decode(b, sg, mu, w, R, mode) for b, sg, mu, w, R in items]
    got, status, rc = decode_batch(pop_ctx(), items, mode)
    assert rc == 0 and not any(status)
    for g, wd in zip(got, want):
        assert np.array_equal(g, wd)
