"""Parity tests proper: the CUDA path, called through the C ABI, against the oracle
(oracle/gmm_oracle.c) on the same seeded inputs, against the committed golden
vectors from the reference's own extension, and -- at full benchmark size --
through size-independent properties. Integer/byte work: bit-exact."""
import ctypes as C
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from flashgmm_b200 import _lib, ans, synth
from oracle import port, ref
from helpers import lut_offset, make_rows

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"
MD5 = json.loads((GOLD / "gmm_md5.json").read_text())


def dev(*arrs):
    """numpy (N,K) views with strides (1,N) -> cuda tensors with the same strides."""
    out = []
    for a in arrs:
        if a.ndim == 2:
            out.append(torch.from_numpy(np.ascontiguousarray(a.T)).cuda().t())
        else:
            out.append(torch.from_numpy(np.ascontiguousarray(a)).cuda())
    return out


def startfreq_gpu(sym, sg, mu, w, mode):
    s, a, m, ww = dev(sym, sg, mu, w)
    out = torch.empty(max(s.numel(), 1), dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib().flashgmm_gmm_startfreq(
        _lib.ctx(0), s.data_ptr(), a.data_ptr(), m.data_ptr(), ww.data_ptr(), s.numel(),
        a.stride(0), a.stride(1), mode, out.data_ptr(), None))
    return out[: s.numel()].cpu().numpy().view(np.uint32)


def oracle_packed(sym, sg, mu, w, mode):
    st, fr, bp = port.gmm_startfreq(sym, sg, mu, w, mode)
    p = st.astype(np.uint32) | (fr.astype(np.uint32) << 16)
    p[bp == 1] = 0xFFFF
    return p


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("kw", [dict(n=200_000, seed=41), dict(n=50_000, seed=42, sigma_span=0.0),
                                dict(n=50_000, seed=43, mean_scale=40.0),
                                dict(n=50_000, seed=44, sigma_span=0.05, offcentre=9.0),
                                dict(n=4099, seed=45), dict(n=3, seed=46), dict(n=1, seed=47)],
                         ids=lambda k: f"n{k['n']}s{k['seed']}")
def test_startfreq_kernel_bit_exact(kw, mode):
    sym, sg, mu, w = synth.gmm_numpy(**kw)
    got = startfreq_gpu(sym, sg, mu, w, mode)
    want = oracle_packed(sym, sg, mu, w, mode)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0, (bad[:5], got[bad[:5]], want[bad[:5]])


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_startfreq_extreme_arguments(mode):
    """|t| from 0 to 1e6, exact zeros, sigma at both clamps, huge symbols."""
    r = np.random.default_rng(50 + mode)
    n = 120_000
    mu = (r.standard_normal((4, n)) * 10 ** r.uniform(-3, 3, (4, n))).astype(np.float32)
    sg = np.clip(10 ** r.uniform(-1, 2.5, (4, n)), 0.11, 256).astype(np.float32)
    w = r.dirichlet(np.ones(4), n).T.astype(np.float32)
    sym = np.round(mu[0] + r.standard_normal(n) * sg[0] * 10 ** r.uniform(-2, 1, n)).astype(np.int64)
    sym[:1000] = r.integers(-2**31, 2**31, 1000)
    sym = np.clip(sym, -2**31, 2**31 - 1).astype(np.int32)
    mu[:, 1000:1100] = sym[1000:1100] - 0.5          # t == 0 exactly at x1
    mu[:, 1100:1200] = sym[1100:1200] + 0.5
    sg, mu, w = (np.ascontiguousarray(x).T for x in (sg, mu, w))
    got = startfreq_gpu(sym, sg, mu, w, mode)
    want = oracle_packed(sym, sg, mu, w, mode)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("layout", ["kmajor", "aos", "strided", "unaligned"])
def test_startfreq_layouts(layout):
    sym, sg, mu, w = synth.gmm_numpy(10_001, seed=48)
    want = oracle_packed(sym, sg, mu, w, 0)
    s = torch.from_numpy(sym).cuda()
    P = [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (sg, mu, w)]   # (N,4) row-major
    if layout == "kmajor":
        P = [p.t().contiguous().t() for p in P]
    elif layout == "strided":
        P = [torch.cat([p, p], 1).repeat_interleave(2, 0)[::2, :4] for p in P]    # strides (16, 1)
    elif layout == "unaligned":
        big = [torch.zeros(4 * 10_001 + 1, device="cuda") for _ in P]
        for b, p in zip(big, P):
            b[1:] = p.t().reshape(-1)
        P = [b[1:].view(4, -1).t() for b in big]                                  # K-major, 4-byte aligned only
    out = torch.empty(s.numel(), dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib().flashgmm_gmm_startfreq(_lib.ctx(0), s.data_ptr(), P[0].data_ptr(), P[1].data_ptr(),
                                                 P[2].data_ptr(), s.numel(), P[0].stride(0), P[0].stride(1), 0,
                                                 out.data_ptr(), None))
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want)


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("R", [1, 5, 27, 31, 32, 200])
def test_cdf_table_kernel_bit_exact(R, mode):
    sym, sg, mu, w = synth.gmm_numpy(3000 if R < 100 else 700, seed=60 + R)
    _, a, m, ww = dev(sym, sg, mu, w)
    L = _lib.lib()
    rs = L.flashgmm_gmm_table_row_stride(R)
    n = a.size(0)
    tab = torch.full((n, rs), 0x7777, dtype=torch.int16, device="cuda")
    _lib.check(L.flashgmm_gmm_cdf_table(_lib.ctx(0), a.data_ptr(), m.data_ptr(), ww.data_ptr(), n, a.stride(0),
                                        a.stride(1), mode, R, tab.data_ptr(), rs, None))
    got = tab.cpu().numpy().view(np.uint16)
    want = port.gmm_cdf_table(sg, mu, w, R, mode)
    assert np.array_equal(got[:, : 2 * R + 2], want)
    rows = make_rows(want, R)
    nonmono = rows[:, 2 * R + 2] != 0
    assert np.array_equal(got[:, 2 * R + 2] != 0, nonmono)
    assert (got[:, 2 * R + 3: lut_offset(R)] == 0).all()
    assert np.array_equal(got[~nonmono, lut_offset(R):], rows[~nonmono, lut_offset(R):])   # jump tables


@pytest.mark.parametrize("key", sorted(MD5))
@pytest.mark.parametrize("where", ["host", "device"])
def test_encode_decode_match_reference_golden(key, where):
    g = MD5[key]
    mode = int(key[-1])
    sym, sg, mu, w = synth.gmm_numpy(**g["synth"])
    if where == "device":
        s, a, m, ww = dev(sym, sg, mu, w)
    else:
        s, a, m, ww = (torch.from_numpy(np.ascontiguousarray(sym)),) + tuple(
            torch.from_numpy(np.ascontiguousarray(x.T)).t() for x in (sg, mu, w))
    b = ans.gmm_encode(s, a, m, ww, mode)
    assert len(b) == g["nbytes"] and hashlib.md5(b).hexdigest() == g["bytes_md5"]
    d = ans.gmm_decode(b, a, m, ww, g["R"], mode)
    assert d.device.type == ("cuda" if where == "device" else "cpu")
    assert hashlib.md5(d.cpu().numpy().tobytes()).hexdigest() == g["decoded_md5"]


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_golden_small_vectors_with_int32_extremes(mode):
    z = np.load(GOLD / "gmm_small.npz")
    sym, sg, mu, w = (torch.from_numpy(z[k]).cuda() for k in ("symbols", "scales", "means", "weights"))
    b = ans.gmm_encode(sym, sg, mu, w, mode)
    assert b == z[f"bytes_mode{mode}"].tobytes()
    d = ans.gmm_decode(b, sg, mu, w, int(z["R"]), mode)
    assert np.array_equal(d.cpu().numpy(), z[f"decoded_mode{mode}"])


def test_chunked_pipeline_equals_single_chunk(monkeypatch):
    """Several chunks per stream (reverse-order encode, forward decode) == oracle bytes."""
    sym, sg, mu, w = synth.gmm_numpy(70_001, seed=71, sigma_span=1.0, offcentre=1.5)
    want = port.gmm_encode(sym, sg, mu, w, 0)
    R = synth.search_radius(sym)
    monkeypatch.setenv("FLASHGMM_ENC_CHUNK", "8192")
    monkeypatch.setenv("FLASHGMM_DEC_CHUNK_BYTES", "65536")
    L = _lib.lib()
    h = C.c_void_p()
    _lib.check(L.flashgmm_ctx_create(0, C.byref(h)))
    try:
        for mem in (_lib.MEM_HOST, _lib.MEM_DEVICE):
            if mem == _lib.MEM_DEVICE:
                s, a, m, ww = dev(sym, sg, mu, w)
            else:
                s, a, m, ww = (torch.from_numpy(sym),) + tuple(torch.from_numpy(np.ascontiguousarray(x)) for x in (sg, mu, w))
            out = np.empty(len(want) + 64, np.uint8)
            n_out = C.c_size_t()
            _lib.check(L.flashgmm_gmm_encode(h, s.data_ptr(), a.data_ptr(), m.data_ptr(), ww.data_ptr(), s.numel(),
                                             a.stride(0), a.stride(1), 0, mem, out.ctypes.data, out.size, C.byref(n_out)))
            assert out[: n_out.value].tobytes() == want
            dec = torch.empty_like(s)
            _lib.check(L.flashgmm_gmm_decode(h, want, len(want), a.data_ptr(), m.data_ptr(), ww.data_ptr(), s.numel(),
                                             a.stride(0), a.stride(1), 0, mem, R, dec.data_ptr()))
            assert np.array_equal(dec.cpu().numpy(), sym)
    finally:
        L.flashgmm_ctx_destroy(h)


def test_batch_of_streams_matches_oracle():
    """ELIC-shaped: 10 streams of different lengths in one call."""
    sizes = [12288, 12288, 12288, 12288, 24576, 24576, 49152, 49152, 147456, 147456]
    data = [synth.gmm_numpy(n, seed=80 + i) for i, n in enumerate(sizes)]
    L = _lib.lib()
    arr = (_lib.Stream * len(sizes))()
    keep = []
    for i, (sym, sg, mu, w) in enumerate(data):
        s, a, m, ww = dev(sym, sg, mu, w)
        out = np.empty(sym.size * 3 + 64, np.uint8)
        keep += [s, a, m, ww, out]
        arr[i].symbols, arr[i].scales, arr[i].means, arr[i].weights = s.data_ptr(), a.data_ptr(), m.data_ptr(), ww.data_ptr()
        arr[i].n, arr[i].stride_n, arr[i].stride_k = sym.size, a.stride(0), a.stride(1)
        arr[i].bytes, arr[i].bytes_cap = out.ctypes.data, out.size
    torch.cuda.synchronize()
    _lib.check(L.flashgmm_gmm_encode_batch(_lib.ctx(0), arr, len(sizes), 0, _lib.MEM_DEVICE))
    encs = []
    for i, (sym, sg, mu, w) in enumerate(data):
        b = keep[5 * i + 4][: arr[i].bytes_len].tobytes()
        assert arr[i].status == 0 and b == port.gmm_encode(sym, sg, mu, w, 0)
        encs.append(np.frombuffer(b, np.uint8))
    outs = []
    for i, (sym, sg, mu, w) in enumerate(data):
        o = torch.empty(sym.size, dtype=torch.int32, device="cuda")
        outs.append(o)
        arr[i].bytes, arr[i].bytes_cap = encs[i].ctypes.data, encs[i].size
        arr[i].symbols_out = o.data_ptr()
        arr[i].max_bs_value = synth.search_radius(sym)
    _lib.check(L.flashgmm_gmm_decode_batch(_lib.ctx(0), arr, len(sizes), 0, _lib.MEM_DEVICE))
    for o, (sym, *_r) in zip(outs, data):
        assert np.array_equal(o.cpu().numpy(), sym)


def test_against_compiled_reference_live():
    """When oracle/_ref travelled to this box: reference extension vs CUDA path directly."""
    if not ref.available():
        pytest.skip("oracle/_ref not present")
    sym, sg, mu, w = synth.gmm_numpy(147_456, seed=90)
    R = synth.search_radius(sym)
    s, a, m, ww = dev(sym, sg, mu, w)
    for mode in (0, 1, 2):
        with ref.RefWorker(mode) as wk:
            b, _ = wk.gmm_encode(sym, sg, mu, w, R)
            d, _ = wk.gmm_decode(b, sg, mu, w, R)
        assert ans.gmm_encode(s, a, m, ww, mode) == b
        assert np.array_equal(ans.gmm_decode(b, a, m, ww, R, mode).cpu().numpy(), d)


def test_full_size_microbench_properties():
    """16 Mi symbols (BASELINE.json config 4): round trip, chunk-order independence of
    the packed records (checksum of checksums), and oracle agreement on a random sample."""
    n = 1 << 24
    sym, sg, mu, w = synth.gmm_torch(n, seed=1234)
    R = int(sym.abs().max().item()) + 2
    b = ans.gmm_encode(sym, sg, mu, w, 0)
    assert 3.0 < len(b) * 8 / n < 4.2                      # ~3.5 bit/symbol
    d = ans.gmm_decode(b, sg, mu, w, R, 0)
    assert torch.equal(d, sym)
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    _lib.check(_lib.lib().flashgmm_gmm_startfreq(_lib.ctx(0), sym.data_ptr(), sg.data_ptr(), mu.data_ptr(),
                                                 w.data_ptr(), n, sg.stride(0), sg.stride(1), 0, out.data_ptr(), None))
    idx = torch.randint(0, n, (50_000,), device="cuda")
    pick = lambda t: np.ascontiguousarray(t[idx].cpu().numpy())
    want = oracle_packed(pick(sym), pick(sg), pick(mu), pick(w), 0)
    assert np.array_equal(out[idx].cpu().numpy().view(np.uint32), want)


def test_truncated_stream_is_reported():
    sym, sg, mu, w = synth.gmm_numpy(20_000, seed=95)
    s, a, m, ww = dev(sym, sg, mu, w)
    b = ans.gmm_encode(s, a, m, ww, 0)
    with pytest.raises(_lib.FlashGMMError) as e:
        ans.gmm_decode(b[: len(b) // 2], a, m, ww, synth.search_radius(sym), 0)
    assert e.value.code == _lib.ETRUNC
