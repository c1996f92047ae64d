"""The CPU restatement (oracle/gmm_oracle.c) is pinned to the reference:
 (1) against golden vectors produced by the reference's compiled extension
     (tests/golden/make_golden.py), (2) directly against that extension when
     oracle/_ref is present, (3) against closed-form sanity properties."""
import hashlib
import json
from pathlib import Path

import numpy as np
import pytest

from flashgmm_b200 import synth
from oracle import port, ref
from helpers import extreme_params

GOLD = Path(__file__).resolve().parent / "golden"
MD5 = json.loads((GOLD / "gmm_md5.json").read_text())


@pytest.mark.parametrize("key", sorted(MD5))
def test_port_matches_golden_md5(key):
    g = MD5[key]
    mode = int(key[-1])
    sym, sg, mu, w = synth.gmm_numpy(**g["synth"])
    b = port.gmm_encode(sym, sg, mu, w, mode)
    assert len(b) == g["nbytes"]
    assert hashlib.md5(b).hexdigest() == g["bytes_md5"]
    d = port.gmm_decode(b, sg, mu, w, g["R"], mode)
    assert hashlib.md5(d.tobytes()).hexdigest() == g["decoded_md5"]
    assert np.array_equal(d, sym) == g["roundtrip"]


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_port_matches_golden_small(mode):
    z = np.load(GOLD / "gmm_small.npz")
    sym, sg, mu, w = z["symbols"], z["scales"], z["means"], z["weights"]
    b = port.gmm_encode(sym, sg, mu, w, mode)
    assert b == z[f"bytes_mode{mode}"].tobytes()
    d = port.gmm_decode(b, sg, mu, w, int(z["R"]), mode)
    assert np.array_equal(d, z[f"decoded_mode{mode}"])
    assert np.array_equal(d, sym)           # incl. INT32_MIN/MAX through the bypass nibbles


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_port_matches_compiled_reference(mode, ref_available):
    if not ref_available:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    with ref.RefWorker(mode) as wk:
        for seed, kw in [(11, {}), (12, dict(sigma_span=0.0)), (13, dict(mean_scale=30.0)),
                         (14, dict(sigma_span=0.02, offcentre=5.0))]:
            sym, sg, mu, w = synth.gmm_numpy(30_000, seed=seed, **kw)
            R = synth.search_radius(sym)
            b, _ = wk.gmm_encode(sym, sg, mu, w, R)
            assert port.gmm_encode(sym, sg, mu, w, mode) == b
            d, _ = wk.gmm_decode(b, sg, mu, w, R)
            assert np.array_equal(port.gmm_decode(b, sg, mu, w, R, mode), d)
            # row-major (N,4) hand-over gives the same stream as the K-major views
            b2, _ = wk.gmm_encode(sym, sg, mu, w, R, kmajor=False)
            assert b2 == b


def test_ref_variants_agree(ref_available):
    """-march=native and -march=x86-64-v3 builds contract the same FMAs."""
    if not ref_available or not (ref.REF_DIR / "ans_native.so").exists() or not (ref.REF_DIR / "ans_v3.so").exists():
        pytest.skip("needs both oracle/_ref variants")
    if ref.pick_variant() != "native":
        pytest.skip("native variant not runnable on this host")
    sym, sg, mu, w = synth.gmm_numpy(20_000, seed=21)
    outs = []
    for v in ("native", "v3"):
        with ref.RefWorker(0, variant=v) as wk:
            outs.append(wk.gmm_encode(sym, sg, mu, w, 0)[0])
    assert outs[0] == outs[1]


def test_exp256_and_phi_sanity():
    x = np.linspace(-20, 5, 101).astype(np.float32)
    e = port.exp256(x)
    assert np.allclose(e, np.exp(x.astype(np.float64)), rtol=3e-7)
    t = np.linspace(-6, 6, 121).astype(np.float32)
    from math import erf, sqrt
    truth = np.array([0.5 * (1 + erf(v / sqrt(2))) for v in t])
    assert np.abs(port.phi(0, t) - truth).max() < 4e-3      # Polya
    assert np.abs(port.phi(1, t) - truth).max() < 1e-6      # A&S 26.2.17 (7.5e-8 + fp32)
    assert np.abs(port.phi(2, t) - truth).max() < 1e-2      # logistic
    assert port.phi(0, np.float32(-0.0)) == np.float32(0.5)


def test_table_coder_roundtrip_and_reference(ref_available):
    r = np.random.default_rng(5)
    n_cdfs, width = 6, 40
    cdfs = np.zeros((n_cdfs, width), np.int32)
    sizes = np.zeros(n_cdfs, np.int32)
    for c in range(n_cdfs):
        m = int(r.integers(3, width - 2))
        pmf = r.integers(1, 1000, m + 1).astype(np.float64)
        q = np.maximum(1, np.floor(pmf / pmf.sum() * 65536)).astype(np.int64)
        q[np.argmax(q)] += 65536 - q.sum()
        cdfs[c, 1:m + 2] = np.cumsum(q)
        sizes[c] = m + 2
    offsets = -(sizes // 2).astype(np.int32)
    n = 5000
    idx = r.integers(0, n_cdfs, n).astype(np.int32)
    sym = (r.normal(0, 6, n)).round().astype(np.int32)       # some out of range -> bypass
    sym[:4] = [100000, -100000, 0, 1]
    b = port.table_encode(sym, idx, cdfs, sizes, offsets)
    assert np.array_equal(port.table_decode(b, idx, cdfs, sizes, offsets), sym)
    if ref_available:
        with ref.RefWorker(0) as wk:
            rows = [cdfs[c, :sizes[c]] for c in range(n_cdfs)]
            assert wk.table_encode(sym, idx, rows, sizes, offsets) == b
            assert np.array_equal(wk.table_decode(b, idx, rows, sizes, offsets), sym)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_phi_is_monotone_on_sampled_windows(mode):
    """Windows of consecutive floats (the exhaustive run over all 2^32 is oracle/check_monotone.c):
    a monotone Phi makes every quantised CDF row non-decreasing, which the interleaved decoder's
    bisection and the compatible decoder's fast path rely on."""
    import ctypes as C
    import struct

    L = port.lib()
    L.fgo_phi_count_decreasing.restype = C.c_int64
    r = np.random.default_rng(3)
    starts = [1e-30, 1e-6, 0.01, 0.4999, 1.0, 2.0, 3.3, 5.2, 5.9, 9.0, 13.2, 40.0] + list(10 ** r.uniform(-3, 1.5, 8))
    for t0 in starts:
        bits = struct.unpack("<I", struct.pack("<f", t0))[0]
        assert L.fgo_phi_count_decreasing(C.c_int(mode), C.c_uint32(bits), C.c_uint32(400_000)) == 0, t0


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_port_matches_compiled_reference_on_extreme_parameters(mode, ref_available):
    if not ref_available:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    with ref.RefWorker(mode) as wk:
        for seed in (201, 202, 203):
            sym, sg, mu, w = extreme_params(seed)
            R = synth.search_radius(sym)
            b, _ = wk.gmm_encode(sym, sg, mu, w, R)
            assert port.gmm_encode(sym, sg, mu, w, mode) == b, seed
            # the search decoder on the same parameters: oracle == reference symbol for symbol (whether or not the
            # stream round-trips -- with non-monotone rows the reference itself may not)
            small = slice(0, 1500)                                      # the reference's search is O(log R) calls per symbol
            b2, _ = wk.gmm_encode(sym[small], sg[small], mu[small], w[small], R)
            d, _ = wk.gmm_decode(b2, sg[small], mu[small], w[small], R)
            assert np.array_equal(port.gmm_decode(b2, sg[small], mu[small], w[small], R, mode), d), seed
