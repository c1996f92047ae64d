"""Interleaved multi-lane rANS (opt-in format, no reference counterpart): the GPU coder must
produce byte for byte the container the CPU statement in oracle/gmm_oracle.c produces, decode
the oracle's containers, round-trip every input, and stay within a bounded overhead of the
reference-compatible single stream."""
import numpy as np
import pytest
import torch

from flashgmm_b200 import ans, ileave, synth
from oracle import port

pytestmark = pytest.mark.gpu


def dev(*arrs):
    out = []
    for a in arrs:
        if a.ndim == 2:
            out.append(torch.from_numpy(np.ascontiguousarray(a.T)).cuda().t())
        else:
            out.append(torch.from_numpy(np.ascontiguousarray(a)).cuda())
    return out


def case(n, seed, outliers=True, **kw):
    sym, sg, mu, w = synth.gmm_numpy(n, seed=seed, **kw)
    sym = sym.copy()
    if outliers and n > 50:
        sym[[3, 40, n - 1]] = [5000, -2 ** 31, 2 ** 31 - 1]
    return sym, sg, mu, w


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("n,sl", [(1, 10), (31, 10), (33, 10), (1024, 10), (1025, 10), (5000, 10), (70_000, 12),
                                  (200_000, 13)])
def test_bytes_match_cpu_statement_and_roundtrip(n, sl, mode):
    sym, sg, mu, w = case(n, 100 + n % 97)
    R = 20
    s, a, m, ww = dev(sym, sg, mu, w)
    blob = ileave.encode(s, a, m, ww, R, mode, sl)
    want = port.ileave_encode(sym, sg, mu, w, R, mode, sl)
    assert blob == want
    assert ileave.info(blob) == {"n": n, "mode": mode, "R": R, "segment_log2": sl}
    d = ileave.decode(blob, a, m, ww)
    assert torch.equal(d, s)
    assert np.array_equal(port.ileave_decode(blob, sg, mu, w), sym)


def test_empty_stream():
    e = np.zeros(0, np.int32)
    p = np.zeros((0, 4), np.float32)
    s, a, m, ww = (torch.from_numpy(x).cuda() for x in (e, p, p, p))
    blob = ileave.encode(s, a, m, ww, 5)
    assert blob == port.ileave_encode(e, p, p, p, 5, 0, 13) and len(blob) == 36
    assert ileave.decode(blob, a, m, ww).numel() == 0


@pytest.mark.parametrize("R", [0, 1, 3, 22, 300])
def test_any_search_radius_roundtrips(R):
    """Symbols outside [-R, R] are escaped: R is a speed / rate knob, never a correctness one."""
    sym, sg, mu, w = case(30_000, 7)
    s, a, m, ww = dev(sym, sg, mu, w)
    blob = ileave.encode(s, a, m, ww, R, 0, 11)
    assert blob == port.ileave_encode(sym, sg, mu, w, R, 0, 11)
    assert torch.equal(ileave.decode(blob, a, m, ww), s)


def test_heavy_bypass_variant():
    """sigma == 0.11 with symbols far from every mean: almost every symbol escapes."""
    sym, sg, mu, w = case(50_000, 9, outliers=False, sigma_span=0.0, offcentre=9.0)
    s, a, m, ww = dev(sym, sg, mu, w)
    blob = ileave.encode(s, a, m, ww, 40, 0, 12)
    assert blob == port.ileave_encode(sym, sg, mu, w, 40, 0, 12)
    assert torch.equal(ileave.decode(blob, a, m, ww), s)


def test_host_memory_inputs_and_row_major_layout():
    sym, sg, mu, w = case(40_000, 11)
    want = port.ileave_encode(sym, sg, mu, w, 22, 0, 13)
    hs = torch.from_numpy(sym)
    kmaj = [torch.from_numpy(np.ascontiguousarray(x.T)).t() for x in (sg, mu, w)]
    assert ileave.encode(hs, *kmaj, 22) == want
    assert torch.equal(ileave.decode(want, *kmaj), hs)
    aos = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (sg, mu, w)]
    assert ileave.encode(hs.cuda(), *aos, 22) == want
    assert torch.equal(ileave.decode(want, *aos).cpu(), hs)


def test_size_overhead_vs_single_stream_is_bounded():
    n = 1 << 20
    sym, sg, mu, w = synth.gmm_numpy(n, seed=1234)
    s, a, m, ww = dev(sym, sg, mu, w)
    single = len(ans.gmm_encode(s, a, m, ww, 0))
    R = synth.search_radius(sym)
    for sl, limit in ((13, 1.04), (16, 1.01)):
        blob = ileave.encode(s, a, m, ww, R, 0, sl)
        assert torch.equal(ileave.decode(blob, a, m, ww), s)
        assert len(blob) <= limit * single, (sl, len(blob), single)


def test_full_size_roundtrip_16mi_symbols():
    """BASELINE.json configs[3] size: property check (round trip) instead of the CPU statement."""
    n = 1 << 24
    s, a, m, ww = synth.gmm_torch(n, seed=1234, device="cuda")
    R = int(s.abs().max().item()) + 2
    blob = ileave.encode(s, a, m, ww, R)
    assert len(blob) < 0.6 * n
    assert torch.equal(ileave.decode(blob, a, m, ww), s)


def test_corrupt_containers_are_rejected():
    sym, sg, mu, w = case(20_000, 13)
    s, a, m, ww = dev(sym, sg, mu, w)
    blob = bytearray(ileave.encode(s, a, m, ww, 22, 0, 10))
    with pytest.raises(Exception):
        ileave.decode(bytes(blob[: len(blob) // 2]), a, m, ww)           # truncated
    bad = bytearray(blob); bad[0] = ord("X")
    with pytest.raises(ValueError):
        ileave.decode(bytes(bad), a, m, ww)                              # wrong magic
    with pytest.raises(ValueError):
        ileave.decode(bytes(blob), a[:-1], m[:-1], ww[:-1])              # n mismatch
    bad = bytearray(blob); bad[32 + 4] ^= 0x02                            # offset no longer 4-byte aligned
    with pytest.raises(ValueError):
        ileave.decode(bytes(bad), a, m, ww)
    bad = bytearray(blob); bad[32 + 4] ^= 0x10                            # segment boundary moved: caught by the kernel
    with pytest.raises(Exception):
        ileave.decode(bytes(bad), a, m, ww)


def test_parameters_that_are_not_a_mixture_still_roundtrip():
    """weights summing above 1 wrap the 16-bit CDF (start + freq > 65535 would overlap the escape slot), negative or
    NaN weights / scales break the row's monotonicity the decoder's search relies on. Such symbols are escaped (raw int32):
    the container round-trips whatever the parameters are, and the GPU still writes the CPU statement's bytes."""
    sym, sg, mu, w = case(20_000, 21, outliers=False)
    sg, mu, w = (np.ascontiguousarray(x).copy() for x in (sg, mu, w))
    w[100:4000] *= 3.0                     # unnormalised: sums to 3
    w[5000:5200, 1] = -0.25                # negative weight
    w[6000:6050, 2] = np.nan
    sg[7000:7100, 0] = 0.0
    sg[7200:7300, 3] = -1.0
    mu[8000:8050, 1] = np.inf
    s, a, m, ww = dev(sym, sg, mu, w)
    for mode in (0, 1, 2):
        blob = ileave.encode(s, a, m, ww, 30, mode, 11)
        assert blob == port.ileave_encode(sym, sg, mu, w, 30, mode, 11)
        assert torch.equal(ileave.decode(blob, a, m, ww), s)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_parameter_corners_roundtrip_and_match_the_cpu_statement(mode):
    """tests/helpers.extreme_params (scales 1e-4 .. 5000, one-hot / vanishing / unnormalised weights, means on interval
    boundaries, off-model symbols): wrapped intervals and insane parameters are escaped, the rest is coded -- same bytes
    as the CPU statement of the format, exact round trip."""
    from helpers import extreme_params
    sym, sg, mu, w = extreme_params(400 + mode, n=20000)
    s, a, m, ww = dev(sym, sg, mu, w)
    for R in (30, 2000):
        blob = ileave.encode(s, a, m, ww, R, mode, 12)
        assert blob == port.ileave_encode(sym, sg, mu, w, R, mode, 12)
        assert torch.equal(ileave.decode(blob, a, m, ww), s)
