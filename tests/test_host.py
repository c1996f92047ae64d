"""CPU-side checks of the product library: ABI completeness and the serial host
halves (rANS push/pop, table coder) against the oracle. No GPU compute is called."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from flashgmm_b200 import _lib, ans, synth
from oracle import port, ref
from helpers import make_rows, row_stride

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol():
    hdr = (ROOT / "include" / "flashgmm.h").read_text()
    declared = set(re.findall(r"\b(flashgmm_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"flashgmm_ctx", "flashgmm_status", "flashgmm_stream"}
    assert len(declared) >= 17
    L = C.CDLL(str(_lib.LIB_PATH))
    missing = [s for s in sorted(declared) if not hasattr(L, s)]
    assert not missing, missing
    assert set(_lib._SIGS) == declared          # the ctypes table binds exactly the header
    assert _lib.lib().flashgmm_version() >= 100


def test_stream_struct_layout_matches_header():
    # 7 pointers/int64 + int32 (+pad) + ptr + 2 size_t + ptr + int (+pad)
    assert C.sizeof(_lib.Stream) == 8 * 7 + 8 + 8 + 16 + 8 + 8


def test_math_constants_are_the_reference_literals():
    src = (ROOT / "flashgmm_b200/csrc/gmm_math.cuh").read_text()
    want = {
        "EXP_HI": np.float32(88.3762626647949), "EXP_LO": np.float32(-88.3762626647949),
        "LOG2EF": np.float32(1.44269504088896341), "EXP_C1": np.float32(0.693359375),
        "EXP_C2": np.float32(-2.12194440e-4), "EXP_P0": np.float32(1.9875691500E-4),
        "EXP_P1": np.float32(1.3981999507E-3), "EXP_P2": np.float32(8.3334519073E-3),
        "EXP_P3": np.float32(4.1665795894E-2), "EXP_P4": np.float32(1.6666665459E-1),
        "EXP_P5": np.float32(5.0000001201E-1),
        "M2_OVER_PI": np.float32(-2.0) / np.float32(3.14159265358979323846),
        "INV_SQRT_2PI": np.float32(0.3989422804014327), "AS_P": np.float32(0.2316419),
        "AS_B1": np.float32(0.319381530), "AS_B2": np.float32(-0.356563782),
        "AS_B3": np.float32(1.781477937), "AS_B4": np.float32(-1.821255978),
        "AS_B5": np.float32(1.330274429), "LOGI_K": np.float32(1.702),
    }
    for k, v in want.items():
        m = re.search(r"#define FGMM_%s\s+FGMM_F\(0x([0-9a-f]{8})u\)" % k, src)
        assert m, k
        assert m.group(1) == "%08x" % np.float32(v).view(np.uint32), k


def _packed_from_oracle(sym, sg, mu, w, mode):
    st, fr, bp = port.gmm_startfreq(sym, sg, mu, w, mode)
    p = st.astype(np.uint32) | (fr.astype(np.uint32) << 16)
    p[bp == 1] = 0xFFFF
    return p


def _push(packed, sym):
    L = _lib.lib()
    cap = packed.size * 48 + 64
    out = np.empty(cap, np.uint8)
    n = C.c_size_t()
    _lib.check(L.flashgmm_rans_push_packed(packed.ctypes.data, sym.ctypes.data if sym is not None else None,
                                           packed.size, out.ctypes.data, cap, C.byref(n)))
    return out[: n.value].tobytes()


CASES = [dict(n=50_000, seed=31), dict(n=20_000, seed=32, sigma_span=0.0),
         dict(n=20_000, seed=33, sigma_span=0.05, offcentre=9.0), dict(n=4099, seed=34), dict(n=1, seed=35),
         dict(n=0, seed=36)]


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("kw", CASES, ids=lambda k: f"n{k['n']}s{k['seed']}")
def test_host_push_and_pop_match_oracle(kw, mode):
    sym, sg, mu, w = synth.gmm_numpy(**kw)
    sym = sym.copy()
    if sym.size > 10:
        sym[3], sym[4] = 2**31 - 1, -2**31          # 8-nibble bypass payloads
    packed = _packed_from_oracle(sym, sg, mu, w, mode)
    want = port.gmm_encode(sym, sg, mu, w, mode)
    assert _push(packed, sym) == want               # reciprocal-multiply push == div/mod push
    # pop on oracle-made rows (radius limited: rows are 2R+2 wide)
    core = np.clip(sym, -60, 60)
    R = synth.search_radius(core)
    want = port.gmm_encode(core, sg, mu, w, mode)
    L = _lib.lib()
    rs = L.flashgmm_gmm_table_row_stride(R)
    assert rs == row_stride(R)
    T = make_rows(port.gmm_cdf_table(sg, mu, w, R, mode), R)
    out = np.empty(core.size, np.int32)
    buf = np.frombuffer(want, np.uint8)
    _lib.check(L.flashgmm_rans_pop_table(buf.ctypes.data, buf.size, T.ctypes.data, rs, core.size, R, out.ctypes.data))
    assert np.array_equal(out, port.gmm_decode(want, sg, mu, w, R, mode))
    assert np.array_equal(out, core)


def test_pop_replays_reference_search_on_non_monotone_rows():
    """Rows that are not non-decreasing must give what the reference's binary search gives."""
    r = np.random.default_rng(7)
    n, R = 4000, 9
    L = _lib.lib()
    rs = L.flashgmm_gmm_table_row_stride(R)
    W = 2 * R + 2
    rows = np.sort(r.integers(0, 65535, (n, W)), axis=1).astype(np.uint16)
    rows[:, 0] = 0
    swap = r.integers(1, W - 1, n)
    ar = np.arange(n)
    a, b = rows[ar, swap].copy(), rows[ar, swap + 1].copy()
    rows[ar, swap], rows[ar, swap + 1] = b, a                      # one inversion per row
    T = make_rows(rows, R)
    assert T[:, W].all()
    # a stream of arbitrary states: random words are a valid (if meaningless) rANS stream
    words = r.integers(0, 2**32, 3 * n + 8, dtype=np.uint64).astype(np.uint32)
    words[1] |= 0x80000000 >> 0
    enc = words.tobytes()
    out = np.empty(n, np.int32)
    rc = L.flashgmm_rans_pop_table(enc, len(enc), T.ctypes.data, rs, n, R, out.ctypes.data)
    assert rc == 0
    # python replay of rans_interface.cpp:793-883 on the same rows
    p = 2
    x = int(words[0]) | (int(words[1]) << 32)
    exp = np.empty(n, np.int32)
    for i in range(n):
        cum = x & 0xFFFF
        if cum == 65535:
            pytest.skip("random stream hit the bypass marker")
        s_, e_, mid = -R, R, 0
        while s_ <= e_:
            mid = s_ + (e_ - s_) // 2 if (e_ - s_) >= 0 else s_
            c1, c2 = int(rows[i, mid + R]), int(rows[i, mid + R + 1])
            if c1 <= cum < c2:
                break
            elif c1 > cum:
                e_ = mid - 1
            else:
                s_ = mid + 1
        c1, c2 = int(rows[i, mid + R]), int(rows[i, mid + R + 1])
        pmf = (c2 - c1) & 0xFFFF or 1
        x = pmf * (x >> 16) + (x & 0xFFFF) - c1
        x &= (1 << 64) - 1
        if x < (1 << 31):
            x = (x << 32) | int(words[p]); p += 1
        exp[i] = mid
    assert np.array_equal(out, exp)


def test_reciprocal_push_equals_division_for_extreme_states():
    """Drive the encoder state across its whole range with every interesting freq."""
    r = np.random.default_rng(3)
    freqs = np.concatenate([np.arange(1, 300), [4095, 4096, 4097, 32767, 32768, 32769, 65534, 65535],
                            r.integers(1, 65536, 3000)]).astype(np.uint32)
    for rep in range(3):
        f = r.permutation(np.tile(freqs, 4))
        start = (r.integers(0, 65536, f.size) % (65536 - f + 1)).astype(np.uint32)
        packed = start | (f << 16)
        # python big-int reference of rans64.h:77-103
        x = 1 << 31
        words = []
        for s_, f_ in zip(start[::-1].tolist(), f[::-1].tolist()):
            if x >= ((1 << 47) * f_):
                words.append(x & 0xFFFFFFFF); x >>= 32
            x = ((x // f_) << 16) + (x % f_) + s_
        words += [x >> 32, x & 0xFFFFFFFF]
        want = np.array(words[::-1], np.uint32).tobytes()
        assert _push(packed, None) == want


def test_table_coder_matches_oracle():
    r = np.random.default_rng(5)
    n_cdfs, width = 5, 64
    cdfs = np.zeros((n_cdfs, width), np.int32)
    sizes = np.zeros(n_cdfs, np.int32)
    for c in range(n_cdfs):
        m = int(r.integers(3, width - 2))
        q = np.maximum(1, np.floor(r.dirichlet(np.ones(m + 1)) * 65000)).astype(np.int64)
        q[np.argmax(q)] += 65536 - q.sum()
        cdfs[c, 1:m + 2] = np.cumsum(q)
        sizes[c] = m + 2
    offsets = -(sizes // 2).astype(np.int32)
    n = 20000
    idx = r.integers(0, n_cdfs, n).astype(np.int32)
    sym = r.normal(0, 8, n).round().astype(np.int32)
    sym[:6] = [2**26, -2**26, 0, 1, -1, 77]
    want = port.table_encode(sym, idx, cdfs, sizes, offsets)
    got = ans.table_encode(sym, idx, [row[:s].tolist() for row, s in zip(cdfs, sizes)], sizes.tolist(), offsets.tolist())
    assert got == want
    assert np.array_equal(ans.table_decode(want, idx, cdfs, sizes, offsets), sym)
    assert ans.RansDecoder().decode_with_indexes(want, idx.tolist(), cdfs.tolist(), sizes.tolist(), offsets.tolist()) == sym.tolist()
    assert ans.table_encode([], [], cdfs, sizes, offsets) == port.table_encode(sym[:0], idx[:0], cdfs, sizes, offsets)


def _toy_cdfs(r, n_cdfs=5, width=64):
    cdfs = np.zeros((n_cdfs, width), np.int32)
    sizes = np.zeros(n_cdfs, np.int32)
    for c in range(n_cdfs):
        m = int(r.integers(3, width - 2))
        q = np.maximum(1, np.floor(r.dirichlet(np.ones(m + 1)) * 65000)).astype(np.int64)
        q[np.argmax(q)] += 65536 - q.sum()
        cdfs[c, 1:m + 2] = np.cumsum(q)
        sizes[c] = m + 2
    return cdfs, sizes, -(sizes // 2).astype(np.int32)


def test_buffered_encoder_and_stream_decoder(ref_available):
    """BufferedRansEncoder (rans_interface.cpp:334-399, 557-585) and set_stream / decode_stream
    (:886-956): several pushes make ONE stream; the decoder's state persists across calls."""
    r = np.random.default_rng(8)
    cdfs, sizes, offsets = _toy_cdfs(r)
    n = 6000
    idx = r.integers(0, len(sizes), n).astype(np.int32)
    sym = r.normal(0, 8, n).round().astype(np.int32)
    sym[:3] = [2**20, -2**20, 0]
    be = ans.BufferedRansEncoder()
    be.encode_with_indexes(sym[:2500], idx[:2500], cdfs, sizes, offsets)
    be.encode_with_indexes(sym[2500:].tolist(), idx[2500:].tolist(), cdfs.tolist(), sizes.tolist(), offsets.tolist())
    b = be.flush()
    assert b == port.table_encode(sym, idx, cdfs, sizes, offsets)          # == one-shot encode of the concatenation
    assert be.flush() == port.table_encode(sym[:0], idx[:0], cdfs, sizes, offsets)   # session cleared: empty stream
    d = ans.RansDecoder()
    with pytest.raises(RuntimeError):
        d.decode_stream(idx[:1], cdfs, sizes, offsets)
    d.set_stream(b)
    a1 = d.decode_stream(idx[:1000], cdfs, sizes, offsets)
    a2 = d.decode_stream(idx[1000:].tolist(), cdfs.tolist(), sizes.tolist(), offsets.tolist())
    assert a1 + a2 == sym.tolist()
    # GMM records pushed through the session (packed words as the startfreq kernel emits them)
    gs, sg, mu, w = synth.gmm_numpy(3000, seed=77, sigma_span=1.0)
    gs = gs.copy(); gs[:2] = [400, -9]
    st, fr, bp = port.gmm_startfreq(gs, sg, mu, w, 0)
    packed = st.astype(np.uint32) | (fr.astype(np.uint32) << 16)
    packed[bp == 1] = 0xFFFF
    L = _lib.lib()
    _lib.check(L.flashgmm_benc_push_gmm_packed(be._h, packed.ctypes.data, gs.ctypes.data, packed.size))
    assert be.flush() == port.gmm_encode(gs, sg, mu, w, 0)
    if ref_available:
        rows = [cdfs[c, :sizes[c]] for c in range(len(sizes))]
        with ref.RefWorker(0) as wk:
            tab = dict(symbols=sym[:2500], indexes=idx[:2500], cdfs=rows, cdf_sizes=sizes, offsets=offsets)
            tab2 = dict(symbols=sym[2500:], indexes=idx[2500:], cdfs=rows, cdf_sizes=sizes, offsets=offsets)
            assert wk.call("buffered", pushes=[("table", tab), ("table", tab2)]) == b
            outs = wk.call("decode_stream", encoded=b, index_lists=[idx[:1000], idx[1000:]], cdfs=rows, cdf_sizes=sizes,
                           offsets=offsets)
            assert np.concatenate(outs).tolist() == sym.tolist()
            # mixed session: table symbols then GMM symbols in one stream
            gm = dict(symbols=gs, scales=sg, means=mu, weights=w)
            mixed_ref = wk.call("buffered", pushes=[("table", tab), ("gmm", gm)])
        be.encode_with_indexes(sym[:2500], idx[:2500], cdfs, sizes, offsets)
        _lib.check(L.flashgmm_benc_push_gmm_packed(be._h, packed.ctypes.data, gs.ctypes.data, packed.size))
        assert be.flush() == mixed_ref


def test_error_behaviour():
    L = _lib.lib()
    out = np.empty(8, np.uint8)
    n = C.c_size_t()
    pk = np.array([0xFFFF], np.uint32)           # bypass record without symbols
    assert L.flashgmm_rans_push_packed(pk.ctypes.data, None, 1, out.ctypes.data, 8, C.byref(n)) == _lib.EINVAL
    assert b"needs symbols" in L.flashgmm_last_error()
    pk = np.array([1 | (5 << 16)] * 100, np.uint32)
    assert L.flashgmm_rans_push_packed(pk.ctypes.data, None, 100, out.ctypes.data, 8, C.byref(n)) == _lib.ENOSPC
    assert n.value > 8
    # truncated stream
    R = 3
    rs = L.flashgmm_gmm_table_row_stride(R)
    T = make_rows(np.tile(np.linspace(0, 65535, 2 * R + 2).astype(np.uint16), (50, 1)), R)
    o = np.empty(50, np.int32)
    enc = np.array([0x12345678, 0x80000000 >> 1], np.uint32).tobytes()   # state only, no renorm words
    assert L.flashgmm_rans_pop_table(enc, len(enc), T.ctypes.data, rs, 50, R, o.ctypes.data) == _lib.ETRUNC
    with pytest.raises(ValueError):
        ans.set_approx_mode(3)
    assert L.flashgmm_max_encoded_bytes(10) == 4 * (100 + 16)


def test_approx_mode_env_parsing(monkeypatch):
    for val, want in [("1", 1), ("2", 2), ("7", 0), ("x", 0), ("-1", 0)]:
        monkeypatch.setenv("APPROX_MODE", val)
        ans._mode = None
        assert ans.get_approx_mode() == want
    monkeypatch.delenv("APPROX_MODE")
    ans._mode = None
    assert ans.get_approx_mode() == 0


# ---- a14: scalar single-Gaussian search coder (host, libm) ------------------------------------------------
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_scalar_gaussian_coder_matches_the_reference_built_here(mode, ref_available):
    """RansEncoder.encode_with_indexes(symbols, scales, max) / RansDecoder.decode_with_indexes(encoded, scales, max)
    (rans_interface.cpp:401-456, 690-764). Parity is unpinned across hosts (glibc expf): the check is against the
    reference compiled on THIS machine, through its own pybind overloads."""
    if not ref_available:
        pytest.skip("compiled reference (oracle/_ref) not present")
    from oracle import ref
    from flashgmm_b200 import ans

    rng = np.random.default_rng(100 + mode)
    n = 20000
    scales = (0.11 + 6.0 * rng.random(n) ** 2).astype(np.float32)
    sym = np.round(rng.standard_normal(n) * scales).astype(np.int32)
    sym[:6] = [0, 900, -900, 70000, -(2 ** 31), 2 ** 31 - 1]            # bypass payloads of 0..8 nibbles
    R = 64
    inside = np.abs(sym) <= R
    with ref.RefWorker(mode=mode) as w:
        want = w.scalar_encode(sym, scales)
        got = ans.scalar_gaussian_encode(sym, scales, mode)
        assert got == want
        dec_ref = w.scalar_decode(want, scales, R)
    dec = ans.scalar_gaussian_decode(got, scales, R, mode)
    assert np.array_equal(dec, dec_ref)
    assert np.array_equal(dec[inside], sym[inside])
    # the pybind-shaped overloads
    ans.set_approx_mode(mode)
    try:
        assert ans.RansEncoder().encode_with_indexes(sym.tolist(), scales.tolist(), R) == want
        assert ans.RansDecoder().decode_with_indexes(want, scales.tolist(), R) == dec_ref.tolist()
    finally:
        ans.set_approx_mode(0)


def test_pop_table_does_not_read_past_the_last_row():
    """Regression (round 2): the 16-entry vector search over-read the end of the table by up to 16 bytes (rows are
    padded to 8 entries); with the table ending at a page boundary that was a segfault. The table is placed so that
    its last byte is the last byte of a readable page followed by a PROT_NONE page."""
    import mmap

    R = 10                                           # 21 candidates -> 2 vectors = 32 entries, row stride 24
    sym, sg, mu, w = synth.gmm_numpy(3000, seed=77, sigma_span=1.0, mean_scale=1.0)
    core = np.clip(sym, -R + 1, R - 1)
    enc = port.gmm_encode(core, sg, mu, w, 0)
    L = _lib.lib()
    rs = L.flashgmm_gmm_table_row_stride(R)
    assert rs == 24
    T = make_rows(port.gmm_cdf_table(sg, mu, w, R, 0), R)
    nbytes = T.nbytes
    page = mmap.PAGESIZE
    span = (nbytes + page - 1) // page * page
    mm = mmap.mmap(-1, span + page)
    base = C.addressof(C.c_char.from_buffer(mm))
    libc = C.CDLL(None, use_errno=True)
    libc.mprotect.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
    assert libc.mprotect(base + span, page, 0) == 0                        # PROT_NONE guard page
    start = base + span - nbytes
    C.memmove(start, T.ctypes.data, nbytes)
    out = np.empty(core.size, np.int32)
    buf = np.frombuffer(enc, np.uint8)
    try:
        _lib.check(L.flashgmm_rans_pop_table(buf.ctypes.data, buf.size, start, rs, core.size, R, out.ctypes.data))
    finally:
        libc.mprotect(base + span, page, 3)
    assert np.array_equal(out, core)


# ---- f3: the container format ----------------------------------------------------------------------------
def test_container_roundtrip_and_corruption():
    import torch
    from flashgmm_b200 import container

    rng = np.random.default_rng(3)
    def stream(c):
        bm = torch.from_numpy((rng.random(c) < 0.7).astype(np.int64))
        return (rng.integers(0, 256, int(rng.integers(8, 4000)), dtype=np.uint8).tobytes(), int(rng.integers(1, 300)), bm)
    # one image, the reference's structure (tuples), ELIC-like shapes
    groups = [16, 16, 32, 64, 192]
    strings = [stream(g) for g in groups for _ in range(2)] + [[b"zzzz-hyper"]]
    shape = {"y": [torch.Size([g, 32, 48]) for g in groups], "hyper": torch.Size([8, 12])}
    blob = container.pack(strings, shape)
    got, shp, info = container.unpack(blob)
    assert shp == {"y": tuple((g, 32, 48) for g in groups), "hyper": (8, 12)}
    assert len(got) == len(strings) and got[-1] == strings[-1]
    for a, b in zip(got[:-1], strings[:-1]):
        assert isinstance(a, tuple) and a[0] == b[0] and a[1] == b[1] and torch.equal(a[2], b[2])
    assert info["images"] == 1 and info["passes"] == 10 and info["payload_bytes"] == sum(len(s[0]) for s in strings[:-1])
    assert info["total_bytes"] == len(blob) and 0 < info["side_info_bytes"] < 10 * (2 + 2 + 24 + 3) + 1
    # a batch of 3 images, checkerboard-like
    strings = [[stream(192) for _ in range(3)] for _ in range(2)] + [[b"a", b"bb", b""]]
    got, shp, info = container.unpack(container.pack(strings, {"y": (192, 32, 48), "hyper": (8, 12)}))
    assert info["images"] == 3 and all(isinstance(e, list) and len(e) == 3 for e in got[:-1]) and got[-1] == [b"a", b"bb", b""]
    for e, f in zip(got[:-1], strings[:-1]):
        for a, b in zip(e, f):
            assert a[0] == b[0] and a[1] == b[1] and torch.equal(a[2], b[2])
    # corruption, truncation, wrong magic / version are refused
    bad = bytearray(blob)
    bad[len(bad) // 2] ^= 1
    for broken in (bytes(bad), blob[:-7], b"XXXX" + blob[4:], blob[:4] + b"\x09" + blob[5:]):
        with pytest.raises(container.ContainerError):
            container.unpack(broken)


def test_exact_math_shortcuts_native_harness(tmp_path):
    """tests/native/check_exact_math.c: (1) the magic-number floor (kept as a measured-and-rejected alternative in
    gmm_math.cuh) equals floorf / truncation for EVERY finite binary32 input; (2) division by a shared, correctly rounded
    reciprocal with two residual corrections (gmm_math.cuh::div_by_shared_rcp, used by the windowed rows kernel) equals
    IEEE division on random (x - mu, sigma) pairs incl. all-ones and power-of-two significands. CPU fmaf == GPU
    __fmaf_rn, so the result transfers. 2 * 10^8 pairs here; 2 * 10^9 were run once for DESIGN.md."""
    import subprocess
    from pathlib import Path

    src = Path(__file__).resolve().parent / "native" / "check_exact_math.c"
    exe = tmp_path / "chk"
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-mfma", "-o", str(exe), str(src), "-lm", "-lpthread"], check=True)
    out = subprocess.run([str(exe), "200000000"], capture_output=True, text=True, timeout=600).stdout
    assert "magic floor: 0 mismatches" in out, out
    assert "2 corrections: 0 mismatches" in out, out
