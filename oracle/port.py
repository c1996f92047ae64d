"""ctypes binding of the C restatement ``oracle/gmm_oracle.c`` (TEST INFRASTRUCTURE).

All functions take numpy arrays. Parameter arrays are ``(N, 4)`` float32 with
arbitrary element strides (the reference hands over ``(N,K)`` views with strides
``(1,N)`` -- entropy_models.py:813-827); strides are passed through, no copy.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = _HERE / "_build" / "libgmm_oracle.so"
        if not so.exists():
            from . import build
            build.build_c_oracle()
        L = C.CDLL(str(so))
        L.fgo_exp256.restype = C.c_float
        L.fgo_exp256.argtypes = [C.c_float]
        L.fgo_phi.restype = C.c_float
        L.fgo_phi.argtypes = [C.c_int, C.c_float]
        L.fgo_gmm_encode.restype = C.c_int64
        L.fgo_table_encode.restype = C.c_int64
        L.fgo_gmm_decode.restype = C.c_int
        L.fgo_table_decode.restype = C.c_int
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _params(scales, means, weights):
    out = []
    sn = sk = None
    for a in (scales, means, weights):
        a = np.asarray(a)
        assert a.dtype == np.float32 and a.ndim == 2 and a.shape[1] == 4, (a.dtype, a.shape)
        s = (a.strides[0] // 4, a.strides[1] // 4)
        if sn is None:
            sn, sk = s
        elif s != (sn, sk):
            a = np.ascontiguousarray(a)
            raise ValueError("parameter arrays must share strides")
        out.append(a)
    return out, sn, sk


def exp256(x: np.ndarray) -> np.ndarray:
    L = lib()
    x = np.asarray(x, np.float32)
    return np.array([L.fgo_exp256(float(v)) for v in x.ravel()], np.float32).reshape(x.shape)


def phi(mode: int, t: np.ndarray) -> np.ndarray:
    L = lib()
    t = np.asarray(t, np.float32)
    return np.array([L.fgo_phi(mode, float(v)) for v in t.ravel()], np.float32).reshape(t.shape)


def gmm_startfreq(symbols, scales, means, weights, mode=0):
    """-> (start u16[N], freq u16[N], bypass u8[N]); bypass rows are (65535, 1)."""
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    sym = np.ascontiguousarray(symbols, np.int32)
    n = sym.shape[0]
    start = np.empty(n, np.uint16)
    freq = np.empty(n, np.uint16)
    byp = np.empty(n, np.uint8)
    lib().fgo_gmm_startfreq(_p(sym), _p(sc), _p(mu), _p(w), C.c_int64(n), C.c_int64(sn), C.c_int64(sk),
                            C.c_int(mode), _p(start), _p(freq), _p(byp))
    return start, freq, byp


def gmm_cdf_table(scales, means, weights, R, mode=0):
    """-> u16[N, 2R+2]: T[i, j] = q16(cdf_i(j - R - 0.5))."""
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    n = sc.shape[0]
    W = 2 * R + 2
    t = np.empty((n, W), np.uint16)
    lib().fgo_gmm_cdf_table(_p(sc), _p(mu), _p(w), C.c_int64(n), C.c_int64(sn), C.c_int64(sk),
                            C.c_int(mode), C.c_int32(R), _p(t), C.c_int64(W))
    return t


def gmm_encode(symbols, scales, means, weights, mode=0) -> bytes:
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    sym = np.ascontiguousarray(symbols, np.int32)
    n = sym.shape[0]
    cap = 4 * (10 * n + 16) + 64
    out = np.empty(cap, np.uint8)
    nb = lib().fgo_gmm_encode(_p(sym), _p(sc), _p(mu), _p(w), C.c_int64(n), C.c_int64(sn), C.c_int64(sk),
                              C.c_int(mode), _p(out), C.c_size_t(cap))
    assert nb >= 0
    return out[:nb].tobytes()


def gmm_decode(encoded: bytes, scales, means, weights, max_bs_value: int, mode=0, return_overrun=False):
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    n = sc.shape[0]
    enc = np.frombuffer(encoded, np.uint8)
    out = np.empty(n, np.int32)
    ov = lib().fgo_gmm_decode(_p(enc), C.c_size_t(enc.size), _p(sc), _p(mu), _p(w), C.c_int64(n),
                              C.c_int64(sn), C.c_int64(sk), C.c_int(mode), C.c_int32(max_bs_value), _p(out))
    return (out, ov) if return_overrun else out


def _cdf_args(cdfs, cdf_sizes, offsets):
    cdfs = np.ascontiguousarray(cdfs, np.int32)
    assert cdfs.ndim == 2
    return cdfs, np.ascontiguousarray(cdf_sizes, np.int32), np.ascontiguousarray(offsets, np.int32)


def table_encode(symbols, indexes, cdfs, cdf_sizes, offsets) -> bytes:
    cdfs, sizes, offs = _cdf_args(cdfs, cdf_sizes, offsets)
    sym = np.ascontiguousarray(symbols, np.int32)
    idx = np.ascontiguousarray(indexes, np.int32)
    n = sym.shape[0]
    cap = 4 * (10 * n + 16) + 64
    out = np.empty(cap, np.uint8)
    nb = lib().fgo_table_encode(_p(sym), _p(idx), C.c_int64(n), _p(cdfs), C.c_int64(cdfs.shape[1]),
                                _p(sizes), _p(offs), _p(out), C.c_size_t(cap))
    assert nb >= 0
    return out[:nb].tobytes()


def table_decode(encoded: bytes, indexes, cdfs, cdf_sizes, offsets):
    cdfs, sizes, offs = _cdf_args(cdfs, cdf_sizes, offsets)
    idx = np.ascontiguousarray(indexes, np.int32)
    n = idx.shape[0]
    enc = np.frombuffer(encoded, np.uint8)
    out = np.empty(n, np.int32)
    lib().fgo_table_decode(_p(enc), C.c_size_t(enc.size), _p(idx), C.c_int64(n), _p(cdfs),
                           C.c_int64(cdfs.shape[1]), _p(sizes), _p(offs), _p(out))
    return out


def gmm_likelihood_f64(x, scales, means, weights, K=4, scale_bound=0.11, lik_bound=1e-9):
    """x: f32[N]; params f32[N, K] (any strides). float64 truth for the likelihood kernels."""
    x = np.ascontiguousarray(x, np.float32)
    sc, mu, w = (np.asarray(a) for a in (scales, means, weights))
    assert sc.dtype == np.float32 and sc.shape == (x.shape[0], K)
    sn, sk = sc.strides[0] // 4, sc.strides[1] // 4
    for a in (mu, w):
        assert (a.strides[0] // 4, a.strides[1] // 4) == (sn, sk)
    out = np.empty(x.shape[0], np.float64)
    lib().fgo_gmm_likelihood_f64(_p(x), _p(sc), _p(mu), _p(w), C.c_int64(x.shape[0]), C.c_int64(sn),
                                 C.c_int64(sk), C.c_int(K), C.c_double(scale_bound), C.c_double(lik_bound), _p(out))
    return out


def ileave_encode(symbols, scales, means, weights, R: int, mode=0, segment_log2=13) -> bytes:
    """CPU statement of the interleaved container (flashgmm_b200/csrc/gmm_ileave.cu)."""
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    sym = np.ascontiguousarray(symbols, np.int32)
    n = sym.shape[0]
    S = 1 << segment_log2
    n_seg = (n + S - 1) >> segment_log2
    cap = 32 + 4 * (n_seg + 1) + n_seg * 2 * (3 * S + 80)
    out = np.empty(cap, np.uint8)
    L = lib()
    L.fgo_ileave_encode.restype = C.c_int64
    nb = L.fgo_ileave_encode(_p(sym), _p(sc), _p(mu), _p(w), C.c_int64(n), C.c_int64(sn), C.c_int64(sk), C.c_int(mode),
                             C.c_int32(R), C.c_int(segment_log2), _p(out), C.c_size_t(cap))
    assert nb >= 0
    return out[:nb].tobytes()


def ileave_decode(encoded: bytes, scales, means, weights):
    (sc, mu, w), sn, sk = _params(scales, means, weights)
    n = sc.shape[0]
    enc = np.frombuffer(encoded, np.uint8)
    out = np.zeros(n, np.int32)
    L = lib()
    L.fgo_ileave_decode.restype = C.c_int
    rc = L.fgo_ileave_decode(_p(enc), C.c_size_t(enc.size), _p(sc), _p(mu), _p(w), C.c_int64(n), C.c_int64(sn),
                             C.c_int64(sk), _p(out))
    if rc:
        raise ValueError("malformed interleaved stream")
    return out
