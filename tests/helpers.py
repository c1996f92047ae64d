"""Shared test helpers: build decode-table rows (layout of gmm_kernels.cu) from the
oracle's quantised CDF values."""
import numpy as np

LUT_BUCKETS, LUT_SHIFT = 32, 11


def lut_offset(R):
    return ((2 * R + 3) + 7) & ~7


def has_lut(R):
    """Rows of <= 64 candidates are searched whole (AVX2) and carry no jump table."""
    return 2 * R + 1 > 64


def row_stride(R):
    return lut_offset(R) + (LUT_BUCKETS if has_lut(R) else 0)


def make_rows(tab: np.ndarray, R: int) -> np.ndarray:
    """tab: u16[n, 2R+2] -> u16[n, row_stride(R)] with flag slot and jump table."""
    n, W = tab.shape
    assert W == 2 * R + 2
    T = np.zeros((n, row_stride(R)), np.uint16)
    T[:, :W] = tab
    T[:, W] = (np.diff(tab.astype(np.int32), axis=1) < 0).any(1)
    cand = tab[:, : 2 * R + 1].astype(np.int64)                   # candidates j in [0, 2R]
    for b in range(LUT_BUCKETS if has_lut(R) else 0):
        cnt = (cand <= (b << LUT_SHIFT)).sum(1)
        T[:, lut_offset(R) + b] = np.maximum(cnt - 1, 0)
    return T


def extreme_params(seed, n=6000):
    """Parameter corners the synthetic mixture never visits: scales far outside the model's clamp (the coder itself does
    not clamp, rans_interface.cpp:491-495), one-hot and vanishing weights, weight sums off 1, means far from the symbols
    (bypass), exact ties at the interval boundaries."""
    r = np.random.default_rng(seed)
    K = 4
    mu = (r.choice([0.5, 3.0, 40.0, 3000.0], (K, n)) * r.standard_normal((K, n))).astype(np.float32)
    sg = r.choice(np.array([1e-4, 0.05, 0.11, 0.5, 1.0, 7.0, 256.0, 5000.0], np.float32), (K, n))
    kind = r.integers(0, 5, n)
    w = r.dirichlet(np.ones(K), n).T.astype(np.float32)
    onehot = np.eye(K, dtype=np.float32)[r.integers(0, K, n)].T
    w = np.where(kind == 1, onehot, w)
    w = np.where(kind == 2, w * np.float32(0.97), w)                    # sums below 1
    w = np.where(kind == 3, w * np.float32(1.0004), w)                  # a hair above 1 (65535 * sum still < 65536)
    w = np.where(kind == 4, np.maximum(w, np.float32(1e-7)) * (w > 0.3), w).astype(np.float32)   # some components exactly 0
    w[:, (w.sum(0) == 0)] = np.float32(0.25)
    c = r.integers(0, K, n)
    ar = np.arange(n)
    sym = np.round(mu[c, ar] + np.minimum(sg[c, ar], 300.0) * r.standard_normal(n)).astype(np.int64)
    sym[::17] += r.integers(-40, 40, sym[::17].size)                    # off-model symbols: bypass
    half = (np.round(mu[0]) + 0.5).astype(np.float32)
    mu[0, ::29] = half[::29]                                            # mean exactly on an interval boundary
    sym = np.clip(sym, -20000, 20000).astype(np.int32)
    return sym, np.ascontiguousarray(sg).T, np.ascontiguousarray(mu).T, np.ascontiguousarray(w).T
