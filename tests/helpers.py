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
