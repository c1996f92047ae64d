"""Drop-in for the reference's pybind module ``compressai.ans``
(/root/reference/compressai/cpp_exts/rans/rans_interface.cpp:961-1036).

Same classes, method names, argument names and return types:

* ``RansEncoder.encode_with_indexes_gmm(symbols, scales, means, weights, max_value) -> bytes``
* ``RansDecoder.decode_with_indexes_gmm(encoded, scales, means, weights, max_bs_value) -> Tensor[int32]``
* ``RansEncoder.encode_with_indexes(symbols, indexes, cdfs, cdfs_sizes, offsets) -> bytes``
* ``RansDecoder.decode_with_indexes(encoded, indexes, cdfs, cdfs_sizes, offsets) -> list[int]``
* ``RansDecoder.set_stream / decode_stream``, ``BufferedRansEncoder``

Differences, all additive: tensors may live on a CUDA device (then nothing but
4 bytes/symbol crosses PCIe on encode) and the decoder returns its result on the
device the parameters are on. ``APPROX_MODE`` (0 Polya, 1 A&S, 2 logistic -- the
reference's code numbering, rans_interface.cpp:98) is read from the environment on
first use like the reference does, or set with :func:`set_approx_mode`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib

_mode = None


def get_approx_mode() -> int:
    """rans_interface.cpp:99-116: unset / unparsable / out of range -> 0."""
    global _mode
    if _mode is None:
        try:
            m = int(os.environ.get("APPROX_MODE", "0"))
        except ValueError:
            m = 0
        _mode = m if 0 <= m <= 2 else 0
    return _mode


def set_approx_mode(mode: int) -> None:
    global _mode
    if mode not in (0, 1, 2):
        raise ValueError("APPROX_MODE must be 0 (Polya), 1 (Abramowitz-Stegun) or 2 (logistic)")
    _mode = int(mode)


def _tensor(x, dtype):
    if isinstance(x, torch.Tensor):
        if x.dtype != dtype:
            raise TypeError(f"expected {dtype}, got {x.dtype}")   # reference: data_ptr<T>() throws
        return x
    t = torch.as_tensor(np.asarray(x))
    if t.dtype != dtype:
        raise TypeError(f"expected {dtype}, got {t.dtype}")
    return t


def _gmm_views(scales, means, weights, n=None):
    """(N,4) float32 views -> same-device tensors sharing strides the C ABI accepts."""
    ts = [_tensor(a, torch.float32) for a in (scales, means, weights)]
    for t in ts:
        if t.dim() != 2 or t.size(1) != 4:
            raise ValueError(f"GMM parameters must be (N, 4) tensors (K=4 is the only bound instantiation); got {tuple(t.shape)}")
    N = ts[0].size(0)
    if any(t.size(0) != N for t in ts) or (n is not None and N != n):
        raise ValueError("symbols / scales / means / weights disagree on N")
    dev = ts[0].device
    if any(t.device != dev for t in ts):
        raise ValueError("scales / means / weights must be on one device")
    strides = ts[0].stride()
    same = all(t.stride() == strides for t in ts)
    on_host = dev.type == "cpu"
    ok_host = strides == (4, 1) or (strides[0] == 1 and strides[1] >= N) or N <= 1
    if not same or (on_host and not ok_host):
        ts = [t.contiguous() for t in ts]
        strides = ts[0].stride()
    if N <= 1:
        ts = [t.contiguous() for t in ts]
        strides = (4, 1)
    return ts, N, strides, dev


def _dev_index(dev):
    if dev.type == "cpu":
        return torch.cuda.current_device() if torch.cuda.is_available() else 0
    return dev.index if dev.index is not None else torch.cuda.current_device()


def gmm_encode(symbols, scales, means, weights, mode=None) -> bytes:
    sym = _tensor(symbols, torch.int32)
    if sym.dim() != 1:
        raise ValueError("symbols must be 1-D")
    (sc, mu, w), N, (sn, sk), dev = _gmm_views(scales, means, weights, sym.numel())
    if sym.device != dev:
        raise ValueError("symbols and parameters must be on one device")
    sym = sym.contiguous()
    mem = _lib.MEM_HOST if dev.type == "cpu" else _lib.MEM_DEVICE
    if mem == _lib.MEM_DEVICE:
        torch.cuda.current_stream(dev).synchronize()   # producers of the inputs ran on torch's stream
    L = _lib.lib()
    h = _lib.ctx(_dev_index(dev))
    cap = max(64, (N * 9) // 4 + 64)
    mode = get_approx_mode() if mode is None else mode
    for _ in range(2):
        out = np.empty(cap, np.uint8)
        n_out = C.c_size_t(0)
        rc = L.flashgmm_gmm_encode(h, sym.data_ptr(), sc.data_ptr(), mu.data_ptr(), w.data_ptr(), N, sn, sk,
                                   mode, mem, out.ctypes.data, cap, C.byref(n_out))
        if rc == _lib.ENOSPC:
            cap = n_out.value
            continue
        _lib.check(rc)
        return out[: n_out.value].tobytes()
    raise RuntimeError("unreachable")


def gmm_decode(encoded: bytes, scales, means, weights, max_bs_value: int, mode=None) -> torch.Tensor:
    (sc, mu, w), N, (sn, sk), dev = _gmm_views(scales, means, weights)
    mem = _lib.MEM_HOST if dev.type == "cpu" else _lib.MEM_DEVICE
    if mem == _lib.MEM_DEVICE:
        torch.cuda.current_stream(dev).synchronize()
    out = torch.empty(N, dtype=torch.int32, device=dev)
    buf = np.frombuffer(encoded, np.uint8)
    mode = get_approx_mode() if mode is None else mode
    _lib.check(_lib.lib().flashgmm_gmm_decode(
        _lib.ctx(_dev_index(dev)), buf.ctypes.data, buf.size, sc.data_ptr(), mu.data_ptr(), w.data_ptr(),
        N, sn, sk, mode, mem, int(max_bs_value), out.data_ptr()))
    return out


def _cdf_matrix(cdfs, cdfs_sizes, offsets):
    sizes = np.ascontiguousarray(cdfs_sizes, np.int32)
    offs = np.ascontiguousarray(offsets, np.int32)
    if isinstance(cdfs, (np.ndarray, torch.Tensor)):
        m = np.ascontiguousarray(np.asarray(cdfs), np.int32)
    else:
        width = max((len(r) for r in cdfs), default=0)
        m = np.zeros((len(cdfs), width), np.int32)
        for i, r in enumerate(cdfs):
            m[i, : len(r)] = r
    if m.ndim != 2 or m.shape[0] != sizes.shape[0] or offs.shape[0] != sizes.shape[0]:
        raise ValueError("cdfs / cdfs_sizes / offsets disagree")
    return m, sizes, offs


def table_encode(symbols, indexes, cdfs, cdfs_sizes, offsets) -> bytes:
    sym = np.ascontiguousarray(symbols, np.int32)
    idx = np.ascontiguousarray(indexes, np.int32)
    if sym.shape != idx.shape or sym.ndim != 1:
        raise ValueError("symbols / indexes must be 1-D and the same length")
    m, sizes, offs = _cdf_matrix(cdfs, cdfs_sizes, offsets)
    cap = max(64, sym.size * 4 + 64)
    L = _lib.lib()
    for _ in range(2):
        out = np.empty(cap, np.uint8)
        n_out = C.c_size_t(0)
        rc = L.flashgmm_rans_encode_table(sym.ctypes.data, idx.ctypes.data, sym.size, m.ctypes.data, m.shape[0],
                                          m.shape[1], sizes.ctypes.data, offs.ctypes.data, out.ctypes.data, cap,
                                          C.byref(n_out))
        if rc == _lib.ENOSPC:
            cap = n_out.value
            continue
        _lib.check(rc)
        return out[: n_out.value].tobytes()
    raise RuntimeError("unreachable")


def table_decode(encoded: bytes, indexes, cdfs, cdfs_sizes, offsets) -> np.ndarray:
    idx = np.ascontiguousarray(indexes, np.int32)
    m, sizes, offs = _cdf_matrix(cdfs, cdfs_sizes, offsets)
    buf = np.frombuffer(encoded, np.uint8)
    out = np.empty(idx.size, np.int32)
    _lib.check(_lib.lib().flashgmm_rans_decode_table(buf.ctypes.data, buf.size, idx.ctypes.data, idx.size,
                                                     m.ctypes.data, m.shape[0], m.shape[1], sizes.ctypes.data,
                                                     offs.ctypes.data, out.ctypes.data))
    return out


def scalar_gaussian_encode(symbols, scales, mode=None) -> bytes:
    """``RansEncoder.encode_with_indexes(symbols, scales, max_value)`` (rans_interface.cpp:401-456): zero-mean
    Gaussian of scale ``scales[i]`` per symbol through the SCALAR libm approximations. Host routine (the
    reference's is scalar CPU code and no shipped model calls it); parity unpinned across hosts."""
    sym = np.ascontiguousarray(symbols, np.int32)
    sc = np.ascontiguousarray(scales, np.float32)
    if sym.ndim != 1 or sym.shape != sc.shape:
        raise ValueError("symbols / scales must be 1-D and the same length")
    mode = get_approx_mode() if mode is None else mode
    cap = max(64, sym.size * 5 + 64)
    out = np.empty(cap, np.uint8)
    n_out = C.c_size_t(0)
    _lib.check(_lib.lib().flashgmm_scalar_gaussian_encode(sym.ctypes.data, sc.ctypes.data, sym.size, mode,
                                                          out.ctypes.data, cap, C.byref(n_out)))
    return out[: n_out.value].tobytes()


def scalar_gaussian_decode(encoded: bytes, scales, max_bs_value: int, mode=None) -> np.ndarray:
    """``RansDecoder.decode_with_indexes(encoded, scales, max_bs_value)`` (rans_interface.cpp:690-764)."""
    sc = np.ascontiguousarray(scales, np.float32)
    buf = np.frombuffer(encoded, np.uint8)
    out = np.empty(sc.size, np.int32)
    mode = get_approx_mode() if mode is None else mode
    _lib.check(_lib.lib().flashgmm_scalar_gaussian_decode(buf.ctypes.data, buf.size, sc.ctypes.data, sc.size, mode,
                                                          int(max_bs_value), out.ctypes.data))
    return out


class RansEncoder:
    """rans_interface.cpp:986-1004."""

    def encode_with_indexes(self, symbols, indexes, cdfs=None, cdfs_sizes=None, offsets=None):
        if cdfs_sizes is None:            # the (symbols, scales, max_value) overload, rans_interface.cpp:600-607
            return scalar_gaussian_encode(symbols, indexes)
        return table_encode(symbols, indexes, cdfs, cdfs_sizes, offsets)

    def encode_with_indexes_gmm(self, symbols, scales, means, weights, max_value=0):
        return gmm_encode(symbols, scales, means, weights)


class RansDecoder:
    """rans_interface.cpp:1006-1035."""

    def __init__(self):
        self._stream = None
        self._pos = 0

    def decode_with_indexes(self, encoded, indexes, cdfs=None, cdfs_sizes=None, offsets=None):
        if cdfs_sizes is None:            # the (encoded, scales, max_bs_value) overload, rans_interface.cpp:690-764
            return scalar_gaussian_decode(encoded, indexes, int(cdfs)).tolist()
        return table_decode(encoded, indexes, cdfs, cdfs_sizes, offsets).tolist()

    def decode_with_indexes_gmm(self, encoded, scales, means, weights, max_bs_value):
        return gmm_decode(encoded, scales, means, weights, max_bs_value)


def _gmm_packed_host(symbols, scales, means, weights, mode=None):
    """(start | freq << 16) records of the startfreq kernel as host arrays (+ host symbols)."""
    sym = _tensor(symbols, torch.int32)
    (sc, mu, w), N, (sn, sk), dev = _gmm_views(scales, means, weights, sym.numel())
    if dev.type == "cpu":      # buffered API with CPU tensors: run the kernel on the current device
        cdev = torch.device("cuda", torch.cuda.current_device())
        sym_d = sym.to(cdev)
        sc, mu, w = (torch.stack([t[:, k] for k in range(4)]).to(cdev).t() for t in (sc, mu, w))
        sn, sk = sc.stride()
        dev = cdev
    else:
        if sym.device != dev:
            raise ValueError("symbols and parameters must be on one device")
        sym_d = sym
    sym_d = sym_d.contiguous()
    out = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev)
    mode = get_approx_mode() if mode is None else mode
    _lib.check(_lib.lib().flashgmm_gmm_startfreq(_lib.ctx(_dev_index(dev)), sym_d.data_ptr(), sc.data_ptr(),
                                                 mu.data_ptr(), w.data_ptr(), N, sn, sk, mode, out.data_ptr(),
                                                 st.cuda_stream))
    packed = out[:N].cpu().numpy().view(np.uint32)
    return packed, sym.cpu().numpy()


class BufferedRansEncoder:
    """rans_interface.cpp:977-984 (+ :334-399, 458-585): several encode calls, one stream."""

    def __init__(self):
        h = C.c_void_p()
        _lib.check(_lib.lib().flashgmm_benc_create(C.byref(h)))
        self._h = h

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.lib().flashgmm_benc_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def encode_with_indexes(self, symbols, indexes, cdfs, cdfs_sizes, offsets) -> None:
        sym = np.ascontiguousarray(symbols, np.int32)
        idx = np.ascontiguousarray(indexes, np.int32)
        if sym.shape != idx.shape or sym.ndim != 1:
            raise ValueError("symbols / indexes must be 1-D and the same length")
        m, sizes, offs = _cdf_matrix(cdfs, cdfs_sizes, offsets)
        _lib.check(_lib.lib().flashgmm_benc_push_table(self._h, sym.ctypes.data, idx.ctypes.data, sym.size,
                                                       m.ctypes.data, m.shape[0], m.shape[1], sizes.ctypes.data,
                                                       offs.ctypes.data))

    def encode_with_indexes_gmm(self, symbols, scales, means, weights, max_value=0) -> None:
        packed, sym = _gmm_packed_host(symbols, scales, means, weights)
        sym = np.ascontiguousarray(sym, np.int32)
        _lib.check(_lib.lib().flashgmm_benc_push_gmm_packed(self._h, packed.ctypes.data, sym.ctypes.data, packed.size))

    def flush(self) -> bytes:
        L = _lib.lib()
        cap = max(64, int(L.flashgmm_benc_entries(self._h)) * 4 + 64)
        out = np.empty(cap, np.uint8)
        n_out = C.c_size_t(0)
        _lib.check(L.flashgmm_benc_flush(self._h, out.ctypes.data, cap, C.byref(n_out)))
        return out[: n_out.value].tobytes()


def _set_stream(self, encoded) -> None:
    """RansDecoder.set_stream, rans_interface.cpp:886-893."""
    buf = np.frombuffer(encoded, np.uint8)
    if getattr(self, "_sdec", None):
        _lib.lib().flashgmm_sdec_destroy(self._sdec)
        self._sdec = None
    h = C.c_void_p()
    _lib.check(_lib.lib().flashgmm_sdec_create(buf.ctypes.data, buf.size, C.byref(h)))
    self._sdec = h


def _decode_stream(self, indexes, cdfs, cdfs_sizes, offsets):
    """RansDecoder.decode_stream, rans_interface.cpp:895-956: continues where the last call stopped."""
    if not getattr(self, "_sdec", None):
        raise RuntimeError("decode_stream called before set_stream")
    idx = np.ascontiguousarray(indexes, np.int32)
    m, sizes, offs = _cdf_matrix(cdfs, cdfs_sizes, offsets)
    out = np.empty(idx.size, np.int32)
    _lib.check(_lib.lib().flashgmm_sdec_decode_table(self._sdec, idx.ctypes.data, idx.size, m.ctypes.data, m.shape[0],
                                                     m.shape[1], sizes.ctypes.data, offs.ctypes.data, out.ctypes.data))
    return out.tolist()


def _dec_del(self):
    try:
        if getattr(self, "_sdec", None):
            _lib.lib().flashgmm_sdec_destroy(self._sdec)
            self._sdec = None
    except Exception:
        pass


RansDecoder.set_stream = _set_stream
RansDecoder.decode_stream = _decode_stream
RansDecoder.__del__ = _dec_del
