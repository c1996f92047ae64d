"""Opt-in interleaved multi-lane rANS coder (north_star item 4): the GPU encodes AND decodes,
only the container's bytes cross PCIe. Own format ("FGIL", ``csrc/gmm_ileave.cu``) -- NOT
readable by the reference; the single-stream coder in :mod:`flashgmm_b200.ans` stays the
reference-compatible default.

    blob = ileave.encode(symbols, scales, means, weights, R)      # tensors on CUDA (or CPU: staged)
    sym  = ileave.decode(blob, scales, means, weights)            # -> int32 tensor on the same device
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .ans import _dev_index, _gmm_views, _tensor, get_approx_mode

DEFAULT_SEGMENT_LOG2 = 13


def info(encoded) -> dict:
    buf = np.frombuffer(encoded, np.uint8)
    n, mode, R, sl = C.c_int64(), C.c_int(), C.c_int32(), C.c_int()
    _lib.check(_lib.lib().flashgmm_gmm_ileave_info(buf.ctypes.data, buf.size, C.byref(n), C.byref(mode), C.byref(R),
                                                   C.byref(sl)))
    return {"n": n.value, "mode": mode.value, "R": R.value, "segment_log2": sl.value}


_zone = None
_zone_lock = __import__("threading").Lock()


def _landing_zone(cap: int):
    """Reused page-locked landing zone for the container (a fresh pageable array per call cost more than the kernel:
    first-touch page faults + a pageable device-to-host copy)."""
    global _zone
    if _zone is None or _zone.numel() < cap:
        _zone = torch.empty(int(cap * 1.25) + 4096, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
    return _zone.numpy()[:cap]


def encode(symbols, scales, means, weights, R: int, mode=None, segment_log2: int = 0, as_view: bool = False):
    """Symbols outside ``[-R, R]`` (and symbols the model gives zero mass) are escaped as raw
    int32, so ``R`` trades search depth in the decoder against escapes: any ``R >= 0`` round-trips.
    -> ``bytes``; with ``as_view=True`` a uint8 numpy view of the page-locked landing zone instead (valid until the next
    ``encode`` call of this process): a fresh 7 MB ``bytes`` object costs more host time than the whole device side."""
    sym = _tensor(symbols, torch.int32)
    if sym.dim() != 1:
        raise ValueError("symbols must be 1-D")
    (sc, mu, w), N, (sn, sk), dev = _gmm_views(scales, means, weights, sym.numel())
    if sym.device != dev:
        raise ValueError("symbols and parameters must be on one device")
    sym = sym.contiguous()
    mem = _lib.MEM_HOST if dev.type == "cpu" else _lib.MEM_DEVICE
    if mem == _lib.MEM_DEVICE:
        torch.cuda.current_stream(dev).synchronize()
    L = _lib.lib()
    mode = get_approx_mode() if mode is None else mode
    # typical streams are ~0.5 B/symbol; retry once with the exact size if escapes blow that up
    cap = max(4096, N + (N >> 13) * 160 + 4096)
    for _ in range(2):
        with _zone_lock:
            out = _landing_zone(cap)
            n_out = C.c_size_t(0)
            rc = L.flashgmm_gmm_ileave_encode(_lib.ctx(_dev_index(dev)), sym.data_ptr(), sc.data_ptr(), mu.data_ptr(),
                                              w.data_ptr(), N, sn, sk, mode, mem, int(R), int(segment_log2),
                                              out.ctypes.data, cap, C.byref(n_out))
            if rc == _lib.ENOSPC:
                cap = n_out.value
                continue
            _lib.check(rc)
            return out[: n_out.value] if as_view else out[: n_out.value].tobytes()
    raise RuntimeError("unreachable")


def decode(encoded, scales, means, weights) -> torch.Tensor:
    (sc, mu, w), N, (sn, sk), dev = _gmm_views(scales, means, weights)
    mem = _lib.MEM_HOST if dev.type == "cpu" else _lib.MEM_DEVICE
    if mem == _lib.MEM_DEVICE:
        torch.cuda.current_stream(dev).synchronize()
    out = torch.empty(N, dtype=torch.int32, device=dev)
    buf = np.frombuffer(encoded, np.uint8)
    _lib.check(_lib.lib().flashgmm_gmm_ileave_decode(_lib.ctx(_dev_index(dev)), buf.ctypes.data, buf.size,
                                                     sc.data_ptr(), mu.data_ptr(), w.data_ptr(), N, sn, sk, mem,
                                                     out.data_ptr()))
    return out
