"""ctypes binding of ``libflashgmm_b200.so`` (the C ABI in ``include/flashgmm.h``).

There is no CPU fallback: if the shared library is missing, or no CUDA device can
be opened, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path



def want_more_hw_queues(n: int = 32) -> None:
    """Opt-in (bench.py and standalone tools call it before CUDA is initialised): ~20 CUDA streams are busy at
    once (worker threads, device encoder / decoder); with the driver's default of 8 hardware queues they share
    queues and a copy queued behind a long kernel stalls its neighbours. The setting is read when the CUDA
    context is created, so it has no effect in a process whose context already exists (the library then runs
    with the default; tests/test_queues_gpu.py covers that configuration). A value set by the user wins."""
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", str(int(n)))


PKG = Path(__file__).resolve().parent
LIB_PATH = PKG / "libflashgmm_b200.so"

OK, EINVAL, ECUDA, ETRUNC, ENOSPC = 0, 1, 2, 3, 4
MEM_HOST, MEM_DEVICE = 0, 1

_vp, _i64, _i32, _sz = C.c_void_p, C.c_int64, C.c_int32, C.c_size_t


class Stream(C.Structure):
    """Mirror of ``flashgmm_stream`` (include/flashgmm.h)."""
    _fields_ = [
        ("symbols", _vp), ("scales", _vp), ("means", _vp), ("weights", _vp),
        ("n", _i64), ("stride_n", _i64), ("stride_k", _i64), ("max_bs_value", _i32),
        ("bytes", _vp), ("bytes_cap", _sz), ("bytes_len", _sz), ("symbols_out", _vp),
        ("status", C.c_int),
    ]


_SIGS = {
    "flashgmm_version": (C.c_int, []),
    "flashgmm_request_hw_queues": (C.c_int, [C.c_int]),
    "flashgmm_last_error": (C.c_char_p, []),
    "flashgmm_launch_count": (C.c_uint64, []),
    "flashgmm_max_encoded_bytes": (_sz, [_i64]),
    "flashgmm_profile_enable": (None, [C.c_int]),
    "flashgmm_profile_reset": (None, []),
    "flashgmm_profile_read": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]),
    "flashgmm_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "flashgmm_ctx_destroy": (None, [_vp]),
    "flashgmm_gmm_startfreq": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, _vp, _vp]),
    "flashgmm_gmm_table_row_stride": (C.c_int, [_i32]),
    "flashgmm_gmm_cdf_table": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, _i32, _vp, _i64, _vp]),
    "flashgmm_gmm_encode": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, C.c_int, _vp, _sz,
                                      C.POINTER(_sz)]),
    "flashgmm_gmm_decode": (C.c_int, [_vp, _vp, _sz, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, C.c_int, _i32, _vp]),
    "flashgmm_gmm_encode_batch": (C.c_int, [_vp, C.POINTER(Stream), C.c_int, C.c_int, C.c_int]),
    "flashgmm_gmm_decode_batch": (C.c_int, [_vp, C.POINTER(Stream), C.c_int, C.c_int, C.c_int]),
    "flashgmm_gmm_ileave_max_bytes": (_sz, [_i64, C.c_int]),
    "flashgmm_gmm_ileave_encode": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, C.c_int, _i32, C.c_int,
                                             _vp, _sz, C.POINTER(_sz)]),
    "flashgmm_gmm_ileave_info": (C.c_int, [_vp, _sz, C.POINTER(_i64), C.POINTER(C.c_int), C.POINTER(_i32),
                                           C.POINTER(C.c_int)]),
    "flashgmm_gmm_ileave_decode": (C.c_int, [_vp, _vp, _sz, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, _vp]),
    "flashgmm_gmm_likelihood_fwd": (C.c_int, [_vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _i64, _i64, C.c_int, C.c_float,
                                              C.c_float, _vp, _vp, _vp]),
    "flashgmm_gmm_likelihood_bwd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, C.c_int, C.c_float,
                                              C.c_float, _vp, _vp, _vp, _vp, _vp]),
    "flashgmm_gmm_prologue_scan": (C.c_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, C.POINTER(C.c_float),
                                             C.POINTER(C.c_float), _vp]),
    "flashgmm_gmm_prologue_gather": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, C.c_int, _i64, _vp, _vp, _vp,
                                               _vp, _vp]),
    "flashgmm_gmm_epilogue_scatter": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "flashgmm_gmm_prologue_scan_batch": (C.c_int, [_vp, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "flashgmm_gmm_prologue_scan_batch_noalloc": (C.c_int, [C.c_int, _vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "flashgmm_gmm_prologue_gather_batch": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, C.c_int, _vp, _vp, _vp,
                                                     _vp, _vp, _vp]),
    "flashgmm_gmm_epilogue_scatter_batch": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "flashgmm_eb_compress_batch": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _i64, _vp,
                                             C.POINTER(Stream), _vp]),
    "flashgmm_eb_decompress_batch": (C.c_int, [_vp, C.POINTER(Stream), _i64, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _vp]),
    "flashgmm_eb_likelihood": (C.c_int, [_vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "flashgmm_scalar_gaussian_encode": (C.c_int, [_vp, _vp, _i64, C.c_int, _vp, _sz, C.POINTER(_sz)]),
    "flashgmm_scalar_gaussian_decode": (C.c_int, [_vp, _sz, _vp, _i64, C.c_int, _i32, _vp]),
    "flashgmm_rans_push_packed": (C.c_int, [_vp, _vp, _i64, _vp, _sz, C.POINTER(_sz)]),
    "flashgmm_rans_pop_table": (C.c_int, [_vp, _sz, _vp, _i64, _i64, _i32, _vp]),
    "flashgmm_rans_encode_table": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _sz, C.POINTER(_sz)]),
    "flashgmm_rans_decode_table": (C.c_int, [_vp, _sz, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _vp]),
    "flashgmm_benc_create": (C.c_int, [C.POINTER(_vp)]),
    "flashgmm_benc_destroy": (None, [_vp]),
    "flashgmm_benc_entries": (_i64, [_vp]),
    "flashgmm_benc_push_table": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp]),
    "flashgmm_benc_push_gmm_packed": (C.c_int, [_vp, _vp, _vp, _i64]),
    "flashgmm_benc_flush": (C.c_int, [_vp, _vp, _sz, C.POINTER(_sz)]),
    "flashgmm_sdec_create": (C.c_int, [_vp, _sz, C.POINTER(_vp)]),
    "flashgmm_sdec_destroy": (None, [_vp]),
    "flashgmm_sdec_decode_table": (C.c_int, [_vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _vp]),
}

_lib = None
_lock = threading.RLock()   # ctx() calls lib() while holding it
_ctxs: dict = {}


class FlashGMMError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"flashgmm status {code}: {msg}")
        self.code = code


def lib():
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not LIB_PATH.exists():
                    raise ImportError(
                        f"{LIB_PATH} is missing: build it with `python -m flashgmm_b200.build` "
                        "(needs nvcc; flashgmm_b200 has no CPU fallback)")
                L = C.CDLL(str(LIB_PATH))
                for name, (res, args) in _SIGS.items():
                    fn = getattr(L, name)   # AttributeError if the ABI and header drift apart
                    fn.restype = res
                    fn.argtypes = args
                _lib = L
    return _lib


def check(code: int):
    if code == OK:
        return
    msg = lib().flashgmm_last_error().decode("utf-8", "replace")
    if code == EINVAL:
        raise ValueError(f"flashgmm: {msg}")
    raise FlashGMMError(code, msg)


_tls = threading.local()


def ctx(device: int = 0):
    """Per-device context handle (created on first use, kept for the process). A thread inside a
    ``flashgmm_b200.Lane`` gets the lane's own ctx instead (``bind_thread_ctx``)."""
    own = getattr(_tls, "ctxs", None)
    if own:
        h = own.get(device)
        if h is not None:
            return h
    h = _ctxs.get(device)
    if h is None:
        with _lock:
            h = _ctxs.get(device)
            if h is None:
                out = _vp()
                check(lib().flashgmm_ctx_create(int(device), C.byref(out)))
                h = _ctxs[device] = out
    return h


def new_ctx(device: int = 0):
    """A further context on ``device`` (own streams, buffers and worker threads): a ctx serialises its calls, so
    concurrent host threads -- e.g. batch k+1's encode beside batch k's decode -- use one ctx each. Freed at exit."""
    out = _vp()
    check(lib().flashgmm_ctx_create(int(device), C.byref(out)))
    _extra_ctxs.append(out)
    return out


_extra_ctxs: list = []


def bind_thread_ctx(device: int, handle) -> object:
    """Make ``ctx(device)`` return ``handle`` on the CALLING thread (``None`` removes the binding).
    -> the previous binding, to restore."""
    own = getattr(_tls, "ctxs", None)
    if own is None:
        own = _tls.ctxs = {}
    prev = own.get(device)
    if handle is None:
        own.pop(device, None)
    else:
        own[device] = handle
    return prev


def launch_count() -> int:
    return int(lib().flashgmm_launch_count())
