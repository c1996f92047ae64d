"""Many independent GMM streams per call (one per checkerboard pass, channel group,
image or tile): ``flashgmm_gmm_encode_batch`` / ``flashgmm_gmm_decode_batch``.

The kernels of all streams are queued back to back and the serial host rANS of
the streams runs on the context's worker threads. Each stream's bytes are
identical to ``RansEncoder.encode_with_indexes_gmm`` on that stream alone.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import List, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .ans import _dev_index, _gmm_views, _tensor, get_approx_mode


_tls = threading.local()


def _arena(nbytes: int) -> np.ndarray:
    """Per-thread landing zone for the bitstreams of a one-shot ``encode_batch`` (its results leave as ``bytes``
    copies, so the next call may write over it): fresh ``np.empty`` buffers would be page-faulted in by the push
    workers on every call."""
    a = getattr(_tls, "arena", None)
    if a is None or a.size < nbytes:
        a = _tls.arena = np.empty(max(nbytes, 1 << 20) * 5 // 4, np.uint8)
    return a


class EncodeJob:
    """Pre-bound encode call over fixed tensors (re-usable: no per-call allocation)."""

    def __init__(self, streams: Sequence[Tuple], mode=None, ctx=None, shared_arena: bool = False):
        self.mode = get_approx_mode() if mode is None else mode
        self.n_streams = len(streams)
        self.arr = (_lib.Stream * max(1, self.n_streams))()
        self._keep = []
        self.outs: List[np.ndarray] = []
        self.device = None
        arena, used = None, 0
        if shared_arena:
            arena = _arena(sum(max(64, (int(st[0].numel()) * 9) // 4 + 64) + 64 for st in streams))
        for i, (sym, sc, mu, w) in enumerate(streams):
            sym = _tensor(sym, torch.int32).contiguous()
            (sc, mu, w), N, (sn, sk), dev = _gmm_views(sc, mu, w, sym.numel())
            if sym.device != dev:
                raise ValueError("symbols and parameters must be on one device")
            if self.device is None:
                self.device = dev
            elif dev != self.device:
                raise ValueError("all streams of a batch must be on one device")
            cap = max(64, (N * 9) // 4 + 64)
            if arena is None:
                out = np.empty(cap, np.uint8)
            else:
                out = arena[used: used + cap]
                used += (cap + 63) & ~63
            self._keep += [sym, sc, mu, w]
            self.outs.append(out)
            a = self.arr[i]
            a.symbols, a.scales, a.means, a.weights = sym.data_ptr(), sc.data_ptr(), mu.data_ptr(), w.data_ptr()
            a.n, a.stride_n, a.stride_k = N, sn, sk
            a.bytes, a.bytes_cap = out.ctypes.data, out.size
        self.mem = _lib.MEM_HOST if (self.device is None or self.device.type == "cpu") else _lib.MEM_DEVICE
        self.ctx = ctx if ctx is not None else _lib.ctx(_dev_index(self.device if self.device is not None else torch.device("cpu")))

    def run(self) -> List[np.ndarray]:
        """-> per-stream uint8 views of the produced bitstreams (valid until the next run)."""
        L = _lib.lib()
        rc = L.flashgmm_gmm_encode_batch(self.ctx, self.arr, self.n_streams, self.mode, self.mem)
        if rc == _lib.ENOSPC:      # pathological bypass-heavy stream: grow and retry once
            for i in range(self.n_streams):
                if self.arr[i].status == _lib.ENOSPC:
                    self.outs[i] = np.empty(self.arr[i].bytes_len + 64, np.uint8)
                    self.arr[i].bytes, self.arr[i].bytes_cap = self.outs[i].ctypes.data, self.outs[i].size
            rc = L.flashgmm_gmm_encode_batch(self.ctx, self.arr, self.n_streams, self.mode, self.mem)
        _lib.check(rc)
        return [self.outs[i][: self.arr[i].bytes_len] for i in range(self.n_streams)]


class DecodeJob:
    """Pre-bound decode call: parameters and output tensors fixed, bitstreams per run."""

    def __init__(self, streams: Sequence[Tuple], mode=None, ctx=None):
        """streams: (scales, means, weights, max_bs_value) per stream."""
        self.mode = get_approx_mode() if mode is None else mode
        self.n_streams = len(streams)
        self.arr = (_lib.Stream * max(1, self.n_streams))()
        self._keep = []
        self.outs: List[torch.Tensor] = []
        self.device = None
        for i, (sc, mu, w, R) in enumerate(streams):
            (sc, mu, w), N, (sn, sk), dev = _gmm_views(sc, mu, w)
            if self.device is None:
                self.device = dev
            elif dev != self.device:
                raise ValueError("all streams of a batch must be on one device")
            out = torch.empty(N, dtype=torch.int32, device=dev, pin_memory=(dev.type == "cpu" and torch.cuda.is_available()))
            self._keep += [sc, mu, w]
            self.outs.append(out)
            a = self.arr[i]
            a.scales, a.means, a.weights = sc.data_ptr(), mu.data_ptr(), w.data_ptr()
            a.n, a.stride_n, a.stride_k, a.max_bs_value = N, sn, sk, int(R)
            a.symbols_out = out.data_ptr()
        self.mem = _lib.MEM_HOST if (self.device is None or self.device.type == "cpu") else _lib.MEM_DEVICE
        self.ctx = ctx if ctx is not None else _lib.ctx(_dev_index(self.device if self.device is not None else torch.device("cpu")))

    def run(self, bitstreams: Sequence) -> List[torch.Tensor]:
        if len(bitstreams) != self.n_streams:
            raise ValueError(f"{len(bitstreams)} bitstreams for a job of {self.n_streams} streams")
        bufs = []
        for i, b in enumerate(bitstreams):
            buf = b if isinstance(b, np.ndarray) else np.frombuffer(b, np.uint8)
            bufs.append(buf)
            self.arr[i].bytes, self.arr[i].bytes_cap = buf.ctypes.data, buf.size
        _lib.check(_lib.lib().flashgmm_gmm_decode_batch(self.ctx, self.arr, self.n_streams, self.mode, self.mem))
        return self.outs


def _sync_inputs(dev):
    if dev is not None and dev.type == "cuda":
        torch.cuda.current_stream(dev).synchronize()


def encode_batch(streams: Sequence[Tuple], mode=None) -> List[bytes]:
    """streams: (symbols, scales, means, weights) per stream -> list of bytes."""
    job = EncodeJob(streams, mode, shared_arena=True)
    _sync_inputs(job.device)
    return [v.tobytes() for v in job.run()]


def decode_batch(streams: Sequence[Tuple], mode=None) -> List[torch.Tensor]:
    """streams: (encoded, scales, means, weights, max_bs_value) per stream -> list of int32 tensors."""
    job = DecodeJob([s[1:] for s in streams], mode)
    _sync_inputs(job.device)
    return job.run([s[0] for s in streams])
