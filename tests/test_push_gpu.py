"""Batch encode against the oracle's restatement of
BufferedRansEncoder::encode_with_indexes_gmm<4> + flush (rans_interface.cpp:458-585), through both
serial halves the batch call can take -- FLASHGMM_PUSH=device: batched startfreq kernel +
one-warp-per-stream push kernel (csrc/gmm_push.cu); FLASHGMM_PUSH=host: batched startfreq waves +
worker-thread pushes (flashgmm_abi.cu::encode_batch_waves) -- and the default split of the two:
byte-identical streams, bypass payloads included, and the ENOSPC contract of the C ABI."""
import ctypes as C

import numpy as np
import pytest
import torch

from flashgmm_b200 import _lib, synth
from oracle import port

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["device", "host"])
def push_ctx(monkeypatch, request):
    """A private ctx whose batch encodes always take the device path / the host waves path."""
    made = []

    def make():
        monkeypatch.setenv("FLASHGMM_PUSH", request.param)
        h = C.c_void_p()
        _lib.check(_lib.lib().flashgmm_ctx_create(0, C.byref(h)))
        made.append(h)
        return h

    yield make
    for h in made:
        _lib.lib().flashgmm_ctx_destroy(h)


def kmajor(a, device):
    t = torch.from_numpy(np.ascontiguousarray(a.T))
    t = t.cuda() if device == "cuda" else t.pin_memory()
    return t.t()


def encode_batch(h, data, mode, device="cuda", cap=None):
    n = len(data)
    arr = (_lib.Stream * max(1, n))()
    keep, outs = [], []
    for i, (sym, sg, mu, w) in enumerate(data):
        s = torch.from_numpy(np.ascontiguousarray(sym))
        s = s.cuda() if device == "cuda" else s.pin_memory()
        a, m, ww = (kmajor(x, device) for x in (sg, mu, w))
        out = np.zeros(cap if cap is not None else 44 * sym.size + 64, np.uint8)
        keep += [s, a, m, ww]
        outs.append(out)
        arr[i].symbols, arr[i].scales, arr[i].means, arr[i].weights = s.data_ptr(), a.data_ptr(), m.data_ptr(), ww.data_ptr()
        arr[i].n, arr[i].stride_n, arr[i].stride_k = sym.size, a.stride(0), a.stride(1)
        arr[i].bytes, arr[i].bytes_cap = out.ctypes.data, out.size
    torch.cuda.synchronize()
    rc = _lib.lib().flashgmm_gmm_encode_batch(h, arr, n, mode, _lib.MEM_DEVICE if device == "cuda" else _lib.MEM_HOST)
    return [outs[i][: arr[i].bytes_len].tobytes() for i in range(n)], [arr[i].status for i in range(n)], \
        [arr[i].bytes_len for i in range(n)], rc


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("device", ["cuda", "cpu"])
def test_elic_shaped_batch_is_byte_identical(push_ctx, mode, device):
    sizes = [12288, 12288, 12288, 12288, 24576, 24576, 49152, 49152, 147456, 147456, 0, 1, 5, 33, 4099]
    data = [synth.gmm_numpy(n, seed=1300 + i) for i, n in enumerate(sizes)]
    got, status, _lens, rc = encode_batch(push_ctx(), data, mode, device)
    assert rc == 0 and not any(status)
    for g, (sym, sg, mu, w) in zip(got, data):
        assert g == port.gmm_encode(sym, sg, mu, w, mode)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_bypass_payloads_and_int32_extremes(push_ctx, mode):
    data = []
    sym, sg, mu, w = synth.gmm_numpy(5000, seed=1401, sigma_span=0.0, offcentre=300.0)      # every symbol bypassed
    sym[:8] = [2**31 - 1, -2**31, 0, 1, -1, 65535, -65536, 123456789]
    data.append((sym, sg, mu, w))
    sym, sg, mu, w = synth.gmm_numpy(20000, seed=1402, sigma_span=0.3, offcentre=2.5)       # a mix
    data.append((sym, sg, mu, w))
    r = np.random.default_rng(1403 + mode)
    n = 30000
    mu2 = (r.standard_normal((4, n)) * 10 ** r.uniform(-3, 2, (4, n))).astype(np.float32)
    sg2 = np.clip(10 ** r.uniform(-1, 2.5, (4, n)), 0.11, 256).astype(np.float32)
    w2 = np.ascontiguousarray(r.dirichlet(np.ones(4), n).T, dtype=np.float32)
    sym2 = np.round(mu2[0] + r.standard_normal(n) * sg2[0]).astype(np.int32)
    data.append((sym2, sg2.T, mu2.T, w2.T))
    data.append(synth.gmm_numpy(7, seed=1404))
    sym, sg, mu, w = synth.gmm_numpy(40000, seed=1405)                                      # a handful of bypasses per wave
    sym[::5003] = 30000
    data.append((sym, sg, mu, w))
    got, status, _lens, rc = encode_batch(push_ctx(), data, mode)
    assert rc == 0 and not any(status)
    for g, (sym, sg, mu, w) in zip(got, data):
        assert g == port.gmm_encode(sym, sg, mu, w, mode)


def test_output_buffer_too_small_reports_the_needed_size(push_ctx):
    data = [synth.gmm_numpy(20000 + i, seed=1500 + i) for i in range(4)]
    want = [port.gmm_encode(*d, 0) for d in data]
    h = push_ctx()
    got, status, lens, rc = encode_batch(h, data, 0, cap=len(want[0]) - 400)
    assert rc == _lib.ENOSPC and all(s == _lib.ENOSPC for s in status)
    assert lens == [len(b) for b in want]                       # the size a retry must offer
    got, status, lens, rc = encode_batch(h, data, 0, cap=max(len(b) for b in want))
    assert rc == 0 and got == want


@pytest.mark.parametrize("device", ["cuda", "cpu"])
def test_default_split_of_an_image_batch(device):
    """Default ctx, two ELIC-shaped images: the cost model splits the streams between the worker
    threads and the push kernel; the bytes are the oracle's."""
    sizes = [12288, 12288, 12288, 12288, 24576, 24576, 49152, 49152, 147456, 147456] * 2
    data = [synth.gmm_numpy(n, seed=1700 + i) for i, n in enumerate(sizes)]
    h = C.c_void_p()
    _lib.check(_lib.lib().flashgmm_ctx_create(0, C.byref(h)))
    try:
        got, status, _lens, rc = encode_batch(h, data, 0, device)
    finally:
        _lib.lib().flashgmm_ctx_destroy(h)
    assert rc == 0 and not any(status)
    for g, (sym, sg, mu, w) in zip(got, data):
        assert g == port.gmm_encode(sym, sg, mu, w, 0)


def test_auto_policy_on_a_small_batch():
    """Default ctx, 6 equal streams: whatever the cost model decides, the bytes are the oracle's."""
    from flashgmm_b200 import batch
    data = [synth.gmm_numpy(9000 + i, seed=1600 + i) for i in range(6)]
    streams = [(torch.from_numpy(sym).cuda(), kmajor(sg, "cuda"), kmajor(mu, "cuda"), kmajor(w, "cuda"))
               for sym, sg, mu, w in data]
    got = batch.encode_batch(streams, 0)
    assert got == [port.gmm_encode(*d, 0) for d in data]


def test_pageable_host_tensors_fall_back_to_the_per_stream_path(push_ctx):
    """Pageable (not page-locked) host tensors cannot be read in place: both batch halves hand the
    streams to the staging path; same bytes."""
    data = [synth.gmm_numpy(6000 + 37 * i, seed=1800 + i) for i in range(5)]
    n = len(data)
    arr = (_lib.Stream * n)()
    keep, outs = [], []
    for i, (sym, sg, mu, w) in enumerate(data):
        s = np.ascontiguousarray(sym)
        a, m, ww = (np.ascontiguousarray(x.T) for x in (sg, mu, w))          # K-major planes, pageable
        out = np.zeros(44 * sym.size + 64, np.uint8)
        keep += [s, a, m, ww]
        outs.append(out)
        arr[i].symbols, arr[i].scales, arr[i].means, arr[i].weights = (x.ctypes.data for x in (s, a, m, ww))
        arr[i].n, arr[i].stride_n, arr[i].stride_k = sym.size, 1, sym.size
        arr[i].bytes, arr[i].bytes_cap = out.ctypes.data, out.size
    rc = _lib.lib().flashgmm_gmm_encode_batch(push_ctx(), arr, n, 0, _lib.MEM_HOST)
    assert rc == 0
    for i, d in enumerate(data):
        assert outs[i][: arr[i].bytes_len].tobytes() == port.gmm_encode(*d, 0)


def test_one_worker_thread(monkeypatch):
    """FLASHGMM_WORKERS=1: the wave tasks of all streams run on a single worker, in order."""
    monkeypatch.setenv("FLASHGMM_WORKERS", "1")
    monkeypatch.setenv("FLASHGMM_PUSH", "host")
    h = C.c_void_p()
    _lib.check(_lib.lib().flashgmm_ctx_create(0, C.byref(h)))
    try:
        data = [synth.gmm_numpy(n, seed=1900 + i) for i, n in enumerate([30000, 1, 9000, 0, 17000, 40000])]
        got, status, _lens, rc = encode_batch(h, data, 0)
    finally:
        _lib.lib().flashgmm_ctx_destroy(h)
    assert rc == 0 and not any(status)
    for g, d in zip(got, data):
        assert g == port.gmm_encode(*d, 0)


@pytest.mark.parametrize("queues", ["8", "32"])
def test_batch_paths_with_either_hardware_queue_count(queues):
    """the library no longer sets CUDA_DEVICE_MAX_CONNECTIONS on load. Under the reference's scripts the
    CUDA context predates the library (8 queues, the default -- which is also what every other test here runs with);
    bench.py asks for 32 before CUDA starts. Both configurations must give the oracle's bytes: smoke() in a child."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parents[1]
    env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS=queues)
    code = ("import torch; torch.zeros(1).cuda(); import os; assert os.environ['CUDA_DEVICE_MAX_CONNECTIONS'] == '%s'; "
            "import __graft_entry__ as g; g.smoke()" % queues)
    r = subprocess.run([sys.executable, "-c", code], cwd=str(root), env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "smoke ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]
