"""flashgmm_gmm_prologue_scan_batch_noalloc: the batch scan with its scratch in a file-scope __device__ array
(include/flashgmm.h; no counterpart in the reference).

Collected last on purpose: called through the C ABI and compared with the existing,
allocating entry point flashgmm_gmm_prologue_scan_batch -- integer / rounding work, so the comparison is bit-exact --
on two calls in a row (the second with different data and a smaller shape: the shared array keeps the first call's
contents, and nothing of them may leak into the second result)."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from flashgmm_b200 import _lib  # noqa: E402

MAX_WORDS = 65536                                            # FLASHGMM_NOALLOC_SCAN_MAX_WORDS


def _latent(B, Cn, H, W, seed, dev):
    g = torch.Generator().manual_seed(seed)
    y = torch.randn(B, Cn, H, W, generator=g) * 6.0
    y[:, 1::5] *= 0.01                                       # channels that round to all zeros
    y[0, 0] = 0.0
    if B > 1:
        y[1] -= 40.0                                         # an image whose maximum is negative
    y[B - 1, Cn - 1, H - 1, W - 1] = 0.5                     # round-half-to-even
    return y.to(dev).contiguous()


def _outputs(B, Cn, H, W, dev):
    return (torch.full((B, Cn, H, W), float("nan"), device=dev), torch.full((B, Cn), -7, dtype=torch.int64, device=dev),
            np.full((B, Cn), -7, np.int32), np.full((B, 2), np.nan, np.float32))


def _scan_existing(y):
    B, Cn, H, W = y.shape
    yq, bm_d, bm_h, ext = _outputs(B, Cn, H, W, y.device)
    st = torch.cuda.current_stream(y.device)
    _lib.check(_lib.lib().flashgmm_gmm_prologue_scan_batch(
        _lib.ctx(y.device.index or 0), y.data_ptr(), B, Cn, H * W, yq.data_ptr(), bm_d.data_ptr(),
        bm_h.ctypes.data, ext.ctypes.data, st.cuda_stream))
    return yq, bm_d, bm_h, ext


def _scan_noalloc(y, outs=None):
    B, Cn, H, W = y.shape
    yq, bm_d, bm_h, ext = outs or _outputs(B, Cn, H, W, y.device)
    st = torch.cuda.current_stream(y.device)
    rc = _lib.lib().flashgmm_gmm_prologue_scan_batch_noalloc(
        int(y.device.index or 0), y.data_ptr(), B, Cn, H * W, yq.data_ptr(), bm_d.data_ptr(),
        bm_h.ctypes.data, ext.ctypes.data, st.cuda_stream)
    return rc, (yq, bm_d, bm_h, ext)


def _same(got, want):
    assert torch.equal(got[0], want[0])                      # round(y), bit for bit
    assert torch.equal(got[1], want[1])                      # int64 bitmap on the device
    assert np.array_equal(got[2], want[2])                   # int32 bitmap on the host
    assert got[3].tobytes() == want[3].tobytes()             # max / min of the unrounded y, bit for bit


@pytest.mark.gpu
def test_noalloc_scan_equals_the_allocating_entry_point_twice_in_a_row():
    dev = torch.device("cuda", 0)
    first = _latent(24, 320, 32, 24, seed=1, dev=dev)        # the ELIC batch of the benchmark: 24 * 322 words of scratch
    second = _latent(3, 192, 16, 12, seed=2, dev=dev) + 100.0  # smaller, other extrema: stale scratch would show
    want = [_scan_existing(first), _scan_existing(second)]
    torch.cuda.synchronize()
    # the outputs are allocated before the calls under test: torch's allocator must not move during them
    outs = [_outputs(*first.shape, dev), _outputs(*second.shape, dev), _outputs(*first.shape, dev)]
    torch.cuda.synchronize()
    reserved = torch.cuda.memory_reserved(dev)
    rc1, got1 = _scan_noalloc(first, outs[0])                # a device's first call also resolves the array's address
    free_before, _total = torch.cuda.mem_get_info(dev)       # (and the driver loads the reset kernel's code)
    rc2, got2 = _scan_noalloc(second, outs[1])
    rc3, got3 = _scan_noalloc(first, outs[2])                # the first data again: same answer as before
    free_after, _total = torch.cuda.mem_get_info(dev)
    reserved_after = torch.cuda.memory_reserved(dev)         # read before the comparisons below allocate temporaries
    assert rc1 == 0 and rc2 == 0 and rc3 == 0, _lib.lib().flashgmm_last_error()
    _same(got1, want[0])
    _same(got2, want[1])
    _same(got3, want[0])
    assert reserved_after == reserved
    assert free_after >= free_before                         # no device memory taken by the calls
    # and against the tensor ops the kernel stands for (entropy_models.py:834-845)
    assert torch.equal(got2[0], torch.round(second))
    assert np.array_equal(got2[2] != 0, (torch.round(second).abs().sum((2, 3)) != 0).cpu().numpy())
    assert np.array_equal(got2[3][:, 0], second.amax((1, 2, 3)).cpu().numpy())
    assert np.array_equal(got2[3][:, 1], second.amin((1, 2, 3)).cpu().numpy())


@pytest.mark.gpu
def test_noalloc_scan_rejects_sizes_beyond_the_static_array():
    dev = torch.device("cuda", 0)
    L = _lib.lib()
    B, Cn = 128, 512                                         # 128 * 514 = 65792 words > 65536
    assert B * (Cn + 2) > MAX_WORDS
    y = torch.zeros(B, Cn, 1, 1, device=dev)
    rc, _ = _scan_noalloc(y)
    assert rc == _lib.EINVAL
    msg = L.flashgmm_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "static maximum" in msg and "65536" in msg
    # the largest shape that fits is served: 127 * 514 + ... <= 65536
    y = _latent(127, 514 - 2, 2, 2, seed=3, dev=dev)
    assert 127 * 514 <= MAX_WORDS
    rc, got = _scan_noalloc(y)
    assert rc == 0, L.flashgmm_last_error()
    _same(got, _scan_existing(y))
    # empty batch: nothing to do, nothing touched
    rc = L.flashgmm_gmm_prologue_scan_batch_noalloc(0, None, 0, 192, 64, None, None, None, None, None)
    assert rc == 0
