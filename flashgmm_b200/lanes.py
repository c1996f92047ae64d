"""Several batches in flight on one GPU, through the MODEL-facing API.

One coding call is a chain of dependent steps (ten pass-sequential decode calls for ELIC, each bounded by its longest
stream's serial rANS walk, with the small networks in between): it leaves most of the GPU idle, and only a second
independent batch can use the rest. A ``Lane`` gives ONE host thread what it needs to run
``latent_codec.compress / decompress`` (or the batch API) beside other threads doing the same on the same model:

* its own library context (a ctx serialises its calls and owns the record buffers, streams and worker threads);
* its own ``NetRunner`` (captured CUDA graphs have static input / output buffers);
* its own torch stream (work of different lanes must not queue behind each other on the default stream).

The model's weights and tables are shared and only read. Encode sessions are per thread already. Usage::

    lanes = [flashgmm_b200.Lane(device=0) for _ in range(4)]
    def worker(lane, batches):
        with lane:                                   # binds ctx, runner and stream to THIS thread
            for y in batches:
                out = net.latent_codec.compress(y)
                ...
    # one thread per lane; run each lane once on its own before going parallel so that its graphs exist

No counterpart in the reference (one image per call, everything on one thread and the default stream).
"""
from __future__ import annotations

import torch

from . import _lib
from .graphs import NetRunner, bind_thread_runner


class Lane:
    def __init__(self, device: int = 0, runner: NetRunner = None):
        self.device = int(device)
        self.ctx = _lib.new_ctx(self.device)
        self.runner = runner or NetRunner()
        self.stream = torch.cuda.Stream(device=self.device)
        self._saved = None

    def __enter__(self) -> "Lane":
        if self._saved is not None:
            raise RuntimeError("a Lane is entered by one thread at a time")
        dev = torch.device("cuda", self.device)
        self.stream.wait_stream(torch.cuda.current_stream(dev))      # inputs produced on the caller's stream so far
        scope = torch.cuda.stream(self.stream)
        scope.__enter__()
        self._saved = (_lib.bind_thread_ctx(self.device, self.ctx), bind_thread_runner(self.runner), scope)
        return self

    def __exit__(self, *exc):
        prev_ctx, prev_runner, scope = self._saved
        self._saved = None
        self.stream.synchronize()
        scope.__exit__(*exc)
        _lib.bind_thread_ctx(self.device, prev_ctx)
        bind_thread_runner(prev_runner)
        return False
