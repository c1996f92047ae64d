"""CUDA-graph replay of the small networks between two coding passes.

Between pass k and pass k+1 of a checkerboard / channel-group codec the reference runs a 5x5 masked
context convolution, three 1x1 convolutions, a softmax and a dozen slicing ops eagerly
(/root/reference/compressai/latent_codecs/checkerboard.py:275-330, channel_groups.py:111-158,
gaussian_mixture_conditional.py:198-202): ~40 launches of a few microseconds of work each, per image
and pass. ``NetRunner`` captures each such step once per input shape in a ``torch.cuda.CUDAGraph``
and replays it, so a pass costs one launch of the whole step.

Two batching modes, because conv / GEMM libraries may pick different kernels (different summation
orders) for different batch sizes, and the decoder must reproduce the encoder's parameters bit for bit:

* ``per_image`` (default): the step is captured at batch 1 and replayed once per image -- image b's
  parameters, and therefore its bitstream, are exactly those of coding image b alone;
* ``batched``: the step runs once on the whole batch (fewer launches; encoder and decoder must then
  use the same batch size).
"""
from __future__ import annotations

import os
import threading
import warnings
from typing import Callable, Dict, List, Sequence, Tuple

import torch
from torch import Tensor


_side_streams: Dict[tuple, List["torch.cuda.Stream"]] = {}


def _branch_streams(dev: torch.device, n: int) -> List["torch.cuda.Stream"]:
    """Side streams the images of a group are captured on (shared by all captures of a device made by one host thread:
    they only carry work while a capture or its warm-up is in progress, and a stream can be part of one capture only)."""
    pool = _side_streams.setdefault((dev, threading.get_ident()), [])
    while len(pool) < n:
        pool.append(torch.cuda.Stream(device=dev))
    return pool[:n]


class _Captured:
    """One captured instance of a step: static inputs / outputs for `group` images, its graph, and the side stream it
    replays on. With group > 1 the graph contains the step `group` times, each on ONE image (batch-1 shapes, so every
    image gets exactly the kernels and summation orders of coding it alone), as `group` PARALLEL branches: the images
    are captured on side streams forked from and joined to the capture stream, so a replay runs `group` independent
    chains of small kernels side by side instead of one chain `group` times as long."""

    def __init__(self, fn: Callable[..., Tuple[Tensor, ...]], example: Sequence[Tensor], group: int = 0):
        dev = example[0].device
        self.group = group
        self.inputs = [t.detach().clone() for t in example]
        self.stream = torch.cuda.Stream(device=dev)
        branches = _branch_streams(dev, group) if group > 1 else []

        def run():
            if group <= 0:
                return fn(*self.inputs)
            if group == 1:
                return fn(*(t[0:1] for t in self.inputs))
            main = torch.cuda.current_stream(dev)
            per = []
            for g, side in enumerate(branches):               # fork: one branch per image
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    per.append(fn(*(t[g:g + 1] for t in self.inputs)))
            for side in branches:                             # join
                main.wait_stream(side)
            return tuple(torch.cat([o[j] for o in per], dim=0) for j in range(len(per[0])))

        self.stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self.stream), torch.no_grad():
            for _ in range(2):                                # library warm-up outside the capture
                run()
        torch.cuda.current_stream(dev).wait_stream(self.stream)
        torch.cuda.synchronize(dev)                           # the branch streams are idle before the capture starts
        self.graph = torch.cuda.CUDAGraph()
        # "thread_local": other host threads (further lanes, the library's workers) may allocate while this one captures
        with torch.no_grad(), torch.cuda.graph(self.graph, stream=self.stream, capture_error_mode="thread_local"):
            self.outputs = run()

    def __call__(self, *tensors: Tensor) -> Tuple[Tensor, ...]:
        n = tensors[0].size(0)
        for dst, src in zip(self.inputs, tensors):
            dst[:n].copy_(src)
        self.graph.replay()
        return self.outputs


class NetRunner:
    """``runner(key, fn, *inputs) -> tuple of tensors``: ``fn`` maps batched tensors to a tuple of batched
    tensors, is a pure function of its inputs and of module weights, and is only used without autograd."""

    def __init__(self, mode: str = None, graphs: bool = None):
        mode = mode or os.environ.get("FLASHGMM_NET_BATCHING", "per_image")
        if mode not in ("per_image", "batched"):
            raise ValueError('net batching mode must be "per_image" or "batched"')
        self.mode = mode
        self.graphs = (os.environ.get("FLASHGMM_GRAPHS", "1") != "0") if graphs is None else bool(graphs)
        self.lanes = max(1, int(os.environ.get("FLASHGMM_NET_LANES", "4")))   # per-image replays in flight at once
        self._cache: Dict[tuple, object] = {}
        self.replays = 0
        self.captures = 0

    def clear(self) -> None:
        self._cache.clear()

    def _one(self, key, fn, tensors):
        if not self.graphs or not tensors[0].is_cuda or torch.is_grad_enabled() and any(t.requires_grad for t in tensors):
            with torch.no_grad():
                return fn(*tensors)
        pool = self._instances(key, fn, tensors, 1)
        if not pool:
            with torch.no_grad():
                return fn(*tensors)
        self.replays += 1
        return pool[0](*tensors)

    def _instances(self, key, fn, tensors, want: int, group: int = 0):
        """Up to `want` captured instances of the step for this input signature (each replays on its own stream)."""
        sig = (key, group, tuple((tuple(t.shape), t.dtype, t.device) for t in tensors))
        pool = self._cache.get(sig)
        if pool is None:
            pool = self._cache[sig] = []
        while pool is not False and len(pool) < want:
            try:
                pool.append(_Captured(fn, tensors, group))
                self.captures += 1
            except Exception as e:                            # capture is an optimisation of the same CUDA work
                warnings.warn(f"flashgmm_b200: CUDA-graph capture failed ({e}); running this step eagerly")
                pool = self._cache[sig] = False
        return pool

    def __call__(self, key, fn: Callable[..., Tuple[Tensor, ...]], *inputs: Tensor) -> Tuple[Tensor, ...]:
        inputs = tuple(t.contiguous() for t in inputs)
        B = inputs[0].size(0)
        if self.mode == "batched" or B == 1:
            outs = self._one(key, fn, inputs)
            return tuple(o.clone() for o in outs) if self.graphs else tuple(outs)
        eager = (not self.graphs or not inputs[0].is_cuda
                 or (torch.is_grad_enabled() and any(t.requires_grad for t in inputs)))
        group = -(-B // self.lanes)                           # images per replay
        n_inst = -(-B // group)
        pool = None if eager else self._instances(key, fn, tuple(t[0:group] for t in inputs), n_inst, group)
        acc = None
        if not pool:                                          # eager, image by image
            for b in range(B):
                with torch.no_grad():
                    outs = fn(*(t[b:b + 1] for t in inputs))
                if acc is None:
                    acc = [o.new_empty((B,) + tuple(o.shape[1:])) for o in outs]
                for a, o in zip(acc, outs):
                    a[b:b + 1].copy_(o)
            return tuple(acc)
        # images [i * group, (i + 1) * group) replay on instance i: the small kernels of several images overlap on the
        # GPU, and one replay (one launch from Python) serves `group` images
        cur = torch.cuda.current_stream(inputs[0].device)
        acc = [o.new_empty((B,) + tuple(o.shape[1:])) for o in pool[0].outputs]
        for i, inst in enumerate(pool[:n_inst]):
            b0, b1 = i * group, min(B, (i + 1) * group)
            inst.stream.wait_stream(cur)
            with torch.cuda.stream(inst.stream):
                outs = inst(*(t[b0:b1] for t in inputs))
                for a, o in zip(acc, outs):
                    a[b0:b1].copy_(o[: b1 - b0])
            self.replays += 1
        for inst in pool[:n_inst]:
            cur.wait_stream(inst.stream)
            for t in list(inputs) + acc:                      # the side streams touch them: keep the allocator from reusing them early
                t.record_stream(inst.stream)
        return tuple(acc)


_default = None
_tls = threading.local()


def default_runner() -> NetRunner:
    """The runner of modules that were not given one by ``mount(net, runner=...)``: the calling thread's
    (``bind_thread_runner``, used by ``flashgmm_b200.Lane``) or the process-wide one. A runner's captured graphs have
    static input / output buffers, so two host threads must never replay through the same runner."""
    own = getattr(_tls, "runner", None)
    if own is not None:
        return own
    global _default
    if _default is None:
        _default = NetRunner()
    return _default


def thread_runner():
    """The runner bound to the calling thread, or None."""
    return getattr(_tls, "runner", None)


def bind_thread_runner(runner):
    """Make ``default_runner()`` return ``runner`` on the CALLING thread (``None`` removes the binding). -> the previous one."""
    prev = getattr(_tls, "runner", None)
    _tls.runner = runner
    return prev
