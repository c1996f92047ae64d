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
import warnings
from typing import Callable, Dict, Sequence, Tuple

import torch
from torch import Tensor


class _Captured:
    def __init__(self, fn: Callable[..., Tuple[Tensor, ...]], example: Sequence[Tensor]):
        self.inputs = [t.detach().clone() for t in example]
        side = torch.cuda.Stream(device=example[0].device)
        side.wait_stream(torch.cuda.current_stream(example[0].device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):                                # library warm-up outside the capture
                fn(*self.inputs)
        torch.cuda.current_stream(example[0].device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.outputs = fn(*self.inputs)

    def __call__(self, *tensors: Tensor) -> Tuple[Tensor, ...]:
        for dst, src in zip(self.inputs, tensors):
            dst.copy_(src)
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
        self._cache: Dict[tuple, object] = {}
        self.replays = 0
        self.captures = 0

    def clear(self) -> None:
        self._cache.clear()

    def _one(self, key, fn, tensors):
        if not self.graphs or not tensors[0].is_cuda or torch.is_grad_enabled() and any(t.requires_grad for t in tensors):
            with torch.no_grad():
                return fn(*tensors)
        sig = (key, tuple((tuple(t.shape), t.dtype, t.device) for t in tensors))
        g = self._cache.get(sig)
        if g is None:
            try:
                g = _Captured(fn, tensors)
                self.captures += 1
            except Exception as e:                            # capture is an optimisation of the same CUDA work
                warnings.warn(f"flashgmm_b200: CUDA-graph capture failed ({e}); running this step eagerly")
                g = False
            self._cache[sig] = g
        if g is False:
            with torch.no_grad():
                return fn(*tensors)
        self.replays += 1
        return g(*tensors)

    def __call__(self, key, fn: Callable[..., Tuple[Tensor, ...]], *inputs: Tensor) -> Tuple[Tensor, ...]:
        inputs = tuple(t.contiguous() for t in inputs)
        B = inputs[0].size(0)
        if self.mode == "batched" or B == 1:
            outs = self._one(key, fn, inputs)
            return tuple(o.clone() for o in outs) if self.graphs else tuple(outs)
        acc = None
        for b in range(B):
            outs = self._one(key, fn, tuple(t[b:b + 1] for t in inputs))
            if acc is None:
                acc = [o.new_empty((B,) + tuple(o.shape[1:])) for o in outs]
            for a, o in zip(acc, outs):
                a[b:b + 1].copy_(o)                           # before the next replay overwrites the graph's outputs
        return tuple(acc)


_default = None


def default_runner() -> NetRunner:
    global _default
    if _default is None:
        _default = NetRunner()
    return _default
