"""``GaussianMixtureConditional`` -- drop-in for the reference's entropy model
(/root/reference/compressai/entropy_models/entropy_models.py:762-910).

Same constructor arguments, attribute / buffer names (so ``state_dict`` keys match:
``scale_table``, ``scale_bound``, ``lower_bound_scale.bound``,
``likelihood_lower_bound.bound``, ``_offset``, ``_quantized_cdf``, ``_cdf_length``),
method names, argument order and return structure:

* ``forward(inputs, scales, means, weights, training=None) -> (outputs, likelihood)``
  -- one fused CUDA kernel forward, one backward (``csrc/gmm_likelihood.cu``), instead of
  K x ~12 elementwise launches (entropy_models.py:774-808);
* ``compress(y, scales, means, weights) -> ((bytes, abs_max, zero_bitmap), y_quantized)``
  (entropy_models.py:833-867) -- byte-identical stream, nothing but 4 B/symbol leaves the GPU;
* ``decompress(strings, abs_max, zero_bitmap, scales, means, weights) -> y_hat``
  (entropy_models.py:872-910);
* ``reshape_entropy_parameters`` (entropy_models.py:810-828).

Deliberate differences: tensors must be CUDA float32 (there is no CPU fallback);
``decompress`` returns ``y_hat`` on the parameters' device instead of the CPU
(every caller assigns it into a CUDA tensor: checkerboard.py:324, channel_groups.py:154).
"""
from __future__ import annotations

import ctypes as C

from typing import Any, List, Optional, Tuple, Union

import torch
import torch.nn as nn
from torch import Tensor

from . import _lib
from . import ans as _ans


# --------------------------------------------------------------------------------------
class LowerBound(nn.Module):
    """``compressai.ops.LowerBound`` (ops/bound_ops.py:36-80): ``max(x, bound)`` whose
    gradient passes where ``x >= bound`` or the gradient pushes ``x`` up."""

    bound: Tensor

    def __init__(self, bound: float):
        super().__init__()
        self.register_buffer("bound", torch.Tensor([float(bound)]))

    def forward(self, x):
        return _LowerBoundFn.apply(x, self.bound)


class _LowerBoundFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bound):
        ctx.save_for_backward(x, bound)
        return torch.max(x, bound)

    @staticmethod
    def backward(ctx, grad_output):
        x, bound = ctx.saved_tensors
        return ((x >= bound) | (grad_output < 0)) * grad_output, None


class _EntropyCoder:
    """Proxy with the reference's attribute names (entropy_models.py:48-82): the model
    code calls ``self.entropy_coder._encoder.encode_with_indexes_gmm(...)``."""

    def __init__(self, method: str = "ans"):
        if method != "ans":
            raise ValueError(f'Unknown entropy coder "{method}" (only "ans" exists on the B200 path)')
        self.name = method
        self._encoder = _ans.RansEncoder()
        self._decoder = _ans.RansDecoder()

    def encode_with_indexes(self, *args, **kwargs):
        return self._encoder.encode_with_indexes(*args, **kwargs)

    def decode_with_indexes(self, *args, **kwargs):
        return self._decoder.decode_with_indexes(*args, **kwargs)


# --------------------------------------------------------------------------------------
def _require_cuda_f32(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("flashgmm_b200 runs on CUDA tensors only (no CPU fallback); got a CPU tensor")
        if t.dtype != torch.float32:
            raise TypeError(f"flashgmm_b200 expects float32 tensors, got {t.dtype}")


class _GMMLikelihoodFn(torch.autograd.Function):
    """(inputs, noise|None, quant, scales, means, weights, K, scale_bound, lik_bound) -> (outputs, likelihood)."""

    @staticmethod
    def forward(ctx, inputs, noise, quant, scales, means, weights, K, scale_bound, lik_bound):
        _require_cuda_f32(inputs, noise, scales, means, weights)
        B = inputs.size(0) if inputs.dim() > 0 else 1
        plane = inputs.numel() // max(B, 1)
        for p in (scales, means, weights):
            if p.numel() != inputs.numel() * K or p.size(0) != B:
                raise ValueError(f"parameter shape {tuple(p.shape)} does not match inputs {tuple(inputs.shape)} with K={K}")
        x = inputs.contiguous()
        sc, mu, w = scales.contiguous(), means.contiguous(), weights.contiguous()
        nz = noise.contiguous() if noise is not None else None
        y = torch.empty_like(x)
        lik = torch.empty_like(x)
        dev = x.device
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib().flashgmm_gmm_likelihood_fwd(
            _lib.ctx(dev.index), x.data_ptr(), nz.data_ptr() if nz is not None else None, int(quant),
            sc.data_ptr(), mu.data_ptr(), w.data_ptr(), B, plane, int(K), float(scale_bound), float(lik_bound),
            y.data_ptr(), lik.data_ptr(), st))
        ctx.save_for_backward(y, sc, mu, w)
        ctx.geom = (B, plane, int(K), float(scale_bound), float(lik_bound), int(quant))
        ctx.shapes = (inputs.shape, scales.shape, means.shape, weights.shape)
        return y.view(inputs.shape), lik.view(inputs.shape)

    @staticmethod
    def backward(ctx, grad_y, grad_lik):
        y, sc, mu, w = ctx.saved_tensors
        B, plane, K, sb, lb, quant = ctx.geom
        dev = y.device
        gl = grad_lik.contiguous() if grad_lik is not None else torch.zeros_like(y)
        gy = torch.empty_like(y)
        gsc, gmu, gw = torch.empty_like(sc), torch.empty_like(mu), torch.empty_like(w)
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(_lib.lib().flashgmm_gmm_likelihood_bwd(
            _lib.ctx(dev.index), y.data_ptr(), sc.data_ptr(), mu.data_ptr(), w.data_ptr(), gl.data_ptr(), B, plane, K,
            sb, lb, gy.data_ptr(), gsc.data_ptr(), gmu.data_ptr(), gw.data_ptr(), st))
        ish, ssh, msh, wsh = ctx.shapes
        if quant == 2:                       # torch.round: zero gradient (entropy_models.py:176)
            gin = torch.zeros(ish, dtype=y.dtype, device=dev)
        else:
            gin = gy.view(ish)
            if grad_y is not None:
                gin = gin + grad_y
        gnoise = None
        if quant == 1 and ctx.needs_input_grad[1]:
            gnoise = gin
        return gin, gnoise, None, gsc.view(ssh), gmu.view(msh), gw.view(wsh), None, None, None


# --------------------------------------------------------------------------------------
class GaussianMixtureConditional(nn.Module):
    """entropy_models.py:762-910 (and the parts of EntropyModel / GaussianConditional it uses)."""

    def __init__(
        self,
        K: int = 3,
        scale_table: Optional[Union[List, Tuple]] = None,
        *args: Any,
        scale_bound: float = 0.11,
        tail_mass: float = 1e-9,
        likelihood_bound: float = 1e-9,
        entropy_coder: Optional[str] = None,
        entropy_coder_precision: int = 16,
        **kwargs: Any,
    ):
        super().__init__()
        if not isinstance(scale_table, (type(None), list, tuple)):
            raise ValueError(f'Invalid type for scale_table "{type(scale_table)}"')
        if isinstance(scale_table, (list, tuple)) and len(scale_table) < 1:
            raise ValueError(f'Invalid scale_table length "{len(scale_table)}"')
        if scale_table and (scale_table != sorted(scale_table) or any(s <= 0 for s in scale_table)):
            raise ValueError(f'Invalid scale_table "({scale_table})"')
        self.K = K
        self.entropy_coder = _EntropyCoder(entropy_coder or "ans")
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.use_likelihood_bound = likelihood_bound > 0
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        self.register_buffer("_offset", torch.IntTensor())
        self.register_buffer("_quantized_cdf", torch.IntTensor())
        self.register_buffer("_cdf_length", torch.IntTensor())
        self.tail_mass = float(tail_mass)
        if scale_bound is None and scale_table:
            scale_bound = scale_table[0]
        if scale_bound is None or scale_bound <= 0:
            raise ValueError("Invalid parameters")
        self.lower_bound_scale = LowerBound(scale_bound)
        self.register_buffer("scale_table",
                             torch.Tensor(tuple(float(s) for s in scale_table)) if scale_table else torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]))
        self._scale_bound_f = float(torch.tensor(float(scale_bound), dtype=torch.float32))
        self._lik_bound_f = float(torch.tensor(float(likelihood_bound), dtype=torch.float32)) if likelihood_bound > 0 else 0.0

    # the reference pickles the coder by name (entropy_models.py:128-135)
    def __getstate__(self):
        attributes = self.__dict__.copy()
        attributes["entropy_coder"] = self.entropy_coder.name
        return attributes

    def __setstate__(self, state):
        self.__dict__ = state
        self.entropy_coder = _EntropyCoder(self.__dict__.pop("entropy_coder"))

    @property
    def offset(self):
        return self._offset

    @property
    def quantized_cdf(self):
        return self._quantized_cdf

    @property
    def cdf_length(self):
        return self._cdf_length

    def update_scale_table(self, scale_table, force=False):
        """Kept for ``CompressionModel.update`` (models/base.py:139-140): the GMM coder
        never reads the scale-table CDFs, so nothing is built."""
        return False

    def update(self):
        return None

    # -- entropy_models.py:157-182 -------------------------------------------------------
    def quantize(self, inputs: Tensor, mode: str, means: Optional[Tensor] = None) -> Tensor:
        if mode not in ("noise", "dequantize", "symbols"):
            raise ValueError(f'Invalid quantization mode: "{mode}"')
        if mode == "noise":
            noise = torch.empty_like(inputs).uniform_(-0.5, 0.5)
            return inputs + noise
        outputs = inputs.clone()
        if means is not None:
            outputs -= means
        outputs = torch.round(outputs)
        if mode == "dequantize":
            if means is not None:
                outputs += means
            return outputs
        return outputs.int()

    @staticmethod
    def dequantize(inputs: Tensor, means: Optional[Tensor] = None, dtype: torch.dtype = torch.float) -> Tensor:
        if means is not None:
            outputs = inputs.type_as(means)
            outputs += means
        else:
            outputs = inputs.type(dtype)
        return outputs

    # -- entropy_models.py:774-790 -------------------------------------------------------
    def _likelihood(self, inputs: Tensor, scales: Tensor, means: Tensor, weights: Tensor) -> Tensor:
        """Unbounded mixture likelihood of already-quantised ``inputs`` (one fused kernel)."""
        _, lik = _GMMLikelihoodFn.apply(inputs, None, 0, scales, means, weights, self.K, self._scale_bound_f, 0.0)
        return lik

    # -- entropy_models.py:792-808 -------------------------------------------------------
    def forward(self, inputs: Tensor, scales: Tensor, means: Tensor, weights: Tensor,
                training: Optional[bool] = None) -> Tuple[Tensor, Tensor]:
        if training is None:
            training = self.training
        lb = self._lik_bound_f if self.use_likelihood_bound else 0.0
        if training:
            # same RNG call as the reference's quantize("noise") (entropy_models.py:163-167)
            noise = torch.empty_like(inputs).uniform_(-0.5, 0.5)
            return _GMMLikelihoodFn.apply(inputs, noise, 1, scales, means, weights, self.K, self._scale_bound_f, lb)
        return _GMMLikelihoodFn.apply(inputs, None, 2, scales, means, weights, self.K, self._scale_bound_f, lb)

    # -- entropy_models.py:810-828 -------------------------------------------------------
    def reshape_entropy_parameters(self, scales, means, weights, nonzero):
        """(B, K*C, H, W) -> three (Nsym, K) views with strides (1, Nsym): K-major planes,
        symbol order = kept channel, then row-major spatial; scales clamped to [0.11, 256]."""
        reshape_size = (scales.size(0), self.K, scales.size(1) // self.K, -1)

        def planes(t):
            return t.reshape(*reshape_size)[:, :, nonzero].permute(1, 0, 2, 3).reshape(self.K, -1).permute(1, 0)

        return planes(scales).clamp(0.11, 256), planes(means), planes(weights)

    y_q = None
    y_p = None

    # -- compress / decompress glue on the device (csrc/gmm_prologue.cu) ------------------
    def _kept_planes(self, y_q, scales, means, weights, bitmap, n_kept):
        """Kept channels' symbols (int32) and parameters as (Nsym, K) views with strides (1, Nsym):
        what reshape_entropy_parameters + the channel compaction of entropy_models.py:846 produce,
        in one launch."""
        C_, HW = bitmap.numel(), scales.size(2) * scales.size(3)
        n_sym = n_kept * HW
        dev = scales.device
        sym = torch.empty(n_sym, dtype=torch.int32, device=dev) if y_q is not None else None
        planes = [torch.empty((self.K, n_sym), dtype=torch.float32, device=dev) for _ in range(3)]
        sc, mu, w = (t.contiguous() for t in (scales, means, weights))
        _lib.check(_lib.lib().flashgmm_gmm_prologue_gather(
            _lib.ctx(dev.index or 0), y_q.data_ptr() if y_q is not None else None, sc.data_ptr(), mu.data_ptr(),
            w.data_ptr(), bitmap.data_ptr(), C_, HW, self.K, n_sym, sym.data_ptr() if sym is not None else None,
            planes[0].data_ptr(), planes[1].data_ptr(), planes[2].data_ptr(),
            torch.cuda.current_stream(dev).cuda_stream))
        return sym, planes[0].t(), planes[1].t(), planes[2].t()

    # -- entropy_models.py:833-867 -------------------------------------------------------
    def compress(self, y, scales, means, weights):
        _require_cuda_f32(y, scales, means, weights)
        if self.K != 4:
            raise ValueError("the GMM rANS coder binds K=4 only (rans_interface.cpp:60, 982)")
        if y.size(0) != 1:
            raise ValueError("compress takes one image at a time (the reference squeezes the batch axis, entropy_models.py:842)")
        dev = y.device
        C_, HW = y.size(1), y.size(2) * y.size(3)
        y = y.contiguous()
        y_quantized = torch.empty_like(y)
        zero_bitmap = torch.empty(C_, dtype=torch.int64, device=dev)
        bitmap_h = (C.c_int32 * max(1, C_))()
        y_max, y_min = C.c_float(), C.c_float()
        # one kernel + one synchronisation: round, zero bitmap, extrema of the UNROUNDED y
        _lib.check(_lib.lib().flashgmm_gmm_prologue_scan(
            _lib.ctx(dev.index or 0), y.data_ptr(), C_, HW, y_quantized.data_ptr(), zero_bitmap.data_ptr(), bitmap_h,
            C.byref(y_max), C.byref(y_min), torch.cuda.current_stream(dev).cuda_stream))
        # abs_max truncated like .abs().int() (entropy_models.py:834-837)
        abs_max = max(int(abs(y_max.value)), int(abs(y_min.value))) + 1
        abs_max = 1 if abs_max < 1 else abs_max
        n_kept = sum(1 for i in range(C_) if bitmap_h[i])
        symbols, scales, means, weights = self._kept_planes(y_quantized, scales, means, weights, zero_bitmap, n_kept)
        if self.y_q is None:
            self.y_q = symbols
        else:
            self.y_p = symbols
        rv = self.entropy_coder._encoder.encode_with_indexes_gmm(symbols, scales, means, weights, abs_max + 1)
        return (rv, abs_max, zero_bitmap), y_quantized

    # -- entropy_models.py:872-910 -------------------------------------------------------
    def decompress(self, strings, abs_max, zero_bitmap, scales, means, weights):
        _require_cuda_f32(scales, means, weights)
        if self.K != 4:
            raise ValueError("the GMM rANS coder binds K=4 only (rans_interface.cpp:60, 982)")
        if scales.size(0) != 1:
            raise ValueError("decompress takes one image at a time (entropy_models.py:842, 905)")
        dev = scales.device
        bitmap = zero_bitmap.to(device=dev, dtype=torch.int64).contiguous()
        C_, HW = bitmap.numel(), scales.size(2) * scales.size(3)
        n_kept = int(torch.count_nonzero(bitmap).item())               # the one sync of this call
        _sym, sc, mu, w = self._kept_planes(None, scales, means, weights, bitmap, n_kept)
        symbols = self.entropy_coder._decoder.decode_with_indexes_gmm(strings, sc, mu, w, abs_max + 1)
        self.y_q = None
        symbols = symbols.to(device=dev, dtype=torch.int32).contiguous()
        y_hat = torch.empty(1, C_, scales.size(2), scales.size(3), device=dev)
        _lib.check(_lib.lib().flashgmm_gmm_epilogue_scatter(
            _lib.ctx(dev.index or 0), symbols.data_ptr() if symbols.numel() else None, bitmap.data_ptr(), C_, HW,
            y_hat.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        return y_hat
