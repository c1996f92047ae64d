"""Entropy models of the GMM hot path, B200-native.

``GaussianMixtureConditional`` -- drop-in for the reference's entropy model
(/root/reference/compressai/entropy_models/entropy_models.py:762-910), and
``EntropyBottleneck`` -- the factorised prior of the hyper-latent z branch (:330-618), whose
compress / decompress run the table coder with z resident on the GPU.

Same constructor arguments, parameter / buffer names (``state_dict`` keys and shapes match:
``scale_table``, ``scale_bound``, ``lower_bound_scale.bound``, ``likelihood_lower_bound.bound``,
``_offset``, ``_quantized_cdf``, ``_cdf_length``; ``_matrix{i}``, ``_bias{i}``, ``_factor{i}``,
``quantiles``, ``target``), method names, argument order and return structure:

* ``forward(inputs, scales, means, weights, training=None) -> (outputs, likelihood)``
  -- one fused CUDA kernel forward, one backward (``csrc/gmm_likelihood.cu``) instead of
  K x ~12 elementwise launches (:774-808);
* ``compress(y, scales, means, weights) -> ((bytes, abs_max, zero_bitmap), y_quantized)``
  (:833-867) -- byte-identical stream, nothing but 4 B/symbol leaves the GPU;
* ``decompress(strings, abs_max, zero_bitmap, scales, means, weights) -> y_hat`` (:872-910);
* ``reshape_entropy_parameters`` (:810-828);
* ``compress_batch / decompress_batch``: the same for B images of one coding pass, every stream of
  the pass in one call of the batched device coder (new; the reference codes one image at a time).

Checkpoints: ``load_state_dict`` accepts state dicts saved before or after ``update()`` (the
table buffers are resized to the checkpoint's, which the reference does in
``CompressionModel.load_state_dict``, models/base.py:94-115).

Deliberate differences: GMM coding needs CUDA float32 tensors (there is no CPU fallback);
``decompress`` returns ``y_hat`` on the parameters' device instead of the CPU (every caller
assigns it into a CUDA tensor: checkerboard.py:324, channel_groups.py:154).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Any, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
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


def pmf_to_quantized_cdf(pmf: Sequence[float], precision: int = 16) -> List[int]:
    """``compressai._CXX.pmf_to_quantized_cdf`` (cpp_exts/ops/ops.cpp:40-108) on the host, in integer
    arithmetic: float32 ``round(p * 2^precision)`` (half away from zero), rescale to the total, prefix
    sum, last entry forced to ``2^precision``, then every empty slot takes one count from the
    smallest slot that can spare it (first such slot on ties). Runs once per model, at ``update()``."""
    one = 1 << precision
    p32 = np.asarray(pmf, dtype=np.float32)
    if p32.ndim != 1 or not np.all(np.isfinite(p32)) or np.any(p32 < 0):
        raise ValueError("Invalid `pmf`, non-finite or negative element found")
    counts = np.floor(p32.astype(np.float64) * one + 0.5).astype(np.int64)   # exact: p * 2^16 < 2^40
    total = int(counts.sum())
    if total == 0:
        raise ValueError("Invalid `pmf`: at least one element must have a non-zero probability.")
    cdf = [0] * (len(counts) + 1)
    acc = 0                                                  # slot 0 of the reference's vector is the leading 0
    for i, c in enumerate(counts.tolist()):
        acc += (one * c) // total
        cdf[i + 1] = acc
    cdf[-1] = one
    n = len(cdf)
    for i in range(n - 1):
        if cdf[i] != cdf[i + 1]:
            continue
        best, donor = None, -1
        for j in range(n - 1):
            f = cdf[j + 1] - cdf[j]
            if f > 1 and (best is None or f < best):
                best, donor = f, j
        if donor < 0:
            raise ValueError("pmf_to_quantized_cdf: no slot can spare a count")
        if donor < i:
            for j in range(donor + 1, i + 1):
                cdf[j] -= 1
        else:
            for j in range(i + 1, donor + 1):
                cdf[j] += 1
    return cdf


# --------------------------------------------------------------------------------------
def _require_cuda_f32(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("flashgmm_b200 runs on CUDA tensors only (no CPU fallback); got a CPU tensor")
        if t.dtype != torch.float32:
            raise TypeError(f"flashgmm_b200 expects float32 tensors, got {t.dtype}")


def _stream_of(dev) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


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
        _lib.check(_lib.lib().flashgmm_gmm_likelihood_fwd(
            _lib.ctx(dev.index), x.data_ptr(), nz.data_ptr() if nz is not None else None, int(quant),
            sc.data_ptr(), mu.data_ptr(), w.data_ptr(), B, plane, int(K), float(scale_bound), float(lik_bound),
            y.data_ptr(), lik.data_ptr(), _stream_of(dev)))
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
        _lib.check(_lib.lib().flashgmm_gmm_likelihood_bwd(
            _lib.ctx(dev.index), y.data_ptr(), sc.data_ptr(), mu.data_ptr(), w.data_ptr(), gl.data_ptr(), B, plane, K,
            sb, lb, gy.data_ptr(), gsc.data_ptr(), gmu.data_ptr(), gw.data_ptr(), _stream_of(dev)))
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
_TABLE_BUFFERS = ("_offset", "_quantized_cdf", "_cdf_length")


class _TableModel(nn.Module):
    """What ``EntropyModel`` (entropy_models.py:101-327) gives its subclasses: the likelihood bound,
    the coder proxy, the three table buffers filled by ``update()``, quantisation helpers.

    The table buffers change shape at ``update()``; ``_load_from_state_dict`` therefore takes the
    checkpoint's shapes for ``_resizable`` buffers (models/base.py:94-115 does this from the outside for
    instances of the reference's own classes; a replacement class has to do it itself)."""

    _resizable: Tuple[str, ...] = _TABLE_BUFFERS

    def __init__(self, likelihood_bound: float = 1e-9, entropy_coder: Optional[str] = None,
                 entropy_coder_precision: int = 16):
        super().__init__()
        self.entropy_coder = _EntropyCoder(entropy_coder or "ans")
        self.entropy_coder_precision = int(entropy_coder_precision)
        self.use_likelihood_bound = likelihood_bound > 0
        if self.use_likelihood_bound:
            self.likelihood_lower_bound = LowerBound(likelihood_bound)
        for name in _TABLE_BUFFERS:
            self.register_buffer(name, torch.IntTensor())

    # the reference pickles the coder by name (entropy_models.py:128-135)
    def __getstate__(self):
        state = dict(self.__dict__)
        state["entropy_coder"] = self.entropy_coder.name
        return state

    def __setstate__(self, state):
        self.__dict__ = state
        self.entropy_coder = _EntropyCoder(self.__dict__.pop("entropy_coder"))

    offset = property(lambda self: self._offset)
    quantized_cdf = property(lambda self: self._quantized_cdf)
    cdf_length = property(lambda self: self._cdf_length)

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        for name in self._resizable:
            src = state_dict.get(prefix + name)
            cur = self._buffers.get(name)
            if src is not None and cur is not None and src.shape != cur.shape:
                self._buffers[name] = torch.empty(src.shape, dtype=cur.dtype, device=cur.device)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
        self._tables_changed()

    def _tables_changed(self):
        """Cached host / device copies of the tables are stale."""

    def quantize(self, inputs: Tensor, mode: str, means: Optional[Tensor] = None) -> Tensor:
        """"noise": + U(-1/2, 1/2); "dequantize": round(x - means) + means; "symbols": int32 round(x - means)
        (entropy_models.py:157-182)."""
        if mode == "noise":
            return inputs + torch.empty_like(inputs).uniform_(-0.5, 0.5)
        if mode not in ("dequantize", "symbols"):
            raise ValueError(f'Invalid quantization mode: "{mode}"')
        centred = inputs if means is None else inputs - means
        rounded = torch.round(centred)
        if mode == "symbols":
            return rounded.int()
        return rounded if means is None else rounded + means

    @staticmethod
    def dequantize(inputs: Tensor, means: Optional[Tensor] = None, dtype: torch.dtype = torch.float) -> Tensor:
        """entropy_models.py:185-194."""
        if means is None:
            return inputs.type(dtype)
        return inputs.type_as(means) + means

    def _pmf_to_cdf(self, pmf: Tensor, tail_mass: Tensor, pmf_length: Tensor, max_length: int) -> Tensor:
        """Rows of quantised CDFs, zero padded to ``max_length + 2`` (entropy_models.py:203-211)."""
        rows = torch.zeros((len(pmf_length), max_length + 2), dtype=torch.int32)
        pmf_h, tail_h, len_h = pmf.detach().float().cpu(), tail_mass.detach().float().cpu(), pmf_length.cpu().tolist()
        for i, n in enumerate(len_h):
            q = pmf_to_quantized_cdf(torch.cat((pmf_h[i, :n], tail_h[i].reshape(-1)[:1])).tolist(),
                                     self.entropy_coder_precision)
            rows[i, : len(q)] = torch.tensor(q, dtype=torch.int32)
        return rows.to(pmf.device)

    def _check_tables(self):
        if self._quantized_cdf.numel() == 0 or self._offset.numel() == 0 or self._cdf_length.numel() == 0:
            raise ValueError("Uninitialized CDFs. Run update() first")
        if self._quantized_cdf.dim() != 2 or self._offset.dim() != 1 or self._cdf_length.dim() != 1:
            raise ValueError("Invalid CDF / offset / length tensor shapes")


# --------------------------------------------------------------------------------------
class EntropyBottleneck(_TableModel):
    """Factorised prior of the hyper-latent z (entropy_models.py:330-618; Balle et al. 2018).

    The density model (cumulative through a chain of positive matrices, biases and tanh gates) and
    ``update()`` are tensor math that runs once per training step / once per model and stay on PyTorch.
    ``compress`` / ``decompress`` are the z branch of every image: the reference marshals five Python
    lists per image (:237-270); here z stays on the GPU -- a kernel looks every symbol up in its
    channel's CDF row (``csrc/gmm_table.cu``), 8-byte records come down in one copy and the ctx's worker
    threads run the serial rANS, one image each. Streams are byte-identical to the reference's."""

    def __init__(self, channels: int, *args: Any, tail_mass: float = 1e-9, init_scale: float = 10,
                 filters: Tuple[int, ...] = (3, 3, 3, 3), **kwargs: Any):
        super().__init__(*args, **kwargs)
        self.channels = int(channels)
        self.filters = tuple(int(f) for f in filters)
        self.init_scale = float(init_scale)
        self.tail_mass = float(tail_mass)
        widths = (1,) + self.filters + (1,)
        depth = len(self.filters) + 1
        scale = self.init_scale ** (1 / depth)
        for i in range(depth):
            fan_out, fan_in = widths[i + 1], widths[i]
            m = torch.full((self.channels, fan_out, fan_in), float(np.log(np.expm1(1 / scale / fan_out))))
            self.register_parameter(f"_matrix{i:d}", nn.Parameter(m))
            b = torch.empty(self.channels, fan_out, 1)
            nn.init.uniform_(b, -0.5, 0.5)
            self.register_parameter(f"_bias{i:d}", nn.Parameter(b))
            if i < depth - 1:
                self.register_parameter(f"_factor{i:d}", nn.Parameter(torch.zeros(self.channels, fan_out, 1)))
        q = torch.tensor([-self.init_scale, 0.0, self.init_scale])
        self.quantiles = nn.Parameter(q.repeat(self.channels, 1, 1))
        t = float(np.log(2 / self.tail_mass - 1))
        self.register_buffer("target", torch.Tensor([-t, 0, t]))
        self._host_tables = None

    def _tables_changed(self):
        self._host_tables = None

    def _get_medians(self) -> Tensor:
        return self.quantiles[:, :, 1:2]

    # -- density model ---------------------------------------------------------------------------
    def _density_parameters(self) -> List[Tensor]:
        """[matrix0, bias0, factor0, matrix1, ..., matrix4, bias4] in layer order."""
        depth = len(self.filters) + 1
        out = []
        for i in range(depth):
            out += [getattr(self, f"_matrix{i:d}"), getattr(self, f"_bias{i:d}")]
            if i < depth - 1:
                out.append(getattr(self, f"_factor{i:d}"))
        return out

    def _logits_cumulative(self, inputs: Tensor, stop_gradient: bool, params: Optional[Sequence[Tensor]] = None) -> Tensor:
        """(C, 1, N) -> (C, 1, N): logits of the cumulative (entropy_models.py:421-440)."""
        logits = inputs
        depth = len(self.filters) + 1
        it = iter(self._density_parameters() if params is None else params)
        for i in range(depth):
            m, b = next(it), next(it)
            if stop_gradient:
                m, b = m.detach(), b.detach()
            logits = torch.matmul(F.softplus(m), logits)
            logits = logits + b
            if i < depth - 1:
                f = next(it)
                if stop_gradient:
                    f = f.detach()
                logits = logits + torch.tanh(f) * torch.tanh(logits)
        return logits

    def _likelihood_fused(self, values: Tensor) -> Tensor:
        """Unbounded likelihood of (C, 1, N) values through csrc/eb_likelihood.cu (filters (3, 3, 3, 3), CUDA float32);
        the backward pass recomputes the tensor-op formula under autograd."""
        return _EBLikelihoodFn.apply(self, values, *self._density_parameters())

    def _likelihood(self, inputs: Tensor, stop_gradient: bool = False):
        lower = self._logits_cumulative(inputs - 0.5, stop_gradient=stop_gradient)
        upper = self._logits_cumulative(inputs + 0.5, stop_gradient=stop_gradient)
        return torch.sigmoid(upper) - torch.sigmoid(lower), lower, upper

    def loss(self) -> Tensor:
        return torch.abs(self._logits_cumulative(self.quantiles, stop_gradient=True) - self.target).sum()

    def forward(self, x: Tensor, training: Optional[bool] = None) -> Tuple[Tensor, Tensor]:
        """(B, C, ...) -> (z_hat, likelihood): channels first, one density per channel (entropy_models.py:463-510)."""
        if training is None:
            training = self.training
        xc = x.transpose(0, 1).contiguous()
        shape = xc.size()
        values = xc.reshape(shape[0], 1, -1)
        outputs = self.quantize(values, "noise" if training else "dequantize", self._get_medians())
        if outputs.is_cuda and outputs.dtype == torch.float32 and self.filters == (3, 3, 3, 3):
            likelihood = self._likelihood_fused(outputs)     # one kernel instead of ~30 small tensor ops
        else:
            likelihood, _, _ = self._likelihood(outputs)
        if self.use_likelihood_bound:
            likelihood = self.likelihood_lower_bound(likelihood)
        outputs = outputs.reshape(shape).transpose(0, 1).contiguous()
        likelihood = likelihood.reshape(shape).transpose(0, 1).contiguous()
        return outputs, likelihood

    # -- tables ----------------------------------------------------------------------------------
    def update(self, force: bool = False, update_quantiles: bool = False) -> bool:
        """Build the per-channel quantised CDFs from the learned density (entropy_models.py:389-419)."""
        if self._offset.numel() > 0 and not force:
            return False
        if update_quantiles:
            self._update_quantiles()
        q = self.quantiles.detach()
        medians = q[:, 0, 1]
        below = torch.clamp(torch.ceil(medians - q[:, 0, 0]).int(), min=0)
        above = torch.clamp(torch.ceil(q[:, 0, 2] - medians).int(), min=0)
        self._offset = -below
        start = medians - below
        length = above + below + 1
        longest = int(length.max().item())
        samples = torch.arange(longest, device=start.device)[None, :] + start[:, None, None]
        pmf, lower, upper = self._likelihood(samples, stop_gradient=True)
        tail = torch.sigmoid(lower[:, 0, :1]) + torch.sigmoid(-upper[:, 0, -1:])
        self._quantized_cdf = self._pmf_to_cdf(pmf[:, 0, :], tail, length, longest)
        self._cdf_length = length + 2
        self._tables_changed()
        return True

    @torch.no_grad()
    def _update_quantiles(self, search_radius=1e5, rtol=1e-4, atol=1e-3):
        """Bisection for the three quantiles of every channel (entropy_models.py:573-604)."""
        dev = self.quantiles.device
        for i in range(len(self.target)):
            lo = torch.full((self.channels, 1, 1), -search_radius, device=dev)
            hi = torch.full((self.channels, 1, 1), search_radius, device=dev)
            tgt = self.target[i]
            f = lambda v: self._logits_cumulative(v, stop_gradient=True)  # noqa: E731
            lo = torch.where(tgt <= f(hi), lo, hi)
            hi = torch.where(f(lo) <= tgt, hi, lo)
            while not torch.isclose(lo, hi, rtol=rtol, atol=atol).all():
                mid = (lo + hi) / 2
                fm = f(mid)
                lo = torch.where(fm <= tgt, mid, lo)
                hi = torch.where(fm >= tgt, mid, hi)
            self.quantiles[:, :, i] = ((lo + hi) / 2)[:, :, 0]

    @staticmethod
    def _build_indexes(size):
        """CDF row of every element: its channel (entropy_models.py:556-567)."""
        n, c = size[0], size[1]
        idx = torch.arange(c, dtype=torch.int32).view(1, c, *([1] * (len(size) - 2)))
        return idx.expand(n, c, *size[2:]).contiguous()

    def _tables_on(self, dev):
        """(cdfs, sizes, offsets) int32 on ``dev`` + contiguous host copies, cached until the tables change."""
        if self._host_tables is None or self._host_tables[0] != dev:
            self._check_tables()
            d = tuple(t.to(device=dev, dtype=torch.int32).contiguous() for t in
                      (self._quantized_cdf, self._cdf_length, self._offset))
            h = tuple(np.ascontiguousarray(t.cpu().numpy(), np.int32) for t in d)
            self._host_tables = (dev, d, h)
        return self._host_tables[1], self._host_tables[2]

    # -- coding ----------------------------------------------------------------------------------
    def compress(self, x: Tensor) -> List[bytes]:
        """(B, C, ...) float32 CUDA -> B byte strings (entropy_models.py:605-611 -> :237-270)."""
        _require_cuda_f32(x)
        if x.dim() < 2 or x.size(1) != self.channels:
            raise ValueError(f"expected (B, {self.channels}, ...), got {tuple(x.shape)}")
        dev = x.device
        (cdfs, sizes, offs), _ = self._tables_on(dev)
        B, Cc = x.size(0), x.size(1)
        HW = x[0, 0].numel()
        z = x.contiguous()
        med = self._get_medians().detach().reshape(-1).to(device=dev, dtype=torch.float32).contiguous()
        arr = (_lib.Stream * max(1, B))()
        bufs = []
        for b in range(B):
            buf = np.empty(max(64, Cc * HW * 5 + 64), np.uint8)   # <= 9 nibbles + marker per symbol
            bufs.append(buf)
            arr[b].bytes, arr[b].bytes_cap = buf.ctypes.data, buf.size
        _lib.check(_lib.lib().flashgmm_eb_compress_batch(
            _lib.ctx(dev.index or 0), z.data_ptr(), med.data_ptr(), cdfs.data_ptr(), sizes.data_ptr(), offs.data_ptr(),
            B, Cc, HW, cdfs.size(1), None, arr, _stream_of(dev)))
        return [bufs[b][: arr[b].bytes_len].tobytes() for b in range(B)]

    def decompress(self, strings: Sequence[bytes], size) -> Tensor:
        """B byte strings -> z_hat (B, C, *size) float32 on the tables' device (entropy_models.py:613-618 -> :272-327)."""
        if not isinstance(strings, (tuple, list)):
            raise ValueError("Invalid `strings` parameter type.")
        dev = self._quantized_cdf.device
        if dev.type != "cuda":
            raise RuntimeError("flashgmm_b200 decodes on a CUDA device only (move the model to cuda)")
        (cdfs, _, _), (h_cdfs, h_sizes, h_offs) = self._tables_on(dev)
        B, Cc = len(strings), self._quantized_cdf.size(0)
        size = tuple(int(s) for s in size)
        HW = int(np.prod(size)) if len(size) else 1
        sym = torch.empty((B, Cc) + size, dtype=torch.int32, device=dev)
        arr = (_lib.Stream * max(1, B))()
        keep = []
        for b, s in enumerate(strings):
            buf = np.frombuffer(s, np.uint8)
            keep.append(buf)
            arr[b].bytes, arr[b].bytes_cap = buf.ctypes.data, buf.size
        _lib.check(_lib.lib().flashgmm_eb_decompress_batch(
            _lib.ctx(dev.index or 0), arr, B, h_cdfs.ctypes.data, h_sizes.ctypes.data, h_offs.ctypes.data, Cc, HW,
            cdfs.size(1), sym.data_ptr(), _stream_of(dev)))
        med = self._get_medians().detach().reshape(1, Cc, *([1] * len(size))).to(dev)
        return self.dequantize(sym, med.expand(B, Cc, *size), med.dtype)


class _EBLikelihoodFn(torch.autograd.Function):
    """(module, values (C, 1, N), *density parameters) -> sigmoid(logits(v + 1/2)) - sigmoid(logits(v - 1/2))."""

    @staticmethod
    def forward(ctx, eb, values, *params):
        C_ = values.size(0)
        n = values.numel() // max(C_, 1)
        packed = []
        it = iter(params)
        depth = len(eb.filters) + 1
        for i in range(depth):
            m, b = next(it), next(it)
            packed += [F.softplus(m.detach()).reshape(C_, -1), b.detach().reshape(C_, -1)]
            if i < depth - 1:
                packed.append(torch.tanh(next(it).detach()).reshape(C_, -1))
        table = torch.cat(packed, dim=1).to(torch.float32).contiguous()
        assert table.size(1) == 58
        v = values.detach().contiguous()
        out = torch.empty_like(v)
        dev = v.device
        _lib.check(_lib.lib().flashgmm_eb_likelihood(_lib.ctx(dev.index or 0), v.data_ptr(), table.data_ptr(), C_, n,
                                                     out.data_ptr(), _stream_of(dev)))
        ctx.eb = eb
        ctx.save_for_backward(values, *params)
        return out

    @staticmethod
    def backward(ctx, grad):
        values, *params = ctx.saved_tensors
        with torch.enable_grad():
            v = values.detach().requires_grad_(True)
            ps = [p.detach().requires_grad_(True) for p in params]
            lower = ctx.eb._logits_cumulative(v - 0.5, False, ps)
            upper = ctx.eb._logits_cumulative(v + 0.5, False, ps)
            lik = torch.sigmoid(upper) - torch.sigmoid(lower)
            grads = torch.autograd.grad(lik, [v] + ps, grad, allow_unused=True)
        return (None, *grads)


# --------------------------------------------------------------------------------------
def _normal_cdf(t: Tensor) -> Tensor:
    """Phi(t) through erfc, the numerically safe form (entropy_models.py:674-678)."""
    return 0.5 * torch.erfc(t * -(2 ** -0.5))


class GaussianMixtureConditional(_TableModel):
    """entropy_models.py:762-910 (and the parts of GaussianConditional, :621-759, it inherits)."""

    _resizable = _TABLE_BUFFERS + ("scale_table",)

    def __init__(self, K: int = 3, scale_table: Optional[Union[List, Tuple]] = None, *args: Any,
                 scale_bound: float = 0.11, tail_mass: float = 1e-9, **kwargs: Any):
        super().__init__(*args, **kwargs)
        if scale_table is not None:
            if not isinstance(scale_table, (list, tuple)):
                raise ValueError(f'Invalid type for scale_table "{type(scale_table)}"')
            if len(scale_table) < 1:
                raise ValueError(f'Invalid scale_table length "{len(scale_table)}"')
            if any(s <= 0 for s in scale_table) or list(scale_table) != sorted(scale_table):
                raise ValueError(f'Invalid scale_table "({scale_table})"')
        self.K = K
        self.tail_mass = float(tail_mass)
        if scale_bound is None:
            scale_bound = scale_table[0] if scale_table else None
        if scale_bound is None or scale_bound <= 0:
            raise ValueError("Invalid parameters")
        self.lower_bound_scale = LowerBound(scale_bound)
        self.register_buffer("scale_table", self._as_table(scale_table) if scale_table else torch.Tensor())
        self.register_buffer("scale_bound", torch.Tensor([float(scale_bound)]))

    @staticmethod
    def _as_table(scale_table) -> Tensor:
        return torch.Tensor(tuple(float(s) for s in scale_table))

    # -- GaussianConditional's table side (unused by the GMM coder, kept so that update() leaves the
    #    same buffers behind as the reference: models/base.py:139-140, entropy_models.py:684-718) ----
    def update_scale_table(self, scale_table, force: bool = False) -> bool:
        if self._offset.numel() > 0 and not force:
            return False
        self.scale_table = self._as_table(scale_table).to(self.scale_table.device)
        self.update()
        return True

    def update(self) -> None:
        multiplier = -_norm_ppf(self.tail_mass / 2)
        centre = torch.ceil(self.scale_table * multiplier).int()
        length = 2 * centre + 1
        longest = int(torch.max(length).item())
        dist = torch.abs(torch.arange(longest, device=centre.device).int() - centre[:, None]).float()
        sigma = self.scale_table.unsqueeze(1).float()
        upper = _normal_cdf((0.5 - dist) / sigma)
        lower = _normal_cdf((-0.5 - dist) / sigma)
        self._quantized_cdf = self._pmf_to_cdf(upper - lower, 2 * lower[:, :1], length, longest)
        self._offset = -centre
        self._cdf_length = length + 2

    # -- bounds are read from the (loadable) buffers at call time -----------------------------------
    def _bounds(self) -> Tuple[float, float]:
        sb = float(self.lower_bound_scale.bound)
        lb = float(self.likelihood_lower_bound.bound) if self.use_likelihood_bound else 0.0
        return sb, lb

    # -- entropy_models.py:774-790 -------------------------------------------------------
    def _likelihood(self, inputs: Tensor, scales: Tensor, means: Tensor, weights: Tensor) -> Tensor:
        """Unbounded mixture likelihood of already-quantised ``inputs`` (one fused kernel)."""
        sb, _ = self._bounds()
        _, lik = _GMMLikelihoodFn.apply(inputs, None, 0, scales, means, weights, self.K, sb, 0.0)
        return lik

    # -- entropy_models.py:792-808 -------------------------------------------------------
    def forward(self, inputs: Tensor, scales: Tensor, means: Tensor, weights: Tensor,
                training: Optional[bool] = None) -> Tuple[Tensor, Tensor]:
        if training is None:
            training = self.training
        sb, lb = self._bounds()
        if training:
            # same RNG call as the reference's quantize("noise") (entropy_models.py:163-167)
            noise = torch.empty_like(inputs).uniform_(-0.5, 0.5)
            return _GMMLikelihoodFn.apply(inputs, noise, 1, scales, means, weights, self.K, sb, lb)
        return _GMMLikelihoodFn.apply(inputs, None, 2, scales, means, weights, self.K, sb, lb)

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
    # test hook: a list here receives ("enc" | "dec", per-image (symbols, scales, means, weights) views, abs_max list)
    # for every coding call, so that a checker can replay the same tensors through the reference
    trace = None

    # -- compress / decompress glue on the device (csrc/gmm_prologue.cu), B images per launch -------
    def _check_coding_args(self, *tensors):
        _require_cuda_f32(*tensors)
        if self.K != 4:
            raise ValueError("the GMM rANS coder binds K=4 only (rans_interface.cpp:60, 982)")

    def _scan(self, y: Tensor):
        """round, zero bitmap, abs_max per image: one kernel, one synchronisation.
        -> (y_quantized, bitmap (B, C) int64 device, bitmap host (B, C) int32, abs_max list)."""
        B, C_, HW = y.size(0), y.size(1), y.size(2) * y.size(3)
        dev = y.device
        y_q = torch.empty_like(y)
        bitmap = torch.empty((B, C_), dtype=torch.int64, device=dev)
        bitmap_h = np.empty((B, max(1, C_)), np.int32)
        ext = np.zeros((B, 2), np.float32)
        _lib.check(_lib.lib().flashgmm_gmm_prologue_scan_batch(
            _lib.ctx(dev.index or 0), y.data_ptr(), B, C_, HW, y_q.data_ptr(), bitmap.data_ptr(),
            bitmap_h.ctypes.data, ext.ctypes.data, _stream_of(dev)))
        # abs_max from the UNROUNDED y, truncated like .abs().int() (entropy_models.py:834-837)
        abs_max = [max(1, max(int(abs(float(ext[b, 0]))), int(abs(float(ext[b, 1])))) + 1) for b in range(B)]
        return y_q, bitmap, bitmap_h[:, :C_], abs_max

    def _gather(self, y_q, scales, means, weights, bitmap, n_kept: Sequence[int]):
        """Kept channels' symbols (int32) and parameters as per-image (Nsym, K) views with strides (1, Nsym):
        what reshape_entropy_parameters + the channel compaction of entropy_models.py:846 produce, for every
        image of the batch in one launch. -> list of (symbols | None, scales, means, weights) per image."""
        B, C_ = bitmap.shape
        HW = scales.size(2) * scales.size(3)
        dev = scales.device
        n_sym = np.asarray([k * HW for k in n_kept], np.int64)
        total = int(n_sym.sum())
        sym = torch.empty(max(total, 1), dtype=torch.int32, device=dev) if y_q is not None else None
        planes = torch.empty((3, self.K * max(total, 1)), dtype=torch.float32, device=dev)
        sc, mu, w = (t.contiguous() for t in (scales, means, weights))
        _lib.check(_lib.lib().flashgmm_gmm_prologue_gather_batch(
            _lib.ctx(dev.index or 0), y_q.data_ptr() if y_q is not None else None, sc.data_ptr(), mu.data_ptr(),
            w.data_ptr(), bitmap.data_ptr(), B, C_, HW, self.K, n_sym.ctypes.data,
            sym.data_ptr() if sym is not None else None, planes[0].data_ptr(), planes[1].data_ptr(),
            planes[2].data_ptr(), _stream_of(dev)))
        out, off = [], 0
        for b in range(B):
            n = int(n_sym[b])
            views = [planes[t, self.K * off: self.K * (off + n)].view(self.K, n).t() for t in range(3)]
            out.append((sym[off: off + n] if sym is not None else None, *views))
            off += n
        return out

    def _scatter(self, symbols: Optional[Tensor], bitmap: Tensor, n_kept: Sequence[int], shape) -> Tensor:
        B, C_, H, W = shape
        dev = bitmap.device
        n_sym = np.asarray([k * H * W for k in n_kept], np.int64)
        y_hat = torch.empty(shape, dtype=torch.float32, device=dev)
        _lib.check(_lib.lib().flashgmm_gmm_epilogue_scatter_batch(
            _lib.ctx(dev.index or 0), symbols.data_ptr() if symbols is not None and symbols.numel() else None,
            bitmap.data_ptr(), n_sym.ctypes.data, B, C_, H * W, y_hat.data_ptr(), _stream_of(dev)))
        return y_hat

    # -- entropy_models.py:833-867 -------------------------------------------------------
    def compress(self, y, scales, means, weights):
        self._check_coding_args(y, scales, means, weights)
        if y.size(0) != 1:
            raise ValueError("compress takes one image at a time (the reference squeezes the batch axis, "
                             "entropy_models.py:842); use compress_batch for B images")
        y = y.contiguous()
        y_quantized, bitmap, bitmap_h, abs_max = self._scan(y)
        ((symbols, sc, mu, w),) = self._gather(y_quantized, scales, means, weights, bitmap, [int(bitmap_h[0].sum())])
        if self.y_q is None:
            self.y_q = symbols
        else:
            self.y_p = symbols
        if self.trace is not None:
            self.trace.append(("enc", [(symbols, sc, mu, w)], list(abs_max)))
        rv = self.entropy_coder._encoder.encode_with_indexes_gmm(symbols, sc, mu, w, abs_max[0] + 1)
        return (rv, abs_max[0], bitmap[0]), y_quantized

    def compress_batch(self, y, scales, means, weights, session: Optional["EncodeSession"] = None):
        """B images of one coding pass: ``y`` (B, C, H, W), parameters (B, K*C, H, W).
        -> ([(bytes, abs_max, zero_bitmap)] * B, y_quantized (B, C, H, W)); image b's tuple is what
        ``compress`` returns for image b alone (byte-identical stream). With a ``session`` the streams are
        only registered -- the tuples hold ``None`` for the bytes until ``session.flush()`` codes every
        registered stream of every pass in ONE batched call (the encoder knows all of y up front)."""
        self._check_coding_args(y, scales, means, weights)
        y = y.contiguous()
        y_quantized, bitmap, bitmap_h, abs_max = self._scan(y)
        per_image = self._gather(y_quantized, scales, means, weights, bitmap, bitmap_h.sum(1).tolist())
        if self.trace is not None:
            self.trace.append(("enc", per_image, list(abs_max)))
        own = session is None
        if own:
            session = EncodeSession()
        results = [session.add(st, abs_max[b], bitmap[b]) for b, st in enumerate(per_image)]
        if own:
            session.flush()
            results = [tuple(r) for r in results]
        return results, y_quantized

    # -- entropy_models.py:872-910 -------------------------------------------------------
    def decompress(self, strings, abs_max, zero_bitmap, scales, means, weights):
        self._check_coding_args(scales, means, weights)
        if scales.size(0) != 1:
            raise ValueError("decompress takes one image at a time (entropy_models.py:842, 905); "
                             "use decompress_batch for B images")
        dev = scales.device
        bitmap = torch.as_tensor(zero_bitmap).to(device=dev, dtype=torch.int64).reshape(1, -1).contiguous()
        n_kept = [int(torch.count_nonzero(bitmap).item())]           # the one sync of this call
        ((_s, sc, mu, w),) = self._gather(None, scales, means, weights, bitmap, n_kept)
        symbols = self.entropy_coder._decoder.decode_with_indexes_gmm(strings, sc, mu, w, abs_max + 1)
        self.y_q = None
        symbols = symbols.to(device=dev, dtype=torch.int32).contiguous()
        if self.trace is not None:
            self.trace.append(("dec", [(symbols, sc, mu, w)], [int(abs_max)]))
        return self._scatter(symbols, bitmap, n_kept, (1, bitmap.size(1), scales.size(2), scales.size(3)))

    def decompress_batch(self, items: Sequence[Tuple], scales, means, weights) -> Tensor:
        """``items[b] = (bytes, abs_max, zero_bitmap)`` for the B images of one coding pass, parameters
        (B, K*C, H, W) -> y_hat (B, C, H, W) on the parameters' device. All B streams decode in one call
        of the batched device decoder (rows kernel -> one pop warp per stream)."""
        from . import batch as _batch

        self._check_coding_args(scales, means, weights)
        B = scales.size(0)
        if len(items) != B:
            raise ValueError(f"{len(items)} coded images for a batch of {B}")
        dev = scales.device
        bm = [torch.as_tensor(it[2]).reshape(-1) for it in items]
        if all(not t.is_cuda for t in bm):
            bitmap_h = torch.stack(bm).to(torch.int64)
            bitmap = bitmap_h.to(dev)
        else:
            bitmap = torch.stack([t.to(dev) for t in bm]).to(torch.int64)
            bitmap_h = bitmap.cpu()                                  # the one sync of this call
        n_kept = (bitmap_h != 0).sum(1).tolist()
        per_image = self._gather(None, scales, means, weights, bitmap.contiguous(), n_kept)
        streams = [(items[b][0], sc, mu, w, int(items[b][1]) + 1) for b, (_s, sc, mu, w) in enumerate(per_image)]
        live = [b for b in range(B) if n_kept[b] > 0]
        outs = _batch.decode_batch([streams[b] for b in live]) if live else []
        if self.trace is not None:
            it = iter(outs)
            self.trace.append(("dec", [(next(it) if n_kept[b] > 0 else None, *per_image[b][1:]) for b in range(B)],
                               [int(items[b][1]) for b in range(B)]))
        symbols = torch.cat(outs) if len(outs) > 1 else (outs[0] if outs else None)
        self.y_q = None
        return self._scatter(symbols, bitmap.contiguous(), n_kept, (B, bitmap.size(1), scales.size(2), scales.size(3)))


def _norm_ppf(q: float) -> float:
    """Quantile of the standard normal (scipy.stats.norm.ppf in the reference, entropy_models.py:680-682)."""
    try:
        from scipy.stats import norm
        return float(norm.ppf(q))
    except Exception:                                        # pragma: no cover - scipy is in the image
        return float(-math.sqrt(2.0) * torch.erfinv(torch.tensor(1.0 - 2.0 * q, dtype=torch.float64)).item())


class PendingStream(list):
    """``[bytes | None, abs_max, zero_bitmap]`` of a stream registered in an ``EncodeSession``; the bytes are
    filled in by ``flush()``."""


class EncodeSession:
    """Streams registered by ``compress_batch`` across the passes of a codec (checkerboard anchors /
    non-anchors, channel groups), coded together by ``flush()`` in one ``flashgmm_gmm_encode_batch``
    call: the encoder knows every pass's symbols and parameters before the first byte is produced, only
    the decoder has to wait for pass k before it can form pass k+1's parameters."""

    def __init__(self):
        self._streams: List[Tuple] = []
        self._slots: List[list] = []

    def add(self, stream: Tuple, abs_max: int, bitmap: Tensor):
        slot = PendingStream([None, abs_max, bitmap])
        sym = stream[0]
        if sym is None or sym.numel() == 0:
            slot[0] = _EMPTY_STREAM                           # nothing kept: Rans64EncFlush of the initial state
        else:
            self._streams.append(stream)
            self._slots.append(slot)
        return slot

    def flush(self) -> None:
        from . import batch as _batch

        if self._streams:
            for slot, b in zip(self._slots, _batch.encode_batch(self._streams)):
                slot[0] = b
        self._streams, self._slots = [], []


_EMPTY_STREAM = bytes.fromhex("0000008000000000")            # state 2^31 as two little-endian words
