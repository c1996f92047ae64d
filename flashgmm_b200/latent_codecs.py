"""Latent codecs of the GMM models with BATCHED, device-resident compress / decompress.

Drop-ins (same constructor arguments, sub-module names -- so checkpoints load unchanged --, and the
dict-valued returns of ``forward / compress / decompress``) for the codecs that the reference's
GMM models are assembled from:

==============================================  =========================================================
here                                            reference (under /root/reference/compressai/latent_codecs/)
==============================================  =========================================================
``GaussianMixtureConditionalLatentCodec``       gaussian_mixture_conditional.py:43-202
``CheckerboardLatentCodec``                     checkerboard.py:61-418 (forward: ``onepass``, the only method
                                                the GMM models use -- models/ckbd_gmm.py:127, elic_gmm.py:203)
``ChannelGroupsLatentCodec``                    channel_groups.py:44-173
``HyperLatentCodec`` / ``HyperpriorLatentCodec``  hyper.py:46-110 / hyperprior.py:45-139
==============================================  =========================================================

What is different from the reference:

* ``compress`` / ``decompress`` take a batch of B images. The reference codes one image per call
  (``n = 1`` at checkerboard.py:307, channel_groups.py:138; ``.squeeze(0)`` in the entropy model); here
  every coding pass hands its B streams to the batched device coder in one call
  (``GaussianMixtureConditional.compress_batch / decompress_batch``). At B = 1 the returned structure
  is the reference's; at B > 1 every per-pass entry of ``strings`` is a list of B
  ``(bytes, abs_max, zero_bitmap)`` tuples, image b's tuple being exactly what coding image b alone
  gives.
* ``compress`` is not pass-sequential: the encoder knows every pass's ``y`` up front, so the passes only
  REGISTER their streams in an ``EncodeSession`` and the outermost codec codes all of them (2 passes x B
  for the checkerboard model, 10 x B for ELIC) in ONE batched call. ``decompress`` has to be
  pass-sequential (pass k+1's parameters need pass k's symbols): one batched decode call per pass.
* the small networks between two passes (masked context conv, 1x1 parameter convs, softmax, the
  checkerboard (un)embedding) are replayed from CUDA graphs (``graphs.NetRunner``).

``mount(net)`` installs all of this into an existing model tree (the reference's own
``Cheng2020AnchorCheckerboardGMMv2`` / ``Elic2022GMM`` instances included): see its docstring.
"""
from __future__ import annotations

import threading
import types
from itertools import accumulate
from typing import Any, Dict, List, Mapping, Optional, Sequence, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

from .entropy_models import EncodeSession, EntropyBottleneck, GaussianMixtureConditional, PendingStream
from .graphs import NetRunner, default_runner, thread_runner


def quantize_ste(x: Tensor) -> Tensor:
    """compressai/ops/ops.py:66: round with a straight-through gradient."""
    return (torch.round(x) - x).detach() + x


# ----------------------------------------------------------------------------------------------
# encode sessions: opened by the outermost codec, used by every GMM codec below it
_tls = threading.local()


class _session_scope:
    """``with _session_scope() as (session, owner)``: owner is True for the outermost scope, which flushes."""

    def __enter__(self):
        cur = getattr(_tls, "session", None)
        self.owner = cur is None
        if self.owner:
            cur = _tls.session = EncodeSession()
        return cur, self.owner

    def __exit__(self, *exc):
        if self.owner:
            _tls.session = None
        return False


def _finish(strings):
    """Pending slots (mutable ``[bytes, abs_max, bitmap]``) -> the reference's tuples, in place of the nesting."""
    out = []
    for s in strings:
        if isinstance(s, PendingStream):
            out.append(tuple(s))
        elif isinstance(s, list) and s and isinstance(s[0], PendingStream):
            out.append([tuple(x) for x in s])
        else:
            out.append(s)
    return out


def _runner_of(module) -> NetRunner:
    """The calling thread's runner inside a ``Lane`` (graphs have static buffers: one runner per host thread), else the
    one ``mount(net, runner=...)`` gave the module, else the process-wide default."""
    return thread_runner() or getattr(module, "_net_runner", None) or default_runner()


def _wkey(module: nn.Module, tag: str):
    p = next(module.parameters(), None)
    return (tag, id(module), p.data_ptr() if p is not None else 0)


# ----------------------------------------------------------------------------------------------
class GaussianMixtureConditionalLatentCodec(nn.Module):
    """Turns the parameter tensor of one pass into (scales, means, weights) and codes y with them.

    ``entropy_parameters`` output is split in three along the channels; the weights get a softmax over the K
    components of every channel (gaussian_mixture_conditional.py:183-202). ``quantizer``: ``"noise"``
    (y_hat = round(y) at coding time) or ``"weighted_mean_ste"`` (y is centred on the mixture's mean before
    rounding, :109-123, 135-145, 167-179)."""

    gaussian_mixture_conditional: GaussianMixtureConditional
    entropy_parameters: nn.Module

    def __init__(self, K: int = 4, scale_table: Optional[Union[List, Tuple]] = None,
                 gaussian_mixture_conditional: Optional[GaussianMixtureConditional] = None,
                 entropy_parameters: Optional[nn.Module] = None, quantizer: str = "noise",
                 chunks: Tuple[str, ...] = ("scales", "means", "weights"), **kwargs):
        super().__init__()
        if quantizer not in ("noise", "weighted_mean_ste"):
            raise AssertionError(f"quantizer {quantizer} not supported")
        if tuple(chunks) != ("scales", "means", "weights"):
            raise ValueError("a Gaussian-mixture codec needs the three chunks (scales, means, weights)")
        self.K = K
        self.quantizer = quantizer
        self.chunks = tuple(chunks)
        self.gaussian_mixture_conditional = gaussian_mixture_conditional or GaussianMixtureConditional(
            K=K, scale_table=scale_table)
        self.entropy_parameters = entropy_parameters or nn.Identity()

    def _mixture(self, ctx_params: Tensor):
        """-> scales, means, weights (softmax over K), and for the ste quantiser the mixture mean."""
        p = self.entropy_parameters(ctx_params)
        scales, means, logits = p.chunk(3, dim=1)
        B, KC, H, W = logits.shape
        weights = F.softmax(logits.reshape(B, self.K, KC // self.K, H, W), dim=1).reshape(B, KC, H, W)
        centre = None
        if self.quantizer == "weighted_mean_ste":
            m = means.reshape(B, self.K, KC // self.K, H, W)
            centre = (m * weights.reshape(B, self.K, KC // self.K, H, W)).sum(dim=1)
            means = (m - centre.unsqueeze(1)).reshape(B, KC, H, W)
        return scales, means, weights, centre

    # reference helper names, kept for callers that use them directly (:183-202)
    def _chunk(self, params: Tensor):
        return params.chunk(3, dim=1)

    def _reshape_gmm_weight(self, weight: Tensor) -> Tensor:
        B, KC, H, W = weight.shape
        return F.softmax(weight.reshape(B, self.K, KC // self.K, H, W), dim=1).reshape(B, KC, H, W)

    def forward(self, y: Tensor, ctx_params: Tensor) -> Dict[str, Any]:
        scales, means, weights, centre = self._mixture(ctx_params)
        if centre is not None:
            y = quantize_ste(y - centre) + centre
        y_hat, lik = self.gaussian_mixture_conditional(y, scales, means, weights)
        return {"likelihoods": {"y": lik}, "y_hat": y_hat}

    def compress(self, y: Tensor, ctx_params: Tensor) -> Dict[str, Any]:
        scales, means, weights, centre = self._mixture(ctx_params)
        if centre is not None:
            y = quantize_ste(y - centre)
        gmc = self.gaussian_mixture_conditional
        if y.size(0) == 1 and getattr(_tls, "session", None) is None:
            y_strings, y_hat = gmc.compress(y, scales, means, weights)            # the reference's call shape
        else:
            items, y_hat = gmc.compress_batch(y, scales, means, weights, session=getattr(_tls, "session", None))
            y_strings = items[0] if y.size(0) == 1 else items
        return {"strings": [y_strings], "shape": y.shape[2:4], "y_hat": y_hat}

    def decompress(self, strings, shape: Tuple[int, int], ctx_params: Tensor, **kwargs) -> Dict[str, Any]:
        (y_strings,) = strings
        scales, means, weights, centre = self._mixture(ctx_params)
        gmc = self.gaussian_mixture_conditional
        if isinstance(y_strings, tuple):                                          # one image, the reference's structure
            y_hat = gmc.decompress(*y_strings, scales, means, weights)
        else:
            y_hat = gmc.decompress_batch(y_strings, scales, means, weights)
        if centre is not None:
            y_hat = y_hat + centre
        if tuple(y_hat.shape[2:4]) != tuple(shape):
            raise AssertionError(f"decoded shape {tuple(y_hat.shape[2:4])} != {tuple(shape)}")
        return {"y_hat": y_hat}


# ----------------------------------------------------------------------------------------------
def _masked_conv(conv: nn.Module, x: Tensor) -> Tensor:
    """``CheckerboardMaskedConv2d.forward`` (layers/layers.py:138-140) without rebinding ``weight.data``
    (the mask is 0/1, so masking again gives the same weights): safe inside a CUDA-graph capture."""
    mask = getattr(conv, "mask", None)
    if mask is None or not isinstance(conv, nn.Conv2d):
        return conv(x)
    return F.conv2d(x, conv.weight * mask, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups)


class CheckerboardLatentCodec(nn.Module):
    """Two-pass checkerboard context model (He et al. 2021): anchors are coded from the side parameters
    alone, non-anchors additionally from a masked convolution of the decoded anchors."""

    def __init__(self, latent_codec: Optional[Mapping[str, nn.Module]] = None,
                 entropy_parameters: Optional[nn.Module] = None, context_prediction: Optional[nn.Module] = None,
                 anchor_parity: str = "even", forward_method: str = "twopass", **kwargs):
        super().__init__()
        if anchor_parity not in ("even", "odd"):
            raise ValueError('anchor_parity must be "even" or "odd"')
        self._kwargs = kwargs
        self.anchor_parity = anchor_parity
        self.non_anchor_parity = "odd" if anchor_parity == "even" else "even"
        self.forward_method = forward_method
        self.entropy_parameters = entropy_parameters or nn.Identity()
        self.context_prediction = context_prediction or nn.Identity()
        latent_codec = dict(latent_codec or {})
        if "y" not in latent_codec:
            latent_codec["y"] = GaussianMixtureConditionalLatentCodec()
        for k, v in latent_codec.items():                     # registered directly, like the reference (save_direct)
            setattr(self, k, v)
        self.latent_codec = latent_codec

    def __getitem__(self, key: str) -> nn.Module:
        return self.latent_codec[key]

    # -- checkerboard (un)embedding: pure data movement ---------------------------------------------
    def _split(self, t: Tensor) -> Tuple[Tensor, Tensor]:
        """(n, c, h, w) -> anchors, non-anchors, each (n, c, h, w/2): row r keeps the columns whose parity
        equals (r + [anchor_parity == "odd"]) % 2 as anchors (checkerboard.py:332-354)."""
        h, w = t.shape[-2:]
        pairs = t.unflatten(-1, (w // 2, 2))
        first, second = pairs[..., 0], pairs[..., 1]
        odd_row = (torch.arange(h, device=t.device) % 2 == 1).view(h, 1)
        a, b = torch.where(odd_row, second, first), torch.where(odd_row, first, second)
        return (a, b) if self.anchor_parity == "even" else (b, a)

    def _join(self, anchors: Tensor, non_anchors: Tensor) -> Tensor:
        """Inverse of ``_split`` (checkerboard.py:356-378)."""
        if self.anchor_parity != "even":
            anchors, non_anchors = non_anchors, anchors
        h = anchors.shape[-2]
        odd_row = (torch.arange(h, device=anchors.device) % 2 == 1).view(h, 1)
        first, second = torch.where(odd_row, non_anchors, anchors), torch.where(odd_row, anchors, non_anchors)
        return torch.stack((first, second), dim=-1).flatten(-2)

    def unembed(self, y: Tensor) -> Tensor:
        return torch.stack(self._split(y))

    def embed(self, y_: Tensor) -> Tensor:
        return self._join(y_[0], y_[1])

    def _keep_non_anchors(self, t: Tensor) -> Tensor:
        a, b = self._split(t)
        return self._join(torch.zeros_like(a), b)

    def merge(self, *args):
        return torch.cat(args, dim=1)

    def quantize(self, y: Tensor) -> Tensor:
        if self.training:
            return y + torch.empty_like(y).uniform_(-0.5, 0.5)
        return torch.round(y)

    # -- training / likelihood evaluation -----------------------------------------------------------
    def forward(self, y: Tensor, side_params: Tensor) -> Dict[str, Any]:
        if self.forward_method != "onepass":
            raise ValueError(f'forward_method "{self.forward_method}": the GMM codecs run "onepass" only '
                             "(models/ckbd_gmm.py:127)")
        # checkerboard.py:154-171: context from the (noisy / rounded) latent at the non-anchor positions
        y_hat = self.quantize(y)
        y_ctx = self._keep_non_anchors(_masked_conv(self.context_prediction, y_hat))
        params = self.entropy_parameters(self.merge(y_ctx, side_params))
        y_out = self.latent_codec["y"](y, params)
        return {"likelihoods": {"y": y_out["likelihoods"]["y"]}, "y_hat": y_hat}

    # -- the step between two passes, replayed from a CUDA graph ------------------------------------
    def _ctx_channels(self, c: int) -> int:
        w = getattr(self.context_prediction, "weight", None)
        return int(w.shape[0]) if w is not None else c

    def _pass_params(self, i: int, anchors_hat: Optional[Tensor], side_half: Tensor, c: int) -> Tensor:
        """Parameters of pass i at the half-resolution positions: entropy_parameters(cat(ctx_i, side_i));
        ctx_0 = 0 (checkerboard.py:283-286: the anchors' context is masked out entirely)."""
        runner = _runner_of(self)
        if i == 0:
            def step0(side):
                n, _, h, w2 = side.shape
                ctx = side.new_zeros((n, self._ctx_channels(c), h, w2))
                return (self.entropy_parameters(torch.cat((ctx, side), dim=1)),)
            return runner(_wkey(self.entropy_parameters, "ckbd0") + (self._ctx_channels(c),), step0, side_half)[0]

        def step1(anc, side):
            full = self._join(anc, torch.zeros_like(anc))
            ctx = self._split(_masked_conv(self.context_prediction, full))[1]
            return (self.entropy_parameters(torch.cat((ctx, side), dim=1)),)
        return runner(_wkey(self.context_prediction, "ckbd1") + _wkey(self.entropy_parameters, "ep"), step1,
                      anchors_hat, side_half)[0]

    # -- checkerboard.py:275-297 ----------------------------------------------------------------------
    def compress(self, y: Tensor, side_params: Tensor) -> Dict[str, Any]:
        with _session_scope() as (session, owner), torch.no_grad():
            y_halves = self._split(y)
            side_halves = self._split(side_params)
            y_hat_halves: List[Tensor] = [None, None]
            strings = [None, None]
            for i in range(2):
                params_i = self._pass_params(i, y_hat_halves[0], side_halves[i].contiguous(), y.size(1))
                out = self.latent_codec["y"].compress(y_halves[i].contiguous(), params_i)
                y_hat_halves[i] = out["y_hat"]
                [strings[i]] = out["strings"]
            y_hat = self._join(y_hat_halves[0], y_hat_halves[1])
            if owner:
                session.flush()
                strings = _finish(strings)
        return {"strings": strings, "shape": y_hat.shape[1:], "y_hat": y_hat}

    # -- checkerboard.py:299-330 ----------------------------------------------------------------------
    def decompress(self, strings, shape: Tuple[int, ...], side_params: Tensor, **kwargs) -> Dict[str, Any]:
        c, h, w = shape
        with torch.no_grad():
            side_halves = self._split(side_params)
            y_hat_halves: List[Tensor] = [None, None]
            for i in range(2):
                params_i = self._pass_params(i, y_hat_halves[0], side_halves[i].contiguous(), c)
                out = self.latent_codec["y"].decompress([strings[i]], (h, w // 2), params_i)
                y_hat_halves[i] = out["y_hat"]
            y_hat = self._join(y_hat_halves[0], y_hat_halves[1])
        return {"y_hat": y_hat}


# ----------------------------------------------------------------------------------------------
class ChannelGroupsLatentCodec(nn.Module):
    """Channel-conditional coding (Minnen & Singh 2020; ELIC's uneven groups): group k's parameters come
    from the side parameters and a convolution stack over the already decoded groups < k."""

    def __init__(self, latent_codec: Optional[Mapping[str, nn.Module]] = None,
                 channel_context: Optional[Mapping[str, nn.Module]] = None, *, groups: List[int], **kwargs):
        super().__init__()
        self._kwargs = kwargs
        self.groups = list(groups)
        self.groups_acc = list(accumulate(self.groups, initial=0))
        self.channel_context = nn.ModuleDict(channel_context)
        self.latent_codec = nn.ModuleDict(latent_codec)

    def __getitem__(self, key: str) -> nn.Module:
        return self.latent_codec[key]

    def merge_y(self, *args):
        return torch.cat(args, dim=1)

    def merge_params(self, *args):
        return torch.cat(args, dim=1)

    def _get_ctx_params(self, k: int, side_params: Tensor, y_hat_: Sequence[Tensor]) -> Tensor:
        if k == 0:
            return side_params
        return self.merge_params(self.channel_context[f"y{k}"](self.merge_y(*y_hat_[:k])), side_params)

    def _ctx_params_graphed(self, k: int, side_params: Tensor, y_hat: Tensor) -> Tensor:
        if k == 0:
            return side_params
        net = self.channel_context[f"y{k}"]
        lo = self.groups_acc[k]

        def step(prev, side):
            return (torch.cat((net(prev), side), dim=1),)
        return _runner_of(self)(_wkey(net, f"chctx{k}"), step, y_hat[:, :lo].contiguous(), side_params)[0]

    def forward(self, y: Tensor, side_params: Tensor) -> Dict[str, Any]:
        y_groups = torch.split(y, self.groups, dim=1)
        y_hats: List[Tensor] = []
        liks: List[Tensor] = []
        for k in range(len(self.groups)):
            out = self.latent_codec[f"y{k}"](y_groups[k], self._get_ctx_params(k, side_params, y_hats))
            y_hats.append(out["y_hat"])
            liks.append(out["likelihoods"]["y"])
        return {"likelihoods": {"y": torch.cat(liks, dim=1)}, "y_hat": torch.cat(y_hats, dim=1)}

    # -- channel_groups.py:111-129 --------------------------------------------------------------------
    def compress(self, y: Tensor, side_params: Tensor) -> Dict[str, Any]:
        with _session_scope() as (session, owner), torch.no_grad():
            y_hat = torch.zeros_like(y)
            strings: List[Any] = []
            shapes = []
            per_group = None
            for k in range(len(self.groups)):
                lo, hi = self.groups_acc[k], self.groups_acc[k + 1]
                out = self.latent_codec[f"y{k}"].compress(y[:, lo:hi].contiguous(),
                                                          self._ctx_params_graphed(k, side_params, y_hat))
                y_hat[:, lo:hi] = out["y_hat"]
                if per_group is None:
                    per_group = len(out["strings"])
                elif per_group != len(out["strings"]):
                    raise AssertionError("every group must produce the same number of streams")
                strings.extend(out["strings"])
                shapes.append(out["shape"])
            if owner:
                session.flush()
                strings = _finish(strings)
        return {"strings": strings, "shape": shapes, "y_hat": y_hat}

    # -- channel_groups.py:131-158 --------------------------------------------------------------------
    def decompress(self, strings, shape, side_params: Tensor, **kwargs) -> Dict[str, Any]:
        per_group = len(strings) // len(self.groups)
        c_total = sum(s[0] for s in shape)
        with torch.no_grad():
            y_hat = torch.zeros((side_params.size(0), c_total, *shape[0][1:]), device=side_params.device)
            for k in range(len(self.groups)):
                lo, hi = self.groups_acc[k], self.groups_acc[k + 1]
                out = self.latent_codec[f"y{k}"].decompress(strings[per_group * k: per_group * (k + 1)], shape[k],
                                                            self._ctx_params_graphed(k, side_params, y_hat))
                y_hat[:, lo:hi] = out["y_hat"]
        return {"y_hat": y_hat}


# ----------------------------------------------------------------------------------------------
class HyperLatentCodec(nn.Module):
    """The side-information branch: z = h_a(y) through the factorised prior, parameters = h_s(z_hat)."""

    def __init__(self, entropy_bottleneck: Optional[nn.Module] = None, h_a: Optional[nn.Module] = None,
                 h_s: Optional[nn.Module] = None, quantizer: str = "noise", **kwargs):
        super().__init__()
        if entropy_bottleneck is None:
            raise AssertionError("HyperLatentCodec needs an entropy_bottleneck")
        self.entropy_bottleneck = entropy_bottleneck
        self.h_a = h_a or nn.Identity()
        self.h_s = h_s or nn.Identity()
        self.quantizer = quantizer

    def forward(self, y: Tensor) -> Dict[str, Any]:
        z = self.h_a(y)
        z_hat, z_lik = self.entropy_bottleneck(z)
        if self.quantizer == "ste":
            med = self.entropy_bottleneck._get_medians()
            z_hat = quantize_ste(z - med) + med
        return {"likelihoods": {"z": z_lik}, "params": self.h_s(z_hat)}

    def compress(self, y: Tensor) -> Dict[str, Any]:
        with torch.no_grad():
            (z,) = _runner_of(self)(_wkey(self.h_a, "h_a"), lambda t: (self.h_a(t),), y)
            shape = z.size()[-2:]
            z_strings = self.entropy_bottleneck.compress(z)
            z_hat = self.entropy_bottleneck.decompress(z_strings, shape)          # what the decoder will see
            (params,) = _runner_of(self)(_wkey(self.h_s, "h_s"), lambda t: (self.h_s(t),), z_hat)
        return {"strings": [z_strings], "shape": shape, "params": params}

    def decompress(self, strings, shape: Tuple[int, int], **kwargs) -> Dict[str, Any]:
        (z_strings,) = strings
        with torch.no_grad():
            z_hat = self.entropy_bottleneck.decompress(z_strings, shape)
            (params,) = _runner_of(self)(_wkey(self.h_s, "h_s"), lambda t: (self.h_s(t),), z_hat)
        return {"params": params}


class HyperpriorLatentCodec(nn.Module):
    """y coded with parameters reconstructed from the hyper branch (Balle et al. 2018)."""

    def __init__(self, latent_codec: Optional[Mapping[str, nn.Module]] = None, **kwargs):
        super().__init__()
        latent_codec = dict(latent_codec or {})
        if "y" not in latent_codec or "hyper" not in latent_codec:
            raise ValueError('HyperpriorLatentCodec needs the "y" and "hyper" codecs')
        for k, v in latent_codec.items():                     # registered directly, like the reference (save_direct)
            setattr(self, k, v)
        self.latent_codec = latent_codec

    def __getitem__(self, key: str) -> nn.Module:
        return self.latent_codec[key]

    def forward(self, y: Tensor) -> Dict[str, Any]:
        hyper = self.latent_codec["hyper"](y)
        y_out = self.latent_codec["y"](y, hyper["params"])
        return {"likelihoods": {"y": y_out["likelihoods"]["y"], "z": hyper["likelihoods"]["z"]}, "y_hat": y_out["y_hat"]}

    def compress(self, y: Tensor) -> Dict[str, Any]:
        with _session_scope() as (session, owner):
            hyper = self.latent_codec["hyper"].compress(y)
            y_out = self.latent_codec["y"].compress(y, hyper["params"])
            [z_strings] = hyper["strings"]
            strings = [*y_out["strings"], z_strings]
            if owner:
                session.flush()
                strings = _finish(strings)
        return {"strings": strings, "shape": {"y": y_out["shape"], "hyper": hyper["shape"]}, "y_hat": y_out["y_hat"]}

    def decompress(self, strings, shape: Dict[str, Any], **kwargs) -> Dict[str, Any]:
        *y_strings, z_strings = strings
        hyper = self.latent_codec["hyper"].decompress([z_strings], shape["hyper"])
        y_out = self.latent_codec["y"].decompress(y_strings, shape["y"], hyper["params"])
        return {"y_hat": y_out["y_hat"]}


# ----------------------------------------------------------------------------------------------
_BY_NAME = {
    "GaussianMixtureConditionalLatentCodec": GaussianMixtureConditionalLatentCodec,
    "CheckerboardLatentCodec": CheckerboardLatentCodec,
    "ChannelGroupsLatentCodec": ChannelGroupsLatentCodec,
    "HyperLatentCodec": HyperLatentCodec,
    "HyperpriorLatentCodec": HyperpriorLatentCodec,
}
_CODING_METHODS = ("compress", "decompress", "_pass_params", "_ctx_params_graphed", "_ctx_channels", "_split", "_join",
                   "_mixture", "_keep_non_anchors")


def mount(net: nn.Module, runner: Optional[NetRunner] = None) -> nn.Module:
    """Install the B200 path into an existing model tree, in place, and return it.

    Works on the reference's own model instances (``compressai.models.ckbd_gmm.
    Cheng2020AnchorCheckerboardGMMv2``, ``compressai.models.elic_gmm.Elic2022GMM``) and on any tree built
    from classes with the same names and attributes; modules are recognised by class NAME, nothing is
    imported from ``compressai``.

    * every ``GaussianMixtureConditional`` / ``EntropyBottleneck`` instance is replaced by this package's
      class, constructed with the old instance's hyper-parameters and loaded with its ``state_dict``
      (before or after ``update()``);
    * every checkerboard / channel-group / hyper / hyperprior / GMM latent codec keeps its identity, its
      sub-modules and its ``forward`` -- only its coding methods (``compress``, ``decompress`` and their
      helpers) are rebound to the batched device implementations above.

    ``state_dict`` keys are unchanged, so checkpoints keep loading into the mounted model.
    """
    def swap_entropy_models(parent: nn.Module):
        for name, child in list(parent.named_children()):
            cls = type(child).__name__
            if cls == "GaussianMixtureConditional" and not isinstance(child, GaussianMixtureConditional):
                new = GaussianMixtureConditional(K=child.K, scale_bound=float(child.lower_bound_scale.bound),
                                                 tail_mass=getattr(child, "tail_mass", 1e-9),
                                                 likelihood_bound=(float(child.likelihood_lower_bound.bound)
                                                                   if getattr(child, "use_likelihood_bound", False) else 0.0),
                                                 entropy_coder_precision=child.entropy_coder_precision)
            elif cls == "EntropyBottleneck" and not isinstance(child, EntropyBottleneck):
                new = EntropyBottleneck(child.channels, tail_mass=child.tail_mass, init_scale=child.init_scale,
                                        filters=child.filters,
                                        likelihood_bound=(float(child.likelihood_lower_bound.bound)
                                                          if getattr(child, "use_likelihood_bound", False) else 0.0),
                                        entropy_coder_precision=child.entropy_coder_precision)
            else:
                swap_entropy_models(child)
                continue
            ref_tensor = next(iter(child.state_dict().values()))
            new.to(ref_tensor.device)
            new.load_state_dict(child.state_dict(), strict=True)
            new.train(child.training)
            setattr(parent, name, new)
            # codecs that keep their children in a plain dict as well (save_direct) must see the new module
            lc = getattr(parent, "latent_codec", None)
            if isinstance(lc, dict) and lc.get(name) is child:
                lc[name] = new

    swap_entropy_models(net)
    for module in net.modules():
        impl = _BY_NAME.get(type(module).__name__)
        if impl is None or isinstance(module, impl):
            if impl is not None and runner is not None:
                module._net_runner = runner
            continue
        for meth in _CODING_METHODS:
            fn = impl.__dict__.get(meth)
            if fn is not None:
                setattr(module, meth, types.MethodType(fn, module))
        if impl is CheckerboardLatentCodec and not hasattr(module, "non_anchor_parity"):
            module.non_anchor_parity = "odd" if module.anchor_parity == "even" else "even"
        if runner is not None:
            module._net_runner = runner
    return net
