"""The two GMM image-compression architectures, assembled from flashgmm_b200's codecs
(TEST INFRASTRUCTURE: ``/root/reference`` and its Python package do not exist on the GPU box, so
the model-level parity tests build the architectures here).

``Cheng2020AnchorCheckerboardGMMv2``  /root/reference/compressai/models/ckbd_gmm.py:61-141
``Elic2022GMM``                       /root/reference/compressai/models/elic_gmm.py:64-235
with the layer blocks they use (compressai/layers/layers.py:147-336, 391-417, layers/gdn.py:41-93,
ops/parametrizers.py:36-63, models/utils.py:150-168, models/base.py:173-213).

Same module tree, attribute names, parameter shapes AND construction order as the reference, so that
(a) ``state_dict`` keys / shapes are the reference's -- its checkpoints load with ``strict=True`` -- and
(b) ``torch.manual_seed(s)`` before the constructor gives bit-identical random-init weights
(``tests/test_model_cpu.py`` checks both against the reference's own classes where it is present).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from flashgmm_b200 import (ChannelGroupsLatentCodec, CheckerboardLatentCodec, EntropyBottleneck,
                           GaussianMixtureConditionalLatentCodec, HyperLatentCodec, HyperpriorLatentCodec, LowerBound)


# ---- layer blocks ---------------------------------------------------------------------------
def conv3x3(i, o, stride=1):
    return nn.Conv2d(i, o, kernel_size=3, stride=stride, padding=1)


def conv1x1(i, o, stride=1):
    return nn.Conv2d(i, o, kernel_size=1, stride=stride)


def subpel_conv3x3(i, o, r=1):
    return nn.Sequential(nn.Conv2d(i, o * r ** 2, kernel_size=3, padding=1), nn.PixelShuffle(r))


def conv(i, o, kernel_size=5, stride=2):
    return nn.Conv2d(i, o, kernel_size=kernel_size, stride=stride, padding=kernel_size // 2)


def deconv(i, o, kernel_size=5, stride=2):
    return nn.ConvTranspose2d(i, o, kernel_size=kernel_size, stride=stride, output_padding=stride - 1,
                              padding=kernel_size // 2)


class NonNegative(nn.Module):
    def __init__(self, minimum=0.0, reparam_offset=2 ** -18):
        super().__init__()
        self.register_buffer("pedestal", torch.Tensor([reparam_offset ** 2]))
        self.lower_bound = LowerBound((minimum + reparam_offset ** 2) ** 0.5)

    def init(self, x):
        return torch.sqrt(torch.max(x + self.pedestal, self.pedestal))

    def forward(self, x):
        return self.lower_bound(x) ** 2 - self.pedestal


class GDN(nn.Module):
    def __init__(self, ch, inverse=False, beta_min=1e-6, gamma_init=0.1):
        super().__init__()
        self.inverse = bool(inverse)
        self.beta_reparam = NonNegative(minimum=float(beta_min))
        self.beta = nn.Parameter(self.beta_reparam.init(torch.ones(ch)))
        self.gamma_reparam = NonNegative()
        self.gamma = nn.Parameter(self.gamma_reparam.init(float(gamma_init) * torch.eye(ch)))

    def forward(self, x):
        c = x.size(1)
        norm = F.conv2d(x ** 2, self.gamma_reparam(self.gamma).reshape(c, c, 1, 1), self.beta_reparam(self.beta))
        return x * (torch.sqrt(norm) if self.inverse else torch.rsqrt(norm))


class ResidualBlockWithStride(nn.Module):
    def __init__(self, i, o, stride=2):
        super().__init__()
        self.conv1 = conv3x3(i, o, stride=stride)
        self.leaky_relu = nn.LeakyReLU(inplace=True)
        self.conv2 = conv3x3(o, o)
        self.gdn = GDN(o)
        self.skip = conv1x1(i, o, stride=stride) if (stride != 1 or i != o) else None

    def forward(self, x):
        out = self.gdn(self.conv2(self.leaky_relu(self.conv1(x))))
        return out + (x if self.skip is None else self.skip(x))


class ResidualBlockUpsample(nn.Module):
    def __init__(self, i, o, upsample=2):
        super().__init__()
        self.subpel_conv = subpel_conv3x3(i, o, upsample)
        self.leaky_relu = nn.LeakyReLU(inplace=True)
        self.conv = conv3x3(o, o)
        self.igdn = GDN(o, inverse=True)
        self.upsample = subpel_conv3x3(i, o, upsample)

    def forward(self, x):
        return self.igdn(self.conv(self.leaky_relu(self.subpel_conv(x)))) + self.upsample(x)


class ResidualBlock(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.conv1 = conv3x3(i, o)
        self.leaky_relu = nn.LeakyReLU(inplace=True)
        self.conv2 = conv3x3(o, o)
        self.skip = conv1x1(i, o) if i != o else None

    def forward(self, x):
        out = self.leaky_relu(self.conv2(self.leaky_relu(self.conv1(x))))
        return out + (x if self.skip is None else self.skip(x))


class ResidualBottleneckBlock(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        mid = min(i, o) // 2
        self.conv1 = conv1x1(i, mid)
        self.relu1 = nn.ReLU(inplace=True)
        self.conv2 = conv3x3(mid, mid)
        self.relu2 = nn.ReLU(inplace=True)
        self.conv3 = conv1x1(mid, o)
        self.skip = conv1x1(i, o) if i != o else nn.Identity()

    def forward(self, x):
        return self.conv3(self.relu2(self.conv2(self.relu1(self.conv1(x))))) + self.skip(x)


class _ResidualUnit(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.conv = nn.Sequential(conv1x1(n, n // 2), nn.ReLU(inplace=True), conv3x3(n // 2, n // 2),
                                  nn.ReLU(inplace=True), conv1x1(n // 2, n))
        self.relu = nn.ReLU(inplace=True)

    def forward(self, x):
        return self.relu(self.conv(x) + x)


class AttentionBlock(nn.Module):
    def __init__(self, n):
        super().__init__()
        self.conv_a = nn.Sequential(_ResidualUnit(n), _ResidualUnit(n), _ResidualUnit(n))
        self.conv_b = nn.Sequential(_ResidualUnit(n), _ResidualUnit(n), _ResidualUnit(n), conv1x1(n, n))

    def forward(self, x):
        return self.conv_a(x) * torch.sigmoid(self.conv_b(x)) + x


class CheckerboardMaskedConv2d(nn.Conv2d):
    """5x5 conv whose kernel only sees the opposite checkerboard colour (layers/layers.py:118-170, mask type A)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        mask = torch.ones_like(self.weight.data)
        mask[:, :, 0::2, 0::2] = 0
        mask[:, :, 1::2, 1::2] = 0
        self.register_buffer("mask", mask)

    def forward(self, x):
        return F.conv2d(x, self.weight * self.mask, self.bias, self.stride, self.padding, self.dilation, self.groups)


def channel_ramp(i, o, *, min_ch, num_layers, act, **conv_kwargs):
    """``sequential_channel_ramp`` (layers/layers.py:391-417): linearly interpolated widths, floored, inner
    widths clipped to ``min_ch``; conv + activation, no activation after the last conv."""
    widths = torch.linspace(i, o, num_layers + 1).floor().int()
    widths[1:-1] = widths[1:-1].clip(min=min_ch)
    widths = widths.tolist()
    mods = []
    for a, b in zip(widths[:-1], widths[1:]):
        mods += [nn.Conv2d(a, b, **conv_kwargs), act()]
    return nn.Sequential(*mods[:-1])


# ---- models -----------------------------------------------------------------------------------
class _VAE(nn.Module):
    """g_a -> latent codec -> g_s (models/base.py:173-213) + update / aux_loss (models/base.py:117-171)."""

    def forward(self, x):
        y_out = self.latent_codec(self.g_a(x))
        return {"x_hat": self.g_s(y_out["y_hat"]), "likelihoods": y_out["likelihoods"]}

    def compress(self, x):
        return self.latent_codec.compress(self.g_a(x))

    def decompress(self, *args, **kwargs):
        y_hat = self.latent_codec.decompress(*args, **kwargs)["y_hat"]
        return {"x_hat": self.g_s(y_hat).clamp_(0, 1)}

    def update(self, scale_table=None, force=False, update_quantiles=False):
        if scale_table is None:
            scale_table = torch.exp(torch.linspace(math.log(0.11), math.log(256), 64))
        updated = False
        for m in self.modules():
            if type(m).__name__ == "EntropyBottleneck":
                updated |= m.update(force=force, update_quantiles=update_quantiles)
            if type(m).__name__ == "GaussianMixtureConditional":
                updated |= m.update_scale_table(scale_table, force=force)
        return updated

    def aux_loss(self):
        return sum(m.loss() for m in self.modules() if type(m).__name__ == "EntropyBottleneck")


class Cheng2020AnchorCheckerboardGMMv2(_VAE):
    def __init__(self, N=192, K=4, quantizer="noise"):
        super().__init__()
        self.K = K
        self.g_a = nn.Sequential(
            ResidualBlockWithStride(3, N, stride=2), ResidualBlock(N, N),
            ResidualBlockWithStride(N, N, stride=2), ResidualBlock(N, N),
            ResidualBlockWithStride(N, N, stride=2), ResidualBlock(N, N),
            conv3x3(N, N, stride=2))
        self.g_s = nn.Sequential(
            ResidualBlock(N, N), ResidualBlockUpsample(N, N, 2),
            ResidualBlock(N, N), ResidualBlockUpsample(N, N, 2),
            ResidualBlock(N, N), ResidualBlockUpsample(N, N, 2),
            ResidualBlock(N, N), subpel_conv3x3(N, 3, 2))
        lrelu = lambda: nn.LeakyReLU(inplace=True)  # noqa: E731
        h_a = nn.Sequential(conv3x3(N, N), lrelu(), conv3x3(N, N), lrelu(), conv3x3(N, N, stride=2), lrelu(),
                            conv3x3(N, N), lrelu(), conv3x3(N, N, stride=2))
        h_s = nn.Sequential(conv3x3(N, N), lrelu(), subpel_conv3x3(N, N, 2), lrelu(), conv3x3(N, N * 3 // 2), lrelu(),
                            subpel_conv3x3(N * 3 // 2, N * 3 // 2, 2), lrelu(), conv3x3(N * 3 // 2, N * 2))
        self.latent_codec = HyperpriorLatentCodec(latent_codec={
            "y": CheckerboardLatentCodec(
                latent_codec={"y": GaussianMixtureConditionalLatentCodec(K=K, quantizer=quantizer)},
                entropy_parameters=nn.Sequential(nn.Conv2d(N * 12 // 3, N * 10 // 3, 1), lrelu(),
                                                 nn.Conv2d(N * 10 // 3, N * 10 // 3, 1), lrelu(),
                                                 nn.Conv2d(N * 10 // 3, 3 * K * N, 1)),
                context_prediction=CheckerboardMaskedConv2d(N, 2 * N, kernel_size=5, stride=1, padding=2),
                forward_method="onepass"),
            "hyper": HyperLatentCodec(entropy_bottleneck=EntropyBottleneck(N), h_a=h_a, h_s=h_s, quantizer="ste"),
        })


class Elic2022GMM(_VAE):
    def __init__(self, N=192, M=320, K=4, quantizer="noise", groups=None):
        super().__init__()
        self.K = K
        self.groups = list(groups) if groups is not None else [16, 16, 32, 64, M - 128]
        if sum(self.groups) != M:
            raise ValueError("groups must sum to M")
        G = self.groups
        rbb3 = lambda: [ResidualBottleneckBlock(N, N) for _ in range(3)]  # noqa: E731
        self.g_a = nn.Sequential(conv(3, N), *rbb3(), conv(N, N), *rbb3(), AttentionBlock(N), conv(N, N), *rbb3(),
                                 conv(N, M), AttentionBlock(M))
        self.g_s = nn.Sequential(AttentionBlock(M), deconv(M, N), *rbb3(), deconv(N, N), AttentionBlock(N), *rbb3(),
                                 deconv(N, N), *rbb3(), deconv(N, 3))
        relu = lambda: nn.ReLU(inplace=True)  # noqa: E731
        h_a = nn.Sequential(conv(M, N, kernel_size=3, stride=1), relu(), conv(N, N), relu(), conv(N, N))
        h_s = nn.Sequential(deconv(N, N), relu(), deconv(N, N * 3 // 2), relu(),
                            deconv(N * 3 // 2, N * 2, kernel_size=3, stride=1))
        channel_context = {
            f"y{t}": channel_ramp(sum(G[:t]), G[t] * 2, min_ch=N, num_layers=3, act=relu, kernel_size=5, stride=1, padding=2)
            for t in range(1, len(G))}
        spatial_context = [CheckerboardMaskedConv2d(G[t], G[t] * 2, kernel_size=5, stride=1, padding=2)
                           for t in range(len(G))]
        param_aggregation = [
            channel_ramp(G[t] * 2 + (t > 0) * G[t] * 2 + N * 2, G[t] * 3 * K, min_ch=N * 2, num_layers=3, act=relu,
                         kernel_size=1, stride=1, padding=0)
            for t in range(len(G))]
        scctx = {
            f"y{t}": CheckerboardLatentCodec(
                latent_codec={"y": GaussianMixtureConditionalLatentCodec(K=K, quantizer=quantizer)},
                context_prediction=spatial_context[t], entropy_parameters=param_aggregation[t], forward_method="onepass")
            for t in range(len(G))}
        self.latent_codec = HyperpriorLatentCodec(latent_codec={
            "y": ChannelGroupsLatentCodec(groups=G, channel_context=channel_context, latent_codec=scctx),
            "hyper": HyperLatentCodec(entropy_bottleneck=EntropyBottleneck(N), h_a=h_a, h_s=h_s, quantizer="ste"),
        })


# ---- plausible parameters for random-init models (tests and bench.py's model_codec block) -------------------
class SyntheticMixtures:
    """Replaces ``_mixture`` of every GMM latent codec by a table of synthetic parameters, one entry per call in
    coding order (the encoder and the decoder make the same calls), and draws y from them."""

    def __init__(self, net, seed=1234):
        self.codecs = [m for m in net.modules() if isinstance(m, GaussianMixtureConditionalLatentCodec)]
        self.seed = seed
        self.calls = {}

    def params(self, codec_i, call_i, B, C, H, W, K, dev):
        g = torch.Generator().manual_seed(self.seed + 1000 * codec_i + call_i)
        means = 3.0 * torch.randn(B, K * C, H, W, generator=g)
        scales = 0.11 + 4.0 * torch.rand(B, K * C, H, W, generator=g)
        weights = torch.softmax(torch.randn(B, K, C, H, W, generator=g), dim=1).reshape(B, K * C, H, W)
        comp = torch.multinomial(weights.reshape(B, K, -1).permute(0, 2, 1).reshape(-1, K), 1, generator=g).reshape(B, 1, C, H, W)
        pick = lambda t: torch.gather(t.reshape(B, K, C, H, W), 1, comp).squeeze(1)  # noqa: E731
        y = pick(means) + pick(scales) * torch.randn(B, C, H, W, generator=g)
        return scales.to(dev), means.to(dev), weights.to(dev), y.to(dev)

    def __enter__(self):
        for i, codec in enumerate(self.codecs):
            def fake(ctx_params, codec=codec, i=i):
                n = self.calls.get(i, 0)
                self.calls[i] = n + 1
                B, _, H, W = ctx_params.shape
                C = self.channels[i]
                sc, mu, w, _ = self.params(i, n % 2, B, C, H, W, codec.K, ctx_params.device)
                return sc, mu, w, None
            codec._mixture = fake
        return self

    def __exit__(self, *exc):
        for codec in self.codecs:
            del codec._mixture
        return False
