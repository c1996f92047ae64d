"""``GaussianMixtureConditionalLatentCodec`` -- drop-in for
/root/reference/compressai/latent_codecs/gaussian_mixture_conditional.py:43-202.

Same constructor arguments, sub-module names (``gaussian_mixture_conditional``,
``entropy_parameters`` -- checkpoints load unchanged), the ``noise`` and
``weighted_mean_ste`` quantisers, and the dict-valued returns of
``forward / compress / decompress``. The reference's debug ``print`` calls
(constructor and per-call timing, :95, :146, :166) are not reproduced.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple, Union

import torch
import torch.nn as nn
from torch import Tensor

from .entropy_models import GaussianMixtureConditional


def quantize_ste(x: Tensor) -> Tensor:
    """compressai/ops/ops.py:66: round with a straight-through gradient."""
    return (torch.round(x) - x).detach() + x


class GaussianMixtureConditionalLatentCodec(nn.Module):
    gaussian_mixture_conditional: GaussianMixtureConditional
    entropy_parameters: nn.Module

    def __init__(
        self,
        K: int = 4,
        scale_table: Optional[Union[List, Tuple]] = None,
        gaussian_mixture_conditional: Optional[GaussianMixtureConditional] = None,
        entropy_parameters: Optional[nn.Module] = None,
        quantizer: str = "noise",
        chunks: Tuple[str] = ("scales", "means", "weights"),
        **kwargs,
    ):
        super().__init__()
        assert quantizer in ["noise", "weighted_mean_ste"], f"quantizer {quantizer} not supported"
        self.K = K
        self.quantizer = quantizer
        self.gaussian_mixture_conditional = (
            gaussian_mixture_conditional
            if gaussian_mixture_conditional is not None
            else GaussianMixtureConditional(K=K, scale_table=scale_table)
        )
        self.entropy_parameters = entropy_parameters or nn.Identity()
        self.chunks = tuple(chunks)

    # gaussian_mixture_conditional.py:109-123 / 135-145 / 167-179
    def _recentre(self, means_hat: Tensor, weights: Tensor):
        """K x C means and weights -> (weighted mean (B,C,H,W), means minus it)."""
        B, _, H, W = means_hat.shape
        m = means_hat.view(B, self.K, -1, H, W)
        w = weights.view(B, self.K, -1, H, W)
        weighted_sum = torch.sum(m * w, dim=1)
        m = m - weighted_sum.unsqueeze(1)
        return weighted_sum, m.reshape(B, -1, H, W)

    def forward(self, y: Tensor, ctx_params: Tensor) -> Dict[str, Any]:
        gaussian_params = self.entropy_parameters(ctx_params)
        scales_hat, means_hat, weights = self._chunk(gaussian_params)
        weights = self._reshape_gmm_weight(weights)
        if self.quantizer == "noise":
            y_hat, y_likelihoods = self.gaussian_mixture_conditional(y, scales_hat, means_hat, weights)
        else:
            weighted_sum, means_hat = self._recentre(means_hat, weights)
            y_hat = quantize_ste(y - weighted_sum) + weighted_sum
            y_hat, y_likelihoods = self.gaussian_mixture_conditional(y_hat, scales_hat, means_hat, weights)
        return {"likelihoods": {"y": y_likelihoods}, "y_hat": y_hat}

    def compress(self, y: Tensor, ctx_params: Tensor) -> Dict[str, Any]:
        gaussian_params = self.entropy_parameters(ctx_params)
        scales_hat, means_hat, weights = self._chunk(gaussian_params)
        weights = self._reshape_gmm_weight(weights)
        if self.quantizer == "noise":
            y_strings, y_hat = self.gaussian_mixture_conditional.compress(y, scales_hat, means_hat, weights)
        else:
            weighted_sum, means_hat = self._recentre(means_hat, weights)
            y_hat = quantize_ste(y - weighted_sum)
            y_strings, y_hat = self.gaussian_mixture_conditional.compress(y_hat, scales_hat, means_hat, weights)
        return {"strings": [y_strings], "shape": y.shape[2:4], "y_hat": y_hat}

    def decompress(self, strings: List[List[bytes]], shape: Tuple[int, int], ctx_params: Tensor,
                   **kwargs) -> Dict[str, Any]:
        (y_strings,) = strings
        gaussian_params = self.entropy_parameters(ctx_params)
        scales_hat, means_hat, weights = self._chunk(gaussian_params)
        weights = self._reshape_gmm_weight(weights)
        if self.quantizer == "noise":
            y_hat = self.gaussian_mixture_conditional.decompress(*y_strings, scales_hat, means_hat, weights)
        else:
            weighted_sum, means_hat = self._recentre(means_hat, weights)
            y_hat = self.gaussian_mixture_conditional.decompress(*y_strings, scales_hat, means_hat, weights)
            y_hat = y_hat + weighted_sum
        assert y_hat.shape[2:4] == shape
        return {"y_hat": y_hat}

    # gaussian_mixture_conditional.py:183-196
    def _chunk(self, params: Tensor):
        scales, means = None, None
        if self.chunks == ("scales",):
            scales = params
        if self.chunks == ("means",):
            means = params
        if self.chunks == ("scales", "means"):
            scales, means = params.chunk(2, 1)
        if self.chunks == ("means", "scales"):
            means, scales = params.chunk(2, 1)
        if self.chunks == ("scales", "means", "weights"):
            scales, means, weights = params.chunk(3, 1)
            return scales, means, weights
        return scales, means

    # gaussian_mixture_conditional.py:198-202
    def _reshape_gmm_weight(self, weight):
        weight = torch.reshape(weight, (weight.size(0), self.K, weight.size(1) // self.K, weight.size(2), -1))
        weight = nn.functional.softmax(weight, dim=1)
        weight = torch.reshape(weight, (weight.size(0), weight.size(1) * weight.size(2), weight.size(3), -1))
        return weight
