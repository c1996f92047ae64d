"""Seeded inputs shared by tests/golden/make_model_golden.py (which feeds them to the
reference's Python modules) and the tests (which feed them to flashgmm_b200). CPU
torch generators only, so the values are the same on every machine."""
import torch

LIK_CASES = {
    # name: (B, M, H, W, K, seed)
    "vec_k4": (2, 8, 4, 4, 4, 11),
    "odd_k4": (2, 6, 5, 7, 4, 12),
    "k3": (1, 5, 3, 4, 3, 13),
    "k1": (1, 4, 4, 4, 1, 14),
}
CODEC_CASES = {
    # name: (C, h, w, quantizer, seed, gain)
    "noise_12x6x8": (12, 6, 8, "noise", 21, 4.0),
    "ste_10x5x6": (10, 5, 6, "weighted_mean_ste", 22, 3.0),
    "allzero": (6, 4, 4, "noise", 23, 0.0),
    "wide_16x8x8": (16, 8, 8, "noise", 24, 40.0),
}


def lik_inputs(B, M, H, W, K, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, M, H, W, generator=g) * 3
    scales = torch.rand(B, K * M, H, W, generator=g) * 3
    scales[:, ::3] *= 0.03                                  # below the 0.11 bound
    means = torch.randn(B, K * M, H, W, generator=g) * 2
    w = torch.softmax(torch.randn(B, K, M, H, W, generator=g), 1).reshape(B, K * M, H, W)
    x.view(-1)[:5] = torch.tensor([60.0, -75.0, 0.0, 0.5, -0.5])   # likelihood bound / ties
    return x, scales, means, w


def codec_inputs(C, h, w, seed, gain, K=4):
    g = torch.Generator().manual_seed(seed)
    y = torch.randn(1, C, h, w, generator=g) * gain
    if gain > 0:
        y[:, 1] = 0.3 * torch.rand(h, w, generator=g)       # rounds to an all-zero channel
        y[:, C - 2] = 0.0
        y.view(-1)[7] = 900.7                               # forces bypass + large abs_max
    params = torch.randn(1, 3 * K * C, h, w, generator=g)
    params[:, : K * C] = params[:, : K * C].abs() * 2       # scales: some below 0.11, none negative
    params[:, K * C: 2 * K * C] *= gain if gain > 0 else 1.0
    return y, params
