"""Seeded synthetic GMM coding workloads (BASELINE.md section 3 / SURVEY.md section 8d).

``means = 3*randn``, ``scales = clamp(0.11 + 4*rand, 0.11, 256)``, ``weights =
softmax(randn)``, symbols drawn from the mixture and rounded: about 3.5 bit/symbol,
almost no bypass. Parameters are produced K-major, ``[K][N]`` planes, and handed out
as ``(N, K)`` views with strides ``(1, N)`` -- what the reference's
``reshape_entropy_parameters`` gives its coder (entropy_models.py:810-828).
"""
from __future__ import annotations

import numpy as np


def gmm_numpy(n: int, seed: int = 1234, K: int = 4, sigma_lo: float = 0.11, sigma_span: float = 4.0,
              mean_scale: float = 3.0, offcentre: float = 0.0):
    """-> sym int32[N], scales/means/weights float32 (N,K) views with strides (1,N)."""
    r = np.random.default_rng(seed)
    mu = (mean_scale * r.standard_normal((K, n))).astype(np.float32)
    sg = np.clip(sigma_lo + sigma_span * r.random((K, n)), 0.11, 256).astype(np.float32)
    lg = r.standard_normal((K, n)).astype(np.float32)
    w = np.exp(lg - lg.max(0, keepdims=True))
    w = (w / w.sum(0, keepdims=True)).astype(np.float32)
    c = (r.random(n)[None, :] > np.cumsum(w, 0)).sum(0).clip(0, K - 1)
    ar = np.arange(n)
    sym = np.round(mu[c, ar] + sg[c, ar] * r.standard_normal(n) + offcentre).astype(np.int32)
    return sym, sg.T, mu.T, w.T


def search_radius(sym) -> int:
    """max|sym| + 2, the `abs_max + 1` the reference's caller passes (entropy_models.py:834-865)."""
    if len(sym) == 0:
        return 2
    return int(np.abs(np.asarray(sym, np.int64)).max()) + 2


def gmm_torch(n: int, seed: int = 1234, device="cuda", K: int = 4, sigma_lo: float = 0.11,
              sigma_span: float = 4.0, mean_scale: float = 3.0):
    """Same distribution generated on `device` with torch (for 16 Mi-symbol benches)."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)
    mu = mean_scale * torch.randn(K, n, device=device, generator=g)
    sg = (sigma_lo + sigma_span * torch.rand(K, n, device=device, generator=g)).clamp_(0.11, 256)
    w = torch.softmax(torch.randn(K, n, device=device, generator=g), dim=0)
    u = torch.rand(n, device=device, generator=g)
    c = (u[None, :] > torch.cumsum(w, 0)).sum(0).clamp_(0, K - 1)
    ar = torch.arange(n, device=device)
    sym = torch.round(mu[c, ar] + sg[c, ar] * torch.randn(n, device=device, generator=g)).to(torch.int32)
    return sym, sg.t(), mu.t(), w.t()
