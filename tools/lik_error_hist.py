"""Error histogram of the fused likelihood kernel (GPU box): relative error of (a) the kernel vs the reference's CPU fp32 goldens,
(b) the kernel vs a float64 evaluation of the same formula, (c) the goldens vs float64 -- i.e. which side carries how much of the
difference. python tools/lik_error_hist.py > profiles/r02_likelihood_error_hist.txt"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np, torch
import flashgmm_b200 as fg
from golden_inputs import LIK_CASES, lik_inputs
GOLD = np.load(ROOT / "tests/golden/model_small.npz")

def lik64(y, sc, mu, w, K, sb=0.11, lb=1e-9):
    y, sc, mu, w = (t.double() for t in (y, sc, mu, w))
    M = y.size(1); out = torch.zeros_like(y); c = -(2 ** -0.5)
    for k in range(K):
        s = torch.clamp_min(sc[:, M*k:M*(k+1)], float(np.float32(sb))); v = torch.abs(y - mu[:, M*k:M*(k+1)])
        out += (0.5*torch.erfc(c*((0.5-v)/s)) - 0.5*torch.erfc(c*((-0.5-v)/s))) * w[:, M*k:M*(k+1)]
    return out.clamp_min(float(np.float32(lb)))

def hist(name, a, b):
    rel = (np.abs(a.astype(np.float64) - b) / np.maximum(np.abs(b), 1e-300)).reshape(-1)
    edges = [0, 1e-8, 1e-7, 3e-7, 1e-6, 3e-6, 1e-5, 1e-4, 1e-3, np.inf]
    h, _ = np.histogram(rel, edges)
    print(f"  {name:34s} max {rel.max():.2e}  p99 {np.quantile(rel, .99):.2e}  counts per bin {edges[1:]}: {h.tolist()}")

for name, (B, M, H, W, K, seed) in sorted(LIK_CASES.items()):
    x, sc, mu, w = lik_inputs(B, M, H, W, K, seed)
    m = fg.GaussianMixtureConditional(K=K).cuda().eval()
    y, lik = m(*(t.cuda() for t in (x, sc, mu, w)))
    truth = lik64(torch.round(x), sc, mu, w, K).numpy()
    print(f"case {name}: {lik.numel()} elements, likelihood range [{truth.min():.2e}, {truth.max():.2e}]")
    hist("kernel vs reference golden (fp32)", lik.cpu().numpy(), GOLD[f"lik/{name}/likelihood"].astype(np.float64))
    hist("kernel vs float64", lik.cpu().numpy(), truth)
    hist("reference golden vs float64", GOLD[f"lik/{name}/likelihood"], truth)
# a wide sweep incl. large scales, where the fp32 difference of two Phi values cancels (both implementations alike)
g = torch.Generator().manual_seed(99)
B, M, H, W, K = 4, 32, 16, 16, 4
x = torch.randn(B, M, H, W, generator=g) * 6
sc = torch.exp(torch.rand(B, K*M, H, W, generator=g) * 7.7 - 2.2)          # 0.11 .. 256
mu = torch.randn(B, K*M, H, W, generator=g) * 3
w = torch.softmax(torch.randn(B, K, M, H, W, generator=g), 1).reshape(B, K*M, H, W)
m = fg.GaussianMixtureConditional(K=K).cuda().eval()
y, lik = m(*(t.cuda() for t in (x, sc, mu, w)))
cpu = fg.GaussianMixtureConditional(K=K)
M_ = M; c = -(2 ** -0.5); ref32 = torch.zeros_like(x); xr = torch.round(x)
for k in range(K):
    s = torch.clamp_min(sc[:, M_*k:M_*(k+1)], 0.11); v = torch.abs(xr - mu[:, M_*k:M_*(k+1)])
    ref32 += (0.5*torch.erfc(c*((0.5-v)/s)) - 0.5*torch.erfc(c*((-0.5-v)/s))) * w[:, M_*k:M_*(k+1)]
ref32 = ref32.clamp_min(1e-9)
truth = lik64(xr, sc, mu, w, K).numpy()
print(f"wide sweep (scales log-uniform in [0.11, 256]): {lik.numel()} elements")
hist("kernel vs torch CPU fp32 (reference ops)", lik.cpu().numpy(), ref32.numpy().astype(np.float64))
hist("kernel vs float64", lik.cpu().numpy(), truth)
hist("torch CPU fp32 vs float64", ref32.numpy(), truth)
