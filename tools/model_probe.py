import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch
from flashgmm_b200 import ans
import gmm_models as gm
def run(cls, kw, modes):
    torch.manual_seed(0)
    net = cls(**kw).eval(); net.update(); net = net.cuda()
    g = torch.Generator().manual_seed(0)
    x = torch.rand(1, 3, 512, 768, generator=g).cuda()
    with torch.no_grad():
        y = 20.0 * net.g_a(x)
    for m in modes:
        ans.set_approx_mode(m)
        enc = net.latent_codec.compress(y)
        dec = net.latent_codec.decompress(enc["strings"], enc["shape"])
        print(cls.__name__, "mode", m, torch.equal(dec["y_hat"], enc["y_hat"]), flush=True)
run(gm.Cheng2020AnchorCheckerboardGMMv2, dict(N=192, K=4), [0, 1, 2])
run(gm.Elic2022GMM, dict(N=192, M=320, K=4), [0, 1, 2])
