"""Launches every kernel the bench line quotes, once or twice at a representative size, for `ncu --set full` (GPU box):
  startfreq (16 Mi symbols, modes 0/1/2), windowed rows + pop (one pass of 24 x 147456, modes 0/1/2), push (device policy),
  likelihood fwd/bwd (16x192x16x16 and 64x320x32x48), interleaved encode/decode (4 Mi symbols), prologue glue, z-branch records.
Plain run first (must exit 0), then under ncu with -k regex filters (tools/gpu_profile.sh)."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch
import flashgmm_b200 as fg
from flashgmm_b200 import _lib, ans, batch, ileave, synth

dev = torch.device("cuda", 0)
L = _lib.lib()
h = _lib.ctx(0)
which = set((sys.argv[1] if len(sys.argv) > 1 else "sf,dec,push,lik,il,glue").split(","))
if "sf" in which:
    n = 1 << 24
    sym, sg, mu, w = synth.gmm_torch(n, seed=1234, device=dev)
    out = torch.empty(n, dtype=torch.int32, device=dev)
    for mode in (0, 1, 2):
        _lib.check(L.flashgmm_gmm_startfreq(h, sym.data_ptr(), sg.data_ptr(), mu.data_ptr(), w.data_ptr(), n, sg.stride(0),
                                            sg.stride(1), mode, out.data_ptr(), None))
    del sym, sg, mu, w, out
if "dec" in which or "push" in which:
    streams = [synth.gmm_torch(147456, seed=900 + i, device=dev) for i in range(24)]
    R = max(int(s[0].abs().max().item()) for s in streams) + 2
    torch.cuda.synchronize()
    if "push" in which:
        os.environ["FLASHGMM_PUSH"] = "device"
        dctx = _lib.new_ctx(0)
        del os.environ["FLASHGMM_PUSH"]
        batch.EncodeJob(streams, 0, ctx=dctx).run()
    if "dec" in which:
        for mode in (0, 1, 2):
            bs = [x.copy() for x in batch.EncodeJob(streams, mode).run()]
            outs = batch.DecodeJob([(a, b, c, R) for (_, a, b, c) in streams], mode).run(bs)
            assert all(torch.equal(o, s[0]) for o, s in zip(outs, streams))
    del streams
if "lik" in which:
    for (B, M, H, W) in ((16, 192, 16, 16), (64, 320, 32, 48)):
        g = torch.Generator(device=dev).manual_seed(7)
        x = (torch.randn(B, M, H, W, device=dev, generator=g) * 3).requires_grad_(True)
        sc = (torch.rand(B, 4 * M, H, W, device=dev, generator=g) * 3 + 0.05).requires_grad_(True)
        mu = torch.randn(B, 4 * M, H, W, device=dev, generator=g).requires_grad_(True)
        w = torch.softmax(torch.randn(B, 4, M, H, W, device=dev, generator=g), 1).reshape(B, 4 * M, H, W).requires_grad_(True)
        m = fg.GaussianMixtureConditional(K=4).to(dev).train()
        y, lik = m(x, sc, mu, w)
        lik.log().sum().backward()
        torch.cuda.synchronize()
if "il" in which:
    n = 1 << 22
    sym, sg, mu, w = synth.gmm_torch(n, seed=77, device=dev)
    R = int(sym.abs().max().item()) + 2
    blob = ileave.encode(sym, sg, mu, w, R, 0)
    assert torch.equal(ileave.decode(blob, sg, mu, w), sym)
if "glue" in which:
    g = torch.Generator().manual_seed(3)
    B, Cc, H, W = 24, 192, 32, 24
    y = (torch.randn(B, Cc, H, W, generator=g) * 3).cuda()
    sc = (torch.rand(B, 4 * Cc, H, W, generator=g) * 3 + 0.05).cuda()
    mu = torch.randn(B, 4 * Cc, H, W, generator=g).cuda()
    wt = torch.softmax(torch.randn(B, 4, Cc, H, W, generator=g), 1).reshape(B, 4 * Cc, H, W).cuda()
    gmc = fg.GaussianMixtureConditional(K=4).cuda()
    items, yq = gmc.compress_batch(y, sc, mu, wt)
    assert torch.equal(gmc.decompress_batch(items, sc, mu, wt), yq)
    eb = fg.EntropyBottleneck(192).cuda()
    eb.update()
    z = (torch.randn(24, 192, 8, 12, generator=g) * 3).cuda()
    s = eb.compress(z)
    eb.decompress(s, (8, 12))
torch.cuda.synchronize()
print("targets ok")
