"""Where the end-to-end (host-buffer) step spends its time: PCIe copy bandwidth, then encode and
decode through the public batch API with pinned host inputs, timed separately."""
import os, sys, json, time, faulthandler
faulthandler.dump_traceback_later(150, exit=True)
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from flashgmm_b200 import _lib, synth
from flashgmm_b200.batch import DecodeJob, EncodeJob
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda", 0)
res = {"B": B}
h = torch.empty(256 << 20, dtype=torch.uint8, pin_memory=True); d = torch.empty_like(h, device=dev)
for name, (a, b) in {"h2d": (d, h), "d2h": (h, d)}.items():
    a.copy_(b, non_blocking=True); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        a.copy_(b, non_blocking=True)
    torch.cuda.synchronize()
    res[name + "_gbs"] = round(4 * h.numel() / (time.perf_counter() - t0) / 1e9, 1)
imgs = [bench.make_image(torch, synth, i, dev) for i in range(B)]
flat = [st for img in imgs for st in img]
radii = [int(st[0].abs().max().item()) + 2 for st in flat]
hflat = [st for img in imgs for st in bench.to_pinned(torch, img)]
nsym = sum(st[0].numel() for st in flat)
for tag, streams in (("dev", flat), ("host", hflat)):
    enc = EncodeJob(streams, 0); dec = DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(streams, radii)], 0)
    bs = [x.copy() for x in enc.run()]
    dec.run(bs)
    if os.environ.get("FLASHGMM_TRACE"):
        L = _lib.lib(); L.flashgmm_profile_enable(1); print("----", tag, file=sys.stderr); dec.run(bs); L.flashgmm_profile_enable(0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5):
        enc.run()
    t1 = time.perf_counter()
    for _ in range(5):
        dec.run(bs)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    res[tag] = {"enc_ms": round((t1 - t0) / 5e-3, 2), "dec_ms": round((t2 - t1) / 5e-3, 2),
                "enc_msym_s": round(nsym / ((t1 - t0) / 5) / 1e6), "dec_msym_s": round(nsym / ((t2 - t1) / 5) / 1e6)}
res["h2d_floor_ms"] = {"enc": round(nsym * 52 / res["h2d_gbs"] / 1e6, 2), "dec": round(nsym * 48 / res["h2d_gbs"] / 1e6, 2)}
print(json.dumps(res))
