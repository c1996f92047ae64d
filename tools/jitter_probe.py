"""Call-time distribution of batch encode / decode (device-resident inputs): median, p95, max."""
import sys, json, time, faulthandler
faulthandler.dump_traceback_later(120, exit=True)
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from flashgmm_b200 import _lib, synth
from flashgmm_b200.batch import DecodeJob, EncodeJob
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 24
N = int(sys.argv[2]) if len(sys.argv) > 2 else 100
dev = torch.device("cuda", 0)
imgs = [bench.make_image(torch, synth, i, dev) for i in range(B)]
flat = [st for img in imgs for st in img]
radii = [int(st[0].abs().max().item()) + 2 for st in flat]
enc = EncodeJob(flat, 0); dec = DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(flat, radii)], 0)
bs = [x.copy() for x in enc.run()]
for _ in range(3):
    enc.run(); dec.run(bs)
res = {"B": B, "calls": N}
for name, fn in (("encode", lambda: enc.run()), ("decode", lambda: dec.run(bs))):
    ts = []
    for _ in range(N):
        torch.cuda.synchronize(); t0 = time.perf_counter(); fn(); ts.append((time.perf_counter() - t0) * 1e3)
    a = np.array(ts)
    res[name] = {"median": round(float(np.median(a)), 3), "p95": round(float(np.percentile(a, 95)), 3),
                 "max": round(float(a.max()), 3), "mean": round(float(a.mean()), 3),
                 "over_1.5x_median": int((a > 1.5 * np.median(a)).sum())}
print(json.dumps(res))
