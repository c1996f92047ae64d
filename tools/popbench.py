import sys, json, ctypes as C, time, faulthandler
faulthandler.dump_traceback_later(50, exit=True)
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from flashgmm_b200 import _lib, synth
from flashgmm_b200.batch import DecodeJob, EncodeJob
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda", 0)
imgs = [bench.make_image(torch, synth, i, dev) for i in range(B)]
flat = [st for img in imgs for st in img]
radii = [int(st[0].abs().max().item()) + 2 for st in flat]
enc = EncodeJob(flat, 0); dec = DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(flat, radii)], 0)
bs = [x.copy() for x in enc.run()]
L = _lib.lib()
for _ in range(3):
    outs = dec.run(bs)
for o, st in zip(outs, flat):
    assert torch.equal(o, st[0])
L.flashgmm_profile_reset(); L.flashgmm_profile_enable(1)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5):
    dec.run(bs)
torch.cuda.synchronize(); t1 = time.perf_counter()
L.flashgmm_profile_enable(0)
r = {}
for kind, name in ((1, "rows"), (4, "pop")):
    ms, n, by = C.c_double(), C.c_uint64(), C.c_double()
    L.flashgmm_profile_read(kind, C.byref(ms), C.byref(n), C.byref(by))
    r[name] = round(ms.value / 5, 3)
nsym = sum(st[0].numel() for st in flat)
enc.run(); enc.run()
L.flashgmm_profile_reset(); L.flashgmm_profile_enable(1)
torch.cuda.synchronize(); t2 = time.perf_counter()
for _ in range(5):
    enc.run()
torch.cuda.synchronize(); t3 = time.perf_counter()
L.flashgmm_profile_enable(0)
for kind, name in ((0, "startfreq"), (5, "push")):
    ms, n, by = C.c_double(), C.c_uint64(), C.c_double()
    L.flashgmm_profile_read(kind, C.byref(ms), C.byref(n), C.byref(by))
    r[name] = round(ms.value / 5, 3)
r["encode_ms"] = round((t3 - t2) / 5 * 1e3, 3)
print(json.dumps({"B": B, "decode_ms": round((t1 - t0) / 5 * 1e3, 3), "msym_s": round(nsym / ((t1 - t0) / 5) / 1e6, 1), "kernel_ms_per_call": r}))
