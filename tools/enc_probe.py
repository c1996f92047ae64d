"""Batch encode timing (device-resident inputs) with kernel profiling off and on."""
import sys, json, time, faulthandler
faulthandler.dump_traceback_later(80, exit=True)
sys.path.insert(0, "/root/repo")
import torch
from flashgmm_b200 import _lib, synth
from flashgmm_b200.batch import EncodeJob
import bench
B = int(sys.argv[1]) if len(sys.argv) > 1 else 24
dev = torch.device("cuda", 0)
imgs = [bench.make_image(torch, synth, i, dev) for i in range(B)]
flat = [st for img in imgs for st in img]
enc = EncodeJob(flat, 0)
L = _lib.lib()
res = {"B": B}
for _ in range(3):
    enc.run()
for tag, on in (("prof_off", 0), ("prof_on", 1), ("prof_off2", 0)):
    L.flashgmm_profile_enable(on)
    enc.run()
    ts = []
    for _ in range(8):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        enc.run()
        ts.append(round((time.perf_counter() - t0) * 1e3, 2))
    res[tag] = ts
L.flashgmm_profile_enable(0)
print(json.dumps(res))
