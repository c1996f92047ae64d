import sys, os, time
sys.path.insert(0, "/root/repo")
import torch
from flashgmm_b200 import _lib, synth
from flashgmm_b200.batch import EncodeJob
import bench
B = int(sys.argv[1])
dev = torch.device("cuda", 0)
imgs = [bench.make_image(torch, synth, i, dev) for i in range(B)]
flat = [st for img in imgs for st in img]
enc = EncodeJob(flat, 0)
for _ in range(4):
    t0 = time.perf_counter(); enc.run(); print("call ms", round((time.perf_counter() - t0) * 1e3, 2), file=sys.stderr)
