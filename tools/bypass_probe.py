"""Dev probe (GPU box): bypass-heavy streams (sigma = 0.11, means far from the symbols: what random-init models
give) through every encode / decode path, against the C oracle. Usage: python tools/bypass_probe.py"""
import os, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

CHILD = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
from flashgmm_b200 import ans, batch, synth
from oracle import port
def make(n, seed):
    sym, sg, mu, w = synth.gmm_numpy(n, seed=seed, sigma_lo=0.11, sigma_span=0.0, mean_scale=40.0)
    return sym, sg, mu, w
for sizes in ([147456, 147456], [147456], [12288]*4 + [49152]*2, [5000]*30):
    data = [make(n, 10+i) for i, n in enumerate(sizes)]
    dev = [(torch.from_numpy(s).cuda(),) + tuple(torch.from_numpy(np.ascontiguousarray(x.T)).cuda().t() for x in (a, b, c)) for s, a, b, c in data]
    want = [port.gmm_encode(s, a, b, c, 0) for s, a, b, c in data]
    got = batch.encode_batch(dev, 0)
    ok = [bytes(g) == w for g, w in zip(got, want)]
    first = [next((i for i, (x, y) in enumerate(zip(bytes(g), w)) if x != y), -1) for g, w in zip(got, want)]
    print("encode_batch", sizes[:3], len(sizes), "ok" if all(ok) else ("MISMATCH lens %%s first diff %%s" %% ([(len(g), len(w)) for g, w in zip(got, want)][:4], first[:4])), flush=True)
    if len(sizes) == 1:
        g1 = ans.gmm_encode(*dev[0], 0)
        print("  single gmm_encode", g1 == want[0], flush=True)
    R = [synth.search_radius(d[0]) for d in data]
    outs = batch.decode_batch([(w_, d[1], d[2], d[3], r) for w_, d, r in zip(want, dev, R)], 0)
    print("  decode_batch", all(np.array_equal(o.cpu().numpy(), d[0]) for o, d in zip(outs, data)), flush=True)
    o1 = ans.gmm_decode(want[0], dev[0][1], dev[0][2], dev[0][3], R[0], 0)
    print("  single gmm_decode", np.array_equal(o1.cpu().numpy(), data[0][0]), flush=True)
''' % str(ROOT)

for env in ({}, {"FLASHGMM_PUSH": "host"}, {"FLASHGMM_PUSH": "device"}, {"FLASHGMM_ENC_WAVES": "0", "FLASHGMM_PUSH": "host"}):
    print("==== env", env, flush=True)
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-X", "faulthandler", "-c", CHILD], env=e, capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-1500:] if r.returncode else "", flush=True)
