"""Dev probe: the model path with several batches in flight (flashgmm_b200.Lane).  python tools/model_lanes_probe.py [B] [lanes ...]"""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 24
lanes = [int(a) for a in sys.argv[2:]] or [2, 4]
for n in lanes:
    code = ("import sys,json; sys.path.insert(0,%r); import torch, bench; "
            "print(json.dumps(bench.model_codec_bench(torch, torch.device('cuda',0), %d, reps=4, lanes=%d)))" % (str(ROOT), B, n))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    print("lanes", n, r.stdout.strip().splitlines()[-1] if r.returncode == 0 and r.stdout.strip() else ("FAILED " + r.stdout[-300:] + r.stderr[-1500:]), flush=True)
