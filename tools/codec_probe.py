"""Dev probe: model_codec block under both net batching modes and with graphs off; python tools/codec_probe.py"""
import os, subprocess, sys, json
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
code = "import sys,json; sys.path.insert(0,%r); import torch, bench; print(json.dumps(bench.model_codec_bench(torch, torch.device('cuda',0), 24)))" % str(ROOT)
for env in ({}, {"FLASHGMM_NET_BATCHING": "batched"}, {"FLASHGMM_GRAPHS": "0"}, {"FLASHGMM_GRAPHS": "0", "FLASHGMM_NET_BATCHING": "batched"}):
    e = dict(os.environ); e.update(env)
    r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=600)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print(env, {k: d[k] for k in ("compress_ms", "decompress_ms", "graph_captures", "graph_replays", "net_batching")}, flush=True)
    except Exception:
        print(env, "FAILED", r.stdout[-500:], r.stderr[-1500:], flush=True)
