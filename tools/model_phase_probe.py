"""Dev probe: where the time of the MODEL path (latent_codec.compress / decompress of ELIC, B images) goes, and how it
moves with the number of replay lanes.   python tools/model_phase_probe.py [B]

Phases are separated by torch.cuda.synchronize(), so their sum exceeds the un-instrumented call (no overlap between
phases); the lanes sweep runs un-instrumented in child processes (FLASHGMM_NET_LANES is read when the runner is built)."""
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def phases(B):
    import torch
    import bench
    import flashgmm_b200.latent_codecs as LC
    import flashgmm_b200.entropy_models as EM
    acc = {}
    state = {"side": "?", "calls": 0}

    def timed(name, fn):
        def wrapper(*a, **k):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = fn(*a, **k)
            torch.cuda.synchronize()
            key = state["side"] + " " + name
            acc[key] = acc.get(key, 0.0) + (time.perf_counter() - t0) * 1e3
            return out
        return wrapper

    def top(side, fn):
        def wrapper(*a, **k):
            if side == "C":
                state["calls"] += 1
                if state["calls"] == 2:                       # the first round trip captured the graphs: start over
                    acc.clear()
            state["side"] = side
            return fn(*a, **k)
        return wrapper

    LC.HyperpriorLatentCodec.compress = top("C", LC.HyperpriorLatentCodec.compress)
    LC.HyperpriorLatentCodec.decompress = top("D", LC.HyperpriorLatentCodec.decompress)

    LC.HyperLatentCodec.compress = timed("hyper.compress", LC.HyperLatentCodec.compress)
    LC.HyperLatentCodec.decompress = timed("hyper.decompress", LC.HyperLatentCodec.decompress)
    EM.EntropyBottleneck.compress = timed("  eb.compress", EM.EntropyBottleneck.compress)
    EM.EntropyBottleneck.decompress = timed("  eb.decompress", EM.EntropyBottleneck.decompress)
    EM.EncodeSession.flush = timed("session.flush (one encode_batch)", EM.EncodeSession.flush)
    EM.GaussianMixtureConditional.compress_batch = timed("gmc.compress_batch (scan+gather, register)", EM.GaussianMixtureConditional.compress_batch)
    EM.GaussianMixtureConditional.decompress_batch = timed("gmc.decompress_batch (gather+decode+scatter)", EM.GaussianMixtureConditional.decompress_batch)
    LC.CheckerboardLatentCodec._pass_params = timed("ckbd nets (_pass_params)", LC.CheckerboardLatentCodec._pass_params)
    LC.ChannelGroupsLatentCodec._ctx_params_graphed = timed("channel ctx nets", LC.ChannelGroupsLatentCodec._ctx_params_graphed)
    reps = 3
    out = bench.model_codec_bench(torch, torch.device("cuda", 0), B, reps=reps)
    n_calls = reps
    print(json.dumps({"B": B, "instrumented_total": out, "per_call_ms": {k: round(v / n_calls, 3) for k, v in acc.items()}}, indent=1))


if __name__ == "__main__":
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    if os.environ.get("_PROBE_CHILD") == "phases":
        phases(B)
        sys.exit(0)
    code = ("import sys,json; sys.path.insert(0,%r); import torch, bench; "
            "d = bench.model_codec_bench(torch, torch.device('cuda',0), %d); "
            "print(json.dumps({k: d[k] for k in ('compress_ms','decompress_ms','graph_replays')}))" % (str(ROOT), B))
    for lanes in ((1, 2, 4, 8) if os.environ.get("PROBE_LANES") else ()):
        e = dict(os.environ, FLASHGMM_NET_LANES=str(lanes))
        r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=600)
        print("lanes", lanes, r.stdout.strip().splitlines()[-1] if r.returncode == 0 else ("FAILED " + r.stderr[-800:]), flush=True)
    e = dict(os.environ, _PROBE_CHILD="phases")
    r = subprocess.run([sys.executable, __file__, str(B)], env=e, capture_output=True, text=True, timeout=600)
    print(r.stdout if r.returncode == 0 else ("FAILED " + r.stderr[-1500:]))
