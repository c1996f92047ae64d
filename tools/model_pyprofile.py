"""Dev probe: cProfile of the model path's host side (latent_codec.compress / decompress, ELIC, B images).
python tools/model_pyprofile.py [B]"""
import cProfile
import io
import pstats
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import bench  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 24
state = {}
orig = bench.time.perf_counter


class Grab(Exception):
    pass


# reuse model_codec_bench's set-up: intercept the codec after its warm-up round trip
import flashgmm_b200.latent_codecs as LC  # noqa: E402

calls = {"n": 0}
real_compress = LC.HyperpriorLatentCodec.compress


def spy(self, y):
    calls["n"] += 1
    if calls["n"] == 2:                                       # the first call was the warm-up
        state["lc"], state["y"] = self, y
        raise Grab()
    return real_compress(self, y)


LC.HyperpriorLatentCodec.compress = spy
try:
    bench.model_codec_bench(torch, torch.device("cuda", 0), B, reps=1)
except Grab:
    pass
LC.HyperpriorLatentCodec.compress = real_compress
lc, y = state["lc"], state["y"]
enc = lc.compress(y)
torch.cuda.synchronize()
for name, fn in (("compress", lambda: lc.compress(y)), ("decompress", lambda: lc.decompress(enc["strings"], enc["shape"]))):
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    pr.disable()
    out = io.StringIO()
    pstats.Stats(pr, stream=out).sort_stats("tottime").print_stats(28)
    print("=" * 30, name, "(3 calls)")
    print("\n".join(line[:170] for line in out.getvalue().splitlines()[:48]))
