"""Generate the golden fixtures from the reference's own compiled extension.

Run in the build container (needs oracle/_ref, built by `python oracle/build.py`
from /root/reference):

    python tests/golden/make_golden.py

Writes tests/golden/gmm_small.npz (full inputs + reference outputs, n=512) and
tests/golden/gmm_md5.json (md5 of reference bitstreams / decoded symbols for larger
seeded cases that tests regenerate with flashgmm_b200.synth).
"""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from flashgmm_b200 import synth  # noqa: E402
from oracle import ref  # noqa: E402

HERE = Path(__file__).resolve().parent

# name -> synth kwargs (+ post-processing tag)
CASES = {
    "default_4k": dict(n=4096, seed=1),
    "default_100k": dict(n=100_000, seed=2),
    "narrow_sigma_20k": dict(n=20_000, seed=3, sigma_lo=0.11, sigma_span=0.0),
    "wide_means_20k": dict(n=20_000, seed=4, mean_scale=40.0),
    "offcentre_bypass_20k": dict(n=20_000, seed=5, sigma_span=0.05, offcentre=9.0),
    "ragged_4099": dict(n=4099, seed=6),
    "single": dict(n=1, seed=7),
    "empty": dict(n=0, seed=8),
}


def small_case():
    sym, sg, mu, w = synth.gmm_numpy(512, seed=99)
    sym = sym.copy()
    sym[:8] = [0, -1, 1, 300, -300, 2**31 - 1, -2**31, 7]        # far outliers -> bypass nibbles 1..8
    sg = np.ascontiguousarray(sg).copy()
    sg[8:16] = 0.11
    sg[16:24] = 256.0
    return sym, sg, np.ascontiguousarray(mu), np.ascontiguousarray(w)


def main():
    md5 = {}
    small = {}
    sym, sg, mu, w = small_case()
    small.update(symbols=sym, scales=sg, means=mu, weights=w)
    R_small = 400
    for mode in (0, 1, 2):
        with ref.RefWorker(mode) as wk:
            b, _ = wk.gmm_encode(sym, sg, mu, w, R_small)
            d, _ = wk.gmm_decode(b, sg, mu, w, R_small)
            small[f"bytes_mode{mode}"] = np.frombuffer(b, np.uint8)
            small[f"decoded_mode{mode}"] = d
            for name, kw in CASES.items():
                s, a, m, ww = synth.gmm_numpy(**kw)
                R = synth.search_radius(s)
                b, _ = wk.gmm_encode(s, a, m, ww, R)
                d, _ = wk.gmm_decode(b, a, m, ww, R)
                md5[f"{name}/mode{mode}"] = dict(
                    synth=kw, R=R, nbytes=len(b), bytes_md5=hashlib.md5(b).hexdigest(),
                    decoded_md5=hashlib.md5(np.ascontiguousarray(d, np.int32).tobytes()).hexdigest(),
                    roundtrip=bool(np.array_equal(d, s)))
    small["R"] = np.int32(R_small)
    np.savez_compressed(HERE / "gmm_small.npz", **small)
    (HERE / "gmm_md5.json").write_text(json.dumps(md5, indent=1, sort_keys=True))
    print("wrote", len(md5), "md5 cases;", {k: v["nbytes"] for k, v in md5.items() if k.endswith("mode0")})
    print("roundtrip flags:", {k: v["roundtrip"] for k, v in md5.items()})


if __name__ == "__main__":
    main()
