"""Golden vectors from the reference's own PYTHON entropy model and latent codec.

Run in the build container (needs /root/reference and oracle/_ref):

    python tests/golden/make_model_golden.py

Imports ``GaussianMixtureConditional`` (compressai/entropy_models/entropy_models.py:762-910)
and ``GaussianMixtureConditionalLatentCodec``
(compressai/latent_codecs/gaussian_mixture_conditional.py:43-202) from the reference
tree unmodified (oracle/ref_py.py) and records, on CPU fp32:

* likelihood cases: ``forward(..., training=False)`` outputs and the gradients of
  ``sum(log(likelihood))`` through ``_likelihood`` + ``LowerBound`` w.r.t. inputs,
  scales, means and weights;
* codec cases: ``compress`` -> (bytes, abs_max, zero_bitmap, y_hat) and ``decompress``
  -> y_hat for both quantisers, one file section per APPROX_MODE (the reference caches
  the mode in a static, so each mode runs in its own subprocess).

Writes tests/golden/model_small.npz.
"""
import io
import os
import subprocess
import sys
from contextlib import redirect_stdout
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
HERE = Path(__file__).resolve().parent

sys.path.insert(0, str(ROOT / "tests"))
from golden_inputs import CODEC_CASES, LIK_CASES, codec_inputs, lik_inputs  # noqa: E402


def run_mode(mode: int):
    """Executed in a subprocess with APPROX_MODE set: returns a dict of arrays."""
    import torch

    from oracle import ref_py

    em, lc = ref_py.load()
    torch.set_num_threads(1)
    out = {}
    if mode == 0:
        for name, (B, M, H, W, K, seed) in LIK_CASES.items():
            x, sc, mu, w = lik_inputs(B, M, H, W, K, seed)
            m = em.GaussianMixtureConditional(K=K).eval()
            y, lik = m(x, sc, mu, w)
            xs = [t.clone().requires_grad_(True) for t in (y, sc, mu, w)]
            l2 = m.likelihood_lower_bound(m._likelihood(*xs))
            torch.log(l2).sum().backward()
            out[f"lik/{name}/y"] = y.numpy()
            out[f"lik/{name}/likelihood"] = lik.detach().numpy()
            for t, nm in zip(xs, ("gy", "gscales", "gmeans", "gweights")):
                out[f"lik/{name}/{nm}"] = t.grad.numpy()
    for name, (C, h, w, quant, seed, gain) in CODEC_CASES.items():
        y, params = codec_inputs(C, h, w, seed, gain)
        codec = lc.GaussianMixtureConditionalLatentCodec(K=4, quantizer=quant)
        seen = {}
        inner = codec.gaussian_mixture_conditional.compress

        def spy(y_, sc_, mu_, w_, _inner=inner, _seen=seen):
            # the exact tensors the entropy model receives (CPU softmax / recentring differ from the
            # GPU's in the last ulp, so entropy-model-level parity is checked on these)
            _seen.update(y=y_.clone(), scales=sc_.clone(), means=mu_.clone(), weights=w_.clone())
            return _inner(y_, sc_, mu_, w_)

        codec.gaussian_mixture_conditional.compress = spy
        with redirect_stdout(io.StringIO()):
            enc = codec.compress(y, params)
            (bts, abs_max, bitmap) = enc["strings"][0]
            dec = codec.decompress(enc["strings"], enc["shape"], params)
        key = f"codec/{name}/mode{mode}"
        out[f"{key}/bytes"] = np.frombuffer(bts, np.uint8)
        out[f"{key}/abs_max"] = np.int64(abs_max)
        out[f"{key}/zero_bitmap"] = bitmap.numpy()
        out[f"{key}/y_hat_enc"] = enc["y_hat"].numpy()
        out[f"{key}/y_hat_dec"] = dec["y_hat"].numpy()
        if mode == 0:
            for nm, t in seen.items():
                out[f"codec/{name}/em_{nm}"] = t.detach().numpy()
    return out


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--mode":
        mode = int(sys.argv[2])
        np.savez(sys.argv[3], **run_mode(mode))
        return
    merged = {}
    for mode in (0, 1, 2):
        tmp = HERE / f"_tmp_mode{mode}.npz"
        env = dict(os.environ, APPROX_MODE=str(mode), USE_SIMD="1", CUDA_VISIBLE_DEVICES="")
        subprocess.run([sys.executable, __file__, "--mode", str(mode), str(tmp)], check=True, env=env, cwd=str(ROOT))
        with np.load(tmp) as z:
            merged.update({k: z[k] for k in z.files})
        tmp.unlink()
    np.savez_compressed(HERE / "model_small.npz", **merged)
    print(f"wrote {HERE / 'model_small.npz'}: {len(merged)} arrays")


if __name__ == "__main__":
    main()
