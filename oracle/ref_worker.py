"""Worker process that runs the reference's compiled extension (TEST INFRASTRUCTURE).

Spawned by :class:`oracle.ref.RefWorker` with APPROX_MODE / USE_SIMD in the
environment. Calls the reference's public pybind API exactly as
``GaussianMixtureConditional.compress/decompress`` does
(entropy_models.py:859-865, 883-889). Timed ops return (result, seconds-per-call).
"""
from __future__ import annotations

import pickle
import struct
import sys
import time


def main():
    import numpy as np
    import torch

    from oracle import ref

    torch.set_num_threads(1)
    ans = ref.load_ans()
    enc = ans.RansEncoder()
    dec = ans.RansDecoder()
    stdin, stdout = sys.stdin.buffer, sys.stdout.buffer
    store = {}

    def params(kw):
        out = []
        for name in ("scales", "means", "weights"):
            a = np.ascontiguousarray(kw[name], np.float32)
            if kw.get("kmajor", True):
                # (K, N) contiguous storage viewed as (N, K) with strides (1, N)
                t = torch.from_numpy(np.ascontiguousarray(a.T)).permute(1, 0)
            else:
                t = torch.from_numpy(a)
            out.append(t)
        return out

    def handle(op, kw):
        if op == "gmm_encode":
            sc, mu, w = params(kw)
            sym = torch.from_numpy(np.ascontiguousarray(kw["symbols"], np.int32))
            rep = kw.get("repeat", 1)
            t0 = time.perf_counter()
            for _ in range(rep):
                b = enc.encode_with_indexes_gmm(sym, sc, mu, w, int(kw["max_value"]))
            dt = (time.perf_counter() - t0) / rep
            return bytes(b), dt
        if op == "gmm_decode":
            sc, mu, w = params(kw)
            rep = kw.get("repeat", 1)
            t0 = time.perf_counter()
            for _ in range(rep):
                o = dec.decode_with_indexes_gmm(kw["encoded"], sc, mu, w, int(kw["max_bs_value"]))
            dt = (time.perf_counter() - t0) / rep
            return o.numpy().copy(), dt
        if op == "table_encode":
            b = enc.encode_with_indexes(list(map(int, kw["symbols"])), list(map(int, kw["indexes"])),
                                        [list(map(int, r)) for r in kw["cdfs"]],
                                        list(map(int, kw["cdf_sizes"])), list(map(int, kw["offsets"])))
            return bytes(b)
        if op == "table_decode":
            o = dec.decode_with_indexes(kw["encoded"], list(map(int, kw["indexes"])),
                                        [list(map(int, r)) for r in kw["cdfs"]],
                                        list(map(int, kw["cdf_sizes"])), list(map(int, kw["offsets"])))
            return np.asarray(o, np.int32)
        if op == "scalar_encode":   # single-Gaussian overload, rans_interface.cpp:401-456, 600-607
            return bytes(enc.encode_with_indexes(list(map(int, kw["symbols"])), list(map(float, kw["scales"])),
                                                 int(kw.get("max_value", 0))))
        if op == "scalar_decode":   # rans_interface.cpp:690-764
            return np.asarray(dec.decode_with_indexes(kw["encoded"], list(map(float, kw["scales"])),
                                                      int(kw["max_bs_value"])), np.int32)
        if op == "buffered":      # BufferedRansEncoder: a list of ("table", ...) / ("gmm", ...) pushes, one flush
            be = ans.BufferedRansEncoder()
            for kind, a in kw["pushes"]:
                if kind == "table":
                    be.encode_with_indexes(list(map(int, a["symbols"])), list(map(int, a["indexes"])),
                                           [list(map(int, r)) for r in a["cdfs"]], list(map(int, a["cdf_sizes"])),
                                           list(map(int, a["offsets"])))
                else:
                    sc, mu, w = params(a)
                    be.encode_with_indexes_gmm(torch.from_numpy(np.ascontiguousarray(a["symbols"], np.int32)), sc, mu, w,
                                               int(a.get("max_value", 0)))
            return bytes(be.flush())
        if op == "decode_stream":  # set_stream once, then one decode_stream call per index list
            d2 = ans.RansDecoder()
            d2.set_stream(kw["encoded"])
            outs = []
            for idx in kw["index_lists"]:
                outs.append(np.asarray(d2.decode_stream(list(map(int, idx)), [list(map(int, r)) for r in kw["cdfs"]],
                                                        list(map(int, kw["cdf_sizes"])), list(map(int, kw["offsets"]))),
                                       np.int32))
            return outs
        if op == "put":           # keep a stream resident in the worker (bench: no pickling inside a step)
            store[kw["key"]] = (torch.from_numpy(np.ascontiguousarray(kw["symbols"], np.int32)), params(kw),
                                int(kw["max_bs_value"]))
            return True
        if op == "roundtrip":     # encode + decode every stored stream named in `keys`
            ok, te, td = True, 0.0, 0.0
            for key in kw["keys"]:
                sym, (sc, mu, w), R = store[key]
                t0 = time.perf_counter()
                b = enc.encode_with_indexes_gmm(sym, sc, mu, w, R)
                t1 = time.perf_counter()
                o = dec.decode_with_indexes_gmm(b, sc, mu, w, R)
                t2 = time.perf_counter()
                ok = ok and bool(torch.equal(o, sym))
                te += t1 - t0
                td += t2 - t1
            return ok, te, td
        if op == "ping":
            return "pong"
        raise ValueError(op)

    while True:
        hdr = stdin.read(8)
        if len(hdr) < 8:
            return
        (n,) = struct.unpack("<Q", hdr)
        op, kw = pickle.loads(stdin.read(n))
        try:
            res = (True, handle(op, kw))
        except Exception as e:  # noqa: BLE001
            res = (False, repr(e))
        blob = pickle.dumps(res, protocol=4)
        stdout.write(struct.pack("<Q", len(blob)))
        stdout.write(blob)
        stdout.flush()


if __name__ == "__main__":
    main()
