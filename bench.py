#!/usr/bin/env python
"""bench.py -- GMM rANS enc/dec throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[1]): the ten GMM streams of one ELIC2022GMM
(M=320, groups 16/16/32/64/192, two checkerboard passes each) 768x512 image --
491 520 symbols, K=4 -- with seeded synthetic entropy parameters (random-init
models code ~100 % bypass). One step = encode all ten streams
to reference-compatible single-stream rANS bytes AND decode them back.
`value` counts every coded symbol once per direction: 2 * 491 520 per step.

Arms
  b200      : flashgmm_b200 through its public API (C ABI underneath).
              value = inputs resident in HBM; e2e = inputs in pinned host memory,
              copies inside the timed region.
  reference : the reference's own compiled CPU extension (oracle/_ref), all host
              threads it can use (one process per stream).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

ELIC_GROUPS = (16, 16, 32, 64, 192)          # elic_gmm.py:92-93
LATENT_HW = (32, 48)                         # 512x768 / 16
N_IMAGES = 8                                 # rotating input sets: 8 x 25.6 MB > 126 MB L2


def elic_stream_sizes():
    h, w = LATENT_HW
    sizes = []
    for g in ELIC_GROUPS:
        sizes += [g * h * w // 2] * 2        # anchor / non-anchor halves
    return sizes


def emit(obj):
    print(json.dumps(obj), flush=True)


# ----------------------------------------------------------------------------
def start_clock_sampler(gpu_index: int):
    q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        return subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                 "-i", str(gpu_index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
    except OSError:
        return None


def stop_clock_sampler(p):
    if p is None:
        return None
    p.terminate()
    try:
        out, _ = p.communicate(timeout=5)
    except subprocess.TimeoutExpired:
        p.kill()
        out, _ = p.communicate()
    sm, mx, reasons = [], 0.0, set()
    for line in out.splitlines():
        f = [x.strip() for x in line.split(",")]
        if len(f) < 9:
            continue
        try:
            sm.append(float(f[1]))
            mx = max(mx, float(f[2]))
        except ValueError:
            continue
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
            if v.lower().startswith("active"):
                reasons.add(name)
    if not sm:
        return None
    return {"sm_mhz": statistics.median(sm), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


# ----------------------------------------------------------------------------
def make_image(torch, synth, seed, device):
    """One image = ten (sym, scales, means, weights) streams, K-major planes on `device`."""
    streams = []
    for i, n in enumerate(elic_stream_sizes()):
        streams.append(synth.gmm_torch(n, seed=seed * 100 + i, device=device))
    return streams


def to_pinned(torch, streams):
    out = []
    for sym, sg, mu, w in streams:
        hs = torch.empty(sym.shape, dtype=sym.dtype, pin_memory=True).copy_(sym)
        ps = []
        for p in (sg, mu, w):       # keep the (N,K) view with strides (1,N): planes [K][N]
            planes = torch.empty((p.size(1), p.size(0)), dtype=p.dtype, pin_memory=True).copy_(p.t())
            ps.append(planes.t())
        out.append((hs, *ps))
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist

    from flashgmm_b200 import _lib, synth
    from flashgmm_b200.batch import DecodeJob, EncodeJob

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    sizes = elic_stream_sizes()
    nsym = sum(sizes)
    images = [make_image(torch, synth, 1000 * rank + i, dev) for i in range(N_IMAGES)]
    radii = [[int(s[0].abs().max().item()) + 2 for s in img] for img in images]
    torch.cuda.synchronize()
    L = _lib.lib()
    mode = 0

    def jobs_for(imgs):
        enc = [EncodeJob(img, mode) for img in imgs]
        dec = [DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(img, rr)], mode)
               for img, rr in zip(imgs, radii)]
        return enc, dec

    def step(enc, dec, i):
        k = i % N_IMAGES
        streams = enc[k].run()
        return dec[k].run(streams), streams

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(enc, dec, steps, profile=False):
        barrier()
        if profile:
            L.flashgmm_profile_reset()
            L.flashgmm_profile_enable(1)
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            step(enc, dec, i)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        ms = max(e0.elapsed_time(e1), 0.0)
        if profile:
            L.flashgmm_profile_enable(0)
        launches = _lib.launch_count() - l0
        barrier()
        t = torch.tensor([max(ms / 1e3, wall)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches

    # ---- device-resident arm ----
    enc_d, dec_d = jobs_for(images)
    for i in range(max(args.warmup, 3)):
        outs, streams = step(enc_d, dec_d, i)
        k = i % N_IMAGES
        for o, (sym, *_r) in zip(outs, images[k]):
            if not torch.equal(o, sym):
                raise SystemExit("round trip mismatch in warm-up: refusing to report a number")
    bits = sum(s.size for s in streams) * 8 / nsym
    sampler = start_clock_sampler(local) if rank == 0 else None
    secs, launches = timed(enc_d, dec_d, args.steps, profile=True)
    clocks = stop_clock_sampler(sampler)

    import ctypes as C
    prof = {}
    for kind, name in ((0, "gmm_startfreq"), (1, "gmm_cdf_table")):
        ms, n, by = C.c_double(), C.c_uint64(), C.c_double()
        L.flashgmm_profile_read(kind, C.byref(ms), C.byref(n), C.byref(by))
        prof[name] = dict(ms=ms.value, launches=n.value, bytes=by.value)

    # encode-only / decode-only split on the same inputs (reported beside the headline)
    bs = [[x.copy() for x in enc_d[k].run()] for k in range(N_IMAGES)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.steps):
        enc_d[i % N_IMAGES].run()
    t1 = time.perf_counter()
    for i in range(args.steps):
        dec_d[i % N_IMAGES].run(bs[i % N_IMAGES])
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    enc_t, dec_t = (t1 - t0) / args.steps, (t2 - t1) / args.steps

    # ---- end-to-end arm: pinned host inputs, copies inside the timed region ----
    host_images = [to_pinned(torch, img) for img in images]
    enc_h, dec_h = [EncodeJob(img, mode) for img in host_images], None
    dec_h = [DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(img, rr)], mode)
             for img, rr in zip(host_images, radii)]
    for i in range(max(args.warmup, 3)):
        outs, _ = step(enc_h, dec_h, i)
        for o, (sym, *_r) in zip(outs, host_images[i % N_IMAGES]):
            if not torch.equal(o, sym):
                raise SystemExit("e2e round trip mismatch in warm-up")
    secs_e2e, _ = timed(enc_h, dec_h, args.steps)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    total_syms = 2 * nsym * args.steps * world
    rs = [L.flashgmm_gmm_table_row_stride(R) for R in radii[0]]
    h2d = sum(n * 52 + n * 48 for n in sizes)
    d2h = sum(n * 4 + 4 + n * r * 2 for n, r in zip(sizes, rs))
    peak, peak_src = measured_peak_gbs()
    dom = max(prof, key=lambda k: prof[k]["ms"])
    def roof(name):
        p = prof[name]
        ach = p["bytes"] / (p["ms"] / 1e3) / 1e9 if p["ms"] > 0 else 0.0
        return {"kernel": name, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                "avg_launch_us": round(1e3 * p["ms"] / max(1, p["launches"]), 2), "launches": p["launches"],
                "share_of_gpu_time": round(p["ms"] / max(1e-9, sum(v["ms"] for v in prof.values())), 3)}
    line = {
        "metric": "GMM rANS enc+dec Msymbols/s (single-stream, reference-compatible bytes)",
        "value": round(total_syms / secs / 1e6, 2), "unit": "Msymbols/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(1e3 * secs / args.steps, 4),
        "ms_per_image": round(1e3 * secs / args.steps, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "elic2022gmm_768x512: 10 GMM streams/image, 491520 symbols, K=4, APPROX_MODE=0",
                   "step": "encode 10 streams + decode 10 streams of one image per GPU",
                   "l2": f"inputs rotate over {N_IMAGES} images = {N_IMAGES * nsym * 52 / 1e6:.0f} MB > 126 MB L2",
                   "bits_per_symbol": round(bits, 3), "search_radius": radii[0][-1],
                   "host_workers": int(os.environ.get("FLASHGMM_WORKERS", "0")) or "auto(min(16, cores-1))",
                   "host_cores": os.cpu_count()},
        "encode_ms_per_image": round(1e3 * enc_t, 4), "decode_ms_per_image": round(1e3 * dec_t, 4),
        "encode_msym_s": round(nsym / enc_t / 1e6, 2), "decode_msym_s": round(nsym / dec_t / 1e6, 2),
        "e2e": {"value": round(total_syms / secs_e2e / 1e6, 2), "unit": "Msymbols/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(1e3 * secs_e2e / args.steps, 4)},
        "gpu_launches": int(launches),
        "roofline": roof(dom),
        "roofline_kernels": {k: roof(k) for k in prof},
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_sample(images[0], radii[0])
    emit(line)
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------
def _np_streams(streams):
    import numpy as np
    out = []
    for sym, sg, mu, w in streams:
        out.append((sym.cpu().numpy(), *(np.ascontiguousarray(p.cpu().numpy()) for p in (sg, mu, w))))
    return out


def cpu_baseline_sample(streams, radii, repeat=8):
    """Reference CPU path (oracle/_ref, else the C port) on one image, one thread."""
    import numpy as np

    from oracle import port, ref
    data = _np_streams(streams)
    nsym = sum(d[0].size for d in data)
    te = td = 0.0
    if ref.available():
        kind = "reference"
        with ref.RefWorker(0) as wk:
            for (sym, sg, mu, w), R in zip(data, radii):
                b, dt = wk.gmm_encode(sym, sg, mu, w, R, repeat=repeat)
                d, dt2 = wk.gmm_decode(b, sg, mu, w, R, repeat=repeat)
                assert np.array_equal(d, sym)
                te += dt
                td += dt2
    else:
        kind = "port"
        for (sym, sg, mu, w), R in zip(data, radii):
            t0 = time.perf_counter()
            b = port.gmm_encode(sym, sg, mu, w, 0)
            t1 = time.perf_counter()
            port.gmm_decode(b, sg, mu, w, R, 0)
            t2 = time.perf_counter()
            te += t1 - t0
            td += t2 - t1
    return {"value": round(2 * nsym / (te + td) / 1e6, 3), "unit": "Msymbols/s", "cores": 1, "kind": kind,
            "sample": f"one image (10 streams, {nsym} symbols) encode+decode, x{repeat if kind == 'reference' else 1} repeats, "
                      "USE_SIMD=1 APPROX_MODE=0, tensors already on the CPU",
            "encode_msym_s": round(nsym / te / 1e6, 3), "decode_msym_s": round(nsym / td / 1e6, 3)}


def run_reference(args):
    """The reference's own CPU implementation, one worker process per stream."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import threading

    import numpy as np
    import torch

    from flashgmm_b200 import synth
    from oracle import port, ref

    sizes = elic_stream_sizes()
    nsym = sum(sizes)
    n_img = 2
    images = [_np_streams(make_image(torch, synth, i, "cpu")) for i in range(n_img)]
    radii = [[synth.search_radius(s[0]) for s in img] for img in images]
    cores = os.cpu_count() or 1
    P = max(1, min(cores, len(sizes)))
    use_ref = ref.available()
    kind = "reference" if use_ref else "port"
    # longest-processing-time assignment of streams to workers
    order = sorted(range(len(sizes)), key=lambda i: -sizes[i])
    load, assign = [0] * P, [[] for _ in range(P)]
    for i in order:
        k = load.index(min(load))
        assign[k].append(i)
        load[k] += sizes[i]
    workers = [ref.RefWorker(0) if use_ref else None for _ in range(P)]

    def work(k, img, rr, ok):
        for i in assign[k]:
            sym, sg, mu, w = img[i]
            if use_ref:
                b, _ = workers[k].gmm_encode(sym, sg, mu, w, rr[i])
                d, _ = workers[k].gmm_decode(b, sg, mu, w, rr[i])
            else:
                b = port.gmm_encode(sym, sg, mu, w, 0)
                d = port.gmm_decode(b, sg, mu, w, rr[i], 0)
            ok[i] = bool(np.array_equal(d, sym))

    def step(i):
        img, rr = images[i % n_img], radii[i % n_img]
        ok = [False] * len(sizes)
        th = [threading.Thread(target=work, args=(k, img, rr, ok)) for k in range(P)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert all(ok)

    for i in range(args.warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
    secs = time.perf_counter() - t0
    for wk in workers:
        if wk is not None:
            wk.close()
    val = round(2 * nsym * args.steps / secs / 1e6, 3)
    emit({
        "impl": "reference", "metric": "GMM rANS enc+dec Msymbols/s (single-stream, reference-compatible bytes)",
        "value": val, "unit": "Msymbols/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(1e3 * secs / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "elic2022gmm_768x512: 10 GMM streams/image, 491520 symbols, K=4, APPROX_MODE=0",
                   "step": "encode 10 streams + decode 10 streams of one image", "host_cores": cores},
        "cpu_baseline": {"value": val, "unit": "Msymbols/s", "cores": P, "kind": kind,
                         "sample": f"whole image per step; {P} worker processes (one per stream, LPT), "
                                   "pickled numpy hand-over outside the reference call but inside the step"},
        "e2e": {"value": val, "unit": "Msymbols/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
