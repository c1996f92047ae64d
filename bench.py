#!/usr/bin/env python
"""bench.py -- GMM rANS enc/dec throughput on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--decode-order pass-sequential|pass-parallel] [--pipeline-depth D]
                    [--workload images24|sharded49] [--images-per-step B] [--mode 0|1|2]

Workload (BASELINE.json configs[1] shape, batch of configs[4]): per GPU and step, the GMM streams of 24 synthetic
768x512 ELIC2022GMM images (M=320, groups 16/16/32/64/192, two checkerboard passes each: 10 streams and 491 520
symbols per image, K=4), seeded synthetic entropy parameters (random-init models code ~100 % bypass).
One step = encode all 240 streams to reference-compatible single-stream rANS bytes AND decode them back, bit-exactly.
`value` counts every coded symbol once per direction: 2 * 491 520 * 24 per step.

Decode order. A real decoder gets pass p+1's parameters from pass p's decoded symbols, so the default
is PASS-SEQUENTIAL: ten dependent decode calls per step, each carrying the 24 streams of one pass, the next one issued
only after the previous has returned. (Round 1 decoded all 240 streams in one call; that number is kept beside the
headline as `decode_pass_parallel_*`, labelled an upper bound.) The encoder knows all of y up front: one call.

Pipeline. Images are independent, so `--pipeline-depth D` host threads (one library ctx each) each take whole steps:
one batch's encode overlaps other batches' pass-sequential decodes on the GPU (the decode chain of a batch is serial
and uses 24 warps; the machine has room for more). Every step still codes its own 24 images completely; K steps are
timed from the first call to the last return. `value_one_batch_in_flight` is the D = 1 figure.

Arms
  b200      : flashgmm_b200 through its public batch API (C ABI underneath).
              value = inputs resident in HBM; e2e = inputs in page-locked host memory, transfers inside the timed
              region (the kernels read the parameters over PCIe in place; bitstreams and symbols come back by copy).
              Encode: startfreq kernels -> serial push on host workers / push kernel. Decode: windowed rows kernel ->
              pop kernel, one warp per stream, entirely on the GPU.
  reference : the reference's own compiled CPU extension (oracle/_ref), all host cores (one process per core).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

ELIC_GROUPS = (16, 16, 32, 64, 192)          # elic_gmm.py:92-93
LATENT_HW = (32, 48)                         # 512x768 / 16
N_SETS = 2                                   # rotating batches of images (each batch > 126 MB L2 for B >= 5)
# Constants from the committed ncu --set full captures of round 2 (profiles/r02_ncu_summary.json, made by
# tools/gpu_profile.sh on the same kernels: startfreq_batched 3 538 944 symbols, windowed rows 786 432, pop 786 432,
# push 3 538 944). `traffic` in the bench line = these DRAM bytes per symbol x the symbols per launch of THIS run: it is a
# constant from the profile, not a counter read during the run (ncu cannot run inside the timed region).
NCU_DRAM_BYTES_PER_SYMBOL = {"gmm_startfreq": 54.9, "gmm_cdf_rows": 245.0, "gmm_pop_rows": 280.0, "gmm_push": 4.3}
NCU_PIPES = {   # issue slots per cycle and scheduler; busiest-pipe utilisation in % of peak (sm__inst_executed_pipe_*)
    "gmm_startfreq": {"issue_active": 0.82, "pipe_fma": 42.8, "pipe_xu": 52.2, "pipe_alu": 42.9, "source": "r02_ncu_summary.json push[0]"},
    "gmm_cdf_rows": {"issue_active": 0.87, "pipe_fma": 49.0, "pipe_xu": 30.5, "pipe_alu": 44.6, "pipe_lsu": 14.1,
                     "warp_insts_per_symbol": 316, "by_mode_us_per_786432_symbols": {"0": 262, "1": 293, "2": 325},
                     "source": "r02_ncu_summary.json rows_final (after the register-resident rewrite)"},
    "gmm_pop_rows": {"issue_active": 0.23, "pipe_fma": 1.8, "pipe_xu": 0.0, "pipe_alu": 5.4, "warps_active_pct": 1.56,
                     "source": "r02_ncu_summary.json pop[2]"},
    "gmm_push": {"issue_active": 0.39, "pipe_fma": 6.0, "pipe_xu": 0.0, "pipe_alu": 9.2, "warps_active_pct": 1.56,
                 "source": "r02_ncu_summary.json push[1]"},
}
ALGO_BYTES_PER_SYMBOL = {"gmm_startfreq": 56.0, "gmm_cdf_rows": 48.0 + 272.0, "gmm_pop_rows": 272.0 + 4.0, "gmm_push": 4.0 + 8.0 * 0.14}


def elic_stream_sizes():
    h, w = LATENT_HW
    sizes = []
    for g in ELIC_GROUPS:
        sizes += [g * h * w // 2] * 2        # anchor / non-anchor halves
    return sizes


def common_config(args):
    """The `config` both arms print (the driver compares them): what is coded per step, not how."""
    B = args.images_per_step
    nsym = sum(elic_stream_sizes())
    return {"workload": f"elic2022gmm_768x512 x{B} images/step: 10 GMM streams/image, 491520 symbols/image, K=4, "
                        f"APPROX_MODE={args.mode}",
            "step": f"encode {10 * B} streams + decode {10 * B} streams ({B} images) per GPU; every step codes its images "
                    "completely, bytes identical to the reference's single-stream coder",
            "images_per_step": B, "approx_mode": args.mode, "decode_order": args.decode_order, "job": args.workload,
            "l2": f"inputs rotate over {N_SETS} batches of {B} images = {N_SETS * B * nsym * 52 / 1e6:.0f} MB "
                  "(one batch alone exceeds the 126 MB L2)"}


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


def bind_to_gpu_cpus(local_rank):
    """Several ranks on one box: run this rank (its worker threads, its page-locked buffers, which
    are first touched here) on the CPUs NVML reports as local to its GPU, so that host<->device
    traffic does not cross the socket interconnect. Returns the number of CPUs bound to, or None."""
    if os.environ.get("FLASHGMM_BENCH_NO_BIND"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        idx = local_rank
        if vis and all(v.strip().isdigit() for v in vis.split(",")):
            idx = int(vis.split(",")[local_rank])
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 64) + 63) // 64)
        cpus = {64 * i + b for i, wd in enumerate(words) for b in range(64) if (wd >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


def run_b200(args):
    import ctypes as C
    import threading

    import torch
    import torch.distributed as dist

    from flashgmm_b200 import _lib, shard, synth
    from flashgmm_b200.batch import DecodeJob, EncodeJob

    _lib.want_more_hw_queues(32)          # before the CUDA context exists (read at context creation)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    # host worker threads (serial push of the encoder) are shared by the ranks of one box
    if world > 1 and "FLASHGMM_WORKERS" not in os.environ:
        lws = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
        per_rank = (os.cpu_count() or 1) // max(1, lws)
        # every pipeline lane has its own ctx and worker pool. With cores to spare the pools may overlap (encodes of
        # different lanes rarely coincide); with few cores per rank (8-GPU box: 4) the pools are cut down so that the cost
        # model sends the push to the GPU (one warp per stream) instead of queueing host threads: measured at N = 8,
        # 3 workers per ctx 10.4 Gsym/s, 1 worker per ctx 13.6 Gsym/s
        workers = per_rank - 1 if per_rank >= 12 else (per_rank - 1) // 4
        os.environ["FLASHGMM_WORKERS"] = str(max(1, min(16, workers)))
    numa_cpus = bind_to_gpu_cpus(local) if world > 1 else None
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B = args.images_per_step
    items_all = None
    if args.workload == "sharded49":
        # BASELINE.json configs[4] as written: 24 images + one 3840x2160 image cut into 5 x 5 tiles of 768x512 (bottom padded
        # to 2560) = 49 independent work items of equal cost, dealt over the ranks by shard.assign: STRONG scaling
        tiles = shard.tile_image(2160, 3840)
        items_all = [("image", i) for i in range(24)] + [("tile", t) for t in range(len(tiles))]
        mine = shard.my_items([1.0] * len(items_all), world, rank)
        B = len(mine)
    sizes = elic_stream_sizes()
    n_pass = len(sizes)
    nsym = sum(sizes)
    mode = args.mode
    order = args.decode_order
    depth = args.pipeline_depth
    if depth <= 0:      # default: 6 batches in flight. Measured (Msym/s, depth 4 -> 6): N=1 2282 -> 2818, N=4 7867 -> 10384,
        depth = 6       # N=8 15810 -> 17253; 8 adds 2 % at N=1
    # two rotating batches of B images; one batch alone (B x 25.6 MB) already exceeds L2 for B >= 5
    if items_all is None:
        batches = [[make_image(torch, synth, 1000 * rank + 100 * s + i, dev) for i in range(B)] for s in range(N_SETS)]
    else:       # the item's index seeds its content: the job codes the same 49 items whatever the number of ranks
        batches = [[make_image(torch, synth, 50000 + 100 * s + it, dev) for it in mine] for s in range(N_SETS)]
    radii = [[[int(st[0].abs().max().item()) + 2 for st in img] for img in bt] for bt in batches]
    torch.cuda.synchronize()
    L = _lib.lib()

    def flat(bt):
        return [st for img in bt for st in img]

    class Lane:
        """One host thread's share of the pipeline: its own ctx (a ctx serialises its calls) and pre-bound jobs.
        A step = encode every stream of the batch in one call (the encoder knows all of y up front), then decode
        pass by pass: pass p's call carries the B streams of pass p, and pass p+1 is not issued before it returns
        (a real decoder forms pass p+1's parameters from pass p's symbols)."""

        def __init__(self, bts, ctx):
            self.ctx = ctx
            self.enc = [EncodeJob(flat(bt), mode, ctx=self.ctx) for bt in bts]
            self.dec = []
            for bt, rb in zip(bts, radii):
                per_stream = [[(sg, mu, w, R) for (_, sg, mu, w), R in zip(img, rr)] for img, rr in zip(bt, rb)]
                if order == "pass-sequential":
                    self.dec.append([DecodeJob([per_stream[i][p] for i in range(B)], mode, ctx=self.ctx) for p in range(n_pass)])
                else:
                    self.dec.append([DecodeJob([st for img in per_stream for st in img], mode, ctx=self.ctx)])

        def encode(self, k):
            return self.enc[k].run()

        def decode(self, k, streams):
            """-> decoded symbols in flat (image, pass) order"""
            if order != "pass-sequential":
                return self.dec[k][0].run(streams)
            outs = [None] * (B * n_pass)
            for p, job in enumerate(self.dec[k]):
                for i, o in enumerate(job.run([streams[i * n_pass + p] for i in range(B)])):
                    outs[i * n_pass + p] = o
            return outs

        def step(self, i):
            k = i % N_SETS
            streams = self.encode(k)
            return self.decode(k, streams), streams

    def check(outs, bt):
        for o, (sym, *_r) in zip(outs, flat(bt)):
            if not torch.equal(o, sym):
                raise SystemExit("round trip mismatch in warm-up: refusing to report a number")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def run_steps(lanes, steps):
        """`steps` steps over the lanes: lane j takes steps j, j + D, ... (independent images); with D > 1 one
        batch's encode and another's pass-sequential decode overlap on the GPU."""
        if len(lanes) == 1:
            for i in range(steps):
                lanes[0].step(i)
            return
        errs = []

        def work(j):
            try:
                torch.cuda.set_device(local)
                for i in range(j, steps, len(lanes)):
                    lanes[j].step(i)
            except BaseException as e:                       # noqa: BLE001
                errs.append(e)
        th = [threading.Thread(target=work, args=(j,)) for j in range(len(lanes))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]

    def timed(lanes, steps, profile=False):
        barrier()
        if profile:
            L.flashgmm_profile_reset()
            L.flashgmm_profile_enable(1)
        l0 = _lib.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        run_steps(lanes, steps)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        ms = max(e0.elapsed_time(e1), 0.0)
        if profile:
            L.flashgmm_profile_enable(0)
        launches = _lib.launch_count() - l0
        barrier()
        # the library's kernels run on its own streams: the step ends when the host call returns,
        # so the wall clock of the synchronous calls is the step time (events on torch's stream agree)
        return shard.max_over_ranks(max(ms / 1e3, wall), device=dev), launches

    W = max(args.warmup, 3)
    # ---- device-resident arm ----
    ctxs = [_lib.ctx(local) if j == 0 else _lib.new_ctx(local) for j in range(depth)]
    # Several batches in flight, inputs in HBM: the lanes' ctxs are created with FLASHGMM_PUSH=device -- the push warps'
    # latency (7.2 ms per batch against 2.7 on 15 host workers) hides behind the other lanes' decodes, and D x 15 worker
    # threads no longer share the 16 cores with the D lane threads (measured, D = 6: 2830-3130 Msym/s against 2500-2800).
    # One batch in flight, the per-kernel profile, the latency figures and the end-to-end arm keep the default policy
    # (cost-model split). With several ranks per host the cost model gets there by itself (one worker per ctx).
    push_device_lanes = world == 1 and depth >= 3 and "FLASHGMM_PUSH" not in os.environ
    if push_device_lanes:
        os.environ["FLASHGMM_PUSH"] = "device"
        try:
            ctxs_pipe = [_lib.new_ctx(local) for _ in range(depth)]
        finally:
            del os.environ["FLASHGMM_PUSH"]
    else:
        ctxs_pipe = ctxs
    lanes_d = [Lane(batches, ctxs_pipe[j]) for j in range(depth)]
    lanes_1 = [Lane(batches, ctxs[0])] if push_device_lanes else lanes_d[:1]
    for ln in lanes_d + (lanes_1 if push_device_lanes else []):
        for i in range(W if ln in (lanes_d[0], lanes_1[0]) else 2):
            outs, streams = ln.step(i)
            check(outs, batches[i % N_SETS])
    bits = sum(s.size for s in streams) * 8 / (nsym * B)
    sampler = start_clock_sampler(local) if rank == 0 else None
    secs, launches = timed(lanes_d, args.steps)
    clocks = stop_clock_sampler(sampler)
    secs_d1, _ = timed(lanes_1, max(2, args.steps // 2)) if depth > 1 else (secs, 0)
    steps_d1 = max(2, args.steps // 2) if depth > 1 else args.steps
    # per-kernel CUDA-event timing in a separate pass on one lane (events add host work to every launch)
    timed(lanes_1, max(2, args.steps // 4), profile=True)
    prof = {}
    for kind, name in ((0, "gmm_startfreq"), (1, "gmm_cdf_rows"), (4, "gmm_pop_rows"), (5, "gmm_push")):
        ms, n, by = C.c_double(), C.c_uint64(), C.c_double()
        L.flashgmm_profile_read(kind, C.byref(ms), C.byref(n), C.byref(by))
        if n.value:
            prof[name] = dict(ms=ms.value, launches=n.value, bytes=by.value)

    # encode-only / decode-only split on the same inputs, one lane (reported beside the headline)
    ln0 = lanes_1[0]
    bs = [[x.copy() for x in ln0.encode(k)] for k in range(N_SETS)]
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.steps):
        ln0.encode(i % N_SETS)
    t1 = time.perf_counter()
    for i in range(args.steps):
        ln0.decode(i % N_SETS, bs[i % N_SETS])
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    enc_t, dec_t = (t1 - t0) / args.steps, (t2 - t1) / args.steps
    # the other decode order beside it, labelled (pass-parallel = all 10 x B streams in ONE call: an upper bound
    # no real decoder reaches, since pass p+1's parameters need pass p's symbols)
    other = DecodeJob([(sg, mu, w, R) for img, rr in zip(batches[0], radii[0]) for (_, sg, mu, w), R in zip(img, rr)], mode)
    other.run(bs[0])
    t0 = time.perf_counter()
    for i in range(max(2, args.steps // 2)):
        other.run(bs[0])
    torch.cuda.synchronize()
    dec_parallel_t = (time.perf_counter() - t0) / max(2, args.steps // 2)
    del other

    # single-image latency (one image's ten streams: one encode call, ten dependent decode calls)
    e1j = EncodeJob(batches[0][0], mode)
    if order == "pass-sequential":
        d1j = [DecodeJob([(sg, mu, w, R)], mode) for (_, sg, mu, w), R in zip(batches[0][0], radii[0][0])]
    else:
        d1j = [DecodeJob([(sg, mu, w, R) for (_, sg, mu, w), R in zip(batches[0][0], radii[0][0])], mode)]
    lat_e, lat_d = [], []
    for i in range(12):
        t0 = time.perf_counter()
        b1 = e1j.run()
        t1 = time.perf_counter()
        if order == "pass-sequential":
            for p, job in enumerate(d1j):
                job.run([b1[p]])
        else:
            d1j[0].run(b1)
        t2 = time.perf_counter()
        if i >= 2:
            lat_e.append(t1 - t0)
            lat_d.append(t2 - t1)
    del e1j, d1j

    # ---- end-to-end arm: pinned host inputs, copies inside the timed region ----
    if args.no_e2e:
        secs_e2e = float("nan")
    else:
        host_batches = [[to_pinned(torch, img) for img in bt] for bt in batches]
        lanes_h = [Lane(host_batches, ctxs[j]) for j in range(depth)]
        for ln in lanes_h:
            for i in range(W if ln is lanes_h[0] else 2):
                outs, _ = ln.step(i)
                check(outs, host_batches[i % N_SETS])
        secs_e2e, _ = timed(lanes_h, args.steps)

    micro = kernel_microbench(torch, synth, L, _lib, dev, mode) if (world == 1 and not args.no_microbench) else None

    item_table = None
    if items_all is not None:
        per_item = [(it, int(sum(streams[i * n_pass + p].size for p in range(n_pass)))) for i, it in enumerate(mine)]
        item_table = shard.gather_item_sizes(len(items_all), per_item, device=dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    total_syms = 2 * nsym * B * args.steps * world
    item_bytes = None
    if items_all is not None:
        total_syms = 2 * nsym * len(items_all) * args.steps
    stream_bytes = sum(s.size for s in streams)                      # last step's bitstreams (whole batch)
    h2d = B * sum(n * 52 + n * 48 for n in sizes) + stream_bytes     # enc: symbols + params; dec: params + bitstreams
    d2h = B * sum(n * 4 + 4 + n * 4 for n in sizes) + 16 * B * len(sizes)   # enc: packed records; dec: symbols + state
    peak, peak_src = measured_peak_gbs()
    dom = max(prof, key=lambda k: prof[k]["ms"])
    issue_peak = 148 * 4 * (clocks["sm_mhz"] if clocks else 1965.0) * 1e6      # warp-instructions / s

    notes = {
        "gmm_startfreq": "FP32-issue-bound by construction (bit-exact IEEE div/sqrt + polynomial exp, 8 Phi per symbol); "
                         "launches here are per stream and small, the 16 Mi-symbol microbench line is the kernel at size",
        "gmm_cdf_rows": "FP32-issue-bound by construction (33 window boundaries x 4 components per symbol whatever the search "
                        "radius, ~47 SASS instructions per Phi); DESIGN.md section 5",
        "gmm_pop_rows": "latency-bound by construction: the rANS state recurrence is serial per stream, one warp per "
                        "stream, ~100 cycles per symbol; HBM fraction is not the figure of merit, ns/symbol/stream is",
        "gmm_push": "latency-bound by construction: the rANS state recurrence is serial per stream, one warp per "
                    "stream, ~95 cycles per symbol; runs beside the host worker threads on the shorter streams",
    }

    def roof(name):
        p = prof[name]
        ach = p["bytes"] / (p["ms"] / 1e3) / 1e9 if p["ms"] > 0 else 0.0
        per_launch = p["bytes"] / max(1, p["launches"])
        dram = NCU_DRAM_BYTES_PER_SYMBOL.get(name)
        return {"kernel": name, "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4),
                "traffic": round(dram * per_launch / ALGO_BYTES_PER_SYMBOL[name]) if dram else None,
                "algorithmic_bytes_per_launch": round(per_launch), "peak_source": peak_src,
                "avg_launch_us": round(1e3 * p["ms"] / max(1, p["launches"]), 2), "launches": p["launches"],
                "share_of_summed_kernel_time": round(p["ms"] / max(1e-9, sum(v["ms"] for v in prof.values())), 3),
                "share_note": "share of the SUM of kernel durations (launches overlap on several CUDA streams), not of wall time",
                "traffic_note": "DRAM bytes per symbol from profiles/r02_ncu_summary.json x symbols per launch of this run",
                "ncu_pipes": NCU_PIPES.get(name),
                "note": notes[name]}

    line = {
        "metric": "GMM rANS enc+dec Msymbols/s (single-stream, reference-compatible bytes)",
        "value": round(total_syms / secs / 1e6, 2), "unit": "Msymbols/s", "n_gpus": world,
        "steps": args.steps, "warmup": W, "ms_per_step": round(1e3 * secs / args.steps, 4),
        "ms_per_image": round(1e3 * secs / args.steps / B, 4),
        "higher_is_better": True, "scaling": "weak" if items_all is None else "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": common_config(args),
        "details": {
                   "bits_per_symbol": round(bits, 3), "search_radius": radii[0][0][-1],
                   "decode_order": order + (f": {n_pass} dependent decode calls of {B} streams per step (pass p+1 is issued only "
                                            "after pass p has returned)" if order == "pass-sequential" else
                                            f": all {n_pass * B} streams in one call (upper bound, not a real decoder)"),
                   "pipeline_depth": depth,
                   "pipeline": f"{depth} independent batches in flight (one host thread + ctx each): one batch's encode overlaps "
                               "another's pass-sequential decode; every step still codes its own 24 images completely",
                   "decode": "device-resident (rows kernel -> pop kernel, one warp per stream)",
                   "encode": "batched startfreq launches -> serial push: the longest streams on host worker threads (records "
                             "come down wave by wave), the rest on the device (push kernel, one warp per stream), "
                             "split by a cost model (FLASHGMM_PUSH)"
                             + ("; the lanes of `value` (several batches in flight, inputs in HBM) use FLASHGMM_PUSH=device: "
                                "every stream on the push kernel" if push_device_lanes else ""),
                   "push_policy_of_value_lanes": "device" if push_device_lanes else "cost model",
                   "host_workers": int(os.environ.get("FLASHGMM_WORKERS", "0")) or min(16, max(1, (os.cpu_count() or 2) - 1)),
                   "host_cores": os.cpu_count(), "cpus_bound_to_gpu_locality": numa_cpus},
        "value_one_batch_in_flight": round(2 * nsym * B * steps_d1 * world / secs_d1 / 1e6, 2),
        "decode_pass_parallel_msym_s": round(nsym * B / dec_parallel_t / 1e6, 2),
        "decode_pass_parallel_ms_per_step": round(1e3 * dec_parallel_t, 4),
        "encode_ms_per_step": round(1e3 * enc_t, 4), "decode_ms_per_step": round(1e3 * dec_t, 4),
        "encode_ms_per_image": round(1e3 * enc_t / B, 4), "decode_ms_per_image": round(1e3 * dec_t / B, 4),
        "encode_msym_s": round(nsym * B / enc_t / 1e6, 2), "decode_msym_s": round(nsym * B / dec_t / 1e6, 2),
        "latency_single_image_ms": {"encode": round(1e3 * statistics.median(lat_e), 4),
                                    "decode": round(1e3 * statistics.median(lat_d), 4)},
        "e2e": {"value": round(total_syms / secs_e2e / 1e6, 2), "unit": "Msymbols/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(1e3 * secs_e2e / args.steps, 4),
                "host_memory_read_gb_s_all_ranks": round(world * h2d * args.steps / secs_e2e / 1e9, 1),
                "note": "per GPU and step; every rank reads its page-locked inputs through the host's memory system: on the "
                        "single-socket 8-GPU box ~160 GB/s in aggregate is the ceiling of this number"},
        "gpu_launches": int(launches),
        "roofline": roof(dom),
        "roofline_kernels": {k: roof(k) for k in prof},
        "issue_peak_gwarpinst_s": round(issue_peak / 1e9, 1),
        "clocks": clocks,
    }
    if item_table is not None:
        line["sharded49"] = {"items": len(item_table), "items_on_rank0": B, "item_bytes": item_table,
                             "total_bytes": int(sum(item_table)),
                             "note": "24 images + 25 tiles (3840x2160 as 5x5 tiles of 768x512), shard.assign over the ranks; "
                                     "per-item coded bytes gathered with shard.gather_item_sizes"}
    if micro:
        line["microbench"] = micro
    if world == 1 and not args.no_microbench:
        # side measurements: a failure in one of them is recorded in its place, the headline line is still printed
        def extra(key, fn):
            try:
                line[key] = fn()
            except (Exception, SystemExit) as e:              # noqa: BLE001
                line[key] = {"error": f"{type(e).__name__}: {e}"[:300]}

        extra("mode_and_radius_sweep", lambda: sweep_modes_and_radius(torch, synth, L, _lib, dev))
        extra("likelihood", lambda: likelihood_bench(torch, L, _lib, dev))
        extra("model_codec", lambda: model_codec_bench(torch, dev, B))
        extra("model_codec_single_image", lambda: model_codec_bench(torch, dev, 1, reps=5))   # what the reference's eval scripts do
        extra("model_codec_two_lanes", lambda: model_codec_bench(torch, dev, B, reps=3, lanes=2))
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_sample(batches[0][0], radii[0][0])
        try:
            full = reference_full_call(torch, dev)
        except Exception as e:                                # noqa: BLE001
            full = {"error": f"{type(e).__name__}: {e}"[:300]}
        if full:
            line["cpu_baseline_full_call"] = full
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def kernel_microbench(torch, synth, L, _lib, dev, mode, n=1 << 24, iters=10):
    """BASELINE.json configs[3]: 16 Mi symbols, K=4 -- the kernels timed alone with CUDA events
    on the launching stream (inputs 872 MB, far larger than L2)."""
    sym, sg, mu, w = synth.gmm_torch(n, seed=1234, device=dev)
    out = torch.empty(n, dtype=torch.int32, device=dev)
    st = torch.cuda.Stream(device=dev)
    h = _lib.ctx(dev.index)
    res = {"n": n}
    with torch.cuda.stream(st):
        def sf():
            _lib.check(L.flashgmm_gmm_startfreq(h, sym.data_ptr(), sg.data_ptr(), mu.data_ptr(), w.data_ptr(), n,
                                                sg.stride(0), sg.stride(1), mode, out.data_ptr(), st.cuda_stream))
        for _ in range(3):
            sf()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(iters):
            sf()
        e1.record(st)
        st.synchronize()
        ms = e0.elapsed_time(e1) / iters
        peak, _src = measured_peak_gbs()
        res["gmm_startfreq"] = {"ms": round(ms, 4), "gsym_s": round(n / ms / 1e6, 2),
                                "gb_s": round(56 * n / ms / 1e6, 1), "hbm_frac": round(56 * n / ms / 1e6 / peak, 4)}
        R = 22
        rs = L.flashgmm_gmm_table_row_stride(R)
        m = 1 << 22
        tab = torch.empty((m, rs), dtype=torch.int16, device=dev)
        def rows():
            _lib.check(L.flashgmm_gmm_cdf_table(h, sg.data_ptr(), mu.data_ptr(), w.data_ptr(), m, sg.stride(0),
                                                sg.stride(1), mode, R, tab.data_ptr(), rs, st.cuda_stream))
        for _ in range(2):
            rows()
        e0.record(st)
        for _ in range(iters):
            rows()
        e1.record(st)
        st.synchronize()
        ms = e0.elapsed_time(e1) / iters
        by = (48 + 2 * rs) * m
        res["gmm_cdf_rows"] = {"n": m, "R": R, "ms": round(ms, 4), "gsym_s": round(m / ms / 1e6, 3),
                               "gb_s": round(by / ms / 1e6, 1), "hbm_frac": round(by / ms / 1e6 / peak, 4)}
    del tab, out
    # single-stream (reference-compatible) vs interleaved (opt-in format) on the same 16 Mi symbols
    import ctypes as C

    from flashgmm_b200 import ans, ileave
    Rs = int(sym.abs().max().item()) + 2
    t0 = time.perf_counter()
    single = ans.gmm_encode(sym, sg, mu, w, mode)
    t1 = time.perf_counter()
    dec = ans.gmm_decode(single, sg, mu, w, Rs, mode)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    assert torch.equal(dec, sym)
    res["single_stream"] = {"bytes": len(single), "encode_ms": round(1e3 * (t1 - t0), 2), "decode_ms": round(1e3 * (t2 - t1), 2),
                            "encode_msym_s": round(n / (t1 - t0) / 1e6, 1), "decode_msym_s": round(n / (t2 - t1) / 1e6, 1),
                            "note": "one stream = one host thread for the serial rANS recurrence"}
    blob = ileave.encode(sym, sg, mu, w, Rs, mode)            # warm-up (buffers)
    assert torch.equal(ileave.decode(blob, sg, mu, w), sym)
    L.flashgmm_profile_reset()
    L.flashgmm_profile_enable(1)
    reps = 3
    t0 = time.perf_counter()
    for _ in range(reps):
        blob = ileave.encode(sym, sg, mu, w, Rs, mode, as_view=True)      # the container in page-locked host memory
    t1 = time.perf_counter()
    blob = bytes(blob)
    for _ in range(reps):
        ileave.decode(blob, sg, mu, w)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    L.flashgmm_profile_enable(0)
    il = {"bytes": len(blob), "size_vs_single": round(len(blob) / len(single), 4), "segment_log2": ileave.info(blob)["segment_log2"],
          "encode_call_ms": round(1e3 * (t1 - t0) / reps, 3), "decode_call_ms": round(1e3 * (t2 - t1) / reps, 3)}
    for kind, name in ((2, "encode"), (3, "decode")):
        msv, nl, by = C.c_double(), C.c_uint64(), C.c_double()
        L.flashgmm_profile_read(kind, C.byref(msv), C.byref(nl), C.byref(by))
        k_ms = msv.value / max(1, nl.value)
        il[f"{name}_kernel_ms"] = round(k_ms, 4)
        il[f"{name}_kernel_gsym_s"] = round(n / k_ms / 1e6, 2) if k_ms > 0 else None
        il[f"{name}_hbm_frac"] = round(52 * n / k_ms / 1e6 / peak, 4) if k_ms > 0 else None
    res["interleaved"] = il
    return res


def _prof_read(L, kind):
    import ctypes as C
    ms, n, by = C.c_double(), C.c_uint64(), C.c_double()
    L.flashgmm_profile_read(kind, C.byref(ms), C.byref(n), C.byref(by))
    return ms.value, n.value, by.value


def sweep_modes_and_radius(torch, synth, L, _lib, dev, n=1 << 22, iters=5):
    """BASELINE.json configs[2] (APPROX_MODE sweep) and VERDICT r1 #5 (cost vs search radius): the startfreq kernel
    timed alone per mode (CUDA events on the launching stream), and one decode pass of 24 x 147 456-symbol streams
    per mode and per search radius with the library's per-kernel event timers (rows and pop kernels)."""
    from flashgmm_b200.batch import DecodeJob, EncodeJob
    peak, _src = measured_peak_gbs()
    sym, sg, mu, w = synth.gmm_torch(n, seed=4321, device=dev)
    out = torch.empty(n, dtype=torch.int32, device=dev)
    st = torch.cuda.Stream(device=dev)
    h = _lib.ctx(dev.index)
    res = {"startfreq": {}, "decode_pass": {}, "radius": {}}
    with torch.cuda.stream(st):
        for mode in (0, 1, 2):
            def sf():
                _lib.check(L.flashgmm_gmm_startfreq(h, sym.data_ptr(), sg.data_ptr(), mu.data_ptr(), w.data_ptr(), n,
                                                    sg.stride(0), sg.stride(1), mode, out.data_ptr(), st.cuda_stream))
            for _ in range(2):
                sf()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(iters):
                sf()
            e1.record(st)
            st.synchronize()
            ms = e0.elapsed_time(e1) / iters
            res["startfreq"][f"mode{mode}"] = {"n": n, "ms": round(ms, 4), "gsym_s": round(n / ms / 1e6, 2),
                                               "hbm_frac": round(56 * n / ms / 1e6 / peak, 4)}
    del out
    streams = [synth.gmm_torch(147456, seed=900 + i, device=dev) for i in range(24)]
    R0 = max(int(s[0].abs().max().item()) for s in streams) + 2
    # a ctx that always takes the device path (the cost model sends passes with small rows to the host workers): these
    # entries are about the rows / pop KERNELS
    old_policy = os.environ.get("FLASHGMM_POP_POLICY")
    os.environ["FLASHGMM_POP_POLICY"] = "device"
    try:
        dctx = _lib.new_ctx(dev.index)
    finally:
        if old_policy is None:
            del os.environ["FLASHGMM_POP_POLICY"]
        else:
            os.environ["FLASHGMM_POP_POLICY"] = old_policy

    def decode_pass(mode, R):
        bs = [x.copy() for x in EncodeJob(streams, mode).run()]
        job = DecodeJob([(a, b, c, R) for (_, a, b, c) in streams], mode, ctx=dctx)
        outs = job.run(bs)
        assert all(torch.equal(o, s[0]) for o, s in zip(outs, streams))
        L.flashgmm_profile_reset()
        L.flashgmm_profile_enable(1)
        t0 = time.perf_counter()
        for _ in range(3):
            job.run(bs)
        torch.cuda.synchronize()
        call = (time.perf_counter() - t0) / 3
        L.flashgmm_profile_enable(0)
        rows_ms, rows_n, _ = _prof_read(L, 1)
        pop_ms, pop_n, _ = _prof_read(L, 4)
        nsym = 24 * 147456
        return {"call_ms": round(1e3 * call, 3), "rows_kernel_ms_per_call": round(rows_ms / 3, 3),
                "rows_gsym_s": round(nsym / (rows_ms / 3) / 1e6, 3), "call_ns_per_symbol_per_stream": round(1e9 * call / 147456, 1),
                "decode_msym_s": round(nsym / call / 1e6, 1)}
    for mode in (0, 1, 2):
        res["decode_pass"][f"mode{mode}"] = decode_pass(mode, R0)
    for R in (R0, 64, 128, 256, 1024):
        res["radius"][f"R{R}"] = decode_pass(0, R)
    # peaky mixtures (sigma in [0.11, 0.9], close means: what trained latents look like): most window boundaries are
    # saturated and the rows kernel skips their evaluation exactly
    broad = streams
    streams = [synth.gmm_torch(147456, seed=950 + i, device=dev, sigma_span=0.8, mean_scale=1.0) for i in range(24)]
    res["peaky"] = {f"mode{m}": decode_pass(m, max(int(s[0].abs().max().item()) for s in streams) + 2) for m in (0, 1, 2)}
    streams = broad
    res["note"] = ("decode_pass = one coding pass of 24 streams x 147456 symbols (device-resident); radius = the same pass "
                   "decoded with a larger search radius than the data needs: the windowed rows kernel evaluates 33 boundaries "
                   "per symbol whatever R is")
    return res


def likelihood_bench(torch, L, _lib, dev, iters=20):
    """north_star item 5: GaussianMixtureConditional.forward (training likelihood) fwd + bwd at the training batch
    (16 x 192 x 16 x 16, Readme.md:52) and at one ELIC image (1 x 320 x 32 x 48); algorithmic bytes 60 / 108 per element."""
    import flashgmm_b200 as fg
    peak, _src = measured_peak_gbs()
    res = {}
    for name, (B, M, H, W) in (("train_16x192x16x16", (16, 192, 16, 16)), ("elic_1x320x32x48", (1, 320, 32, 48)),
                               ("large_64x320x32x48", (64, 320, 32, 48))):
        g = torch.Generator(device=dev).manual_seed(7)
        x = (torch.randn(B, M, H, W, device=dev, generator=g) * 3).requires_grad_(True)
        sc = (torch.rand(B, 4 * M, H, W, device=dev, generator=g) * 3 + 0.05).requires_grad_(True)
        mu = torch.randn(B, 4 * M, H, W, device=dev, generator=g).requires_grad_(True)
        w = torch.softmax(torch.randn(B, 4, M, H, W, device=dev, generator=g), 1).reshape(B, 4 * M, H, W).requires_grad_(True)
        m = fg.GaussianMixtureConditional(K=4).to(dev).train()
        n = x.numel()
        def fwd():
            return m(x, sc, mu, w)
        for _ in range(3):
            y, lik = fwd()
            lik.log().sum().backward()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        tf, tbs = 0.0, []
        for _ in range(iters):
            y, lik = fwd()
            go = torch.ones_like(lik)
            torch.cuda.synchronize()
            e[1].record()
            lik.backward(go)
            e[2].record()
            torch.cuda.synchronize()
            tbs.append(e[1].elapsed_time(e[2]))
        # forward alone, back to back
        torch.cuda.synchronize()
        with torch.no_grad():
            e[0].record()
            for _ in range(iters):
                fwd()
            e[1].record()
            torch.cuda.synchronize()
            tf = e[0].elapsed_time(e[1]) / iters
        tb = statistics.median(tbs)
        res[name] = {"elements": n, "fwd_ms": round(tf, 4), "fwd_gb_s": round(60 * n / tf / 1e6, 1), "fwd_hbm_frac": round(60 * n / tf / 1e6 / peak, 4),
                     "bwd_ms": round(tb, 4), "bwd_gb_s": round(108 * n / tb / 1e6, 1), "bwd_hbm_frac": round(108 * n / tb / 1e6 / peak, 4)}
    res["note"] = ("fwd includes the noise draw (one torch kernel) and the fused likelihood kernel; bwd is the fused backward kernel "
                   "plus autograd bookkeeping; the two small shapes are launch-latency-bound (a few microseconds of HBM time)")
    return res


def model_codec_bench(torch, dev, B=24, reps=3, lanes=1):
    """The MODEL-facing path: `latent_codec.compress / decompress` of the ELIC2022GMM architecture for a
    batch of B images through flashgmm_b200's pass codecs -- hyper branch (h_a, z table coder, h_s), 5 channel groups x 2
    checkerboard passes, every network step replayed from a CUDA graph per image, scan / gather / scatter kernels per pass,
    ONE batched encode call for all 10 x B streams, ten batched pass-sequential decode calls. The architecture comes from
    tests/gmm_models.py (random-init; the reference's Python package does not exist on the GPU box); random-init networks
    clamp every sigma to 0.11 (~100 % bypass), so each pass's parameters are REPLACED by the synthetic mixture of config 4
    after the networks have run, and y is drawn from it -- the coder sees plausible streams, the orchestration does all its work."""
    sys.path.insert(0, str(ROOT / "tests"))
    import gmm_models as gm
    import flashgmm_b200 as fg
    torch.manual_seed(0)
    net = gm.Elic2022GMM(N=192, M=320, K=4).eval()
    net.update()
    net = net.to(dev)
    ycodec = net.latent_codec.latent_codec["y"]
    groups, (H, W), K = list(ycodec.groups), LATENT_HW, 4
    g = torch.Generator(device=dev).manual_seed(11)
    tables, parts = {}, []
    for i, C in enumerate(groups):
        halves = []
        for p in range(2):
            shp = (B, K, C, H, W // 2)
            mu = 3.0 * torch.randn(shp, device=dev, generator=g)
            sc = 0.11 + 4.0 * torch.rand(shp, device=dev, generator=g)
            w = torch.softmax(torch.randn(shp, device=dev, generator=g), dim=1)
            u = torch.rand((B, 1, C, H, W // 2), device=dev, generator=g)
            comp = (u > torch.cumsum(w, 1)).sum(1, keepdim=True).clamp_(0, K - 1)
            y = torch.gather(mu, 1, comp).squeeze(1) + torch.gather(sc, 1, comp).squeeze(1) * torch.randn((B, C, H, W // 2), device=dev, generator=g)
            tables[(i, p)] = tuple(t.reshape(B, K * C, H, W // 2).contiguous() for t in (sc, mu, w))
            halves.append(y)
        parts.append(ycodec.latent_codec[f"y{i}"]._join(halves[0], halves[1]))
    y = torch.cat(parts, dim=1).contiguous()
    calls = {}
    codecs = [ycodec.latent_codec[f"y{i}"].latent_codec["y"] for i in range(len(groups))]
    for i, codec in enumerate(codecs):
        def fake(ctx_params, codec=codec, i=i, real=codec._mixture):
            real(ctx_params)                                  # the entropy-parameter head and the softmax still run
            key = (threading.get_ident(), i)                  # passes alternate per codec and per calling thread
            n = calls.get(key, 0)
            calls[key] = n + 1
            return (*tables[(i, n % 2)], None)
        codec._mixture = fake
    lc = net.latent_codec
    if lanes > 1:
        return _model_codec_lanes(torch, dev, lc, y, lanes, reps)
    enc = lc.compress(y)
    dec = lc.decompress(enc["strings"], enc["shape"])
    if not torch.equal(dec["y_hat"], torch.round(y)):
        raise SystemExit("model codec round trip mismatch: refusing to report a number")
    te = td = 0.0
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        enc = lc.compress(y)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        lc.decompress(enc["strings"], enc["shape"])
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        te += (t1 - t0) / reps
        td += (t2 - t1) / reps
    nsym = y.numel()
    payload = sum(len(s[0]) for e in enc["strings"][:-1] for s in ([e] if isinstance(e, tuple) else e))
    from flashgmm_b200.graphs import default_runner
    r = default_runner()
    return {"model": "Elic2022GMM(N=192, M=320, K=4) latent codec, random-init, synthetic-mixture parameters", "images": B,
            "symbols": nsym, "compress_ms": round(1e3 * te, 2), "decompress_ms": round(1e3 * td, 2),
            "compress_ms_per_image": round(1e3 * te / B, 3), "decompress_ms_per_image": round(1e3 * td / B, 3),
            "value_msym_s": round(2 * nsym / (te + td) / 1e6, 1), "bits_per_symbol": round(8 * payload / nsym, 3),
            "graph_captures": r.captures, "graph_replays": r.replays, "net_batching": r.mode,
            "note": "host wall clock of latent_codec.compress / decompress incl. h_a / h_s, z coder, context and parameter "
                    "networks (CUDA-graph replays per image), glue kernels, batched coder; one batch in flight"}


def _model_codec_lanes(torch, dev, lc, y, n_lanes, reps):
    """The same model path with `n_lanes` batches in flight: one host thread per flashgmm_b200.Lane (own library ctx, own
    CUDA graphs, own stream), every thread compresses and decompresses its own batch of B images `reps` times; the model
    is shared. Wall clock over all threads, after each lane has run once on its own (graph captures)."""
    import flashgmm_b200 as fg
    lanes = [fg.Lane(device=dev.index or 0) for _ in range(n_lanes)]
    want = torch.round(y)
    for lane in lanes:                                        # captures, buffers; also the correctness gate per lane
        with lane:
            enc = lc.compress(y)
            dec = lc.decompress(enc["strings"], enc["shape"])
            ok = torch.equal(dec["y_hat"], want)
        if not ok:
            raise SystemExit("model codec (lanes) round trip mismatch: refusing to report a number")
    torch.cuda.synchronize()
    errors, t_c, t_d = [], [0.0] * n_lanes, [0.0] * n_lanes
    gate = threading.Barrier(n_lanes + 1)

    def work(k):
        try:
            with lanes[k]:
                gate.wait()
                for _ in range(reps):
                    t0 = time.perf_counter()
                    enc = lc.compress(y)
                    torch.cuda.current_stream().synchronize()
                    t1 = time.perf_counter()
                    dec = lc.decompress(enc["strings"], enc["shape"])
                    torch.cuda.current_stream().synchronize()
                    t_c[k] += (t1 - t0) / reps
                    t_d[k] += (time.perf_counter() - t1) / reps
                if not torch.equal(dec["y_hat"], want):
                    errors.append(f"lane {k}: round trip mismatch")
        except Exception as e:                                # noqa: BLE001
            errors.append(f"lane {k}: {type(e).__name__}: {e}")
            try:
                gate.abort()
            except Exception:
                pass

    # a thread that comes back from a library call (GIL released) has to get the GIL back from a thread that is busy
    # issuing small torch ops and gives it up only every `switch interval` (5 ms by default): dozens of such hand-overs
    # per coding call
    old_interval = sys.getswitchinterval()
    sys.setswitchinterval(float(os.environ.get("FLASHGMM_BENCH_SWITCH_INTERVAL", "2e-4")))
    threads = [threading.Thread(target=work, args=(k,)) for k in range(n_lanes)]
    for t in threads:
        t.start()
    gate.wait()
    t0 = time.perf_counter()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    interval = sys.getswitchinterval()
    sys.setswitchinterval(old_interval)
    if errors:
        raise SystemExit("model codec (lanes): " + "; ".join(errors))
    B, nsym = y.size(0), y.numel()
    return {"lanes": n_lanes, "images_per_batch": B, "batches": n_lanes * reps,
            "ms_per_batch": round(1e3 * wall / (n_lanes * reps), 2),
            "ms_per_image": round(1e3 * wall / (n_lanes * reps * B), 3),
            "value_msym_s": round(2 * nsym * n_lanes * reps / wall / 1e6, 1),
            "compress_ms_in_lane": round(1e3 * sum(t_c) / n_lanes, 2), "decompress_ms_in_lane": round(1e3 * sum(t_d) / n_lanes, 2),
            "python_switch_interval_s": interval,
            "note": "compress + decompress of every batch through latent_codec.*, one host thread per Lane; wall clock over all lanes"}


def reference_full_call(torch, dev, repeat=3):
    """BASELINE.md 3 'Calls timed': the reference's GaussianMixtureConditional.compress / decompress AS THE REFERENCE RUNS
    THEM with the model on the GPU -- its tensor ops on the device, its 2 .item() + 1 .tolist() syncs and 4 D2H copies per
    pass (entropy_models.py:833-867), its compiled CPU coder (oracle/_ref, loaded in this process), its CPU scatter and the
    H2D of y_hat (:872-910, checkerboard.py:324) -- on one ELIC image (10 passes), parameters already on the GPU."""
    from oracle import ref
    if not ref.available():
        return None
    from flashgmm_b200 import synth
    ans_ref = ref.load_ans()
    enc, dec = ans_ref.RansEncoder(), ans_ref.RansDecoder()
    K, (H, W) = 4, LATENT_HW
    passes = []
    for i, g in enumerate(ELIC_GROUPS):
        for half in range(2):
            n = g * H * W // 2
            sym, sg, mu, w = synth.gmm_torch(n, seed=7000 + 2 * i + half, device=dev)
            y = sym.float().reshape(1, g, H, W // 2)
            to4 = lambda p: p.t().reshape(K, g, H, W // 2).reshape(1, K * g, H, W // 2).contiguous()  # noqa: E731
            passes.append((y, to4(sg), to4(mu), to4(w)))

    def views(t, nz):
        return t.reshape(1, K, t.size(1) // K, -1)[:, :, nz].permute(1, 0, 2, 3).reshape(K, -1).permute(1, 0)

    def compress(y, sc, mu, w):
        abs_max = max(y.max().abs().int().item(), y.min().abs().int().item()) + 1
        yq = torch.round(y)
        bitmap = torch.where(torch.sum(torch.abs(yq), (3, 2)).squeeze(0) == 0, 0, 1)
        nz = torch.nonzero(bitmap).flatten().tolist()
        sym = yq[:, nz].reshape(-1).int()
        b = enc.encode_with_indexes_gmm(sym.to("cpu"), views(sc, nz).clamp(0.11, 256).to("cpu"), views(mu, nz).to("cpu"),
                                        views(w, nz).to("cpu"), abs_max + 1)
        return b, abs_max, bitmap

    def decompress(b, abs_max, bitmap, sc, mu, w):
        nz = torch.nonzero(bitmap).flatten().tolist()
        sym = dec.decode_with_indexes_gmm(b, views(sc, nz).clamp(0.11, 256).to("cpu"), views(mu, nz).to("cpu"),
                                          views(w, nz).to("cpu"), abs_max + 1)
        y_hat = torch.zeros(1, bitmap.numel(), sc.size(2), sc.size(3))
        y_hat[:, nz] = sym.reshape(1, len(nz), sc.size(2), sc.size(3)).float()
        return y_hat.to(dev)

    nsym = sum(p[0].numel() for p in passes)
    te = td = 0.0
    for r in range(repeat + 1):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        coded = [compress(*p) for p in passes]
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        outs = [decompress(*c, *p[1:]) for c, p in zip(coded, passes)]
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        if r:
            te += t1 - t0
            td += t2 - t1
    assert all(torch.equal(o, p[0]) for o, p in zip(outs, passes))
    te /= repeat
    td /= repeat
    return {"value": round(2 * nsym / (te + td) / 1e6, 3), "unit": "Msymbols/s", "cores": 1, "kind": "reference",
            "sample": f"one ELIC image (10 passes, {nsym} symbols), x{repeat}: the reference's compress/decompress call sequence "
                      "(its torch ops on the GPU, 3 syncs + 4 D2H copies per pass, compiled CPU coder oracle/_ref, CPU scatter, H2D of y_hat)",
            "compress_ms_per_image": round(1e3 * te, 2), "decompress_ms_per_image": round(1e3 * td, 2),
            "encode_msym_s": round(nsym / te / 1e6, 3), "decode_msym_s": round(nsym / td / 1e6, 3)}


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
    """The reference's own CPU implementation on the same batch, one worker process per host
    core (the reference is single-threaded and holds the GIL; independent streams are the only
    parallelism it offers). Streams are resident in the workers before the timed region."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import threading

    import torch

    from flashgmm_b200 import synth
    from oracle import port, ref

    B = args.images_per_step
    sizes = elic_stream_sizes()
    nsym = sum(sizes)
    mode = args.mode
    images = [_np_streams(make_image(torch, synth, i, "cpu")) for i in range(B)]
    streams = [st for img in images for st in img]
    radii = [synth.search_radius(st[0]) for st in streams]
    cores = os.cpu_count() or 1
    P = max(1, min(cores, len(streams)))
    use_ref = ref.available()
    kind = "reference" if use_ref else "port"
    # longest-processing-time assignment of streams to workers
    order = sorted(range(len(streams)), key=lambda i: -streams[i][0].size)
    load, assign = [0] * P, [[] for _ in range(P)]
    for i in order:
        k = load.index(min(load))
        assign[k].append(i)
        load[k] += streams[i][0].size
    workers = [ref.RefWorker(mode) if use_ref else None for _ in range(P)]
    if use_ref:
        for k in range(P):
            for i in assign[k]:
                sym, sg, mu, w = streams[i]
                workers[k].call("put", key=i, symbols=sym, scales=sg, means=mu, weights=w, max_bs_value=radii[i])

    def work(k, ok):
        if use_ref:
            ok[k] = workers[k].call("roundtrip", keys=assign[k])[0]
        else:
            good = True
            for i in assign[k]:
                sym, sg, mu, w = streams[i]
                b = port.gmm_encode(sym, sg, mu, w, mode)
                good = good and bool((port.gmm_decode(b, sg, mu, w, radii[i], mode) == sym).all())
            ok[k] = good

    def step(_i):
        ok = [False] * P
        th = [threading.Thread(target=work, args=(k, ok)) for k in range(P)]
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
    val = round(2 * nsym * B * args.steps / secs / 1e6, 3)
    emit({
        "impl": "reference", "metric": "GMM rANS enc+dec Msymbols/s (single-stream, reference-compatible bytes)",
        "value": val, "unit": "Msymbols/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(1e3 * secs / args.steps, 3), "ms_per_image": round(1e3 * secs / args.steps / B, 3),
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": common_config(args), "details": {"host_cores": cores},
        "cpu_baseline": {"value": val, "unit": "Msymbols/s", "cores": P, "kind": kind,
                         "sample": f"whole batch per step; {P} worker processes (USE_SIMD=1), streams dealt "
                                   "longest-first, tensors resident in the workers"},
        "e2e": {"value": val, "unit": "Msymbols/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-microbench", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="development only: skip the end-to-end arm (the line then has no e2e value)")
    ap.add_argument("--images-per-step", type=int, default=24,
                    help="images per step and per GPU (BASELINE.json configs[4] works on 24 images)")
    ap.add_argument("--mode", type=int, default=0, choices=[0, 1, 2], help="APPROX_MODE (reference code numbering)")
    ap.add_argument("--workload", default="images24", choices=["images24", "sharded49"],
                    help="images24: --images-per-step images per GPU and step (weak scaling, the headline); sharded49: "
                         "BASELINE configs[4] as written, 24 images + a 3840x2160 image as 25 tiles = 49 items dealt over "
                         "the ranks (strong scaling)")
    ap.add_argument("--decode-order", default="pass-sequential", choices=["pass-sequential", "pass-parallel"],
                    help="pass-sequential (default, what a real decoder can do): one decode call per coding pass, "
                         "issued after the previous pass has returned; pass-parallel: every stream in one call")
    ap.add_argument("--pipeline-depth", type=int, default=0,
                    help="independent batches in flight (host threads, one ctx each); 1 = strictly one after the other; "
                         "0 (default) = 6")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
