"""Summarise a tools/gpu_round.sh visit into profiles/: launch list -> per-kernel totals,
each `ncu --set full` report -> raw CSV + a JSON of the metrics DESIGN.md / bench.py quote.
usage: python tools/ncu_summarise.py TAG   (reads gpurun_out/TAG_*, writes profiles/r01_TAG_*)"""
import csv, io, json, subprocess, sys, collections, shutil, os
T = sys.argv[1]
O, P = "gpurun_out", "profiles"
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.per_cycle_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__average_warp_latency_per_inst_issued.ratio",
        "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "pcie__read_bytes.sum", "pcie__write_bytes.sum"]
for f in ("bench.json", "bench_ref.json"):
    if os.path.exists(f"{O}/{T}_{f}"):
        shutil.copy(f"{O}/{T}_{f}", f"{P}/r01_{T}_{f.replace('bench_ref', 'bench_reference').replace('bench.json', 'bench_b200.json')}")
# launch list
path = f"{O}/{T}_launches.csv"
if os.path.exists(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    agg = collections.OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"].split("(")[0]
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
        a = agg.setdefault(name, [0, 0.0, 0.0])
        a[0] += 1; a[1] += us; a[2] = max(a[2], us)
    tot = sum(a[1] for a in agg.values())
    with open(f"{P}/r01_{T}_ncu_launch_summary.csv", "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gmm_ of: python bench.py --no-microbench --no-cpu-baseline --steps 2 --warmup 3\n")
        f.write("# (serialised, cold-cache launch times: the SHARES are comparable with bench.py's live CUDA-event shares, the absolutes are not)\n")
        f.write("kernel,launches,total_us,avg_us,max_us,share\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{a[0]},{a[1]:.1f},{a[1] / a[0]:.2f},{a[2]:.1f},{a[1] / tot:.4f}\n")
    print(open(f"{P}/r01_{T}_ncu_launch_summary.csv").read())
summ = []
for name in ("pop", "rows", "sfb", "push", "sf"):
    rep = f"{O}/{T}_{name}.ncu-rep"
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open(f"{P}/r01_{T}_ncu_full_{name}.csv", "w").write(raw)
    rd = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rd[0], rd[1], rd[2]
    d = {"kernel": vals[hdr.index("Kernel Name")]}
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            d[k] = f"{vals[i]} {units[i]}".strip()
    summ.append(d)
json.dump(summ, open(f"{P}/r01_{T}_ncu_full_summary.json", "w"), indent=1)
print(json.dumps(summ, indent=1))
