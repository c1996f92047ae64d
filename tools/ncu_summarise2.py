"""gpurun_out/r02_*.ncu-rep + r02_launches.csv -> profiles/r02_ncu_summary.json (per captured launch: duration, DRAM bytes,
issue rate, FMA / ALU / XU pipe utilisation, warps active, instructions) and profiles/r02_ncu_launch_summary.csv.
usage: python tools/ncu_summarise2.py"""
import collections, csv, io, json, os, subprocess
O, P = "gpurun_out", "profiles"
KEYS = {
    "gpu__time_duration.sum": "duration",
    "launch__grid_size": "grid", "launch__block_size": "block", "launch__registers_per_thread": "regs",
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
    "smsp__inst_executed.sum": "warp_insts",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "smsp__issue_active.avg.per_cycle_active": "issue_active_per_cycle",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "pipe_fma_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "pipe_xu_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "pipe_alu_pct",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "pipe_fma_inst_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "pipe_lsu_pct",
    "smsp__average_warp_latency_per_inst_issued.ratio": "warp_latency_per_inst",
}
out = {}
for name in ("sf", "rows", "pop", "push", "lik", "il", "glue", "rows_final", "lik_final"):
    # *_final: recaptured after the register-resident rewrite of the windowed rows kernel (all three modes) and with the
    # EntropyBottleneck likelihood kernel (tools/gpu_final.sh writes them as r02b_*)
    rep = f"{O}/r02b_{name[:-6]}.ncu-rep" if name.endswith("_final") else f"{O}/r02_{name}.ncu-rep"
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(raw)))
    hdr, units = rd[0], rd[1]
    rows = []
    for vals in rd[2:]:
        if len(vals) != len(hdr):
            continue
        d = {"kernel": vals[hdr.index("Kernel Name")].split("(")[0][:90]}
        for k, short in KEYS.items():
            if k in hdr:
                i = hdr.index(k)
                try:
                    d[short] = float(vals[i].replace(",", ""))
                    if short in ("duration", "dram_read", "dram_write"):
                        d[short + "_unit"] = units[i]
                except ValueError:
                    d[short] = vals[i]
        rows.append(d)
    out[name] = rows
json.dump(out, open(f"{P}/r02_ncu_summary.json", "w"), indent=1)
for name, rows in out.items():
    for d in rows:
        print(name, json.dumps(d))
path = f"{O}/r02_launches.csv"
if os.path.exists(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for r in csv.DictReader(io.StringIO("".join(lines))):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        nm = r["Kernel Name"].split("(")[0]
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        us = v / 1e3 if u in ("ns", "nsecond") else v if u in ("us", "usecond") else v * 1e3
        a = agg.setdefault(nm, [0, 0.0, 0.0])
        a[0] += 1; a[1] += us; a[2] = max(a[2], us)
    tot = sum(a[1] for a in agg.values())
    with open(f"{P}/r02_ncu_launch_summary.csv", "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gmm_|eb_ of: python bench.py --steps 2 --warmup 3 --no-microbench --no-cpu-baseline --pipeline-depth 1\n")
        f.write("# (serialised, cold-cache launch times: the SHARES are comparable with bench.py's live CUDA-event shares, the absolutes are not)\n")
        f.write("kernel,launches,total_us,avg_us,max_us,share\n")
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k},{a[0]},{a[1]:.1f},{a[1] / a[0]:.2f},{a[2]:.1f},{a[1] / tot:.4f}\n")
    print(open(f"{P}/r02_ncu_launch_summary.csv").read())
