"""Summarise Nsight Compute captures into profiles/ (text + traffic.json).

    python tools/ncu_summary.py full   gpurun_out/prof.ncu-rep      profiles/r01_k_fused_se_ws_full.txt
    python tools/ncu_summary.py launch gpurun_out/launches.csv      profiles/r01_bench_launches.txt
"""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum.per_second", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tma.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def full(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    lines = [f"# ncu --set full --clock-control none --import-source on : {os.path.basename(rep)}"]
    traffic = {}
    for vals in rows[2:]:
        d = dict(zip(hdr, vals)); u = dict(zip(hdr, units))
        name = d.get("Kernel Name", "?")
        lines.append(f"\n## {name}")
        for k in KEYS:
            if k in d:
                lines.append(f"{k:75s} {d[k]:>16s} {u[k]}")
        for k in hdr:
            if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio"):
                lines.append(f"{k:75s} {d[k]:>16s}")
        try:
            scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}
            t = float(d["dram__bytes_read.sum"]) * scale[u["dram__bytes_read.sum"]] + \
                float(d["dram__bytes_write.sum"]) * scale[u["dram__bytes_write.sum"]]
            traffic[name.split("(")[0].split("::")[-1]] = t
        except (KeyError, ValueError):
            pass
    open(out, "w").write("\n".join(lines) + "\n")
    tj = os.path.join(os.path.dirname(out), "traffic.json")
    old = json.load(open(tj)) if os.path.exists(tj) else {}
    # traffic.json is keyed by configuration, then kernel (bench.py reads [cfg][kernel]); NCU_CFG names the configuration
    old.setdefault(os.environ.get("NCU_CFG", "other"), {}).update(traffic)
    json.dump(old, open(tj, "w"), indent=1, sort_keys=True)
    print(open(out).read())


def launch(csvfile, out):
    rows = [r for r in csv.reader(open(csvfile)) if r and r[0].isdigit()]
    hdr = next(r for r in csv.reader(open(csvfile)) if r and r[0] == "ID")
    iK, iV, iU = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = defaultdict(lambda: [0, 0.0])
    for r in rows:
        v = float(r[iV].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[iU], 1e-6)
        k = r[iK].split("(")[0]
        agg[k][0] += 1; agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    lines = [f"# ncu --metrics gpu__time_duration.sum --clock-control none : {os.path.basename(csvfile)}",
             "# per-launch times are cold-cache and serialised: compare SHARES, not absolutes",
             f"{'kernel':60s} {'launches':>9s} {'total ms':>12s} {'avg ms':>10s} {'share':>8s}"]
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"{k[:60]:60s} {n:9d} {ms:12.3f} {ms / n:10.4f} {100 * ms / tot:7.2f}%")
    open(out, "w").write("\n".join(lines) + "\n")
    print(open(out).read())


if __name__ == "__main__":
    {"full": full, "launch": launch}[sys.argv[1]](sys.argv[2], sys.argv[3])
