"""Summarise ncu outputs brought back in gpurun_out/ into small tracked files under profiles/.
    python tools/ncu_summarize.py launches gpurun_out/launches.csv profiles/rNN_launches_summary.csv "<command>"
    python tools/ncu_summarize.py full gpurun_out/prof.ncu-rep profiles/rNN_ncu_full_summary.csv
"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max", "l1tex__data_bank_conflicts_pipe_lsu.sum",
    "smsp__inst_executed.sum",
]


def short(name):
    name = re.sub(r"\(.*$", "", name)
    return name.replace("void ", "").replace("ddimb::", "").strip()


def launches(src, dst, cmd):
    rows = [l for l in open(src) if l.startswith('"')]
    rd = csv.DictReader(io.StringIO("".join(rows)))
    agg = defaultdict(lambda: [0, 0.0])
    for r in rd:
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = short(r["Kernel Name"])
        agg[k][0] += 1
        agg[k][1] += float(r["Metric Value"].replace(",", "")) * 1e-6
    tot = sum(v[1] for v in agg.values())
    with open(dst, "w") as f:
        f.write(f"# ncu launch list summary: {cmd}\n# gpu__time_duration.sum, --clock-control none; per-kernel totals over {sum(v[0] for v in agg.values())} launches (cold-cache, serialised: compare shares)\n")
        f.write("kernel,launches,total_ms,share\n")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"\"{k}\",{v[0]},{v[1]:.3f},{v[1] / tot:.4f}\n")


def full(src, dst):
    out = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rd[0], rd[1], rd[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in KEEP if c in idx]
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none (raw page), one row per captured launch\n")
        f.write("id,kernel,grid,block," + ",".join(f"{c} [{units[idx[c]]}]" for c in cols) + "\n")
        for r in data:
            f.write(",".join([r[idx["ID"]], '"' + short(r[idx["Kernel Name"]]) + '"', '"' + r[idx["Grid Size"]] + '"', '"' + r[idx["Block Size"]] + '"'] + [r[idx[c]].replace(",", "") for c in cols]) + "\n")


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else "")
    else:
        full(sys.argv[2], sys.argv[3])
