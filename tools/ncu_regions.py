"""Summarise an ncu report's source page by program region (warp-role code is contiguous): samples, instructions, top stalls.
    python tools/ncu_regions.py gpurun_out/x.ncu-rep [bucket]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 150
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, r = rows[0], rows[2]
for k in ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct",
          "gpu__dram_throughput.avg.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct", "smsp__inst_executed.sum",
          "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct", "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "lts__t_bytes.sum ", "l1tex__m_xbar2l1tex_read_bytes.sum "]:
    for h, v in zip(hdr, r):
        if h.startswith(k) and "per_second" not in h:
            print(h, v)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]; idx = {h: i for i, h in enumerate(hdr)}
b = [r for r in rows[2:] if len(r) > 10 and r[0].startswith("0x")]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
acc = collections.OrderedDict()
for i, r in enumerate(b):
    n = int(r[idx["Instructions Executed"]]); s = int(r[idx["# Samples"]])
    a = acc.setdefault(i // bucket, [0, 0, collections.Counter(), collections.Counter()])
    a[0] += s; a[1] += n; a[2][n] += 1
    for st in stalls:
        a[3][st] += int(r[idx[st]])
tot = sum(a[0] for a in acc.values())
print("total samples", tot)
for k, a in acc.items():
    if a[0] > tot * 0.003:
        print(k * bucket, "samples", a[0], "inst", a[1], a[2].most_common(2), [(x[6:], y) for x, y in a[3].most_common(3)])
