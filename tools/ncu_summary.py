"""Summarise an .ncu-rep here (no GPU): key raw metrics + hot SASS regions of one kernel of the capture.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.txt] [kernel index, default 0]"""
import re, csv
import io
import subprocess
import sys

rep = sys.argv[1]
out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
KI = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hdr, units, vals = rows[0], rows[1], rows[2 + KI]
KEYS = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum.per_cycle_elapsed",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum"]
for h, u, v in zip(hdr, units, vals):
    if h in KEYS or h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        print("%-82s %-16s %s" % (h, u, v), file=out)
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]      # one block per captured launch
name = vals[hdr.index("Kernel Name")]                    # the source page groups launches by kernel, the raw page by time
base = re.search(r"(\w+)\s*(<|\()", name.split("::")[-1]).group(1)
pick = next((j for j in range(len(starts) - 1) if re.search(r"\b%s\b" % base, rows[starts[j]][1])), KI)
rows = rows[starts[pick]:starts[pick + 1]]
hdr, data = rows[1], [r for r in rows[2:] if len(r) > 10]
ia, ie, it, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
tot = sum(int(r[ie]) for r in data)
ts = sum(int(r[isamp]) for r in data)
print("\nSASS instructions %d, warp-instructions executed %d, thread-instructions %d" %
      (len(data), tot, sum(int(r[it]) for r in data)), file=out)
runs = []
for k, r in enumerate(data):
    e, t, s = int(r[ie]), int(r[it]), int(r[isamp])
    if runs and abs(e - runs[-1]["e0"]) <= 0.03 * max(e, runs[-1]["e0"]):
        R = runs[-1]; R["n"] += 1; R["e"] += e; R["t"] += t; R["s"] += s; R["end"] = k
    else:
        runs.append(dict(start=k, end=k, n=1, e0=e, e=e, t=t, s=s))
print("hot regions (runs of equal execution count): first-last #instr Mexec/instr share% samples% lanes/inst opcodes", file=out)
for R in runs:
    if R["e"] / tot > 0.004:
        ops = {}
        for r in data[R["start"]:R["end"] + 1]:
            w = r[ia].split()
            op = (w[1] if w[0].startswith("@") else w[0]).split(".")[0]
            ops[op] = ops.get(op, 0) + 1
        print("%4d-%4d %4d %9.1f %6.2f %6.2f %5.1f  %s" % (R["start"], R["end"], R["n"], R["e0"] / 1e6, 100 * R["e"] / tot,
              100 * R["s"] / max(ts, 1), R["t"] / max(R["e"], 1),
              " ".join("%s:%d" % kv for kv in sorted(ops.items(), key=lambda x: -x[1])[:8])), file=out)
