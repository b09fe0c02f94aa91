"""Per-kernel figures of an .ncu-rep as JSON (read here, no GPU): what bench.py attaches to its
line as roofline.measured.  One entry per kernel name (the last captured launch of each).
usage: python tools/ncu_json.py gpurun_out/r2_frame.ncu-rep profiles/r2_ncu_frame.json "how it was captured" """
import csv
import io
import json
import re
import subprocess
import sys

rep, out_path = sys.argv[1], sys.argv[2]
source = sys.argv[3] if len(sys.argv) > 3 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]


def num(row, key, default=None):
    if key not in hdr:
        return default
    v = row[hdr.index(key)].replace(",", "")
    try:
        return float(v)
    except ValueError:
        return default


def scaled(row, key):
    """bytes / time columns come with a unit row (Kbyte, Mbyte, us, ms ...)"""
    v = num(row, key)
    if v is None:
        return None
    u = units[hdr.index(key)].lower()
    mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
            "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}.get(u, 1)
    return v * mult


kernels = {}
for row in rows[2:]:
    if len(row) < len(hdr):
        continue
    name = row[hdr.index("Kernel Name")]
    m = re.search(r"(k_\w+)", name)
    short = m.group(1) if m else name
    stalls = {}
    for h in hdr:
        if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
            v = num(row, h)
            if v is not None and v >= 0.3:
                stalls[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(v, 2)
    rd, wr = scaled(row, "dram__bytes_read.sum") or 0.0, scaled(row, "dram__bytes_write.sum") or 0.0
    kernels[short] = {
        "duration_ms_under_ncu": scaled(row, "gpu__time_duration.sum"),
        "grid": num(row, "launch__grid_size"), "block": num(row, "launch__block_size"),
        "registers": num(row, "launch__registers_per_thread"),
        "issue_active_pct": num(row, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "sm_throughput_pct": num(row, "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        "warps_active_pct": num(row, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        "lanes_per_inst": num(row, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        "warp_instructions": num(row, "smsp__inst_executed.sum"),
        "dram_bytes": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr,
        "dram_pct_of_peak": num(row, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "l1_hit_pct": num(row, "l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": num(row, "lts__t_sector_hit_rate.pct"),
        "local_loads": num(row, "smsp__sass_inst_executed_op_local_ld.sum"), "local_stores": num(row, "smsp__sass_inst_executed_op_local_st.sum"),
        "stalls_per_issue": stalls,
    }
json.dump({"source": source, "kernels": kernels}, open(out_path, "w"), indent=1)
print(json.dumps({k: {x: v[x] for x in ("duration_ms_under_ncu", "issue_active_pct", "lanes_per_inst", "dram_bytes")} for k, v in kernels.items()}, indent=1))
