#!/usr/bin/env python
"""Summarises an .ncu-rep (raw + source pages) into the numbers DESIGN.md/profiles quote.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [launch_index]"""
import collections, csv, io, re, subprocess, sys

rep = sys.argv[1]
li = int(sys.argv[2]) if len(sys.argv) > 2 else 1
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
r = data[min(li, len(data) - 1)]
print("kernel:", r[ix["Kernel Name"]][:90], "grid", r[ix["Grid Size"]], "block", r[ix["Block Size"]])
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__cycles_active.avg", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
for w in want:
    if w in ix:
        print("  %-70s %s %s" % (w, r[ix[w]], units[ix[w]]))
st = []
for h in hdr:
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
        try:
            st.append((float(r[ix[h]]), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
        except ValueError:
            pass
print("stalls (warps per issue):", ", ".join("%s %.2f" % (n, v) for v, n in sorted(st, reverse=True)[:8]))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h2 = None
ops = collections.Counter(); tot = 0
for row in rows:
    if "Instructions Executed" in row:
        h2 = {h: i for i, h in enumerate(row)}
        if tot: break          # first kernel only
        continue
    if h2 is None or len(row) <= h2["Instructions Executed"] or not row[h2["Instructions Executed"]].isdigit():
        continue
    n = int(row[h2["Instructions Executed"]])
    t = row[h2["Source"]].split()
    if not t: continue
    op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    parts = op.split(".")
    key = parts[0] + ("." + parts[1] if parts[0] in ("IMAD", "LDS", "STS", "LDG", "STG") and len(parts) > 1 else "")
    ops[key] += n; tot += n
print("executed warp-instructions (first kernel in report): %d" % tot)
print("  " + ", ".join("%s %.1f%%" % (k, 100.0 * v / tot) for k, v in ops.most_common(16)))
fp64 = sum(v for k, v in ops.items() if k in ("DFMA", "DMUL", "DADD", "DSETP"))
print("  FP64-pipe share of instructions: %.1f%%" % (100.0 * fp64 / tot))
