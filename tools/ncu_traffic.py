#!/usr/bin/env python
"""Writes one workload's entry of profiles/traffic.json from an `ncu --set full` capture of bench.py:

    python tools/ncu_traffic.py <workload> <kernel name> <rep.ncu-rep> <regex over mangled names of the captured instantiations> [first launch index]

The three launches starting at the index (default 0) must be stages 1, 2, 3 of one time step of the dominant
kernel.  Recorded per launch: DRAM bytes read + written, executed warp instructions, FP64 pipe activity and
duration; from the source page of the first captured launch the FP64 share of the executed instructions.  The
entry also records the hash of the kernel family's SASS in the library as built NOW (the capture must come from
this build): bench.py uses the entry only while the built library still has that hash."""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from sass_hashes import family_hash, hashes  # noqa: E402


def main():
    workload, family, rep, mangled_re = sys.argv[1:5]
    first = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    sel = [r for r in data if family in r[ix["Kernel Name"]]][first:first + 3]
    assert len(sel) == 3, "need three launches of %s in %s" % (family, rep)

    def col(name, scale=1.0):
        return [float(r[ix[name]].replace(",", "")) * scale for r in sel]

    def to_bytes(name):
        mul = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[units[ix[name]].lower()]
        return col(name, mul)
    dram = [a + b for a, b in zip(to_bytes("dram__bytes_read.sum"), to_bytes("dram__bytes_write.sum"))]
    du = units[ix["gpu__time_duration.sum"]].lower()
    dur = col("gpu__time_duration.sum", {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(du, 1e-9))
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    h2, ops, tot = None, collections.Counter(), 0
    for row in csv.reader(io.StringIO(src)):
        if "Instructions Executed" in row:
            if tot:
                break
            h2 = {h: i for i, h in enumerate(row)}
            continue
        if h2 is None or len(row) <= h2["Instructions Executed"] or not row[h2["Instructions Executed"]].isdigit():
            continue
        t = row[h2["Source"]].split()
        if not t:
            continue
        op = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
        n = int(row[h2["Instructions Executed"]])
        ops[op.split(".")[0]] += n
        tot += n
    fp64 = sum(v for k, v in ops.items() if k in ("DFMA", "DMUL", "DADD", "DSETP"))
    entry = {
        "dram_bytes_per_launch_mean": sum(dram) / 3.0,
        "dram_bytes_per_launch_stage123": dram,
        "warp_inst_per_launch_stage123": col("smsp__inst_executed.sum"),
        "fp64_share_of_inst_stage2": fp64 / tot if tot else None,
        "fp64_pipe_active_pct_stage123": col("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
        "ncu_duration_s_stage123": dur,
        "kernel_family": mangled_re,
        "sass_md5": family_hash(hashes(), mangled_re),
        "source": "%s (ncu --set full --clock-control none; 3 consecutive launches = stages 1,2,3 of one time step; FP64 share from the source page of the first captured launch)" % os.path.relpath(rep, ROOT),
    }
    p = os.path.join(ROOT, "profiles", "traffic.json")
    d = json.load(open(p)) if os.path.exists(p) else {}
    d[workload] = entry
    json.dump(d, open(p, "w"), indent=1)
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()
