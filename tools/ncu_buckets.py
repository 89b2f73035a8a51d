#!/usr/bin/env python
"""Summarise an ncu report: per-kernel headline metrics, and SASS instruction / stall-sample buckets for one kernel."""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else None)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
idx = [hdr.index(w) if w in hdr else None for w in want]
print(" | ".join(w.split(".")[0][-28:] for w in want))
for r in rows[2:]:
    print(" | ".join((r[i][-30:] if i is not None else "NA") for i in idx))
if kern:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    i0 = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    i1 = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    end = i1[1] - 2 if len(i1) > 1 else len(rows)
    hdr, body = rows[i0], [r for r in rows[i0 + 1:end] if len(r) > 6]
    ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
    tot = sum(int(r[ie]) for r in body)
    st = sum(int(r[ss]) for r in body)
    print("total warp instr", tot, "sass lines", len(body), "samples", st)
    step = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    for b in range(0, len(body), step):
        seg = body[b:b + step]
        s = sum(int(r[ie]) for r in seg)
        sm = sum(int(r[ss]) for r in seg)
        ops = " ".join(sorted(set((r[1].split()[1] if r[1].strip().startswith("@") else r[1].split()[0]) for r in seg)))[:100]
        print(f"{b:4d} {100 * s / max(tot, 1):5.1f}% instr {100 * sm / max(st, 1):5.1f}% samples  {ops}")
