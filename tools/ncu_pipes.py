#!/usr/bin/env python
"""Pipe utilisation of the matcher / kNN kernels from ncu --set full reports: one markdown row per kernel (the median-duration
launch of those captured), with the integer pipes the Hamming work runs on (ALU: LOP3 / IADD / compare, XU: POPC) next to issue
slots, occupancy and the stall reasons. usage: ncu_pipes.py <out.md> <title> <report.ncu-rep> [...]"""
import csv
import subprocess
import sys

out, title, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
KEYS = [("gpu__time_duration.sum", "us"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu (popc) %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lsu data pipe %"),
        ("smsp__inst_executed.sum", "warp instr"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("launch__local_mem_size_per_thread" if False else "smsp__inst_executed_op_local_ld.sum", "local ld instr")]
by_kernel = {}
for rep in reps:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    u = dict(zip(hdr, units))
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d["Kernel Name"].split("(")[0].split("::")[-1]
        rec = {"kernel": name, "grid": d.get("Grid Size", ""), "block": d.get("Block Size", "")}
        for k, label in KEYS:
            v = d.get(k, "")
            try:
                v = float(v)
                if k == "gpu__time_duration.sum":
                    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u.get(k, "us"), 1.0)
                if "bytes" in k:
                    v = "%.2f MB" % (v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u.get(k, "Mbyte"), 1.0))
            except ValueError:
                pass
            rec[label] = v
        stalls = sorted(((float(v), k.split("issue_stalled_")[1].split("_per_")[0]) for k, v in d.items()
                         if "average_warps_issue_stalled" in k and k.endswith(".ratio") and v not in ("", "n/a")), reverse=True)[:3]
        rec["top stalls (warps per issue)"] = ", ".join(f"{n} {v:.2f}" for v, n in stalls)
        by_kernel.setdefault(name, []).append(rec)
cols = ["kernel", "launches captured", "grid", "block"] + [l for _, l in KEYS] + ["top stalls (warps per issue)"]
with open(out, "w") as f:
    f.write(title + "\n\n| " + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
    for name, recs in by_kernel.items():
        recs.sort(key=lambda r: r["us"] if isinstance(r["us"], float) else 0.0)
        r = dict(recs[len(recs) // 2])
        r["launches captured"] = len(recs)
        f.write("| " + " | ".join((f"{r[c]:.1f}" if isinstance(r[c], float) else str(r[c])) for c in cols) + " |\n")
print(open(out).read())
