#!/usr/bin/env python
"""Condense ncu --set full reports into profiles/: a markdown table per kernel (duration, DRAM bytes, issue utilisation,
pipe utilisation, top stall reasons) and traffic.json (DRAM read + write bytes per launch, what bench.py reports as
roofline.traffic). usage: ncu_summary.py <out_prefix> <report.ncu-rep> [...]"""
import csv
import json
import os
import subprocess
import sys

out_prefix, reps = sys.argv[1], sys.argv[2:]
KEYS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram rd MB"), ("dram__bytes_write.sum", "dram wr MB"),
        ("smsp__inst_executed.sum", "warp instr"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"), ("launch__registers_per_thread", "regs"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu %"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "lsu data pipe %"),
        ("l1tex__t_sector_hit_rate.pct", "L1 hit %"), ("lts__t_sector_hit_rate.pct", "L2 hit %")]
rows_out, traffic = [], {}
for rep in reps:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        name = d["Kernel Name"].split("(")[0].split("::")[-1]
        rec = {"kernel": name, "grid": d.get("Grid Size", ""), "block": d.get("Block Size", "")}
        for k, label in KEYS:
            v = d.get(k, "")
            try:
                v = float(v)
                if "bytes" in k:
                    scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u.get(k, "Mbyte"), 1.0)
                    v *= scale
                if k == "gpu__time_duration.sum":
                    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u.get(k, "us"), 1.0)
            except ValueError:
                pass
            rec[label] = v
        stalls = sorted(((float(v), k.split("issue_stalled_")[1].split("_per_")[0]) for k, v in d.items()
                         if "average_warps_issue_stalled" in k and k.endswith(".ratio") and v not in ("", "n/a")), reverse=True)[:3]
        rec["top stalls (warps per issue)"] = ", ".join(f"{n} {v:.2f}" for v, n in stalls)
        rows_out.append(rec)
        # DRAM bytes of ALL launches of the kernel in the captured step (the capture holds exactly one step: the seven pyramid
        # launches add up), so that the figure describes the same thing as bench.py's algorithmic bytes per step
        traffic[name] = traffic.get(name, 0) + int((rec["dram rd MB"] + rec["dram wr MB"]) * 1e6)
cols = ["kernel", "grid", "block"] + [l for _, l in KEYS] + ["top stalls (warps per issue)"]
with open(out_prefix + "_ncu_summary.md", "w") as f:
    f.write("ncu --set full --clock-control none, one step of 32 stereo pairs (64 frames of 1280x720), tools/profile_step.py\n\n")
    f.write("| " + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
    for r in rows_out:
        f.write("| " + " | ".join((f"{r[c]:.1f}" if isinstance(r[c], float) else str(r[c])) for c in cols) + " |\n")
tj = os.path.join(os.path.dirname(out_prefix), "traffic.json")
old = json.load(open(tj)) if os.path.exists(tj) else {}
old.update(traffic)
old["_source"] = "dram__bytes_read.sum + dram__bytes_write.sum summed over every launch of the kernel in ONE step, " + os.path.basename(out_prefix) + "_ncu_summary.md"
json.dump(old, open(tj, "w"), indent=1, sort_keys=True)
print(open(out_prefix + "_ncu_summary.md").read())
