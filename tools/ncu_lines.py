#!/usr/bin/env python
"""Per-source-line totals (warp instructions executed, stall samples) of one kernel from an ncu report captured with
--import-source on: python tools/ncu_lines.py report.ncu-rep kernel_regex [min_share]"""
import collections
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--kernel-name",
                      "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[h]
h = -1
ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
inst, samp, text = collections.Counter(), collections.Counter(), {}
cur, fname = None, ""
for r in rows[h + 1:]:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if len(r) <= ie or r[0] == "Line No":
        continue
    if r[0].strip():  # a CUDA source line row; SASS rows below it have an empty line number
        cur = (fname, int(r[0]))
        text[cur] = r[1]
        continue
    if cur is None:
        continue
    try:
        inst[cur] += int(r[ie] or 0)
        samp[cur] += int(r[ss] or 0)
    except ValueError:
        pass
ti, ts = sum(inst.values()), sum(samp.values())
print("total warp instr %d, samples %d" % (ti, ts))
for ln in sorted(text):
    if inst[ln] >= min_share * ti or samp[ln] >= min_share * ts:
        print("%s:%4d %6.2f%% instr %6.2f%% samp | %s" % (ln[0][:12], ln[1], 100.0 * inst[ln] / max(ti, 1), 100.0 * samp[ln] / max(ts, 1), text[ln][:110]))
