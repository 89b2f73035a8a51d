#!/usr/bin/env python
"""Kernel shares of the ncu launch list (gpu__time_duration.sum per launch) beside the stage shares bench.py reports.
usage: launch_shares.py <launches.csv> <bench.json> > profiles/rN_launch_shares.md"""
import collections
import csv
import json
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
bench = json.loads(open(sys.argv[2]).read())
t = collections.defaultdict(float)
n = collections.Counter()
for r in rows:
    name = r[4].split("(")[0].split("::")[-1].replace("unnamed>", "").strip(":> ")
    t[name] += float(r[14]) / 1e3
    n[name] += 1
tot = sum(t.values())
stage_of = {"k_resize_walk": "pyramid", "k_pad_borders": "blur", "k_gauss7": "blur", "k_gauss7_tma": "blur", "k_copy2d": "pyramid", "k_fast_cells": "fast", "k_quadtree": "quadtree",
            "k_offsets": "describe", "k_describe": "describe", "k_stereo_prep": "describe", "k_stereo_rows": "describe",
            "k_stereo_match": "describe", "k_stereo_median": "describe"}
st = collections.defaultdict(float)
for k, v in t.items():
    st[stage_of.get(k, "other")] += v
sm = bench["roofline"]["stage_ms"]
stot = sum(sm.values())
print("ncu launch list (`--metrics gpu__time_duration.sum --clock-control none`, cold caches, serialised launches) of")
print("`python bench.py --steps 2 --warmup 1`, first %d launches, against the stage times of the un-profiled bench run.\n" % len(rows))
print("| kernel | launches | total us | share |\n|---|---|---|---|")
for k, v in sorted(t.items(), key=lambda kv: -kv[1]):
    print(f"| {k} | {n[k]} | {v:.1f} | {100 * v / tot:.1f} % |")
print("\n| stage | share under ncu | share in bench.py stage_ms |\n|---|---|---|")
for k in ("pyramid", "fast", "quadtree", "blur", "describe"):
    print(f"| {k} | {100 * st[k] / tot:.1f} % | {100 * sm[k] / stot:.1f} % |")
if st.get("other"):
    print(f"| other | {100 * st['other'] / tot:.1f} % | - |")
