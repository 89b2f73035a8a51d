#!/usr/bin/env python
"""Device time and achieved HBM GB/s of the pre-processing kernels on a batch of 64 colour 1280x720 frames (CUDA events on
the stream the kernels run on). Algorithmic bytes: gray 3+1 B/px, remap 1+1 B/px (+ the shared 8 B/px maps once),
imadjust 1 (histogram) + 1 + 1 B/px. usage: preproc_bench.py [frames=64] [reps=20]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lpslam_b200 import preproc  # noqa: E402,F401
from lpslam_b200._lib import check, lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
W, H = 1280, 720
L = lib()
g = torch.Generator(device="cuda").manual_seed(1)
color = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, device="cuda", generator=g)
gray = torch.empty((N, H, W), dtype=torch.uint8, device="cuda")
rect = torch.empty_like(gray)
adj = torch.empty_like(gray)
ys, xs = torch.meshgrid(torch.arange(H, device="cuda", dtype=torch.float32), torch.arange(W, device="cuda", dtype=torch.float32), indexing="ij")
r2 = ((xs - W / 2) / W) ** 2 + ((ys - H / 2) / W) ** 2
mx = (W / 2 + (xs - W / 2) * (1 + 0.2 * r2)).contiguous()
my = (H / 2 + (ys - H / 2) * (1 + 0.2 * r2)).contiguous()
scratch = torch.empty(L.lpslam_imadjust_scratch_bytes(N) + 16, dtype=torch.uint8, device="cuda")
sp = (scratch.data_ptr() + 15) // 16 * 16
st = torch.cuda.current_stream().cuda_stream


def k_gray():
    check(L.lpslam_gray_dev(color.data_ptr(), W, H, 3 * W, 3 * W * H, 3, 1, N, gray.data_ptr(), W, W * H, st))


def k_remap():
    check(L.lpslam_remap_linear_dev(gray.data_ptr(), W, H, W, W * H, mx.data_ptr(), my.data_ptr(), W, rect.data_ptr(), W, H, W, W * H, N, st))


def k_adjust():
    check(L.lpslam_imadjust_dev(rect.data_ptr(), W, H, W, W * H, N, 1, 0.0, 1.0, -0.3, 1.4, adj.data_ptr(), W, W * H, sp, st))


peak = 6545.3
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except OSError:
    pass
px = N * W * H
out = {}
for name, fn, nbytes in (("k_gray", k_gray, 4 * px), ("k_remap", k_remap, 2 * px + 8 * W * H), ("imadjust (k_hist+k_lut+k_apply_lut)", k_adjust, 3 * px)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out[name] = {"ms": round(ms, 4), "alg_bytes": nbytes, "achieved_gbs": round(nbytes / ms / 1e6, 1), "frac_of_hbm_peak": round(nbytes / ms / 1e6 / peak, 3)}
print(json.dumps({"frames": N, "shape": [H, W], "peak_gbs": peak, "kernels": out}))
