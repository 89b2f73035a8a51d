#!/usr/bin/env python
"""Runs N device-resident stereo steps (no CPU baseline, no end-to-end leg) so that ncu sees a predictable launch order:
15 kernels per step — 7 x k_resize_linear, k_fast_cells, k_quadtree, k_gauss7, k_offsets, k_describe, k_stereo_prep,
k_stereo_match, k_stereo_median."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lpslam_b200 import match, orb  # noqa: E402
from lpslam_b200._lib import check, lib  # noqa: E402
from lpslam_b200.synth import synth_stereo_pair  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
W, H = 1280, 720
L = lib()
e = orb.orb_extractor(max_num_keypts=2000, max_batch=2 * B)
pairs = [synth_stereo_pair(H, W, 2000 + k) for k in range(min(B, 4))]
match.stereo_frames(e, [pairs[k % len(pairs)][0] for k in range(B)], [pairs[k % len(pairs)][1] for k in range(B)], 84.0, 0.12)
dptr, dpitch, dstride = C.c_void_p(), C.c_size_t(), C.c_size_t()
check(L.lpslam_orb_input_dev(e._h, W, H, C.byref(dptr), C.byref(dpitch), C.byref(dstride)))
for _ in range(steps):
    check(L.lpslam_stereo_frames_dev(e._h, dptr, B, W, H, dpitch.value, 84.0, 0.12))
check(L.lpslam_orb_sync(e._h))
print("done")
