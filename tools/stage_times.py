#!/usr/bin/env python
"""Per-stage device time (median of N device-resident stereo steps, CUDA events inside the library) for kernel tuning.
usage: stage_times.py [pairs=32] [reps=20]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lpslam_b200 import match, orb  # noqa: E402
from lpslam_b200._lib import check, lib  # noqa: E402
from lpslam_b200.synth import synth_stereo_pair  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
W, H = 1280, 720
L = lib()
e = orb.orb_extractor(max_num_keypts=2000, max_batch=2 * B)
pairs = [synth_stereo_pair(H, W, 2000 + k) for k in range(min(B, 4))]
match.stereo_frames(e, [pairs[k % len(pairs)][0] for k in range(B)], [pairs[k % len(pairs)][1] for k in range(B)], 84.0, 0.12)
dptr, dpitch, dstride = C.c_void_p(), C.c_size_t(), C.c_size_t()
check(L.lpslam_orb_input_dev(e._h, W, H, C.byref(dptr), C.byref(dpitch), C.byref(dstride)))
e.enable_timing(True)
acc = {}
for _ in range(reps + 3):
    check(L.lpslam_stereo_frames_dev(e._h, dptr, B, W, H, dpitch.value, 84.0, 0.12))
    st = e.stage_times()
    for k, v in st.items():
        acc.setdefault(k, []).append(v)
print(os.environ.get("LPSLAM_NVCC_FLAGS", ""), {k: round(float(np.median(v[3:])), 4) for k, v in acc.items()})
