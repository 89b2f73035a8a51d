"""median wall time of lpslam_stereo_frames with one 720p pair per call (LPSLAM_CUDA_GRAPH=0 / 1)"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.getcwd())
from lpslam_b200 import match, orb, replay  # noqa: E402,F401
from lpslam_b200._lib import KP_DTYPE, check, lib  # noqa: E402
from lpslam_b200.synth import synth_stereo_pair  # noqa: E402

L = lib()
host = replay.pinned_array((4, 720, 1280))
for p in range(2):
    host[2 * p], host[2 * p + 1] = synth_stereo_pair(720, 1280, 100 + p)
e = orb.orb_extractor(max_num_keypts=2000, max_batch=2)
cap = 8192
kp, desc = np.empty(cap, KP_DTYPE), np.empty((cap, 32), np.uint8)
xr, dep, offs = np.empty(cap, np.float32), np.empty(cap, np.float32), np.zeros(3, np.uint32)
ts = []
for i in range(220):
    p = i % 2
    pl, pr = (C.c_void_p * 1)(host[2 * p].ctypes.data), (C.c_void_p * 1)(host[2 * p + 1].ctypes.data)
    t = time.perf_counter()
    check(L.lpslam_stereo_frames(e._h, pl, pr, 1, 1280, 720, 1280, 84.0, 0.12, kp.ctypes.data, desc.ctypes.data, cap, offs.ctypes.data,
                                 xr.ctypes.data, dep.ctypes.data))
    ts.append((time.perf_counter() - t) * 1e3)
ts = np.sort(ts[20:])
print("graph", os.environ.get("LPSLAM_CUDA_GRAPH", "default"), "median ms", round(float(np.median(ts)), 4), "p99", round(float(ts[197]), 4),
      "keypoints", int(offs[2]))
