#!/usr/bin/env python
"""Phase clocks of the level-0 quadtree block of frame 0 (library built with LPSLAM_NVCC_FLAGS=-DLPS_QT_PROFILE)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lpslam_b200 import match, orb  # noqa: E402
from lpslam_b200._lib import lib  # noqa: E402
from lpslam_b200.synth import synth_stereo_pair  # noqa: E402

B = 32
e = orb.orb_extractor(max_num_keypts=2000, max_batch=2 * B)
pairs = [synth_stereo_pair(720, 1280, 2000 + k) for k in range(4)]
for _ in range(3):
    match.stereo_frames(e, [pairs[k % 4][0] for k in range(B)], [pairs[k % 4][1] for k in range(B)], 84.0, 0.12)
out = (C.c_ulonglong * 16)()
L = lib()
L.lpslam_debug_qt_profile.argtypes = [C.POINTER(C.c_ulonglong)]
print("rc", L.lpslam_debug_qt_profile(out))
t = list(out)
names = ["A list+keys", "B radix sort", "C initial nodes", "(quota)", "D breadth passes", "E fullest-first", "F winners"]
for i in range(6):
    print(f"{names[i]:18s} {t[i + 1] - t[i]:9d} clk")
print("total", t[6] - t[0], "clk; N =", t[8], "m =", t[9], "keybits =", t[10], "quota =", t[11])
# deep trees: clustered corners force splits below the prefix-ordered depth (fallback counter, profile build only)
import numpy as np  # noqa: E402
rng = np.random.default_rng(1)
img = np.full((480, 640), 120, np.uint8)
for _ in range(3):
    y0, x0 = int(rng.integers(40, 380)), int(rng.integers(40, 540))
    img[y0:y0 + 48, x0:x0 + 56] = rng.integers(0, 256, (48, 56), dtype=np.uint8)
L.lpslam_debug_qt_profile(out)
before = out[12]
orb.orb_extractor(max_num_keypts=3000).extract(img)
L.lpslam_debug_qt_profile(out)
print("complete-sort fallbacks on the clustered image:", out[12] - before, "(of 8 levels); during the stereo batches:", before)
