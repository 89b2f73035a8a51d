#!/usr/bin/env python
"""Wall-clock latency per call of the per-frame matchers through the host-buffer C ABI (H2D + kernels + D2H, what a tracker
thread sees) next to the CPU oracle on the same inputs, at BASELINE configs[1] sizes (about 2400 keypoints per frame).
These are the 'replicas only' workloads of DESIGN.md section 6: microseconds of arithmetic, so the figure that matters is
the call latency, not a roofline. gpu_ms goes through the Python mirror of the matcher classes (numpy marshalling included),
c_abi_ms calls the C entry point with the arguments already marshalled (what the C++ drop-in bodies pay), median and p99.
usage: matcher_latency.py [reps=200]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C  # noqa: E402

from lpslam_b200 import _lib, match  # noqa: E402
from lpslam_b200.synth import synth_image  # noqa: E402
from oracle import pyoracle  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
o = pyoracle.port()
rng = np.random.default_rng(0)
ext = o.extractor(max_kpts=2000)
kp, desc = ext.extract(synth_image(720, 1280, 2000))
kp2, desc2 = o.extractor(max_kpts=2000).extract(np.roll(synth_image(720, 1280, 2000), (2, 5), (0, 1)))
n = len(kp)
sf = ext.table(0)
args = dict(fx=700.0, fy=700.0, cx=640.0, cy=360.0, cols=1280, rows=720)
ocam, cam = pyoracle.make_camera(**args), match.make_camera(**args)


def timeit(fn, k=None):
    fn()
    fn()
    t0 = time.perf_counter()
    for _ in range(k or min(reps, 30)):
        fn()
    return (time.perf_counter() - t0) / (k or min(reps, 30)) * 1e3


def abi(fn):
    """median / p99 of `reps` single calls (ms)"""
    for _ in range(5):
        fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append((time.perf_counter() - t0) * 1e3)
    t.sort()
    return {"c_abi_ms": round(t[len(t) // 2], 4), "c_abi_p99_ms": round(t[min(len(t) - 1, int(len(t) * 0.99))], 4)}


L = _lib.lib()


out = {}
# brute force (robust.cc:215-321)
noisy = desc ^ np.packbits(rng.random((n, 256)) < 0.05, axis=1, bitorder="little")
out["brute_force_match"] = {"gpu_ms": timeit(lambda: match.robust(0.8, False).brute_force_match(desc, noisy)),
                            "cpu_ms": timeit(lambda: o.match_bruteforce(desc, noisy, None, None, None, 0.8, False))}
bf_out, bf_n = np.full(n, -1, np.int32), C.c_uint32(0)
out["brute_force_match"].update(abi(lambda: L.lpslam_match_bruteforce(desc.ctypes.data, n, noisy.ctypes.data, n, None, None, None, 0.8, 0,
                                                                      bf_out.ctypes.data, C.byref(bf_n))))
# local-map search (projection.cc:11-83): 3 landmarks per keypoint
src = np.concatenate([np.arange(n)] * 3)
reproj = np.stack([kp["x"][src], kp["y"][src]], 1).astype(np.float64) + rng.normal(0, 2, (len(src), 2))
level = np.clip(kp["octave"][src] + rng.integers(0, 2, len(src)), 0, 7).astype(np.int32)
ldesc = desc[src] ^ np.packbits(rng.random((len(src), 256)) < 0.05, axis=1, bitorder="little")
act = np.ones(len(src), np.uint8)
lxr = np.zeros(len(src), np.float32)
xr = np.full(n, -1, np.float32)
free = np.zeros(n, np.uint8)
out["match_frame_and_landmarks"] = {
    "rows": int(len(src)),
    "gpu_ms": timeit(lambda: match.projection(0.8, True).match_frame_and_landmarks(cam, act, reproj, lxr, level, ldesc, kp, desc, xr, free, sf, 5.0)),
    "cpu_ms": timeit(lambda: o.match_landmarks(ocam, act, reproj, lxr, level, ldesc, kp, desc, xr, free, sf, 5.0, 0.8))}
lm_out, lm_n = np.full(n, -1, np.int32), C.c_uint32(0)
sf32 = np.ascontiguousarray(sf, np.float32)
out["match_frame_and_landmarks"].update(abi(lambda: L.lpslam_match_landmarks(
    C.byref(cam), len(src), act.ctypes.data, reproj.ctypes.data, lxr.ctypes.data, level.ctypes.data, ldesc.ctypes.data, None, n, kp.ctypes.data,
    desc.ctypes.data, xr.ctypes.data, free.ctypes.data, sf32.ctypes.data, 5.0, 0.8, lm_out.ctypes.data, C.byref(lm_n))))
# initializer (area.cc:8-112)
prev = np.stack([kp["x"], kp["y"]], 1).astype(np.float32)
out["match_in_consistent_area"] = {"gpu_ms": timeit(lambda: match.area(0.9, True).match_in_consistent_area(cam, kp, desc, kp2, desc2, prev, 100)),
                                   "cpu_ms": timeit(lambda: o.match_area(ocam, kp, desc, prev, kp2, desc2, 100, 0.9, True))}
ar_out, ar_n, ar_prev = np.full(n, -1, np.int32), C.c_uint32(0), prev.copy()
out["match_in_consistent_area"].update(abi(lambda: L.lpslam_match_area(C.byref(cam), n, kp.ctypes.data, desc.ctypes.data, ar_prev.ctypes.data, len(kp2),
                                                                        kp2.ctypes.data, desc2.ctypes.data, 100, 0.9, 1, ar_out.ctypes.data,
                                                                        C.byref(ar_n))))
# frame post-processing (frame.cc:37-51)
model = match.Perspective(700.0, 700.0, 640.0, 360.0, (match.C.c_double * 5)(-0.17, 0.027, 0.0005, -0.0003, 0.0))


def cpu_post():
    u = o.undistort_keypoints(700.0, 700.0, 640.0, 360.0, [-0.17, 0.027, 0.0005, -0.0003, 0.0], kp)
    o.keypoints_to_bearings(700.0, 700.0, 640.0, 360.0, u)
    o.assign_grid(ocam, u)


out["frame_postprocess"] = {"gpu_ms": timeit(lambda: match.frame_postprocess(model, cam, kp)), "cpu_ms": timeit(cpu_post)}
for v in out.values():
    v["gpu_ms"], v["cpu_ms"] = round(v["gpu_ms"], 3), round(v["cpu_ms"], 3)
print(json.dumps({"keypoints": int(n), "reps": reps,
                  "note": "wall clock per call; gpu_ms / cpu_ms incl. numpy marshalling on both sides, c_abi_ms = the C entry point alone "
                          "(pinned staging, one copy in, kernels, one copy out, on the calling thread's stream); CPU = oracle port, 1 thread",
                  "calls": out}))
