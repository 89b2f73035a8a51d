import os, sys, numpy as np
sys.path.insert(0, os.getcwd())
exec(open("tools/matcher_latency.py").read().split("out = {}")[0])
src = np.concatenate([np.arange(n)] * 3)
reproj = np.stack([kp["x"][src], kp["y"][src]], 1).astype(np.float64) + rng.normal(0, 2, (len(src), 2))
level = np.clip(kp["octave"][src] + rng.integers(0, 2, len(src)), 0, 7).astype(np.int32)
ldesc = desc[src] ^ np.packbits(rng.random((len(src), 256)) < 0.05, axis=1, bitorder="little")
act = np.ones(len(src), np.uint8); lxr = np.zeros(len(src), np.float32); xr = np.full(n, -1, np.float32); free = np.zeros(n, np.uint8)
m = match.projection(0.8, True)
for _ in range(5):
    m.match_frame_and_landmarks(cam, act, reproj, lxr, level, ldesc, kp, desc, xr, free, sf, 5.0)
os.environ[sys.argv[1] if len(sys.argv) > 1 else "LPS_MATCH_TIMING"] = "1"
for _ in range(3):
    m.match_frame_and_landmarks(cam, act, reproj, lxr, level, ldesc, kp, desc, xr, free, sf, 5.0)
