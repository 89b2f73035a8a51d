"""Matcher cases shared by tools/gen_golden_match.py (reference outputs -> tests/golden/match_*.npz) and
tests/test_match_golden_gpu.py (CUDA against those outputs). Each entry of CASES is run(matcher, port): `matcher` answers the
matcher calls (the reference build, the port, or the CUDA mirror wrapped to the same method names), `port` supplies the
extractor that builds keypoints for the geometric cases. Sizes follow BASELINE configs C1 / C2 / C3 (1000 / 2000 / 4000
keypoints)."""
import numpy as np

import test_match_gpu as tm
from oracle import pyoracle


def _bf(n, seed, chk, clustered=False):
    def run(m, port):
        rng = np.random.default_rng(seed + 100)
        if clustered:
            protos = rng.integers(0, 256, (6, 32), dtype=np.uint8)
            d1, d2 = tm.clustered(rng, n, protos, 0.03), tm.clustered(rng, n + n // 8, protos, 0.03)
        else:
            d1, d2 = tm.frame_pair(n, n, seed)
        valid = (rng.random(len(d2)) < 0.8).astype(np.uint8)
        a1 = rng.uniform(0, 360, len(d1)).astype(np.float32)
        a2 = rng.uniform(0, 360, len(d2)).astype(np.float32)
        out, num = m.match_bruteforce(d1, d2, valid, a1, a2, 0.8, chk)
        return dict(match_2_in_1=out, num=np.int64(num))
    return run


def _proj(seed, stereo, temps, chk, margin):
    def run(m, port):
        args, Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, obs = tm.projection_case(port, seed, stereo, temps)
        out, num = m.match_projection(pyoracle.make_camera(**args), Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, margin,
                                      chk, obs)
        return dict(curr_match_last=out, num=np.int64(num))
    return run


def _lm(seed, stereo, temps, margin, ratio):
    def run(m, port):
        args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = tm.landmarks_case(port, seed, stereo, temps)
        out, num = m.match_landmarks(pyoracle.make_camera(**args), active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, margin, ratio, obs)
        return dict(kp_match_lm=out, num=np.int64(num))
    return run


def _area(seed, chk):
    def run(m, port):
        from lpslam_b200.synth import synth_image
        o = port.extractor(max_kpts=1000)
        k1, d1 = o.extract(synth_image(480, 640, 50 + seed))
        k2, d2 = o.extract(np.roll(synth_image(480, 640, 50 + seed), (3, 7), axis=(0, 1)))
        prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
        cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
        out, num, pp = m.match_area(cam, k1, d1, prev, k2, d2, 100, 0.9, chk)
        return dict(matched_2_in_1=out, num=np.int64(num), prev=pp)
    return run


CASES = {
    "bf_c1_1000": _bf(1000, 1, False), "bf_c1_1000_orient": _bf(1000, 1, True),
    "bf_c2_2000": _bf(2000, 2, False), "bf_c2_2000_orient": _bf(2000, 2, True),
    "bf_c3_4000": _bf(4000, 3, False), "bf_c3_4000_orient": _bf(4000, 3, True),
    "bf_clustered_1500": _bf(1500, 31, False, True),
    "proj_mono": _proj(1, False, False, True, 20.0), "proj_mono_noorient": _proj(1, False, False, False, 40.0),
    "proj_stereo": _proj(2, True, False, True, 10.0), "proj_stereo_temporaries": _proj(3, True, True, True, 20.0),
    "landmarks_mono": _lm(1, False, False, 5.0, 0.8), "landmarks_stereo_temporaries": _lm(3, True, True, 20.0, 0.6),
    "area_orient": _area(1, True), "area": _area(2, False),
}
