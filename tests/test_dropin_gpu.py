"""GPU: the C++ drop-in layer (cpp/openvslam/feature/orb_extractor.h + orb_extractor_cuda.cc, match/stereo_cuda.cc) driven
exactly as the reference drives its own classes (tracking_module.cc:78-94, data/frame.cc:55-93), compared with the oracle.
The harness binary is built by cpp/Makefile in the build container (it needs the reference headers) and travels with gpurun."""
import os
import subprocess

import numpy as np
import pytest

from lpslam_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "cpp", "build", "dropin_smoke")
FX_B, B = 700.0 * 0.12, 0.12
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def test_dropin_classes_match_oracle(oracle, tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("cpp/build/dropin_smoke not built (needs /root/reference at build time)")
    h, w, n = 360, 640, 800
    left, right = synth_stereo_pair(h, w, 4242)
    lp, rp, out = tmp_path / "l.raw", tmp_path / "r.raw", tmp_path / "out.bin"
    left.tofile(lp)
    right.tofile(rp)
    r = subprocess.run([BIN, str(w), str(h), str(n), str(lp), str(rp), str(out), repr(FX_B), repr(B)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = open(out, "rb").read()
    nl, nr = np.frombuffer(raw, np.uint32, 2)
    o = 8
    kl = np.frombuffer(raw, KP, nl, o); o += 28 * nl
    dl = np.frombuffer(raw, np.uint8, 32 * nl, o).reshape(-1, 32); o += 32 * nl
    kr = np.frombuffer(raw, KP, nr, o); o += 28 * nr
    dr = np.frombuffer(raw, np.uint8, 32 * nr, o).reshape(-1, 32); o += 32 * nr
    xr = np.frombuffer(raw, np.float32, nl, o); o += 4 * nl
    dep = np.frombuffer(raw, np.float32, nl, o); o += 4 * nl
    pw, ph = np.frombuffer(raw, np.int32, 2, o); o += 8
    p3 = np.frombuffer(raw, np.uint8, pw * ph, o).reshape(ph, pw); o += pw * ph
    same_after_copy = np.frombuffer(raw, np.int32, 1, o)[0]; o += 4
    copy_levels = []
    for _ in range(1, 8):
        mw, mh = np.frombuffer(raw, np.int32, 2, o); o += 8
        copy_levels.append(np.frombuffer(raw, np.uint8, mw * mh, o).reshape(mh, mw)); o += mw * mh
    empty_image_untouched = np.frombuffer(raw, np.int32, 1, o)[0]; o += 4
    assert o == len(raw)

    ol, orr = oracle.extractor(max_kpts=n), oracle.extractor(max_kpts=n)
    okl, odl = ol.extract(left)
    okr, odr = orr.extract(right)
    assert nl == len(okl) and nr == len(okr)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kl[f], okl[f]) and np.array_equal(kr[f], okr[f]), f
    assert np.abs(kl["angle"] - okl["angle"]).max() <= 1e-4 * 180 / np.pi
    assert (dl == odl).all(axis=1).mean() >= 0.999 and (dr == odr).all(axis=1).mean() >= 0.999
    assert np.array_equal(p3, ol.level(3))  # image_pyramid_ side effect (orb_extractor.h:76)
    assert empty_image_untouched == 1  # orb_extractor.cc:73-75 through the drop-in class: outputs keep what they held
    assert same_after_copy == 1  # a copy-constructed extractor (own device context) extracts the same keypoints and descriptors
    for level, m in enumerate(copy_levels, 1):  # levels 1..7 arrive in one copy into one allocation
        assert np.array_equal(m, ol.level(level)), level
    oxr, odep = oracle.match_stereo([ol.level(l) for l in range(8)], [orr.level(l) for l in range(8)], okl, okr, odl, odr,
                                    ol.table(0), ol.table(1), FX_B, B)
    assert (dl == odl).all() and (dr == odr).all()
    assert np.array_equal(xr, oxr) and np.array_equal(dep, odep)
    assert (xr >= 0).sum() > 50


def test_dropin_matcher_classes_match_reference(oracle):
    """cpp/openvslam/match/robust_cuda.cc and projection_cuda.cc: the class methods robust::brute_force_match and
    projection::match_current_and_last_frames, called on data::frame / data::keyframe / data::landmark objects exactly as the
    reference's own classes are in oracle/_ref/libmatch_ref.so (same glue, same object construction), here running on the GPU.
    Held against the port, which conftest.PinnedOracle holds against the reference build."""
    import test_match_gpu as tm
    from oracle import pymatchref, pyoracle
    lib = os.path.join(ROOT, "cpp", "build", "libdropin_match.so")
    if not os.path.exists(lib):
        pytest.skip("cpp/build/libdropin_match.so not built (needs /root/reference at build time)")
    d = pymatchref.MatchRef(lib)
    for n1, n2, seed in ((1000, 1000, 1), (2000, 2000, 2), (37, 500, 4)):
        d1, d2 = tm.frame_pair(n1, n2, seed)
        rng = np.random.default_rng(seed + 100)
        valid = (rng.random(n2) < 0.8).astype(np.uint8)
        a1, a2 = rng.uniform(0, 360, n1).astype(np.float32), rng.uniform(0, 360, n2).astype(np.float32)
        for chk in (False, True):
            want, nw = oracle.match_bruteforce(d1, d2, valid, a1, a2, 0.8, chk)
            got, ng = d.match_bruteforce(d1, d2, valid, a1, a2, 0.8, chk)
            assert np.array_equal(got, want) and ng == nw
    for seed, stereo, temps in ((1, False, False), (2, True, False), (3, True, True)):
        args, Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, obs = tm.projection_case(oracle, seed, stereo, temps)
        for margin, chk in ((10.0, True), (20.0, False)):
            cam = pyoracle.make_camera(**args)
            want, nw = oracle.match_projection(cam, Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, margin, chk, obs)
            got, ng = d.match_projection(cam, Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, margin, chk, obs)
            assert np.array_equal(got, want) and ng == nw and nw > 50
