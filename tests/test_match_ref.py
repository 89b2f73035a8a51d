"""CPU pin of the matcher oracle: every case of tests/test_match_gpu.py is replayed here WITHOUT a GPU, with the CUDA mirror
(lpslam_b200.match) replaced by a stand-in that answers from the port. The `oracle` fixture is conftest.PinnedOracle, which runs
every matcher call through the reference's own sources as well (oracle/_ref/libmatch_ref.so = match/robust.cc, projection.cc,
area.cc, fuse.cc, bow_tree.cc, data/common.cc, data/landmark.cc, camera/perspective.cc compiled unmodified) and fails on any
difference. What the GPU tests then prove (CUDA == port on these cases) is CUDA == reference.

Also reproduces the known-answer vectors the reference's own unit tests hold for this path:
test/openvslam/match/angle_checker.cc, test/openvslam/data/common_get_cell_indices.cc, test/openvslam/match/base.cc."""
import ctypes as C
import os
import sys
import types

import numpy as np
import pytest

from lpslam_b200._lib import KP_DTYPE

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def _fake_match(port):
    """lpslam_b200.match's classes and functions, answered by the port (same argument order as the mirror)"""
    from oracle import pyoracle
    m = types.ModuleType("lpslam_b200.match")
    m.KP_DTYPE = KP_DTYPE
    m.C = C
    m.make_camera = pyoracle.make_camera

    class Perspective:
        def __init__(self, fx, fy, cx, cy, dist):
            self.fx, self.fy, self.cx, self.cy, self.dist = fx, fy, cx, cy, list(dist)

    class ReprojMatchParams:
        def __init__(self, margin, lowe_ratio, hamm_dist_thr, level_lo, level_hi, use_ratio, use_x_right):
            self.v = (margin, lowe_ratio, hamm_dist_thr, level_lo, level_hi, use_ratio, use_x_right)

    def frame_postprocess(model, cam, kp):
        und = port.undistort_keypoints(model.fx, model.fy, model.cx, model.cy, model.dist, kp)
        start, items = port.assign_grid(cam, und)
        return und, port.keypoints_to_bearings(model.fx, model.fy, model.cx, model.cy, und), start, items

    def match_reprojected(cam, prm, act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, obs=None, ang=None):
        margin, ratio, thr, lo, hi, ur, ux = prm.v
        lxr = np.zeros(len(act), np.float32) if lxr is None else lxr
        xr = np.full(len(kp), -1.0, np.float32) if xr is None else xr
        return port.match_reprojected(cam, margin, ratio, thr, lo, hi, ur, ux, act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, obs, ang)

    def compute_landmark_descriptors(descs, offsets):
        n = len(offsets) - 1
        best, out = np.zeros(n, np.uint32), np.zeros((n, 32), np.uint8)
        for l in range(n):
            d = descs[offsets[l]:offsets[l + 1]]
            if len(d):
                best[l] = port.landmark_descriptor(d)
                out[l] = d[best[l]]
        return best, out

    class robust:
        def __init__(self, lowe_ratio=0.8, check_orientation=False):
            self.r, self.c = lowe_ratio, check_orientation

        def brute_force_match(self, d1, d2, valid=None, a1=None, a2=None):
            out, n = port.match_bruteforce(d1, d2, valid, a1, a2, self.r, self.c)
            i1 = np.nonzero(out >= 0)[0]
            return np.stack([i1, out[i1]], axis=1).astype(np.int32), out

        def match_for_triangulation(self, v1, s1, v2, s2, E, epi):
            out, n = port.match_triangulation(v1, v2, s1, s2, E, epi, self.c)
            i1 = np.nonzero(out >= 0)[0]
            return np.stack([i1, out[i1]], axis=1).astype(np.int32), out

    class projection:
        def __init__(self, lowe_ratio=0.6, check_orientation=True):
            self.r, self.c = lowe_ratio, check_orientation

        def match_current_and_last_frames(self, cam, Tc, Tl, has, outl, pos, ld, loct, lang, kp, cd, xr, tk, sf, margin, obs=None):
            return port.match_projection(cam, Tc, Tl, has, outl, pos, ld, loct, lang, kp, cd, xr, tk, sf, margin, self.c, obs)

        def match_frame_and_landmarks(self, cam, act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, margin=5.0, obs=None):
            if len(kp) == 0 or len(act) == 0:
                return np.full(len(kp), -1, np.int32), 0
            return port.match_landmarks(cam, act, rp, lxr, lvl, ld, kp, cd, xr, tk, sf, margin, self.r, obs)

        def match_frame_and_keyframe(self, cam, act, rp, lvl, ld, ang, kp, cd, tk, sf, margin, thr):
            return port.match_reprojected(cam, margin, 0.0, thr, -1, 1, 0, 0, act, rp, np.zeros(len(act), np.float32), lvl, ld, kp, cd,
                                          np.full(len(kp), -1.0, np.float32), tk, sf, None, ang if self.c else None)

        def match_by_Sim3_transform(self, cam, act, rp, lvl, ld, kp, cd, tk, sf, margin):
            return port.match_sim3(cam, margin, act, rp, lvl, ld, kp, cd, tk, sf)

        def match_keyframes_mutually(self, c1, c2, s1, s2, margin):
            return port.match_mutual(c1, c2, margin, s1, s2)

    class fuse:
        def __init__(self, lowe_ratio=0.6):
            pass

        def detect_duplication(self, cam, act, rp, lvl, ld, kp, cd, sf, margin):
            if len(kp) == 0:
                return np.full(len(act), -1, np.int32), np.full(len(act), 256, np.uint32), 0
            z = np.zeros(len(act), np.float32)
            return port.match_fuse(cam, margin, False, False, act, rp, z, lvl, ld, kp, cd, np.full(len(kp), -1.0, np.float32), sf,
                                   np.ones(len(sf), np.float32))

        def replace_duplication(self, cam, act, rp, lxr, lvl, ld, kp, cd, xr, sf, isq, margin=3.0):
            return port.match_fuse(cam, margin, True, True, act, rp, lxr, lvl, ld, kp, cd, xr, sf, isq)

    class bow_tree:
        def __init__(self, lowe_ratio=0.6, check_orientation=True):
            self.r, self.c = lowe_ratio, check_orientation

        def match_frame_and_keyframe(self, kvec, valid, kd, kang, fvec, fd, fang):
            return port.match_bow(kvec, fvec, valid, kd, kang, fd, fang, self.r, self.c)

        def match_keyframes(self, v1, val1, d1, a1, v2, val2, d2, a2):
            return port.match_bow_keyframes(v1, v2, val1, d1, a1, val2, d2, a2, self.r, self.c)

    class area:
        def __init__(self, lowe_ratio=0.9, check_orientation=True):
            self.r, self.c = lowe_ratio, check_orientation

        def match_in_consistent_area(self, cam, k1, d1, k2, d2, prev, margin=100):
            return port.match_area(cam, k1, d1, prev, k2, d2, margin, self.r, self.c)

    def hamming(a, b):
        a2, b2 = np.atleast_2d(a), np.atleast_2d(b)
        out = np.array([port.hamming(x, y) for x, y in zip(a2, b2)], np.uint32)
        return out if np.ndim(a) == 2 else int(out[0])

    for k, v in dict(Perspective=Perspective, ReprojMatchParams=ReprojMatchParams, frame_postprocess=frame_postprocess,
                     match_reprojected=match_reprojected, compute_landmark_descriptors=compute_landmark_descriptors, robust=robust,
                     projection=projection, fuse=fuse, bow_tree=bow_tree, area=area, compute_descriptor_distance_32=hamming).items():
        setattr(m, k, v)
    return m


# cases that drive a device-resident extractor context itself (nothing to replay through the port)
DEVICE_ONLY = {"test_match_consecutive_frames"}


def _gpu_cases():
    import test_match_gpu as tm  # its tests carry the gpu mark; here they are replayed on the CPU
    out = []
    for name in sorted(dir(tm)):
        fn = getattr(tm, name)
        if not name.startswith("test_") or not callable(fn) or name in DEVICE_ONLY:
            continue
        marks = [mk for mk in getattr(fn, "pytestmark", []) if mk.name == "parametrize"]
        if not marks:
            out.append(pytest.param(name, None, None, id=name))
            continue
        names = [s.strip() for s in marks[0].args[0].split(",")]
        for vals in marks[0].args[1]:
            vals = vals if isinstance(vals, (tuple, list)) and len(names) > 1 else (vals,)
            out.append(pytest.param(name, names, tuple(vals), id=f"{name}[{'-'.join(map(str, vals))}]"))
    return out


@pytest.mark.parametrize("name,argnames,argvals", _gpu_cases())
def test_replay_gpu_case_on_port_and_reference(oracle, match_ref, monkeypatch, name, argnames, argvals):
    import inspect

    import lpslam_b200
    import test_match_gpu as tm
    fake = _fake_match(oracle._port)
    monkeypatch.setitem(sys.modules, "lpslam_b200.match", fake)
    monkeypatch.setattr(lpslam_b200, "match", fake, raising=False)
    fn = getattr(tm, name)
    kwargs = dict(zip(argnames or [], argvals or []))
    if "oracle" in inspect.signature(fn).parameters:
        kwargs["oracle"] = oracle
    before = oracle.pinned_calls
    fn(**kwargs)
    if "oracle" in kwargs and name not in ("test_bruteforce_on_extracted_frames",):
        assert oracle.pinned_calls > before, "the case never reached the reference build"


# ---- known answers of the reference's own unit tests -------------------------------------------------------------------

def test_angle_checker_reference_cases(oracle, match_ref):
    """test/openvslam/match/angle_checker.cc:12-180, all five cases (constructor, top 1 / 2 / 3 bins, all bins), on the reference's own
    header and on the port."""
    deltas = [34.8, 34.9, 35.0, 35.1, 35.2, 323.8, 323.9, 324.1, 324.2, 126.9, 127.0, 127.1, 0.0, 90.0, 180.0, 270.0]
    ids = [35] * 5 + [324] * 4 + [127] * 3 + [0, 60, 180, 270]
    # check_assert (:5-8): thresholds 3 and 30 are constructible
    for thr in (3, 30):
        match_ref.angle_check(deltas, ids, thr)
    for thr, keep in ((1, {35}), (2, {35, 324}), (3, {35, 324, 127}), (30, {35, 324, 127, 0, 60, 180, 270})):
        inv, val = match_ref.angle_check(deltas, ids, thr)
        assert len(val) > 0 and set(val.tolist()) <= keep, thr                   # :40-45, :81-86, :123-128, :164-168
        assert set(inv.tolist()) <= set(ids) - keep and (len(inv) > 0) == (thr != 30), thr  # :47-51, :88-93, :130-135, :170-171
        assert sorted(val.tolist() + inv.tolist()) == sorted(ids)
        inv_port = oracle._port.angle_check(deltas, ids, thr)
        assert np.array_equal(inv_port, inv), thr                                  # same ids in the same order
    rng = np.random.default_rng(9)
    for _ in range(300):  # random histograms with many equal bin sizes: the std::sort tie handling of :126-134
        n = int(rng.integers(0, 60))
        d = (rng.integers(-12, 24, n) * 30.0 + rng.uniform(-14, 14, n)).astype(np.float32)
        i = np.arange(n, dtype=np.int32)
        thr = int(rng.choice([1, 2, 3, 5, 30]))
        assert np.array_equal(oracle._port.angle_check(d, i, thr), match_ref.angle_check(d, i, thr)[0])


def test_get_cell_indices_reference_cases(oracle, match_ref):
    """test/openvslam/data/common_get_cell_indices.cc:16-144, its three cases (valid_cases_1 with distortion, valid_cases_2 on every
    cell corner, invalid_cases), on the reference's get_cell_indices and on the port's grid assignment.

    The lpslam fork's get_cell_indices ROUNDS (data/common.h:60-61, cvRound) while this unit test of the fork still expects the
    truncating cell index of upstream OpenVSLAM: a point in the right / lower half of a cell lands in the next cell. The reference's
    own function, compiled here, therefore disagrees with its own test on those vectors (counted below). Parity is with the CODE:
    on every vector the port must equal the reference's function; the vectors that do satisfy the test's expectation are asserted too."""
    cols, rows = 2000, 1000

    def camera(k1=0.0, k2=0.0):  # create_perspective_camera (:8-14): fx = fy = rows, principal point at the centre, k1 / k2 as FLOAT literals
        return match_ref.camera_info(float(rows), float(rows), cols / 2.0, rows / 2.0, [float(np.float32(k1)), float(np.float32(k2)), 0, 0, 0], cols,
                                     rows, is_monocular=True)

    def check(cam, pts, want):
        kp = np.zeros(len(pts), KP_DTYPE)
        kp["x"], kp["y"] = [np.float32(p[0]) for p in pts], [np.float32(p[1]) for p in pts]
        start, items = oracle.assign_grid(cam, kp)  # PinnedOracle: port == reference's assign_keypoints_to_grid
        where = {}
        for c in range(cam.grid_cols * cam.grid_rows):
            for i in items[start[c]:start[c + 1]]:
                where[int(i)] = (c // cam.grid_rows, c % cam.grid_rows)
        as_expected = 0
        for i, (p, w) in enumerate(zip(pts, want)):
            found, cx, cy = match_ref.get_cell_indices(cam, np.float32(p[0]), np.float32(p[1]))
            assert (i in where) == found and (not found or where[i] == (cx, cy)), p  # port's grid == reference's function
            as_expected += (not found) if w is None else (found and (cx, cy) == w)
        return as_expected

    eps = np.float32(0.01)
    cam = camera(-0.1, 0.1)
    gc, gr = cam.grid_cols, cam.grid_rows
    lo_x, hi_x, lo_y, hi_y = np.float32(cam.min_x), np.float32(cam.max_x), np.float32(cam.min_y), np.float32(cam.max_y)
    # valid_cases_1 (:16-63): corners and edge centres of the undistorted bounds
    pts = [(lo_x, lo_y), (hi_x - eps, lo_y), (lo_x, hi_y - eps), (hi_x - eps, hi_y - eps), (cols / 2.0, lo_y), (cols / 2.0, hi_y - eps),
           (lo_x, rows / 2.0), (hi_x - eps, rows / 2.0)]
    want = [(0, 0), (gc - 1, 0), (0, gr - 1), (gc - 1, gr - 1), (gc // 2 - 1, 0), (gc // 2 - 1, gr - 1), (0, gr // 2 - 1), (gc - 1, gr // 2 - 1)]
    assert check(cam, pts, want) == 1  # only the top-left corner: the seven others sit in the upper half of their cell and round up
    # invalid_cases (:117-144)
    pts = [(lo_x - eps, lo_y - eps), (hi_x, lo_y - eps), (lo_x - eps, hi_y), (hi_x, hi_y), (cols / 2.0, lo_y - eps), (cols / 2.0, hi_y),
           (lo_x - eps, rows / 2.0), (hi_x, rows / 2.0)]
    assert check(cam, pts, [None] * len(pts)) == 5  # three points a hundredth outside the lower bound round INTO cell 0
    # valid_cases_2 (:65-115): the four corners of every cell of an undistorted camera
    cam = camera()
    cw, ch = 1.0 / cam.inv_cell_width, 1.0 / cam.inv_cell_height
    pts, want = [], []
    for ix in range(cam.grid_cols):
        for iy in range(cam.grid_rows):
            for dx, dy in ((0, 0), (1, 0), (0, 1), (1, 1)):
                x = np.float32((ix + dx) * cw + (-0.01 if dx else 0.01))
                y = np.float32((iy + dy) * ch + (-0.01 if dy else 0.01))
                pts.append((x, y))
                want.append((ix, iy))
    assert check(cam, pts, want) == len(pts) // 4  # the top-left corner of every cell


def test_keypoints_in_cell_port_vs_reference(oracle, match_ref):
    """data/common.cc:127-186 on random keypoints: the candidate list (order included) of the reference equals what the port's
    matchers iterate — checked through a matcher whose result lists every candidate (fuse search with a 256-bit threshold would
    hide order), so compare the lists directly via the port's grid + a brute-force filter in the reference's visiting order."""
    from oracle import pyoracle
    rng = np.random.default_rng(3)
    cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    n = 3000
    kp = np.zeros(n, KP_DTYPE)
    kp["x"], kp["y"] = rng.uniform(-5, 645, n).astype(np.float32), rng.uniform(-5, 485, n).astype(np.float32)
    kp["octave"] = rng.integers(0, 8, n)
    start, items = oracle.assign_grid(cam, kp)
    for _ in range(200):
        rx, ry = np.float32(rng.uniform(-30, 670)), np.float32(rng.uniform(-30, 510))
        margin = np.float32(rng.choice([3.0, 7.5, 20.0, 64.0]))
        lo, hi = rng.choice([(-1, -1), (0, 0), (2, 4), (-1, 3), (5, -1)])
        got = match_ref.keypoints_in_cell(cam, kp, rx, ry, margin, lo, hi)
        # restatement of :127-186 over the port's CSR grid
        icw, ich = cam.inv_cell_width, cam.inv_cell_height
        x0 = max(0, int(np.floor((float(rx) - cam.min_x - float(margin)) * icw)))
        x1 = min(cam.grid_cols - 1, int(np.ceil((float(rx) - cam.min_x + float(margin)) * icw)))
        y0 = max(0, int(np.floor((float(ry) - cam.min_y - float(margin)) * ich)))
        y1 = min(cam.grid_rows - 1, int(np.ceil((float(ry) - cam.min_y + float(margin)) * ich)))
        want = []
        if x0 < cam.grid_cols and x1 >= 0 and y0 < cam.grid_rows and y1 >= 0:
            chk = (0 < lo) or (0 <= hi)
            for cx in range(x0, x1 + 1):
                for cy in range(y0, y1 + 1):
                    for i in items[start[cx * cam.grid_rows + cy]:start[cx * cam.grid_rows + cy + 1]]:
                        if chk and (kp["octave"][i] < lo or (0 <= hi and hi < kp["octave"][i])):
                            continue
                        if abs(np.float32(kp["x"][i] - rx)) < margin and abs(np.float32(kp["y"][i] - ry)) < margin:
                            want.append(i)
        assert list(got) == want


def test_hamming_reference_cases(oracle, match_ref):
    # test/openvslam/match/base.cc:11-57 on the reference's own compute_descriptor_distance_32 / _64
    a = np.zeros(32, np.uint8)
    for v in (32, 64):
        assert match_ref.hamming(a, a, v) == 0
        assert match_ref.hamming(a, ~a, v) == 256
        assert match_ref.hamming(np.full(32, 0x66, np.uint8), np.full(32, 0x3C, np.uint8), v) == 128
    rng = np.random.default_rng(1)
    for _ in range(200):
        x, y = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
        assert oracle.hamming(x, y) == int(np.unpackbits(x ^ y).sum())  # PinnedOracle: port == reference


def test_camera_and_reprojection_reference(match_ref):
    """camera/perspective.cc:20-31 (bounds / grid constants with and without distortion) and :151-169 (reprojection) against the
    flat camera description the C ABI takes (lpslam_b200.match.make_camera == pyoracle.make_camera)."""
    from oracle import pyoracle
    ref_cam = match_ref.camera_info(500.0, 500.0, 320.0, 240.0, [0, 0, 0, 0, 0], 640, 480, 50.0)
    cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480, focal_x_baseline=50.0, true_baseline=0.1)
    for f in ("min_x", "max_x", "min_y", "max_y", "grid_cols", "grid_rows", "inv_cell_width", "inv_cell_height", "focal_x_baseline",
              "true_baseline"):
        assert getattr(ref_cam, f) == getattr(cam, f), f
    dist_cam = match_ref.camera_info(700.0, 705.0, 640.0, 360.0, [-0.17, 0.027, 0.0005, -0.0003, 0.0], 1280, 720)
    assert dist_cam.min_x < 0 and dist_cam.max_x > 1280  # barrel distortion: undistorted corners lie outside the image


def test_descriptor_json_reference(match_ref):
    """data/common.cc:71-95 against lpslam_b200.mapio's format code (io/map_database_io.cc): same JSON numbers, same bytes back."""
    import json

    from lpslam_b200 import mapio
    rng = np.random.default_rng(4)
    d = rng.integers(0, 256, (50, 32), dtype=np.uint8)
    text = match_ref.descriptors_to_json(d)
    assert json.loads(text) == mapio.convert_descriptors_to_json(d)
    assert np.array_equal(match_ref.json_to_descriptors(json.dumps(mapio.convert_descriptors_to_json(d))), d)
    assert np.array_equal(mapio.convert_json_to_descriptors(json.loads(text)), d)
