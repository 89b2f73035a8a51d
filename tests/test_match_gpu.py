"""GPU parity of the Hamming matchers against the oracle: match indices and distances bit-exact given identical descriptors."""
import numpy as np
import pytest

from lpslam_b200.synth import synth_descriptors, synth_image

pytestmark = pytest.mark.gpu


def test_hamming_known_answers():
    # test/openvslam/match/base.cc:11-57
    from lpslam_b200 import match
    a = np.zeros(32, np.uint8)
    assert match.compute_descriptor_distance_32(a, a) == 0
    assert match.compute_descriptor_distance_32(a, ~a) == 256
    assert match.compute_descriptor_distance_32(np.full(32, 0x66, np.uint8), np.full(32, 0x3C, np.uint8)) == 128
    rng = np.random.default_rng(1)
    x = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    y = rng.integers(0, 256, (500, 32), dtype=np.uint8)
    want = np.unpackbits(x ^ y, axis=1).sum(axis=1)
    assert np.array_equal(match.compute_descriptor_distance_32(x, y), want)


def frame_pair(n1, n2, seed, flip=0.06):
    """keyframe descriptors = noisy copies of frame descriptors (so matches exist), shuffled, plus duplicates to force
    claim conflicts."""
    rng = np.random.default_rng(seed)
    d1 = rng.integers(0, 256, (n1, 32), dtype=np.uint8)
    src = rng.integers(0, n1, n2)
    src[: n2 // 5] = rng.integers(0, max(n1 // 20, 1), n2 // 5)   # many rows fight for a few frame keypoints
    flips = rng.random((n2, 256)) < flip
    d2 = d1[src] ^ np.packbits(flips, axis=1, bitorder="little")
    return d1, d2


@pytest.mark.parametrize("n1,n2,seed", [(1000, 1000, 1), (2000, 2000, 2), (4000, 4000, 3), (37, 500, 4), (500, 3, 5), (3, 40, 6)])
def test_bruteforce(oracle, n1, n2, seed):
    from lpslam_b200 import match
    d1, d2 = frame_pair(n1, n2, seed)
    rng = np.random.default_rng(seed + 100)
    valid = (rng.random(n2) < 0.8).astype(np.uint8)
    a1 = rng.uniform(0, 360, n1).astype(np.float32)
    a2 = rng.uniform(0, 360, n2).astype(np.float32)
    for chk in (False, True):
        want, nw = oracle.match_bruteforce(d1, d2, valid, a1, a2, 0.8, chk)
        pairs, got = match.robust(0.8, chk).brute_force_match(d1, d2, valid, a1, a2)
        assert np.array_equal(got, want)
        assert len(pairs) == nw
    if n1 >= 1000:
        assert nw > 10


def test_bruteforce_exhaustion(oracle):
    """every keyframe row is a near copy of the same 3 frame descriptors: top-4 lists run dry and the exact rescan runs"""
    from lpslam_b200 import match
    rng = np.random.default_rng(7)
    d1 = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    d1[3:40] = d1[0] ^ np.packbits(rng.random((37, 256)) < 0.03, axis=1, bitorder="little")
    d2 = d1[rng.integers(0, 3, 200)] ^ np.packbits(rng.random((200, 256)) < 0.02, axis=1, bitorder="little")
    for ratio in (0.8, 1.0, 3.0):
        want, _ = oracle.match_bruteforce(d1, d2, None, None, None, ratio, False)
        _, got = match.robust(ratio, False).brute_force_match(d1, d2)
        assert np.array_equal(got, want)


def test_bruteforce_on_extracted_frames(oracle):
    from lpslam_b200 import match
    a = synth_image(480, 640, 50)
    b = np.roll(a, (3, 5), axis=(0, 1))
    o = oracle.extractor(max_kpts=1000)
    ka, da = o.extract(a)
    kb, db = o.extract(b)
    want, nw = oracle.match_bruteforce(da, db, None, ka["angle"], kb["angle"], 0.8, True)
    _, got = match.robust(0.8, True).brute_force_match(da, db, None, ka["angle"], kb["angle"])
    assert nw > 50 and np.array_equal(got, want)


def projection_case(oracle, seed, stereo, with_temporaries):
    from oracle import pyoracle
    rng = np.random.default_rng(seed)
    img = synth_image(480, 640, 60 + seed)
    o = oracle.extractor(max_kpts=1000)
    kp, desc = o.extract(img)
    n = len(kp)
    fx = fy = 500.0
    cx, cy = 320.0, 240.0
    args = dict(fx=fx, fy=fy, cx=cx, cy=cy, cols=640, rows=480, focal_x_baseline=50.0 if stereo else 0.0, true_baseline=0.1,
                is_monocular=not stereo)
    depth = rng.uniform(2.0, 8.0, n)
    pos = np.stack([(kp["x"] - cx) / fx * depth, (kp["y"] - cy) / fy * depth, depth], axis=1)  # world = last camera
    Tl = np.eye(4)
    Tc = np.eye(4)
    Tc[:3, 3] = [0.02, -0.01, -0.3 if stereo else 0.05]   # forward motion for the stereo case
    has = (rng.random(n) < 0.85).astype(np.uint8)
    outl = (rng.random(n) < 0.05).astype(np.uint8)
    ldesc = desc ^ np.packbits(rng.random((n, 256)) < 0.05, axis=1, bitorder="little")
    xr = np.where(rng.random(n) < 0.5, kp["x"] - 50.0 / depth, -1.0).astype(np.float32) if stereo else np.full(n, -1.0, np.float32)
    taken = (rng.random(n) < 0.1).astype(np.uint8)
    obs = (rng.random(n) < 0.5).astype(np.uint8) if with_temporaries else None
    sf = o.table(0)
    return args, Tc, Tl, has, outl, pos, ldesc, kp["octave"], kp["angle"], kp, desc, xr, taken, sf, obs


@pytest.mark.parametrize("seed,stereo,temps", [(1, False, False), (2, True, False), (3, True, True), (4, False, True)])
def test_projection(oracle, seed, stereo, temps):
    from lpslam_b200 import match
    from oracle import pyoracle
    args, Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc, xr, taken, sf, obs = projection_case(oracle, seed, stereo, temps)
    for margin in (10.0, 20.0, 40.0):
        for chk in (True, False):
            want, nw = oracle.match_projection(pyoracle.make_camera(**args), Tc, Tl, has, outl, pos, ldesc, loct, lang, kp, desc,
                                               xr, taken, sf, margin, chk, obs)
            got, ng = match.projection(0.9, chk).match_current_and_last_frames(match.make_camera(**args), Tc, Tl, has, outl, pos,
                                                                              ldesc, loct, lang, kp, desc, xr, taken, sf, margin,
                                                                              obs)
            assert np.array_equal(got, want) and ng == nw
    assert nw > 50


def landmarks_case(oracle, seed, stereo, with_temporaries):
    """local map for match_frame_and_landmarks: up to three landmarks per frame keypoint (so that landmarks fight for
    keypoints and the ratio test sees near-duplicates), reprojections jittered around the keypoints, levels at or above
    the keypoint octave, plus unrelated landmarks anywhere in the image"""
    rng = np.random.default_rng(seed)
    img = synth_image(480, 640, 80 + seed)
    o = oracle.extractor(max_kpts=1000)
    kp, desc = o.extract(img)
    n = len(kp)
    args = dict(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480, focal_x_baseline=50.0 if stereo else 0.0,
                true_baseline=0.1, is_monocular=not stereo)
    src = np.concatenate([np.arange(n), rng.integers(0, n, n), rng.integers(0, n // 10, n // 2)])
    rng.shuffle(src)
    m = len(src)
    reproj = np.stack([kp["x"][src], kp["y"][src]], axis=1).astype(np.float64) + rng.normal(0, 2.0, (m, 2))
    level = np.clip(kp["octave"][src] + rng.integers(0, 2, m), 0, 7).astype(np.int32)
    ldesc = desc[src] ^ np.packbits(rng.random((m, 256)) < 0.06, axis=1, bitorder="little")
    extra = 200
    reproj = np.concatenate([reproj, rng.uniform([-20, -20], [660, 500], (extra, 2))])
    level = np.concatenate([level, rng.integers(0, 8, extra).astype(np.int32)])
    ldesc = np.concatenate([ldesc, rng.integers(0, 256, (extra, 32), dtype=np.uint8)])
    m += extra
    active = (rng.random(m) < 0.9).astype(np.uint8)
    depth = rng.uniform(2.0, 8.0, n)
    xr = np.where(rng.random(n) < 0.5, kp["x"] - 50.0 / depth, -1.0).astype(np.float32) if stereo else np.full(n, -1.0, np.float32)
    lxr = (reproj[:, 0] - rng.uniform(3.0, 30.0, m)).astype(np.float32)
    lxr[: len(src)] = np.where(xr[src] > 0, xr[src] + rng.normal(0, 3.0, len(src)), lxr[: len(src)]).astype(np.float32)
    taken = (rng.random(n) < 0.1).astype(np.uint8)
    obs = (rng.random(m) < 0.5).astype(np.uint8) if with_temporaries else None
    return args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, o.table(0), obs


@pytest.mark.parametrize("seed,stereo,temps", [(1, False, False), (2, True, False), (3, True, True), (4, False, True)])
def test_frame_and_landmarks(oracle, seed, stereo, temps):
    # match/projection.cc:11-83; margins as the tracker uses them (module/tracking_module.cc search_local_landmarks: 5 / 20)
    from lpslam_b200 import match
    from oracle import pyoracle
    args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = landmarks_case(oracle, seed, stereo, temps)
    total = 0
    for margin in (5.0, 20.0):
        for ratio in (0.8, 0.6):
            want, nw = oracle.match_landmarks(pyoracle.make_camera(**args), active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf,
                                              margin, ratio, obs)
            got, ng = match.projection(ratio, True).match_frame_and_landmarks(match.make_camera(**args), active, reproj, lxr, level,
                                                                              ldesc, kp, desc, xr, taken, sf, margin, obs)
            assert np.array_equal(got, want) and ng == nw
            total += nw
    assert total > 400


def test_frame_and_landmarks_edges():
    from lpslam_b200 import match
    cam = match.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    sf = (1.2 ** np.arange(8)).astype(np.float32)
    kp = np.zeros(0, match.KP_DTYPE)
    out, n = match.projection(0.8, True).match_frame_and_landmarks(cam, np.ones(3, np.uint8), np.zeros((3, 2)), np.zeros(3, np.float32),
                                                                   np.zeros(3, np.int32), np.zeros((3, 32), np.uint8), kp,
                                                                   np.zeros((0, 32), np.uint8), np.zeros(0, np.float32),
                                                                   np.zeros(0, np.uint8), sf, 5.0)
    assert n == 0 and len(out) == 0
    kp = np.zeros(2, match.KP_DTYPE)
    kp["x"], kp["y"] = [100.0, 300.0], [100.0, 200.0]
    out, n = match.projection(0.8, True).match_frame_and_landmarks(cam, np.zeros(0, np.uint8), np.zeros((0, 2)), np.zeros(0, np.float32),
                                                                   np.zeros(0, np.int32), np.zeros((0, 32), np.uint8), kp,
                                                                   np.zeros((2, 32), np.uint8), np.full(2, -1, np.float32),
                                                                   np.zeros(2, np.uint8), sf, 5.0)
    assert n == 0 and (out == -1).all()


@pytest.mark.parametrize("dist", [(-0.17, 0.027, 0.0005, -0.0003, 0.0), (0.05, -0.01, 0.0, 0.0, 0.002), (0.0, 0.0, 0.0, 0.0, 0.0)])
def test_frame_postprocess(oracle, dist):
    # data/frame.cc:37-51: undistort (cv::undistortPoints arithmetic, oracle pinned against cv2), bearings, grid
    from lpslam_b200 import match
    from oracle import pyoracle
    o = oracle.extractor(max_kpts=2000)
    kp, _ = o.extract(synth_image(720, 1280, 91))
    fx, fy, cx, cy = 700.0, 705.0, 640.0, 360.0
    want_u = oracle.undistort_keypoints(fx, fy, cx, cy, dist, kp)
    bounds = dict(min_x=float(want_u["x"].min()) + 5.0, max_x=float(want_u["x"].max()) - 5.0, min_y=-3.0, max_y=723.0)  # some keypoints fall off the grid
    args = dict(fx=fx, fy=fy, cx=cx, cy=cy, cols=1280, rows=720)
    ocam, cam = pyoracle.make_camera(**args), match.make_camera(**args)
    for c in (ocam, cam):
        c.min_x, c.max_x, c.min_y, c.max_y = bounds["min_x"], bounds["max_x"], bounds["min_y"], bounds["max_y"]
        c.inv_cell_width = c.grid_cols / (c.max_x - c.min_x)
        c.inv_cell_height = c.grid_rows / (c.max_y - c.min_y)
    want_b = oracle.keypoints_to_bearings(fx, fy, cx, cy, want_u)
    want_s, want_i = oracle.assign_grid(ocam, want_u)
    model = match.Perspective(fx, fy, cx, cy, (match.C.c_double * 5)(*dist))
    und, bearings, start, items = match.frame_postprocess(model, cam, kp)
    for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(und[f], want_u[f]), f
    assert np.array_equal(bearings, want_b)          # double, bit-exact
    assert np.array_equal(start, want_s) and np.array_equal(items, want_i)
    assert 0 < len(items) < len(kp)
    und0, b0, s0, i0 = match.frame_postprocess(model, cam, kp[:0])
    assert len(und0) == 0 and s0[-1] == 0 and len(i0) == 0


def test_landmark_descriptors(oracle):
    # data/landmark.cc:127-183: observation lists of 0..300 descriptors, clustered so that medians tie often
    from lpslam_b200 import match
    rng = np.random.default_rng(12)
    sizes = list(rng.integers(1, 12, 300)) + [0, 1, 2, 3, 32, 33, 64, 65, 100, 300, 0]
    chunks = []
    for n in sizes:
        base = rng.integers(0, 256, (1, 32), dtype=np.uint8)
        flips = rng.random((n, 256)) < rng.choice([0.0, 0.02, 0.1])
        chunks.append(base.repeat(n, 0) ^ np.packbits(flips, axis=1, bitorder="little"))
    descs = np.concatenate(chunks)
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
    best, out = match.compute_landmark_descriptors(descs, offsets)
    for l, n in enumerate(sizes):
        if n == 0:
            assert best[l] == 0 and not out[l].any()
            continue
        want = oracle.landmark_descriptor(descs[offsets[l]:offsets[l + 1]])
        assert best[l] == want, (l, n)
        assert np.array_equal(out[l], descs[offsets[l] + want])


@pytest.mark.parametrize("seed,chk", [(1, True), (2, False), (3, True)])
def test_area(oracle, seed, chk):
    # match/area.cc:8-112 (monocular initializer): frame 2 = frame 1 shifted, plus decoys that fight for the same keypoints
    from lpslam_b200 import match
    from oracle import pyoracle
    rng = np.random.default_rng(seed)
    a = synth_image(480, 640, 50 + seed)
    b = np.roll(a, (3, 7), axis=(0, 1))
    o = oracle.extractor(max_kpts=1000)
    k1, d1 = o.extract(a)
    k2, d2 = o.extract(b)
    args = dict(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)   # initializer: previous match = own position
    total = 0
    for margin in (20, 100):
        for ratio in (0.9, 0.7):
            want, nw, pw = oracle.match_area(pyoracle.make_camera(**args), k1, d1, prev, k2, d2, margin, ratio, chk)
            got, ng, pg = match.area(ratio, chk).match_in_consistent_area(match.make_camera(**args), k1, d1, k2, d2, prev, margin)
            assert np.array_equal(got, want) and ng == nw and np.array_equal(pg, pw)
            total += nw
    assert total > 100
    # second round from the updated positions, with near-duplicate descriptors in frame 1 (take-overs, area.cc:76-80)
    d1b = d1.copy()
    d1b[1::2] = d1b[::2][: len(d1b[1::2])] ^ np.packbits(rng.random((len(d1b[1::2]), 256)) < 0.01, axis=1, bitorder="little")
    k1b = k1.copy()
    k1b["x"][1::2], k1b["y"][1::2] = k1b["x"][::2][: len(k1b[1::2])], k1b["y"][::2][: len(k1b[1::2])]
    k1b["octave"][1::2] = k1b["octave"][::2][: len(k1b[1::2])]
    prevb = np.stack([k1b["x"], k1b["y"]], axis=1).astype(np.float32)
    want, nw, pw = oracle.match_area(pyoracle.make_camera(**args), k1b, d1b, prevb, k2, d2, 50, 0.95, chk)
    got, ng, pg = match.area(0.95, chk).match_in_consistent_area(match.make_camera(**args), k1b, d1b, k2, d2, prevb, 50)
    assert np.array_equal(got, want) and ng == nw and np.array_equal(pg, pw) and nw > 20


@pytest.mark.parametrize("seed,stereo", [(5, False), (6, True)])
def test_frame_and_keyframe_and_knobs(oracle, seed, stereo):
    # match/projection.cc:210-312 (window level-1..level+1, caller's threshold, no ratio, angle check) and every other
    # combination of the shared loop's knobs
    from lpslam_b200 import match
    from oracle import pyoracle
    args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = landmarks_case(oracle, seed, stereo, True)
    rng = np.random.default_rng(seed)
    lm_angle = rng.uniform(0, 360, len(active)).astype(np.float32)
    near = rng.random(len(active)) < 0.7
    src_angle = np.interp(np.arange(len(active)), np.arange(len(kp)), kp["angle"]).astype(np.float32)
    lm_angle = np.where(near, src_angle, lm_angle).astype(np.float32)
    ocam, cam = pyoracle.make_camera(**args), match.make_camera(**args)
    total = 0
    for margin, thr in ((10.0, 100), (15.0, 50)):
        for chk in (True, False):
            want, nw = oracle.match_reprojected(ocam, margin, 0.0, thr, -1, 1, 0, 0, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf,
                                                None, lm_angle if chk else None)
            got, ng = match.projection(0.6, chk).match_frame_and_keyframe(cam, active, reproj, level, ldesc, lm_angle, kp, desc, taken, sf,
                                                                          margin, thr)
            assert np.array_equal(got, want) and ng == nw
            total += nw
    assert total > 200
    for lo, hi, use_ratio, use_xr, thr in ((0, 0, 1, 0, 80), (-2, 2, 0, 1, 100), (-1, 0, 1, 1, 30), (0, 1, 1, 1, 256)):
        prm = match.ReprojMatchParams(8.0, 0.75, thr, lo, hi, use_ratio, use_xr)
        want, nw = oracle.match_reprojected(ocam, 8.0, 0.75, thr, lo, hi, use_ratio, use_xr, active, reproj, lxr, level, ldesc, kp, desc, xr,
                                            taken, sf, obs, lm_angle)
        got, ng = match.match_reprojected(cam, prm, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs, lm_angle)
        assert np.array_equal(got, want) and ng == nw


def bow_vector(rng, n, n_nodes):
    """random vocabulary-node assignment of n keypoints as (node_ids ascending, node_start, items in ascending keypoint order)"""
    node = rng.integers(0, n_nodes, n)
    ids = np.unique(node)
    items = np.concatenate([np.nonzero(node == v)[0] for v in ids]).astype(np.uint32)
    start = np.concatenate([[0], np.cumsum([(node == v).sum() for v in ids])]).astype(np.uint32)
    return (ids.astype(np.uint32), start, items), node


@pytest.mark.parametrize("seed,chk", [(1, True), (2, False), (3, True)])
def test_bow_tree(oracle, seed, chk):
    # match/bow_tree.cc:16-126: keyframe = frame descriptors with noise, nodes mostly shared so that matches exist, few
    # nodes so that rows fight for keypoints
    from lpslam_b200 import match
    rng = np.random.default_rng(seed)
    o = oracle.extractor(max_kpts=1000)
    kp, desc = o.extract(synth_image(480, 640, 70 + seed))
    n = len(kp)
    perm = rng.permutation(n)
    kdesc = desc[perm] ^ np.packbits(rng.random((n, 256)) < 0.04, axis=1, bitorder="little")
    kang = (kp["angle"][perm] + rng.normal(0, 3.0, n)).astype(np.float32) % 360
    fvec, fnode = bow_vector(rng, n, 60)
    knode = np.where(rng.random(n) < 0.85, fnode[perm], rng.integers(0, 80, n))
    ids = np.unique(knode)
    kvec = (ids.astype(np.uint32), np.concatenate([[0], np.cumsum([(knode == v).sum() for v in ids])]).astype(np.uint32),
            np.concatenate([np.nonzero(knode == v)[0] for v in ids]).astype(np.uint32))
    valid = (rng.random(n) < 0.8).astype(np.uint8)
    total = 0
    for ratio in (0.6, 0.75, 0.9):
        want, nw = oracle.match_bow(kvec, fvec, valid, kdesc, kang, desc, kp["angle"], ratio, chk)
        got, ng = match.bow_tree(ratio, chk).match_frame_and_keyframe(kvec, valid, kdesc, kang, fvec, desc, kp["angle"])
        assert np.array_equal(got, want) and ng == nw
        total += nw
    assert total > 300


def clustered(rng, n, protos, flip):
    """descriptors drawn from a few prototypes with a handful of flipped bits: small distances everywhere, ties and near
    ties, so that the kept four candidates run dry, the lower bounds do not settle rows and the exact rescans run"""
    return protos[rng.integers(0, len(protos), n)] ^ np.packbits(rng.random((n, 256)) < flip, axis=1, bitorder="little")


@pytest.mark.parametrize("seed", [11, 12])
def test_frame_and_landmarks_clustered(oracle, seed):
    from lpslam_b200 import match
    from oracle import pyoracle
    args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = landmarks_case(oracle, seed, False, seed == 12)
    rng = np.random.default_rng(seed)
    protos = rng.integers(0, 256, (5, 32), dtype=np.uint8)
    desc = clustered(rng, len(kp), protos, 0.03)
    ldesc = clustered(rng, len(ldesc), protos, 0.03)
    level = np.clip(kp["octave"][rng.integers(0, len(kp), len(level))], 0, 7).astype(np.int32)
    total = 0
    for margin in (5.0, 30.0):
        for ratio in (0.8, 1.0, 2.0):
            want, nw = oracle.match_landmarks(pyoracle.make_camera(**args), active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf,
                                              margin, ratio, obs)
            got, ng = match.projection(ratio, True).match_frame_and_landmarks(match.make_camera(**args), active, reproj, lxr, level,
                                                                              ldesc, kp, desc, xr, taken, sf, margin, obs)
            assert np.array_equal(got, want) and ng == nw
            total += nw
    assert total > 400


@pytest.mark.parametrize("seed", [21, 22])
def test_area_clustered(oracle, seed):
    from lpslam_b200 import match
    from oracle import pyoracle
    rng = np.random.default_rng(seed)
    o = oracle.extractor(max_kpts=1500)
    k1, _ = o.extract(synth_image(480, 640, seed))
    k2, _ = o.extract(synth_image(480, 640, seed + 1))
    k1["octave"][rng.random(len(k1)) < 0.7] = 0
    k2["octave"][rng.random(len(k2)) < 0.7] = 0
    protos = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    d1, d2 = clustered(rng, len(k1), protos, 0.03), clustered(rng, len(k2), protos, 0.03)
    args = dict(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
    total = 0
    for margin in (30, 120):
        for ratio in (0.9, 1.0, 1.5):
            want, nw, pw = oracle.match_area(pyoracle.make_camera(**args), k1, d1, prev, k2, d2, margin, ratio, True)
            got, ng, pg = match.area(ratio, True).match_in_consistent_area(match.make_camera(**args), k1, d1, k2, d2, prev, margin)
            assert np.array_equal(got, want) and ng == nw and np.array_equal(pg, pw)
            total += nw
    assert total > 200


def test_bruteforce_clustered(oracle):
    from lpslam_b200 import match
    rng = np.random.default_rng(31)
    protos = rng.integers(0, 256, (6, 32), dtype=np.uint8)
    d1, d2 = clustered(rng, 1500, protos, 0.03), clustered(rng, 1700, protos, 0.03)
    for ratio in (0.8, 1.0, 1.6):
        want, nw = oracle.match_bruteforce(d1, d2, None, None, None, ratio, False)
        _, got = match.robust(ratio, False).brute_force_match(d1, d2)
        assert np.array_equal(got, want)
    assert nw > 100


@pytest.mark.parametrize("seed,stereo", [(1, False), (2, True), (3, True)])
def test_fuse(oracle, seed, stereo):
    # match/fuse.cc:69-101 (detect_duplication) and :176-238 (replace_duplication: chi-square gate, unsigned level quirk)
    from lpslam_b200 import match
    from oracle import pyoracle
    args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = landmarks_case(oracle, seed, stereo, False)
    isq = (1.0 / (np.asarray(sf, np.float32) ** 2)).astype(np.float32)
    total = 0
    for margin in (3.0, 4.0, 10.0):
        want, wd, nw = oracle.match_fuse(pyoracle.make_camera(**args), margin, False, False, active, reproj, lxr, level, ldesc, kp, desc, xr,
                                         sf, isq)
        got, gd, ng = match.fuse().detect_duplication(match.make_camera(**args), active, reproj, level, ldesc, kp, desc, sf, margin)
        assert np.array_equal(got, want) and np.array_equal(gd, wd) and ng == nw
        total += nw
        want, wd, nw = oracle.match_fuse(pyoracle.make_camera(**args), margin, True, True, active, reproj, lxr, level, ldesc, kp, desc, xr,
                                         sf, isq)
        got, gd, ng = match.fuse().replace_duplication(match.make_camera(**args), active, reproj, lxr, level, ldesc, kp, desc, xr, sf, isq,
                                                       margin)
        assert np.array_equal(got, want) and np.array_equal(gd, wd) and ng == nw
        assert (got[np.asarray(level) == 0] == -1).all()  # fuse.cc:193 on unsigned levels
        total += nw
    assert total > 500


def test_fuse_edges():
    from lpslam_b200 import match
    cam = match.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    sf = np.array([1.2 ** i for i in range(8)], np.float32)
    kp = np.zeros(0, match.KP_DTYPE)
    best, dist, n = match.fuse().detect_duplication(cam, np.ones(3, np.uint8), np.zeros((3, 2)), np.zeros(3, np.int32),
                                                    np.zeros((3, 32), np.uint8), kp, np.zeros((0, 32), np.uint8), sf, 3.0)
    assert n == 0 and (best == -1).all() and (dist == 256).all()


@pytest.mark.parametrize("seed,chk", [(1, True), (2, False), (3, True)])
def test_bow_tree_match_keyframes(oracle, seed, chk):
    # match/bow_tree.cc:128-250: both sides keyframes, only keypoints with live landmarks take part, result per keyframe-1 keypoint
    from lpslam_b200 import match
    rng = np.random.default_rng(seed)
    o = oracle.extractor(max_kpts=1000)
    kp, desc = o.extract(synth_image(480, 640, 90 + seed))
    n = len(kp)
    perm = rng.permutation(n)
    d1 = desc[perm] ^ np.packbits(rng.random((n, 256)) < 0.04, axis=1, bitorder="little")
    a1 = (kp["angle"][perm] + rng.normal(0, 3.0, n)).astype(np.float32) % 360
    vec2, node2 = bow_vector(rng, n, 50)
    node1 = np.where(rng.random(n) < 0.85, node2[perm], rng.integers(0, 70, n))
    ids = np.unique(node1)
    vec1 = (ids.astype(np.uint32), np.concatenate([[0], np.cumsum([(node1 == v).sum() for v in ids])]).astype(np.uint32),
            np.concatenate([np.nonzero(node1 == v)[0] for v in ids]).astype(np.uint32))
    v1 = (rng.random(n) < 0.8).astype(np.uint8)
    v2 = (rng.random(n) < 0.7).astype(np.uint8)
    total = 0
    for ratio in (0.6, 0.75, 0.95):
        want, nw = oracle.match_bow_keyframes(vec1, vec2, v1, d1, a1, v2, desc, kp["angle"], ratio, chk)
        got, ng = match.bow_tree(ratio, chk).match_keyframes(vec1, v1, d1, a1, vec2, v2, desc, kp["angle"])
        assert np.array_equal(got, want) and ng == nw
        assert (got[v1 == 0] == -1).all() and v2[got[got >= 0]].all()
        total += nw
    assert total > 200


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_match_by_sim3_transform(oracle, seed):
    # match/projection.cc:317-413 (search loop :372-410): landmarks fight for keypoints, unsigned level comparison
    from lpslam_b200 import match
    from oracle import pyoracle
    args, active, reproj, lxr, level, ldesc, kp, desc, xr, taken, sf, obs = landmarks_case(oracle, seed, False, False)
    total = 0
    for margin in (3.0, 7.5, 10.0):
        want, nw = oracle.match_sim3(pyoracle.make_camera(**args), margin, active, reproj, level, ldesc, kp, desc, taken, sf)
        got, ng = match.projection().match_by_Sim3_transform(match.make_camera(**args), active, reproj, level, ldesc, kp, desc, taken,
                                                             sf, margin)
        assert np.array_equal(got, want) and ng == nw
        assert (got[taken != 0] == -1).all()
        hit = got[got >= 0]
        assert (np.asarray(level)[hit] > 0).all()  # :385-390 on unsigned levels: a predicted level of 0 never matches
        total += nw
    assert total > 500


def mutual_side(rng, kp_self, kp_other, desc_other, n_other_levels=8):
    """landmarks of one keyframe (a slot per keypoint): reprojected near a random keypoint of the other keyframe, descriptor =
    that keypoint's with noise, level at or above its octave"""
    n = len(kp_self)
    tgt = rng.integers(0, len(kp_other), n)
    reproj = np.stack([kp_other["x"][tgt], kp_other["y"][tgt]], axis=1).astype(np.float64) + rng.normal(0, 1.5, (n, 2))
    level = np.clip(kp_other["octave"][tgt] + rng.integers(0, 2, n), 0, n_other_levels - 1).astype(np.int32)
    ldesc = desc_other[tgt] ^ np.packbits(rng.random((n, 256)) < 0.08, axis=1, bitorder="little")
    active = (rng.random(n) < 0.8).astype(np.uint8)
    return active, reproj, level, ldesc, tgt


@pytest.mark.parametrize("seed", [1, 2])
def test_match_keyframes_mutually(oracle, seed):
    # match/projection.cc:415-636: both one-way searches (threshold HIGH, unsigned levels) and the cross-check
    from lpslam_b200 import match
    from oracle import pyoracle
    rng = np.random.default_rng(seed)
    o1, o2 = oracle.extractor(max_kpts=1000), oracle.extractor(max_kpts=800)
    kp1, d1 = o1.extract(synth_image(480, 640, 90 + seed))
    kp2, d2 = o2.extract(synth_image(480, 640, 95 + seed))
    sf = o1.table(0)
    a1, r1, l1, ld1, t1 = mutual_side(rng, kp1, kp2, d2)
    a2, r2, l2, ld2, t2 = mutual_side(rng, kp2, kp1, d1)
    # make a share of the pairs mutual: landmark of keypoint t1[i] in keyframe 2 points back at i
    back = rng.random(len(kp1)) < 0.6
    idx = np.nonzero(back)[0]
    r2[t1[idx]] = np.stack([kp1["x"][idx], kp1["y"][idx]], axis=1) + rng.normal(0, 1.0, (len(idx), 2))
    l2[t1[idx]] = np.clip(kp1["octave"][idx] + rng.integers(0, 2, len(idx)), 0, 7)
    ld2[t1[idx]] = d1[idx] ^ np.packbits(rng.random((len(idx), 256)) < 0.05, axis=1, bitorder="little")
    args = dict(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    total = 0
    for margin in (4.0, 7.5):
        want, nw = oracle.match_mutual(pyoracle.make_camera(**args), pyoracle.make_camera(**args), margin, (a1, r1, l1, ld1, kp1, d1, sf),
                                       (a2, r2, l2, ld2, kp2, d2, sf))
        got, ng = match.projection().match_keyframes_mutually(match.make_camera(**args), match.make_camera(**args),
                                                              (a1, r1, l1, ld1, kp1, d1, sf), (a2, r2, l2, ld2, kp2, d2, sf), margin)
        assert np.array_equal(got, want) and ng == nw
        total += nw
    assert total > 100


def triangulation_case(oracle, seed, stereo):
    """two keyframes related by a sideways translation: bearings of keyframe 2 = bearings of keyframe 1 pushed through the
    relative pose with random depths (so that E_12 holds for the true pairs), descriptors with noise, shared BoW nodes"""
    rng = np.random.default_rng(seed)
    o = oracle.extractor(max_kpts=1000)
    kp1, d1 = o.extract(synth_image(480, 640, 60 + seed))
    n = len(kp1)
    fx = fy = 500.0
    cx, cy = 320.0, 240.0
    b1 = np.stack([(kp1["x"] - cx) / fx, (kp1["y"] - cy) / fy, np.ones(n)], axis=1).astype(np.float64)
    b1 /= np.linalg.norm(b1, axis=1, keepdims=True)
    ang = 0.05
    R21 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t21 = np.array([0.3, 0.02, 0.05])
    depth = rng.uniform(2.0, 10.0, n)
    p2 = (R21 @ (b1 * depth[:, None]).T).T + t21
    perm = rng.permutation(n)
    b2 = p2[perm] / np.linalg.norm(p2[perm], axis=1, keepdims=True)
    b2 += rng.normal(0, 1.5e-3, b2.shape)  # some pairs fall outside the 0.2 degree x scale band
    b2 /= np.linalg.norm(b2, axis=1, keepdims=True)
    kp2 = kp1[perm].copy()
    kp2["angle"] = (kp2["angle"] + rng.normal(0, 4.0, n)).astype(np.float32) % 360
    d2 = d1[perm] ^ np.packbits(rng.random((n, 256)) < 0.05, axis=1, bitorder="little")
    # E_12 = [t_12]x R_12 with x_1 = R_12 x_2 + t_12
    R12, t12 = R21.T, -R21.T @ t21
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    E12 = tx @ R12
    c1_in_2 = t21 / np.linalg.norm(t21)  # centre of keyframe 1 seen from keyframe 2, as a bearing
    vec1, node1 = bow_vector(rng, n, 50)
    node2 = np.where(rng.random(n) < 0.9, node1[perm], rng.integers(0, 60, n))
    ids = np.unique(node2)
    vec2 = (ids.astype(np.uint32), np.concatenate([[0], np.cumsum([(node2 == v).sum() for v in ids])]).astype(np.uint32),
            np.concatenate([np.nonzero(node2 == v)[0] for v in ids]).astype(np.uint32))
    h1 = (rng.random(n) < 0.3).astype(np.uint8)
    h2 = (rng.random(n) < 0.3).astype(np.uint8)
    x1 = np.where(rng.random(n) < 0.5, kp1["x"] - 5.0, -1.0).astype(np.float32) if stereo else np.full(n, -1.0, np.float32)
    x2 = np.where(rng.random(n) < 0.5, kp2["x"] - 5.0, -1.0).astype(np.float32) if stereo else np.full(n, -1.0, np.float32)
    return vec1, (kp1, d1, b1, x1, h1, o.table(0)), vec2, (kp2, d2, b2, x2, h2), E12, c1_in_2


@pytest.mark.parametrize("seed,stereo,chk", [(1, False, True), (2, True, False), (3, True, True)])
def test_match_for_triangulation(oracle, seed, stereo, chk):
    # match/robust.cc:18-178: epipole gate, epipolar constraint, `<=` replacement of the best, keypoints consumed in order
    from lpslam_b200 import match
    vec1, s1, vec2, s2, E12, epi = triangulation_case(oracle, seed, stereo)
    want, nw = oracle.match_triangulation(vec1, vec2, s1, s2, E12, epi, chk)
    pairs, got = match.robust(0.6, chk).match_for_triangulation(vec1, s1, vec2, s2, E12, epi)
    assert np.array_equal(got, want) and len(pairs) == nw
    assert nw > 150
    assert (got[s1[4] != 0] == -1).all() and (s2[4][got[got >= 0]] == 0).all()  # only keypoints without landmarks
    # the epipole itself as a bearing: nothing near it may match when neither side is a stereo keypoint
    want2, nw2 = oracle.match_triangulation(vec1, vec2, s1, s2, E12, s2[2][0], chk)
    _, got2 = match.robust(0.6, chk).match_for_triangulation(vec1, s1, vec2, s2, E12, s2[2][0])
    assert np.array_equal(got2, want2)


def test_match_for_triangulation_last_equal_wins(oracle):
    # robust.cc:108: `best_hamm_dist < hamm_dist` lets a later candidate at the SAME distance replace the best
    from lpslam_b200 import match
    from lpslam_b200.orb import KP_DTYPE
    rng = np.random.default_rng(5)
    n2 = 12
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    d1 = base[None].copy()
    d2 = np.repeat(base[None], n2, axis=0)
    for i in range(n2):  # distances 3, 2, 2, 3, 2, ... : the LAST of the distance-2 candidates must win
        flips = 2 if i % 3 else 3
        d2[i, :flips] ^= 1
    kp1 = np.zeros(1, KP_DTYPE)
    kp2 = np.zeros(n2, KP_DTYPE)
    b1 = np.array([[0.0, 0.0, 1.0]])
    b2 = np.repeat(np.array([[0.0, 0.0, 1.0]]), n2, axis=0)
    E = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])  # t = x: bearings along z satisfy it exactly
    vec1 = (np.array([7], np.uint32), np.array([0, 1], np.uint32), np.array([0], np.uint32))
    vec2 = (np.array([7], np.uint32), np.array([0, n2], np.uint32), np.arange(n2, dtype=np.uint32))
    s1 = (kp1, d1, b1, np.array([-1.0], np.float32), np.zeros(1, np.uint8), np.ones(8, np.float32))
    s2 = (kp2, d2, b2, np.full(n2, -1.0, np.float32), np.zeros(n2, np.uint8))
    epi = np.array([1.0, 0.0, 0.0])
    want, nw = oracle.match_triangulation(vec1, vec2, s1, s2, E, epi, False)
    _, got = match.robust(0.6, False).match_for_triangulation(vec1, s1, vec2, s2, E, epi)
    assert nw == 1 and want[0] == 11 and np.array_equal(got, want)


@pytest.mark.parametrize("seed", [31, 32])
def test_match_for_triangulation_clustered(oracle, seed):
    """few prototypes and few vocabulary nodes: long candidate lists full of equal distances, so that the kept four
    candidates run dry, the exact rescans run and the `<=` replacement order decides most rows"""
    from lpslam_b200 import match
    vec1, s1, vec2, s2, E12, epi = triangulation_case(oracle, seed, seed % 2 == 0)
    rng = np.random.default_rng(seed)
    n = len(s1[0])
    protos = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    d1, d2 = clustered(rng, n, protos, 0.01), clustered(rng, n, protos, 0.01)
    vec1, _ = bow_vector(rng, n, 6)
    vec2, _ = bow_vector(rng, n, 6)
    # bearings all close to the axis and E_12 of a pure x translation: most pairs are inliers, the descriptor order decides
    b = np.stack([rng.normal(0, 0.2, n), rng.normal(0, 2e-3, n), np.ones(n)], axis=1)
    b /= np.linalg.norm(b, axis=1, keepdims=True)
    b2 = b[rng.permutation(n)]
    E = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
    s1 = (s1[0], d1, b, s1[3], (rng.random(n) < 0.1).astype(np.uint8), s1[5])
    s2 = (s2[0], d2, b2, s2[3], (rng.random(n) < 0.1).astype(np.uint8))
    want, nw = oracle.match_triangulation(vec1, vec2, s1, s2, E, np.array([1.0, 0.0, 0.0]), True)
    _, got = match.robust(0.6, True).match_for_triangulation(vec1, s1, vec2, s2, E, np.array([1.0, 0.0, 0.0]))
    assert np.array_equal(got, want)
    assert nw > 200


def test_match_consecutive_frames(oracle):
    # robust::brute_force_match between consecutive frames of a device-resident batch == the host-buffer matcher per pair
    from lpslam_b200 import match, orb
    imgs = [np.roll(synth_image(480, 640, 77), (2 * k, 3 * k), axis=(0, 1)) for k in range(5)]
    e = orb.orb_extractor(max_num_keypts=1000, max_batch=5)
    res = e.extract_batch(imgs)
    n_rows = sum(len(k) for k, _ in res)
    got, cnt = match.match_consecutive_frames(e, n_rows, len(imgs), 0.8)
    off = np.concatenate([[0], np.cumsum([len(k) for k, _ in res])])
    assert (got[: off[1]] == -1).all() and cnt[0] == 0
    for f in range(1, len(imgs)):
        want, nw = oracle.match_bruteforce(res[f][1], res[f - 1][1], None, None, None, 0.8, False)
        assert np.array_equal(got[off[f]:off[f + 1]], want) and cnt[f] == nw and nw > 100
