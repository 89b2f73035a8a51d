"""CPU: the oracle port (oracle/*.cc) — and the reference build (oracle/_ref) where it exists — against the committed
golden vectors, which tools/gen_golden.py produced from the reference's own sources. Also the known-answer cases the
reference's unit tests hold for this path (test/openvslam/match/base.cc, match/angle_checker.cc, util/trigonometric.cc,
feature/orb_extractor.cc)."""
import math

import numpy as np
import pytest

from golden_util import EXTRACT_FIXTURES, STEREO_FIXTURES, case_input, crc, load, stereo_input

FX_B, B = 700.0 * 0.12, 0.12


def check_extract(o, name, is_port):
    g = load(name)
    img, mask, params = case_input(g)
    e = o.extractor(**params)
    kp, desc = e.extract(img, mask)
    assert kp.tobytes() == g["kp"].tobytes(), "keypoints differ from the reference"
    assert np.array_equal(desc.reshape(-1, 32), g["desc"].reshape(-1, 32)), "descriptors differ from the reference"
    for l in range(params["num_levels"]):
        assert crc(e.level(l)) == int(g["pyr_crc"][l]), f"pyramid level {l}"
        assert np.array_equal(e.selected(l), g[f"sel_l{l}"]), f"quadtree selection level {l}"
        if is_port:
            c = e.candidates(l)
            assert len(c) == int(g["cand_n"][l]) and crc(c.astype(np.int32)) == int(g["cand_crc"][l]), f"FAST candidates level {l}"
            if int(g["blur_crc"][l]) >= 0:
                assert crc(e.level(l, blurred=True)) == int(g["blur_crc"][l]), f"blurred level {l}"


@pytest.mark.parametrize("name", EXTRACT_FIXTURES)
def test_port_extract(oracle, name):
    check_extract(oracle, name, True)


@pytest.mark.parametrize("name", EXTRACT_FIXTURES)
def test_reference_build_extract(ref_oracle, name):
    check_extract(ref_oracle, name, False)


def check_stereo(o, name):
    g = load(name)
    (left, right), n = stereo_input(g)
    el, er = o.extractor(max_kpts=n), o.extractor(max_kpts=n)
    kl, dl = el.extract(left)
    kr, dr = er.extract(right)
    assert np.array_equal(np.stack([kl["x"], kl["y"]], 1), g["kp_l_xy"]) and np.array_equal(np.stack([kr["x"], kr["y"]], 1), g["kp_r_xy"])
    assert crc(dl) == int(g["desc_l_crc"]) and crc(dr) == int(g["desc_r_crc"])
    xr, dep = o.match_stereo([el.level(l) for l in range(8)], [er.level(l) for l in range(8)], kl, kr, dl, dr,
                             el.table(0), el.table(1), FX_B, B)
    assert np.array_equal(xr, g["stereo_x_right"]) and np.array_equal(dep, g["depths"])


@pytest.mark.parametrize("name", STEREO_FIXTURES)
def test_port_stereo(oracle, name):
    check_stereo(oracle, name)


@pytest.mark.parametrize("name", STEREO_FIXTURES[:1])
def test_reference_build_stereo(ref_oracle, name):
    check_stereo(ref_oracle, name)


def test_toy_corner_property():
    # test/openvslam/feature/orb_extractor.cc:24-49: every keypoint within 2 * scale of the single corner (300, 300)
    g = load("toy_corner_600x600")
    kp = g["kp"]
    assert len(kp) > 0
    sf = np.float32(1.2) ** kp["octave"].astype(np.float32)
    assert (np.abs(kp["x"] - 300) <= 2.0 * sf + 1e-3).all() and (np.abs(kp["y"] - 300) <= 2.0 * sf + 1e-3).all()


def test_mask_properties():
    # test/openvslam/feature/orb_extractor.cc:116-332: no keypoint inside a masked region
    g = load("imgmask_640x480")
    kp = g["kp"]
    inside = (kp["x"] >= 200) & (kp["x"] < 500) & (kp["y"] >= 100) & (kp["y"] < 300)
    assert len(kp) > 0 and not inside.any()
    g = load("rectmask_640x480")
    kp = g["kp"]
    for x0, x1, y0, y1 in g["p_mask_rects"]:
        inside = (kp["x"] > 640 * x0 + 1) & (kp["x"] < 640 * x1 - 1) & (kp["y"] > 480 * y0 + 1) & (kp["y"] < 480 * y1 - 1)
        assert not inside.any()


def test_hamming_known_answers(oracle):
    # test/openvslam/match/base.cc:11-57
    cases = [(0b01010101, 0b01010101, 0), (0b01010101, 0b10101010, 256), (0b01100110, 0b00111100, 128)]
    for a, b, want in cases:
        da, db = np.full(32, a, np.uint8), np.full(32, b, np.uint8)
        assert oracle.hamming(da, db, 32) == want and oracle.hamming(da, db, 64) == want


def test_trigonometric_tolerance(oracle):
    # test/openvslam/util/trigonometric.cc:8-20
    for deg in range(0, 3601):
        rad = float(np.float32(np.float32(deg / 10.0) * np.float32(math.pi) / np.float32(180.0)))
        assert abs(math.cos(rad) - oracle.lib.orc_cos(rad)) < 1e-3
        assert abs(math.sin(rad) - oracle.lib.orc_sin(rad)) < 1e-3


def test_angle_checker_top3(oracle):
    # test/openvslam/match/angle_checker.cc:106-145 (histogram length 30, three fullest bins kept)
    deltas = [34.8, 34.9, 35.0, 35.1, 35.2, 323.8, 323.9, 324.1, 324.2, 126.9, 127.0, 127.1, 0.0, 90.0, 180.0, 270.0]
    ids = [35] * 5 + [324] * 4 + [127] * 3 + [0, 60, 180, 270]
    inv = oracle.angle_invalid(deltas, ids)
    assert sorted(inv.tolist()) == [0, 60, 180, 270]


def test_landmark_matcher_semantics(oracle):
    """match/projection.cc:11-83 hand-built cases for the port: strict '<' updates, the ratio test only between
    candidates of the same scale level, keypoints claimed by earlier landmarks, the x_right gate."""
    from oracle import pyoracle
    cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    kp = np.zeros(3, pyoracle.KP_DTYPE)
    kp["x"], kp["y"], kp["octave"] = [100.0, 102.0, 300.0], [100.0, 101.0, 200.0], [1, 1, 0]
    d = np.zeros((3, 32), np.uint8)
    d[1, 0] = 0x03          # keypoint 1 is 2 bits from keypoint 0
    d[2] = 0xFF
    none = np.full(3, -1.0, np.float32)
    free = np.zeros(3, np.uint8)

    def run(lm_desc, reproj, level, ratio=0.8, taken=free, xr=none, lxr=None, active=None, obs=None):
        m = len(level)
        return oracle.match_landmarks(cam, np.ones(m, np.uint8) if active is None else active, reproj, np.zeros(m, np.float32) if lxr is None else lxr,
                                      level, lm_desc, kp, d, xr, taken, sf, 5.0, ratio, obs)

    one = np.zeros((1, 32), np.uint8)
    # best = keypoint 0 (distance 0), second = keypoint 1 (distance 2) on the same level: 0 > 0.8 * 2 is false -> match
    out, n = run(one, [[100.0, 100.0]], [1])
    assert n == 1 and list(out) == [0, -1, -1]
    # landmark 1 bit from keypoint 0 and 1 bit from keypoint 1 -> tie, first visited wins, ratio 1 > 0.8 * 1 rejects
    tie = one.copy()
    tie[0, 0] = 0x01
    out, n = run(tie, [[100.0, 100.0]], [1])
    assert n == 0
    # same, but the second best sits on another level: no ratio test
    kp["octave"][1] = 0
    out, n = run(tie, [[100.0, 100.0]], [1])
    assert n == 1 and out[0] == 0
    kp["octave"][1] = 1
    # two identical landmarks: the first claims keypoint 0, the second falls to keypoint 1
    out, n = run(np.zeros((2, 32), np.uint8), [[100.0, 100.0], [100.0, 100.0]], [1, 1])
    assert n == 2 and list(out) == [0, 1, -1]
    # ... unless the first landmark has no observation: then the second one overwrites keypoint 0
    out, n = run(np.zeros((2, 32), np.uint8), [[100.0, 100.0], [100.0, 100.0]], [1, 1], obs=np.array([0, 1], np.uint8))
    assert n == 2 and list(out) == [1, -1, -1]
    # level window is [level - 1, level]: a level-3 landmark does not see octave-1 keypoints
    out, n = run(one, [[100.0, 100.0]], [3])
    assert n == 0
    # x_right gate: |x_right_lm - x_right_kp| > margin * scale drops keypoint 0, keypoint 1 (no stereo) stays
    xr = np.array([90.0, -1.0, -1.0], np.float32)
    out, n = run(one, [[100.0, 100.0]], [1], xr=xr, lxr=np.array([70.0], np.float32))
    assert n == 1 and out[1] == 0
    # inactive landmarks are skipped
    out, n = run(one, [[100.0, 100.0]], [1], active=np.zeros(1, np.uint8))
    assert n == 0


def test_fuse_search_semantics(oracle):
    """match/fuse.cc:69-101 / :176-238 hand-built cases for the port: rows are independent (two landmarks may name the same
    keypoint), octave window [level - 1, level], HAMMING_DIST_THR_LOW, replace_duplication's chi-square gate and its
    unsigned level arithmetic."""
    from oracle import pyoracle
    cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    isq = (1.0 / sf ** 2).astype(np.float32)
    kp = np.zeros(3, pyoracle.KP_DTYPE)
    kp["x"], kp["y"], kp["octave"] = [100.0, 102.0, 300.0], [100.0, 101.0, 200.0], [1, 0, 2]
    d = np.zeros((3, 32), np.uint8)
    d[1, 0] = 0x03
    d[2] = 0xFF
    xr = np.array([-1.0, 90.0, -1.0], np.float32)

    def run(lm_desc, reproj, level, chi=False, uns=False, lxr=None):
        m = len(level)
        return oracle.match_fuse(cam, 3.0, chi, uns, np.ones(m, np.uint8), reproj, np.zeros(m, np.float32) if lxr is None else lxr, level,
                                 lm_desc, kp, d, xr, sf, isq)

    two = np.zeros((2, 32), np.uint8)
    best, dist, n = run(two, [[100.0, 100.0], [100.5, 100.0]], [1, 1])
    assert n == 2 and list(best) == [0, 0] and list(dist) == [0, 0]          # nothing is marked as taken
    best, dist, n = run(two[:1], [[100.0, 100.0]], [3])
    assert n == 0 and best[0] == -1 and dist[0] == 256                      # octaves 1 and 0 are outside [2, 3]
    far = np.full((1, 32), 0x0F, np.uint8)                                   # 128 bits from keypoint 0, 126 from keypoint 1
    best, dist, n = run(far, [[100.0, 100.0]], [1])
    assert n == 0 and best[0] == -1 and dist[0] == 126                      # best_dist (keypoint 1) is reported, the match is not
    # chi-square gate: keypoint 0 (mono, 2 dof) sits 2.9 px away: 8.41 / 1.44 = 5.84 < 5.99 passes, 3.0 px -> 6.25 rejects
    best, _, _ = run(two[:1], [[102.9, 100.0]], [1], chi=True)
    assert best[0] == 0
    best, dist, _ = run(two[:1], [[103.0, 100.0]], [1], chi=True)
    assert best[0] != 0
    # keypoint 1 has a stereo match (3 dof, x_right 90): an x_right error of 2.5 px alone gives 6.25 < 7.81 at level 0
    best, _, _ = run(two[:1], [[102.0, 101.0]], [0], chi=True, lxr=np.array([92.5], np.float32))
    assert best[0] == 1
    best, _, _ = run(two[:1], [[102.0, 101.0]], [0], chi=True, lxr=np.array([93.0], np.float32))
    assert best[0] == -1
    # replace_duplication compares unsigned levels: a predicted level of 0 rejects every candidate (fuse.cc:193)
    best, _, n = run(two[:1], [[102.0, 101.0]], [0], chi=True, uns=True, lxr=np.array([90.0], np.float32))
    assert n == 0 and best[0] == -1


def test_loop_closing_search_semantics(oracle):
    """match/projection.cc:317-413 (match_by_Sim3_transform) and :415-636 (match_keyframes_mutually), hand-built cases for the
    port: keypoints are consumed in row order, HAMMING_DIST_THR_LOW vs _HIGH, the unsigned level comparison, the cross-check."""
    from oracle import pyoracle
    cam = pyoracle.make_camera(fx=500.0, fy=500.0, cx=320.0, cy=240.0, cols=640, rows=480)
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    kp = np.zeros(3, pyoracle.KP_DTYPE)
    kp["x"], kp["y"], kp["octave"] = [100.0, 101.0, 300.0], [100.0, 100.0, 200.0], [1, 1, 0]
    d = np.zeros((3, 32), np.uint8)
    d[1, 0] = 0x01
    free = np.zeros(3, np.uint8)
    two = np.zeros((2, 32), np.uint8)
    # both rows prefer keypoint 0; the second finds it matched and falls back to keypoint 1 (distance 1)
    out, n = oracle.match_sim3(cam, 5.0, np.ones(2, np.uint8), [[100.0, 100.0], [100.0, 100.0]], [1, 1], two, kp, d, free, sf)
    assert n == 2 and list(out) == [0, 1, -1]
    # a keypoint that already holds a matched landmark is never a candidate
    out, n = oracle.match_sim3(cam, 5.0, np.ones(1, np.uint8), [[100.0, 100.0]], [1], two[:1], kp, d, np.array([1, 0, 0], np.uint8), sf)
    assert n == 1 and list(out) == [-1, 0, -1]
    # threshold LOW = 50: 51 differing bits are rejected
    far = np.zeros((1, 32), np.uint8)
    far[0, :6] = 0xFF
    far[0, 6] = 0x07
    out, n = oracle.match_sim3(cam, 5.0, np.ones(1, np.uint8), [[100.0, 100.0]], [1], far, kp, d, np.array([0, 1, 0], np.uint8), sf)
    assert n == 0
    # predicted level 0: `scale_level < 0u - 1` holds for every keypoint, nothing matches (keypoint 2 has octave 0)
    out, n = oracle.match_sim3(cam, 5.0, np.ones(1, np.uint8), [[300.0, 200.0]], [0], two[:1], kp, d, free, sf)
    assert n == 0
    out, n = oracle.match_sim3(cam, 5.0, np.ones(1, np.uint8), [[300.0, 200.0]], [1], two[:1], kp, d, free, sf)
    assert n == 1 and out[2] == 0

    # mutual: keyframe 1 has 2 keypoints, keyframe 2 has 2; 1/0 <-> 2/1 agree, 1/1 -> 2/1 is one-sided
    kp1 = np.zeros(2, pyoracle.KP_DTYPE)
    kp1["x"], kp1["y"], kp1["octave"] = [50.0, 400.0], [60.0, 300.0], [1, 1]
    kp2 = np.zeros(2, pyoracle.KP_DTYPE)
    kp2["x"], kp2["y"], kp2["octave"] = [500.0, 200.0], [100.0, 150.0], [1, 1]
    d1 = np.zeros((2, 32), np.uint8)
    d2 = np.zeros((2, 32), np.uint8)
    side_1 = (np.ones(2, np.uint8), [[200.0, 150.0], [200.5, 150.0]], [1, 1], np.zeros((2, 32), np.uint8), kp1, d1, sf)
    side_2 = (np.ones(2, np.uint8), [[0.0, 0.0], [50.0, 60.0]], [1, 1], np.zeros((2, 32), np.uint8), kp2, d2, sf)
    out, n = oracle.match_mutual(cam, cam, 5.0, side_1, side_2)
    assert n == 1 and list(out) == [1, -1]
    # threshold HIGH = 100: a landmark descriptor 100 bits away still matches, 101 does not
    lm = np.zeros((2, 32), np.uint8)
    lm[0, :12] = 0xFF
    lm[0, 12] = 0x0F
    out, n = oracle.match_mutual(cam, cam, 5.0, (side_1[0], side_1[1], side_1[2], lm, kp1, d1, sf), side_2)
    assert n == 1
    lm[0, 12] = 0x1F
    out, n = oracle.match_mutual(cam, cam, 5.0, (side_1[0], side_1[1], side_1[2], lm, kp1, d1, sf), side_2)
    assert n == 0


def test_triangulation_search_semantics(oracle):
    """match/robust.cc:18-178 hand-built cases for the port: an equal distance replaces the best (:108), the 0.2 degree x scale
    band of check_epipolar_constraint with its one-sided residual (:329-340), the epipole gate for mono pairs (:112-123)."""
    from oracle import pyoracle
    kp1 = np.zeros(1, pyoracle.KP_DTYPE)
    kp2 = np.zeros(3, pyoracle.KP_DTYPE)
    d1 = np.zeros((1, 32), np.uint8)
    d2 = np.zeros((3, 32), np.uint8)
    d2[0, 0] = 0x01
    d2[1, 0] = 0x02
    d2[2, 0] = 0x07
    z = np.array([0.0, 0.0, 1.0])
    E = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])  # [t]x with t = x, R = I
    vec1 = (np.array([3], np.uint32), np.array([0, 1], np.uint32), np.array([0], np.uint32))
    vec2 = (np.array([3], np.uint32), np.array([0, 3], np.uint32), np.arange(3, dtype=np.uint32))
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    mono1, mono2 = np.array([-1.0], np.float32), np.full(3, -1.0, np.float32)

    def run(b1, b2, epi=(1.0, 0.0, 0.0), x1=mono1, x2=mono2, h2=np.zeros(3, np.uint8)):
        return oracle.match_triangulation(vec1, vec2, (kp1, d1, np.asarray([b1], np.float64), x1, np.zeros(1, np.uint8), sf),
                                          (kp2, d2, np.asarray(b2, np.float64), x2, h2), E, np.asarray(epi, np.float64), False)

    out, n = run(z, [z, z, z])
    assert n == 1 and out[0] == 1                      # distances 1, 1, 3: the later of the two equal ones
    out, n = run(z, [z, z, z], h2=np.array([0, 1, 0], np.uint8))
    assert out[0] == 0                                 # a keypoint with a landmark is not a candidate
    # epipolar plane normal of bearing_2 = z is E z = (0, -1, 0); bearing_1 tilted towards +y by 0.3 degree: the residual
    # pi/2 - acos(-sin) is negative and passes, tilted towards -y it is +0.3 degree > 0.2 degree and fails
    a = np.deg2rad(0.3)
    out, n = run([0.0, np.sin(a), np.cos(a)], [z, z, z])
    assert n == 1
    out, n = run([0.0, -np.sin(a), np.cos(a)], [z, z, z])
    assert n == 0
    a = np.deg2rad(0.15)
    out, n = run([0.0, -np.sin(a), np.cos(a)], [z, z, z])
    assert n == 1
    # epipole along the candidates' bearing: mono-mono pairs within 3 degrees of it are refused, a stereo keypoint is not
    out, n = run(z, [z, z, z], epi=(0.0, 0.0, 1.0))
    assert n == 0
    out, n = run(z, [z, z, z], epi=(0.0, 0.0, 1.0), x2=np.array([-1.0, -1.0, 5.0], np.float32))
    assert n == 1 and out[0] == 2


def test_scale_tables_reference_cases(oracle, ref_oracle):
    """test/openvslam/feature/orb_params.cc:168-213 (calc_scale_factors / calc_inv_scale_factors / calc_level_sigma_sq /
    calc_inv_level_sigma_sq with 10 levels and scale 1.26): on the reference build with EXPECT_FLOAT_EQ's 4-ulp bar, and
    port == reference build bit for bit."""
    def ulps(a, b):
        a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
        return np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64)).max()

    L, sf = 10, np.float32(1.26)
    r, p = ref_oracle.extractor(max_kpts=1000, scale_factor=float(sf), num_levels=L), oracle.extractor(max_kpts=1000, scale_factor=float(sf), num_levels=L)
    for which in range(5):
        assert np.array_equal(r.table(which), p.table(which)), which
    lv = np.arange(L)
    assert ulps(r.table(0), np.power(sf, lv, dtype=np.float32)) <= 4
    assert ulps(r.table(1), np.power(np.float32(1.0) / sf, lv, dtype=np.float32)) <= 4
    s = np.cumprod(np.concatenate([[np.float32(1.0)], np.full(L - 1, sf, np.float32)]), dtype=np.float32)
    assert ulps(r.table(2), s * s) <= 4
    assert ulps(r.table(3), np.float32(1.0) / (s * s)) <= 4
