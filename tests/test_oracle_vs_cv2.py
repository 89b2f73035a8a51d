"""Pin the oracle's restated OpenCV primitives bit-for-bit against cv2 (the third-party arithmetic the reference calls:
cv::resize orb_extractor.cc:273, cv::GaussianBlur :136, cv::FAST :343-350, cv::fastAtan2 :633).
cv2 is only present in the build container; on the GPU box these tests skip and the committed golden vectors
(tests/golden, test_oracle_golden.py) carry the pin."""
import numpy as np
import pytest

from lpslam_b200.synth import noise_image, synth_image

cv2 = pytest.importorskip("cv2")

SIZES = [(480, 640), (720, 1280), (333, 444), (97, 131)]


def level_sizes(h, w, n=8, sf=1.2):
    s = np.float32(1.0)
    out = [(h, w)]
    for _ in range(1, n):
        s = np.float32(sf) * s
        out.append((int(round(h * 1.0 / float(s))), int(round(w * 1.0 / float(s)))))
    return out


@pytest.mark.parametrize("h,w", SIZES)
def test_resize_chain(oracle, h, w):
    img = synth_image(h, w, 1)
    prev = img
    for (lh, lw) in level_sizes(h, w)[1:]:
        if lh < 8 or lw < 8:
            break
        want = cv2.resize(prev, (lw, lh), interpolation=cv2.INTER_LINEAR)
        got = oracle.resize(prev, lw, lh)
        assert np.array_equal(want, got), (lh, lw, int((want != got).sum()))
        prev = want


def test_resize_half(oracle):
    img = noise_image(200, 320, 3)
    want = cv2.resize(img, (160, 100), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(want, oracle.resize(img, 160, 100))


@pytest.mark.parametrize("h,w", SIZES + [(40, 64), (7, 9)])
def test_blur(oracle, h, w):
    for img in (synth_image(h, w, 2), noise_image(h, w, 4)):
        want = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(want, oracle.blur(img))


@pytest.mark.parametrize("thr", [20, 7])
def test_fast(oracle, thr):
    det = cv2.FastFeatureDetector_create(threshold=thr, nonmaxSuppression=True, type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    for img in (synth_image(240, 320, 5), noise_image(70, 70, 6), noise_image(33, 70, 7), synth_image(70, 70, 8)):
        kps = det.detect(img)
        want = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)
        got = oracle.fast(img, thr)
        assert np.array_equal(want, got)


def test_fast_atan2(oracle):
    rng = np.random.default_rng(9)
    ys = rng.integers(-3_000_000, 3_000_000, 5000)
    xs = rng.integers(-3_000_000, 3_000_000, 5000)
    for y, x in list(zip(ys, xs)) + [(0, 0), (0, 5), (5, 0), (-5, 0), (0, -5), (7, 7), (-7, 7)]:
        want = cv2.fastAtan2(float(y), float(x))
        got = oracle.atan2(float(y), float(x))
        assert np.float32(want) == np.float32(got), (y, x, want, got)


def test_rect_fill_matches_line_aa(oracle):
    m = np.full((100, 120), 255, np.uint8)
    cv2.rectangle(m, (12, 20), (60, 71), 0, -1, cv2.LINE_AA)
    # LINE_AA leaves a non-zero anti-aliased fringe; the reference only tests mask == 0 (orb_extractor.cc:281-283),
    # so the set of zero pixels is what matters: exactly the inclusive rectangle.
    want = np.zeros((100, 120), bool)
    want[20:72, 12:61] = True
    assert np.array_equal(m == 0, want)


def test_undistort_points_vs_cv2():
    """camera/perspective.cc:98-130 -> cv::undistortPoints(K, dist, P = K, (EPS | MAX_ITER, 20, 1e-6)): the restatement in
    oracle/frame_oracle.cc is bit-identical to cv2 on lens models whose iteration converges inside the image. (Outside of
    that, e.g. k1 = -0.3 at the image corners, OpenCV releases differ among themselves; the port keeps the loop of the 4.2
    series the reference builds against.)"""
    from oracle import pyoracle
    o = pyoracle.port()
    rng = np.random.default_rng(3)
    K = np.array([[700.0, 0, 640.0], [0, 705.0, 360.0], [0, 0, 1]])
    crit = (cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_MAX_ITER, 20, 1e-6)
    for dist in ([-0.17, 0.027, 0.0005, -0.0003, 0.0], [0.05, -0.01, 0, 0, 0.002], [0.1, 0.0, -0.002, 0.001, 0.0], [0, 0, 0, 0, 0]):
        kp = np.zeros(20000, pyoracle.KP_DTYPE)
        kp["x"], kp["y"] = rng.uniform(0, 1280, 20000), rng.uniform(0, 720, 20000)
        kp["octave"], kp["angle"], kp["size"] = rng.integers(0, 8, 20000), rng.uniform(0, 360, 20000), 31.0
        got = o.undistort_keypoints(700.0, 705.0, 640.0, 360.0, dist, kp)
        src = np.stack([kp["x"], kp["y"]], 1).astype(np.float32).reshape(-1, 1, 2)
        # camera/perspective.cc:21-22: the reference hands OpenCV FLOAT matrices (cv::Mat_<float>) for K and the distortion vector
        Kf = K.astype(np.float32)
        want = cv2.undistortPointsIter(src, Kf, np.array(dist, np.float32), None, Kf, crit).reshape(-1, 2)
        assert np.array_equal(got["x"], want[:, 0]) and np.array_equal(got["y"], want[:, 1]), dist
        assert np.array_equal(got["octave"], kp["octave"]) and np.array_equal(got["angle"], kp["angle"])
        assert (got["response"] == 0).all() and (got["class_id"] == -1).all()


def _undistort_maps(h, w):
    K = np.array([[0.9 * w, 0, w / 2.0], [0, 0.9 * w, h / 2.0], [0, 0, 1]])
    dist = np.array([-0.25, 0.08, 0.001, -0.002, 0.0])
    R = cv2.Rodrigues(np.array([0.01, -0.02, 0.005]))[0]
    P = np.array([[0.85 * w, 0, w / 2.0 - 8, 0], [0, 0.85 * w, h / 2.0 + 5, 0], [0, 0, 1, 0]])
    return cv2.initUndistortRectifyMap(K, dist, R, P, (w, h), cv2.CV_32FC1)


def test_preprocessing_vs_cv2():
    """util/image_converter.cc:8-39 (cvtColor), LpSlam ImageProcessing.h:258-370 (calcHist + convertTo) and :245-249
    (remap INTER_LINEAR, CV_32FC1 maps) against cv2."""
    from oracle import pyoracle
    o = pyoracle.port()
    rng = np.random.default_rng(11)
    for h, w in ((480, 640), (97, 131)):
        c3 = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        c4 = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        assert np.array_equal(o.gray(c3, "BGR"), cv2.cvtColor(c3, cv2.COLOR_BGR2GRAY))
        assert np.array_equal(o.gray(c3, "RGB"), cv2.cvtColor(c3, cv2.COLOR_RGB2GRAY))
        assert np.array_equal(o.gray(c4, "BGR"), cv2.cvtColor(c4, cv2.COLOR_BGRA2GRAY))
        assert np.array_equal(o.gray(c4, "RGB"), cv2.cvtColor(c4, cv2.COLOR_RGBA2GRAY))
        img = synth_image(h, w, 5)
        mx, my = _undistort_maps(h, w)
        mx[0, :10], my[5, :7], mx[7, 3], my[9, 9], mx[10, 10], mx[11, 11] = -5.3, h + 3.2, w - 0.6, -0.4, -0.97, -1.0
        assert np.array_equal(o.remap(img, mx, my), cv2.remap(img, mx, my, cv2.INTER_LINEAR))
        assert np.array_equal(o.remap(noise_image(h, w, 6), mx, my), cv2.remap(noise_image(h, w, 6), mx, my, cv2.INTER_LINEAR))
    # imadjust: histogram limits against cv2.calcHist + the reference's loops, value map against convertScaleAbs on the
    # non-negative side (Python has no Mat::convertTo; convertScaleAbs shares its float kernel) with dispatch off
    was = cv2.useOptimized()
    cv2.setUseOptimized(False)
    try:
        for seed in range(6):
            img = np.clip(synth_image(240, 320, 20 + seed).astype(np.int32) // (1 + seed % 3) + 10 * seed, 0, 255).astype(np.uint8)
            dst, lut, lim = o.imadjust(img, None, None, -0.3, 1.4)
            hist = cv2.calcHist([img], [0], None, [256], [0, 256]).ravel()
            total, s, lo, hi = img.size, 0, 0.0, 1.0
            for i in range(256):
                if s >= int(0.01 * total):
                    lo = i
                    break
                s += int(hist[i])
            s = 0
            for i in range(255, -1, -1):
                if s >= int((1 - 0.99) * total):
                    hi = i
                    break
                s += int(hist[i])
            assert lim[0] == lo / 255 and lim[1] == hi / 255
            alpha = (1.4 + 0.3) / (lim[1] - lim[0])
            beta = (1.4 - lim[1] * alpha) * 255
            ramp = np.arange(256, dtype=np.uint8).reshape(1, 256).repeat(16, 0)
            ref = cv2.convertScaleAbs(ramp, alpha=alpha, beta=beta)[0]
            pos = (np.arange(256, dtype=np.float32) * np.float32(alpha) + np.float32(beta)) >= 0
            assert np.array_equal(lut[pos], ref[pos]) and (lut[~pos] == 0).all()
            assert np.array_equal(dst, lut[img])
    finally:
        cv2.setUseOptimized(was)
