"""GPU parity of the image pre-processing kernels (gray conversion, imadjust, undistortion remap) against the oracle, which
is pinned against cv2 in test_oracle_vs_cv2.py. Bit-exact."""
import ctypes as C

import numpy as np
import pytest

from lpslam_b200.synth import noise_image, synth_image

pytestmark = pytest.mark.gpu


def synth_maps(h, w, seed):
    """smooth radial + shear displacement field, with out-of-image regions and exact-integer / half positions"""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    cx, cy = w / 2.0, h / 2.0
    r2 = ((xs - cx) / w) ** 2 + ((ys - cy) / w) ** 2
    mx = (cx + (xs - cx) * (1 + 0.3 * r2) + 0.01 * (ys - cy) - 3.0).astype(np.float32)
    my = (cy + (ys - cy) * (1 + 0.3 * r2) + 2.5).astype(np.float32)
    mx[:3] = np.round(mx[:3])
    my[3:6] = np.round(my[3:6]) + 0.5
    mx[h // 2, : w // 4] = rng.uniform(-40, w + 40, w // 4)
    my[h // 2 + 1, : w // 4] = rng.uniform(-40, h + 40, w // 4)
    return mx, my


@pytest.mark.parametrize("h,w", [(480, 640), (720, 1280), (97, 131), (33, 5)])
def test_gray(oracle, h, w):
    from lpslam_b200 import preproc
    rng = np.random.default_rng(h + w)
    for ch in (3, 4):
        img = rng.integers(0, 256, (h, w, ch), dtype=np.uint8)
        for order in ("RGB", "BGR"):
            assert np.array_equal(preproc.convert_to_grayscale(img, order), oracle.gray(img, order))
    g = rng.integers(0, 256, (h, w), dtype=np.uint8)
    assert preproc.convert_to_grayscale(g, "Gray") is g or np.array_equal(preproc.convert_to_grayscale(g, "Gray"), g)


@pytest.mark.parametrize("h,w", [(480, 640), (720, 1280), (97, 131)])
def test_imadjust(oracle, h, w):
    from lpslam_b200 import preproc
    cases = [synth_image(h, w, 3), noise_image(h, w, 4), (synth_image(h, w, 5) // 3 + 40).astype(np.uint8),
             np.full((h, w), 255, np.uint8), np.full((h, w), 7, np.uint8)]
    for img in cases:
        want, lut, lim = oracle.imadjust(img, None, None, -0.3, 1.4)
        got, glim = preproc.imadjust(img, None, None, -0.3, 1.4, return_limits=True)
        assert np.array_equal(glim, lim)
        assert np.array_equal(got, want)
        want2, _, _ = oracle.imadjust(img, 0.1, 0.8, 0.0, 1.0)
        assert np.array_equal(preproc.imadjust(img, 0.1, 0.8, 0.0, 1.0), want2)
    a, b = preproc.AdjustIntensityProcessor().processImage(cases[0], cases[1])
    assert np.array_equal(a, oracle.imadjust(cases[0], None, None, -0.3, 1.4)[0])
    assert np.array_equal(b, oracle.imadjust(cases[1], None, None, -0.3, 1.4)[0])


@pytest.mark.parametrize("h,w", [(480, 640), (720, 1280), (97, 131)])
def test_remap(oracle, h, w):
    from lpslam_b200 import preproc
    mx, my = synth_maps(h, w, 9)
    for img in (synth_image(h, w, 6), noise_image(h, w, 7)):
        assert np.array_equal(preproc.Undistort(mx, my).undistort(img), oracle.remap(img, mx, my))
    # destination size differs from the source size
    img = noise_image(h, w, 8)
    sub_x, sub_y = mx[: h // 2, : w // 2 + 1].copy(), my[: h // 2, : w // 2 + 1].copy()
    assert np.array_equal(preproc.Undistort(sub_x, sub_y).undistort(img), oracle.remap(img, sub_x, sub_y))


def test_device_chain_into_extractor(oracle):
    """colour frames -> gray -> remap -> imadjust on the device, written into the extractor's own level-0 buffer, then
    extraction: equals the oracle run step by step on the host."""
    import torch
    from lpslam_b200 import orb, preproc
    from lpslam_b200._lib import check, lib
    h, w, n = 360, 640, 3
    rng = np.random.default_rng(5)
    base = [synth_image(h, w, 30 + k) for k in range(n)]
    color = np.stack([np.stack([np.roll(b, 1, 1), b, np.roll(b, 2, 0)], axis=2) for b in base])  # n x h x w x 3, BGR
    color = (color // 2 + 30).astype(np.uint8)
    mx, my = synth_maps(h, w, 1)
    L = lib()
    e = orb.orb_extractor(max_num_keypts=500, max_batch=n)
    dptr, dpitch, dstride = e.input_dev(w, h)
    st = L.lpslam_orb_stream(e._h)
    d_color = torch.from_numpy(color).cuda()
    d_gray = torch.empty((n, h, w), dtype=torch.uint8, device="cuda")
    d_rect = torch.empty((n, h, w), dtype=torch.uint8, device="cuda")
    d_mx, d_my = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda()
    scratch = torch.empty(L.lpslam_imadjust_scratch_bytes(n) + 16, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    check(L.lpslam_gray_dev(d_color.data_ptr(), w, h, 3 * w, 3 * w * h, 3, 1, n, d_gray.data_ptr(), w, w * h, st))
    check(L.lpslam_remap_linear_dev(d_gray.data_ptr(), w, h, w, w * h, d_mx.data_ptr(), d_my.data_ptr(), w, d_rect.data_ptr(), w, h, w,
                                    w * h, n, st))
    sp = (scratch.data_ptr() + 15) // 16 * 16
    check(L.lpslam_imadjust_dev(d_rect.data_ptr(), w, h, w, w * h, n, 1, 0.0, 1.0, -0.3, 1.4, dptr, dpitch, dstride,
                                sp, st))
    res = e.extract_batch_dev(n, w, h)
    for k in range(n):
        img = oracle.imadjust(oracle.remap(oracle.gray(color[k], "BGR"), mx, my), None, None, -0.3, 1.4)[0]
        kp, desc = oracle.extractor(max_kpts=500).extract(img)
        assert np.array_equal(res[k][0]["x"], kp["x"]) and np.array_equal(res[k][0]["y"], kp["y"])
        assert (res[k][1] == desc).all(axis=1).mean() >= 0.999
