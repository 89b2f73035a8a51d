"""lpslam_imadjust_dev_static: the intensity adjustment for hosts that must not allocate device memory while they run (a
real-time camera callback, where a cudaMalloc can stall the device for an unbounded time; a plugin inside a process whose own
pool took the device memory at start-up, where a cudaMalloc fails). Its scratch is a fixed array inside the library, so the
caller brings images only. Checked here through the C ABI against lpslam_imadjust_dev, which works in a scratch buffer the
caller allocates: same pixels and same limits, bit for bit (both run the same integer / float arithmetic).

The array is shared by all calls and keeps the last call's histograms, so the test makes two calls in a row with different
images and image counts: the second one only passes when the entry point clears what it must. This file is named to be
collected last: a failure here must not hide the other GPU tests behind `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

from lpslam_b200.synth import noise_image, synth_image

pytestmark = pytest.mark.gpu


def _adjust_both(L, check, torch, imgs, auto, stream=None):
    """runs the caller-scratch entry point and the static-scratch one on the same device images -> (out, limits) of each"""
    from lpslam_b200 import preproc  # noqa: F401  (registers the signatures)
    n, h, w = imgs.shape
    d_src = torch.from_numpy(imgs).cuda()
    outs, lims = [], []
    for static in (False, True):
        d_dst = torch.zeros_like(d_src)
        if static:
            d_lim = torch.full((2 * n,), -1.0, dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            check(L.lpslam_imadjust_dev_static(d_src.data_ptr(), w, h, w, w * h, n, int(auto), 0.1, 0.8, -0.3, 1.4, d_dst.data_ptr(),
                                               w, w * h, d_lim.data_ptr(), stream))
            torch.cuda.synchronize()
            lims.append(d_lim.cpu().numpy())
        else:
            scratch = torch.zeros(L.lpslam_imadjust_scratch_bytes(n) + 16, dtype=torch.uint8, device="cuda")
            sp = (scratch.data_ptr() + 15) // 16 * 16
            torch.cuda.synchronize()
            check(L.lpslam_imadjust_dev(d_src.data_ptr(), w, h, w, w * h, n, int(auto), 0.1, 0.8, -0.3, 1.4, d_dst.data_ptr(), w,
                                        w * h, sp, stream))
            torch.cuda.synchronize()
            off = sp - scratch.data_ptr()
            lims.append(scratch[off:off + 16 * n].cpu().numpy().view(np.float64))  # limits sit at the head of the caller's scratch
        outs.append(d_dst.cpu().numpy())
    return outs, lims


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    import torch
    from lpslam_b200._lib import check, lib
    L = lib()
    h, w = 360, 640
    first = np.stack([synth_image(h, w, 70), noise_image(h, w, 71), synth_image(h, w, 72) // 2 + 40]).astype(np.uint8)
    # second call: fewer images, other content, darker -> any histogram count left over from the first call shifts the limits
    second = np.stack([(noise_image(h, w, 73) // 3).astype(np.uint8), synth_image(h, w, 74)]).astype(np.uint8)
    for imgs in (first, second):
        (want, got), (want_lim, got_lim) = _adjust_both(L, check, torch, imgs, auto=True)
        assert np.array_equal(got, want)
        assert np.array_equal(got_lim, want_lim)
        assert not np.array_equal(got, imgs)  # the adjustment did something
    # given limits (no histogram pass) right after the automatic ones, odd width (the unaligned store path)
    odd = np.ascontiguousarray(first[:, :, :637])
    (want, got), (want_lim, got_lim) = _adjust_both(L, check, torch, odd, auto=False)
    assert np.array_equal(got, want) and np.array_equal(got_lim, want_lim)
    assert np.allclose(got_lim.reshape(-1, 2), [0.1, 0.8])


def test_static_scratch_rejects_more_images_than_it_holds():
    import torch
    from lpslam_b200 import preproc
    from lpslam_b200._lib import lib, ERR_ARG
    L = lib()
    n = preproc.IMADJUST_STATIC_MAX_IMAGES + 1
    d = torch.zeros((n, 16, 32), dtype=torch.uint8, device="cuda")
    rc = L.lpslam_imadjust_dev_static(d.data_ptr(), 32, 16, 32, 32 * 16, n, 1, 0.0, 1.0, 0.0, 1.0, d.data_ptr(), 32, 32 * 16, None, None)
    assert rc == ERR_ARG
    msg = L.lpslam_cuda_last_error().decode()
    assert str(preproc.IMADJUST_STATIC_MAX_IMAGES) in msg and "LPSLAM_IMADJUST_STATIC_MAX_IMAGES" in msg
    # the largest batch it does hold works
    n = preproc.IMADJUST_STATIC_MAX_IMAGES
    rng = np.random.default_rng(3)
    imgs = rng.integers(0, 256, (n, 48, 80), dtype=np.uint8)
    from lpslam_b200._lib import check
    (want, got), (want_lim, got_lim) = _adjust_both(L, check, torch, imgs, auto=True)
    assert np.array_equal(got, want) and np.array_equal(got_lim, want_lim)
