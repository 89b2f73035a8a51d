"""Python mirror of the reference's matcher classes (match/stereo.h, match/robust.h, match/projection.h) over the C ABI.
Constructor arguments and method names follow the reference; frames / keyframes are passed as flat numpy arrays."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check, lib, register, u8p, f32p, szp

register("lpslam_match_stereo", C.c_int, [C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                          szp, szp, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                          _lib.u32p])
register("lpslam_stereo_frames", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_uint32, C.c_int,
                                           C.c_int, C.c_size_t, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_uint32,
                                           C.c_void_p, C.c_void_p, C.c_void_p])
register("lpslam_stereo_frames_dev", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_size_t, C.c_float,
                                               C.c_float])
register("lpslam_stereo_result_dev", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)])

HAMMING_DIST_THR_LOW, HAMMING_DIST_THR_HIGH, MAX_HAMMING_DIST = 50, 100, 256  # match/base.h:13-15


class stereo:
    """match::stereo (match/stereo.h:18-29): same constructor arguments, compute() returns (stereo_x_right, depths)."""

    def __init__(self, left_image_pyramid, right_image_pyramid, keypts_left, keypts_right, descs_left, descs_right,
                 scale_factors, inv_scale_factors, focal_x_baseline, true_baseline):
        self.pl = [np.ascontiguousarray(p, np.uint8) for p in left_image_pyramid]
        self.pr = [np.ascontiguousarray(p, np.uint8) for p in right_image_pyramid]
        self.kl = np.ascontiguousarray(keypts_left, KP_DTYPE)
        self.kr = np.ascontiguousarray(keypts_right, KP_DTYPE)
        self.dl = np.ascontiguousarray(descs_left, np.uint8)
        self.dr = np.ascontiguousarray(descs_right, np.uint8)
        self.sf = np.ascontiguousarray(scale_factors, np.float32)
        self.isf = np.ascontiguousarray(inv_scale_factors, np.float32)
        self.fx_b, self.b = float(focal_x_baseline), float(true_baseline)

    def compute(self):
        n = len(self.pl)
        pl = (C.c_void_p * n)(*[p.ctypes.data for p in self.pl])
        pr = (C.c_void_p * n)(*[p.ctypes.data for p in self.pr])
        lw = (C.c_int * n)(*[p.shape[1] for p in self.pl])
        lh = (C.c_int * n)(*[p.shape[0] for p in self.pl])
        ppl = (C.c_size_t * n)(*[p.strides[0] for p in self.pl])
        ppr = (C.c_size_t * n)(*[p.strides[0] for p in self.pr])
        xr = np.full(len(self.kl), -1.0, np.float32)
        dep = np.full(len(self.kl), -1.0, np.float32)
        nm = C.c_uint32(0)
        check(lib().lpslam_match_stereo(pl, pr, lw, lh, ppl, ppr, n, self.kl.ctypes.data, len(self.kl), self.kr.ctypes.data,
                                        len(self.kr), self.dl.ctypes.data, self.dr.ctypes.data, self.sf.ctypes.data,
                                        self.isf.ctypes.data, self.fx_b, self.b, xr.ctypes.data, dep.ctypes.data,
                                        C.byref(nm)))
        return xr, dep


def stereo_frames(extractor, lefts, rights, focal_x_baseline, true_baseline, cap_per_frame=None):
    """data::frame's stereo constructor for a batch of pairs (data/frame.cc:55-93) on one extractor context.
    Returns a list of dicts: keypts_left / descs_left / keypts_right / descs_right / stereo_x_right / depths."""
    n = len(lefts)
    lefts = [np.ascontiguousarray(a, np.uint8) for a in lefts]
    rights = [np.ascontiguousarray(a, np.uint8) for a in rights]
    h, w = lefts[0].shape
    cap = 2 * n * (int(cap_per_frame) if cap_per_frame else 8 * extractor.orb_params_.max_num_keypts_ + 256)
    kp = np.empty(cap, KP_DTYPE)
    desc = np.empty((cap, 32), np.uint8)
    xr = np.empty(cap, np.float32)
    dep = np.empty(cap, np.float32)
    offs = np.zeros(2 * n + 1, np.uint32)
    pl = (C.c_void_p * n)(*[a.ctypes.data for a in lefts])
    pr = (C.c_void_p * n)(*[a.ctypes.data for a in rights])
    check(lib().lpslam_stereo_frames(extractor._h, pl, pr, n, w, h, w, focal_x_baseline, true_baseline, kp.ctypes.data,
                                     desc.ctypes.data, cap, offs.ctypes.data, xr.ctypes.data, dep.ctypes.data))
    out = []
    for p in range(n):
        a, b, c = int(offs[2 * p]), int(offs[2 * p + 1]), int(offs[2 * p + 2])
        out.append(dict(keypts_left=kp[a:b].copy(), descs_left=desc[a:b].copy(), keypts_right=kp[b:c].copy(),
                        descs_right=desc[b:c].copy(), stereo_x_right=xr[a:b].copy(), depths=dep[a:b].copy()))
    return out
