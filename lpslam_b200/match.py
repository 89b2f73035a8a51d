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
register("lpslam_match_stereo_ctx", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p,
                                              C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, _lib.u32p])
register("lpslam_stereo_frames", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_uint32, C.c_int,
                                           C.c_int, C.c_size_t, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_uint32,
                                           C.c_void_p, C.c_void_p, C.c_void_p])
register("lpslam_stereo_frames_submit", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_uint32, C.c_int, C.c_int,
                                                  C.c_size_t, C.c_float, C.c_float, C.POINTER(C.c_uint64)])
register("lpslam_stereo_frames_collect", C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                                   C.c_void_p])
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


def stereo_on_contexts(extractor_left, extractor_right, keypts_left, keypts_right, descs_left, descs_right, focal_x_baseline,
                       true_baseline):
    """match::stereo::compute (match/stereo.cc:18-249) on the pyramids the two extractors still hold on the device from their last
    extraction (lpslam_match_stereo_ctx: what the drop-in match::stereo does with two drop-in extractors, data/frame.cc:78)."""
    kl, kr = np.ascontiguousarray(keypts_left, KP_DTYPE), np.ascontiguousarray(keypts_right, KP_DTYPE)
    dl, dr = np.ascontiguousarray(descs_left, np.uint8), np.ascontiguousarray(descs_right, np.uint8)
    xr = np.full(len(kl), -1.0, np.float32)
    dep = np.full(len(kl), -1.0, np.float32)
    nm = C.c_uint32(0)
    check(lib().lpslam_match_stereo_ctx(extractor_left._h, extractor_right._h, kl.ctypes.data, len(kl), kr.ctypes.data, len(kr),
                                        dl.ctypes.data, dr.ctypes.data, float(focal_x_baseline), float(true_baseline), xr.ctypes.data,
                                        dep.ctypes.data, C.byref(nm)))
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


def stereo_frames_submit(extractor, lefts, rights, focal_x_baseline, true_baseline):
    """lpslam_stereo_frames_submit: enqueue a batch of pairs, return (ticket, keep_alive). The arrays in keep_alive are the buffers
    the copies read from: hold on to them until the ticket is collected."""
    n = len(lefts)
    lefts = [np.ascontiguousarray(a, np.uint8) for a in lefts]
    rights = [np.ascontiguousarray(a, np.uint8) for a in rights]
    h, w = lefts[0].shape
    pl = (C.c_void_p * n)(*[a.ctypes.data for a in lefts])
    pr = (C.c_void_p * n)(*[a.ctypes.data for a in rights])
    t = C.c_uint64(0)
    check(lib().lpslam_stereo_frames_submit(extractor._h, pl, pr, n, w, h, w, focal_x_baseline, true_baseline, C.byref(t)))
    return t.value, (lefts, rights, n)


def stereo_frames_collect(extractor, ticket, keep_alive, cap_per_frame=None):
    """lpslam_stereo_frames_collect: wait for a submitted batch and return the same list of dicts as stereo_frames."""
    n = keep_alive[2]
    cap = 2 * n * (int(cap_per_frame) if cap_per_frame else 8 * extractor.orb_params_.max_num_keypts_ + 256)
    kp = np.empty(cap, KP_DTYPE)
    desc = np.empty((cap, 32), np.uint8)
    xr = np.empty(cap, np.float32)
    dep = np.empty(cap, np.float32)
    offs = np.zeros(2 * n + 1, np.uint32)
    check(lib().lpslam_stereo_frames_collect(extractor._h, ticket, kp.ctypes.data, desc.ctypes.data, cap, offs.ctypes.data, xr.ctypes.data,
                                             dep.ctypes.data))
    out = []
    for p in range(n):
        a, b, c = int(offs[2 * p]), int(offs[2 * p + 1]), int(offs[2 * p + 2])
        out.append(dict(keypts_left=kp[a:b].copy(), descs_left=desc[a:b].copy(), keypts_right=kp[b:c].copy(),
                        descs_right=desc[b:c].copy(), stereo_x_right=xr[a:b].copy(), depths=dep[a:b].copy()))
    return out


register("lpslam_match_bruteforce", C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_float, C.c_int, C.c_void_p, _lib.u32p])
register("lpslam_match_consecutive_frames", C.c_int, [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p])
register("lpslam_hamming_pairs", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p])
register("lpslam_match_projection", C.c_int, [C.POINTER(_lib.Camera), C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p,
                                              _lib.u32p])
register("lpslam_match_landmarks", C.c_int, [C.POINTER(_lib.Camera), C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_float, C.c_float, C.c_void_p, _lib.u32p])


class Perspective(C.Structure):
    """camera::perspective intrinsics + distortion (camera/perspective.h), lpslam_perspective in the C ABI"""
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double), ("dist", C.c_double * 5)]


register("lpslam_frame_postprocess", C.c_int, [C.POINTER(Perspective), C.POINTER(_lib.Camera), C.c_uint32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, _lib.u32p])


def frame_postprocess(model, cam, keypts):
    """data/frame.cc:37-51: (undist_keypts, bearings, cell_start, cell_items) of a frame's keypoints on the device."""
    kp = np.ascontiguousarray(keypts, KP_DTYPE)
    und = np.empty_like(kp)
    bearings = np.empty((len(kp), 3), np.float64)
    ncell = cam.grid_cols * cam.grid_rows
    start = np.zeros(ncell + 1, np.uint32)
    items = np.empty(max(len(kp), 1), np.uint32)
    n = C.c_uint32(0)
    check(lib().lpslam_frame_postprocess(C.byref(model), C.byref(cam), len(kp), kp.ctypes.data, und.ctypes.data, bearings.ctypes.data,
                                         start.ctypes.data, items.ctypes.data, C.byref(n)))
    return und, bearings, start, items[: n.value]
class ReprojMatchParams(C.Structure):
    """lpslam_reproj_match_params"""
    _fields_ = [("margin", C.c_float), ("lowe_ratio", C.c_float), ("hamm_dist_thr", C.c_uint32), ("level_lo", C.c_int32),
                ("level_hi", C.c_int32), ("use_ratio", C.c_int32), ("use_x_right", C.c_int32)]


register("lpslam_match_reprojected", C.c_int, [C.POINTER(_lib.Camera), C.POINTER(ReprojMatchParams), C.c_uint32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                               C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, _lib.u32p])


def match_reprojected(cam, params, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_descs, undist_keypts, descs, stereo_x_right,
                      has_landmark, scale_factors, lm_has_observation=None, lm_angle=None):
    """lpslam_match_reprojected over numpy arrays: returns (kp_match_lm, num_matches)."""
    act = np.ascontiguousarray(lm_active, np.uint8)
    rp = np.ascontiguousarray(lm_reproj, np.float64).reshape(-1, 2)
    lxr = None if lm_x_right is None else np.ascontiguousarray(lm_x_right, np.float32)
    lvl = np.ascontiguousarray(lm_scale_level, np.int32)
    ld = np.ascontiguousarray(lm_descs, np.uint8).reshape(-1, 32)
    obs = None if lm_has_observation is None else np.ascontiguousarray(lm_has_observation, np.uint8)
    ang = None if lm_angle is None else np.ascontiguousarray(lm_angle, np.float32)
    kp = np.ascontiguousarray(undist_keypts, KP_DTYPE)
    cd = np.ascontiguousarray(descs, np.uint8)
    xr = None if stereo_x_right is None else np.ascontiguousarray(stereo_x_right, np.float32)
    tk = np.ascontiguousarray(has_landmark, np.uint8)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    out = np.full(len(kp), -1, np.int32)
    n = C.c_uint32(0)
    check(lib().lpslam_match_reprojected(C.byref(cam), C.byref(params), len(act), act.ctypes.data, rp.ctypes.data, _lib.ptr(lxr),
                                         lvl.ctypes.data, ld.ctypes.data, _lib.ptr(obs), _lib.ptr(ang), len(kp), kp.ctypes.data,
                                         cd.ctypes.data, _lib.ptr(xr), tk.ctypes.data, sf.ctypes.data, out.ctypes.data, C.byref(n)))
    return out, int(n.value)


register("lpslam_match_fuse", C.c_int, [C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, _lib.u32p])


class fuse:
    """match::fuse (match/fuse.h): the keypoint search of detect_duplication / replace_duplication. The caller keeps the
    per-landmark geometry and the association step (they work on its landmark / keyframe objects); rows are the landmarks
    that passed the visibility gates, with their reprojection, x_right and predicted scale level."""

    def __init__(self, lowe_ratio=0.6):
        self.lowe_ratio_ = float(lowe_ratio)  # the fuse matcher carries it (match/base.h) but never applies it

    @staticmethod
    def _search(cam, margin, chi, unsigned_levels, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_descs, undist_keypts, descs,
                stereo_x_right, scale_factors, inv_level_sigma_sq):
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64).reshape(-1, 2)
        lxr = None if lm_x_right is None else np.ascontiguousarray(lm_x_right, np.float32)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_descs, np.uint8).reshape(-1, 32)
        kp = np.ascontiguousarray(undist_keypts, KP_DTYPE)
        cd = np.ascontiguousarray(descs, np.uint8)
        xr = None if stereo_x_right is None else np.ascontiguousarray(stereo_x_right, np.float32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        isq = None if inv_level_sigma_sq is None else np.ascontiguousarray(inv_level_sigma_sq, np.float32)
        best = np.full(len(act), -1, np.int32)
        dist = np.full(len(act), 256, np.uint32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_fuse(C.byref(cam), float(margin), int(chi), int(unsigned_levels), len(act), act.ctypes.data, rp.ctypes.data,
                                      _lib.ptr(lxr), lvl.ctypes.data, ld.ctypes.data, len(kp), kp.ctypes.data, cd.ctypes.data,
                                      _lib.ptr(xr), sf.ctypes.data, _lib.ptr(isq), best.ctypes.data, dist.ctypes.data, C.byref(n)))
        return best, dist, int(n.value)

    def detect_duplication(self, cam, lm_active, lm_reproj, lm_scale_level, lm_descs, undist_keypts, descs, scale_factors, margin):
        """match/fuse.cc:12-120 (search loop :69-101): returns (best_idx, best_dist, num_fused)"""
        return self._search(cam, margin, False, False, lm_active, lm_reproj, None, lm_scale_level, lm_descs, undist_keypts, descs, None,
                            scale_factors, None)

    def replace_duplication(self, cam, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_descs, undist_keypts, descs, stereo_x_right,
                            scale_factors, inv_level_sigma_sq, margin=3.0):
        """match/fuse.cc:122-266 (search loop :176-238, with the reprojection-error gate and the unsigned level arithmetic)"""
        return self._search(cam, margin, True, True, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_descs, undist_keypts, descs,
                            stereo_x_right, scale_factors, inv_level_sigma_sq)


register("lpslam_match_candidates", C.c_int, [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_int, C.c_void_p, _lib.u32p])


class bow_tree:
    """match::bow_tree (match/bow_tree.h): BoW feature vectors are passed as (node_ids, node_start, items) uint32 arrays, node
    ids ascending (the iteration order of DBoW2::FeatureVector / fbow::BoWFeatVector)."""

    def __init__(self, lowe_ratio=0.6, check_orientation=True):
        self.lowe_ratio_, self.check_orientation_ = float(lowe_ratio), bool(check_orientation)

    def match_frame_and_keyframe(self, keyfrm_feat_vec, keyfrm_has_live_lm, keyfrm_descs, keyfrm_angles, frm_feat_vec, frm_descs,
                                 frm_angles):
        """match/bow_tree.cc:16-126: returns (frm_match_keyfrm, num_matches); frm_match_keyfrm[i] = keyframe keypoint whose
        landmark frame keypoint i receives (matched_lms_in_frm.at(i) = keyfrm_lms.at(that)), or -1."""
        kid, kst, kit = [np.ascontiguousarray(x, np.uint32) for x in keyfrm_feat_vec]
        fid, fst, fit = [np.ascontiguousarray(x, np.uint32) for x in frm_feat_vec]
        valid = np.ascontiguousarray(keyfrm_has_live_lm, np.uint8)
        kd = np.ascontiguousarray(keyfrm_descs, np.uint8)
        fd = np.ascontiguousarray(frm_descs, np.uint8)
        # the merge walk of bow_tree.cc:36-109 over the two node lists: one row per keyframe keypoint with a live landmark
        rows, begin, end = [], [], []
        a = b = 0
        while a < len(kid) and b < len(fid):
            if kid[a] == fid[b]:
                for ki in kit[kst[a]:kst[a + 1]]:
                    if valid[ki]:
                        rows.append(ki)
                        begin.append(fst[b])
                        end.append(fst[b + 1])
                a += 1
                b += 1
            elif kid[a] < fid[b]:
                a += 1
            else:
                b += 1
        rows = np.asarray(rows, np.int64)
        out = np.full(len(fd), -1, np.int32)
        n = C.c_uint32(0)
        rd = np.ascontiguousarray(kd[rows]) if len(rows) else np.zeros((0, 32), np.uint8)
        rb, re_ = np.asarray(begin, np.uint32), np.asarray(end, np.uint32)
        ra = np.ascontiguousarray(np.asarray(keyfrm_angles, np.float32)[rows]) if self.check_orientation_ else None
        fa = np.ascontiguousarray(frm_angles, np.float32) if self.check_orientation_ else None
        check(lib().lpslam_match_candidates(len(rows), rd.ctypes.data, rb.ctypes.data, re_.ctypes.data, fit.ctypes.data, len(fit),
                                            _lib.ptr(ra), len(fd), fd.ctypes.data, _lib.ptr(fa), None, HAMMING_DIST_THR_LOW,
                                            self.lowe_ratio_, 1, out.ctypes.data, C.byref(n)))
        res = np.where(out >= 0, rows[np.clip(out, 0, max(len(rows) - 1, 0))] if len(rows) else -1, -1).astype(np.int32)
        return res, int(n.value)

    def match_keyframes(self, keyfrm_1_feat_vec, keyfrm_1_has_live_lm, keyfrm_1_descs, keyfrm_1_angles, keyfrm_2_feat_vec,
                        keyfrm_2_has_live_lm, keyfrm_2_descs, keyfrm_2_angles):
        """match/bow_tree.cc:128-250 (loop detection, before the Sim3 solver): returns (match_2_of_1, num_matches);
        match_2_of_1[i] = keyframe-2 keypoint whose landmark keyframe-1 keypoint i receives
        (matched_lms_in_keyfrm_1.at(i) = keyfrm_2_lms.at(that)), or -1. Same search as match_frame_and_keyframe: a
        keyframe-2 keypoint without a live landmark is never a candidate, which is what 'occupied beforehand' means."""
        id1, st1, it1 = [np.ascontiguousarray(x, np.uint32) for x in keyfrm_1_feat_vec]
        id2, st2, it2 = [np.ascontiguousarray(x, np.uint32) for x in keyfrm_2_feat_vec]
        v1 = np.ascontiguousarray(keyfrm_1_has_live_lm, np.uint8)
        v2 = np.ascontiguousarray(keyfrm_2_has_live_lm, np.uint8)
        d1 = np.ascontiguousarray(keyfrm_1_descs, np.uint8)
        d2 = np.ascontiguousarray(keyfrm_2_descs, np.uint8)
        rows, begin, end = [], [], []
        a = b = 0
        while a < len(id1) and b < len(id2):  # the merge walk of bow_tree.cc:150-236
            if id1[a] == id2[b]:
                for i1 in it1[st1[a]:st1[a + 1]]:
                    if v1[i1]:
                        rows.append(i1)
                        begin.append(st2[b])
                        end.append(st2[b + 1])
                a += 1
                b += 1
            elif id1[a] < id2[b]:
                a += 1
            else:
                b += 1
        rows = np.asarray(rows, np.int64)
        out = np.full(len(d2), -1, np.int32)
        n = C.c_uint32(0)
        rd = np.ascontiguousarray(d1[rows]) if len(rows) else np.zeros((0, 32), np.uint8)
        rb, re_ = np.asarray(begin, np.uint32), np.asarray(end, np.uint32)
        ra = np.ascontiguousarray(np.asarray(keyfrm_1_angles, np.float32)[rows]) if self.check_orientation_ else None
        a2 = np.ascontiguousarray(keyfrm_2_angles, np.float32) if self.check_orientation_ else None
        taken = np.ascontiguousarray(1 - np.minimum(v2, 1), np.uint8)
        check(lib().lpslam_match_candidates(len(rows), rd.ctypes.data, rb.ctypes.data, re_.ctypes.data, it2.ctypes.data, len(it2),
                                            _lib.ptr(ra), len(d2), d2.ctypes.data, _lib.ptr(a2), taken.ctypes.data, HAMMING_DIST_THR_LOW,
                                            self.lowe_ratio_, 1, out.ctypes.data, C.byref(n)))
        res = np.full(len(d1), -1, np.int32)
        hit = np.nonzero(out >= 0)[0]
        res[rows[out[hit]]] = hit.astype(np.int32)
        return res, int(n.value)


register("lpslam_match_sim3", C.c_int, [C.POINTER(_lib.Camera), C.c_float, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, _lib.u32p])
_MUTUAL_SIDE = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
register("lpslam_match_mutual", C.c_int, [C.POINTER(_lib.Camera), C.POINTER(_lib.Camera), C.c_float] + _MUTUAL_SIDE + _MUTUAL_SIDE
         + [C.c_void_p, _lib.u32p])


class _projection_keyframes:
    """The keyframe-to-keyframe members of match::projection (match/projection.h) used by loop closing. The per-landmark
    geometry in front of the search loops (reprojection under the Sim3, valid-distance / viewing-angle gates,
    predict_scale_level) stays the caller's; rows come with its results."""

    def __init__(self, lowe_ratio=0.6, check_orientation=True):
        self.lowe_ratio_, self.check_orientation_ = float(lowe_ratio), bool(check_orientation)  # carried, not used by these two

    def match_by_Sim3_transform(self, cam, lm_active, lm_reproj, lm_scale_level, lm_descs, undist_keypts, descs,
                                has_matched_lm, scale_factors, margin):
        """match/projection.cc:317-413: returns (kp_match_lm, num_matches); kp_match_lm[i] = row (landmark) that keypoint i
        of the keyframe received (matched_lms_in_keyfrm.at(i) = lm), or -1."""
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64).reshape(-1, 2)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_descs, np.uint8).reshape(-1, 32)
        kp = np.ascontiguousarray(undist_keypts, KP_DTYPE)
        cd = np.ascontiguousarray(descs, np.uint8)
        tk = np.ascontiguousarray(has_matched_lm, np.uint8)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        out = np.full(len(kp), -1, np.int32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_sim3(C.byref(cam), float(margin), len(act), act.ctypes.data, rp.ctypes.data, lvl.ctypes.data,
                                      ld.ctypes.data, len(kp), kp.ctypes.data, cd.ctypes.data, tk.ctypes.data, sf.ctypes.data,
                                      out.ctypes.data, C.byref(n)))
        return out, int(n.value)

    def match_keyframes_mutually(self, cam_1, cam_2, side_1, side_2, margin):
        """match/projection.cc:415-636. side_k = (lm_active, lm_reproj_in_other, lm_level_in_other, lm_descs, undist_keypts,
        descs, scale_factors) of keyframe k, one landmark slot per keypoint. Returns (matched_idx_2_in_1, num_matches)."""
        def prep(side):
            act, rp, lvl, ld, kp, kd, sc = side
            return [np.ascontiguousarray(act, np.uint8), np.ascontiguousarray(rp, np.float64).reshape(-1, 2),
                    np.ascontiguousarray(lvl, np.int32), np.ascontiguousarray(ld, np.uint8).reshape(-1, 32),
                    np.ascontiguousarray(kp, KP_DTYPE), np.ascontiguousarray(kd, np.uint8), np.ascontiguousarray(sc, np.float32)]

        a, b = prep(side_1), prep(side_2)
        out = np.full(len(a[4]), -1, np.int32)
        n = C.c_uint32(0)

        def args(x):
            return [len(x[4]), x[0].ctypes.data, x[1].ctypes.data, x[2].ctypes.data, x[3].ctypes.data, x[4].ctypes.data,
                    x[5].ctypes.data, x[6].ctypes.data]

        check(lib().lpslam_match_mutual(C.byref(cam_1), C.byref(cam_2), float(margin), *args(a), *args(b), out.ctypes.data, C.byref(n)))
        return out, int(n.value)


register("lpslam_match_triangulation", C.c_int, [C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                 C.c_void_p, _lib.u32p])


def match_for_triangulation(feat_vec_1, side_1, feat_vec_2, side_2, E_12, epiplane_in_keyfrm_2, check_orientation=True):
    """match::robust::match_for_triangulation (match/robust.cc:18-178). feat_vec_k = (node_ids, node_start, items);
    side_1 = (undist_keypts, descs, bearings, stereo_x_right, has_landmark, scale_factors), side_2 the same without the
    scale factors. Returns (matched_idx_2_in_1, num_matches); the pairs of :172-180 are the entries >= 0 in index order."""
    id1, st1, it1 = [np.ascontiguousarray(x, np.uint32) for x in feat_vec_1]
    id2, st2, it2 = [np.ascontiguousarray(x, np.uint32) for x in feat_vec_2]
    kp1, d1, b1, x1, h1, sc1 = side_1
    kp2, d2, b2, x2, h2 = side_2[:5]
    kp1, kp2 = np.ascontiguousarray(kp1, KP_DTYPE), np.ascontiguousarray(kp2, KP_DTYPE)
    d1, d2 = np.ascontiguousarray(d1, np.uint8), np.ascontiguousarray(d2, np.uint8)
    b1, b2 = np.ascontiguousarray(b1, np.float64).reshape(-1, 3), np.ascontiguousarray(b2, np.float64).reshape(-1, 3)
    x1, x2 = np.ascontiguousarray(x1, np.float32), np.ascontiguousarray(x2, np.float32)
    h1, h2 = np.ascontiguousarray(h1, np.uint8), np.ascontiguousarray(h2, np.uint8)
    sc1 = np.ascontiguousarray(sc1, np.float32)
    E = np.ascontiguousarray(E_12, np.float64).reshape(9)
    epi = np.ascontiguousarray(epiplane_in_keyfrm_2, np.float64).reshape(3)
    rows, begin, end = [], [], []
    a = b = 0
    while a < len(id1) and b < len(id2):  # the merge walk of robust.cc:55-161
        if id1[a] == id2[b]:
            for i1 in it1[st1[a]:st1[a + 1]]:
                if not h1[i1]:
                    rows.append(i1)
                    begin.append(st2[b])
                    end.append(st2[b + 1])
            a += 1
            b += 1
        elif id1[a] < id2[b]:
            a += 1
        else:
            b += 1
    ri = np.asarray(rows, np.uint32)
    rb, re_ = np.asarray(begin, np.uint32), np.asarray(end, np.uint32)
    out = np.full(len(kp1), -1, np.int32)
    n = C.c_uint32(0)
    check(lib().lpslam_match_triangulation(len(ri), ri.ctypes.data, rb.ctypes.data, re_.ctypes.data, it2.ctypes.data, len(it2),
                                           len(kp1), kp1.ctypes.data, d1.ctypes.data, b1.ctypes.data, x1.ctypes.data, sc1.ctypes.data,
                                           len(kp2), kp2.ctypes.data, d2.ctypes.data, b2.ctypes.data, x2.ctypes.data, h2.ctypes.data,
                                           E.ctypes.data, epi.ctypes.data, int(check_orientation), out.ctypes.data, C.byref(n)))
    return out, int(n.value)


register("lpslam_match_area", C.c_int, [C.POINTER(_lib.Camera), C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p,
                                        C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p, _lib.u32p])


class area:
    """match::area (match/area.h): the matcher of the monocular initializer"""

    def __init__(self, lowe_ratio=0.9, check_orientation=True):
        self.lowe_ratio_, self.check_orientation_ = float(lowe_ratio), bool(check_orientation)

    def match_in_consistent_area(self, cam, undist_keypts_1, descs_1, undist_keypts_2, descs_2, prev_matched_pts, margin=100):
        """match/area.cc:8-112 over flat arrays: returns (matched_indices_2_in_frm_1, num_matches, prev_matched_pts)."""
        k1 = np.ascontiguousarray(undist_keypts_1, KP_DTYPE)
        k2 = np.ascontiguousarray(undist_keypts_2, KP_DTYPE)
        d1 = np.ascontiguousarray(descs_1, np.uint8)
        d2 = np.ascontiguousarray(descs_2, np.uint8)
        pp = np.array(prev_matched_pts, np.float32).reshape(-1, 2).copy()
        out = np.full(len(k1), -1, np.int32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_area(C.byref(cam), len(k1), k1.ctypes.data, d1.ctypes.data, pp.ctypes.data, len(k2), k2.ctypes.data,
                                      d2.ctypes.data, int(margin), self.lowe_ratio_, int(self.check_orientation_), out.ctypes.data,
                                      C.byref(n)))
        return out, int(n.value), pp


register("lpslam_landmark_descriptors", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p])


def compute_landmark_descriptors(obs_descs, offsets):
    """data::landmark::compute_descriptor (data/landmark.cc:127-183) for a batch of landmarks: obs_descs[offsets[l]:offsets[l+1]]
    are the descriptors of landmark l's observations. Returns (best_idx, descriptors)."""
    d = np.ascontiguousarray(obs_descs, np.uint8).reshape(-1, 32)
    off = np.ascontiguousarray(offsets, np.uint32)
    n = len(off) - 1
    best = np.zeros(n, np.uint32)
    out = np.zeros((n, 32), np.uint8)
    check(lib().lpslam_landmark_descriptors(d.ctypes.data, off.ctypes.data, n, best.ctypes.data, out.ctypes.data))
    return best, out


def match_consecutive_frames(extractor, n_rows, n_frames, lowe_ratio=0.8):
    """lpslam_match_consecutive_frames on the batch `extractor` extracted last: (match_prev_in_cur[n_rows], n_matches[n_frames])"""
    out = np.full(max(n_rows, 1), -1, np.int32)
    cnt = np.zeros(n_frames, np.uint32)
    check(lib().lpslam_match_consecutive_frames(extractor._h, float(lowe_ratio), out.ctypes.data, cnt.ctypes.data))
    return out[:n_rows], cnt


def compute_descriptor_distance_32(desc_1, desc_2):
    """match/base.h:18-40 on the device, row-wise over (N, 32) uint8 arrays (or a single pair)."""
    a = np.ascontiguousarray(np.atleast_2d(desc_1), np.uint8)
    b = np.ascontiguousarray(np.atleast_2d(desc_2), np.uint8)
    assert a.shape == b.shape and a.shape[1] == 32
    out = np.zeros(len(a), np.uint32)
    check(lib().lpslam_hamming_pairs(a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data))
    return out if np.ndim(desc_1) == 2 else int(out[0])


compute_descriptor_distance_64 = compute_descriptor_distance_32  # match/base.h:43-63: same value, 64-bit words


class robust:
    """match::robust (match/robust.h:19-30)."""

    def __init__(self, lowe_ratio=0.8, check_orientation=False):
        self.lowe_ratio_, self.check_orientation_ = float(lowe_ratio), bool(check_orientation)

    def brute_force_match(self, frm_descs, keyfrm_descs, keyfrm_valid=None, frm_angles=None, keyfrm_angles=None):
        """match/robust.cc:215-321 over flat arrays: returns the (idx_1, idx_2) pairs sorted by idx_1."""
        d1 = np.ascontiguousarray(frm_descs, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(keyfrm_descs, np.uint8).reshape(-1, 32)
        valid = None if keyfrm_valid is None else np.ascontiguousarray(keyfrm_valid, np.uint8)
        a1 = None if frm_angles is None else np.ascontiguousarray(frm_angles, np.float32)
        a2 = None if keyfrm_angles is None else np.ascontiguousarray(keyfrm_angles, np.float32)
        out = np.full(len(d1), -1, np.int32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_bruteforce(d1.ctypes.data, len(d1), d2.ctypes.data, len(d2), _lib.ptr(valid), _lib.ptr(a1),
                                            _lib.ptr(a2), self.lowe_ratio_, int(self.check_orientation_), out.ctypes.data,
                                            C.byref(n)))
        idx_1 = np.nonzero(out >= 0)[0]
        matches = np.stack([idx_1, out[idx_1]], axis=1).astype(np.int32)
        assert len(matches) == n.value
        return matches, out

    def match_for_triangulation(self, feat_vec_1, side_1, feat_vec_2, side_2, E_12, epiplane_in_keyfrm_2):
        """match/robust.cc:18-178: returns the (idx_1, idx_2) pairs in order of idx_1 and matched_idx_2_in_1."""
        out, n = match_for_triangulation(feat_vec_1, side_1, feat_vec_2, side_2, E_12, epiplane_in_keyfrm_2, self.check_orientation_)
        idx_1 = np.nonzero(out >= 0)[0]
        assert len(idx_1) == n
        return np.stack([idx_1, out[idx_1]], axis=1).astype(np.int32), out


class projection(_projection_keyframes):
    """match::projection (match/projection.h:21-30); match_by_Sim3_transform / match_keyframes_mutually come from the base."""

    def __init__(self, lowe_ratio=0.6, check_orientation=True):
        self.lowe_ratio_, self.check_orientation_ = float(lowe_ratio), bool(check_orientation)

    def match_current_and_last_frames(self, cam, pose_cw_curr, pose_cw_last, last_has_lm, last_outlier, last_pos_w, last_lm_desc,
                                      last_octave, last_angle, curr_undist_keypts, curr_descs, curr_stereo_x_right,
                                      curr_has_observed_lm, scale_factors, margin, last_lm_has_observation=None):
        """match/projection.cc:85-208 over flat arrays: returns (curr_match_last, num_matches)."""
        Tc = np.ascontiguousarray(pose_cw_curr, np.float64)
        Tl = np.ascontiguousarray(pose_cw_last, np.float64)
        has = np.ascontiguousarray(last_has_lm, np.uint8)
        outl = np.ascontiguousarray(last_outlier, np.uint8)
        pos = np.ascontiguousarray(last_pos_w, np.float64)
        ld = np.ascontiguousarray(last_lm_desc, np.uint8)
        loct = np.ascontiguousarray(last_octave, np.int32)
        lang = np.ascontiguousarray(last_angle, np.float32)
        obs = None if last_lm_has_observation is None else np.ascontiguousarray(last_lm_has_observation, np.uint8)
        kp = np.ascontiguousarray(curr_undist_keypts, KP_DTYPE)
        cd = np.ascontiguousarray(curr_descs, np.uint8)
        xr = np.ascontiguousarray(curr_stereo_x_right, np.float32)
        tk = np.ascontiguousarray(curr_has_observed_lm, np.uint8)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        out = np.full(len(kp), -1, np.int32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_projection(C.byref(cam), Tc.ctypes.data, Tl.ctypes.data, len(has), has.ctypes.data,
                                            outl.ctypes.data, pos.ctypes.data, ld.ctypes.data, loct.ctypes.data, lang.ctypes.data,
                                            _lib.ptr(obs), len(kp), kp.ctypes.data, cd.ctypes.data, xr.ctypes.data,
                                            tk.ctypes.data, sf.ctypes.data, float(margin), int(self.check_orientation_),
                                            out.ctypes.data, C.byref(n)))
        return out, int(n.value)

    def match_frame_and_keyframe(self, cam, lm_active, lm_reproj, lm_scale_level, lm_descs, keyfrm_angles, undist_keypts, descs,
                                 has_landmark, scale_factors, margin, hamm_dist_thr):
        """match/projection.cc:210-312 after the caller's geometry (reproject_to_image, valid-distance gate, predict_scale_level):
        lm_active = landmark exists, is not erased, is not in already_matched_lms, reprojects into the image and passes the
        distance gate. Returns (kp_match_lm, num_matches)."""
        prm = ReprojMatchParams(float(margin), 0.0, int(hamm_dist_thr), -1, 1, 0, 0)
        return match_reprojected(cam, prm, lm_active, lm_reproj, None, lm_scale_level, lm_descs, undist_keypts, descs, None, has_landmark,
                                 scale_factors, None, keyfrm_angles if self.check_orientation_ else None)

    def match_frame_and_landmarks(self, cam, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_descs, undist_keypts, descs,
                                  stereo_x_right, has_observed_lm, scale_factors, margin=5.0, lm_has_observation=None):
        """match/projection.cc:11-83 over flat arrays: returns (kp_match_lm, num_matches). lm_active is
        is_observable_in_tracking_ && !will_be_erased(); lm_reproj / lm_x_right / lm_scale_level are the values cached on the
        landmarks by the tracker (data/landmark.h:112-115)."""
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64).reshape(-1, 2)
        lxr = np.ascontiguousarray(lm_x_right, np.float32)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_descs, np.uint8).reshape(-1, 32)
        obs = None if lm_has_observation is None else np.ascontiguousarray(lm_has_observation, np.uint8)
        kp = np.ascontiguousarray(undist_keypts, KP_DTYPE)
        cd = np.ascontiguousarray(descs, np.uint8)
        xr = np.ascontiguousarray(stereo_x_right, np.float32)
        tk = np.ascontiguousarray(has_observed_lm, np.uint8)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        out = np.full(len(kp), -1, np.int32)
        n = C.c_uint32(0)
        check(lib().lpslam_match_landmarks(C.byref(cam), len(act), act.ctypes.data, rp.ctypes.data, lxr.ctypes.data, lvl.ctypes.data,
                                           ld.ctypes.data, _lib.ptr(obs), len(kp), kp.ctypes.data, cd.ctypes.data, xr.ctypes.data,
                                           tk.ctypes.data, sf.ctypes.data, float(margin), self.lowe_ratio_, out.ctypes.data,
                                           C.byref(n)))
        return out, int(n.value)


def make_camera(fx, fy, cx, cy, cols, rows, focal_x_baseline=0.0, true_baseline=1.0, is_monocular=False, num_levels=8,
                grid_cols=64, grid_rows=48):
    """camera::perspective without distortion (camera/perspective.cc:20-31, camera/base.h:62,118-131)."""
    cam = _lib.Camera()
    cam.fx, cam.fy, cam.cx, cam.cy = fx, fy, cx, cy
    cam.focal_x_baseline, cam.true_baseline = focal_x_baseline, true_baseline
    cam.min_x, cam.max_x, cam.min_y, cam.max_y = 0.0, float(cols), 0.0, float(rows)
    cam.grid_cols, cam.grid_rows = grid_cols, grid_rows
    cam.inv_cell_width = float(grid_cols) / float(np.float32(cols) - np.float32(0))  # double division of a float difference (perspective.cc:29-30)
    cam.inv_cell_height = float(grid_rows) / float(np.float32(rows) - np.float32(0))
    cam.is_monocular = int(is_monocular)
    cam.num_scale_levels = num_levels
    return cam
