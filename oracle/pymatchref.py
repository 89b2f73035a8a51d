"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libmatch_ref.so: the reference's OWN matcher sources (match/robust.cc,
projection.cc, area.cc, fuse.cc, bow_tree.cc, angle_checker.h, base.h), data/common.cc, data/landmark.cc and camera/perspective.cc
compiled unmodified from /root/reference (oracle/Makefile, oracle/ref_match_glue.cc). Method names and argument order are those
of pyoracle.Oracle's matcher methods, so a test can run the port and the reference build over the same inputs.

Only tests/ and tools/gen_golden*.py may import this module. The product (lpslam_b200) never does.
"""
import ctypes as C
import os

import numpy as np

from .pyoracle import KP_DTYPE, _CameraStruct

HERE = os.path.dirname(os.path.abspath(__file__))
MATCH_REF_SO = os.path.join(HERE, "_ref", "libmatch_ref.so")

_vp = C.c_void_p


def _a(x, dt):
    return None if x is None else np.ascontiguousarray(x, dt)


def _p(x):
    return None if x is None else C.c_void_p(x.ctypes.data)


def _cam(cam):
    """accepts pyoracle's camera struct or lpslam_b200._lib.Camera (same layout)"""
    return C.cast(C.byref(cam), _vp)


class MatchRef:
    def __init__(self, path=MATCH_REF_SO):
        self.lib = C.CDLL(path)
        L = self.lib
        for name in ("hamming_32", "hamming_64", "landmark_descriptor"):
            getattr(L, "ref_" + name).restype = C.c_uint32
        L.ref_camera_info.restype = None
        L.ref_undistort_keypoints.restype = None
        L.ref_keypoints_to_bearings.restype = None
        L.ref_assign_grid.restype = None
        L.ref_landmark_descriptor_bytes.restype = None

    # ---- match/base.h, match/angle_checker.h ----
    def hamming(self, a, b, variant=32):
        a, b = _a(a, np.uint8), _a(b, np.uint8)
        fn = self.lib.ref_hamming_32 if variant == 32 else self.lib.ref_hamming_64
        return int(fn(_p(a), _p(b)))

    def angle_invalid(self, deltas, ids):
        d, i = _a(deltas, np.float32), _a(ids, np.int32)
        out = np.empty(max(len(i), 1), np.int32)
        n = self.lib.ref_angle_invalid(_p(d), _p(i), len(i), _p(out))
        return out[:n].copy()

    def angle_valid(self, deltas, ids):
        d, i = _a(deltas, np.float32), _a(ids, np.int32)
        out = np.empty(max(len(i), 1), np.int32)
        n = self.lib.ref_angle_valid(_p(d), _p(i), len(i), _p(out))
        return out[:n].copy()

    def angle_check(self, deltas, ids, num_bins_thr):
        """angle_checker<int>(30, num_bins_thr): (invalid ids, valid ids)"""
        d, i = _a(deltas, np.float32), _a(ids, np.int32)
        inv, val = np.empty(max(len(i), 1), np.int32), np.empty(max(len(i), 1), np.int32)
        nv = C.c_int(0)
        n = self.lib.ref_angle_check(_p(d), _p(i), len(i), C.c_uint(num_bins_thr), _p(inv), _p(val), C.byref(nv))
        return inv[:n].copy(), val[: nv.value].copy()

    # ---- match/robust.cc ----
    def match_bruteforce(self, desc_1, desc_2, valid_2=None, angle_1=None, angle_2=None, lowe_ratio=0.8, check_orientation=False):
        d1, d2 = _a(desc_1, np.uint8), _a(desc_2, np.uint8)
        n1, n2 = len(d1), len(d2)
        v = _a(valid_2, np.uint8)
        a1 = np.zeros(n1, np.float32) if angle_1 is None else _a(angle_1, np.float32)
        a2 = np.zeros(n2, np.float32) if angle_2 is None else _a(angle_2, np.float32)
        out = np.empty(n1, np.int32)
        n = self.lib.ref_match_bruteforce(_p(d1), C.c_uint32(n1), _p(d2), C.c_uint32(n2), _p(v), _p(a1), _p(a2), C.c_float(lowe_ratio),
                                          int(check_orientation), _p(out))
        return out, n

    def match_triangulation(self, vec_1, vec_2, side_1, side_2, E_12, epiplane_in_keyfrm_2, check_orientation):
        a = [_a(x, np.uint32) for x in vec_1]
        b = [_a(x, np.uint32) for x in vec_2]
        kp1, d1, b1, x1, h1, sc1 = side_1
        kp2, d2, b2, x2, h2 = side_2[:5]
        kp1, kp2 = _a(kp1, KP_DTYPE), _a(kp2, KP_DTYPE)
        d1, d2 = _a(d1, np.uint8), _a(d2, np.uint8)
        b1, b2 = _a(b1, np.float64), _a(b2, np.float64)
        x1, x2 = _a(x1, np.float32), _a(x2, np.float32)
        h1, h2 = _a(h1, np.uint8), _a(h2, np.uint8)
        sc1 = _a(sc1, np.float32)
        E = _a(E_12, np.float64).reshape(9)
        epi = _a(epiplane_in_keyfrm_2, np.float64).reshape(3)
        out = np.empty(len(kp1), np.int32)
        n = self.lib.ref_match_triangulation(C.c_uint32(len(a[0])), _p(a[0]), _p(a[1]), _p(a[2]), C.c_uint32(len(b[0])), _p(b[0]), _p(b[1]),
                                             _p(b[2]), C.c_uint32(len(kp1)), _p(kp1), _p(d1), _p(b1), _p(x1), _p(h1), _p(sc1),
                                             C.c_uint32(len(sc1)), C.c_uint32(len(kp2)), _p(kp2), _p(d2), _p(b2), _p(x2), _p(h2), _p(E),
                                             _p(epi), int(check_orientation), _p(out))
        return out, n

    # ---- match/projection.cc ----
    def match_projection(self, cam, pose_curr, pose_last, last_has_lm, last_outlier, last_pos_w, last_lm_desc, last_octave, last_angle,
                         curr_kp, curr_desc, curr_x_right, curr_has_observed_lm, scale, margin, check_orientation=True,
                         last_lm_has_observation=None):
        pc, pl = _a(pose_curr, np.float64), _a(pose_last, np.float64)
        has, outl = _a(last_has_lm, np.uint8), _a(last_outlier, np.uint8)
        pos, ld = _a(last_pos_w, np.float64), _a(last_lm_desc, np.uint8)
        loct, lang = _a(last_octave, np.int32), _a(last_angle, np.float32)
        kp, cd = _a(curr_kp, KP_DTYPE), _a(curr_desc, np.uint8)
        xr, tk = _a(curr_x_right, np.float32), _a(curr_has_observed_lm, np.uint8)
        sc = _a(scale, np.float32)
        obs = _a(last_lm_has_observation, np.uint8)
        out = np.empty(len(kp), np.int32)
        n = self.lib.ref_match_projection(_cam(cam), _p(pc), _p(pl), C.c_uint32(len(has)), _p(has), _p(outl), _p(pos), _p(ld), _p(loct),
                                          _p(lang), C.c_uint32(len(kp)), _p(kp), _p(cd), _p(xr), _p(tk), _p(sc), C.c_float(margin),
                                          int(check_orientation), _p(obs), _p(out))
        return out, n

    def match_landmarks(self, cam, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_desc, kp, kp_desc, kp_x_right, kp_has_observed_lm,
                        scale, margin, lowe_ratio, lm_has_observation=None):
        act, rp = _a(lm_active, np.uint8), _a(lm_reproj, np.float64)
        lxr, lvl, ld = _a(lm_x_right, np.float32), _a(lm_scale_level, np.int32), _a(lm_desc, np.uint8)
        kp, kd = _a(kp, KP_DTYPE), _a(kp_desc, np.uint8)
        xr, tk, sc = _a(kp_x_right, np.float32), _a(kp_has_observed_lm, np.uint8), _a(scale, np.float32)
        obs = _a(lm_has_observation, np.uint8)
        out = np.empty(len(kp), np.int32)
        n = self.lib.ref_match_landmarks(_cam(cam), C.c_uint32(len(act)), _p(act), _p(rp), _p(lxr), _p(lvl), _p(ld), _p(obs),
                                         C.c_uint32(len(kp)), _p(kp), _p(kd), _p(xr), _p(tk), _p(sc), C.c_float(margin),
                                         C.c_float(lowe_ratio), _p(out))
        return out, n

    def match_frame_and_keyframe(self, cam, margin, hamm_thr, lm_active, lm_reproj, lm_scale_level, lm_desc, lm_angle, kp, kp_desc, kp_taken,
                                 scale):
        """projection::match_frame_and_keyframe (:210-312) = the port's match_reprojected with (level_lo, level_hi, use_ratio,
        use_x_right) = (-1, +1, 0, 0); lm_angle None switches the orientation check off"""
        act, rp = _a(lm_active, np.uint8), _a(lm_reproj, np.float64)
        lvl, ld, ang = _a(lm_scale_level, np.int32), _a(lm_desc, np.uint8), _a(lm_angle, np.float32)
        kp, kd, tk, sc = _a(kp, KP_DTYPE), _a(kp_desc, np.uint8), _a(kp_taken, np.uint8), _a(scale, np.float32)
        out = np.empty(len(kp), np.int32)
        n = self.lib.ref_match_frame_and_keyframe(_cam(cam), C.c_float(margin), C.c_uint32(hamm_thr), C.c_uint32(len(act)), _p(act), _p(rp),
                                                  _p(lvl), _p(ld), _p(ang), C.c_uint32(len(kp)), _p(kp), _p(kd), _p(tk), _p(sc), _p(out))
        return out, n

    def match_sim3(self, cam, margin, lm_active, lm_reproj, lm_scale_level, lm_desc, kp, kp_desc, kp_has_lm, scale):
        act, rp = _a(lm_active, np.uint8), _a(lm_reproj, np.float64)
        lvl, ld = _a(lm_scale_level, np.int32), _a(lm_desc, np.uint8)
        kp, kd, tk, sc = _a(kp, KP_DTYPE), _a(kp_desc, np.uint8), _a(kp_has_lm, np.uint8), _a(scale, np.float32)
        out = np.empty(len(kp), np.int32)
        n = self.lib.ref_match_sim3(_cam(cam), C.c_float(margin), C.c_uint32(len(act)), _p(act), _p(rp), _p(lvl), _p(ld), C.c_uint32(len(kp)),
                                    _p(kp), _p(kd), _p(tk), _p(sc), _p(out))
        return out, n

    def match_mutual(self, cam_1, cam_2, margin, side_1, side_2):
        def prep(side):
            act, rp, lvl, ld, kp, kd, sc = side
            return [_a(act, np.uint8), _a(rp, np.float64), _a(lvl, np.int32), _a(ld, np.uint8), _a(kp, KP_DTYPE), _a(kd, np.uint8),
                    _a(sc, np.float32)]

        a, b = prep(side_1), prep(side_2)
        out = np.empty(len(a[4]), np.int32)

        def args(x):
            return [C.c_uint32(len(x[4])), _p(x[0]), _p(x[1]), _p(x[2]), _p(x[3]), _p(x[4]), _p(x[5]), _p(x[6])]

        n = self.lib.ref_match_mutual(_cam(cam_1), _cam(cam_2), C.c_float(margin), *args(a), *args(b), _p(out))
        return out, n

    # ---- match/fuse.cc ----
    def match_fuse(self, cam, margin, chi_gate, unsigned_levels, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_desc, kp, kp_desc,
                   kp_x_right, scale, inv_level_sigma_sq):
        """chi_gate False = detect_duplication, True = replace_duplication (whose level arithmetic is unsigned): (best_idx, None, n)"""
        assert bool(chi_gate) == bool(unsigned_levels), "the reference has these two combinations only"
        act, rp = _a(lm_active, np.uint8), _a(lm_reproj, np.float64)
        lxr, lvl, ld = _a(lm_x_right, np.float32), _a(lm_scale_level, np.int32), _a(lm_desc, np.uint8)
        kp, kd = _a(kp, KP_DTYPE), _a(kp_desc, np.uint8)
        xr, sc, isq = _a(kp_x_right, np.float32), _a(scale, np.float32), _a(inv_level_sigma_sq, np.float32)
        best = np.empty(len(act), np.int32)
        n = self.lib.ref_match_fuse(_cam(cam), C.c_float(margin), int(chi_gate), C.c_uint32(len(act)), _p(act), _p(rp), _p(lxr), _p(lvl),
                                    _p(ld), C.c_uint32(len(kp)), _p(kp), _p(kd), _p(xr), _p(sc), _p(isq), _p(best))
        return best, None, n

    # ---- match/bow_tree.cc ----
    def match_bow(self, kf_vec, f_vec, kf_valid, kf_desc, kf_angle, f_desc, f_angle, lowe_ratio, check_orientation):
        ka = [_a(x, np.uint32) for x in kf_vec]
        fa = [_a(x, np.uint32) for x in f_vec]
        kv, kd, kang = _a(kf_valid, np.uint8), _a(kf_desc, np.uint8), _a(kf_angle, np.float32)
        fd, fang = _a(f_desc, np.uint8), _a(f_angle, np.float32)
        out = np.empty(len(fd), np.int32)
        n = self.lib.ref_match_bow(C.c_uint32(len(ka[0])), _p(ka[0]), _p(ka[1]), _p(ka[2]), C.c_uint32(len(fa[0])), _p(fa[0]), _p(fa[1]),
                                   _p(fa[2]), C.c_uint32(len(kd)), _p(kv), _p(kd), _p(kang), C.c_uint32(len(fd)), _p(fd), _p(fang),
                                   C.c_float(lowe_ratio), int(check_orientation), _p(out))
        return out, n

    def match_bow_keyframes(self, vec_1, vec_2, valid_1, desc_1, angle_1, valid_2, desc_2, angle_2, lowe_ratio, check_orientation):
        a = [_a(x, np.uint32) for x in vec_1]
        b = [_a(x, np.uint32) for x in vec_2]
        v1, v2 = _a(valid_1, np.uint8), _a(valid_2, np.uint8)
        d1, d2 = _a(desc_1, np.uint8), _a(desc_2, np.uint8)
        a1, a2 = _a(angle_1, np.float32), _a(angle_2, np.float32)
        out = np.empty(len(d1), np.int32)
        n = self.lib.ref_match_bow_keyframes(C.c_uint32(len(a[0])), _p(a[0]), _p(a[1]), _p(a[2]), C.c_uint32(len(b[0])), _p(b[0]), _p(b[1]),
                                             _p(b[2]), C.c_uint32(len(d1)), _p(v1), _p(d1), _p(a1), C.c_uint32(len(d2)), _p(v2), _p(d2),
                                             _p(a2), C.c_float(lowe_ratio), int(check_orientation), _p(out))
        return out, n

    # ---- match/area.cc ----
    def match_area(self, cam, kp1, desc1, prev_pts, kp2, desc2, margin, lowe_ratio, check_orientation):
        kp1, kp2 = _a(kp1, KP_DTYPE), _a(kp2, KP_DTYPE)
        d1, d2 = _a(desc1, np.uint8), _a(desc2, np.uint8)
        pp = np.array(prev_pts, np.float32).reshape(-1, 2).copy()
        out = np.empty(len(kp1), np.int32)
        n = self.lib.ref_match_area(_cam(cam), C.c_uint32(len(kp1)), _p(kp1), _p(d1), _p(pp), C.c_uint32(len(kp2)), _p(kp2), _p(d2), int(margin),
                                    C.c_float(lowe_ratio), int(check_orientation), _p(out))
        return out, n, pp

    # ---- data/landmark.cc ----
    def landmark_descriptor(self, descs):
        d = _a(descs, np.uint8).reshape(-1, 32)
        return int(self.lib.ref_landmark_descriptor(_p(d), C.c_uint32(len(d))))

    def landmark_descriptor_bytes(self, descs):
        d = _a(descs, np.uint8).reshape(-1, 32)
        out = np.empty(32, np.uint8)
        self.lib.ref_landmark_descriptor_bytes(_p(d), C.c_uint32(len(d)), _p(out))
        return out

    # ---- data/common.cc, camera/perspective.cc ----
    def assign_grid(self, cam, kp):
        kp = _a(kp, KP_DTYPE)
        ncell = cam.grid_cols * cam.grid_rows
        start = np.zeros(ncell + 1, np.uint32)
        items = np.empty(max(len(kp), 1), np.uint32)
        self.lib.ref_assign_grid(_cam(cam), C.c_uint32(len(kp)), _p(kp), _p(start), _p(items))
        return start, items[: start[-1]]

    def get_cell_indices(self, cam, x, y):
        cx, cy = C.c_int(0), C.c_int(0)
        ok = self.lib.ref_get_cell_indices(_cam(cam), C.c_float(x), C.c_float(y), C.byref(cx), C.byref(cy))
        return bool(ok), cx.value, cy.value

    def keypoints_in_cell(self, cam, kp, ref_x, ref_y, margin, min_level=-1, max_level=-1):
        kp = _a(kp, KP_DTYPE)
        out = np.empty(max(len(kp), 1), np.uint32)
        n = self.lib.ref_keypoints_in_cell(_cam(cam), C.c_uint32(len(kp)), _p(kp), C.c_float(ref_x), C.c_float(ref_y), C.c_float(margin),
                                           int(min_level), int(max_level), _p(out))
        return out[:n].copy()

    def camera_info(self, fx, fy, cx, cy, dist5, cols, rows, focal_x_baseline=0.0, is_monocular=False, num_levels=8):
        """camera::perspective's derived members (image bounds incl. distortion, grid constants) as a camera struct"""
        cam = _CameraStruct()
        d = _a(dist5, np.float64)
        self.lib.ref_camera_info(C.c_double(fx), C.c_double(fy), C.c_double(cx), C.c_double(cy), _p(d), C.c_uint32(cols), C.c_uint32(rows),
                                 C.c_double(focal_x_baseline), C.byref(cam))
        cam.is_monocular, cam.num_scale_levels = int(is_monocular), num_levels
        return cam

    def undistort_keypoints(self, fx, fy, cx, cy, dist5, kp):
        kp = _a(kp, KP_DTYPE)
        d = _a(dist5, np.float64)
        out = np.empty_like(kp)
        self.lib.ref_undistort_keypoints(C.c_double(fx), C.c_double(fy), C.c_double(cx), C.c_double(cy), _p(d), C.c_uint32(len(kp)), _p(kp),
                                         _p(out))
        return out

    def keypoints_to_bearings(self, fx, fy, cx, cy, kp):
        kp = _a(kp, KP_DTYPE)
        out = np.empty((len(kp), 3), np.float64)
        self.lib.ref_keypoints_to_bearings(C.c_double(fx), C.c_double(fy), C.c_double(cx), C.c_double(cy), C.c_uint32(len(kp)), _p(kp), _p(out))
        return out

    def reproject_to_image(self, cam, pose_cw, pos_w):
        p, w = _a(pose_cw, np.float64), _a(pos_w, np.float64)
        rp = np.empty(2, np.float64)
        xr = C.c_float(0)
        ok = self.lib.ref_reproject_to_image(_cam(cam), _p(p), _p(w), _p(rp), C.byref(xr))
        return bool(ok), rp, xr.value

    def descriptors_to_json(self, desc):
        d = _a(desc, np.uint8).reshape(-1, 32)
        buf = C.create_string_buffer(len(d) * 100 + 16)
        n = self.lib.ref_descriptors_to_json(_p(d), C.c_uint32(len(d)), buf, len(buf))
        assert n >= 0
        return buf.value.decode()

    def json_to_descriptors(self, text, cap_rows=1 << 16):
        out = np.empty((cap_rows, 32), np.uint8)
        n = self.lib.ref_json_to_descriptors(text.encode(), _p(out), C.c_uint32(cap_rows))
        assert n >= 0
        return out[:n].copy()


_cache = {}


def match_ref():
    """None when oracle/_ref/libmatch_ref.so was never built (a clone without /root/reference and without the shipped artefact)."""
    if "m" not in _cache:
        _cache["m"] = MatchRef() if os.path.exists(MATCH_REF_SO) else None
    return _cache["m"]
