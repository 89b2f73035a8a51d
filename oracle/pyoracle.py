"""TEST INFRASTRUCTURE — ctypes binding of the CPU oracle (oracle/liborb_oracle.so and, when built, oracle/_ref/liborb_ref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product (lpslam_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liborb_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "liborb_ref.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)


def build(force=False):
    """Compile the port (always) and the reference build (when /root/reference is present)."""
    if force or not os.path.exists(PORT_SO):
        subprocess.check_call(["make", "-s", "-f", os.path.join(HERE, "Makefile"), PORT_SO])
    if os.path.isdir("/root/reference/external/openvslam/src/openvslam/feature") and os.path.exists(os.path.join(HERE, "ref_glue.cc")):  # liborb_ref.so + libmatch_ref.so
        subprocess.check_call(["make", "-s", "-f", os.path.join(HERE, "Makefile"), "ref"])


def _ptr(a, t):
    return a.ctypes.data_as(t)


class _CameraStruct(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("focal_x_baseline", C.c_double), ("true_baseline", C.c_double),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("grid_cols", C.c_uint32), ("grid_rows", C.c_uint32),
                ("inv_cell_width", C.c_double), ("inv_cell_height", C.c_double),
                ("is_monocular", C.c_int32), ("num_scale_levels", C.c_uint32)]


def make_camera(fx, fy, cx, cy, cols, rows, focal_x_baseline=0.0, true_baseline=1.0, is_monocular=False, num_levels=8,
                grid_cols=64, grid_rows=48):
    """Perspective camera without distortion: image bounds are the image rectangle (camera/perspective.cc:20-31)."""
    cam = _CameraStruct()
    cam.fx, cam.fy, cam.cx, cam.cy = fx, fy, cx, cy
    cam.focal_x_baseline, cam.true_baseline = focal_x_baseline, true_baseline
    cam.min_x, cam.max_x, cam.min_y, cam.max_y = 0.0, float(cols), 0.0, float(rows)
    cam.grid_cols, cam.grid_rows = grid_cols, grid_rows
    cam.inv_cell_width = float(grid_cols) / float(np.float32(cols) - np.float32(0))  # double division of a float difference (perspective.cc:29-30)
    cam.inv_cell_height = float(grid_rows) / float(np.float32(rows) - np.float32(0))
    cam.is_monocular = int(is_monocular)
    cam.num_scale_levels = num_levels
    return cam


class Oracle:
    """One loaded oracle library (port or reference build) with a numpy-facing API."""

    def __init__(self, path=PORT_SO, prefix="orc"):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        self.prefix = prefix
        L = self.lib
        self._f = lambda name: getattr(L, f"{prefix}_{name}")
        f = self._f
        f("create").restype = C.c_void_p
        f("create").argtypes = [C.c_uint32, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, f32p, C.c_uint32]
        f("destroy").argtypes = [C.c_void_p]
        f("set_max_kpts").argtypes = [C.c_void_p, C.c_uint32]
        f("get_table").argtypes = [C.c_void_p, C.c_int, f32p, C.c_int]
        f("extract").argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_size_t, C.c_void_p, u8p,
                                 C.c_uint32, C.POINTER(C.c_uint32)]
        f("level_size").argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        f("level_copy").argtypes = [C.c_void_p, C.c_int, C.c_int, u8p, C.c_size_t]
        f("level_candidates").argtypes = [C.c_void_p, C.c_int, i32p, C.c_int]
        f("level_selected").argtypes = [C.c_void_p, C.c_int, i32p, C.c_int]
        f("match_stereo").argtypes = [C.POINTER(u8p), C.POINTER(u8p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                      C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.c_int,
                                      C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, u8p, u8p, f32p, f32p,
                                      C.c_float, C.c_float, f32p, f32p]
        if prefix == "orc":
            L.orc_resize_linear_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_int, C.c_int, C.c_size_t]
            L.orc_gaussian7_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_size_t]
            L.orc_fast9_nms.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, i32p, C.c_int]
            L.orc_fast_atan2.restype = C.c_float
            L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
            L.orc_cos.restype = C.c_float
            L.orc_cos.argtypes = [C.c_float]
            L.orc_sin.restype = C.c_float
            L.orc_sin.argtypes = [C.c_float]
            L.orc_hamming_32.restype = C.c_uint32
            L.orc_hamming_32.argtypes = [u8p, u8p]
            L.orc_hamming_64.restype = C.c_uint32
            L.orc_hamming_64.argtypes = [u8p, u8p]
            L.orc_angle_invalid.argtypes = [f32p, i32p, C.c_int, i32p]
            L.orc_match_bruteforce.argtypes = [u8p, C.c_uint32, u8p, C.c_uint32, u8p, f32p, f32p, C.c_float, C.c_int, i32p]
            L.orc_match_projection.argtypes = [C.POINTER(_CameraStruct), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                               C.c_uint32, u8p, u8p, C.POINTER(C.c_double), u8p, i32p, f32p,
                                               C.c_uint32, C.c_void_p, u8p, f32p, u8p, f32p, C.c_float, C.c_int, u8p, i32p]
            L.orc_match_landmarks.argtypes = [C.POINTER(_CameraStruct), C.c_uint32, u8p, C.POINTER(C.c_double), f32p, i32p, u8p, u8p,
                                              C.c_uint32, C.c_void_p, u8p, f32p, u8p, f32p, C.c_float, C.c_float, i32p]
            L.orc_undistort_keypoints.argtypes = [C.c_double] * 4 + [C.POINTER(C.c_double), C.c_uint32, C.c_void_p, C.c_void_p]
            L.orc_undistort_keypoints.restype = None
            L.orc_keypoints_to_bearings.argtypes = [C.c_double] * 4 + [C.c_uint32, C.c_void_p, C.POINTER(C.c_double)]
            L.orc_keypoints_to_bearings.restype = None
            L.orc_assign_grid.argtypes = [C.POINTER(_CameraStruct), C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
            L.orc_assign_grid.restype = None
            L.orc_gray_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, u8p, C.c_size_t]
            L.orc_gray_u8.restype = None
            L.orc_imadjust_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                          u8p, C.c_size_t, u8p, C.POINTER(C.c_double)]
            L.orc_imadjust_u8.restype = None
            L.orc_remap_linear_u8.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, f32p, f32p, C.c_size_t, u8p, C.c_int, C.c_int, C.c_size_t]
            L.orc_remap_linear_u8.restype = None
            L.orc_match_reprojected.argtypes = [C.POINTER(_CameraStruct), C.c_float, C.c_float, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int,
                                                C.c_uint32, u8p, C.POINTER(C.c_double), f32p, i32p, u8p, u8p, f32p,
                                                C.c_uint32, C.c_void_p, u8p, f32p, u8p, f32p, i32p]
            u32p = C.POINTER(C.c_uint32)
            L.orc_match_bow.argtypes = [C.c_uint32, u32p, u32p, u32p, C.c_uint32, u32p, u32p, u32p, u8p, u8p, f32p, C.c_uint32, u8p, f32p,
                                        C.c_float, C.c_int, i32p]
            L.orc_match_area.argtypes = [C.POINTER(_CameraStruct), C.c_uint32, C.c_void_p, u8p, f32p, C.c_uint32, C.c_void_p, u8p, C.c_int,
                                         C.c_float, C.c_int, i32p]
            L.orc_landmark_descriptor.argtypes = [u8p, C.c_uint32]
            L.orc_landmark_descriptor.restype = C.c_uint32
            L.orc_knn_top2.argtypes = [u8p, C.c_uint32, u8p, C.c_uint64, C.POINTER(C.c_int64), C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_uint32)]

    # ---- primitives ----
    def resize(self, src, dw, dh):
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.empty((dh, dw), np.uint8)
        self.lib.orc_resize_linear_u8(_ptr(src, u8p), src.shape[1], src.shape[0], src.strides[0], _ptr(dst, u8p), dw, dh, dw)
        return dst

    def blur(self, src):
        src = np.ascontiguousarray(src, np.uint8)
        dst = np.empty_like(src)
        self.lib.orc_gaussian7_u8(_ptr(src, u8p), src.shape[1], src.shape[0], src.strides[0], _ptr(dst, u8p), dst.strides[0])
        return dst

    def fast(self, img, thr):
        img = np.ascontiguousarray(img, np.uint8)
        cap = img.size // 4 + 16
        out = np.empty((cap, 3), np.int32)
        n = self.lib.orc_fast9_nms(_ptr(img, u8p), img.shape[1], img.shape[0], img.strides[0], thr, _ptr(out, i32p), cap)
        return out[:n].copy()

    def atan2(self, y, x):
        return self.lib.orc_fast_atan2(float(y), float(x))

    def hamming(self, a, b, variant=32):
        a = np.ascontiguousarray(a, np.uint8)
        b = np.ascontiguousarray(b, np.uint8)
        fn = self.lib.orc_hamming_32 if variant == 32 else self.lib.orc_hamming_64
        return int(fn(_ptr(a, u8p), _ptr(b, u8p)))

    # ---- extractor ----
    def extractor(self, max_kpts=2000, scale_factor=1.2, num_levels=8, ini_fast_thr=20, min_fast_thr=7, mask_rects=()):
        return Extractor(self, max_kpts, scale_factor, num_levels, ini_fast_thr, min_fast_thr, mask_rects)

    # ---- matchers ----
    def angle_invalid(self, deltas, ids):
        deltas = np.ascontiguousarray(deltas, np.float32)
        ids = np.ascontiguousarray(ids, np.int32)
        out = np.empty(len(ids), np.int32)
        n = self.lib.orc_angle_invalid(_ptr(deltas, f32p), _ptr(ids, i32p), len(ids), _ptr(out, i32p))
        return out[:n].copy()

    def angle_check(self, deltas, ids, num_bins_thr):
        deltas = np.ascontiguousarray(deltas, np.float32)
        ids = np.ascontiguousarray(ids, np.int32)
        out = np.empty(max(len(ids), 1), np.int32)
        self.lib.orc_angle_check.argtypes = [f32p, i32p, C.c_int, C.c_uint, i32p]
        n = self.lib.orc_angle_check(_ptr(deltas, f32p), _ptr(ids, i32p), len(ids), num_bins_thr, _ptr(out, i32p))
        return out[:n].copy()

    def match_bruteforce(self, desc_1, desc_2, valid_2=None, angle_1=None, angle_2=None, lowe_ratio=0.8,
                         check_orientation=False):
        desc_1 = np.ascontiguousarray(desc_1, np.uint8)
        desc_2 = np.ascontiguousarray(desc_2, np.uint8)
        n1, n2 = len(desc_1), len(desc_2)
        valid = None if valid_2 is None else np.ascontiguousarray(valid_2, np.uint8)
        a1 = np.zeros(n1, np.float32) if angle_1 is None else np.ascontiguousarray(angle_1, np.float32)
        a2 = np.zeros(n2, np.float32) if angle_2 is None else np.ascontiguousarray(angle_2, np.float32)
        out = np.empty(n1, np.int32)
        n = self.lib.orc_match_bruteforce(_ptr(desc_1, u8p), n1, _ptr(desc_2, u8p), n2,
                                          None if valid is None else _ptr(valid, u8p), _ptr(a1, f32p), _ptr(a2, f32p),
                                          lowe_ratio, int(check_orientation), _ptr(out, i32p))
        return out, n

    def match_stereo(self, pyr_l, pyr_r, kp_l, kp_r, desc_l, desc_r, scale, inv_scale, fx_b, b):
        nlv = len(pyr_l)
        pyr_l = [np.ascontiguousarray(p, np.uint8) for p in pyr_l]
        pyr_r = [np.ascontiguousarray(p, np.uint8) for p in pyr_r]
        pl = (u8p * nlv)(*[_ptr(p, u8p) for p in pyr_l])
        pr = (u8p * nlv)(*[_ptr(p, u8p) for p in pyr_r])
        lw = (C.c_int * nlv)(*[p.shape[1] for p in pyr_l])
        lh = (C.c_int * nlv)(*[p.shape[0] for p in pyr_l])
        ppl = (C.c_size_t * nlv)(*[p.strides[0] for p in pyr_l])
        ppr = (C.c_size_t * nlv)(*[p.strides[0] for p in pyr_r])
        kp_l = np.ascontiguousarray(kp_l, KP_DTYPE)
        kp_r = np.ascontiguousarray(kp_r, KP_DTYPE)
        desc_l = np.ascontiguousarray(desc_l, np.uint8)
        desc_r = np.ascontiguousarray(desc_r, np.uint8)
        scale = np.ascontiguousarray(scale, np.float32)
        inv_scale = np.ascontiguousarray(inv_scale, np.float32)
        xr = np.empty(len(kp_l), np.float32)
        depth = np.empty(len(kp_l), np.float32)
        self._f("match_stereo")(pl, pr, lw, lh, ppl, ppr, nlv, kp_l.ctypes.data, len(kp_l), kp_r.ctypes.data, len(kp_r),
                                _ptr(desc_l, u8p), _ptr(desc_r, u8p), _ptr(scale, f32p), _ptr(inv_scale, f32p),
                                fx_b, b, _ptr(xr, f32p), _ptr(depth, f32p))
        return xr, depth

    def match_projection(self, cam, pose_curr, pose_last, last_has_lm, last_outlier, last_pos_w, last_lm_desc, last_octave,
                         last_angle, curr_kp, curr_desc, curr_x_right, curr_has_observed_lm, scale, margin,
                         check_orientation=True, last_lm_has_observation=None):
        dp = C.POINTER(C.c_double)
        pose_curr = np.ascontiguousarray(pose_curr, np.float64)
        pose_last = np.ascontiguousarray(pose_last, np.float64)
        last_has_lm = np.ascontiguousarray(last_has_lm, np.uint8)
        last_outlier = np.ascontiguousarray(last_outlier, np.uint8)
        last_pos_w = np.ascontiguousarray(last_pos_w, np.float64)
        last_lm_desc = np.ascontiguousarray(last_lm_desc, np.uint8)
        last_octave = np.ascontiguousarray(last_octave, np.int32)
        last_angle = np.ascontiguousarray(last_angle, np.float32)
        curr_kp = np.ascontiguousarray(curr_kp, KP_DTYPE)
        curr_desc = np.ascontiguousarray(curr_desc, np.uint8)
        curr_x_right = np.ascontiguousarray(curr_x_right, np.float32)
        curr_has = np.ascontiguousarray(curr_has_observed_lm, np.uint8)
        scale = np.ascontiguousarray(scale, np.float32)
        out = np.empty(len(curr_kp), np.int32)
        n = self.lib.orc_match_projection(C.byref(cam), _ptr(pose_curr, dp), _ptr(pose_last, dp), len(last_has_lm),
                                          _ptr(last_has_lm, u8p), _ptr(last_outlier, u8p), _ptr(last_pos_w, dp),
                                          _ptr(last_lm_desc, u8p), _ptr(last_octave, i32p), _ptr(last_angle, f32p),
                                          len(curr_kp), curr_kp.ctypes.data, _ptr(curr_desc, u8p),
                                          _ptr(curr_x_right, f32p), _ptr(curr_has, u8p), _ptr(scale, f32p),
                                          margin, int(check_orientation),
                                          None if last_lm_has_observation is None else _ptr(np.ascontiguousarray(last_lm_has_observation, np.uint8), u8p),
                                          _ptr(out, i32p))
        return out, n

    def match_landmarks(self, cam, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_desc, kp, kp_desc, kp_x_right,
                        kp_has_observed_lm, scale, margin, lowe_ratio, lm_has_observation=None):
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64)
        lxr = np.ascontiguousarray(lm_x_right, np.float32)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_desc, np.uint8)
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        kd = np.ascontiguousarray(kp_desc, np.uint8)
        xr = np.ascontiguousarray(kp_x_right, np.float32)
        tk = np.ascontiguousarray(kp_has_observed_lm, np.uint8)
        scale = np.ascontiguousarray(scale, np.float32)
        out = np.empty(len(kp), np.int32)
        n = self.lib.orc_match_landmarks(C.byref(cam), len(act), _ptr(act, u8p), _ptr(rp, C.POINTER(C.c_double)), _ptr(lxr, f32p),
                                         _ptr(lvl, i32p), _ptr(ld, u8p),
                                         None if lm_has_observation is None else _ptr(np.ascontiguousarray(lm_has_observation, np.uint8), u8p),
                                         len(kp), kp.ctypes.data, _ptr(kd, u8p), _ptr(xr, f32p), _ptr(tk, u8p), _ptr(scale, f32p),
                                         margin, lowe_ratio, _ptr(out, i32p))
        return out, n

    def undistort_keypoints(self, fx, fy, cx, cy, dist5, kp):
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        d = np.ascontiguousarray(dist5, np.float64)
        out = np.empty_like(kp)
        self.lib.orc_undistort_keypoints(fx, fy, cx, cy, _ptr(d, C.POINTER(C.c_double)), len(kp), kp.ctypes.data, out.ctypes.data)
        return out

    def keypoints_to_bearings(self, fx, fy, cx, cy, kp):
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        out = np.empty((len(kp), 3), np.float64)
        self.lib.orc_keypoints_to_bearings(fx, fy, cx, cy, len(kp), kp.ctypes.data, _ptr(out, C.POINTER(C.c_double)))
        return out

    def assign_grid(self, cam, kp):
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        ncell = cam.grid_cols * cam.grid_rows
        start = np.empty(ncell + 1, np.uint32)
        items = np.empty(len(kp), np.uint32)
        self.lib.orc_assign_grid(C.byref(cam), len(kp), kp.ctypes.data, _ptr(start, C.POINTER(C.c_uint32)), _ptr(items, C.POINTER(C.c_uint32)))
        return start, items[: start[-1]]

    def gray(self, img, order="BGR"):
        img = np.ascontiguousarray(img, np.uint8)
        h, w, ch = img.shape
        dst = np.empty((h, w), np.uint8)
        self.lib.orc_gray_u8(_ptr(img, u8p), w, h, img.strides[0], ch, 0 if order == "RGB" else 1, _ptr(dst, u8p), w)
        return dst

    def imadjust(self, img, low_in=None, high_in=None, low_out=0.0, high_out=1.0):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        dst = np.empty_like(img)
        lut = np.empty(256, np.uint8)
        lim = np.empty(2, np.float64)
        auto = low_in is None or high_in is None
        self.lib.orc_imadjust_u8(_ptr(img, u8p), w, h, img.strides[0], int(auto), 0.0 if auto else low_in, 1.0 if auto else high_in,
                                 low_out, high_out, _ptr(dst, u8p), w, _ptr(lut, u8p), _ptr(lim, C.POINTER(C.c_double)))
        return dst, lut, lim

    def remap(self, img, map_x, map_y):
        img = np.ascontiguousarray(img, np.uint8)
        mx = np.ascontiguousarray(map_x, np.float32)
        my = np.ascontiguousarray(map_y, np.float32)
        dh, dw = mx.shape
        dst = np.empty((dh, dw), np.uint8)
        self.lib.orc_remap_linear_u8(_ptr(img, u8p), img.shape[1], img.shape[0], img.strides[0], _ptr(mx, f32p), _ptr(my, f32p), dw,
                                     _ptr(dst, u8p), dw, dh, dw)
        return dst

    def match_reprojected(self, cam, margin, lowe_ratio, hamm_thr, level_lo, level_hi, use_ratio, use_x_right, lm_active, lm_reproj,
                          lm_x_right, lm_scale_level, lm_desc, kp, kp_desc, kp_x_right, kp_taken, scale, lm_has_observation=None,
                          lm_angle=None):
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64)
        lxr = np.ascontiguousarray(lm_x_right, np.float32)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_desc, np.uint8)
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        kd = np.ascontiguousarray(kp_desc, np.uint8)
        xr = np.ascontiguousarray(kp_x_right, np.float32)
        tk = np.ascontiguousarray(kp_taken, np.uint8)
        scale = np.ascontiguousarray(scale, np.float32)
        out = np.empty(len(kp), np.int32)
        n = self.lib.orc_match_reprojected(C.byref(cam), margin, lowe_ratio, hamm_thr, level_lo, level_hi, int(use_ratio), int(use_x_right),
                                           len(act), _ptr(act, u8p), _ptr(rp, C.POINTER(C.c_double)), _ptr(lxr, f32p), _ptr(lvl, i32p),
                                           _ptr(ld, u8p),
                                           None if lm_has_observation is None else _ptr(np.ascontiguousarray(lm_has_observation, np.uint8), u8p),
                                           None if lm_angle is None else _ptr(np.ascontiguousarray(lm_angle, np.float32), f32p),
                                           len(kp), kp.ctypes.data, _ptr(kd, u8p), _ptr(xr, f32p), _ptr(tk, u8p), _ptr(scale, f32p),
                                           _ptr(out, i32p))
        return out, n

    def match_fuse(self, cam, margin, chi_gate, unsigned_levels, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_desc, kp, kp_desc,
                   kp_x_right, scale, inv_level_sigma_sq):
        """search loop of fuse::detect_duplication / replace_duplication: (best_idx, best_dist, n_fused)"""
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64)
        lxr = np.ascontiguousarray(lm_x_right, np.float32)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_desc, np.uint8)
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        kd = np.ascontiguousarray(kp_desc, np.uint8)
        xr = np.ascontiguousarray(kp_x_right, np.float32)
        scale = np.ascontiguousarray(scale, np.float32)
        isq = np.ascontiguousarray(inv_level_sigma_sq, np.float32)
        best = np.empty(len(act), np.int32)
        dist = np.empty(len(act), np.uint32)
        self.lib.orc_match_fuse.restype = C.c_int
        self.lib.orc_match_fuse.argtypes = [C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_uint32, u8p, C.POINTER(C.c_double), f32p, i32p, u8p,
                                            C.c_uint32, C.c_void_p, u8p, f32p, f32p, f32p, i32p, C.POINTER(C.c_uint32)]
        n = self.lib.orc_match_fuse(C.cast(C.byref(cam), C.c_void_p), margin, int(chi_gate), int(unsigned_levels), len(act), _ptr(act, u8p),
                                    _ptr(rp, C.POINTER(C.c_double)), _ptr(lxr, f32p), _ptr(lvl, i32p), _ptr(ld, u8p), len(kp),
                                    kp.ctypes.data, _ptr(kd, u8p), _ptr(xr, f32p), _ptr(scale, f32p), _ptr(isq, f32p), _ptr(best, i32p),
                                    _ptr(dist, C.POINTER(C.c_uint32)))
        return best, dist, n

    def match_bow(self, kf_vec, f_vec, kf_valid, kf_desc, kf_angle, f_desc, f_angle, lowe_ratio, check_orientation):
        """kf_vec / f_vec: (node_ids, node_start, items) uint32 arrays"""
        u32p = C.POINTER(C.c_uint32)
        ka = [np.ascontiguousarray(x, np.uint32) for x in kf_vec]
        fa = [np.ascontiguousarray(x, np.uint32) for x in f_vec]
        kv = np.ascontiguousarray(kf_valid, np.uint8)
        kd = np.ascontiguousarray(kf_desc, np.uint8)
        kang = np.ascontiguousarray(kf_angle, np.float32)
        fd = np.ascontiguousarray(f_desc, np.uint8)
        fang = np.ascontiguousarray(f_angle, np.float32)
        out = np.empty(len(fd), np.int32)
        n = self.lib.orc_match_bow(len(ka[0]), _ptr(ka[0], u32p), _ptr(ka[1], u32p), _ptr(ka[2], u32p), len(fa[0]), _ptr(fa[0], u32p),
                                   _ptr(fa[1], u32p), _ptr(fa[2], u32p), _ptr(kv, u8p), _ptr(kd, u8p), _ptr(kang, f32p), len(fd),
                                   _ptr(fd, u8p), _ptr(fang, f32p), lowe_ratio, int(check_orientation), _ptr(out, i32p))
        return out, n

    def match_bow_keyframes(self, vec_1, vec_2, valid_1, desc_1, angle_1, valid_2, desc_2, angle_2, lowe_ratio, check_orientation):
        """bow_tree::match_keyframes: (match_2_of_1, n)"""
        u32p = C.POINTER(C.c_uint32)
        a = [np.ascontiguousarray(x, np.uint32) for x in vec_1]
        b = [np.ascontiguousarray(x, np.uint32) for x in vec_2]
        v1, v2 = np.ascontiguousarray(valid_1, np.uint8), np.ascontiguousarray(valid_2, np.uint8)
        d1, d2 = np.ascontiguousarray(desc_1, np.uint8), np.ascontiguousarray(desc_2, np.uint8)
        a1, a2 = np.ascontiguousarray(angle_1, np.float32), np.ascontiguousarray(angle_2, np.float32)
        out = np.empty(len(d1), np.int32)
        f = self.lib.orc_match_bow_keyframes
        f.restype = C.c_int
        f.argtypes = [C.c_uint32, u32p, u32p, u32p, C.c_uint32, u32p, u32p, u32p, C.c_uint32, u8p, u8p, f32p, C.c_uint32, u8p, u8p, f32p,
                      C.c_float, C.c_int, i32p]
        n = f(len(a[0]), _ptr(a[0], u32p), _ptr(a[1], u32p), _ptr(a[2], u32p), len(b[0]), _ptr(b[0], u32p), _ptr(b[1], u32p),
              _ptr(b[2], u32p), len(d1), _ptr(v1, u8p), _ptr(d1, u8p), _ptr(a1, f32p), len(d2), _ptr(v2, u8p), _ptr(d2, u8p),
              _ptr(a2, f32p), lowe_ratio, int(check_orientation), _ptr(out, i32p))
        return out, n

    def match_sim3(self, cam, margin, lm_active, lm_reproj, lm_scale_level, lm_desc, kp, kp_desc, kp_has_lm, scale):
        """search loop of projection::match_by_Sim3_transform: (kp_match_lm, num_matches)"""
        f64p = C.POINTER(C.c_double)
        act = np.ascontiguousarray(lm_active, np.uint8)
        rp = np.ascontiguousarray(lm_reproj, np.float64)
        lvl = np.ascontiguousarray(lm_scale_level, np.int32)
        ld = np.ascontiguousarray(lm_desc, np.uint8)
        kp = np.ascontiguousarray(kp, KP_DTYPE)
        kd = np.ascontiguousarray(kp_desc, np.uint8)
        tk = np.ascontiguousarray(kp_has_lm, np.uint8)
        scale = np.ascontiguousarray(scale, np.float32)
        out = np.empty(len(kp), np.int32)
        f = self.lib.orc_match_sim3
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_float, C.c_uint32, u8p, f64p, i32p, u8p, C.c_uint32, C.c_void_p, u8p, u8p, f32p, i32p]
        n = f(C.cast(C.byref(cam), C.c_void_p), margin, len(act), _ptr(act, u8p), _ptr(rp, f64p), _ptr(lvl, i32p), _ptr(ld, u8p),
              len(kp), kp.ctypes.data, _ptr(kd, u8p), _ptr(tk, u8p), _ptr(scale, f32p), _ptr(out, i32p))
        return out, n

    def match_mutual(self, cam_1, cam_2, margin, side_1, side_2):
        """projection::match_keyframes_mutually; side = (lm_active, lm_reproj_in_other, lm_level_in_other, lm_desc, kp, kp_desc,
        scale_factors): (matched_idx_2_in_1, num_matches)"""
        f64p = C.POINTER(C.c_double)

        def prep(side):
            act, rp, lvl, ld, kp, kd, sc = side
            return [np.ascontiguousarray(act, np.uint8), np.ascontiguousarray(rp, np.float64), np.ascontiguousarray(lvl, np.int32),
                    np.ascontiguousarray(ld, np.uint8), np.ascontiguousarray(kp, KP_DTYPE), np.ascontiguousarray(kd, np.uint8),
                    np.ascontiguousarray(sc, np.float32)]

        a, b = prep(side_1), prep(side_2)
        out = np.empty(len(a[4]), np.int32)
        f = self.lib.orc_match_mutual
        f.restype = C.c_int
        one = [C.c_uint32, u8p, f64p, i32p, u8p, C.c_void_p, u8p, f32p]
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_float] + one + one + [i32p]

        def args(x):
            return [len(x[4]), _ptr(x[0], u8p), _ptr(x[1], f64p), _ptr(x[2], i32p), _ptr(x[3], u8p), x[4].ctypes.data, _ptr(x[5], u8p),
                    _ptr(x[6], f32p)]

        n = f(C.cast(C.byref(cam_1), C.c_void_p), C.cast(C.byref(cam_2), C.c_void_p), margin, *args(a), *args(b), _ptr(out, i32p))
        return out, n

    def match_triangulation(self, vec_1, vec_2, side_1, side_2, E_12, epiplane_in_keyfrm_2, check_orientation):
        """robust::match_for_triangulation; side = (kp, desc, bearings, x_right, has_lm[, scale_factors for side 1]):
        (matched_idx_2_in_1, num_matches)"""
        u32p = C.POINTER(C.c_uint32)
        f64p = C.POINTER(C.c_double)
        a = [np.ascontiguousarray(x, np.uint32) for x in vec_1]
        b = [np.ascontiguousarray(x, np.uint32) for x in vec_2]
        kp1, d1, b1, x1, h1, sc1 = side_1
        kp2, d2, b2, x2, h2 = side_2[:5]
        kp1, kp2 = np.ascontiguousarray(kp1, KP_DTYPE), np.ascontiguousarray(kp2, KP_DTYPE)
        d1, d2 = np.ascontiguousarray(d1, np.uint8), np.ascontiguousarray(d2, np.uint8)
        b1, b2 = np.ascontiguousarray(b1, np.float64), np.ascontiguousarray(b2, np.float64)
        x1, x2 = np.ascontiguousarray(x1, np.float32), np.ascontiguousarray(x2, np.float32)
        h1, h2 = np.ascontiguousarray(h1, np.uint8), np.ascontiguousarray(h2, np.uint8)
        sc1 = np.ascontiguousarray(sc1, np.float32)
        E = np.ascontiguousarray(E_12, np.float64).reshape(9)
        epi = np.ascontiguousarray(epiplane_in_keyfrm_2, np.float64).reshape(3)
        out = np.empty(len(kp1), np.int32)
        f = self.lib.orc_match_triangulation
        f.restype = C.c_int
        f.argtypes = [C.c_uint32, u32p, u32p, u32p, C.c_uint32, u32p, u32p, u32p, C.c_uint32, C.c_void_p, u8p, f64p, f32p, u8p, f32p,
                      C.c_uint32, C.c_void_p, u8p, f64p, f32p, u8p, f64p, f64p, C.c_int, i32p]
        n = f(len(a[0]), _ptr(a[0], u32p), _ptr(a[1], u32p), _ptr(a[2], u32p), len(b[0]), _ptr(b[0], u32p), _ptr(b[1], u32p),
              _ptr(b[2], u32p), len(kp1), kp1.ctypes.data, _ptr(d1, u8p), _ptr(b1, f64p), _ptr(x1, f32p), _ptr(h1, u8p), _ptr(sc1, f32p),
              len(kp2), kp2.ctypes.data, _ptr(d2, u8p), _ptr(b2, f64p), _ptr(x2, f32p), _ptr(h2, u8p), _ptr(E, f64p), _ptr(epi, f64p),
              int(check_orientation), _ptr(out, i32p))
        return out, n

    def match_area(self, cam, kp1, desc1, prev_pts, kp2, desc2, margin, lowe_ratio, check_orientation):
        kp1 = np.ascontiguousarray(kp1, KP_DTYPE)
        kp2 = np.ascontiguousarray(kp2, KP_DTYPE)
        d1 = np.ascontiguousarray(desc1, np.uint8)
        d2 = np.ascontiguousarray(desc2, np.uint8)
        pp = np.array(prev_pts, np.float32).reshape(-1, 2).copy()
        out = np.empty(len(kp1), np.int32)
        n = self.lib.orc_match_area(C.byref(cam), len(kp1), kp1.ctypes.data, _ptr(d1, u8p), _ptr(pp, f32p), len(kp2), kp2.ctypes.data,
                                    _ptr(d2, u8p), int(margin), lowe_ratio, int(check_orientation), _ptr(out, i32p))
        return out, n, pp

    def landmark_descriptor(self, descs):
        d = np.ascontiguousarray(descs, np.uint8).reshape(-1, 32)
        return int(self.lib.orc_landmark_descriptor(_ptr(d, u8p), len(d)))

    def knn_top2(self, q, db):
        q = np.ascontiguousarray(q, np.uint8)
        db = np.ascontiguousarray(db, np.uint8)
        bi = np.empty(len(q), np.int64)
        bd = np.empty(len(q), np.uint32)
        sd = np.empty(len(q), np.uint32)
        self.lib.orc_knn_top2(_ptr(q, u8p), len(q), _ptr(db, u8p), len(db), _ptr(bi, C.POINTER(C.c_int64)),
                              _ptr(bd, C.POINTER(C.c_uint32)), _ptr(sd, C.POINTER(C.c_uint32)))
        return bi, bd, sd


class Extractor:
    def __init__(self, oracle, max_kpts, scale_factor, num_levels, ini_fast_thr, min_fast_thr, mask_rects):
        self.o = oracle
        self.num_levels = num_levels
        rects = np.ascontiguousarray(np.asarray(mask_rects, np.float32).reshape(-1, 4))
        self.h = oracle._f("create")(max_kpts, scale_factor, num_levels, ini_fast_thr, min_fast_thr,
                                     _ptr(rects, f32p) if len(rects) else None, len(rects))
        if not self.h:
            raise RuntimeError("oracle extractor creation failed (bad parameters)")

    def __del__(self):
        if getattr(self, "h", None):
            self.o._f("destroy")(self.h)
            self.h = None

    def set_max_kpts(self, n):
        self.o._f("set_max_kpts")(self.h, n)

    def table(self, which):
        out = np.empty(self.num_levels, np.float32)
        n = self.o._f("get_table")(self.h, which, _ptr(out, f32p), self.num_levels)
        assert n == self.num_levels
        return out

    def extract(self, img, mask=None, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        if cap is None:
            cap = 65536
        kp = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_uint32(0)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        rc = self.o._f("extract")(self.h, _ptr(img, u8p), w, h, img.strides[0],
                                  None if mask is None else _ptr(mask, u8p), 0 if mask is None else mask.strides[0],
                                  kp.ctypes.data, _ptr(desc, u8p), cap, C.byref(n))
        if rc != 0:
            raise RuntimeError(f"oracle extract failed rc={rc} n={n.value}")
        return kp[:n.value].copy(), desc[:n.value].copy()

    def level(self, level, blurred=False):
        w, h = C.c_int(), C.c_int()
        assert self.o._f("level_size")(self.h, level, C.byref(w), C.byref(h)) == 0
        out = np.empty((h.value, w.value), np.uint8)
        rc = self.o._f("level_copy")(self.h, level, int(blurred), _ptr(out, u8p), out.strides[0] if out.size else 0)
        if rc != 0:
            return None
        return out

    def _xys(self, name, level):
        cap = 1 << 20
        out = np.empty((cap, 3), np.int32)
        n = self.o._f(name)(self.h, level, _ptr(out, i32p), cap)
        assert 0 <= n <= cap
        return out[:n].copy()

    def candidates(self, level):
        return self._xys("level_candidates", level)

    def selected(self, level):
        return self._xys("level_selected", level)


_cache = {}


def port():
    if "port" not in _cache:
        build()
        _cache["port"] = Oracle(PORT_SO, "orc")
    return _cache["port"]


def ref():
    """The reference's own sources compiled against the shim; None if never built (e.g. a fresh clone without /root/reference)."""
    if "ref" not in _cache:
        _cache["ref"] = Oracle(REF_SO, "ref") if os.path.exists(REF_SO) else None
    return _cache["ref"]
