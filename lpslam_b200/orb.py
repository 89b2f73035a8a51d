"""Python mirror of openvslam::feature::orb_extractor (feature/orb_extractor.h:13-151) over the C ABI.

Same member names, argument meaning and error behaviour as the reference class; keypoints come back as a structured
numpy array with cv::KeyPoint's fields, descriptors as an (N, 32) uint8 array.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, OrbParams, check, lib


class orb_params:
    """feature/orb_params.h:9-46 (defaults: 2000 keypoints, scale 1.2, 8 levels, FAST 20 / 7)."""

    def __init__(self, max_num_keypts=2000, ini_max_num_keypts=None, scale_factor=1.2, num_levels=8, ini_fast_thr=20,
                 min_fast_thr=7, mask_rects=()):
        self.max_num_keypts_ = int(max_num_keypts)
        self.ini_max_num_keypts_ = int(2 * max_num_keypts if ini_max_num_keypts is None else ini_max_num_keypts)
        self.scale_factor_ = float(scale_factor)
        self.num_levels_ = int(num_levels)
        self.ini_fast_thr_ = int(ini_fast_thr)
        self.min_fast_thr = int(min_fast_thr)
        self.mask_rects_ = [list(map(float, r)) for r in mask_rects]
        for r in self.mask_rects_:  # feature/orb_params.cc:16-26
            if len(r) != 4:
                raise RuntimeError("mask rectangle must contain four parameters")
            if r[0] >= r[1]:
                raise RuntimeError("x_max must be greater than x_min")
            if r[2] >= r[3]:
                raise RuntimeError("y_max must be greater than x_min")


class orb_extractor:
    def __init__(self, max_num_keypts=2000, ini_max_num_keypts=None, scale_factor=1.2, num_levels=8, ini_fast_thr=20,
                 min_fast_thr=7, mask_rects=(), device=0, max_batch=1, params=None):
        p = params if params is not None else orb_params(max_num_keypts, ini_max_num_keypts, scale_factor, num_levels,
                                                         ini_fast_thr, min_fast_thr, mask_rects)
        self.orb_params_ = p
        self.device = device
        self.max_batch = max_batch
        rects = np.ascontiguousarray(np.asarray(p.mask_rects_, np.float32).reshape(-1, 4))
        cp = OrbParams(p.max_num_keypts_, p.ini_max_num_keypts_, p.scale_factor_, p.num_levels_, p.ini_fast_thr_,
                       p.min_fast_thr, rects.ctypes.data_as(_lib.f32p) if len(rects) else None, len(rects))
        h = C.c_void_p()
        self._h = None
        check(lib().lpslam_orb_create(C.byref(h), C.byref(cp), device, max_batch))
        self._h = h
        self.image_pyramid_ = _Pyramid(self)

    def close(self):
        if getattr(self, "_h", None) and lib is not None:  # `lib` is already gone at interpreter shutdown
            lib().lpslam_orb_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- extract (feature/orb_extractor.cc:71-147) ----
    def extract(self, in_image, in_image_mask=None, cap=None):
        img = np.asarray(in_image)
        if img.size == 0:  # :73-75, outputs untouched
            return np.zeros(0, KP_DTYPE), None
        assert img.dtype == np.uint8 and img.ndim == 2  # :79 assert(image.type() == CV_8UC1)
        if img.strides[1] != 1:
            img = np.ascontiguousarray(img)
        h, w = img.shape
        mask = None
        if in_image_mask is not None and np.asarray(in_image_mask).size:
            mask = np.asarray(in_image_mask)
            assert mask.dtype == np.uint8 and mask.shape == img.shape  # :96
            if mask.strides[1] != 1:
                mask = np.ascontiguousarray(mask)
        cap = int(cap) if cap else 8 * self.orb_params_.max_num_keypts_ + 256
        kp = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = C.c_uint32(0)
        check(lib().lpslam_orb_extract(self._h, img.ctypes.data, w, h, img.strides[0],
                                       None if mask is None else mask.ctypes.data, 0 if mask is None else mask.strides[0],
                                       kp.ctypes.data, desc.ctypes.data, cap, C.byref(n)))
        if n.value == 0:
            return np.zeros(0, KP_DTYPE), None  # :115-117 out_descriptors.release()
        return kp[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, images, cap_per_frame=None):
        """Frame-parallel replay: a list / (B, H, W) array of same-size 8UC1 frames -> list of (keypts, descriptors)."""
        imgs = [np.ascontiguousarray(im, np.uint8) for im in images]
        n = len(imgs)
        h, w = imgs[0].shape
        cap = n * (int(cap_per_frame) if cap_per_frame else 8 * self.orb_params_.max_num_keypts_ + 256)
        kp = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        offs = np.zeros(n + 1, np.uint32)
        ptrs = (C.c_void_p * n)(*[im.ctypes.data for im in imgs])
        check(lib().lpslam_orb_extract_batch(self._h, ptrs, n, w, h, w, kp.ctypes.data, desc.ctypes.data, cap,
                                             offs.ctypes.data))
        return [(kp[offs[f]:offs[f + 1]].copy(), desc[offs[f]:offs[f + 1]].copy()) for f in range(n)]

    def input_dev(self, width, height):
        """(ptr, pitch, frame_stride) of the context's own level-0 batch buffer: producers on the device write frames here"""
        ptr, pitch, stride = C.c_void_p(), C.c_size_t(), C.c_size_t()
        check(lib().lpslam_orb_input_dev(self._h, width, height, C.byref(ptr), C.byref(pitch), C.byref(stride)))
        return ptr.value, pitch.value, stride.value

    def extract_batch_dev(self, n, width, height, cap_per_frame=None):
        """extraction of the n frames a device producer left in input_dev(); returns a list of (keypts, descriptors)"""
        ptr, pitch, _ = self.input_dev(width, height)
        check(lib().lpslam_orb_extract_batch_dev(self._h, ptr, n, width, height, pitch))
        cap = n * (int(cap_per_frame) if cap_per_frame else 8 * self.orb_params_.max_num_keypts_ + 256)
        kp = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        offs = np.zeros(n + 1, np.uint32)
        check(lib().lpslam_orb_fetch_result(self._h, kp.ctypes.data, desc.ctypes.data, cap, offs.ctypes.data))
        return [(kp[offs[f]:offs[f + 1]].copy(), desc[offs[f]:offs[f + 1]].copy()) for f in range(n)]

    # ---- getters / setters (feature/orb_extractor.cc:149-206) ----
    def _get(self, which):
        v = C.c_double()
        check(lib().lpslam_orb_get_param(self._h, which, C.byref(v)))
        return v.value

    def get_max_num_keypoints(self):
        return int(self._get(0))

    def set_max_num_keypoints(self, max_num_keypts):
        check(lib().lpslam_orb_set_param(self._h, 0, float(max_num_keypts)))
        self.orb_params_.max_num_keypts_ = int(max_num_keypts)

    def get_scale_factor(self):
        return float(np.float32(self._get(1)))

    def set_scale_factor(self, scale_factor):
        check(lib().lpslam_orb_set_param(self._h, 1, float(scale_factor)))
        self.orb_params_.scale_factor_ = float(scale_factor)

    def get_num_scale_levels(self):
        return int(self._get(2))

    def set_num_scale_levels(self, num_levels):
        check(lib().lpslam_orb_set_param(self._h, 2, float(num_levels)))
        self.orb_params_.num_levels_ = int(num_levels)

    def get_initial_fast_threshold(self):
        return int(self._get(3))

    def set_initial_fast_threshold(self, initial_fast_threshold):
        check(lib().lpslam_orb_set_param(self._h, 3, float(initial_fast_threshold)))
        self.orb_params_.ini_fast_thr_ = int(initial_fast_threshold)

    def get_minimum_fast_threshold(self):
        return int(self._get(4))

    def set_minimum_fast_threshold(self, minimum_fast_threshold):
        check(lib().lpslam_orb_set_param(self._h, 4, float(minimum_fast_threshold)))
        self.orb_params_.min_fast_thr = int(minimum_fast_threshold)

    def _table(self, which):
        n = self.get_num_scale_levels()
        out = np.empty(n, np.float32)
        check(lib().lpslam_orb_get_table(self._h, which, out.ctypes.data_as(_lib.f32p), n))
        return out

    def get_scale_factors(self):
        return self._table(0)

    def get_inv_scale_factors(self):
        return self._table(1)

    def get_level_sigma_sq(self):
        return self._table(2)

    def get_inv_level_sigma_sq(self):
        return self._table(3)

    def get_num_keypts_per_level(self):
        return self._table(4).astype(np.uint32)

    # ---- intermediates (test hooks) ----
    def level(self, level, blurred=False, frame=0):
        w, h = C.c_int(), C.c_int()
        check(lib().lpslam_orb_level_size(self._h, level, C.byref(w), C.byref(h)))
        out = np.empty((h.value, w.value), np.uint8)
        check(lib().lpslam_orb_pyramid(self._h, frame, level, int(blurred), out.ctypes.data, out.strides[0]))
        return out

    def _points(self, fn, level, frame):
        cap = 1 << 16
        while True:
            out = np.empty((cap, 3), np.int32)
            n = C.c_uint32(0)
            rc = fn(self._h, frame, level, out.ctypes.data, cap, C.byref(n))
            if rc == _lib.ERR_CAPACITY:
                cap = int(n.value)
                continue
            check(rc)
            return out[:n.value].copy()

    def candidates(self, level, frame=0):
        return self._points(lib().lpslam_orb_debug_candidates, level, frame)

    def selected(self, level, frame=0):
        return self._points(lib().lpslam_orb_debug_selected, level, frame)

    def enable_timing(self, on=True):
        check(lib().lpslam_orb_enable_timing(self._h, int(on)))

    def stage_times(self):
        ms = np.zeros(7, np.float32)
        check(lib().lpslam_orb_stage_times(self._h, ms.ctypes.data_as(_lib.f32p), 7))
        return dict(zip(["h2d", "pyramid", "fast", "quadtree", "blur", "describe", "d2h"], ms.tolist()))


class _Pyramid:
    """image_pyramid_ (feature/orb_extractor.h:76): levels are fetched lazily from the device on access."""

    def __init__(self, ext):
        self._e = ext

    def __len__(self):
        return self._e.get_num_scale_levels()

    def __getitem__(self, level):
        if level < 0:
            level += len(self)
        if not 0 <= level < len(self):
            raise IndexError(level)
        return self._e.level(level)

    at = __getitem__
