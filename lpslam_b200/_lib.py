"""ctypes binding of liblpslam_cuda.so (include/lpslam_cuda.h). The library is CUDA only: loading it without a built
.so raises, and every compute call fails with LPSLAM_ERR_CUDA when no device is present — there is no CPU path."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblpslam_cuda.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

OK, ERR_ARG, ERR_CAPACITY, ERR_CUDA, ERR_STATE = 0, -1, -2, -3, -4

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)
f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)
szp = C.POINTER(C.c_size_t)


class OrbParams(C.Structure):
    _fields_ = [("max_num_keypts", C.c_uint32), ("ini_max_num_keypts", C.c_uint32), ("scale_factor", C.c_float),
                ("num_levels", C.c_uint32), ("ini_fast_thr", C.c_uint32), ("min_fast_thr", C.c_uint32),
                ("mask_rects", f32p), ("num_mask_rects", C.c_uint32)]


class ResultDev(C.Structure):
    _fields_ = [("kp", C.c_void_p), ("desc", C.c_void_p), ("offsets", C.c_void_p), ("capacity", C.c_uint32)]


class Camera(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double),
                ("focal_x_baseline", C.c_double), ("true_baseline", C.c_double),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("grid_cols", C.c_uint32), ("grid_rows", C.c_uint32),
                ("inv_cell_width", C.c_double), ("inv_cell_height", C.c_double),
                ("is_monocular", C.c_int32), ("num_scale_levels", C.c_uint32)]


class LpslamCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"liblpslam_cuda error {code}: {msg}")
        self.code = code


_lib = None

# name -> (restype, argtypes); every symbol include/lpslam_cuda.h declares
SIGNATURES = {
    "lpslam_cuda_abi_version": (C.c_int, []),
    "lpslam_cuda_last_error": (C.c_char_p, []),
    "lpslam_cuda_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "lpslam_cuda_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "lpslam_cuda_host_free": (C.c_int, [C.c_void_p]),
    "lpslam_orb_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(OrbParams), C.c_int, C.c_uint32]),
    "lpslam_orb_destroy": (None, [C.c_void_p]),
    "lpslam_orb_set_param": (C.c_int, [C.c_void_p, C.c_int, C.c_double]),
    "lpslam_orb_get_param": (C.c_int, [C.c_void_p, C.c_int, f64p]),
    "lpslam_orb_get_table": (C.c_int, [C.c_void_p, C.c_int, f32p, C.c_uint32]),
    "lpslam_orb_extract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_size_t,
                                     C.c_void_p, C.c_void_p, C.c_uint32, u32p]),
    "lpslam_orb_extract_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_uint32, C.c_int, C.c_int, C.c_size_t,
                                           C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "lpslam_orb_extract_batch_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_size_t]),
    "lpslam_orb_get_result_dev": (C.c_int, [C.c_void_p, C.POINTER(ResultDev)]),
    "lpslam_orb_fetch_result": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "lpslam_orb_input_dev": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), szp, szp]),
    "lpslam_orb_sync": (C.c_int, [C.c_void_p]),
    "lpslam_orb_stream": (C.c_void_p, [C.c_void_p]),
    "lpslam_orb_level_size": (C.c_int, [C.c_void_p, C.c_uint32, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "lpslam_orb_pyramid": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p, C.c_size_t]),
    "lpslam_orb_debug_candidates": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, u32p]),
    "lpslam_orb_debug_selected": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32, u32p]),
    "lpslam_orb_stage_times": (C.c_int, [C.c_void_p, f32p, C.c_uint32]),
    "lpslam_orb_enable_timing": (C.c_int, [C.c_void_p, C.c_int]),
}


def register(name, restype, argtypes):
    SIGNATURES[name] = (restype, argtypes)
    if _lib is not None:  # library already loaded: bind the late registration now
        fn = getattr(_lib, name)
        fn.restype = restype
        fn.argtypes = argtypes


def lib():
    """Load the CUDA library (building it first when nvcc and the sources are present). Fails loudly otherwise."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from . import build as _build
            _build.build()
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc):
    if rc < 0:
        raise LpslamCudaError(rc, lib().lpslam_cuda_last_error().decode(errors="replace"))
    return rc


def ptr(a):
    return None if a is None else a.ctypes.data
