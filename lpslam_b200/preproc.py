"""Python mirror of the image pre-processing that precedes extraction in the reference, over the C ABI:
util::convert_to_grayscale (util/image_converter.cc:8-39), LpSlam::ImageProcessing::imadjust (src/Utils/ImageProcessing.h:335-370,
the "AdjustIntensity" processor) and LpSlam::ImageProcessing::Undistort (ImageProcessing.h:130-250, cv::remap with cached maps)."""
import ctypes as C

import numpy as np

from ._lib import check, lib, register

register("lpslam_gray", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t])
register("lpslam_gray_dev", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_uint32, C.c_void_p,
                                      C.c_size_t, C.c_size_t, C.c_void_p])
register("lpslam_imadjust", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                                      C.c_void_p, C.c_size_t, C.c_void_p])
register("lpslam_imadjust_scratch_bytes", C.c_size_t, [C.c_uint32])
register("lpslam_imadjust_dev", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_uint32, C.c_int, C.c_double,
                                          C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p])
# no reference counterpart: imadjust with the scratch in a fixed array inside the library, for hosts that may not allocate
register("lpslam_imadjust_dev_static", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_uint32, C.c_int, C.c_double,
                                                 C.c_double, C.c_double, C.c_double, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
                                                 C.c_void_p])
IMADJUST_STATIC_MAX_IMAGES = 64  # LPSLAM_IMADJUST_STATIC_MAX_IMAGES in include/lpslam_cuda.h
register("lpslam_remap_linear", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p,
                                          C.c_int, C.c_int, C.c_size_t])
register("lpslam_remap_linear_dev", C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p, C.c_size_t,
                                              C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_uint32, C.c_void_p])

COLOR_ORDER = {"RGB": 0, "BGR": 1}


def convert_to_grayscale(img, in_color_order="BGR"):
    """util/image_converter.cc:8-39: 3- or 4-channel 8-bit image -> gray; single-channel / 'Gray' input is returned as is."""
    img = np.ascontiguousarray(img, np.uint8)
    if img.ndim == 2 or in_color_order == "Gray":
        return img
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    check(lib().lpslam_gray(img.ctypes.data, w, h, img.strides[0], ch, COLOR_ORDER[in_color_order], out.ctypes.data, w))
    return out


def imadjust(src, low_in=None, high_in=None, low_out=0.0, high_out=1.0, return_limits=False):
    """ImageProcessing.h:335-370 for CV_8UC1; low_in / high_in None = std::nullopt (stretchlim, 1 % / 99 %)."""
    src = np.ascontiguousarray(src, np.uint8)
    h, w = src.shape
    out = np.empty_like(src)
    lim = np.zeros(2, np.float64)
    auto = low_in is None or high_in is None
    check(lib().lpslam_imadjust(src.ctypes.data, w, h, src.strides[0], int(auto), 0.0 if auto else float(low_in),
                                1.0 if auto else float(high_in), float(low_out), float(high_out), out.ctypes.data, w, lim.ctypes.data))
    return (out, lim) if return_limits else out


class AdjustIntensityProcessor:
    """src/Processor/AdjustIntensityProcessor.h:17-40: imadjust(image, nullopt, nullopt, -0.3, 1.4)"""

    def type(self):
        return "AdjustIntensity"

    def processImage(self, image, image_second=None):
        out = imadjust(image, None, None, -0.3, 1.4)
        return out if image_second is None else (out, imadjust(image_second, None, None, -0.3, 1.4))


class Undistort:
    """ImageProcessing.h:130-250: holds the undistortion maps (built once on the host by the caller, e.g. with
    cv::initUndistortRectifyMap as the reference does) and remaps every image with INTER_LINEAR."""

    def __init__(self, map_x, map_y):
        self.map_x = np.ascontiguousarray(map_x, np.float32)
        self.map_y = np.ascontiguousarray(map_y, np.float32)
        assert self.map_x.shape == self.map_y.shape and self.map_x.ndim == 2

    def undistort(self, image):
        image = np.ascontiguousarray(image, np.uint8)
        dh, dw = self.map_x.shape
        out = np.empty((dh, dw), np.uint8)
        check(lib().lpslam_remap_linear(image.ctypes.data, image.shape[1], image.shape[0], image.strides[0], self.map_x.ctypes.data,
                                        self.map_y.ctypes.data, dw, out.ctypes.data, dw, dh, dw))
        return out
