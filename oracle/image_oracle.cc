// TEST INFRASTRUCTURE — NOT PRODUCT CODE. See orb_oracle.h.
// Restatement of the image pre-processing that runs before extraction:
//   * util::convert_to_grayscale  (/root/reference/external/openvslam/src/openvslam/util/image_converter.cc:8-39)
//     -> cv::cvtColor(RGB/BGR/RGBA/BGRA -> GRAY), 8-bit: 15-bit fixed point (R 9798, G 19235, B 3735), round half up
//   * LpSlam::ImageProcessing::imadjust + stretchlim (/root/reference/src/Utils/ImageProcessing.h:258-370), used by the
//     "AdjustIntensity" processor (/root/reference/src/Processor/AdjustIntensityProcessor.h:17-40) with low_out -0.3,
//     high_out 1.4 -> cv::calcHist + Mat::convertTo(8U, alpha, beta) (float multiply, float add, cvRound, saturate)
//   * LpSlam::ImageProcessing::Undistort::undistort (ImageProcessing.h:130-250) -> cv::remap(INTER_LINEAR) with cached
//     CV_32FC1 maps, BORDER_CONSTANT 0: 5-bit sub-pixel positions, 15-bit weights, round half up
// The arithmetic of all three lives in OpenCV (not vendored); the restatements are PINNED bit-for-bit against cv2 4.13 in
// tests/test_oracle_vs_cv2.py (convertTo against cv2 with setUseOptimized(false): OpenCV's AVX2 dispatch fuses the
// multiply-add and differs from its own baseline in about one (alpha, beta, value) triple per million).
#include "orb_oracle.h"
#include "cvprim.h"

#include <cmath>
#include <cstring>

extern "C" {

// order: 0 = RGB(A), 1 = BGR(A)
void orc_gray_u8(const uint8_t* src, int w, int h, size_t pitch, int channels, int order, uint8_t* dst, size_t dpitch) {
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint8_t* p = src + static_cast<size_t>(y) * pitch + static_cast<size_t>(x) * channels;
            const int r = order == 0 ? p[0] : p[2], g = p[1], b = order == 0 ? p[2] : p[0];
            dst[static_cast<size_t>(y) * dpitch + x] = static_cast<uint8_t>((b * 3735 + g * 19235 + r * 9798 + (1 << 14)) >> 15);
        }
}

// ImageProcessing.h:258-288 on a 256-bin histogram held as float (cv::calcHist output type)
static void stretchlim_from_hist(const float* hist, double* low_value, double* high_value, double low_fract, double high_fract,
                                 unsigned hist_sum) {
    unsigned sum = 0;
    const unsigned low_count = static_cast<unsigned>(low_fract * hist_sum);
    for (unsigned i = 0; i < 256; ++i) {
        if (sum >= low_count) {
            *low_value = i;
            break;
        }
        sum += hist[i];  // unsigned += float: evaluated in float, converted back (exact below 2^24 pixels)
    }
    const unsigned high_count = static_cast<unsigned>((1 - high_fract) * hist_sum);
    sum = 0;
    for (int i = 255; i >= 0; --i) {
        if (sum >= high_count) {
            *high_value = i;
            break;
        }
        sum += hist[i];
    }
}

// ImageProcessing.h:335-370 for CV_8UC1 -> CV_8UC1. auto_limits != 0: stretchlim with the default 1 % / 99 % fractions.
// lut_out (256 entries, nullable) receives the value map; limits_out (2 doubles, nullable) the low / high input limits.
void orc_imadjust_u8(const uint8_t* src, int w, int h, size_t pitch, int auto_limits, double low_in, double high_in,
                     double low_out, double high_out, uint8_t* dst, size_t dpitch, uint8_t* lut_out, double* limits_out) {
    double use_low = 0.0, use_high = 1.0;
    if (auto_limits) {
        // cv::calcHist counts in integers and converts each bin to float once
        int counts[256];
        float hist[256];
        std::memset(counts, 0, sizeof(counts));
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) ++counts[src[static_cast<size_t>(y) * pitch + x]];
        for (int i = 0; i < 256; ++i) hist[i] = static_cast<float>(counts[i]);
        stretchlim_from_hist(hist, &use_low, &use_high, 0.01, 0.99, static_cast<unsigned>(h * w));
        use_low /= 255u;
        use_high /= 255u;
    }
    else {
        use_low = low_in;
        use_high = high_in;
    }
    double alpha = (high_out - low_out) / (use_high - use_low);
    double beta = high_out - use_high * alpha;
    alpha *= 255 / 255;  // max_out / max_in, integer division
    beta *= 255;
    const float a = static_cast<float>(alpha), b = static_cast<float>(beta);
    uint8_t lut[256];
    for (int v = 0; v < 256; ++v) {
        const float f = static_cast<float>(v) * a + b;  // -ffp-contract=off: two roundings, OpenCV's baseline path
        int r;
        if (!std::isfinite(f)) r = 0;  // cvRound -> INT_MIN on x86 -> saturates to 0
        else r = cvprim::round_half_even(f);
        lut[v] = static_cast<uint8_t>(r < 0 ? 0 : (r > 255 ? 255 : r));
    }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) dst[static_cast<size_t>(y) * dpitch + x] = lut[src[static_cast<size_t>(y) * pitch + x]];
    if (lut_out) std::memcpy(lut_out, lut, 256);
    if (limits_out) {
        limits_out[0] = use_low;
        limits_out[1] = use_high;
    }
}

// cv::remap(src, dst, map_x, map_y, INTER_LINEAR, BORDER_CONSTANT, 0) with CV_32FC1 maps
void orc_remap_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch, const float* map_x, const float* map_y,
                         size_t map_pitch /*floats*/, uint8_t* dst, int dw, int dh, size_t dpitch) {
    auto sat16 = [](int v) { return v < -32768 ? -32768 : (v > 32767 ? 32767 : v); };
    auto px = [&](int y, int x) -> int {
        return (x >= 0 && x < sw && y >= 0 && y < sh) ? src[static_cast<size_t>(y) * spitch + x] : 0;
    };
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            const int sx = cvprim::round_half_even(map_x[static_cast<size_t>(y) * map_pitch + x] * 32.0f);
            const int sy = cvprim::round_half_even(map_y[static_cast<size_t>(y) * map_pitch + x] * 32.0f);
            const int ix = sat16(sx >> 5), iy = sat16(sy >> 5), fx = sx & 31, fy = sy & 31;
            const int w0 = (32 - fy) * (32 - fx) * 32, w1 = (32 - fy) * fx * 32, w2 = fy * (32 - fx) * 32, w3 = fy * fx * 32;
            const int v = (px(iy, ix) * w0 + px(iy, ix + 1) * w1 + px(iy + 1, ix) * w2 + px(iy + 1, ix + 1) * w3 + (1 << 14)) >> 15;
            dst[static_cast<size_t>(y) * dpitch + x] = static_cast<uint8_t>(v);
        }
}

}  // extern "C"
