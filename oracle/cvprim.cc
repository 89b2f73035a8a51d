// TEST INFRASTRUCTURE — NOT PRODUCT CODE. See cvprim.h for the provenance of every function.
// Build without -ffast-math and without -march=native / -mfma so that no FMA contraction happens
// (the reference builds with BUILD_WITH_MARCH_NATIVE OFF, /root/reference/CMakeLists.txt:155).
#include "cvprim.h"

#include <limits>

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace cvprim {

int round_half_even(double v) { return static_cast<int>(std::lrint(v)); }
int floor_int(double v) {
    int i = static_cast<int>(v);
    return i - (v < i);
}
int ceil_int(double v) {
    int i = static_cast<int>(v);
    return i + (v > i);
}

int reflect101(int p, int len) {
    if (static_cast<unsigned>(p) < static_cast<unsigned>(len)) return p;
    if (len == 1) return 0;
    do {
        if (p < 0)
            p = -p;
        else
            p = 2 * (len - 1) - p;
    } while (static_cast<unsigned>(p) >= static_cast<unsigned>(len));
    return p;
}

// ---------------------------------------------------------------------------------------------
// resize, INTER_LINEAR, 8UC1 (OpenCV imgproc/resize.cpp: resizeGeneric_ with HResizeLinear<uchar,int,short,2048>
// and VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>). SURVEY Appendix A.1.
// ---------------------------------------------------------------------------------------------
namespace {
struct AxisTab {
    std::vector<int> ofs;     // source index
    std::vector<short> c0, c1; // 11-bit coefficients
};

AxisTab make_axis_tab(int sn, int dn, bool clamp_reset) {
    AxisTab t;
    t.ofs.resize(dn);
    t.c0.resize(dn);
    t.c1.resize(dn);
    const double scale = static_cast<double>(sn) / dn;
    for (int d = 0; d < dn; ++d) {
        float f = static_cast<float>((d + 0.5) * scale - 0.5);
        int s = floor_int(f);
        f -= s;
        if (clamp_reset) {
            // horizontal axis: coefficient reset at both ends (resize.cpp: "if (sx < 0) fx = 0, sx = 0" ...)
            if (s < 0) {
                s = 0;
                f = 0.f;
            }
            if (s >= sn - 1) {
                s = sn - 1;
                f = 0.f;
            }
        }
        t.ofs[d] = s;
        const float w0 = (1.f - f) * 2048.f;
        const float w1 = f * 2048.f;
        t.c0[d] = static_cast<short>(round_half_even(w0));
        t.c1[d] = static_cast<short>(round_half_even(w1));
    }
    return t;
}
} // namespace

void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch,
                      uint8_t* dst, int dw, int dh, size_t dpitch) {
    if (sw == dw && sh == dh) {
        for (int y = 0; y < dh; ++y) std::memcpy(dst + y * dpitch, src + y * spitch, dw);
        return;
    }
    // resize.cpp: INTER_LINEAR with an exact 2x2 decimation is routed to the INTER_AREA fast path.
    if (sw == 2 * dw && sh == 2 * dh) {
        for (int y = 0; y < dh; ++y) {
            const uint8_t* s0 = src + (2 * y) * spitch;
            const uint8_t* s1 = s0 + spitch;
            uint8_t* d = dst + y * dpitch;
            for (int x = 0; x < dw; ++x)
                d[x] = static_cast<uint8_t>((s0[2 * x] + s0[2 * x + 1] + s1[2 * x] + s1[2 * x + 1] + 2) >> 2);
        }
        return;
    }
    const AxisTab tx = make_axis_tab(sw, dw, true);
    const AxisTab ty = make_axis_tab(sh, dh, false);
    std::vector<int> row0(dw), row1(dw);
    auto hpass = [&](int sy, std::vector<int>& out) {
        sy = std::min(std::max(sy, 0), sh - 1);
        const uint8_t* s = src + static_cast<size_t>(sy) * spitch;
        for (int d = 0; d < dw; ++d) {
            const int sx = tx.ofs[d];
            const int sx1 = std::min(sx + 1, sw - 1);
            out[d] = s[sx] * tx.c0[d] + s[sx1] * tx.c1[d];
        }
    };
    for (int y = 0; y < dh; ++y) {
        const int sy = ty.ofs[y];
        hpass(sy, row0);
        hpass(sy + 1, row1);
        const int b0 = ty.c0[y], b1 = ty.c1[y];
        uint8_t* d = dst + static_cast<size_t>(y) * dpitch;
        for (int x = 0; x < dw; ++x) {
            const int v = (((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2;
            d[x] = static_cast<uint8_t>(std::min(std::max(v, 0), 255));
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur 7x7, sigma 2, 8UC1 (OpenCV imgproc/smooth.dispatch.cpp: getGaussianKernelBitExact ->
// ufixedpoint16 kernel [18 34 48 56 48 34 18]/256, GaussianBlurFixedPoint). SURVEY Appendix A.2.
// ---------------------------------------------------------------------------------------------
void gaussian7_u8(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch) {
    static const int k[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> hbuf(static_cast<size_t>(w) * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + static_cast<size_t>(y) * spitch;
        uint16_t* o = hbuf.data() + static_cast<size_t>(y) * w;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int j = 0; j < 7; ++j) acc += k[j] * s[reflect101(x + j - 3, w)];
            o[x] = static_cast<uint16_t>(acc);
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* rows[7];
        for (int i = 0; i < 7; ++i) rows[i] = hbuf.data() + static_cast<size_t>(reflect101(y + i - 3, h)) * w;
        uint8_t* d = dst + static_cast<size_t>(y) * dpitch;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int i = 0; i < 7; ++i) acc += static_cast<uint32_t>(k[i]) * rows[i][x];
            d[x] = static_cast<uint8_t>((acc + 32768u) >> 16);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// FAST-9/16 with non-max suppression (OpenCV features2d/fast.cpp FAST_t<16>, fast_score.cpp
// cornerScore<16>). SURVEY Appendix A.3.
// ---------------------------------------------------------------------------------------------
int fast_best(const uint8_t* p, ptrdiff_t pitch) {
    static const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    static const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    int d[16];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[dy[k] * pitch + dx[k]];
    int a = -256, b = -256;
    for (int s = 0; s < 16; ++s) {
        int mn = 256, mx = -256;
        for (int j = 0; j < 9; ++j) {
            const int t = d[(s + j) & 15];
            mn = std::min(mn, t);
            mx = std::max(mx, t);
        }
        a = std::max(a, mn);
        b = std::max(b, -mx);
    }
    return std::max(a, b);
}

void fast9_nms(const uint8_t* img, int w, int h, size_t pitch, int thr, std::vector<FastPoint>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    // score = best-1 for corners (best > thr), 0 elsewhere, including the 3-px frame of the sub-image
    std::vector<int> score(static_cast<size_t>(w) * h, 0);
    std::vector<uint8_t> corner(static_cast<size_t>(w) * h, 0);
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const int best = fast_best(img + static_cast<size_t>(y) * pitch + x, static_cast<ptrdiff_t>(pitch));
            if (best > thr) {
                score[static_cast<size_t>(y) * w + x] = best - 1;
                corner[static_cast<size_t>(y) * w + x] = 1;
            }
        }
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            if (!corner[static_cast<size_t>(y) * w + x]) continue;
            const int* c = &score[static_cast<size_t>(y) * w + x];
            const int s = c[0];
            if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] && s > c[w - 1] && s > c[w]
                && s > c[w + 1])
                out.push_back({x, y, s});
        }
}

// ---------------------------------------------------------------------------------------------
// fastAtan2 (OpenCV core/mathfuncs_core.simd.hpp atan_f32 / atanImpl). SURVEY Appendix A.4.
// ---------------------------------------------------------------------------------------------
float fast_atan2(float y, float x) {
    static const float scale = static_cast<float>(180.0 / 3.1415926535897932384626433832795);
    static const float p1 = 0.9997878412794807f * scale;
    static const float p3 = -0.3258083974640975f * scale;
    static const float p5 = 0.1555786518463281f * scale;
    static const float p7 = -0.04432655554792128f * scale;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + static_cast<float>(DBL_EPSILON));
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    else {
        c = ax / (ay + static_cast<float>(DBL_EPSILON));
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

void fill_rect_zero(uint8_t* img, int w, int h, size_t pitch, int x0, int y0, int x1, int y1) {
    if (x0 > x1) std::swap(x0, x1);
    if (y0 > y1) std::swap(y0, y1);
    x0 = std::max(x0, 0);
    y0 = std::max(y0, 0);
    x1 = std::min(x1, w - 1);
    y1 = std::min(y1, h - 1);
    for (int y = y0; y <= y1; ++y)
        for (int x = x0; x <= x1; ++x) img[static_cast<size_t>(y) * pitch + x] = 0;
}

// cv::undistortPoints (calib3d, cvUndistortPointsInternal) for one point of a 5-coefficient Brown-Conrady model; the
// rational and thin-prism terms of the general model are zero and are kept as literal zeros so that the floating-point
// expression trees stay those of OpenCV. Pinned bit-for-bit against cv2 4.13 in tests/test_oracle_vs_cv2.py.
void undistort_point(const double K[4], const double dist[5], const double P[4], int max_count, double eps, float uf, float vf,
                     float* out_x, float* out_y) {
    const double fx = K[0], fy = K[1], cx = K[2], cy = K[3];
    const double k0 = dist[0], k1 = dist[1], p1 = dist[2], p2 = dist[3], k4 = dist[4];
    const double ifx = 1.0 / fx, ify = 1.0 / fy;
    const double u = uf, v = vf;
    double x = (u - cx) * ifx, y = (v - cy) * ify;
    const double x0 = x, y0 = y;
    double error = std::numeric_limits<double>::max();
    for (int j = 0;; ++j) {
        if (j >= max_count) break;
        if (error < eps) break;
        double r2 = x * x + y * y;
        const double icdist = (1 + ((0.0 * r2 + 0.0) * r2 + 0.0) * r2) / (1 + ((k4 * r2 + k1) * r2 + k0) * r2);
        if (icdist < 0) {
            x = (u - cx) * ifx;
            y = (v - cy) * ify;
            break;
        }
        const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x) + 0.0 * r2 + 0.0 * r2 * r2;
        const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y + 0.0 * r2 + 0.0 * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
        {
            r2 = x * x + y * y;
            const double r4 = r2 * r2, r6 = r4 * r2;
            const double a1 = 2 * x * y, a2 = r2 + 2 * x * x, a3 = r2 + 2 * y * y;
            const double cdist = 1 + k0 * r2 + k1 * r4 + k4 * r6;
            const double icdist2 = 1. / (1 + 0.0 * r2 + 0.0 * r4 + 0.0 * r6);
            const double xd0 = x * cdist * icdist2 + p1 * a1 + p2 * a2 + 0.0 * r2 + 0.0 * r4;
            const double yd0 = y * cdist * icdist2 + p1 * a3 + p2 * a1 + 0.0 * r2 + 0.0 * r4;
            const double x_proj = xd0 * fx + cx, y_proj = yd0 * fy + cy;
            error = std::sqrt((x_proj - u) * (x_proj - u) + (y_proj - v) * (y_proj - v));
        }
    }
    // new camera matrix P, R = identity: RR = P * I
    const double xx = P[0] * x + 0.0 * y + P[2], yy = 0.0 * x + P[1] * y + P[3], ww = 1. / (0.0 * x + 0.0 * y + 1.0);
    *out_x = static_cast<float>(xx * ww);
    *out_y = static_cast<float>(yy * ww);
}

} // namespace cvprim
