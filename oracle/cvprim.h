// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// CPU restatement of the OpenCV primitives that the reference's ORB front end calls.
// OpenCV is an external dependency of the reference (find_package(OpenCV 4.2 REQUIRED) at
// /root/reference/CMakeLists.txt:53-55; not vendored), so the arithmetic is restated here from
// OpenCV 4.x's published algorithms and pinned bit-for-bit against cv2 4.13 by
// tests/test_oracle_vs_cv2.py. Call sites in the reference:
//   cv::resize        external/openvslam/src/openvslam/feature/orb_extractor.cc:273
//   cv::GaussianBlur  external/openvslam/src/openvslam/feature/orb_extractor.cc:136
//   cv::FAST          external/openvslam/src/openvslam/feature/orb_extractor.cc:343,348
//   cv::fastAtan2     external/openvslam/src/openvslam/feature/orb_extractor.cc:633
//   cvRound/cvFloor/cvCeil  orb_extractor.cc:614,649,680-681; orb_extractor_node.cc:8-9
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this code.
#ifndef LPSLAM_ORACLE_CVPRIM_H
#define LPSLAM_ORACLE_CVPRIM_H

#include <cstdint>
#include <cstddef>
#include <vector>

namespace cvprim {

// cvRound == lrint under the default rounding mode (round-half-even).
int round_half_even(double v);
int floor_int(double v);
int ceil_int(double v);

// cv::borderInterpolate(p, len, BORDER_REFLECT_101)
int reflect101(int p, int len);

// cv::resize(src, dst, Size(dw,dh), 0, 0, INTER_LINEAR) for CV_8UC1.
void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch,
                      uint8_t* dst, int dw, int dh, size_t dpitch);

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 (src may alias dst).
void gaussian7_u8(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch);

struct FastPoint {
    int x, y, score;
};

// FAST-9/16 corner strength of one pixel ("best" in SURVEY Appendix A.3): the pixel is a corner at threshold t
// iff best > t, and then its score is best - 1. Caller guarantees a 3-px margin around (x, y).
int fast_best(const uint8_t* p, ptrdiff_t pitch);

// cv::FAST(img, kps, thr, /*nonmaxSuppression=*/true), TYPE_9_16, on a w x h sub-image. Output in raster order.
void fast9_nms(const uint8_t* img, int w, int h, size_t pitch, int thr, std::vector<FastPoint>& out);

// cv::fastAtan2(y, x) in degrees, float32, no FMA contraction.
float fast_atan2(float y, float x);

// cv::rectangle(mask, (x0,y0), (x1,y1), Scalar(0), -1, LINE_AA) on an 8UC1 image: for an axis-aligned integer
// rectangle the anti-aliased fill degenerates to an exact inclusive fill (checked against cv2 in the tests).
void fill_rect_zero(uint8_t* img, int w, int h, size_t pitch, int x0, int y0, int x1, int y1);

// cv::undistortPoints for one point: K = (fx, fy, cx, cy), dist = (k1, k2, p1, p2, k3), new camera matrix P = (fx', fy', cx', cy'),
// R = identity, termination after max_count iterations or a reprojection error below eps (calib3d cvUndistortPointsInternal).
void undistort_point(const double K[4], const double dist[5], const double P[4], int max_count, double eps, float u, float v,
                     float* out_x, float* out_y);

} // namespace cvprim

#endif
