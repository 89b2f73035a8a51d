// TEST INFRASTRUCTURE - cv::undistortPoints for the one call shape of camera/perspective.{h,cc}: N two-channel float points,
// 3x3 float camera matrix, 5 x 1 float distortion vector, R empty, P = camera matrix, explicit termination criteria. The
// arithmetic is cvprim::undistort_points (restated from calib3d, pinned bit-for-bit against cv2 in tests/test_oracle_vs_cv2.py).
#pragma once
#include <opencv2/core.hpp>
namespace cv {
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& /*R*/, const Mat& P, TermCriteria crit) {
    assert(src.type() == CV_32F && src.channels() == 2 && K.type() == CV_32F && D.type() == CV_32F && D.rows * D.cols == 5);
    const double k[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    const double p[4] = {P.at<float>(0, 0), P.at<float>(1, 1), P.at<float>(0, 2), P.at<float>(1, 2)};
    double d[5];
    for (int i = 0; i < 5; ++i) d[i] = D.rows == 5 ? D.at<float>(i, 0) : D.at<float>(0, i);
    Mat out = src.data == dst.data ? src : Mat(src.rows, src.cols * 2, CV_32F).reshape(2);
    for (int r = 0; r < src.rows; ++r)
        for (int c = 0; c < src.cols; ++c) {
            const float* s = src.ptr<float>(r) + 2 * c;
            float* o = out.ptr<float>(r) + 2 * c;
            cvprim::undistort_point(k, d, p, crit.maxCount, crit.epsilon, s[0], s[1], &o[0], &o[1]);
        }
    dst = out;
}
}  // namespace cv
