// TEST INFRASTRUCTURE — NOT PRODUCT CODE. See orb_oracle.h.
// Restatement of the per-frame post-processing that follows extraction in data::frame's constructors
// (/root/reference/external/openvslam/src/openvslam/data/frame.cc:37-51,75-92): keypoint undistortion, bearing
// conversion and grid assignment for the perspective camera model (camera/perspective.cc:98-141, data/common.cc:97-118).
// The undistortion arithmetic lives in OpenCV (cv::undistortPoints, not vendored): it is restated here from the published
// algorithm (iterative inversion of the Brown-Conrady model, calib3d undistort: cvUndistortPointsInternal) and PINNED
// bit-for-bit against cv2 4.13 in tests/test_oracle_vs_cv2.py.
#include "orb_oracle.h"
#include "cvprim.h"

#include <cmath>
#include <limits>
#include <vector>

extern "C" {

// camera/perspective.cc:98-130: cv::undistortPoints(mat, mat, K, dist(k1,k2,p1,p2,k3), Mat(), K,
// TermCriteria(EPS | MAX_ITER, 20, 1e-6)); pt replaced, angle / size / octave copied, response / class_id left at the
// cv::KeyPoint defaults (0, -1).
void orc_undistort_keypoints(double fx, double fy, double cx, double cy, const double* dist5, uint32_t n, const orc_keypoint* in,
                             orc_keypoint* out) {
    // camera/perspective.cc:21-22: cv_cam_matrix_ and cv_dist_params_ are cv::Mat_<float>, so the intrinsics reach
    // cv::undistortPoints rounded to float (pinned by the reference build, tests/test_match_ref.py)
    const double K[4] = {(double)(float)fx, (double)(float)fy, (double)(float)cx, (double)(float)cy};
    double D[5];
    for (int i = 0; i < 5; ++i) D[i] = (double)(float)dist5[i];
    for (uint32_t i = 0; i < n; ++i) {
        out[i] = in[i];
        cvprim::undistort_point(K, D, K, 20, 1e-6, in[i].x, in[i].y, &out[i].x, &out[i].y);
        out[i].response = 0.0f;
        out[i].class_id = -1;
    }
}

// camera/perspective.cc:132-141
void orc_keypoints_to_bearings(double fx, double fy, double cx, double cy, uint32_t n, const orc_keypoint* kp, double* bearings) {
    for (uint32_t i = 0; i < n; ++i) {
        const double xn = (kp[i].x - cx) / fx, yn = (kp[i].y - cy) / fy;
        const double l2 = std::sqrt(xn * xn + yn * yn + 1.0);
        bearings[3 * i] = xn / l2;
        bearings[3 * i + 1] = yn / l2;
        bearings[3 * i + 2] = 1.0 / l2;
    }
}

// data/common.cc:97-118 as CSR over cell = cx * grid_rows + cy, indices in insertion (ascending) order
void orc_assign_grid(const orc_camera* cam, uint32_t n, const orc_keypoint* kp, uint32_t* cell_start, uint32_t* items) {
    const int gc = static_cast<int>(cam->grid_cols), gr = static_cast<int>(cam->grid_rows);
    std::vector<std::vector<uint32_t>> grid(static_cast<size_t>(gc) * gr);
    for (uint32_t i = 0; i < n; ++i) {
        const int cx = cvprim::round_half_even((kp[i].x - cam->min_x) * cam->inv_cell_width);
        const int cy = cvprim::round_half_even((kp[i].y - cam->min_y) * cam->inv_cell_height);
        if (0 <= cx && cx < gc && 0 <= cy && cy < gr) grid[static_cast<size_t>(cx) * gr + cy].push_back(i);
    }
    uint32_t pos = 0;
    for (size_t c = 0; c < grid.size(); ++c) {
        cell_start[c] = pos;
        for (uint32_t i : grid[c]) items[pos++] = i;
    }
    cell_start[grid.size()] = pos;
}

}  // extern "C"
