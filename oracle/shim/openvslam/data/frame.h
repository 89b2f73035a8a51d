// TEST INFRASTRUCTURE - stand-in for data/frame.h (the real one pulls in g2o, the BoW vocabulary, the extractor and the
// navigation state). Same member names and types as /root/reference/external/openvslam/src/openvslam/data/frame.h:177-268
// for everything the matchers of match/*.cc and data/landmark.cc read or write; get_keypoints_in_cell forwards to the
// reference's own data/common.cc exactly as data/frame.cc:207-209 does.
#pragma once
#include "openvslam/type.h"
#include "openvslam/camera/base.h"
#include "openvslam/data/common.h"

#include <fbow/fbow.h>
#include <opencv2/core.hpp>
#include <vector>

namespace openvslam {
namespace data {

class keyframe;
class landmark;

class frame {
public:
    frame() = default;
    std::vector<unsigned int> get_keypoints_in_cell(const float ref_x, const float ref_y, const float margin, const int min_level = -1,
                                                    const int max_level = -1) const {
        return data::get_keypoints_in_cell(camera_, undist_keypts_, keypt_indices_in_cells_, ref_x, ref_y, margin, min_level, max_level);
    }
    unsigned int id_ = 0;
    camera::base* camera_ = nullptr;
    unsigned int num_keypts_ = 0;
    std::vector<cv::KeyPoint> keypts_;
    std::vector<cv::KeyPoint> undist_keypts_;
    eigen_alloc_vector<Vec3_t> bearings_;
    std::vector<float> stereo_x_right_;
    std::vector<float> depths_;
    fbow::BoWVector bow_vec_;
    fbow::BoWFeatVector bow_feat_vec_;
    cv::Mat descriptors_;
    std::vector<landmark*> landmarks_;
    std::vector<bool> outlier_flags_;
    std::vector<std::vector<std::vector<unsigned int>>> keypt_indices_in_cells_;
    Mat44_t cam_pose_cw_;
    unsigned int num_scale_levels_ = 0;
    float scale_factor_ = 0, log_scale_factor_ = 0;
    std::vector<float> scale_factors_, inv_scale_factors_, level_sigma_sq_, inv_level_sigma_sq_;
};

}  // namespace data
}  // namespace openvslam
