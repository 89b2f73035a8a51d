// TEST INFRASTRUCTURE - stand-in for data/keyframe.h (the real one pulls in g2o, the graph node, the BoW database).
// Member names, types and the accessor bodies follow /root/reference/external/openvslam/src/openvslam/data/keyframe.h and
// keyframe.cc for everything the matchers of match/*.cc and data/landmark.cc touch: get_landmarks / get_valid_landmarks /
// get_landmark / add_landmark / erase_landmark_with_index / replace_landmark (keyframe.cc:199-262), the pose accessors
// (:142-176) and get_keypoints_in_cell (:287-289, forwarding to the reference's data/common.cc).
#pragma once
#include "openvslam/type.h"
#include "openvslam/camera/base.h"
#include "openvslam/data/common.h"
#include "openvslam/data/landmark.h"

#include <fbow/fbow.h>
#include <opencv2/core.hpp>
#include <set>
#include <vector>

namespace openvslam {
namespace data {

class keyframe {
public:
    keyframe() = default;
    void set_cam_pose(const Mat44_t& cam_pose_cw) {  // keyframe.cc:119-140
        cam_pose_cw_ = cam_pose_cw;
        const Mat33_t rot_cw = cam_pose_cw_.block<3, 3>(0, 0);
        const Vec3_t trans_cw = cam_pose_cw_.block<3, 1>(0, 3);
        const Mat33_t rot_wc = rot_cw.transpose();
        cam_center_ = -rot_wc * trans_cw;
    }
    Vec3_t get_cam_center() const { return cam_center_; }
    Mat33_t get_rotation() const { return cam_pose_cw_.block<3, 3>(0, 0); }
    Vec3_t get_translation() const { return cam_pose_cw_.block<3, 1>(0, 3); }

    void add_landmark(landmark* lm, const unsigned int idx) { landmarks_.at(idx) = lm; }
    void erase_landmark_with_index(const unsigned int idx) { landmarks_.at(idx) = nullptr; }
    void replace_landmark(landmark* lm, const unsigned int idx) { landmarks_.at(idx) = lm; }
    std::vector<landmark*> get_landmarks() const { return landmarks_; }
    std::set<landmark*> get_valid_landmarks() const {
        std::set<landmark*> valid;
        for (const auto lm : landmarks_) {
            if (!lm) continue;
            if (lm->will_be_erased()) continue;
            valid.insert(lm);
        }
        return valid;
    }
    landmark* get_landmark(const unsigned int idx) const { return landmarks_.at(idx); }
    std::vector<unsigned int> get_keypoints_in_cell(const float ref_x, const float ref_y, const float margin) const {
        return data::get_keypoints_in_cell(camera_, undist_keypts_, keypt_indices_in_cells_, ref_x, ref_y, margin);
    }
    bool will_be_erased() { return will_be_erased_; }

    unsigned int id_ = 0;
    camera::base* camera_ = nullptr;
    unsigned int num_keypts_ = 0;
    std::vector<cv::KeyPoint> keypts_;
    std::vector<cv::KeyPoint> undist_keypts_;
    eigen_alloc_vector<Vec3_t> bearings_;
    std::vector<std::vector<std::vector<unsigned int>>> keypt_indices_in_cells_;
    std::vector<float> stereo_x_right_;
    std::vector<float> depths_;
    cv::Mat descriptors_;
    fbow::BoWVector bow_vec_;
    fbow::BoWFeatVector bow_feat_vec_;
    unsigned int num_scale_levels_ = 0;
    float scale_factor_ = 0, log_scale_factor_ = 0;
    std::vector<float> scale_factors_, level_sigma_sq_, inv_level_sigma_sq_;
    std::vector<landmark*> landmarks_;
    bool will_be_erased_ = false;

private:
    Mat44_t cam_pose_cw_;
    Vec3_t cam_center_;
};

}  // namespace data
}  // namespace openvslam
