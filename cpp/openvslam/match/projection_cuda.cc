// Drop-in body of openvslam::match::projection::match_current_and_last_frames (reference: match/projection.h:21-30,
// match/projection.cc:85-208) for the perspective camera model (camera/perspective.cc:151-169). The reference's headers are used
// UNCHANGED. Build: compile this file and compile the reference's match/projection.cc with -DLPSLAM_CUDA_PROJECTION, with
// `#ifndef LPSLAM_CUDA_PROJECTION` / `#endif` put around its definition of match_current_and_last_frames.
// A frame whose camera is not a perspective model cannot take this path: the function throws (fisheye / equirectangular /
// radial-division stay on the reference's CPU body; do not define LPSLAM_CUDA_PROJECTION for those set-ups).
#include "openvslam/camera/perspective.h"
#include "openvslam/data/frame.h"
#include "openvslam/data/landmark.h"
#include "openvslam/match/projection.h"

#include "lpslam_cuda.h"

#include <cstring>
#include <stdexcept>
#include <string>

namespace openvslam {
namespace match {

unsigned int projection::match_current_and_last_frames(data::frame& curr_frm, const data::frame& last_frm, const float margin) const {
    if (curr_frm.camera_->model_type_ != camera::model_type_t::Perspective)
        throw std::runtime_error("liblpslam_cuda: match_current_and_last_frames covers the perspective camera model only");
    const auto* cam = static_cast<const camera::perspective*>(curr_frm.camera_);
    lpslam_camera c;
    c.fx = cam->fx_, c.fy = cam->fy_, c.cx = cam->cx_, c.cy = cam->cy_;
    c.focal_x_baseline = cam->focal_x_baseline_, c.true_baseline = cam->true_baseline_;
    c.min_x = cam->img_bounds_.min_x_, c.max_x = cam->img_bounds_.max_x_, c.min_y = cam->img_bounds_.min_y_, c.max_y = cam->img_bounds_.max_y_;
    c.grid_cols = cam->num_grid_cols_, c.grid_rows = cam->num_grid_rows_;
    c.inv_cell_width = cam->inv_cell_width_, c.inv_cell_height = cam->inv_cell_height_;
    c.is_monocular = cam->setup_type_ == camera::setup_type_t::Monocular;
    c.num_scale_levels = last_frm.num_scale_levels_;  // projection.cc:147: the forward-motion window ends at last_frm's top level
    const unsigned int nl = last_frm.num_keypts_, nc = curr_frm.num_keypts_;
    std::vector<uint8_t> has(nl), outl(nl), desc(32 * static_cast<size_t>(nl)), obs(nl), taken(nc);
    std::vector<double> pos(3 * static_cast<size_t>(nl));
    std::vector<int32_t> oct(nl);
    std::vector<float> ang(nl);
    for (unsigned int i = 0; i < nl; ++i) {  // projection.cc:112-127
        auto* lm = last_frm.landmarks_.at(i);
        has[i] = lm != nullptr;
        outl[i] = last_frm.outlier_flags_.at(i);
        oct[i] = last_frm.keypts_.at(i).octave;          // :137
        ang[i] = last_frm.undist_keypts_.at(i).angle;    // :194
        if (!lm) continue;
        const Vec3_t p = lm->get_pos_in_world();
        pos[3 * i] = p(0), pos[3 * i + 1] = p(1), pos[3 * i + 2] = p(2);
        const cv::Mat d = lm->get_descriptor();
        if (!d.empty()) std::memcpy(&desc[32 * static_cast<size_t>(i)], d.data, 32);
        obs[i] = lm->has_observation();
    }
    for (unsigned int i = 0; i < nc; ++i)  // projection.cc:163-165
        taken[i] = curr_frm.landmarks_.at(i) && curr_frm.landmarks_.at(i)->has_observation();
    double Tc[16], Tl[16];  // row-major copies of the column-major Eigen poses
    for (int r = 0; r < 4; ++r)
        for (int k = 0; k < 4; ++k) {
            Tc[4 * r + k] = curr_frm.cam_pose_cw_(r, k);
            Tl[4 * r + k] = last_frm.cam_pose_cw_(r, k);
        }
    const cv::Mat cd = curr_frm.descriptors_.step == 32 ? curr_frm.descriptors_ : curr_frm.descriptors_.clone();
    static_assert(sizeof(cv::KeyPoint) == sizeof(lpslam_keypoint), "cv::KeyPoint layout");
    std::vector<int32_t> m(nc, -1);
    uint32_t n = 0;
    if (lpslam_match_projection(&c, Tc, Tl, nl, has.data(), outl.data(), pos.data(), desc.data(), oct.data(), ang.data(), obs.data(), nc,
                                reinterpret_cast<const lpslam_keypoint*>(curr_frm.undist_keypts_.data()), cd.data,
                                curr_frm.stereo_x_right_.data(), taken.data(), curr_frm.scale_factors_.data(), margin, check_orientation_,
                                m.data(), &n) < 0)
        throw std::runtime_error(std::string("liblpslam_cuda: lpslam_match_projection: ") + lpslam_cuda_last_error());
    // projection.cc:189,199-202: matched keypoints take the landmark. (The tracker clears curr_frm.landmarks_ before this call,
    // module/frame_tracker.cc:36-38, so a keypoint whose match the orientation check removed is empty either way.)
    for (unsigned int i = 0; i < nc; ++i)
        if (m[i] >= 0) curr_frm.landmarks_.at(i) = last_frm.landmarks_.at(m[i]);
    return n;
}

} // namespace match
} // namespace openvslam
