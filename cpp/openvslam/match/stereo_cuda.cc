// Drop-in body of openvslam::match::stereo (reference: match/stereo.h:18-29, match/stereo.cc:18-249). The reference's
// stereo.h is used UNCHANGED; compile this file instead of match/stereo.cc. compute() flattens the members the reference
// constructor stored and calls lpslam_match_stereo; the private helpers of the reference class are never called.
#include "openvslam/match/stereo.h"
#include "openvslam/feature/orb_device_pyramid.h"

#include "lpslam_cuda.h"

#include <cstdlib>

#include <stdexcept>
#include <string>

namespace openvslam {
namespace match {

stereo::stereo(const std::vector<cv::Mat>& left_image_pyramid, const std::vector<cv::Mat>& right_image_pyramid,
               const std::vector<cv::KeyPoint>& keypts_left, const std::vector<cv::KeyPoint>& keypts_right,
               const cv::Mat& descs_left, const cv::Mat& descs_right,
               const std::vector<float>& scale_factors, const std::vector<float>& inv_scale_factors,
               const float focal_x_baseline, const float true_baseline)
    : left_image_pyramid_(left_image_pyramid), right_image_pyramid_(right_image_pyramid),
      num_keypts_(keypts_left.size()), keypts_left_(keypts_left), keypts_right_(keypts_right),
      descs_left_(descs_left), descs_right_(descs_right),
      scale_factors_(scale_factors), inv_scale_factors_(inv_scale_factors),
      focal_x_baseline_(focal_x_baseline), true_baseline_(true_baseline),
      min_disp_(0.0f), max_disp_(focal_x_baseline_ / true_baseline_) {}

void stereo::compute(std::vector<float>& stereo_x_right, std::vector<float>& depths) const {
    const uint32_t levels = static_cast<uint32_t>(left_image_pyramid_.size());
    std::vector<const uint8_t*> pl(levels), pr(levels);
    std::vector<int> lw(levels), lh(levels);
    std::vector<size_t> sl(levels), sr(levels);
    for (uint32_t l = 0; l < levels; ++l) {
        pl[l] = left_image_pyramid_[l].data;
        pr[l] = right_image_pyramid_[l].data;
        lw[l] = left_image_pyramid_[l].cols;
        lh[l] = left_image_pyramid_[l].rows;
        sl[l] = left_image_pyramid_[l].step;
        sr[l] = right_image_pyramid_[l].step;
    }
    stereo_x_right.assign(num_keypts_, -1.0f);
    depths.assign(num_keypts_, -1.0f);
    static_assert(sizeof(cv::KeyPoint) == sizeof(lpslam_keypoint), "cv::KeyPoint layout");
    uint32_t n_matches = 0;
    // both pyramids still on the device (they are the image_pyramid_ of two drop-in extractors' last extract): match there
    const char* force_upload = std::getenv("LPSLAM_CUDA_STEREO_UPLOAD");
    if (levels >= 2 && !(force_upload && force_upload[0] == '1')) {
        lpslam_orb_ctx* cl = feature::cuda_detail::resident_pyramid(pl[0], pl[1]);
        lpslam_orb_ctx* cr = feature::cuda_detail::resident_pyramid(pr[0], pr[1]);
        if (cl && cr && cl != cr) {
            const int rc = lpslam_match_stereo_ctx(cl, cr, reinterpret_cast<const lpslam_keypoint*>(keypts_left_.data()),
                                                   static_cast<uint32_t>(keypts_left_.size()),
                                                   reinterpret_cast<const lpslam_keypoint*>(keypts_right_.data()),
                                                   static_cast<uint32_t>(keypts_right_.size()), descs_left_.data, descs_right_.data,
                                                   focal_x_baseline_, true_baseline_, stereo_x_right.data(), depths.data(), &n_matches);
            if (rc < 0) throw std::runtime_error(std::string("liblpslam_cuda: lpslam_match_stereo_ctx: ") + lpslam_cuda_last_error());
            return;
        }
    }
    // descriptor matrices are n x 32 CV_8U and continuous (created by orb_extractor::extract)
    const int rc = lpslam_match_stereo(pl.data(), pr.data(), lw.data(), lh.data(), sl.data(), sr.data(), levels,
                                       reinterpret_cast<const lpslam_keypoint*>(keypts_left_.data()), static_cast<uint32_t>(keypts_left_.size()),
                                       reinterpret_cast<const lpslam_keypoint*>(keypts_right_.data()), static_cast<uint32_t>(keypts_right_.size()),
                                       descs_left_.data, descs_right_.data, scale_factors_.data(), inv_scale_factors_.data(),
                                       focal_x_baseline_, true_baseline_, stereo_x_right.data(), depths.data(), &n_matches);
    if (rc < 0) throw std::runtime_error(std::string("liblpslam_cuda: lpslam_match_stereo: ") + lpslam_cuda_last_error());
}

} // namespace match
} // namespace openvslam
