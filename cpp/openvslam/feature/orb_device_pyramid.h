// Internal to the CUDA drop-in layer (not part of the reference's interface): which extractor context still holds, on the
// device, the pyramid that a std::vector<cv::Mat> handed to match::stereo is the host copy of. orb_extractor::extract publishes
// (level 0 data pointer, level 1 data pointer) -> context after it filled image_pyramid_; match::stereo::compute looks both eyes
// up and, when both pyramids are still resident, matches on the device copies instead of sending 2 x 3.3 MB back
// (lpslam_match_stereo_ctx). Identity of the cv::Mat buffers stands for identity of their contents: a caller that edits
// image_pyramid_ in place between extract and match::stereo must set LPSLAM_CUDA_STEREO_UPLOAD=1 to get the host copies used.
#ifndef OPENVSLAM_FEATURE_ORB_DEVICE_PYRAMID_H
#define OPENVSLAM_FEATURE_ORB_DEVICE_PYRAMID_H

struct lpslam_orb_ctx;

namespace openvslam {
namespace feature {
namespace cuda_detail {

void publish_pyramid(lpslam_orb_ctx* ctx, const unsigned char* level0, const unsigned char* level1);
void retire_pyramid(lpslam_orb_ctx* ctx);
lpslam_orb_ctx* resident_pyramid(const unsigned char* level0, const unsigned char* level1);

} // namespace cuda_detail
} // namespace feature
} // namespace openvslam

#endif
