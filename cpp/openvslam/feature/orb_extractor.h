// Drop-in replacement for openvslam/feature/orb_extractor.h (reference: external/openvslam/src/openvslam/feature/
// orb_extractor.h:13-151). Same namespace, class name, public members and signatures; put this directory BEFORE the
// reference's src/ on the include path (and compile orb_extractor_cuda.cc instead of orb_extractor.cc +
// orb_extractor_node.cc) and tracking_module.cc:78-94, data/frame.cc:37-92 and every other user build unchanged.
// The body forwards to the C ABI of liblpslam_cuda (include/lpslam_cuda.h); nothing is computed on the host.
#ifndef OPENVSLAM_FEATURE_ORB_EXTRACTOR_H
#define OPENVSLAM_FEATURE_ORB_EXTRACTOR_H

#include "openvslam/feature/orb_params.h"

#include <opencv2/core/mat.hpp>
#include <opencv2/core/types.hpp>

#include <vector>

struct lpslam_orb_ctx;

namespace openvslam {
namespace feature {

class orb_extractor {
public:
    orb_extractor() = delete;

    //! Constructor (orb_extractor.cc:46-61)
    orb_extractor(const unsigned int max_num_keypts, const unsigned int ini_min_num_keypts,
                  const float scale_factor, const unsigned int num_levels,
                  const unsigned int ini_fast_thr, const unsigned int min_fast_thr,
                  const std::vector<std::vector<float>>& mask_rects = {});

    //! Constructor (orb_extractor.cc:63-69)
    explicit orb_extractor(const orb_params& orb_params);

    //! Destructor: releases the CUDA context (the reference's is defaulted; ownership by raw pointer is unchanged)
    virtual ~orb_extractor();

    //! Copyable like the reference's (implicitly defaulted) class: a copy gets its own device context with the same parameters
    //! and shares the image_pyramid_ headers of the source, as copied cv::Mat members do
    orb_extractor(const orb_extractor& other);
    orb_extractor& operator=(const orb_extractor& other);

    //! Extract keypoints and each descriptor of them (orb_extractor.cc:71-147)
    void extract(const cv::_InputArray& in_image, const cv::_InputArray& in_image_mask,
                 std::vector<cv::KeyPoint>& keypts, const cv::_OutputArray& out_descriptors);

    unsigned int get_max_num_keypoints() const;
    void set_max_num_keypoints(const unsigned int max_num_keypts);
    float get_scale_factor() const;
    void set_scale_factor(const float scale_factor);
    unsigned int get_num_scale_levels() const;
    void set_num_scale_levels(const unsigned int num_levels);
    unsigned int get_initial_fast_threshold() const;
    void set_initial_fast_threshold(const unsigned int initial_fast_threshold);
    unsigned int get_minimum_fast_threshold() const;
    void set_minimum_fast_threshold(const unsigned int minimum_fast_threshold);

    std::vector<float> get_scale_factors() const;
    std::vector<float> get_inv_scale_factors() const;
    std::vector<float> get_level_sigma_sq() const;
    std::vector<float> get_inv_level_sigma_sq() const;

    //! Image pyramid (orb_extractor.h:76): host copies of the device pyramid of the last extract; level 0 aliases the
    //! caller's image as in the reference (orb_extractor.cc:267). Levels 1.. arrive in ONE device-to-host copy into a single
    //! allocation that the cv::Mat headers share; as in the reference, the next extract of the same image size overwrites them
    //! in place. Filled by extract() because match::stereo and other users read it (data/frame.cc:78); set
    //! LPSLAM_CUDA_LAZY_PYRAMID=1 to skip the copy when nobody does.
    std::vector<cv::Mat> image_pyramid_;

    //! CUDA device the instance runs on (default 0, or the LPSLAM_CUDA_DEVICE environment variable)
    int cuda_device() const { return device_; }

private:
    void create_context();
    std::vector<float> table(int which) const;

    orb_params orb_params_;
    cv::Mat pyramid_storage_;  //!< levels 1.. of image_pyramid_, packed
    lpslam_orb_ctx* ctx_ = nullptr;
    int device_ = 0;
};

} // namespace feature
} // namespace openvslam

#endif // OPENVSLAM_FEATURE_ORB_EXTRACTOR_H
