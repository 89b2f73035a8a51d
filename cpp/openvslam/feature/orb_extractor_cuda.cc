// Host side of the drop-in openvslam::feature::orb_extractor: flattens the cv:: arguments and calls liblpslam_cuda.
// Replaces orb_extractor.cc + orb_extractor_node.cc of the reference build. No pixel is touched on the host.
#include "openvslam/feature/orb_extractor.h"
#include "openvslam/feature/orb_device_pyramid.h"

#include "lpslam_cuda.h"

#include <mutex>

#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

namespace openvslam {
namespace feature {

namespace {
void check(int rc, const char* what) {
    if (rc < 0) throw std::runtime_error(std::string("liblpslam_cuda: ") + what + ": " + lpslam_cuda_last_error());
}

struct resident {
    lpslam_orb_ctx* ctx;
    const unsigned char *level0, *level1;
};
std::mutex resident_mutex;
std::vector<resident> residents;  // one entry per live extractor that has extracted: a handful
} // namespace

namespace cuda_detail {
void publish_pyramid(lpslam_orb_ctx* ctx, const unsigned char* level0, const unsigned char* level1) {
    std::lock_guard<std::mutex> lock(resident_mutex);
    for (auto& r : residents)
        if (r.ctx == ctx) {
            r.level0 = level0;
            r.level1 = level1;
            return;
        }
    residents.push_back({ctx, level0, level1});
}
void retire_pyramid(lpslam_orb_ctx* ctx) {
    std::lock_guard<std::mutex> lock(resident_mutex);
    for (size_t i = 0; i < residents.size(); ++i)
        if (residents[i].ctx == ctx) {
            residents.erase(residents.begin() + static_cast<long>(i));
            return;
        }
}
lpslam_orb_ctx* resident_pyramid(const unsigned char* level0, const unsigned char* level1) {
    std::lock_guard<std::mutex> lock(resident_mutex);
    for (const auto& r : residents)
        if (r.level0 == level0 && r.level1 == level1) return r.ctx;
    return nullptr;
}
} // namespace cuda_detail

orb_extractor::orb_extractor(const unsigned int max_num_keypts, const unsigned int ini_min_num_keypts,
                             const float scale_factor, const unsigned int num_levels,
                             const unsigned int ini_fast_thr, const unsigned int min_fast_thr,
                             const std::vector<std::vector<float>>& mask_rects)
    : orb_extractor(orb_params(max_num_keypts, ini_min_num_keypts, scale_factor, num_levels, ini_fast_thr, min_fast_thr, mask_rects)) {}

orb_extractor::orb_extractor(const orb_params& orb_params)
    : orb_params_(orb_params) {
    if (const char* d = std::getenv("LPSLAM_CUDA_DEVICE")) device_ = std::atoi(d);
    create_context();
}

orb_extractor::~orb_extractor() {
    cuda_detail::retire_pyramid(ctx_);
    lpslam_orb_destroy(ctx_);
}

orb_extractor::orb_extractor(const orb_extractor& other)
    : image_pyramid_(other.image_pyramid_), orb_params_(other.orb_params_), device_(other.device_) {
    create_context();
}

orb_extractor& orb_extractor::operator=(const orb_extractor& other) {
    if (this != &other) {
        cuda_detail::retire_pyramid(ctx_);
        lpslam_orb_destroy(ctx_);
        ctx_ = nullptr;
        image_pyramid_ = other.image_pyramid_;
        orb_params_ = other.orb_params_;
        device_ = other.device_;
        create_context();
    }
    return *this;
}

void orb_extractor::create_context() {
    std::vector<float> rects;
    for (const auto& r : orb_params_.mask_rects_) rects.insert(rects.end(), r.begin(), r.end());  // validated by orb_params
    lpslam_orb_params p;
    p.max_num_keypts = orb_params_.max_num_keypts_;
    p.ini_max_num_keypts = orb_params_.ini_max_num_keypts_;
    p.scale_factor = orb_params_.scale_factor_;
    p.num_levels = orb_params_.num_levels_;
    p.ini_fast_thr = orb_params_.ini_fast_thr_;
    p.min_fast_thr = orb_params_.min_fast_thr;
    p.mask_rects = rects.empty() ? nullptr : rects.data();
    p.num_mask_rects = static_cast<uint32_t>(orb_params_.mask_rects_.size());
    check(lpslam_orb_create(&ctx_, &p, device_, 1), "lpslam_orb_create");
}

void orb_extractor::extract(const cv::_InputArray& in_image, const cv::_InputArray& in_image_mask,
                            std::vector<cv::KeyPoint>& keypts, const cv::_OutputArray& out_descriptors) {
    if (in_image.empty()) {
        return;  // orb_extractor.cc:73-75: outputs untouched
    }
    const cv::Mat image = in_image.getMat();
    assert(image.type() == CV_8UC1);
    cuda_detail::retire_pyramid(ctx_);  // the device pyramid is about to be replaced
    cv::Mat mask;
    if (!in_image_mask.empty()) {
        mask = in_image_mask.getMat();  // orb_extractor.cc:93-97: the image mask has priority over the rectangle mask
        assert(mask.type() == CV_8UC1);
    }
    static_assert(sizeof(cv::KeyPoint) == sizeof(lpslam_keypoint), "cv::KeyPoint layout");

    // the keypoint count is not bounded by max_num_keypts (a breadth pass of the quadtree is always finished)
    uint32_t cap = 4 * orb_params_.max_num_keypts_ + 1024, n = 0;
    // landing buffers of the calling thread, kept between calls: a fresh std::vector of `cap` keypoints and cap x 32 descriptor
    // bytes would be value-initialised (0.5 MB of stores) on every frame only to be overwritten by the copy from the device
    thread_local std::vector<lpslam_keypoint> kp_land;
    thread_local std::vector<uint8_t> desc_land;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (kp_land.size() < cap) kp_land.resize(cap);
        if (desc_land.size() < static_cast<size_t>(cap) * 32) desc_land.resize(static_cast<size_t>(cap) * 32);
        const int rc = lpslam_orb_extract(ctx_, image.data, image.cols, image.rows, image.step,
                                          mask.empty() ? nullptr : mask.data, mask.empty() ? 0 : mask.step,
                                          kp_land.data(), desc_land.data(), cap, &n);
        if (rc == LPSLAM_ERR_CAPACITY && n > cap && attempt == 0) {
            cap = n;
            continue;
        }
        check(rc, "lpslam_orb_extract");
        break;
    }
    keypts.resize(n);
    if (n) std::memcpy(static_cast<void*>(keypts.data()), kp_land.data(), sizeof(lpslam_keypoint) * static_cast<size_t>(n));
    if (n == 0) {
        out_descriptors.release();  // orb_extractor.cc:115-117
    }
    else {
        out_descriptors.create(static_cast<int>(n), 32, CV_8U);
        cv::Mat d = out_descriptors.getMat();
        if (d.isContinuous()) {
            std::memcpy(d.data, desc_land.data(), static_cast<size_t>(n) * 32);
        }
        else {
            for (uint32_t i = 0; i < n; ++i) std::memcpy(d.ptr(static_cast<int>(i)), desc_land.data() + static_cast<size_t>(i) * 32, 32);
        }
    }

    // image_pyramid_ (orb_extractor.h:76): level 0 aliases the input, the other levels are copied back from the device
    image_pyramid_.resize(orb_params_.num_levels_);
    image_pyramid_.at(0) = image;
    const char* lazy = std::getenv("LPSLAM_CUDA_LAZY_PYRAMID");
    if (!(lazy && lazy[0] == '1') && orb_params_.num_levels_ > 1) {
        const unsigned int n = orb_params_.num_levels_ - 1;
        std::vector<size_t> offsets(n);
        size_t total = 0;
        lpslam_orb_pyramid_levels(ctx_, 0, 1, nullptr, 0, offsets.data(), &total);  // sizes only
        // like the reference, which resizes into the existing image_pyramid_ matrices (orb_extractor.cc:270-273), the levels of the
        // previous extract are overwritten in place while the image size stays the same
        if (pyramid_storage_.cols != static_cast<int>(total)) pyramid_storage_.create(1, static_cast<int>(total), CV_8UC1);
        cv::Mat& storage = pyramid_storage_;
        check(lpslam_orb_pyramid_levels(ctx_, 0, 1, storage.data, total, offsets.data(), &total), "lpslam_orb_pyramid_levels");
        for (unsigned int level = 1; level <= n; ++level) {
            int w = 0, h = 0;
            check(lpslam_orb_level_size(ctx_, level, &w, &h), "lpslam_orb_level_size");
            const int begin = static_cast<int>(offsets[level - 1]);
            image_pyramid_.at(level) = storage.colRange(begin, begin + w * h).reshape(1, h);  // shares storage's reference count
        }
        cuda_detail::publish_pyramid(ctx_, image.data, image_pyramid_.at(1).data);
    }
}

unsigned int orb_extractor::get_max_num_keypoints() const { return orb_params_.max_num_keypts_; }
void orb_extractor::set_max_num_keypoints(const unsigned int max_num_keypts) {
    orb_params_.max_num_keypts_ = max_num_keypts;
    check(lpslam_orb_set_param(ctx_, LPSLAM_ORB_MAX_NUM_KEYPTS, max_num_keypts), "set_max_num_keypoints");
}
float orb_extractor::get_scale_factor() const { return orb_params_.scale_factor_; }
void orb_extractor::set_scale_factor(const float scale_factor) {
    orb_params_.scale_factor_ = scale_factor;
    check(lpslam_orb_set_param(ctx_, LPSLAM_ORB_SCALE_FACTOR, scale_factor), "set_scale_factor");
}
unsigned int orb_extractor::get_num_scale_levels() const { return orb_params_.num_levels_; }
void orb_extractor::set_num_scale_levels(const unsigned int num_levels) {
    orb_params_.num_levels_ = num_levels;
    check(lpslam_orb_set_param(ctx_, LPSLAM_ORB_NUM_LEVELS, num_levels), "set_num_scale_levels");
}
unsigned int orb_extractor::get_initial_fast_threshold() const { return orb_params_.ini_fast_thr_; }
void orb_extractor::set_initial_fast_threshold(const unsigned int initial_fast_threshold) {
    orb_params_.ini_fast_thr_ = initial_fast_threshold;
    check(lpslam_orb_set_param(ctx_, LPSLAM_ORB_INI_FAST_THR, initial_fast_threshold), "set_initial_fast_threshold");
}
unsigned int orb_extractor::get_minimum_fast_threshold() const { return orb_params_.min_fast_thr; }
void orb_extractor::set_minimum_fast_threshold(const unsigned int minimum_fast_threshold) {
    orb_params_.min_fast_thr = minimum_fast_threshold;
    check(lpslam_orb_set_param(ctx_, LPSLAM_ORB_MIN_FAST_THR, minimum_fast_threshold), "set_minimum_fast_threshold");
}

std::vector<float> orb_extractor::table(int which) const {
    std::vector<float> t(orb_params_.num_levels_);
    check(lpslam_orb_get_table(ctx_, which, t.data(), static_cast<uint32_t>(t.size())), "lpslam_orb_get_table");
    return t;
}
std::vector<float> orb_extractor::get_scale_factors() const { return table(0); }
std::vector<float> orb_extractor::get_inv_scale_factors() const { return table(1); }
std::vector<float> orb_extractor::get_level_sigma_sq() const { return table(2); }
std::vector<float> orb_extractor::get_inv_level_sigma_sq() const { return table(3); }

} // namespace feature
} // namespace openvslam
