// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// C glue around the reference's OWN matcher sources, compiled UNMODIFIED from /root/reference against oracle/shim into
// oracle/_ref/libmatch_ref.so (oracle/Makefile):
//   match/robust.cc, match/projection.cc, match/area.cc, match/fuse.cc, match/bow_tree.cc, match/angle_checker.h, match/base.h,
//   data/common.cc, data/landmark.cc (with the real data/landmark.h), camera/base.cc, camera/perspective.cc, and the real
//   openvslam/type.h over a minimal Eigen.
// data::frame and data::keyframe are stand-ins with the reference's member names (oracle/shim/openvslam/data): the real ones
// need g2o, the BoW vocabulary and the extractor. Every entry point below takes the same flat arrays as the port's orc_match_*
// (orb_oracle.h), builds frame / keyframe / landmark objects from them and calls the reference's member function, so the port
// and the CUDA library can be held against the reference line for line.
//
// Where the reference computes geometry in front of its search loop (reprojection, valid-distance gate, predicted scale level)
// and the flat interface carries those results instead, the objects are arranged so that the reference code arrives at
// exactly the given values: a camera::base subclass answers reproject_to_image from a table indexed by the landmark's
// world x coordinate, and a landmark's max_valid_dist_ is set half a level inside the wanted level, so that the
// reference's own predict_scale_level (landmark.cc:239-276) returns it. Poses are identities.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include <opencv2/core.hpp>
#include <nlohmann/json.hpp>

#define private public
#include "openvslam/data/landmark.h"
#undef private
#include "openvslam/camera/perspective.h"
#include "openvslam/data/common.h"
#include "openvslam/data/frame.h"
#include "openvslam/data/keyframe.h"
#include "openvslam/data/map_database.h"
#include "openvslam/match/angle_checker.h"
#include "openvslam/match/base.h"
#ifndef DROPIN_SUBSET
#include "openvslam/match/area.h"
#include "openvslam/match/bow_tree.h"
#include "openvslam/match/fuse.h"
#endif
#include "openvslam/match/projection.h"
#include "openvslam/match/robust.h"

#include "orb_oracle.h"

using namespace openvslam;

namespace {

// answers from tables; everything else of camera::base is never called by the matchers
class table_camera final : public camera::base {
public:
    explicit table_camera(const orc_camera* c)
        : camera::base("table", c->is_monocular ? camera::setup_type_t::Monocular : camera::setup_type_t::Stereo,
                       camera::model_type_t::Perspective, camera::color_order_t::Gray, static_cast<unsigned>(c->max_x),
                       static_cast<unsigned>(c->max_y), 30.0, c->focal_x_baseline, c->true_baseline, c->grid_cols, c->grid_rows) {
        img_bounds_ = camera::image_bounds{c->min_x, c->max_x, c->min_y, c->max_y};
        inv_cell_width_ = c->inv_cell_width;
        inv_cell_height_ = c->inv_cell_height;
    }
    std::vector<double> reproj;   // 2 per row
    std::vector<float> x_right;   // 1 per row
    std::vector<uint8_t> in_image;
    Vec3_t bearing_answer = Vec3_t::Zero();

    void show_parameters() const override {}
    camera::image_bounds compute_image_bounds() const override { return img_bounds_; }
    cv::KeyPoint undistort_keypoint(const cv::KeyPoint& k) const override { return k; }
    void undistort_keypoints(const std::vector<cv::KeyPoint>& a, std::vector<cv::KeyPoint>& b) const override { b = a; }
    Vec3_t convert_keypoint_to_bearing(const cv::KeyPoint&) const override { return Vec3_t::Zero(); }
    void convert_keypoints_to_bearings(const std::vector<cv::KeyPoint>&, eigen_alloc_vector<Vec3_t>&) const override {}
    cv::KeyPoint convert_bearing_to_keypoint(const Vec3_t&) const override { return cv::KeyPoint(); }
    void convert_bearings_to_keypoints(const eigen_alloc_vector<Vec3_t>&, std::vector<cv::KeyPoint>&) const override {}
    bool reproject_to_image(const Mat33_t&, const Vec3_t&, const Vec3_t& pos_w, Vec2_t& rp, float& xr) const override {
        const size_t i = static_cast<size_t>(pos_w(0));
        rp(0) = reproj.at(2 * i);
        rp(1) = reproj.at(2 * i + 1);
        xr = x_right.empty() ? 0.0f : x_right.at(i);
        return in_image.empty() ? true : in_image.at(i) != 0;
    }
    bool reproject_to_bearing(const Mat33_t&, const Vec3_t&, const Vec3_t&, Vec3_t& out) const override {
        out = bearing_answer;
        return true;
    }
    nlohmann::json to_json() const override { return nlohmann::json(); }
};

std::unique_ptr<camera::perspective> real_camera(const orc_camera* c) {
    auto cam = std::make_unique<camera::perspective>(
        "ref", c->is_monocular ? camera::setup_type_t::Monocular : camera::setup_type_t::Stereo, camera::color_order_t::Gray,
        static_cast<unsigned>(c->max_x), static_cast<unsigned>(c->max_y), 30.0, c->fx, c->fy, c->cx, c->cy, 0.0, 0.0, 0.0, 0.0, 0.0,
        c->focal_x_baseline);
    // the flat interface carries these explicitly; a no-distortion camera derives the same values (perspective.cc:27-31)
    cam->img_bounds_ = camera::image_bounds{c->min_x, c->max_x, c->min_y, c->max_y};
    cam->inv_cell_width_ = c->inv_cell_width;
    cam->inv_cell_height_ = c->inv_cell_height;
    const_cast<double&>(cam->true_baseline_) = c->true_baseline;
    return cam;
}

std::vector<cv::KeyPoint> to_cv(const orc_keypoint* kp, uint32_t n) {
    std::vector<cv::KeyPoint> v(n);
    if (n) std::memcpy(static_cast<void*>(v.data()), kp, sizeof(orc_keypoint) * n);
    return v;
}
cv::Mat desc_mat(const uint8_t* d, uint32_t n) {
    return cv::Mat(static_cast<int>(n), 32, CV_8UC1, const_cast<uint8_t*>(d), 32).clone();
}
cv::Mat desc_row(const uint8_t* d) { return cv::Mat(1, 32, CV_8UC1, const_cast<uint8_t*>(d), 32).clone(); }

struct Scene {  // owns everything a call creates
    data::map_database map_db;
    data::keyframe anchor;  // reference keyframe of every landmark (landmark.cc:15 reads its id)
    std::vector<std::unique_ptr<data::landmark>> lms;
    data::landmark* add_lm(const Vec3_t& pos = Vec3_t::Zero()) {
        lms.emplace_back(new data::landmark(pos, &anchor, &map_db));
        return lms.back().get();
    }
};

void fill_scales(const float* sf, uint32_t n_levels, unsigned& num_levels, float& scale_factor, float& log_scale_factor,
                 std::vector<float>& scale_factors) {
    num_levels = n_levels;
    scale_factors.assign(sf, sf + n_levels);
    scale_factor = n_levels > 1 ? sf[1] / sf[0] : 1.2f;
    log_scale_factor = std::log(scale_factor);
}

void fill_frame(data::frame& f, camera::base* cam, uint32_t n, const orc_keypoint* kp, const uint8_t* desc, const float* x_right,
                const float* sf, uint32_t n_levels) {
    f.camera_ = cam;
    f.num_keypts_ = n;
    f.keypts_ = f.undist_keypts_ = to_cv(kp, n);
    if (desc) f.descriptors_ = desc_mat(desc, n);
    f.stereo_x_right_.assign(n, -1.0f);
    if (x_right) f.stereo_x_right_.assign(x_right, x_right + n);
    f.landmarks_.assign(n, nullptr);
    f.outlier_flags_.assign(n, false);
    f.cam_pose_cw_ = Mat44_t::Identity();
    if (sf) fill_scales(sf, n_levels, f.num_scale_levels_, f.scale_factor_, f.log_scale_factor_, f.scale_factors_);
    if (cam) data::assign_keypoints_to_grid(cam, f.undist_keypts_, f.keypt_indices_in_cells_);
}

void fill_keyframe(data::keyframe& k, camera::base* cam, uint32_t n, const orc_keypoint* kp, const uint8_t* desc, const float* x_right,
                   const float* sf, uint32_t n_levels) {
    k.camera_ = cam;
    k.num_keypts_ = n;
    k.keypts_ = k.undist_keypts_ = to_cv(kp, n);
    if (desc) k.descriptors_ = desc_mat(desc, n);
    k.stereo_x_right_.assign(n, -1.0f);
    if (x_right) k.stereo_x_right_.assign(x_right, x_right + n);
    k.landmarks_.assign(n, nullptr);
    k.set_cam_pose(Mat44_t::Identity());
    if (sf) fill_scales(sf, n_levels, k.num_scale_levels_, k.scale_factor_, k.log_scale_factor_, k.scale_factors_);
    if (cam) data::assign_keypoints_to_grid(cam, k.undist_keypts_, k.keypt_indices_in_cells_);
}

std::vector<orc_keypoint> plain_kp(uint32_t n, const float* angle) {
    std::vector<orc_keypoint> v(n);
    for (uint32_t i = 0; i < n; ++i) v[i] = orc_keypoint{0, 0, 31, angle ? angle[i] : -1.0f, 0, 0, -1};
    return v;
}

fbow::BoWFeatVector feat_vec(uint32_t n_nodes, const uint32_t* id, const uint32_t* start, const uint32_t* items) {
    fbow::BoWFeatVector v;
    for (uint32_t a = 0; a < n_nodes; ++a) v[id[a]] = std::vector<uint32_t>(items + start[a], items + start[a + 1]);
    return v;
}

// world position of table row i: x carries the row, z keeps the point far in front of the identity camera
Vec3_t row_pos(size_t i) { return Vec3_t{static_cast<double>(i), 0.0, 1000.0}; }

// makes predict_scale_level(|pos|, scale factor sf) return `level` and the valid-distance gate pass
void aim_level(data::landmark* lm, const Vec3_t& pos, int level, float scale_factor) {
    const float dist = static_cast<float>(pos.norm());
    lm->max_valid_dist_ = dist * std::pow(scale_factor, static_cast<float>(level) - 0.5f);
    lm->min_valid_dist_ = 0.0f;
    lm->mean_normal_ = pos.normalized();
}

template<typename T>
std::map<const void*, int> index_of(const std::vector<T*>& v) {
    std::map<const void*, int> m;
    for (size_t i = 0; i < v.size(); ++i)
        if (v[i]) m[v[i]] = static_cast<int>(i);
    return m;
}

}  // namespace

extern "C" {

uint32_t ref_hamming_32(const uint8_t* a, const uint8_t* b) { return match::compute_descriptor_distance_32(desc_row(a), desc_row(b)); }
uint32_t ref_hamming_64(const uint8_t* a, const uint8_t* b) { return match::compute_descriptor_distance_64(desc_row(a), desc_row(b)); }

// match/angle_checker.h: invalid ids in the order get_invalid_matches returns them
int ref_angle_invalid(const float* delta_angles, const int32_t* ids, int n, int32_t* invalid_out) {
    match::angle_checker<int> chk;
    for (int i = 0; i < n; ++i) chk.append_delta_angle(delta_angles[i], ids[i]);
    const auto inv = chk.get_invalid_matches();
    for (size_t i = 0; i < inv.size(); ++i) invalid_out[i] = inv[i];
    return static_cast<int>(inv.size());
}
// angle_checker<int>(30, num_bins_thr) as test/openvslam/match/angle_checker.cc constructs it; returns the invalid count,
// valid ids go to valid_out (count in *n_valid)
int ref_angle_check(const float* delta_angles, const int32_t* ids, int n, unsigned num_bins_thr, int32_t* invalid_out, int32_t* valid_out,
                    int* n_valid) {
    match::angle_checker<int> chk(30, num_bins_thr);
    for (int i = 0; i < n; ++i) chk.append_delta_angle(delta_angles[i], ids[i]);
    const auto inv = chk.get_invalid_matches();
    const auto val = chk.get_valid_matches();
    for (size_t i = 0; i < inv.size(); ++i) invalid_out[i] = inv[i];
    for (size_t i = 0; i < val.size(); ++i) valid_out[i] = val[i];
    *n_valid = static_cast<int>(val.size());
    return static_cast<int>(inv.size());
}
int ref_angle_valid(const float* delta_angles, const int32_t* ids, int n, int32_t* valid_out) {
    match::angle_checker<int> chk;
    for (int i = 0; i < n; ++i) chk.append_delta_angle(delta_angles[i], ids[i]);
    const auto v = chk.get_valid_matches();
    for (size_t i = 0; i < v.size(); ++i) valid_out[i] = v[i];
    return static_cast<int>(v.size());
}

// match/robust.cc:215-321
int ref_match_bruteforce(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint8_t* valid_2,
                         const float* angle_1, const float* angle_2, float lowe_ratio, int check_orientation, int32_t* match_2_in_1) {
    Scene sc;
    data::frame frm;
    data::keyframe kf;
    const auto k1 = plain_kp(n1, angle_1), k2 = plain_kp(n2, angle_2);
    fill_frame(frm, nullptr, n1, k1.data(), desc_1, nullptr, nullptr, 0);
    fill_keyframe(kf, nullptr, n2, k2.data(), desc_2, nullptr, nullptr, 0);
    for (uint32_t i = 0; i < n2; ++i)
        if (!valid_2 || valid_2[i]) kf.landmarks_[i] = sc.add_lm();
    match::robust m(lowe_ratio, check_orientation != 0);
    std::vector<std::pair<int, int>> matches;
    const unsigned num = m.brute_force_match(frm, &kf, matches);
    for (uint32_t i = 0; i < n1; ++i) match_2_in_1[i] = -1;
    for (const auto& p : matches) match_2_in_1[p.first] = p.second;
    return static_cast<int>(num);
}

// match/projection.cc:85-208 with the real perspective camera (camera/perspective.cc:151-169) and the real grid (data/common.cc)
int ref_match_projection(const orc_camera* cam, const double* pose_cw_curr, const double* pose_cw_last, uint32_t n_last,
                         const uint8_t* last_has_lm, const uint8_t* last_outlier, const double* last_pos_w, const uint8_t* last_lm_desc,
                         const int32_t* last_octave, const float* last_angle, uint32_t n_curr, const orc_keypoint* curr_undist_kp,
                         const uint8_t* curr_desc, const float* curr_x_right, const uint8_t* curr_has_observed_lm,
                         const float* scale_factors, float margin, int check_orientation, const uint8_t* last_lm_has_observation,
                         int32_t* curr_match_last) {
    Scene sc;
    auto camera = real_camera(cam);
    data::frame curr, last;
    fill_frame(curr, camera.get(), n_curr, curr_undist_kp, curr_desc, curr_x_right, scale_factors, cam->num_scale_levels);
    std::vector<orc_keypoint> lk(n_last);
    for (uint32_t i = 0; i < n_last; ++i) lk[i] = orc_keypoint{0, 0, 31, last_angle[i], 0, last_octave[i], -1};
    fill_frame(last, camera.get(), n_last, lk.data(), nullptr, nullptr, scale_factors, cam->num_scale_levels);
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) {
            curr.cam_pose_cw_(r, c) = pose_cw_curr[4 * r + c];
            last.cam_pose_cw_(r, c) = pose_cw_last[4 * r + c];
        }
    for (uint32_t i = 0; i < n_last; ++i) {
        last.outlier_flags_[i] = last_outlier[i] != 0;
        if (!last_has_lm[i]) continue;
        auto* lm = sc.add_lm(Vec3_t{last_pos_w[3 * i], last_pos_w[3 * i + 1], last_pos_w[3 * i + 2]});
        lm->descriptor_ = desc_row(last_lm_desc + 32 * static_cast<size_t>(i));
        lm->num_observations_ = (!last_lm_has_observation || last_lm_has_observation[i]) ? 1 : 0;
        last.landmarks_[i] = lm;
    }
    for (uint32_t i = 0; i < n_curr; ++i)
        if (curr_has_observed_lm[i]) {
            curr.landmarks_[i] = sc.add_lm();
            curr.landmarks_[i]->num_observations_ = 1;
        }
    match::projection m(0.6f, check_orientation != 0);
    const unsigned num = m.match_current_and_last_frames(curr, last, margin);
    const auto where = index_of(last.landmarks_);
    for (uint32_t i = 0; i < n_curr; ++i) {
        const auto it = where.find(curr.landmarks_[i]);
        curr_match_last[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

#ifndef DROPIN_SUBSET
// match/projection.cc:11-83
int ref_match_landmarks(const orc_camera* cam, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj, const float* lm_x_right,
                        const int32_t* lm_scale_level, const uint8_t* lm_desc, const uint8_t* lm_has_observation, uint32_t n_kp,
                        const orc_keypoint* kp_undist, const uint8_t* kp_desc, const float* kp_x_right,
                        const uint8_t* kp_has_observed_lm, const float* scale_factors, float margin, float lowe_ratio,
                        int32_t* kp_match_lm) {
    Scene sc;
    table_camera camera(cam);
    data::frame frm;
    fill_frame(frm, &camera, n_kp, kp_undist, kp_desc, kp_x_right, scale_factors, cam->num_scale_levels);
    for (uint32_t i = 0; i < n_kp; ++i)
        if (kp_has_observed_lm[i]) {
            frm.landmarks_[i] = sc.add_lm();
            frm.landmarks_[i]->num_observations_ = 1;
        }
    std::vector<data::landmark*> local(n_lm);
    for (uint32_t j = 0; j < n_lm; ++j) {
        auto* lm = sc.add_lm();
        lm->is_observable_in_tracking_ = lm_active[j] != 0;
        lm->reproj_in_tracking_ = Vec2_t{lm_reproj[2 * j], lm_reproj[2 * j + 1]};
        lm->x_right_in_tracking_ = lm_x_right[j];
        lm->scale_level_in_tracking_ = lm_scale_level[j];
        lm->descriptor_ = desc_row(lm_desc + 32 * static_cast<size_t>(j));
        lm->num_observations_ = (!lm_has_observation || lm_has_observation[j]) ? 1 : 0;
        local[j] = lm;
    }
    match::projection m(lowe_ratio, true);
    const unsigned num = m.match_frame_and_landmarks(frm, local, margin);
    const auto where = index_of(local);
    for (uint32_t i = 0; i < n_kp; ++i) {
        const auto it = where.find(frm.landmarks_[i]);
        kp_match_lm[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/projection.cc:210-312 match_frame_and_keyframe: rows = keyframe keypoints with their landmarks
int ref_match_frame_and_keyframe(const orc_camera* cam, float margin, uint32_t hamm_thr, uint32_t n_lm, const uint8_t* lm_active,
                                 const double* lm_reproj, const int32_t* lm_scale_level, const uint8_t* lm_desc, const float* lm_angle,
                                 uint32_t n_kp, const orc_keypoint* kp, const uint8_t* kp_desc, const uint8_t* kp_taken,
                                 const float* scale_factors, int32_t* kp_match_lm) {
    Scene sc;
    table_camera camera(cam);
    camera.reproj.assign(lm_reproj, lm_reproj + 2 * static_cast<size_t>(n_lm));
    data::frame frm;
    fill_frame(frm, &camera, n_kp, kp, kp_desc, nullptr, scale_factors, cam->num_scale_levels);
    for (uint32_t i = 0; i < n_kp; ++i)
        if (kp_taken[i]) frm.landmarks_[i] = sc.add_lm();
    data::keyframe kf;
    const auto kk = plain_kp(n_lm, lm_angle);
    fill_keyframe(kf, nullptr, n_lm, kk.data(), nullptr, nullptr, scale_factors, cam->num_scale_levels);
    for (uint32_t j = 0; j < n_lm; ++j) {
        if (!lm_active[j]) continue;  // no landmark on this keyframe keypoint (or erased / already matched / out of image / out of range)
        auto* lm = sc.add_lm(row_pos(j));
        lm->descriptor_ = desc_row(lm_desc + 32 * static_cast<size_t>(j));
        aim_level(lm, row_pos(j), lm_scale_level[j], frm.scale_factor_);
        kf.landmarks_[j] = lm;
    }
    match::projection m(0.6f, lm_angle != nullptr);
    const unsigned num = m.match_frame_and_keyframe(frm, &kf, std::set<data::landmark*>(), margin, hamm_thr);
    const auto where = index_of(kf.landmarks_);
    for (uint32_t i = 0; i < n_kp; ++i) {
        const auto it = where.find(frm.landmarks_[i]);
        kp_match_lm[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/projection.cc:317-413
int ref_match_sim3(const orc_camera* cam, float margin, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                   const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp, const orc_keypoint* kp, const uint8_t* kp_desc,
                   const uint8_t* kp_has_lm, const float* scale_factors, int32_t* kp_match_lm) {
    Scene sc;
    table_camera camera(cam);
    camera.reproj.assign(lm_reproj, lm_reproj + 2 * static_cast<size_t>(n_lm));
    data::keyframe kf;
    fill_keyframe(kf, &camera, n_kp, kp, kp_desc, nullptr, scale_factors, cam->num_scale_levels);
    std::vector<data::landmark*> matched(n_kp, nullptr), cand(n_lm);
    for (uint32_t i = 0; i < n_kp; ++i)
        if (kp_has_lm[i]) matched[i] = sc.add_lm();
    for (uint32_t j = 0; j < n_lm; ++j) {
        auto* lm = sc.add_lm(row_pos(j));
        lm->descriptor_ = desc_row(lm_desc + 32 * static_cast<size_t>(j));
        aim_level(lm, row_pos(j), lm_scale_level[j], kf.scale_factor_);
        lm->will_be_erased_ = !lm_active[j];
        cand[j] = lm;
    }
    match::projection m(0.6f, true);
    const unsigned num = m.match_by_Sim3_transform(&kf, Mat44_t::Identity(), cand, matched, margin);
    const auto where = index_of(cand);
    for (uint32_t i = 0; i < n_kp; ++i) {
        const auto it = where.find(matched[i]);
        kp_match_lm[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/projection.cc:415-636
int ref_match_mutual(const orc_camera* cam_1, const orc_camera* cam_2, float margin, uint32_t n_1, const uint8_t* lm1_active,
                     const double* lm1_reproj_in_2, const int32_t* lm1_level_in_2, const uint8_t* lm1_desc, const orc_keypoint* kp1,
                     const uint8_t* desc1, const float* scale_1, uint32_t n_2, const uint8_t* lm2_active, const double* lm2_reproj_in_1,
                     const int32_t* lm2_level_in_1, const uint8_t* lm2_desc, const orc_keypoint* kp2, const uint8_t* desc2,
                     const float* scale_2, int32_t* matched_2_in_1) {
    Scene sc;
    table_camera c1(cam_1), c2(cam_2);
    // the reference projects BOTH directions through keyfrm_2->camera_ (projection.cc:487,562): one table, rows of keyframe 1 first
    c2.reproj.assign(lm1_reproj_in_2, lm1_reproj_in_2 + 2 * static_cast<size_t>(n_1));
    c2.reproj.insert(c2.reproj.end(), lm2_reproj_in_1, lm2_reproj_in_1 + 2 * static_cast<size_t>(n_2));
    data::keyframe k1, k2;
    fill_keyframe(k1, &c1, n_1, kp1, desc1, nullptr, scale_1, cam_1->num_scale_levels);
    fill_keyframe(k2, &c2, n_2, kp2, desc2, nullptr, scale_2, cam_2->num_scale_levels);
    for (uint32_t j = 0; j < n_1; ++j) {
        if (!lm1_active[j]) continue;
        auto* lm = sc.add_lm(row_pos(j));
        lm->descriptor_ = desc_row(lm1_desc + 32 * static_cast<size_t>(j));
        aim_level(lm, row_pos(j), lm1_level_in_2[j], k2.scale_factor_);
        k1.landmarks_[j] = lm;
    }
    for (uint32_t j = 0; j < n_2; ++j) {
        if (!lm2_active[j]) continue;
        auto* lm = sc.add_lm(row_pos(n_1 + j));
        lm->descriptor_ = desc_row(lm2_desc + 32 * static_cast<size_t>(j));
        aim_level(lm, row_pos(n_1 + j), lm2_level_in_1[j], k1.scale_factor_);
        k2.landmarks_[j] = lm;
    }
    std::vector<data::landmark*> matched(n_1, nullptr);
    match::projection m(0.6f, true);
    const float s_12 = 1.0f;
    const unsigned num = m.match_keyframes_mutually(&k1, &k2, matched, s_12, Mat33_t::Identity(), Vec3_t::Zero(), margin);
    const auto where = index_of(k2.landmarks_);
    for (uint32_t i = 0; i < n_1; ++i) {
        const auto it = where.find(matched[i]);
        matched_2_in_1[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/fuse.cc:12-120 (chi_gate == 0: detect_duplication) and :122-266 (chi_gate != 0: replace_duplication). Every keyframe
// keypoint holds a live landmark with more observations than the rows, so that the keypoint a row selects can be read back
// (duplicated_lms_in_keyfrm / landmark::get_replaced) and rows stay independent. best_idx[j] = keypoint or -1.
int ref_match_fuse(const orc_camera* cam, float margin, int chi_gate, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                   const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp, const orc_keypoint* kp,
                   const uint8_t* kp_desc, const float* kp_x_right, const float* scale_factors, const float* inv_level_sigma_sq,
                   int32_t* best_idx) {
    Scene sc;
    table_camera camera(cam);
    camera.reproj.assign(lm_reproj, lm_reproj + 2 * static_cast<size_t>(n_lm));
    if (lm_x_right) camera.x_right.assign(lm_x_right, lm_x_right + n_lm);
    data::keyframe kf;
    fill_keyframe(kf, &camera, n_kp, kp, kp_desc, kp_x_right, scale_factors, cam->num_scale_levels);
    if (inv_level_sigma_sq) kf.inv_level_sigma_sq_.assign(inv_level_sigma_sq, inv_level_sigma_sq + cam->num_scale_levels);
    for (uint32_t i = 0; i < n_kp; ++i) {
        kf.landmarks_[i] = sc.add_lm();
        kf.landmarks_[i]->num_observations_ = 10;
    }
    std::vector<data::landmark*> cand(n_lm);
    for (uint32_t j = 0; j < n_lm; ++j) {
        auto* lm = sc.add_lm(row_pos(j));
        lm->descriptor_ = desc_row(lm_desc + 32 * static_cast<size_t>(j));
        aim_level(lm, row_pos(j), lm_scale_level[j], kf.scale_factor_);
        lm->will_be_erased_ = !lm_active[j];
        lm->num_observations_ = 1;
        cand[j] = lm;
    }
    const auto where = index_of(kf.landmarks_);
    match::fuse m(0.6f);
    unsigned num;
    if (!chi_gate) {
        std::vector<data::landmark*> dup;
        num = m.detect_duplication(&kf, Mat44_t::Identity(), cand, margin, dup);
        for (uint32_t j = 0; j < n_lm; ++j) {
            const auto it = where.find(dup[j]);
            best_idx[j] = it == where.end() ? -1 : it->second;
        }
    }
    else {
        num = m.replace_duplication(&kf, cand, margin);
        for (uint32_t j = 0; j < n_lm; ++j) {
            const auto it = where.find(cand[j]->get_replaced());
            best_idx[j] = it == where.end() ? -1 : it->second;
        }
    }
    return static_cast<int>(num);
}

// match/bow_tree.cc:16-126
int ref_match_bow(uint32_t n_nodes_kf, const uint32_t* kf_node_id, const uint32_t* kf_node_start, const uint32_t* kf_items,
                  uint32_t n_nodes_f, const uint32_t* f_node_id, const uint32_t* f_node_start, const uint32_t* f_items, uint32_t n_kf,
                  const uint8_t* kf_valid, const uint8_t* kf_desc, const float* kf_angle, uint32_t n_f, const uint8_t* f_desc,
                  const float* f_angle, float lowe_ratio, int check_orientation, int32_t* f_match_kf) {
    Scene sc;
    data::keyframe kf;
    data::frame frm;
    const auto kk = plain_kp(n_kf, kf_angle), fk = plain_kp(n_f, f_angle);
    fill_keyframe(kf, nullptr, n_kf, kk.data(), kf_desc, nullptr, nullptr, 0);
    fill_frame(frm, nullptr, n_f, fk.data(), f_desc, nullptr, nullptr, 0);
    kf.bow_feat_vec_ = feat_vec(n_nodes_kf, kf_node_id, kf_node_start, kf_items);
    frm.bow_feat_vec_ = feat_vec(n_nodes_f, f_node_id, f_node_start, f_items);
    for (uint32_t i = 0; i < n_kf; ++i)
        if (kf_valid[i]) kf.landmarks_[i] = sc.add_lm();
    match::bow_tree m(lowe_ratio, check_orientation != 0);
    std::vector<data::landmark*> matched;
    const unsigned num = m.match_frame_and_keyframe(&kf, frm, matched);
    const auto where = index_of(kf.landmarks_);
    for (uint32_t i = 0; i < n_f; ++i) {
        const auto it = where.find(matched[i]);
        f_match_kf[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/bow_tree.cc:128-250
int ref_match_bow_keyframes(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const uint8_t* valid_1, const uint8_t* desc_1, const float* angle_1, uint32_t n2,
                            const uint8_t* valid_2, const uint8_t* desc_2, const float* angle_2, float lowe_ratio, int check_orientation,
                            int32_t* match_2_of_1) {
    Scene sc;
    data::keyframe k1, k2;
    const auto kk1 = plain_kp(n1, angle_1), kk2 = plain_kp(n2, angle_2);
    fill_keyframe(k1, nullptr, n1, kk1.data(), desc_1, nullptr, nullptr, 0);
    fill_keyframe(k2, nullptr, n2, kk2.data(), desc_2, nullptr, nullptr, 0);
    k1.bow_feat_vec_ = feat_vec(n_nodes_1, node_id_1, node_start_1, items_1);
    k2.bow_feat_vec_ = feat_vec(n_nodes_2, node_id_2, node_start_2, items_2);
    for (uint32_t i = 0; i < n1; ++i)
        if (valid_1[i]) k1.landmarks_[i] = sc.add_lm();
    for (uint32_t i = 0; i < n2; ++i)
        if (valid_2[i]) k2.landmarks_[i] = sc.add_lm();
    match::bow_tree m(lowe_ratio, check_orientation != 0);
    std::vector<data::landmark*> matched;
    const unsigned num = m.match_keyframes(&k1, &k2, matched);
    const auto where = index_of(k2.landmarks_);
    for (uint32_t i = 0; i < n1; ++i) {
        const auto it = where.find(matched[i]);
        match_2_of_1[i] = it == where.end() ? -1 : it->second;
    }
    return static_cast<int>(num);
}

// match/robust.cc:18-178 (the epipole bearing of :27-30 is handed in, as in the flat interface)
int ref_match_triangulation(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc_1, const double* bearings_1, const float* x_right_1,
                            const uint8_t* has_lm_1, const float* scale_1, uint32_t n_levels, uint32_t n2, const orc_keypoint* kp2,
                            const uint8_t* desc_2, const double* bearings_2, const float* x_right_2, const uint8_t* has_lm_2,
                            const double* E_12, const double* epiplane_in_keyfrm_2, int check_orientation, int32_t* matched_2_in_1) {
    Scene sc;
    orc_camera oc{};
    oc.max_x = oc.max_y = 1;
    oc.grid_cols = oc.grid_rows = 1;
    table_camera camera(&oc);
    camera.bearing_answer = Vec3_t{epiplane_in_keyfrm_2[0], epiplane_in_keyfrm_2[1], epiplane_in_keyfrm_2[2]};
    data::keyframe k1, k2;
    fill_keyframe(k1, nullptr, n1, kp1, desc_1, x_right_1, scale_1, n_levels);
    fill_keyframe(k2, nullptr, n2, kp2, desc_2, x_right_2, scale_1, n_levels);
    k2.camera_ = &camera;
    k1.bearings_.resize(n1);
    k2.bearings_.resize(n2);
    for (uint32_t i = 0; i < n1; ++i) k1.bearings_[i] = Vec3_t{bearings_1[3 * i], bearings_1[3 * i + 1], bearings_1[3 * i + 2]};
    for (uint32_t i = 0; i < n2; ++i) k2.bearings_[i] = Vec3_t{bearings_2[3 * i], bearings_2[3 * i + 1], bearings_2[3 * i + 2]};
    k1.bow_feat_vec_ = feat_vec(n_nodes_1, node_id_1, node_start_1, items_1);
    k2.bow_feat_vec_ = feat_vec(n_nodes_2, node_id_2, node_start_2, items_2);
    for (uint32_t i = 0; i < n1; ++i)
        if (has_lm_1[i]) k1.landmarks_[i] = sc.add_lm();
    for (uint32_t i = 0; i < n2; ++i)
        if (has_lm_2[i]) k2.landmarks_[i] = sc.add_lm();
    Mat33_t E;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) E(r, c) = E_12[3 * r + c];
    match::robust m(0.6f, check_orientation != 0);
    std::vector<std::pair<unsigned, unsigned>> pairs;
    const unsigned num = m.match_for_triangulation(&k1, &k2, E, pairs);
    for (uint32_t i = 0; i < n1; ++i) matched_2_in_1[i] = -1;
    for (const auto& p : pairs) matched_2_in_1[p.first] = static_cast<int32_t>(p.second);
    return static_cast<int>(num);
}

// match/area.cc:8-112
int ref_match_area(const orc_camera* cam, uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc1, float* prev_pts, uint32_t n2,
                   const orc_keypoint* kp2, const uint8_t* desc2, int margin, float lowe_ratio, int check_orientation,
                   int32_t* matched_2_in_1) {
    table_camera camera(cam);
    data::frame f1, f2;
    fill_frame(f1, &camera, n1, kp1, desc1, nullptr, nullptr, 0);
    fill_frame(f2, &camera, n2, kp2, desc2, nullptr, nullptr, 0);
    std::vector<cv::Point2f> prev(n1);
    for (uint32_t i = 0; i < n1; ++i) prev[i] = cv::Point2f(prev_pts[2 * i], prev_pts[2 * i + 1]);
    std::vector<int> matched;
    match::area m(lowe_ratio, check_orientation != 0);
    const unsigned num = m.match_in_consistent_area(f1, f2, prev, matched, margin);
    for (uint32_t i = 0; i < n1; ++i) {
        matched_2_in_1[i] = matched[i];
        prev_pts[2 * i] = prev[i].x;
        prev_pts[2 * i + 1] = prev[i].y;
    }
    return static_cast<int>(num);
}

#endif  // DROPIN_SUBSET
// data/landmark.cc:127-183: index (in the given order) of the representative descriptor. observations_ is a std::map keyed
// by keyframe ADDRESS: the keyframes sit in one array, so address order = given order.
uint32_t ref_landmark_descriptor(const uint8_t* descs, uint32_t n) {
    Scene sc;
    std::vector<data::keyframe> kfs(n);
    auto* lm = sc.add_lm();
    for (uint32_t i = 0; i < n; ++i) {
        kfs[i].descriptors_ = desc_mat(descs + 32 * static_cast<size_t>(i), 1);
        kfs[i].stereo_x_right_.assign(1, -1.0f);
        lm->add_observation(&kfs[i], 0);
    }
    lm->compute_descriptor();
    const cv::Mat d = lm->get_descriptor();
    for (uint32_t i = 0; i < n; ++i)
        if (std::memcmp(d.ptr(0), descs + 32 * static_cast<size_t>(i), 32) == 0) return i;  // first row with these bytes
    return 0xFFFFFFFFu;
}
// the descriptor bytes themselves (rows with equal bytes are indistinguishable by index)
void ref_landmark_descriptor_bytes(const uint8_t* descs, uint32_t n, uint8_t* out32) {
    Scene sc;
    std::vector<data::keyframe> kfs(n);
    auto* lm = sc.add_lm();
    for (uint32_t i = 0; i < n; ++i) {
        kfs[i].descriptors_ = desc_mat(descs + 32 * static_cast<size_t>(i), 1);
        kfs[i].stereo_x_right_.assign(1, -1.0f);
        lm->add_observation(&kfs[i], 0);
    }
    lm->compute_descriptor();
    const cv::Mat d = lm->get_descriptor();
    std::memcpy(out32, d.ptr(0), 32);
}

// data/common.cc:97-186 + data/common.h:59-64 with the real perspective camera
void ref_assign_grid(const orc_camera* cam, uint32_t n, const orc_keypoint* kp, uint32_t* cell_start, uint32_t* items) {
    table_camera camera(cam);
    const auto cells = data::assign_keypoints_to_grid(&camera, to_cv(kp, n));
    uint32_t pos = 0;
    size_t c = 0;
    for (const auto& col : cells)
        for (const auto& cell : col) {
            cell_start[c++] = pos;
            for (unsigned i : cell) items[pos++] = i;
        }
    cell_start[c] = pos;
}
int ref_get_cell_indices(const orc_camera* cam, float x, float y, int* cx, int* cy) {
    table_camera camera(cam);
    cv::KeyPoint k;
    k.pt.x = x;
    k.pt.y = y;
    return data::get_cell_indices(&camera, k, *cx, *cy) ? 1 : 0;
}
int ref_keypoints_in_cell(const orc_camera* cam, uint32_t n, const orc_keypoint* kp, float ref_x, float ref_y, float margin, int min_level,
                          int max_level, uint32_t* out) {
    table_camera camera(cam);
    const auto kps = to_cv(kp, n);
    const auto cells = data::assign_keypoints_to_grid(&camera, kps);
    const auto idx = data::get_keypoints_in_cell(&camera, kps, cells, ref_x, ref_y, margin, min_level, max_level);
    for (size_t i = 0; i < idx.size(); ++i) out[i] = idx[i];
    return static_cast<int>(idx.size());
}

// camera/perspective.cc:20-31 (bounds + grid constants), :98-141 (undistortion, bearings), :151-169 (reprojection)
void ref_camera_info(double fx, double fy, double cx, double cy, const double* dist5, uint32_t cols, uint32_t rows, double focal_x_baseline,
                     orc_camera* out) {
    camera::perspective c("ref", camera::setup_type_t::Stereo, camera::color_order_t::Gray, cols, rows, 30.0, fx, fy, cx, cy, dist5[0],
                          dist5[1], dist5[2], dist5[3], dist5[4], focal_x_baseline);
    out->fx = fx, out->fy = fy, out->cx = cx, out->cy = cy;
    out->focal_x_baseline = c.focal_x_baseline_;
    out->true_baseline = c.true_baseline_;
    out->min_x = c.img_bounds_.min_x_, out->max_x = c.img_bounds_.max_x_;
    out->min_y = c.img_bounds_.min_y_, out->max_y = c.img_bounds_.max_y_;
    out->grid_cols = c.num_grid_cols_, out->grid_rows = c.num_grid_rows_;
    out->inv_cell_width = c.inv_cell_width_, out->inv_cell_height = c.inv_cell_height_;
}
void ref_undistort_keypoints(double fx, double fy, double cx, double cy, const double* dist5, uint32_t n, const orc_keypoint* in,
                             orc_keypoint* out) {
    camera::perspective c("ref", camera::setup_type_t::Monocular, camera::color_order_t::Gray, 640, 480, 30.0, fx, fy, cx, cy, dist5[0],
                          dist5[1], dist5[2], dist5[3], dist5[4]);
    std::vector<cv::KeyPoint> und;
    c.undistort_keypoints(to_cv(in, n), und);
    if (n) std::memcpy(static_cast<void*>(out), und.data(), sizeof(orc_keypoint) * n);
}
void ref_keypoints_to_bearings(double fx, double fy, double cx, double cy, uint32_t n, const orc_keypoint* kp, double* bearings) {
    camera::perspective c("ref", camera::setup_type_t::Monocular, camera::color_order_t::Gray, 640, 480, 30.0, fx, fy, cx, cy, 0, 0, 0, 0, 0);
    eigen_alloc_vector<Vec3_t> b;
    c.convert_keypoints_to_bearings(to_cv(kp, n), b);
    for (uint32_t i = 0; i < n; ++i)
        for (int k = 0; k < 3; ++k) bearings[3 * i + k] = b[i](k);
}
int ref_reproject_to_image(const orc_camera* cam, const double* pose_cw, const double* pos_w, double* reproj, float* x_right) {
    auto c = real_camera(cam);
    Mat33_t rot;
    Vec3_t trans;
    for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < 3; ++k) rot(r, k) = pose_cw[4 * r + k];
        trans(r) = pose_cw[4 * r + 3];
    }
    Vec2_t rp;
    const bool in = c->reproject_to_image(rot, trans, Vec3_t{pos_w[0], pos_w[1], pos_w[2]}, rp, *x_right);
    reproj[0] = rp(0);
    reproj[1] = rp(1);
    return in ? 1 : 0;
}

// data/common.cc:71-95: descriptors <-> JSON (the map file format of io/map_database_io.cc)
int ref_descriptors_to_json(const uint8_t* desc, uint32_t n, char* out, int cap) {
    const std::string s = data::convert_descriptors_to_json(desc_mat(desc, n)).dump();
    if (static_cast<int>(s.size()) + 1 > cap) return -static_cast<int>(s.size()) - 1;
    std::memcpy(out, s.c_str(), s.size() + 1);
    return static_cast<int>(s.size());
}
int ref_json_to_descriptors(const char* json, uint8_t* out, uint32_t cap_rows) {
    const cv::Mat d = data::convert_json_to_descriptors(nlohmann::json::parse(json));
    if (static_cast<uint32_t>(d.rows) > cap_rows) return -d.rows;
    for (int i = 0; i < d.rows; ++i) std::memcpy(out + 32 * static_cast<size_t>(i), d.ptr(i), 32);
    return d.rows;
}

}  // extern "C"
