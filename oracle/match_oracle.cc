// TEST INFRASTRUCTURE — NOT PRODUCT CODE. See orb_oracle.h.
// Restatement over flat arrays of the reference's descriptor matchers. Paths relative to
// /root/reference/external/openvslam/src/openvslam/.
// Parity status: match/base.h known answers (test/openvslam/match/base.cc:11-57) and match/stereo.cc (compiled unmodified
// into oracle/_ref) pin the Hamming and stereo paths; robust.cc / projection.cc depend on Eigen + g2o + DBoW2 and cannot be
// compiled here, so those two restatements are "parity unpinned" beyond line-by-line review and the angle_checker cases.
#include "orb_oracle.h"
#include "cvprim.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <numeric>
#include <unordered_set>
#include <utility>
#include <vector>

namespace {

constexpr unsigned HAMMING_DIST_THR_LOW = 50;   // match/base.h:13
constexpr unsigned HAMMING_DIST_THR_HIGH = 100; // match/base.h:14
constexpr unsigned MAX_HAMMING_DIST = 256;      // match/base.h:15

// match/base.h:18-40 — SWAR popcount on 8 x 32-bit words
unsigned hamming_32(const uint8_t* a, const uint8_t* b) {
    const uint32_t* pa = reinterpret_cast<const uint32_t*>(a);
    const uint32_t* pb = reinterpret_cast<const uint32_t*>(b);
    unsigned dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t v = pa[i] ^ pb[i];
        v -= (v >> 1) & 0x55555555U;
        v = (v & 0x33333333U) + ((v >> 2) & 0x33333333U);
        dist += (((v + (v >> 4)) & 0x0F0F0F0FU) * 0x01010101U) >> 24;
    }
    return dist;
}

// match/base.h:43-63 — SWAR popcount on 4 x 64-bit words
unsigned hamming_64(const uint8_t* a, const uint8_t* b) {
    const uint64_t* pa = reinterpret_cast<const uint64_t*>(a);
    const uint64_t* pb = reinterpret_cast<const uint64_t*>(b);
    unsigned dist = 0;
    for (int i = 0; i < 4; ++i) {
        uint64_t v = pa[i] ^ pb[i];
        v -= (v >> 1) & 0x5555555555555555ULL;
        v = (v & 0x3333333333333333ULL) + ((v >> 2) & 0x3333333333333333ULL);
        dist += static_cast<unsigned>((((v + (v >> 4)) & 0x0F0F0F0F0F0F0F0FULL) * 0x0101010101010101ULL) >> 56);
    }
    return dist;
}

// match/angle_checker.h:60-134, histogram_length = 30, num_bins_thr = 3
struct AngleHistogram {
    static constexpr unsigned kLen = 30;
    unsigned kTop = 3;
    std::vector<std::vector<int>> bins = std::vector<std::vector<int>>(kLen);
    const float inv_len = 1.0f / kLen;

    void append(float delta, int id) {
        if (delta < 0.0) delta += 360.0;
        if (360.0 <= delta) delta -= 360.0;
        const unsigned bin = static_cast<unsigned>(cvprim::round_half_even(delta * inv_len));
        bins.at(bin).push_back(id);
    }
    std::vector<int> invalid() const {
        // angle_checker.h:126-134: the same std::sort call on the same container type, so that libstdc++'s handling of
        // equal bin sizes is reproduced exactly
        std::vector<unsigned int> order(bins.size());
        std::iota(order.begin(), order.end(), 0);
        std::sort(order.begin(), order.end(),
                  [this](const unsigned int a, const unsigned int b) { return bins.at(a).size() > bins.at(b).size(); });
        std::vector<int> out;
        for (unsigned bin = 0; bin < kLen; ++bin) {
            const bool keep = std::any_of(order.begin(), order.begin() + kTop, [bin](unsigned i) { return bin == i; });
            if (!keep) out.insert(out.end(), bins[bin].begin(), bins[bin].end());
        }
        return out;
    }
};

} // namespace

extern "C" {

uint32_t orc_hamming_32(const uint8_t* a, const uint8_t* b) { return hamming_32(a, b); }
uint32_t orc_hamming_64(const uint8_t* a, const uint8_t* b) { return hamming_64(a, b); }

int orc_angle_invalid(const float* delta_angles, const int32_t* ids, int n, int32_t* invalid_out) {
    AngleHistogram h;
    for (int i = 0; i < n; ++i) h.append(delta_angles[i], ids[i]);
    const auto inv = h.invalid();
    std::copy(inv.begin(), inv.end(), invalid_out);
    return static_cast<int>(inv.size());
}

// angle_checker<int>(30, num_bins_thr): valid and invalid ids (test/openvslam/match/angle_checker.cc runs thresholds 1, 2, 3, 30)
int orc_angle_check(const float* delta_angles, const int32_t* ids, int n, unsigned num_bins_thr, int32_t* invalid_out) {
    AngleHistogram h;
    h.kTop = num_bins_thr;
    for (int i = 0; i < n; ++i) h.append(delta_angles[i], ids[i]);
    const auto inv = h.invalid();
    std::copy(inv.begin(), inv.end(), invalid_out);
    return static_cast<int>(inv.size());
}

// match/robust.cc:215-321
int orc_match_bruteforce(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint8_t* valid_2,
                         const float* angle_1, const float* angle_2, float lowe_ratio, int check_orientation,
                         int32_t* match_2_in_1) {
    int num = 0;
    AngleHistogram hist;
    std::fill(match_2_in_1, match_2_in_1 + n1, -1);
    std::vector<uint8_t> claimed(n1, 0);
    for (uint32_t i2 = 0; i2 < n2; ++i2) {
        if (valid_2 && !valid_2[i2]) continue; // robust.cc:240-246: no landmark / will be erased
        const uint8_t* d2 = desc_2 + static_cast<size_t>(i2) * 32;
        unsigned best = MAX_HAMMING_DIST, second = MAX_HAMMING_DIST;
        int best_i1 = -1;
        for (uint32_t i1 = 0; i1 < n1; ++i1) {
            if (claimed[i1]) continue; // robust.cc:258
            const unsigned d = hamming_32(d2, desc_1 + static_cast<size_t>(i1) * 32);
            if (d < best) {
                second = best;
                best = d;
                best_i1 = static_cast<int>(i1);
            }
            else if (d < second) {
                second = d;
            }
        }
        if (HAMMING_DIST_THR_LOW < best) continue;
        if (best_i1 < 0) continue;
        if (lowe_ratio * second < static_cast<float>(best)) continue; // robust.cc:285
        match_2_in_1[best_i1] = static_cast<int32_t>(i2);
        claimed[best_i1] = 1;
        if (check_orientation) hist.append(angle_1[best_i1] - angle_2[i2], best_i1);
        ++num;
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            match_2_in_1[id] = -1;
            --num;
        }
    }
    return num;
}

// match/stereo.cc:18-249
int orc_match_stereo(const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* lw, const int* lh,
                     const size_t* lpitch_l, const size_t* lpitch_r, int num_levels,
                     const orc_keypoint* kp_l, uint32_t nl, const orc_keypoint* kp_r, uint32_t nr,
                     const uint8_t* desc_l, const uint8_t* desc_r, const float* scale_factors,
                     const float* inv_scale_factors, float focal_x_baseline, float true_baseline,
                     float* stereo_x_right, float* depths) {
    (void)num_levels;
    (void)lh;
    constexpr unsigned hamm_thr = (HAMMING_DIST_THR_HIGH + HAMMING_DIST_THR_LOW) / 2; // stereo.h:99
    const float min_disp = 0.0f, max_disp = focal_x_baseline / true_baseline;      // stereo.cc:16
    // stereo.cc:114-141: right keypoints bucketed by every row within +-2*scale of their y
    const unsigned rows = static_cast<unsigned>(lh[0]);
    std::vector<std::vector<unsigned>> in_row(rows);
    for (unsigned ir = 0; ir < nr; ++ir) {
        const float y = kp_r[ir].y;
        const float r = 2.0f * scale_factors[kp_r[ir].octave];
        const int hi = cvprim::ceil_int(y + r);
        const int lo = cvprim::floor_int(y - r);
        for (int row = lo; row <= hi; ++row) in_row.at(row).push_back(ir);
    }
    std::fill(stereo_x_right, stereo_x_right + nl, -1.0f);
    std::fill(depths, depths + nl, -1.0f);
    std::vector<std::pair<int, int>> corr_idx;
    for (unsigned il = 0; il < nl; ++il) {
        const orc_keypoint& kl = kp_l[il];
        const int oct = kl.octave;
        const auto& cands = in_row.at(static_cast<size_t>(kl.y));
        if (cands.empty()) continue;
        const float min_xr = kl.x - max_disp;
        const float max_xr = kl.x - min_disp;
        if (max_xr < 0) continue;
        // stereo.cc:143-176
        unsigned best_ir = 0, best_d = hamm_thr;
        for (unsigned ir : cands) {
            const orc_keypoint& kr = kp_r[ir];
            if (kr.octave < oct - 1 || kr.octave > oct + 1) continue;
            if (kr.x < min_xr || max_xr < kr.x) continue;
            const unsigned d = hamming_32(desc_l + static_cast<size_t>(il) * 32, desc_r + static_cast<size_t>(ir) * 32);
            if (d < best_d) {
                best_ir = ir;
                best_d = d;
            }
        }
        if (hamm_thr <= best_d) continue;
        // stereo.cc:178-249 sub-pixel refinement on the pyramid level of the left keypoint
        const float xr = kp_r[best_ir].x;
        const float inv_sf = inv_scale_factors[oct];
        const int sxl = cvprim::round_half_even(kl.x * inv_sf);
        const int syl = cvprim::round_half_even(kl.y * inv_sf);
        const int sxr = cvprim::round_half_even(xr * inv_sf);
        constexpr int win = 5, slide = 5;
        if (sxr - slide - win < 0 || lw[oct] <= sxr + slide + win) continue;
        const uint8_t* L = pyr_l[oct];
        const uint8_t* R = pyr_r[oct];
        const size_t pl = lpitch_l[oct], pr = lpitch_r[oct];
        float best_corr = static_cast<float>(UINT_MAX);
        int best_off = 0;
        float corr[2 * slide + 1];
        const float cl = L[static_cast<size_t>(syl) * pl + sxl];
        for (int off = -slide; off <= slide; ++off) {
            const float cr = R[static_cast<size_t>(syl) * pr + sxr + off];
            double sum = 0.0; // cv::norm(NORM_L1) on CV_32F accumulates in double; all terms are integers
            for (int dy = -win; dy <= win; ++dy)
                for (int dx = -win; dx <= win; ++dx) {
                    const float a = static_cast<float>(L[static_cast<size_t>(syl + dy) * pl + sxl + dx]) - cl;
                    const float b = static_cast<float>(R[static_cast<size_t>(syl + dy) * pr + sxr + off + dx]) - cr;
                    sum += std::fabs(a - b);
                }
            const float c = static_cast<float>(sum);
            if (c < best_corr) {
                best_corr = c;
                best_off = off;
            }
            corr[slide + off] = c;
        }
        if (best_off == -slide || best_off == slide) continue;
        const float c1 = corr[slide + best_off - 1], c2 = corr[slide + best_off], c3 = corr[slide + best_off + 1];
        const float x_delta = static_cast<float>((c1 - c3) / (2.0 * (c1 + c3) - 4.0 * c2));
        if (x_delta < -1.0 || 1.0 < x_delta) continue;
        float best_xr = scale_factors[oct] * (static_cast<float>(sxr + best_off) + x_delta);
        float best_disp = kl.x - best_xr;
        if (best_disp < min_disp || max_disp <= best_disp) continue;
        if (best_disp <= 0.0f) {
            best_disp = 0.01f;
            best_xr = kl.x - best_disp;
        }
        depths[il] = focal_x_baseline / best_disp;
        stereo_x_right[il] = best_xr;
        corr_idx.emplace_back(static_cast<int>(best_corr), static_cast<int>(il));
    }
    // stereo.cc:94-111: drop matches whose SAD exceeds twice the median
    std::sort(corr_idx.begin(), corr_idx.end());
    const size_t mid = corr_idx.size() / 2;
    const float median = corr_idx.empty() ? 0.0f : static_cast<float>(corr_idx[mid].first);
    const float thr = static_cast<float>(2.0 * median);
    int kept = static_cast<int>(corr_idx.size());
    for (size_t i = mid; i < corr_idx.size(); ++i) {
        if (thr < corr_idx[i].first) {
            stereo_x_right[corr_idx[i].second] = -1;
            depths[corr_idx[i].second] = -1;
            --kept;
        }
    }
    return kept;
}

// data/common.cc:97-118: cell (x, y) -> keypoint indices in insertion order
static std::vector<std::vector<unsigned>> assign_grid(const orc_camera* cam, uint32_t n, const orc_keypoint* kp) {
    const int gc = static_cast<int>(cam->grid_cols), gr = static_cast<int>(cam->grid_rows);
    std::vector<std::vector<unsigned>> grid(static_cast<size_t>(gc) * gr);
    for (unsigned i = 0; i < n; ++i) {
        const int cx = cvprim::round_half_even((kp[i].x - cam->min_x) * cam->inv_cell_width);
        const int cy = cvprim::round_half_even((kp[i].y - cam->min_y) * cam->inv_cell_height);
        if (0 <= cx && cx < gc && 0 <= cy && cy < gr) grid[static_cast<size_t>(cx) * gr + cy].push_back(i);
    }
    return grid;
}

// data/common.cc:127-186 get_keypoints_in_cell
static void keypoints_in_cell(const orc_camera* cam, const std::vector<std::vector<unsigned>>& grid, const orc_keypoint* kp,
                              float ref_x, float ref_y, float m, int min_level, int max_level, std::vector<unsigned>& idx) {
    idx.clear();
    const int gc = static_cast<int>(cam->grid_cols), gr = static_cast<int>(cam->grid_rows);
    const int min_cx = std::max(0, cvprim::floor_int((ref_x - cam->min_x - m) * cam->inv_cell_width));
    if (gc <= min_cx) return;
    const int max_cx = std::min(gc - 1, cvprim::ceil_int((ref_x - cam->min_x + m) * cam->inv_cell_width));
    if (max_cx < 0) return;
    const int min_cy = std::max(0, cvprim::floor_int((ref_y - cam->min_y - m) * cam->inv_cell_height));
    if (gr <= min_cy) return;
    const int max_cy = std::min(gr - 1, cvprim::ceil_int((ref_y - cam->min_y + m) * cam->inv_cell_height));
    if (max_cy < 0) return;
    const bool check_level = (0 < min_level) || (0 <= max_level);
    for (int cx = min_cx; cx <= max_cx; ++cx)
        for (int cy = min_cy; cy <= max_cy; ++cy)
            for (unsigned i : grid[static_cast<size_t>(cx) * gr + cy]) {
                if (check_level) {
                    if (kp[i].octave < min_level) continue;
                    if (0 <= max_level && max_level < kp[i].octave) continue;
                }
                const float ddx = kp[i].x - ref_x, ddy = kp[i].y - ref_y;
                if (std::abs(ddx) < m && std::abs(ddy) < m) idx.push_back(i);
            }
}

// The inner loop shared by the reprojection matchers of match/projection.cc, with the knobs in which they differ:
//   match_frame_and_landmarks (:11-83)   window (level - 1, level), thr HAMMING_DIST_THR_HIGH, ratio test, stereo gate
//   match_frame_and_keyframe (:210-312)  window (level - 1, level + 1), thr = hamm_dist_thr, no ratio test, no stereo gate,
//                                        'if (curr_frm.landmarks_.at(curr_idx)) continue', angle check
// The rows come with the reprojection / predicted level the caller computed (data/landmark.h:112-115 caches them for the
// local map; match_frame_and_keyframe computes them inline with camera/perspective.cc:151-169 and landmark.cc:239-256).
int orc_match_reprojected(const orc_camera* cam, float margin, float lowe_ratio, uint32_t hamm_thr, int level_lo, int level_hi,
                          int use_ratio, int use_x_right, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                          const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                          const uint8_t* lm_has_observation, const float* lm_angle, uint32_t n_kp, const orc_keypoint* kp,
                          const uint8_t* kp_desc, const float* kp_x_right, const uint8_t* kp_taken, const float* scale_factors,
                          int32_t* kp_match_lm) {
    int num = 0;
    AngleHistogram hist;
    const auto grid = assign_grid(cam, n_kp, kp);
    std::vector<uint8_t> taken(kp_taken, kp_taken + n_kp);
    std::fill(kp_match_lm, kp_match_lm + n_kp, -1);
    std::vector<unsigned> idx;
    for (unsigned j = 0; j < n_lm; ++j) {
        if (!lm_active[j]) continue;
        const int level = lm_scale_level[j];
        const float m = margin * scale_factors[level];
        keypoints_in_cell(cam, grid, kp, static_cast<float>(lm_reproj[2 * j]), static_cast<float>(lm_reproj[2 * j + 1]), m,
                          level + level_lo, level + level_hi, idx);
        if (idx.empty()) continue;
        const uint8_t* ld = lm_desc + static_cast<size_t>(j) * 32;
        unsigned best = MAX_HAMMING_DIST, second = MAX_HAMMING_DIST;
        int best_level = -1, second_level = -1, best_i = -1;
        for (unsigned i : idx) {
            if (taken[i]) continue;  // sees the assignments of earlier rows
            if (use_x_right && 0 < kp_x_right[i]) {
                const float err = std::abs(lm_x_right[j] - kp_x_right[i]);
                if (m < err) continue;
            }
            const unsigned d = hamming_32(ld, kp_desc + static_cast<size_t>(i) * 32);
            if (d < best) {
                second = best;
                best = d;
                second_level = best_level;
                best_level = kp[i].octave;
                best_i = static_cast<int>(i);
            }
            else if (d < second) {
                second_level = kp[i].octave;
                second = d;
            }
        }
        if (best <= hamm_thr && best_i >= 0) {
            if (use_ratio && best_level == second_level && best > lowe_ratio * second) continue;
            kp_match_lm[best_i] = static_cast<int32_t>(j);
            taken[best_i] = lm_has_observation ? lm_has_observation[j] : 1;
            ++num;
            if (lm_angle) hist.append(lm_angle[j] - kp[best_i].angle, best_i);
        }
    }
    if (lm_angle) {
        for (int id : hist.invalid()) {
            kp_match_lm[id] = -1;
            --num;
        }
    }
    return num;
}

// match/projection.cc:11-83 match_frame_and_landmarks (the local-map search of every tracked frame)
int orc_match_landmarks(const orc_camera* cam, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj /*n_lm x 2*/,
                        const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                        const uint8_t* lm_has_observation /* nullable = all 1 */, uint32_t n_kp, const orc_keypoint* kp,
                        const uint8_t* kp_desc, const float* kp_x_right, const uint8_t* kp_has_observed_lm,
                        const float* scale_factors, float margin, float lowe_ratio, int32_t* kp_match_lm) {
    return orc_match_reprojected(cam, margin, lowe_ratio, HAMMING_DIST_THR_HIGH, -1, 0, 1, 1, n_lm, lm_active, lm_reproj, lm_x_right,
                                 lm_scale_level, lm_desc, lm_has_observation, nullptr, n_kp, kp, kp_desc, kp_x_right,
                                 kp_has_observed_lm, scale_factors, kp_match_lm);
}

// The search loop shared by fuse::detect_duplication (match/fuse.cc:69-101) and fuse::replace_duplication
// (match/fuse.cc:176-238): every row (a landmark that passed the caller's visibility gates, with its reprojection, x_right
// and predicted scale level) takes the keypoint with the smallest Hamming distance (strict '<', first wins) among the
// keypoints inside margin * scale_factors[level] whose octave is level - 1 or level; rows are independent (nothing is
// marked as taken). chi_gate != 0 adds replace_duplication's reprojection-error gate (3 degrees of freedom for keypoints
// with a stereo match, 2 otherwise, fuse.cc:199-222). unsigned_levels != 0 reproduces replace_duplication's unsigned
// arithmetic: `scale_level < pred_scale_level - 1` wraps for a predicted level of 0 and rejects every candidate (fuse.cc:193).
// best_idx[j] = keypoint of row j or -1 (also when the best distance exceeds HAMMING_DIST_THR_LOW), best_dist[j] = its distance.
int orc_match_fuse(const orc_camera* cam, float margin, int chi_gate, int unsigned_levels, uint32_t n_lm, const uint8_t* lm_active,
                   const double* lm_reproj, const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                   uint32_t n_kp, const orc_keypoint* kp, const uint8_t* kp_desc, const float* kp_x_right,
                   const float* scale_factors, const float* inv_level_sigma_sq, int32_t* best_idx, uint32_t* best_dist) {
    int num = 0;
    const auto grid = assign_grid(cam, n_kp, kp);
    std::vector<unsigned> idx;
    for (unsigned j = 0; j < n_lm; ++j) {
        best_idx[j] = -1;
        best_dist[j] = MAX_HAMMING_DIST;
        if (!lm_active[j]) continue;
        const int pred = lm_scale_level[j];
        keypoints_in_cell(cam, grid, kp, static_cast<float>(lm_reproj[2 * j]), static_cast<float>(lm_reproj[2 * j + 1]),
                          margin * scale_factors[pred], -1, -1, idx);
        if (idx.empty()) continue;
        const uint8_t* ld = lm_desc + static_cast<size_t>(j) * 32;
        unsigned best = MAX_HAMMING_DIST;
        int best_i = -1;
        for (unsigned i : idx) {
            if (unsigned_levels) {
                const unsigned scale_level = static_cast<unsigned>(kp[i].octave), upred = static_cast<unsigned>(pred);
                if (scale_level < upred - 1 || upred < scale_level) continue;
            }
            else {
                const int scale_level = kp[i].octave;
                if (scale_level < pred - 1 || pred < scale_level) continue;
            }
            if (chi_gate) {
                if (kp_x_right[i] >= 0) {
                    const auto e_x = lm_reproj[2 * j] - kp[i].x;
                    const auto e_y = lm_reproj[2 * j + 1] - kp[i].y;
                    const auto e_x_right = lm_x_right[j] - kp_x_right[i];
                    const auto reproj_error_sq = e_x * e_x + e_y * e_y + e_x_right * e_x_right;
                    constexpr float chi_sq_3D = 7.81473;
                    if (chi_sq_3D < reproj_error_sq * inv_level_sigma_sq[kp[i].octave]) continue;
                }
                else {
                    const auto e_x = lm_reproj[2 * j] - kp[i].x;
                    const auto e_y = lm_reproj[2 * j + 1] - kp[i].y;
                    const auto reproj_error_sq = e_x * e_x + e_y * e_y;
                    constexpr float chi_sq_2D = 5.99146;
                    if (chi_sq_2D < reproj_error_sq * inv_level_sigma_sq[kp[i].octave]) continue;
                }
            }
            const unsigned d = hamming_32(ld, kp_desc + static_cast<size_t>(i) * 32);
            if (d < best) {
                best = d;
                best_i = static_cast<int>(i);
            }
        }
        best_dist[j] = best;
        if (HAMMING_DIST_THR_LOW < best) continue;
        best_idx[j] = best_i;
        ++num;
    }
    return num;
}

// match/bow_tree.cc:16-126 match_frame_and_keyframe with the feature vectors given as node -> index lists (sorted by
// node id, the iteration order of both vocabulary libraries' containers). kf_valid = keyframe keypoint holds a live landmark.
int orc_match_bow(uint32_t n_nodes_kf, const uint32_t* kf_node_id, const uint32_t* kf_node_start, const uint32_t* kf_items,
                  uint32_t n_nodes_f, const uint32_t* f_node_id, const uint32_t* f_node_start, const uint32_t* f_items,
                  const uint8_t* kf_valid, const uint8_t* kf_desc, const float* kf_angle, uint32_t n_f, const uint8_t* f_desc,
                  const float* f_angle, float lowe_ratio, int check_orientation, int32_t* f_match_kf) {
    int num = 0;
    AngleHistogram hist;
    std::fill(f_match_kf, f_match_kf + n_f, -1);
    uint32_t a = 0, b = 0;
    while (a < n_nodes_kf && b < n_nodes_f) {
        if (kf_node_id[a] == f_node_id[b]) {
            for (uint32_t qa = kf_node_start[a]; qa < kf_node_start[a + 1]; ++qa) {
                const uint32_t ki = kf_items[qa];
                if (!kf_valid[ki]) continue;
                const uint8_t* kd = kf_desc + static_cast<size_t>(ki) * 32;
                unsigned best = MAX_HAMMING_DIST, second = MAX_HAMMING_DIST;
                int best_f = -1;
                for (uint32_t qb = f_node_start[b]; qb < f_node_start[b + 1]; ++qb) {
                    const uint32_t fi = f_items[qb];
                    if (f_match_kf[fi] >= 0) continue;  // matched_lms_in_frm.at(frm_idx) != nullptr
                    const unsigned d = hamming_32(kd, f_desc + static_cast<size_t>(fi) * 32);
                    if (d < best) {
                        second = best;
                        best = d;
                        best_f = static_cast<int>(fi);
                    }
                    else if (d < second) {
                        second = d;
                    }
                }
                if (HAMMING_DIST_THR_LOW < best) continue;
                if (lowe_ratio * second < static_cast<float>(best)) continue;
                f_match_kf[best_f] = static_cast<int32_t>(ki);
                if (check_orientation) hist.append(kf_angle[ki] - f_angle[best_f], best_f);
                ++num;
            }
            ++a;
            ++b;
        }
        else if (kf_node_id[a] < f_node_id[b]) {
            while (a < n_nodes_kf && kf_node_id[a] < f_node_id[b]) ++a;  // lower_bound
        }
        else {
            while (b < n_nodes_f && f_node_id[b] < kf_node_id[a]) ++b;
        }
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            f_match_kf[id] = -1;
            --num;
        }
    }
    return num;
}

// match/bow_tree.cc:128-250 match_keyframes (Sim3 estimation between loop candidates): both sides are keyframes, only
// keypoints that hold a live landmark take part (valid_1 / valid_2), the result is indexed by keyframe-1 keypoint.
int orc_match_bow_keyframes(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const uint8_t* valid_1, const uint8_t* desc_1, const float* angle_1, uint32_t n2,
                            const uint8_t* valid_2, const uint8_t* desc_2, const float* angle_2, float lowe_ratio,
                            int check_orientation, int32_t* match_2_of_1) {
    int num = 0;
    AngleHistogram hist;
    std::fill(match_2_of_1, match_2_of_1 + n1, -1);
    std::vector<uint8_t> already_2(n2, 0);
    uint32_t a = 0, b = 0;
    while (a < n_nodes_1 && b < n_nodes_2) {
        if (node_id_1[a] == node_id_2[b]) {
            for (uint32_t qa = node_start_1[a]; qa < node_start_1[a + 1]; ++qa) {
                const uint32_t i1 = items_1[qa];
                if (!valid_1[i1]) continue;
                const uint8_t* d1 = desc_1 + static_cast<size_t>(i1) * 32;
                unsigned best = MAX_HAMMING_DIST, second = MAX_HAMMING_DIST;
                int best_2 = -1;
                for (uint32_t qb = node_start_2[b]; qb < node_start_2[b + 1]; ++qb) {
                    const uint32_t i2 = items_2[qb];
                    if (!valid_2[i2]) continue;
                    if (already_2[i2]) continue;
                    const unsigned d = hamming_32(d1, desc_2 + static_cast<size_t>(i2) * 32);
                    if (d < best) {
                        second = best;
                        best = d;
                        best_2 = static_cast<int>(i2);
                    }
                    else if (d < second) {
                        second = d;
                    }
                }
                if (HAMMING_DIST_THR_LOW < best) continue;
                if (lowe_ratio * second < static_cast<float>(best)) continue;
                match_2_of_1[i1] = best_2;
                already_2[best_2] = 1;
                ++num;
                if (check_orientation) hist.append(angle_1[i1] - angle_2[best_2], static_cast<int>(i1));
            }
            ++a;
            ++b;
        }
        else if (node_id_1[a] < node_id_2[b]) {
            while (a < n_nodes_1 && node_id_1[a] < node_id_2[b]) ++a;  // lower_bound
        }
        else {
            while (b < n_nodes_2 && node_id_2[b] < node_id_1[a]) ++b;
        }
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            match_2_of_1[id] = -1;
            --num;
        }
    }
    return num;
}

// match/projection.cc:317-413 match_by_Sim3_transform, the search loop (:372-410). Rows = the landmarks that passed the
// caller's gates (:333-370) with their reprojection into the keyframe and predicted scale level; kp_has_lm =
// matched_lms_in_keyfrm.at(idx) != nullptr on entry. kp_match_lm[i] = row whose landmark keypoint i received, or -1.
int orc_match_sim3(const orc_camera* cam, float margin, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                   const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp, const orc_keypoint* kp,
                   const uint8_t* kp_desc, const uint8_t* kp_has_lm, const float* scale_factors, int32_t* kp_match_lm) {
    int num = 0;
    const auto grid = assign_grid(cam, n_kp, kp);
    std::vector<uint8_t> matched(kp_has_lm, kp_has_lm + n_kp);
    std::fill(kp_match_lm, kp_match_lm + n_kp, -1);
    std::vector<unsigned> indices;
    for (unsigned j = 0; j < n_lm; ++j) {
        if (!lm_active[j]) continue;
        const unsigned pred_scale_level = static_cast<unsigned>(lm_scale_level[j]);
        keypoints_in_cell(cam, grid, kp, static_cast<float>(lm_reproj[2 * j]), static_cast<float>(lm_reproj[2 * j + 1]),
                          margin * scale_factors[pred_scale_level], -1, -1, indices);
        if (indices.empty()) continue;
        const uint8_t* ld = lm_desc + static_cast<size_t>(j) * 32;
        unsigned best_dist = MAX_HAMMING_DIST;
        int best_idx = -1;
        for (const auto idx : indices) {
            if (matched[idx]) continue;
            const unsigned scale_level = static_cast<unsigned>(kp[idx].octave);
            if (scale_level < pred_scale_level - 1 || pred_scale_level < scale_level) continue;  // unsigned, :385-390
            const unsigned d = hamming_32(ld, kp_desc + static_cast<size_t>(idx) * 32);
            if (d < best_dist) {
                best_dist = d;
                best_idx = static_cast<int>(idx);
            }
        }
        if (HAMMING_DIST_THR_LOW < best_dist) continue;
        matched[best_idx] = 1;
        kp_match_lm[best_idx] = static_cast<int32_t>(j);
        ++num;
    }
    return num;
}

// one direction of match/projection.cc:453-534 / :536-616: rows of keyframe A against the keypoints of keyframe B
static void mutual_one_way(const orc_camera* cam_b, float margin, uint32_t n_a, const uint8_t* active, const double* reproj,
                           const int32_t* level, const uint8_t* lm_desc, uint32_t n_b, const orc_keypoint* kp_b,
                           const uint8_t* desc_b, const float* scale_b, std::vector<int>& out) {
    out.assign(n_a, -1);
    const auto grid = assign_grid(cam_b, n_b, kp_b);
    std::vector<unsigned> indices;
    for (unsigned a = 0; a < n_a; ++a) {
        if (!active[a]) continue;
        const unsigned pred_scale_level = static_cast<unsigned>(level[a]);
        keypoints_in_cell(cam_b, grid, kp_b, static_cast<float>(reproj[2 * a]), static_cast<float>(reproj[2 * a + 1]),
                          margin * scale_b[pred_scale_level], -1, -1, indices);
        if (indices.empty()) continue;
        const uint8_t* ld = lm_desc + static_cast<size_t>(a) * 32;
        unsigned best_hamm_dist = MAX_HAMMING_DIST;
        int best_idx = -1;
        for (const auto idx : indices) {
            const unsigned scale_level = static_cast<unsigned>(kp_b[idx].octave);
            if (scale_level < pred_scale_level - 1 || pred_scale_level < scale_level) continue;
            const unsigned d = hamming_32(ld, desc_b + static_cast<size_t>(idx) * 32);
            if (d < best_hamm_dist) {
                best_hamm_dist = d;
                best_idx = static_cast<int>(idx);
            }
        }
        if (best_hamm_dist <= HAMMING_DIST_THR_HIGH) out[a] = best_idx;
    }
}

// match/projection.cc:415-636 match_keyframes_mutually: both search loops and the cross-check (:618-631)
int orc_match_mutual(const orc_camera* cam_1, const orc_camera* cam_2, float margin, uint32_t n_1, const uint8_t* lm1_active,
                     const double* lm1_reproj_in_2, const int32_t* lm1_level_in_2, const uint8_t* lm1_desc, const orc_keypoint* kp1,
                     const uint8_t* desc1, const float* scale_1, uint32_t n_2, const uint8_t* lm2_active,
                     const double* lm2_reproj_in_1, const int32_t* lm2_level_in_1, const uint8_t* lm2_desc, const orc_keypoint* kp2,
                     const uint8_t* desc2, const float* scale_2, int32_t* matched_2_in_1) {
    std::vector<int> two_in_1, one_in_2;
    mutual_one_way(cam_2, margin, n_1, lm1_active, lm1_reproj_in_2, lm1_level_in_2, lm1_desc, n_2, kp2, desc2, scale_2, two_in_1);
    mutual_one_way(cam_1, margin, n_2, lm2_active, lm2_reproj_in_1, lm2_level_in_1, lm2_desc, n_1, kp1, desc1, scale_1, one_in_2);
    int num = 0;
    for (unsigned i = 0; i < n_1; ++i) {
        matched_2_in_1[i] = -1;
        const int idx_2 = two_in_1[i];
        if (idx_2 < 0) continue;
        if (one_in_2[idx_2] == static_cast<int>(i)) {
            matched_2_in_1[i] = idx_2;
            ++num;
        }
    }
    return num;
}

// match/robust.cc:323-341 check_epipolar_constraint (E_12 row-major; products summed left to right)
static bool check_epipolar_constraint(const double* b1, const double* b2, const double* E, float bearing_1_scale_factor) {
    double epi[3];
    for (int r = 0; r < 3; ++r) epi[r] = E[3 * r] * b2[0] + E[3 * r + 1] * b2[1] + E[3 * r + 2] * b2[2];
    const double norm = std::sqrt(epi[0] * epi[0] + epi[1] * epi[1] + epi[2] * epi[2]);
    const double cos_residual = (epi[0] * b1[0] + epi[1] * b1[1] + epi[2] * b1[2]) / norm;
    const double residual_rad = M_PI / 2.0 - std::abs(std::acos(cos_residual));
    constexpr double residual_deg_thr = 0.2;
    constexpr double residual_rad_thr = residual_deg_thr * M_PI / 180.0;
    return residual_rad < residual_rad_thr * bearing_1_scale_factor;
}

// match/robust.cc:18-178 match_for_triangulation; feature vectors as sorted node ids + CSR index lists
int orc_match_triangulation(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc_1, const double* bearings_1,
                            const float* x_right_1, const uint8_t* has_lm_1, const float* scale_1, uint32_t n2,
                            const orc_keypoint* kp2, const uint8_t* desc_2, const double* bearings_2, const float* x_right_2,
                            const uint8_t* has_lm_2, const double* E_12, const double* epiplane_in_keyfrm_2, int check_orientation,
                            int32_t* matched_2_in_1) {
    int num = 0;
    AngleHistogram hist;
    std::vector<uint8_t> is_already_matched_in_keyfrm_2(n2, 0);
    std::fill(matched_2_in_1, matched_2_in_1 + n1, -1);
    uint32_t a = 0, b = 0;
    while (a < n_nodes_1 && b < n_nodes_2) {
        if (node_id_1[a] == node_id_2[b]) {
            for (uint32_t qa = node_start_1[a]; qa < node_start_1[a + 1]; ++qa) {
                const uint32_t idx_1 = items_1[qa];
                if (has_lm_1[idx_1]) continue;
                const bool is_stereo_keypt_1 = 0 <= x_right_1[idx_1];
                const double* bearing_1 = bearings_1 + 3 * static_cast<size_t>(idx_1);
                const uint8_t* d1 = desc_1 + static_cast<size_t>(idx_1) * 32;
                unsigned best_hamm_dist = HAMMING_DIST_THR_LOW;
                int best_idx_2 = -1;
                for (uint32_t qb = node_start_2[b]; qb < node_start_2[b + 1]; ++qb) {
                    const uint32_t idx_2 = items_2[qb];
                    if (has_lm_2[idx_2]) continue;
                    if (is_already_matched_in_keyfrm_2[idx_2]) continue;
                    const bool is_stereo_keypt_2 = 0 <= x_right_2[idx_2];
                    const double* bearing_2 = bearings_2 + 3 * static_cast<size_t>(idx_2);
                    const unsigned hamm_dist = hamming_32(d1, desc_2 + static_cast<size_t>(idx_2) * 32);
                    if (HAMMING_DIST_THR_LOW < hamm_dist || best_hamm_dist < hamm_dist) continue;
                    if (!is_stereo_keypt_1 && !is_stereo_keypt_2) {
                        const double cos_dist = epiplane_in_keyfrm_2[0] * bearing_2[0] + epiplane_in_keyfrm_2[1] * bearing_2[1]
                                                + epiplane_in_keyfrm_2[2] * bearing_2[2];
                        constexpr double cos_dist_thr = 0.99862953475;
                        if (cos_dist_thr < cos_dist) continue;
                    }
                    if (check_epipolar_constraint(bearing_1, bearing_2, E_12, scale_1[kp1[idx_1].octave])) {
                        best_idx_2 = static_cast<int>(idx_2);
                        best_hamm_dist = hamm_dist;
                    }
                }
                if (best_idx_2 < 0) continue;
                is_already_matched_in_keyfrm_2[best_idx_2] = 1;
                matched_2_in_1[idx_1] = best_idx_2;
                ++num;
                if (check_orientation) hist.append(kp1[idx_1].angle - kp2[best_idx_2].angle, static_cast<int>(idx_1));
            }
            ++a;
            ++b;
        }
        else if (node_id_1[a] < node_id_2[b]) {
            while (a < n_nodes_1 && node_id_1[a] < node_id_2[b]) ++a;  // lower_bound
        }
        else {
            while (b < n_nodes_2 && node_id_2[b] < node_id_1[a]) ++b;
        }
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            matched_2_in_1[id] = -1;
            --num;
        }
    }
    return num;
}

// match/area.cc:8-112 match_in_consistent_area over flat arrays; prev_pts (n1 x 2) is updated in place
int orc_match_area(const orc_camera* cam, uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc1, float* prev_pts, uint32_t n2,
                   const orc_keypoint* kp2, const uint8_t* desc2, int margin, float lowe_ratio, int check_orientation,
                   int32_t* matched_2_in_1) {
    int num = 0;
    AngleHistogram hist;
    const auto grid = assign_grid(cam, n2, kp2);
    std::fill(matched_2_in_1, matched_2_in_1 + n1, -1);
    std::vector<unsigned> matched_dists_2(n2, MAX_HAMMING_DIST);
    std::vector<int> matched_1_in_2(n2, -1);
    std::vector<unsigned> idx;
    for (unsigned i1 = 0; i1 < n1; ++i1) {
        const int level = kp1[i1].octave;
        if (0 < level) continue;
        keypoints_in_cell(cam, grid, kp2, prev_pts[2 * i1], prev_pts[2 * i1 + 1], static_cast<float>(margin), level, level, idx);
        if (idx.empty()) continue;
        const uint8_t* d1 = desc1 + static_cast<size_t>(i1) * 32;
        unsigned best = MAX_HAMMING_DIST, second = MAX_HAMMING_DIST;
        int best_2 = -1;
        for (unsigned i2 : idx) {
            const unsigned d = hamming_32(d1, desc2 + static_cast<size_t>(i2) * 32);
            if (matched_dists_2[i2] <= d) continue;
            if (d < best) {
                second = best;
                best = d;
                best_2 = static_cast<int>(i2);
            }
            else if (d < second) {
                second = d;
            }
        }
        if (HAMMING_DIST_THR_LOW < best) continue;
        if (second * lowe_ratio < static_cast<float>(best)) continue;
        const int prev_1 = matched_1_in_2[best_2];
        if (0 <= prev_1) {
            matched_2_in_1[prev_1] = -1;
            --num;
        }
        matched_2_in_1[i1] = best_2;
        matched_1_in_2[best_2] = static_cast<int>(i1);
        matched_dists_2[best_2] = best;
        ++num;
        if (check_orientation) hist.append(kp1[i1].angle - kp2[best_2].angle, static_cast<int>(i1));
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            if (0 <= matched_2_in_1[id]) {
                matched_2_in_1[id] = -1;
                --num;
            }
        }
    }
    for (unsigned i1 = 0; i1 < n1; ++i1)
        if (0 <= matched_2_in_1[i1]) {
            prev_pts[2 * i1] = kp2[matched_2_in_1[i1]].x;
            prev_pts[2 * i1 + 1] = kp2[matched_2_in_1[i1]].y;
        }
    return num;
}

// data/landmark.cc:151-177 for one landmark: index of the representative descriptor among its n observations
uint32_t orc_landmark_descriptor(const uint8_t* descs, uint32_t n) {
    std::vector<std::vector<unsigned>> hamm(n, std::vector<unsigned>(n));
    for (unsigned i = 0; i < n; ++i) {
        hamm[i][i] = 0;
        for (unsigned j = i + 1; j < n; ++j) {
            const unsigned d = hamming_32(descs + static_cast<size_t>(i) * 32, descs + static_cast<size_t>(j) * 32);
            hamm[i][j] = d;
            hamm[j][i] = d;
        }
    }
    unsigned best_median = MAX_HAMMING_DIST, best = 0;
    for (unsigned idx = 0; idx < n; ++idx) {
        std::vector<unsigned> part(hamm[idx].begin(), hamm[idx].begin() + n);
        std::sort(part.begin(), part.end());
        const unsigned median = part.at(static_cast<unsigned>(0.5 * (n - 1)));
        if (median < best_median) {
            best_median = median;
            best = idx;
        }
    }
    return best;
}

// match/projection.cc:85-208 with data/common.cc:97-186 and camera/perspective.cc:151-169
int orc_match_projection(const orc_camera* cam, const double* Tc, const double* Tl, uint32_t n_last,
                         const uint8_t* last_has_lm, const uint8_t* last_outlier, const double* last_pos_w,
                         const uint8_t* last_lm_desc, const int32_t* last_octave, const float* last_angle,
                         uint32_t n_curr, const orc_keypoint* kp, const uint8_t* curr_desc, const float* curr_x_right,
                         const uint8_t* curr_has_observed_lm, const float* scale_factors, float margin,
                         int check_orientation, const uint8_t* last_lm_has_observation, int32_t* curr_match_last) {
    int num = 0;
    AngleHistogram hist;
    // data/common.cc:97-118 grid assignment, cell (x, y) -> indices in insertion order
    const int gc = static_cast<int>(cam->grid_cols), gr = static_cast<int>(cam->grid_rows);
    std::vector<std::vector<unsigned>> grid(static_cast<size_t>(gc) * gr);
    for (unsigned i = 0; i < n_curr; ++i) {
        const int cx = cvprim::round_half_even((kp[i].x - cam->min_x) * cam->inv_cell_width);
        const int cy = cvprim::round_half_even((kp[i].y - cam->min_y) * cam->inv_cell_height);
        if (0 <= cx && cx < gc && 0 <= cy && cy < gr) grid[static_cast<size_t>(cx) * gr + cy].push_back(i);
    }
    std::vector<uint8_t> taken(curr_has_observed_lm, curr_has_observed_lm + n_curr);
    std::fill(curr_match_last, curr_match_last + n_curr, -1);

    auto R = [](const double* T, int r, int c) { return T[r * 4 + c]; };
    // trans_wc = -R_cw^T t_cw ; trans_lc = R_lw trans_wc + t_lw  (projection.cc:90-99)
    double twc[3];
    for (int i = 0; i < 3; ++i) twc[i] = -((R(Tc, 0, i) * Tc[3] + R(Tc, 1, i) * Tc[7]) + R(Tc, 2, i) * Tc[11]);
    const double tlc_z = ((R(Tl, 2, 0) * twc[0] + R(Tl, 2, 1) * twc[1]) + R(Tl, 2, 2) * twc[2]) + Tl[11];
    const bool forward = cam->is_monocular ? false : tlc_z > cam->true_baseline;
    const bool backward = cam->is_monocular ? false : -tlc_z > cam->true_baseline;

    std::vector<unsigned> idx;
    for (unsigned il = 0; il < n_last; ++il) {
        if (!last_has_lm[il]) continue;
        if (last_outlier[il]) continue;
        const double* pw = last_pos_w + 3 * il;
        // camera/perspective.cc:151-169
        double pc[3];
        for (int i = 0; i < 3; ++i) pc[i] = ((R(Tc, i, 0) * pw[0] + R(Tc, i, 1) * pw[1]) + R(Tc, i, 2) * pw[2]) + Tc[i * 4 + 3];
        if (pc[2] <= 0.0) continue;
        const double z_inv = 1.0 / pc[2];
        const double u = cam->fx * pc[0] * z_inv + cam->cx;
        const double v = cam->fy * pc[1] * z_inv + cam->cy;
        const float x_right = static_cast<float>(u - cam->focal_x_baseline * z_inv);
        if (!(cam->min_x < u && u < cam->max_x && cam->min_y < v && v < cam->max_y)) continue;

        const int oct = last_octave[il];
        int min_level, max_level;
        if (forward) {
            min_level = oct;
            max_level = static_cast<int>(cam->num_scale_levels) - 1;
        }
        else if (backward) {
            min_level = 0;
            max_level = oct;
        }
        else {
            min_level = oct - 1;
            max_level = oct + 1;
        }
        const float ref_x = static_cast<float>(u), ref_y = static_cast<float>(v);
        const float m = margin * scale_factors[oct];
        // data/common.cc:127-186
        idx.clear();
        do {
            const int min_cx = std::max(0, cvprim::floor_int((ref_x - cam->min_x - m) * cam->inv_cell_width));
            if (gc <= min_cx) break;
            const int max_cx = std::min(gc - 1, cvprim::ceil_int((ref_x - cam->min_x + m) * cam->inv_cell_width));
            if (max_cx < 0) break;
            const int min_cy = std::max(0, cvprim::floor_int((ref_y - cam->min_y - m) * cam->inv_cell_height));
            if (gr <= min_cy) break;
            const int max_cy = std::min(gr - 1, cvprim::ceil_int((ref_y - cam->min_y + m) * cam->inv_cell_height));
            if (max_cy < 0) break;
            const bool check_level = (0 < min_level) || (0 <= max_level);
            for (int cx = min_cx; cx <= max_cx; ++cx)
                for (int cy = min_cy; cy <= max_cy; ++cy)
                    for (unsigned i : grid[static_cast<size_t>(cx) * gr + cy]) {
                        if (check_level) {
                            if (kp[i].octave < min_level) continue;
                            if (0 <= max_level && max_level < kp[i].octave) continue;
                        }
                        const float ddx = kp[i].x - ref_x, ddy = kp[i].y - ref_y;
                        if (std::abs(ddx) < m && std::abs(ddy) < m) idx.push_back(i);
                    }
        } while (false);
        if (idx.empty()) continue;

        const uint8_t* ld = last_lm_desc + static_cast<size_t>(il) * 32;
        unsigned best = MAX_HAMMING_DIST;
        int best_i = -1;
        for (unsigned i : idx) {
            if (taken[i]) continue; // projection.cc:163-165 (includes assignments made by earlier iterations)
            if (curr_x_right[i] > 0) {
                const float err = std::fabs(x_right - curr_x_right[i]);
                if (m < err) continue;
            }
            const unsigned d = hamming_32(ld, curr_desc + static_cast<size_t>(i) * 32);
            if (d < best) {
                best = d;
                best_i = static_cast<int>(i);
            }
        }
        if (HAMMING_DIST_THR_HIGH < best) continue;
        curr_match_last[best_i] = static_cast<int32_t>(il);
        // projection.cc:163-165 re-tests has_observation() of whatever landmark the keypoint holds NOW: a temporary
        // (stereo VO) landmark without observations does not block later rows
        taken[best_i] = last_lm_has_observation ? last_lm_has_observation[il] : 1;
        ++num;
        if (check_orientation) hist.append(last_angle[il] - kp[best_i].angle, best_i);
    }
    if (check_orientation) {
        for (int id : hist.invalid()) {
            curr_match_last[id] = -1;
            --num;
        }
    }
    return num;
}

void orc_knn_top2(const uint8_t* q, uint32_t nq, const uint8_t* db, uint64_t ndb, int64_t* best_idx, uint32_t* best_d,
                  uint32_t* second_d) {
    for (uint32_t i = 0; i < nq; ++i) {
        const uint8_t* qi = q + static_cast<size_t>(i) * 32;
        unsigned b = MAX_HAMMING_DIST, s = MAX_HAMMING_DIST;
        int64_t bi = -1;
        for (uint64_t j = 0; j < ndb; ++j) {
            const unsigned d = hamming_64(qi, db + j * 32);
            if (d < b) {
                s = b;
                b = d;
                bi = static_cast<int64_t>(j);
            }
            else if (d < s) {
                s = d;
            }
        }
        best_idx[i] = bi;
        best_d[i] = b;
        second_d[i] = s;
    }
}

} // extern "C"
