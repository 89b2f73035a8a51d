/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 * C interface of the CPU oracle ("port"): a restatement over flat arrays of the reference's ORB front end
 * (/root/reference/external/openvslam/src/openvslam/feature/orb_extractor.cc, orb_extractor_node.cc, orb_params.cc,
 * match/base.h, match/stereo.cc, match/robust.cc:215-321, match/projection.cc:85-208, data/common.cc:97-186,
 * match/angle_checker.h). Parity status: the OpenCV primitives are pinned bit-for-bit against cv2 4.13
 * (tests/test_oracle_vs_cv2.py + tests/golden/), the control flow is pinned against the reference's own sources compiled
 * unmodified into oracle/_ref (tests/test_oracle_vs_ref.py) and the golden fixtures generated from it.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 */
#ifndef LPSLAM_ORB_ORACLE_H
#define LPSLAM_ORB_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orc_keypoint; /* == cv::KeyPoint, 28 bytes */

typedef struct orc_extractor orc_extractor;

/* ---- primitives (cvprim.h) ---- */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch, uint8_t* dst, int dw, int dh, size_t dpitch);
void orc_gaussian7_u8(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch);
/* returns the number of keypoints; xys = (x, y, score) int triplets, capacity cap triplets */
int orc_fast9_nms(const uint8_t* img, int w, int h, size_t pitch, int thr, int32_t* xys, int cap);
float orc_fast_atan2(float y, float x);
float orc_cos(float v); /* util/trigonometric.h:26-42 */
float orc_sin(float v); /* util/trigonometric.h:44-46 */

/* ---- extractor (orb_extractor.cc) ---- */
orc_extractor* orc_create(uint32_t max_kpts, float scale_factor, uint32_t num_levels, uint32_t ini_fast_thr,
                          uint32_t min_fast_thr, const float* mask_rects, uint32_t num_mask_rects);
void orc_destroy(orc_extractor*);
void orc_set_max_kpts(orc_extractor*, uint32_t);
/* tables: which = 0 scale, 1 inv scale, 2 sigma^2, 3 inv sigma^2 (float[num_levels]); 4 = per-level quota (as float) */
int orc_get_table(const orc_extractor*, int which, float* out, int cap);
/* returns 0 on success, -2 if cap is too small (n_out still holds the needed count) */
int orc_extract(orc_extractor*, const uint8_t* img, int w, int h, size_t pitch, const uint8_t* mask, size_t mask_pitch,
                orc_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* n_out);
/* intermediates of the last extract call */
int orc_level_size(const orc_extractor*, int level, int* w, int* h);
int orc_level_copy(const orc_extractor*, int level, int blurred, uint8_t* dst, size_t pitch);
int orc_level_candidates(const orc_extractor*, int level, int32_t* xys, int cap); /* before the quadtree, border-relative */
int orc_level_selected(const orc_extractor*, int level, int32_t* xys, int cap);   /* after the quadtree, border-relative */

/* ---- matchers ---- */
uint32_t orc_hamming_32(const uint8_t* a, const uint8_t* b); /* match/base.h:18-40 */
uint32_t orc_hamming_64(const uint8_t* a, const uint8_t* b); /* match/base.h:43-63 */

/* match/angle_checker.h: ids whose bin is not among the num_bins_thr fullest; returns count, writes ids in bin order */
int orc_angle_invalid(const float* delta_angles, const int32_t* ids, int n, int32_t* invalid_out);
int orc_angle_check(const float* delta_angles, const int32_t* ids, int n, unsigned num_bins_thr, int32_t* invalid_out);

/* match/robust.cc:215-321 over flat arrays. Rows of desc_2 with valid_2[i]==0 are skipped (no live landmark).
 * match_2_in_1[nq] receives the keyframe index matched to each frame keypoint or -1. Returns the number of matches. */
int orc_match_bruteforce(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint8_t* valid_2,
                         const float* angle_1, const float* angle_2, float lowe_ratio, int check_orientation,
                         int32_t* match_2_in_1);

/* match/stereo.cc:18-249. Pyramids are arrays of level pointers with per-level w/h/pitch. */
int orc_match_stereo(const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* lw, const int* lh,
                     const size_t* lpitch_l, const size_t* lpitch_r, int num_levels,
                     const orc_keypoint* kp_l, uint32_t nl, const orc_keypoint* kp_r, uint32_t nr,
                     const uint8_t* desc_l, const uint8_t* desc_r, const float* scale_factors,
                     const float* inv_scale_factors, float focal_x_baseline, float true_baseline,
                     float* stereo_x_right, float* depths);

/* match/projection.cc:85-208 + data/common.cc:97-186 + camera/perspective.cc:151-169 over flat arrays. */
typedef struct {
    double fx, fy, cx, cy, focal_x_baseline, true_baseline;
    float min_x, max_x, min_y, max_y; /* image bounds (camera/base.h image_bounds) */
    uint32_t grid_cols, grid_rows;    /* 64 x 48 (camera/base.h:62) */
    double inv_cell_width, inv_cell_height;
    int32_t is_monocular;
    uint32_t num_scale_levels;
} orc_camera;

int orc_match_projection(const orc_camera* cam, const double* pose_cw_curr /*4x4 row-major*/,
                         const double* pose_cw_last /*4x4 row-major*/,
                         /* last frame */
                         uint32_t n_last, const uint8_t* last_has_lm, const uint8_t* last_outlier,
                         const double* last_pos_w /*n_last x 3*/, const uint8_t* last_lm_desc /*n_last x 32*/,
                         const int32_t* last_octave, const float* last_angle,
                         /* current frame */
                         uint32_t n_curr, const orc_keypoint* curr_undist_kp, const uint8_t* curr_desc,
                         const float* curr_x_right, const uint8_t* curr_has_observed_lm, const float* scale_factors,
                         float margin, int check_orientation,
                         const uint8_t* last_lm_has_observation /* n_last, NULL = all landmarks have observations */,
                         int32_t* curr_match_last /* n_curr: index of the last-frame landmark or -1 */);

/* match/projection.cc:11-83 match_frame_and_landmarks over flat arrays. lm_active = is_observable_in_tracking_ &&
 * !will_be_erased(); kp_match_lm[i] = index of the landmark assigned to keypoint i by this call, or -1. Returns num_matches. */
int orc_match_landmarks(const orc_camera* cam, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj /*n_lm x 2*/,
                        const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc /*n_lm x 32*/,
                        const uint8_t* lm_has_observation /* nullable = all 1 */, uint32_t n_kp,
                        const orc_keypoint* kp_undist, const uint8_t* kp_desc, const float* kp_x_right,
                        const uint8_t* kp_has_observed_lm, const float* scale_factors, float margin, float lowe_ratio,
                        int32_t* kp_match_lm);

/* data/frame.cc:37-51,75-92 post-processing for the perspective model (camera/perspective.cc:98-141, data/common.cc:97-118) */
void orc_undistort_keypoints(double fx, double fy, double cx, double cy, const double* dist5 /*k1 k2 p1 p2 k3*/, uint32_t n,
                             const orc_keypoint* in, orc_keypoint* out);
void orc_keypoints_to_bearings(double fx, double fy, double cx, double cy, uint32_t n, const orc_keypoint* kp, double* bearings);
void orc_assign_grid(const orc_camera* cam, uint32_t n, const orc_keypoint* kp, uint32_t* cell_start /*cells + 1*/, uint32_t* items /*n*/);

/* image pre-processing ahead of extraction (image_oracle.cc): cvtColor -> GRAY, imadjust / stretchlim, remap INTER_LINEAR */
void orc_gray_u8(const uint8_t* src, int w, int h, size_t pitch, int channels /*3|4*/, int order /*0 RGB, 1 BGR*/, uint8_t* dst,
                 size_t dpitch);
void orc_imadjust_u8(const uint8_t* src, int w, int h, size_t pitch, int auto_limits, double low_in, double high_in,
                     double low_out, double high_out, uint8_t* dst, size_t dpitch, uint8_t* lut_out /*256, nullable*/,
                     double* limits_out /*2, nullable*/);
void orc_remap_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch, const float* map_x, const float* map_y,
                         size_t map_pitch /*floats*/, uint8_t* dst, int dw, int dh, size_t dpitch);

/* shared inner loop of the reprojection matchers of match/projection.cc (match_oracle.cc) */
int orc_match_reprojected(const orc_camera* cam, float margin, float lowe_ratio, uint32_t hamm_thr, int level_lo, int level_hi,
                          int use_ratio, int use_x_right, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                          const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                          const uint8_t* lm_has_observation, const float* lm_angle, uint32_t n_kp, const orc_keypoint* kp,
                          const uint8_t* kp_desc, const float* kp_x_right, const uint8_t* kp_taken, const float* scale_factors,
                          int32_t* kp_match_lm);

/* match/bow_tree.cc:16-126 match_frame_and_keyframe; feature vectors as sorted node ids + CSR index lists.
 * f_match_kf[i] = keyframe keypoint whose landmark frame keypoint i received, or -1. Returns num_matches. */
int orc_match_fuse(const orc_camera* cam, float margin, int chi_gate, int unsigned_levels, uint32_t n_lm, const uint8_t* lm_active,
                   const double* lm_reproj, const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                   uint32_t n_kp, const orc_keypoint* kp, const uint8_t* kp_desc, const float* kp_x_right,
                   const float* scale_factors, const float* inv_level_sigma_sq, int32_t* best_idx, uint32_t* best_dist);
/* match/projection.cc:317-413 match_by_Sim3_transform search loop; :415-636 match_keyframes_mutually (both loops + cross-check);
 * match/robust.cc:18-178 match_for_triangulation (E_12 row-major). Each returns num_matches. */
int orc_match_sim3(const orc_camera* cam, float margin, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                   const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp, const orc_keypoint* kp,
                   const uint8_t* kp_desc, const uint8_t* kp_has_lm, const float* scale_factors, int32_t* kp_match_lm);
int orc_match_mutual(const orc_camera* cam_1, const orc_camera* cam_2, float margin, uint32_t n_1, const uint8_t* lm1_active,
                     const double* lm1_reproj_in_2, const int32_t* lm1_level_in_2, const uint8_t* lm1_desc, const orc_keypoint* kp1,
                     const uint8_t* desc1, const float* scale_1, uint32_t n_2, const uint8_t* lm2_active,
                     const double* lm2_reproj_in_1, const int32_t* lm2_level_in_1, const uint8_t* lm2_desc, const orc_keypoint* kp2,
                     const uint8_t* desc2, const float* scale_2, int32_t* matched_2_in_1);
int orc_match_triangulation(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc_1, const double* bearings_1,
                            const float* x_right_1, const uint8_t* has_lm_1, const float* scale_1, uint32_t n2,
                            const orc_keypoint* kp2, const uint8_t* desc_2, const double* bearings_2, const float* x_right_2,
                            const uint8_t* has_lm_2, const double* E_12, const double* epiplane_in_keyfrm_2, int check_orientation,
                            int32_t* matched_2_in_1);
int orc_match_bow_keyframes(uint32_t n_nodes_1, const uint32_t* node_id_1, const uint32_t* node_start_1, const uint32_t* items_1,
                            uint32_t n_nodes_2, const uint32_t* node_id_2, const uint32_t* node_start_2, const uint32_t* items_2,
                            uint32_t n1, const uint8_t* valid_1, const uint8_t* desc_1, const float* angle_1, uint32_t n2,
                            const uint8_t* valid_2, const uint8_t* desc_2, const float* angle_2, float lowe_ratio,
                            int check_orientation, int32_t* match_2_of_1);
int orc_match_bow(uint32_t n_nodes_kf, const uint32_t* kf_node_id, const uint32_t* kf_node_start, const uint32_t* kf_items,
                  uint32_t n_nodes_f, const uint32_t* f_node_id, const uint32_t* f_node_start, const uint32_t* f_items,
                  const uint8_t* kf_valid, const uint8_t* kf_desc, const float* kf_angle, uint32_t n_f, const uint8_t* f_desc,
                  const float* f_angle, float lowe_ratio, int check_orientation, int32_t* f_match_kf);

/* match/area.cc:8-112 match_in_consistent_area over flat arrays (prev_pts n1 x 2 floats, in / out). Returns num_matches. */
int orc_match_area(const orc_camera* cam, uint32_t n1, const orc_keypoint* kp1, const uint8_t* desc1, float* prev_pts, uint32_t n2,
                   const orc_keypoint* kp2, const uint8_t* desc2, int margin, float lowe_ratio, int check_orientation,
                   int32_t* matched_2_in_1);

/* data/landmark.cc:151-177: representative descriptor (index) among n observation descriptors */
uint32_t orc_landmark_descriptor(const uint8_t* descs, uint32_t n);

/* exact brute-force top-2 (reference semantics of a sequential strict-< scan): per query best index, best and second distance */
void orc_knn_top2(const uint8_t* q, uint32_t nq, const uint8_t* db, uint64_t ndb, int64_t* best_idx, uint32_t* best_d,
                  uint32_t* second_d);

#ifdef __cplusplus
}
#endif
#endif
