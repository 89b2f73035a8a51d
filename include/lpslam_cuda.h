/* liblpslam_cuda — C ABI of the B200 (sm_100a) ORB front end for lpslam / OpenVSLAM.
 *
 * Every entry point replaces one piece of the reference's CPU path; the reference interface it stands in for is cited
 * beside it (paths relative to /root/reference/external/openvslam/src/openvslam/). The host-side C++ drop-in classes in
 * cpp/ (same names and signatures as the reference's) and the Python mirror in lpslam_b200/ both call only this ABI.
 * No torch / OpenCV / STL types cross it: plain pointers, sizes and POD structs.
 *
 * Conventions
 *   - every function returns 0 on success or a negative lpslam_status; lpslam_cuda_last_error() gives the text
 *     (thread-local). There is NO CPU fallback: without a CUDA device the calls fail with LPSLAM_ERR_CUDA.
 *   - "host" pointers may be pageable or pinned (lpslam_cuda_host_alloc gives pinned memory); "_dev" entry points take
 *     device pointers on the context's device and enqueue on the context's stream without synchronising.
 *   - an extractor context may be used by one thread at a time; distinct contexts are independent (own stream, own
 *     device buffers), which is what the reference's two-thread stereo extraction needs (data/frame.cc:65-68).
 */
#ifndef LPSLAM_CUDA_H
#define LPSLAM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LPSLAM_CUDA_ABI_VERSION 1

typedef enum {
    LPSLAM_OK = 0,
    LPSLAM_ERR_ARG = -1,      /* null pointer, zero size, unsupported parameter */
    LPSLAM_ERR_CAPACITY = -2, /* caller buffer too small; the needed count is still written to *n_out */
    LPSLAM_ERR_CUDA = -3,     /* CUDA runtime error (no device, launch failure, out of memory) */
    LPSLAM_ERR_STATE = -4     /* call order violated (e.g. pyramid requested before any extract) */
} lpslam_status;

/* == cv::KeyPoint (28 bytes): pt.x, pt.y, size, angle, response, octave, class_id */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} lpslam_keypoint;

/* feature/orb_params.h:9-46. mask_rects = num_mask_rects x (x_min, x_max, y_min, y_max) as fractions of the image. */
typedef struct {
    uint32_t max_num_keypts;   /* default 2000 */
    uint32_t ini_max_num_keypts;
    float scale_factor;        /* default 1.2 */
    uint32_t num_levels;       /* default 8 */
    uint32_t ini_fast_thr;     /* default 20 */
    uint32_t min_fast_thr;     /* default 7 */
    const float* mask_rects;
    uint32_t num_mask_rects;
} lpslam_orb_params;

typedef struct lpslam_orb_ctx lpslam_orb_ctx; /* one per orb_extractor instance (feature/orb_extractor.h:13-151) */

/* ---------------------------------------------------------------- library ---------------------------------------- */
int lpslam_cuda_abi_version(void);
const char* lpslam_cuda_last_error(void);
int lpslam_cuda_device_count(int* count);
/* pinned host memory for images / outputs (optional; any host memory is accepted by the host entry points) */
int lpslam_cuda_host_alloc(void** ptr, size_t bytes);
int lpslam_cuda_host_free(void* ptr);

/* ---------------------------------------------------------------- ORB extractor ---------------------------------- */
/* orb_extractor::orb_extractor (feature/orb_extractor.cc:46-69) + initialize() (:208-242). max_batch = number of frames a
 * single lpslam_orb_extract_batch call may carry (1 for the live tracker). Throws-equivalent: malformed mask
 * rectangles (feature/orb_params.cc:16-26) return LPSLAM_ERR_ARG. */
int lpslam_orb_create(lpslam_orb_ctx** ctx, const lpslam_orb_params* params, int device, uint32_t max_batch);
void lpslam_orb_destroy(lpslam_orb_ctx* ctx);

/* the five setters of feature/orb_extractor.cc:149-206 (max keypoints / scale factor / levels re-run initialize()) */
typedef enum {
    LPSLAM_ORB_MAX_NUM_KEYPTS = 0,
    LPSLAM_ORB_SCALE_FACTOR = 1,
    LPSLAM_ORB_NUM_LEVELS = 2,
    LPSLAM_ORB_INI_FAST_THR = 3,
    LPSLAM_ORB_MIN_FAST_THR = 4,
    /* read only, no reference counterpart: how many times this context replayed its one-pair CUDA graph (lpslam_stereo_frames) */
    LPSLAM_ORB_DEBUG_GRAPH_LAUNCHES = 100
} lpslam_orb_param_id;
int lpslam_orb_set_param(lpslam_orb_ctx* ctx, int which, double value);
int lpslam_orb_get_param(const lpslam_orb_ctx* ctx, int which, double* value);

/* get_scale_factors / get_inv_scale_factors / get_level_sigma_sq / get_inv_level_sigma_sq (feature/orb_extractor.cc:
 * 176-206 of the getters) and the per-level keypoint quota (:217-226). which: 0..3 as listed, 4 = quota (as float). */
int lpslam_orb_get_table(const lpslam_orb_ctx* ctx, int which, float* out, uint32_t cap);

/* orb_extractor::extract (feature/orb_extractor.cc:71-147). img: 8UC1, row pitch in bytes. mask: optional 8UC1 image
 * mask of the same size (0 = masked), NULL selects the rectangle mask of the parameters, if any.
 * Writes *n_out keypoints (octave order, quadtree order inside an octave) and n_out x 32 descriptor bytes.
 * An empty image (w or h == 0) returns LPSLAM_OK with *n_out = 0 and outputs untouched (:73-75). */
int lpslam_orb_extract(lpslam_orb_ctx* ctx, const uint8_t* img, int width, int height, size_t pitch,
                       const uint8_t* mask, size_t mask_pitch,
                       lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* n_out);

/* batched replay (config "4096 frames frame-parallel"): n_frames <= max_batch images of identical size, each frame
 * extracted exactly as by lpslam_orb_extract. Outputs are packed: frame f owns rows [offsets[f], offsets[f+1]).
 * offsets has n_frames + 1 entries. cap = capacity of kp_out / desc_out in keypoints. */
int lpslam_orb_extract_batch(lpslam_orb_ctx* ctx, const uint8_t* const* imgs, uint32_t n_frames, int width, int height,
                             size_t pitch, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets);

/* device-resident variant: imgs_dev = n_frames images stored back to back with row pitch `pitch` and frame stride
 * pitch*height in device memory; results stay on the device (see lpslam_orb_result_dev). Enqueues only. */
int lpslam_orb_extract_batch_dev(lpslam_orb_ctx* ctx, const uint8_t* imgs_dev, uint32_t n_frames, int width, int height,
                                 size_t pitch);
/* device pointers to the packed results of the last *_dev call (valid until the next call on this context) */
typedef struct {
    const lpslam_keypoint* kp;  /* packed keypoints */
    const uint8_t* desc;        /* packed descriptors, 32 B per row */
    const uint32_t* offsets;    /* n_frames + 1 */
    uint32_t capacity;          /* rows available in kp / desc */
} lpslam_orb_result_dev;
int lpslam_orb_get_result_dev(lpslam_orb_ctx* ctx, lpslam_orb_result_dev* out);
/* blocking copy of the packed results of the last *_dev call to the host (same layout as lpslam_orb_extract_batch;
 * offsets has n_frames + 1 entries) */
int lpslam_orb_fetch_result(lpslam_orb_ctx* ctx, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets);
/* the context's own level-0 batch buffer (max_batch frames, row pitch *pitch, frame stride *frame_stride) for the
 * given image size: a producer that writes frames straight into it (decoder, undistortion kernel) and then passes the
 * same pointer to lpslam_orb_extract_batch_dev avoids the ingest copy. */
int lpslam_orb_input_dev(lpslam_orb_ctx* ctx, int width, int height, uint8_t** ptr, size_t* pitch, size_t* frame_stride);
int lpslam_orb_sync(lpslam_orb_ctx* ctx);
void* lpslam_orb_stream(lpslam_orb_ctx* ctx); /* cudaStream_t of the context */

/* image_pyramid_ (public member, feature/orb_extractor.h:76; consumed by match::stereo, data/frame.cc:78).
 * Level geometry of the last extract, and a lazy device-to-host copy of one level of frame `frame` of the last call.
 * blurred != 0 returns the Gaussian-blurred level the descriptors were sampled from (test hook). */
int lpslam_orb_level_size(const lpslam_orb_ctx* ctx, uint32_t level, int* width, int* height);
int lpslam_orb_pyramid(lpslam_orb_ctx* ctx, uint32_t frame, uint32_t level, int blurred, uint8_t* dst, size_t pitch);
/* Levels first_level .. num_levels - 1 of frame `frame` in ONE device-to-host copy: every level tightly packed (row pitch =
 * level width) at dst + offsets[i], i = level - first_level; *total = bytes needed. LPSLAM_ERR_CAPACITY (with *total and
 * offsets set) when cap is too small or dst is null. What the drop-in extractor fills image_pyramid_ levels 1.. from
 * (feature/orb_extractor.cc:266-275 leaves them in image_pyramid_; data/frame.cc:78 hands them to match::stereo). */
int lpslam_orb_pyramid_levels(lpslam_orb_ctx* ctx, uint32_t frame, uint32_t first_level, uint8_t* dst, size_t cap, size_t* offsets,
                              size_t* total);

/* test hooks: intermediates of the last extract for frame `frame` — FAST candidates handed to the quadtree
 * (feature/orb_extractor.cc:352-369) and the keypoints it selected (:378), as (x, y, score) int32 triplets in
 * border-relative level coordinates, reference order. Returns the count through *n_out. */
int lpslam_orb_debug_candidates(lpslam_orb_ctx* ctx, uint32_t frame, uint32_t level, int32_t* xys, uint32_t cap,
                                uint32_t* n_out);
int lpslam_orb_debug_selected(lpslam_orb_ctx* ctx, uint32_t frame, uint32_t level, int32_t* xys, uint32_t cap,
                              uint32_t* n_out);

/* per-stage device time (ms, CUDA events on the context's stream) of the last extract call:
 * 0 H2D, 1 pyramid, 2 FAST, 3 quadtree, 4 blur, 5 orientation+descriptors, 6 D2H. Returns the number of stages. */
int lpslam_orb_stage_times(lpslam_orb_ctx* ctx, float* ms, uint32_t cap);
int lpslam_orb_enable_timing(lpslam_orb_ctx* ctx, int on);

/* ---------------------------------------------------------------- matchers --------------------------------------- */
/* match::robust::brute_force_match (match/robust.cc:215-321) over flat arrays. desc_1 = frame descriptors (n1 x 32),
 * desc_2 = keyframe descriptors (n2 x 32); valid_2[i] == 0 marks keyframe keypoints without a live landmark (:240-246),
 * NULL = all valid. angle_1/angle_2 (degrees) are read only when check_orientation != 0 (match/angle_checker.h).
 * match_2_in_1[n1] receives the keyframe index matched to each frame keypoint or -1; *n_matches the pair count. */
int lpslam_match_bruteforce(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint8_t* valid_2,
                            const float* angle_1, const float* angle_2, float lowe_ratio, int check_orientation,
                            int32_t* match_2_in_1, uint32_t* n_matches);

/* robust::brute_force_match between CONSECUTIVE frames of the batch a context extracted last (module/frame_tracker.cc:162 calls
 * it with ratio 0.8 and the orientation check off, once per tracked frame): frame f is the "frame" side, frame f - 1 the
 * "keyframe" side with every keypoint holding a landmark. Descriptors never leave the device. match_prev_in_cur has one entry per
 * packed keypoint row of the batch (frame 0: -1): index, inside frame f - 1, of the keypoint matched to that row, or -1.
 * n_matches has one entry per frame (entry 0 = 0). */
int lpslam_match_consecutive_frames(lpslam_orb_ctx* ctx, float lowe_ratio, int32_t* match_prev_in_cur, uint32_t* n_matches);

/* match::stereo::compute (match/stereo.cc:18-249; ctor match/stereo.h:18-29). Pyramids are arrays of num_levels host
 * level pointers with per-level width/height/pitch (the extractors' image_pyramid_). Outputs: stereo_x_right[nl],
 * depths[nl], -1 = no match. */
int lpslam_match_stereo(const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* level_w, const int* level_h,
                        const size_t* pitch_l, const size_t* pitch_r, uint32_t num_levels,
                        const lpslam_keypoint* kp_l, uint32_t nl, const lpslam_keypoint* kp_r, uint32_t nr,
                        const uint8_t* desc_l, const uint8_t* desc_r, const float* scale_factors,
                        const float* inv_scale_factors, float focal_x_baseline, float true_baseline,
                        float* stereo_x_right, float* depths, uint32_t* n_matches);

/* The same on the pyramids two extractor contexts still hold on the device from their last lpslam_orb_extract (left eye on
 * ctx_l, right eye on ctx_r; same device, same image size and level count): nothing but keypoints and descriptors travels.
 * The scale tables are the contexts' own. Used by the drop-in match::stereo when it is handed the image_pyramid_ of two
 * drop-in extractors (data/frame.cc:78). LPSLAM_ERR_STATE when a context has not extracted yet. */
int lpslam_match_stereo_ctx(lpslam_orb_ctx* ctx_l, lpslam_orb_ctx* ctx_r, const lpslam_keypoint* kp_l, uint32_t nl,
                            const lpslam_keypoint* kp_r, uint32_t nr, const uint8_t* desc_l, const uint8_t* desc_r,
                            float focal_x_baseline, float true_baseline, float* stereo_x_right, float* depths, uint32_t* n_matches);

/* fused stereo frames — data::frame's stereo constructor (data/frame.cc:55-93: two extractions + match::stereo) for
 * n_pairs stereo pairs in one pass, without the pyramid round trip: both eyes are extracted on the device by ONE context
 * (max_batch >= 2 * n_pairs; frame 2p = left eye, 2p + 1 = right eye of pair p) and the row matcher reads the
 * device-resident pyramids. Outputs are packed like lpslam_orb_extract_batch (offsets has 2 * n_pairs + 1 entries);
 * stereo_x_right / depths have one float per packed keypoint row and are meaningful on left-eye rows (-1 = no match). */
int lpslam_stereo_frames(lpslam_orb_ctx* ctx, const uint8_t* const* imgs_l, const uint8_t* const* imgs_r, uint32_t n_pairs,
                         int width, int height, size_t pitch, float focal_x_baseline, float true_baseline,
                         lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets,
                         float* stereo_x_right, float* depths);
/* the same, asynchronous: TWO batches in flight on one context, so that ONE host thread keeps the host link and the SMs busy
 * (what LP/Manager/ReplayEngine.h:40-53 does with a reader thread in front of the tracker: the next chunk is read while the
 * current one is processed). submit enqueues the host -> device copies of the batch (pinned host images; images that follow one
 * another in memory go out as one copy), the whole pipeline and a device-side copy of its results, and returns a ticket without
 * waiting; collect blocks until that batch is done and reads its results back. The image buffers must stay valid and unchanged
 * until their ticket is collected. At most two tickets may be outstanding (LPSLAM_ERR_STATE otherwise); collect in submit order. */
int lpslam_stereo_frames_submit(lpslam_orb_ctx* ctx, const uint8_t* const* imgs_l, const uint8_t* const* imgs_r, uint32_t n_pairs,
                                int width, int height, size_t pitch, float focal_x_baseline, float true_baseline, uint64_t* ticket);
int lpslam_stereo_frames_collect(lpslam_orb_ctx* ctx, uint64_t ticket, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap,
                                 uint32_t* offsets, float* stereo_x_right, float* depths);
/* extraction only, same protocol and the same two slots (frame-parallel replay of a log: LP/Manager/ReplayEngine.h:40-53) */
int lpslam_orb_extract_batch_submit(lpslam_orb_ctx* ctx, const uint8_t* const* imgs, uint32_t n_frames, int width, int height,
                                    size_t pitch, uint64_t* ticket);
int lpslam_orb_extract_batch_collect(lpslam_orb_ctx* ctx, uint64_t ticket, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap,
                                     uint32_t* offsets);
/* device-resident variant: imgs_dev holds 2 * n_pairs frames, left / right interleaved. Enqueues only. */
int lpslam_stereo_frames_dev(lpslam_orb_ctx* ctx, const uint8_t* imgs_dev, uint32_t n_pairs, int width, int height,
                             size_t pitch, float focal_x_baseline, float true_baseline);
/* device pointers to stereo_x_right / depths of the last lpslam_stereo_frames_dev call (one float per packed row) */
int lpslam_stereo_result_dev(lpslam_orb_ctx* ctx, const float** stereo_x_right_dev, const float** depths_dev);

/* match::projection::match_current_and_last_frames (match/projection.cc:85-208) with the grid query of
 * data/common.cc:97-186 and the perspective projection of camera/perspective.cc:151-169, over flat arrays. */
typedef struct {
    double fx, fy, cx, cy, focal_x_baseline, true_baseline;
    float min_x, max_x, min_y, max_y; /* camera/base.h image bounds */
    uint32_t grid_cols, grid_rows;    /* 64 x 48, camera/base.h:62 */
    double inv_cell_width, inv_cell_height;
    int32_t is_monocular;
    uint32_t num_scale_levels;
} lpslam_camera;
int lpslam_match_projection(const lpslam_camera* cam, const double* pose_cw_curr /*4x4 row-major*/,
                            const double* pose_cw_last /*4x4 row-major*/,
                            uint32_t n_last, const uint8_t* last_has_lm, const uint8_t* last_outlier,
                            const double* last_pos_w /*n_last x 3*/, const uint8_t* last_lm_desc /*n_last x 32*/,
                            const int32_t* last_octave, const float* last_angle,
                            const uint8_t* last_lm_has_observation /* nullable = all 1: a matched keypoint blocks later
                               landmarks only if the landmark it now holds has observations (projection.cc:163-165) */,
                            uint32_t n_curr, const lpslam_keypoint* curr_undist_kp, const uint8_t* curr_desc,
                            const float* curr_x_right, const uint8_t* curr_has_observed_lm, const float* scale_factors,
                            float margin, int check_orientation, int32_t* curr_match_last, uint32_t* n_matches);

/* match::projection::match_frame_and_landmarks (match/projection.cc:11-83), the local-map search run on every tracked
 * frame (module/local_map_updater + tracking_module::search_local_landmarks), over flat arrays. Each local landmark brings
 * the reprojection the tracker cached on it (data/landmark.h:112-115: reproj_in_tracking_, x_right_in_tracking_,
 * scale_level_in_tracking_); lm_active = is_observable_in_tracking_ && !will_be_erased(). kp_has_observed_lm[i] != 0 when
 * frm.landmarks_[i] exists and has observations (projection.cc:41-43); a keypoint matched by this call blocks later
 * landmarks only if the landmark it now holds has observations (lm_has_observation, nullable = all 1).
 * kp_match_lm[i] = index of the landmark written to frm.landmarks_[i] by this call, or -1 (unchanged). */
int lpslam_match_landmarks(const lpslam_camera* cam, uint32_t n_lm, const uint8_t* lm_active,
                           const double* lm_reproj /*n_lm x 2*/, const float* lm_x_right, const int32_t* lm_scale_level,
                           const uint8_t* lm_desc /*n_lm x 32*/, const uint8_t* lm_has_observation,
                           uint32_t n_kp, const lpslam_keypoint* kp_undist, const uint8_t* kp_desc, const float* kp_x_right,
                           const uint8_t* kp_has_observed_lm, const float* scale_factors, float margin, float lowe_ratio,
                           int32_t* kp_match_lm, uint32_t* n_matches);

/* Per-frame post-processing that follows extraction in data::frame's constructors (data/frame.cc:37-51,75-92) for the
 * perspective camera model: undistort_keypoints (camera/perspective.cc:98-130, cv::undistortPoints with K, (k1 k2 p1 p2
 * k3), P = K, 20 iterations / 1e-6), convert_keypoints_to_bearings (:132-141) and assign_keypoints_to_grid
 * (data/common.cc:97-118). The grid comes back as CSR over cell = cx * grid_rows + cy (cell_start has grid_cols *
 * grid_rows + 1 entries), keypoint indices ascending inside a cell = the reference's push_back order. */
typedef struct {
    double fx, fy, cx, cy;
    double dist[5]; /* k1 k2 p1 p2 k3, camera/perspective.h */
} lpslam_perspective;
int lpslam_frame_postprocess(const lpslam_perspective* model, const lpslam_camera* cam, uint32_t n,
                             const lpslam_keypoint* keypts, lpslam_keypoint* undist_keypts, double* bearings /*n x 3*/,
                             uint32_t* cell_start, uint32_t* cell_items /*n*/, uint32_t* n_in_grid);

/* ------------------------------------------------- image pre-processing ahead of the extractor (device kernels) ----- */
/* util::convert_to_grayscale (util/image_converter.cc:8-39): cv::cvtColor(RGB | BGR | RGBA | BGRA -> GRAY) on 8-bit images.
 * order: 0 = RGB(A), 1 = BGR(A). The *_dev forms take device pointers, work on n_images frames frame_stride bytes apart and
 * only enqueue on `stream` (a cudaStream_t, NULL = default stream): their output can be the extractor's own level-0
 * buffer (lpslam_orb_input_dev), so that a colour frame never returns to the host between decode and extraction. */
int lpslam_gray(const uint8_t* src, int width, int height, size_t pitch, int channels, int order, uint8_t* dst, size_t dst_pitch);
int lpslam_gray_dev(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride, int channels, int order,
                    uint32_t n_images, uint8_t* dst_dev, size_t dst_pitch, size_t dst_frame_stride, void* stream);
/* LpSlam::ImageProcessing::imadjust with stretchlim (src/Utils/ImageProcessing.h:258-370), the body of the "AdjustIntensity"
 * processor (src/Processor/AdjustIntensityProcessor.h:17-40, low_out -0.3, high_out 1.4) for 8UC1 images. auto_limits != 0
 * = the std::nullopt form (1 % / 99 % histogram limits); limits_out (nullable) receives the low / high input limits. */
int lpslam_imadjust(const uint8_t* src, int width, int height, size_t pitch, int auto_limits, double low_in, double high_in,
                    double low_out, double high_out, uint8_t* dst, size_t dst_pitch, double* limits_out);
size_t lpslam_imadjust_scratch_bytes(uint32_t n_images);
int lpslam_imadjust_dev(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride, uint32_t n_images,
                        int auto_limits, double low_in, double high_in, double low_out, double high_out, uint8_t* dst_dev,
                        size_t dst_pitch, size_t dst_frame_stride, void* scratch_dev, void* stream);
/* No counterpart in the reference: lpslam_imadjust_dev for hosts that may not allocate device memory while they run. That is
 * a camera callback on a real-time tracking thread (a cudaMalloc takes driver locks and can stall the whole device for an
 * unbounded time), and a plugin in a process whose own pool claimed the device memory at start-up (a later cudaMalloc there
 * fails). Such a caller can neither make the scratch_dev buffer of lpslam_imadjust_dev nor use lpslam_imadjust, which
 * allocates for it. Here the histograms and value maps live in a fixed array inside the library image: 80 KiB on every GPU
 * that loads the library, held until the process exits. The call allocates and frees nothing; it only enqueues on `stream`.
 *  - Same arithmetic, arguments and output as lpslam_imadjust_dev (without scratch_dev). limits_dev (nullable, device,
 *    2 x n_images doubles) receives low / high input limits per image.
 *  - n_images <= LPSLAM_IMADJUST_STATIC_MAX_IMAGES, else LPSLAM_ERR_ARG with the limit named in lpslam_cuda_last_error().
 *  - It works on the device that is current in the calling thread. The first call on each device looks the array's device
 *    address up (once, kept afterwards): make that first call during start-up, before any stream capture begins.
 *  - All calls on a device share the one array and it keeps what the last call left; the call clears the histogram part it is
 *    about to use itself. Two calls that overlap on the GPU (different streams without an ordering between them, or two host
 *    threads) race on it and produce wrong images: order them, or give all but one of them lpslam_imadjust_dev. */
#define LPSLAM_IMADJUST_STATIC_MAX_IMAGES 64
int lpslam_imadjust_dev_static(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride, uint32_t n_images,
                               int auto_limits, double low_in, double high_in, double low_out, double high_out, uint8_t* dst_dev,
                               size_t dst_pitch, size_t dst_frame_stride, double* limits_dev, void* stream);
/* LpSlam::ImageProcessing::Undistort::undistort (src/Utils/ImageProcessing.h:245-249): cv::remap(INTER_LINEAR,
 * BORDER_CONSTANT 0) with the cached CV_32FC1 maps of cv::initUndistortRectifyMap (map_pitch in floats; the maps have the
 * destination size and are shared by all n_images frames). */
int lpslam_remap_linear(const uint8_t* src, int src_width, int src_height, size_t src_pitch, const float* map_x, const float* map_y,
                        size_t map_pitch, uint8_t* dst, int dst_width, int dst_height, size_t dst_pitch);
int lpslam_remap_linear_dev(const uint8_t* src_dev, int src_width, int src_height, size_t src_pitch, size_t src_frame_stride,
                            const float* map_x_dev, const float* map_y_dev, size_t map_pitch, uint8_t* dst_dev, int dst_width,
                            int dst_height, size_t dst_pitch, size_t dst_frame_stride, uint32_t n_images, void* stream);

/* data::landmark::compute_descriptor (data/landmark.cc:127-183), batched: the observations' descriptors of landmark l
 * are rows [offsets[l], offsets[l + 1]) of obs_desc (keyframes that will be erased already left out, :147-149).
 * best_idx[l] = index inside the landmark's own list of the observation whose median Hamming distance to the others is
 * smallest (first one on ties; 0 for an empty list); desc_out (nullable, n_landmarks x 32) receives that descriptor
 * (zeros for an empty list). Feeds the kNN database from a map (lpslam_knn_add). */
int lpslam_landmark_descriptors(const uint8_t* obs_desc, const uint32_t* offsets /*n_landmarks + 1*/, uint32_t n_landmarks,
                                uint32_t* best_idx, uint8_t* desc_out);

/* The reprojection matchers of match/projection.cc share one inner loop: every row (a landmark with its reprojection into
 * the frame, its predicted scale level and its descriptor) searches the frame's grid inside margin * scale_factors[level]
 * for the free keypoint with the smallest Hamming distance. lpslam_match_reprojected is that loop with the knobs in which
 * the reference's functions differ; the geometry in front of it (reproject_to_image, the valid-distance gate,
 * predict_scale_level: a few flops per landmark in Eigen) stays in the caller, so it is the reference's own code:
 *   match_frame_and_landmarks  (projection.cc:11-83)    level window (-1, 0), thr 100, ratio on,  stereo gate on  = lpslam_match_landmarks
 *   match_frame_and_keyframe   (projection.cc:210-312)  level window (-1,+1), thr = hamm_dist_thr, ratio off, gate off,
 *                                                       orientation check against the keyframe keypoints' angles (lm_angle)
 * lm_angle != NULL runs the angle_checker over (lm_angle[row] - keypoint angle) and clears the rejected keypoints. */
typedef struct {
    float margin;
    float lowe_ratio;
    uint32_t hamm_dist_thr;
    int32_t level_lo, level_hi; /* octave window = [level + level_lo, level + level_hi] */
    int32_t use_ratio;          /* Lowe ratio between best and second best on the same scale level (projection.cc:69-72) */
    int32_t use_x_right;        /* stereo reprojection gate (projection.cc:45-50); off: lm_x_right / kp_x_right may be NULL */
} lpslam_reproj_match_params;
int lpslam_match_reprojected(const lpslam_camera* cam, const lpslam_reproj_match_params* params, uint32_t n_lm,
                             const uint8_t* lm_active, const double* lm_reproj /*n_lm x 2*/, const float* lm_x_right,
                             const int32_t* lm_scale_level, const uint8_t* lm_desc, const uint8_t* lm_has_observation,
                             const float* lm_angle, uint32_t n_kp, const lpslam_keypoint* kp_undist, const uint8_t* kp_desc,
                             const float* kp_x_right, const uint8_t* kp_has_landmark, const float* scale_factors,
                             int32_t* kp_match_lm, uint32_t* n_matches);

/* The search loop shared by match::fuse::detect_duplication (match/fuse.cc:12-120, loop closing: landmarks of the loop
 * candidate's covisibilities into a keyframe under a Sim3) and fuse::replace_duplication (match/fuse.cc:122-266, local
 * mapping: fusing a keyframe's landmarks into its neighbours). The caller keeps the per-landmark geometry (will_be_erased,
 * reproject_to_image, the valid-distance and viewing-angle gates, predict_scale_level: fuse.cc:28-66 / :134-172) and
 * passes every surviving landmark as a row; every row independently takes the keypoint with the smallest Hamming distance
 * (strict '<', first wins) inside margin * scale_factors[level] whose octave is level - 1 or level. best_idx[row] = that
 * keypoint when its distance is <= HAMMING_DIST_THR_LOW (50), else -1; best_dist (nullable) = the distance (256 when the
 * row had no candidate). The association step behind it (duplicate / add_observation, fuse.cc:103-116 / :240-261) walks
 * best_idx in row order on the caller's landmark objects.
 *   chi_square_gate != 0  replace_duplication's reprojection-error gate (fuse.cc:199-222: 3 degrees of freedom for
 *                         keypoints with kp_x_right >= 0, 2 otherwise, against inv_level_sigma_sq[octave])
 *   unsigned_levels != 0  replace_duplication compares unsigned levels, so `octave < level - 1` wraps for level 0 and such
 *                         rows never match (fuse.cc:193); detect_duplication uses int (fuse.cc:60,84) */
int lpslam_match_fuse(const lpslam_camera* cam, float margin, int chi_square_gate, int unsigned_levels, uint32_t n_lm,
                      const uint8_t* lm_active, const double* lm_reproj /*n_lm x 2*/, const float* lm_x_right,
                      const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp, const lpslam_keypoint* kp_undist,
                      const uint8_t* kp_desc, const float* kp_x_right, const float* scale_factors,
                      const float* inv_level_sigma_sq, int32_t* best_idx, uint32_t* best_dist, uint32_t* n_fused);

/* The search loop of match::projection::match_by_Sim3_transform (match/projection.cc:317-413; loop closing: the landmarks
 * seen from the loop candidates reprojected into the current keyframe under the estimated Sim3). The caller keeps the
 * per-landmark geometry (will_be_erased, already_matched, reproject_to_image, the valid-distance and viewing-angle gates,
 * predict_scale_level: :333-370) and passes the surviving landmarks as rows, in the order of `landmarks`. Every row takes
 * the keypoint without a matched landmark (kp_has_matched_lm = matched_lms_in_keyfrm.at(idx) != nullptr, updated as rows
 * match, :381-383 / :409) with the smallest distance (strict '<') inside margin * scale_factors[level], octave in
 * [level - 1, level] compared as UNSIGNED (:385-390: a predicted level of 0 rejects every candidate), accepted when
 * best <= HAMMING_DIST_THR_LOW (:405). kp_match_lm[i] = row now matched to keypoint i, or -1. */
int lpslam_match_sim3(const lpslam_camera* cam, float margin, uint32_t n_lm, const uint8_t* lm_active,
                      const double* lm_reproj /*n_lm x 2*/, const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp,
                      const lpslam_keypoint* kp_undist, const uint8_t* kp_desc, const uint8_t* kp_has_matched_lm,
                      const float* scale_factors, int32_t* kp_match_lm, uint32_t* n_matches);

/* The two search loops and the cross-check of match::projection::match_keyframes_mutually (match/projection.cc:415-636; loop
 * closing, after the Sim3 solver). Rows of side 1 = keyframe 1's keypoints (n_1 = landmarks_1.size()): lm1_active[i] = the
 * keypoint holds a live landmark that is not matched already and passed the caller's geometry (:462-495), lm1_reproj_in_2 /
 * lm1_level_in_2 = its reprojection into keyframe 2 and predicted level there, lm1_desc = the landmark's descriptor;
 * side 2 likewise into keyframe 1. Every row independently takes the closest keypoint of the other keyframe (strict '<',
 * octave in [level - 1, level] as unsigned, best <= HAMMING_DIST_THR_HIGH); matched_idx_2_in_1[i] = the keyframe-2 keypoint
 * when both directions agree (:618-631: matched_lms_in_keyfrm_1.at(i) = landmarks_2.at(that)), else -1. */
int lpslam_match_mutual(const lpslam_camera* cam_1, const lpslam_camera* cam_2, float margin, uint32_t n_1,
                        const uint8_t* lm1_active, const double* lm1_reproj_in_2, const int32_t* lm1_level_in_2,
                        const uint8_t* lm1_desc, const lpslam_keypoint* kp1_undist, const uint8_t* kp1_desc,
                        const float* scale_factors_1, uint32_t n_2, const uint8_t* lm2_active, const double* lm2_reproj_in_1,
                        const int32_t* lm2_level_in_1, const uint8_t* lm2_desc, const lpslam_keypoint* kp2_undist,
                        const uint8_t* kp2_desc, const float* scale_factors_2, int32_t* matched_idx_2_in_1, uint32_t* n_matches);

/* The search loop of match::robust::match_for_triangulation (match/robust.cc:18-178; local mapping, before new landmarks
 * are triangulated). The caller walks the two BoW feature vectors and emits, in iteration order, one row per keyframe-1
 * keypoint WITHOUT a landmark (row_idx_1, :64-69) with the keyframe-2 keypoints of the same vocabulary node as its
 * candidate list. A candidate counts when it holds no landmark (kp2_has_lm), is not matched yet, its distance is
 * <= HAMMING_DIST_THR_LOW, it is not within 3 degrees of the epipole (only when neither keypoint has a stereo match,
 * :112-123) and passes check_epipolar_constraint (:323-341, E_12 row-major, 0.2 degree x scale_factors_1[octave]); among
 * those the smallest distance wins and an EQUAL distance replaces the best (:108 `best_hamm_dist < hamm_dist`).
 * epiplane_in_keyfrm_2 = reproject_to_bearing of keyframe 1's centre (:26-30). check_orientation runs the angle_checker
 * (:143-147, :164-170). matched_idx_2_in_1[i] (n_1) = keyframe-2 keypoint or -1 (the pairs of :172-180 in order of i). */
int lpslam_match_triangulation(uint32_t n_rows, const uint32_t* row_idx_1, const uint32_t* row_cand_begin,
                               const uint32_t* row_cand_end, const uint32_t* cand_items, uint32_t n_items, uint32_t n_1,
                               const lpslam_keypoint* kp1_undist, const uint8_t* kp1_desc, const double* bearings_1 /*n_1 x 3*/,
                               const float* x_right_1, const float* scale_factors_1, uint32_t n_2,
                               const lpslam_keypoint* kp2_undist, const uint8_t* kp2_desc, const double* bearings_2,
                               const float* x_right_2, const uint8_t* kp2_has_lm, const double* E_12 /*9*/,
                               const double* epiplane_in_keyfrm_2 /*3*/, int check_orientation, int32_t* matched_idx_2_in_1,
                               uint32_t* n_matches);

/* The search loop of match::bow_tree::match_frame_and_keyframe (match/bow_tree.cc:16-126), the matcher of BoW-based
 * tracking and relocalisation: the caller walks the two BoW feature vectors (the vocabulary library's containers) and
 * emits, in iteration order, one row per keyframe keypoint that holds a live landmark (bow_tree.cc:45-57) with its
 * descriptor and the frame keypoints of the same vocabulary node as its candidate list (cand_items[begin .. end)). Every
 * row takes the free candidate with the smallest distance (strict '<', first wins), accepted when best <= hamm_dist_thr
 * (HAMMING_DIST_THR_LOW = 50 there) and, with use_ratio, lowe_ratio * second_best >= best; a matched keypoint is not
 * available to later rows. row_angle != NULL runs the orientation check (bow_tree.cc:91-95,111-117).
 * kp_match_row[i] = row matched to frame keypoint i, or -1; kp_taken (nullable) marks keypoints occupied beforehand. */
int lpslam_match_candidates(uint32_t n_rows, const uint8_t* row_desc, const uint32_t* row_cand_begin,
                            const uint32_t* row_cand_end, const uint32_t* cand_items, uint32_t n_items,
                            const float* row_angle, uint32_t n_kp, const uint8_t* kp_desc, const float* kp_angle,
                            const uint8_t* kp_taken, uint32_t hamm_dist_thr, float lowe_ratio, int use_ratio,
                            int32_t* kp_match_row, uint32_t* n_matches);

/* match::area::match_in_consistent_area (match/area.cc:8-112), the matcher of the monocular initializer
 * (initialize/... via module/initializer.cc): frame-1 keypoints of scale level 0 search frame 2 inside a window of
 * +-margin around prev_matched_pts (n1 x 2 floats, updated in place with the matched frame-2 positions, area.cc:104-108).
 * matched_indices_2_in_frm_1[i] = frame-2 keypoint matched to frame-1 keypoint i, or -1. */
int lpslam_match_area(const lpslam_camera* cam, uint32_t n1, const lpslam_keypoint* kp1_undist, const uint8_t* desc1,
                      float* prev_matched_pts, uint32_t n2, const lpslam_keypoint* kp2_undist, const uint8_t* desc2,
                      int margin, float lowe_ratio, int check_orientation, int32_t* matched_indices_2_in_frm_1,
                      uint32_t* n_matches);

/* ---------------------------------------------------------------- kNN database (relocalisation / loop query) ----- */
/* Brute-force Hamming top-2 of query descriptors against a device-resident descriptor database (generalises the
 * best / second-best scan of match/robust.cc:252-274 and the BoW-candidate search of module/relocalizer.cc:31-46).
 * A database is one shard; index_base is the global row index of its first row. */
typedef struct lpslam_knn_db lpslam_knn_db;
int lpslam_knn_create(lpslam_knn_db** db, int device, uint64_t capacity_rows, uint64_t index_base);
void lpslam_knn_destroy(lpslam_knn_db* db);
int lpslam_knn_add(lpslam_knn_db* db, const uint8_t* desc_host, uint64_t n_rows);      /* append rows (H2D) */
int lpslam_knn_add_dev(lpslam_knn_db* db, const uint8_t* desc_dev, uint64_t n_rows);   /* append rows (D2D) */
int lpslam_knn_size(const lpslam_knn_db* db, uint64_t* n_rows);
/* host query: best_idx (global row index, -1 if the shard is empty), best_dist, second_dist; ties -> lowest index,
 * exactly what a sequential strict-< scan gives. */
int lpslam_knn_query(lpslam_knn_db* db, const uint8_t* q_host, uint32_t nq, int64_t* best_idx, uint32_t* best_dist,
                     uint32_t* second_dist);
/* device query: writes per query two packed keys (dist << 40 | global_index), best then second, into keys_dev[nq*2];
 * empty slots hold LPSLAM_KNN_EMPTY_KEY. Enqueues on `stream` (a cudaStream_t, NULL = default stream). */
#define LPSLAM_KNN_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
int lpslam_knn_query_dev(lpslam_knn_db* db, const uint8_t* q_dev, uint32_t nq, uint64_t* keys_dev, void* stream);
/* merge of n_shards per-shard key blocks (gathered over NCCL): gathered_dev[n_shards][nq][2] -> merged_dev[nq][2] */
int lpslam_knn_merge_dev(const uint64_t* gathered_dev, uint32_t n_shards, uint32_t nq, uint64_t* merged_dev, void* stream);
/* fused exchange + merge over NVLink peer memory: every rank passes the device pointers of ALL ranks' key blocks
 * (peer-mapped); the kernel reads the peers' blocks directly and merges. */
int lpslam_knn_merge_peer_dev(const uint64_t* const* shard_keys_dev, uint32_t n_shards, uint32_t nq, uint64_t* merged_dev,
                              void* stream);
/* the whole sharded query below the C ABI (BASELINE configs[4]): every rank of the NCCL communicator calls this with the same
 * queries; per-shard top-2 on the rank's own shard, ncclAllGather of the (nq, 2) keys into gathered_dev[world][nq][2], merge
 * into merged_dev[nq][2] (the same on every rank). nccl_comm is the caller's ncclComm_t; ncclAllGather is resolved at run time
 * from the libnccl the process has loaded, so liblpslam_cuda links against no particular NCCL. Enqueues on `stream`. */
int lpslam_knn_sharded_query_dev(lpslam_knn_db* db, void* nccl_comm, uint32_t world, const uint8_t* q_dev, uint32_t nq,
                                 uint64_t* gathered_dev, uint64_t* merged_dev, void* stream);
/* unpack merged keys on the device into the host-facing triplets */
int lpslam_knn_unpack(const uint64_t* merged_dev, uint32_t nq, int64_t* best_idx, uint32_t* best_dist,
                      uint32_t* second_dist, void* stream);

/* match/base.h:18-63 compute_descriptor_distance_32 / _64 on the device, n pairs (test hook + landmark.cc:153-183) */
int lpslam_hamming_pairs(const uint8_t* a, const uint8_t* b, uint32_t n, uint32_t* dist);

#ifdef __cplusplus
}
#endif
#endif /* LPSLAM_CUDA_H */
