// Extractor context: host-side state behind lpslam_orb_ctx (one per reference orb_extractor instance).
#pragma once

#include "common.cuh"

#include <array>

struct lpslam_orb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    // side stream: level-0 apron + blur run beside FAST and the quadtree (they only depend on the pyramid)
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    // marks "everything enqueued on `stream` so far" for a reader on another stream (lpslam_match_stereo_ctx)
    cudaEvent_t ev_reader = nullptr;
    // the pyramid of a batch is built as up to 4 independent chains of frames on their own streams, so that the small
    // levels of one chain run beside the large levels of another (each level launch alone leaves most SMs idle at its tail)
    cudaStream_t chain[3] = {};
    cudaEvent_t ev_chain_fork = nullptr, ev_chain_join[3] = {};
    int sub_batches = 1;  // FAST + quadtree sub-batches on own streams (LPSLAM_SUBBATCH); measured on B200: 2 -> no gain, 3-4 -> 8-10 % slower
    int pyr_chains = 1;  // measured on B200: 2 chains 0.132 ms vs 0.135 ms, 3-4 chains slower (LPSLAM_PYR_CHAINS)
    uint32_t max_batch = 1;

    // feature/orb_params.h
    uint32_t max_kpts = 2000, ini_max_kpts = 2000, num_levels = 8, ini_thr = 20, min_thr = 7;
    float scale_factor = 1.2f;
    std::vector<std::array<float, 4>> mask_rects;
    // feature/orb_params.cc:38-72, orb_extractor.cc:208-226
    std::vector<float> scale, inv_scale, sigma_sq, inv_sigma_sq;
    std::vector<unsigned> quota;

    // geometry of the configured image size
    int w = 0, h = 0;
    lps::LevelGeom lv[lps::kMaxLevels];
    lps::LevelGeom* lv_dev = nullptr;
    lps::PyramidPlan* pyr_plan = nullptr;  // single-launch TMA pyramid of this geometry (nullptr: per-level k_resize_walk chain)
    lps::BlurPlan* blur_plan = nullptr;  // TMA-staged blur of this geometry (nullptr: LPS_BLUR_V1 forces the ld.global version)
    std::vector<lps::CellDesc> cells;
    lps::CellDesc* cells_dev = nullptr;
    int num_cells = 0;
    size_t cand_stride = 0, node_stride = 0, sel_stride = 0;
    uint32_t out_cap = 0;  // packed output rows available on the device

    // device arena (one allocation per geometry)
    uint8_t* arena = nullptr;
    size_t arena_bytes = 0;
    uint32_t* cell_counts = nullptr;
    uint32_t* cell_slots = nullptr;
    lps::TreeArgs tree{};
    lps::DescribeArgs desc{};
    lps::StereoArgs stereo{};
    uint8_t* mask_dev = nullptr;      // user image mask (uploaded per call)
    uint8_t* rect_mask_dev = nullptr; // rectangle mask, drawn once (orb_extractor.cc:85-88,251-264)
    bool rect_mask_ready = false;
    int* tables_dev = nullptr;

    // pinned host staging
    uint32_t* h_frame_off = nullptr;  // max_batch + 2 (last = overflow flag)
    // lpslam_orb_pyramid_levels: packed levels on the device and their page-locked landing buffer. Owned by the context, not the
    // calling thread: data::frame runs every extract on a thread of its own (frame.cc:60-63), thread-local buffers would be
    // allocated and freed per frame
    // lpslam_stereo_frames with one pair: the whole per-frame pipeline (pyramid .. stereo matcher + the offsets read-back) as one
    // CUDA graph. The first call of a (geometry, parameters, stereo constants) setting runs launch by launch (tables and masks are
    // built on first use), the second is captured, later ones relaunch the graph. LPSLAM_CUDA_GRAPH=0 turns it off.
    cudaGraphExec_t pair_graph = nullptr;
    float pair_graph_fxb = 0.f, pair_graph_tb = 0.f;
    int pair_graph_warm = 0;
    bool pair_graph_off = false;
    uint64_t pair_graph_launches = 0;  // cudaGraphLaunch calls over the context's life (LPSLAM_ORB_DEBUG_GRAPH_LAUNCHES)
    void drop_pair_graph();
    uint8_t* pack_dev = nullptr;
    uint8_t* pack_host = nullptr;
    size_t pack_cap = 0;
    uint32_t last_batch = 0;
    bool has_result = false;

    // asynchronous stereo frames (lpslam_stereo_frames_submit / _collect): two batches in flight on one context. A slot owns a
    // device landing buffer for the next batch's images (filled by the copy-in stream while the previous batch computes) and a
    // copy of a finished batch's packed results (read back by the copy-out stream while the next batch computes).
    struct AsyncSlot {
        uint8_t* landing = nullptr;
        lpslam_keypoint* kp = nullptr;
        uint8_t* desc = nullptr;
        float* xr = nullptr;
        float* dep = nullptr;
        uint32_t* off_dev = nullptr;   // max_batch + 2 (last = overflow flag)
        uint32_t* h_off = nullptr;     // pinned mirror
        cudaEvent_t ev_in = nullptr, ev_landing_free = nullptr, ev_done = nullptr;
        bool busy = false, landing_used = false;
        uint32_t n_pairs = 0, n_frames = 0;
    };
    AsyncSlot slot[2];
    size_t landing_bytes = 0;
    bool async_ready = false;
    cudaStream_t copy_in = nullptr, copy_out = nullptr;
    uint64_t seq = 0;
    void release_async();

    // optional stage timing
    bool timing = false;
    cudaEvent_t ev[8] = {};
    float stage_ms[7] = {};

    void initialize_tables();
    int configure(int width, int height);
    void release_geometry();
    int run_pipeline(uint32_t n_frames, const uint8_t* mask_dev_ptr, size_t mask_pitch);
};
