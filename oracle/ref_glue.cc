// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// C glue around the reference's OWN sources (feature/orb_extractor.cc, orb_extractor_node.cc, orb_params.cc,
// match/stereo.cc), compiled unmodified from /root/reference against oracle/shim. Exports the same entry points as the
// port (orb_oracle.h) under the prefix ref_, so that tests can hold the port against the reference line for line.
//
// Tie-break pin: the reference sorts std::pair<int, orb_extractor_node*> (orb_extractor.cc:408,455), i.e. equal keypoint
// counts are ordered by HEAP ADDRESS of the std::list nodes. The operator new below hands out list nodes of exactly that
// size from monotonic chunks (rewound between calls only when empty), so within one call a later-created node always has
// the higher address: the documented pin "later created first" that both the port and the CUDA kernel implement.
#include <algorithm>
#include <array>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <list>
#include <new>
#include <vector>

#include <opencv2/core.hpp>

#define private public
#include "openvslam/feature/orb_extractor.h"
#undef private
#include "openvslam/match/stereo.h"

#include "orb_oracle.h"

namespace {

constexpr size_t kNodeSize = sizeof(std::_List_node<openvslam::feature::orb_extractor_node>);
constexpr size_t kChunkBytes = size_t(64) << 20;
constexpr size_t kLowWater = size_t(8) << 20;

// Monotonic chunks. Anything of exactly kNodeSize bytes lands here (list nodes, but also e.g. a 3-element
// vector<vector<KeyPoint>>), so a chunk is only rewound when nothing allocated from it is still alive; otherwise the
// offset keeps growing and a fresh chunk is started when the current one runs low (the old one stays valid).
struct Chunk {
    char* base;
    size_t off;
    long live;
};
Chunk g_chunks[64];
int g_nchunks = 0;

inline Chunk* chunk_of(void* p) {
    for (int i = g_nchunks - 1; i >= 0; --i)
        if (p >= g_chunks[i].base && p < g_chunks[i].base + kChunkBytes) return &g_chunks[i];
    return nullptr;
}

// called at the start of every reference call that builds a quadtree
void arena_begin_call() {
    if (g_nchunks == 0) return;
    Chunk& c = g_chunks[g_nchunks - 1];
    if (c.live == 0) c.off = 0;
    else if (kChunkBytes - c.off < kLowWater && g_nchunks < 64) {
        char* b = static_cast<char*>(std::malloc(kChunkBytes));
        if (b) g_chunks[g_nchunks++] = Chunk{b, 0, 0};
    }
}

}  // namespace

void* operator new(size_t n) {
    if (n == kNodeSize) {
        if (g_nchunks == 0) {
            char* b = static_cast<char*>(std::malloc(kChunkBytes));
            if (b) g_chunks[g_nchunks++] = Chunk{b, 0, 0};
        }
        const size_t need = (n + 15) & ~size_t(15);
        if (g_nchunks) {
            Chunk& c = g_chunks[g_nchunks - 1];
            if (c.off + need <= kChunkBytes) {
                void* p = c.base + c.off;
                c.off += need;
                ++c.live;
                return p;
            }
        }
    }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void operator delete(void* p) noexcept {
    if (!p) return;
    if (Chunk* c = chunk_of(p)) {
        --c->live;
        return;
    }
    std::free(p);
}
void operator delete(void* p, size_t) noexcept { operator delete(p); }

using openvslam::feature::orb_extractor;

struct ref_extractor {
    orb_extractor* ex = nullptr;
    cv::Mat image, mask;  // copies of the last inputs
    bool have_sel = false;
    std::vector<std::vector<cv::KeyPoint>> sel;
};

extern "C" {

ref_extractor* ref_create(uint32_t max_kpts, float scale_factor, uint32_t num_levels, uint32_t ini_fast_thr,
                          uint32_t min_fast_thr, const float* mask_rects, uint32_t num_mask_rects) {
    std::vector<std::vector<float>> rects;
    for (uint32_t i = 0; i < num_mask_rects; ++i) rects.push_back({mask_rects[4 * i], mask_rects[4 * i + 1], mask_rects[4 * i + 2], mask_rects[4 * i + 3]});
    try {
        auto* e = new ref_extractor;
        e->ex = new orb_extractor(max_kpts, 2 * max_kpts, scale_factor, num_levels, ini_fast_thr, min_fast_thr, rects);
        return e;
    }
    catch (const std::exception&) {
        return nullptr;
    }
}
void ref_destroy(ref_extractor* e) {
    if (!e) return;
    delete e->ex;
    delete e;
}
void ref_set_max_kpts(ref_extractor* e, uint32_t n) { e->ex->set_max_num_keypoints(n); }

int ref_get_table(const ref_extractor* e, int which, float* out, int cap) {
    const int n = static_cast<int>(e->ex->get_num_scale_levels());
    if (cap < n) return -2;
    std::vector<float> t;
    switch (which) {
        case 0: t = e->ex->get_scale_factors(); break;
        case 1: t = e->ex->get_inv_scale_factors(); break;
        case 2: t = e->ex->get_level_sigma_sq(); break;
        case 3: t = e->ex->get_inv_level_sigma_sq(); break;
        case 4:
            for (unsigned v : e->ex->num_keypts_per_level_) t.push_back(static_cast<float>(v));
            break;
        default: return -1;
    }
    std::copy(t.begin(), t.end(), out);
    return n;
}

int ref_extract(ref_extractor* e, const uint8_t* img, int w, int h, size_t pitch, const uint8_t* mask, size_t mask_pitch,
                orc_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* n_out) {
    *n_out = 0;
    arena_begin_call();
    e->have_sel = false;
    if (!img || w <= 0 || h <= 0) return 0;
    e->image = cv::Mat(h, w, CV_8UC1, const_cast<uint8_t*>(img), pitch).clone();
    e->mask = mask ? cv::Mat(h, w, CV_8UC1, const_cast<uint8_t*>(mask), mask_pitch).clone() : cv::Mat();
    std::vector<cv::KeyPoint> kps;
    cv::Mat desc;
    e->ex->extract(e->image, e->mask, kps, desc);
    *n_out = static_cast<uint32_t>(kps.size());
    if (kps.size() > cap) return -2;
    for (size_t i = 0; i < kps.size(); ++i) {
        std::memcpy(&kp_out[i], &kps[i], sizeof(orc_keypoint));
        std::memcpy(desc_out + i * 32, desc.ptr(static_cast<int>(i)), 32);
    }
    return 0;
}

int ref_level_size(const ref_extractor* e, int level, int* w, int* h) {
    if (level < 0 || level >= static_cast<int>(e->ex->image_pyramid_.size())) return -1;
    *w = e->ex->image_pyramid_[level].cols;
    *h = e->ex->image_pyramid_[level].rows;
    return 0;
}
int ref_level_copy(const ref_extractor* e, int level, int blurred, uint8_t* dst, size_t pitch) {
    if (level < 0 || level >= static_cast<int>(e->ex->image_pyramid_.size())) return -1;
    if (blurred) return -3;  // a local of extract() in the reference, never exposed
    const cv::Mat& m = e->ex->image_pyramid_[level];
    for (int y = 0; y < m.rows; ++y) std::memcpy(dst + static_cast<size_t>(y) * pitch, m.ptr(y), m.cols);
    return 0;
}
int ref_level_candidates(const ref_extractor*, int, int32_t*, int) { return -1; }  // internal to compute_fast_keypoints

int ref_level_selected(ref_extractor* e, int level, int32_t* xys, int cap) {
    if (level < 0 || level >= static_cast<int>(e->ex->get_num_scale_levels())) return -1;
    if (!e->have_sel) {
        // re-run the private FAST + quadtree step on the pyramid of the last extract (same mask priority as :93-107)
        arena_begin_call();
        e->sel.clear();
        if (!e->mask.empty()) e->ex->compute_fast_keypoints(e->sel, e->mask);
        else if (!e->ex->rect_mask_.empty()) e->ex->compute_fast_keypoints(e->sel, e->ex->rect_mask_);
        else e->ex->compute_fast_keypoints(e->sel, cv::Mat());
        e->have_sel = true;
    }
    const auto& v = e->sel[level];
    const int n = static_cast<int>(v.size());
    for (int i = 0; i < std::min(n, cap); ++i) {
        xys[3 * i] = static_cast<int>(v[i].pt.x) - 19;  // back to border-relative coordinates (orb_extractor.cc:385-388)
        xys[3 * i + 1] = static_cast<int>(v[i].pt.y) - 19;
        xys[3 * i + 2] = static_cast<int>(v[i].response);
    }
    return n;
}

int ref_match_stereo(const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* lw, const int* lh,
                     const size_t* lpitch_l, const size_t* lpitch_r, int num_levels, const orc_keypoint* kp_l, uint32_t nl,
                     const orc_keypoint* kp_r, uint32_t nr, const uint8_t* desc_l, const uint8_t* desc_r,
                     const float* scale_factors, const float* inv_scale_factors, float focal_x_baseline, float true_baseline,
                     float* stereo_x_right, float* depths) {
    std::vector<cv::Mat> pl, pr;
    for (int l = 0; l < num_levels; ++l) {
        pl.emplace_back(lh[l], lw[l], CV_8UC1, const_cast<uint8_t*>(pyr_l[l]), lpitch_l[l]);
        pr.emplace_back(lh[l], lw[l], CV_8UC1, const_cast<uint8_t*>(pyr_r[l]), lpitch_r[l]);
    }
    std::vector<cv::KeyPoint> kl(nl), kr(nr);
    if (nl) std::memcpy(static_cast<void*>(kl.data()), kp_l, sizeof(orc_keypoint) * nl);
    if (nr) std::memcpy(static_cast<void*>(kr.data()), kp_r, sizeof(orc_keypoint) * nr);
    cv::Mat dl(static_cast<int>(nl), 32, CV_8UC1, const_cast<uint8_t*>(desc_l), 32);
    cv::Mat dr(static_cast<int>(nr), 32, CV_8UC1, const_cast<uint8_t*>(desc_r), 32);
    const std::vector<float> sf(scale_factors, scale_factors + num_levels), isf(inv_scale_factors, inv_scale_factors + num_levels);
    openvslam::match::stereo m(pl, pr, kl, kr, dl, dr, sf, isf, focal_x_baseline, true_baseline);
    std::vector<float> xr, dep;
    m.compute(xr, dep);
    int kept = 0;
    for (uint32_t i = 0; i < nl; ++i) {
        stereo_x_right[i] = xr[i];
        depths[i] = dep[i];
        kept += dep[i] != -1.0f;
    }
    return kept;
}

}  // extern "C"
