// TEST INFRASTRUCTURE — NOT PRODUCT CODE. See orb_oracle.h.
// Restatement of the reference's ORB extraction over flat arrays. Every function cites the reference lines it follows
// (paths relative to /root/reference/external/openvslam/src/openvslam/).
// Build: g++ -O3 -ffp-contract=off, no -march=native (reference default: /root/reference/CMakeLists.txt:155).
#include "orb_oracle.h"
#include "cvprim.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>
#include <list>
#include <vector>

namespace {

constexpr int kPatch = 31;      // feature/orb_extractor.h:128 fast_patch_size_
constexpr int kHalfPatch = 15;  // feature/orb_extractor.h:130
constexpr int kBorder = 19;     // feature/orb_extractor.h:133 orb_patch_radius_
constexpr int kCell = 64;       // feature/orb_extractor.cc:286
constexpr int kOverlap = 6;     // feature/orb_extractor.cc:285

const signed char kPattern[1024] = {
#include "rbrief_pattern.inc"
};

struct Cand {
    int x, y, score; // border-relative integer coordinates, FAST score
};

// ---- util/trigonometric.h:11-46 ----
constexpr float PI_F = 3.14159265358979f;
constexpr float PI_2_F = PI_F / 2.0f;
constexpr float TWO_PI_F = 2.0f * PI_F;
constexpr float INV_TWO_PI_F = 1.0f / TWO_PI_F;
constexpr float THREE_PI_2_F = 3.0f * PI_2_F;

float poly_cos(float v) {
    const float v2 = v * v;
    return 0.99940307f + v2 * (-0.49558072f + 0.03679168f * v2);
}
float ref_cos(float v) {
    v = v - cvprim::floor_int(v * INV_TWO_PI_F) * TWO_PI_F;
    v = (0.0f < v) ? v : -v;
    if (v < PI_2_F) return poly_cos(v);
    if (v < PI_F) return -poly_cos(PI_F - v);
    if (v < THREE_PI_2_F) return -poly_cos(v - PI_F);
    return poly_cos(TWO_PI_F - v);
}
float ref_sin(float v) { return ref_cos(PI_2_F - v); }

struct Image {
    int w = 0, h = 0;
    std::vector<uint8_t> px;
    uint8_t* row(int y) { return px.data() + static_cast<size_t>(y) * w; }
    const uint8_t* row(int y) const { return px.data() + static_cast<size_t>(y) * w; }
    void alloc(int w_, int h_) {
        w = w_;
        h = h_;
        px.assign(static_cast<size_t>(w) * h, 0);
    }
};

} // namespace

struct orc_extractor {
    uint32_t max_kpts, num_levels, ini_thr, min_thr;
    float scale_factor;
    std::vector<std::array<float, 4>> mask_rects;
    std::vector<float> scale, inv_scale, sigma_sq, inv_sigma_sq;
    std::vector<unsigned> quota;
    std::vector<int> u_max;
    bool rect_mask_ready = false;
    Image rect_mask;
    // state of the last extract call
    std::vector<Image> pyr, blur;
    std::vector<std::vector<Cand>> cand, sel;

    void initialize();
};

// feature/orb_params.cc:38-72 and feature/orb_extractor.cc:208-242
void orc_extractor::initialize() {
    scale.assign(num_levels, 1.0f);
    inv_scale.assign(num_levels, 1.0f);
    sigma_sq.assign(num_levels, 1.0f);
    inv_sigma_sq.assign(num_levels, 1.0f);
    float s = 1.0f;
    for (uint32_t l = 1; l < num_levels; ++l) {
        scale[l] = scale_factor * scale[l - 1];
        inv_scale[l] = (1.0f / scale_factor) * inv_scale[l - 1];
        s = scale_factor * s;
        sigma_sq[l] = s * s;
        inv_sigma_sq[l] = 1.0f / (s * s);
    }
    quota.assign(num_levels, 0);
    double desired = max_kpts * (1.0 - 1.0 / scale_factor) / (1.0 - std::pow(1.0 / scale_factor, static_cast<double>(num_levels)));
    unsigned total = 0;
    for (uint32_t l = 0; l + 1 < num_levels; ++l) {
        quota[l] = static_cast<unsigned>(std::round(desired));
        total += quota[l];
        desired *= 1.0 / scale_factor;
    }
    quota[num_levels - 1] = static_cast<unsigned>(std::max(static_cast<int>(max_kpts) - static_cast<int>(total), 0));

    u_max.assign(kHalfPatch + 1, 0);
    const unsigned vmax = static_cast<unsigned>(std::floor(kHalfPatch * std::sqrt(2.0) / 2 + 1));
    const unsigned vmin = static_cast<unsigned>(std::ceil(kHalfPatch * std::sqrt(2.0) / 2));
    for (unsigned v = 0; v <= vmax; ++v) u_max[v] = static_cast<int>(std::round(std::sqrt(double(kHalfPatch * kHalfPatch) - double(v * v))));
    for (unsigned v = kHalfPatch, v0 = 0; vmin <= v; --v) {
        while (u_max[v0] == u_max[v0 + 1]) ++v0;
        u_max[v] = static_cast<int>(v0);
        ++v0;
    }
}

namespace {

// feature/orb_extractor.cc:281-283 — mask lookup with float scaling and truncation
bool in_mask(const uint8_t* mask, size_t mpitch, unsigned y, unsigned x, float sf) {
    const int my = static_cast<int>(y * sf);
    const int mx = static_cast<int>(x * sf);
    return mask[static_cast<size_t>(my) * mpitch + mx] == 0;
}

// ---- quadtree: feature/orb_extractor.cc:401-593, feature/orb_extractor_node.cc:6-50 ----
struct Node {
    std::vector<Cand> pts;
    int bx = 0, by = 0, ex = 0, ey = 0;
    bool leaf = false;
    long seq = 0; // creation order; stands in for the heap address used by the reference's pool sort
    std::list<Node>::iterator self;
};

std::array<Node, 4> split(const Node& n) {
    // orb_extractor_node.cc:8-9: half extents round up
    const int hx = cvprim::ceil_int((n.ex - n.bx) / 2.0);
    const int hy = cvprim::ceil_int((n.ey - n.by) / 2.0);
    std::array<Node, 4> c;
    c[0].bx = n.bx;      c[0].by = n.by;      c[0].ex = n.bx + hx; c[0].ey = n.by + hy;
    c[1].bx = n.bx + hx; c[1].by = n.by;      c[1].ex = n.ex;      c[1].ey = n.by + hy;
    c[2].bx = n.bx;      c[2].by = n.by + hy; c[2].ex = n.bx + hx; c[2].ey = n.ey;
    c[3].bx = n.bx + hx; c[3].by = n.by + hy; c[3].ex = n.ex;      c[3].ey = n.ey;
    for (const Cand& p : n.pts) {
        unsigned q = 0;
        if (n.bx + hx <= p.x) q += 1;
        if (n.by + hy <= p.y) q += 2;
        c[q].pts.push_back(p);
    }
    return c;
}

using Pool = std::vector<std::pair<int, Node*>>;

void adopt(std::array<Node, 4>& kids, std::list<Node>& nodes, Pool& pool, long& seq) {
    // orb_extractor.cc:555-569: non-empty children go to the FRONT; only children with >1 keypoint are poolable;
    // one-keypoint children are not flagged leaf.
    for (Node& k : kids) {
        if (k.pts.empty()) continue;
        k.seq = seq++;
        nodes.push_front(std::move(k));
        Node& placed = nodes.front();
        placed.self = nodes.begin();
        if (placed.pts.size() == 1) continue;
        pool.emplace_back(static_cast<int>(placed.pts.size()), &placed);
    }
}

std::vector<Cand> distribute(const std::vector<Cand>& in, int min_x, int max_x, int min_y, int max_y, unsigned quota) {
    // orb_extractor.cc:480-553 initial nodes
    const double ratio = static_cast<double>(max_x - min_x) / (max_y - min_y);
    double dx, dy;
    unsigned nx, ny;
    if (ratio > 1) {
        nx = static_cast<unsigned>(std::round(ratio));
        ny = 1;
        dx = static_cast<double>(max_x - min_x) / nx;
        dy = max_y - min_y;
    }
    else {
        nx = 1;
        ny = static_cast<unsigned>(std::round(1 / ratio));
        dx = max_x - min_y; // sic, orb_extractor.cc:500
        dy = static_cast<double>(max_y - min_y) / ny;
    }
    std::list<Node> nodes;
    std::vector<Node*> init(nx * ny);
    long seq = 0;
    for (unsigned i = 0; i < nx * ny; ++i) {
        const unsigned ix = i % nx, iy = i / nx;
        Node n;
        n.bx = static_cast<int>(dx * ix);
        n.by = static_cast<int>(dy * iy);
        n.ex = static_cast<int>(dx * (ix + 1));
        n.ey = static_cast<int>(dy * (iy + 1));
        n.seq = seq++;
        nodes.push_back(std::move(n));
        init[i] = &nodes.back();
    }
    for (const Cand& p : in) {
        const unsigned ix = static_cast<unsigned>(static_cast<float>(p.x) / dx);
        const unsigned iy = static_cast<unsigned>(static_cast<float>(p.y) / dy);
        init.at(ix + iy * nx)->pts.push_back(p);
    }
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->pts.empty()) {
            it = nodes.erase(it);
            continue;
        }
        it->leaf = (it->pts.size() == 1);
        ++it;
    }

    Pool pool;
    bool filled = false;
    // phase 1 (orb_extractor.cc:414-445): breadth passes
    while (true) {
        const size_t prev = nodes.size();
        pool.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->leaf) {
                ++it;
                continue;
            }
            auto kids = split(*it);
            adopt(kids, nodes, pool, seq);
            it = nodes.erase(it);
        }
        if (quota <= nodes.size() || nodes.size() == prev) {
            filled = true;
            break;
        }
        if (quota < nodes.size() + pool.size()) break;
    }
    // phase 2 (orb_extractor.cc:447-475): split the fullest nodes first; ties by creation order, later first
    while (!filled) {
        const size_t prev = nodes.size();
        Pool prev_pool = pool;
        pool.clear();
        std::sort(prev_pool.begin(), prev_pool.end(), [](const std::pair<int, Node*>& a, const std::pair<int, Node*>& b) {
            if (a.first != b.first) return a.first > b.first;
            return a.second->seq > b.second->seq;
        });
        for (const auto& e : prev_pool) {
            auto kids = split(*e.second);
            adopt(kids, nodes, pool, seq);
            nodes.erase(e.second->self);
            if (quota <= nodes.size()) {
                filled = true;
                break;
            }
        }
        if (filled || quota <= nodes.size() || nodes.size() == prev) break;
    }
    // orb_extractor.cc:571-593: first keypoint with the strictly greatest response per node, in list order
    std::vector<Cand> out;
    out.reserve(nodes.size());
    for (const Node& n : nodes) {
        Cand best = n.pts[0];
        for (size_t k = 1; k < n.pts.size(); ++k)
            if (n.pts[k].score > best.score) best = n.pts[k];
        out.push_back(best);
    }
    return out;
}

// feature/orb_extractor.cc:611-634
float ic_angle(const Image& im, int cx, int cy, const std::vector<int>& u_max) {
    int m01 = 0, m10 = 0;
    const uint8_t* c = im.row(cy) + cx;
    const int step = im.w;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = u_max[v];
        for (int u = -d; u <= d; ++u) {
            const int p = c[u + v * step], m = c[u - v * step];
            vsum += p - m;
            m10 += u * (p + m);
        }
        m01 += v * vsum;
    }
    return cvprim::fast_atan2(static_cast<float>(m01), static_cast<float>(m10));
}

// feature/orb_extractor.cc:644-705 (non-SSE branch)
void rbrief(const Image& im, int cx, int cy, float angle_deg, uint8_t* desc) {
    const float angle = static_cast<float>(angle_deg * M_PI / 180.0);
    const float ca = ref_cos(angle);
    const float sa = ref_sin(angle);
    const uint8_t* c = im.row(cy) + cx;
    const int step = im.w;
    auto sample = [&](int idx) {
        const float px = kPattern[idx], py = kPattern[idx + 1];
        const int r = cvprim::round_half_even(px * sa + py * ca);
        const int q = cvprim::round_half_even(px * ca - py * sa);
        return c[r * step + q];
    };
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int b = 0; b < 8; ++b) {
            const int base = (i * 8 + b) * 4;
            val |= (sample(base) < sample(base + 2)) << b;
        }
        desc[i] = static_cast<uint8_t>(val);
    }
}

} // namespace

extern "C" {

void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t spitch, uint8_t* dst, int dw, int dh, size_t dpitch) {
    cvprim::resize_linear_u8(src, sw, sh, spitch, dst, dw, dh, dpitch);
}
void orc_gaussian7_u8(const uint8_t* src, int w, int h, size_t spitch, uint8_t* dst, size_t dpitch) {
    cvprim::gaussian7_u8(src, w, h, spitch, dst, dpitch);
}
int orc_fast9_nms(const uint8_t* img, int w, int h, size_t pitch, int thr, int32_t* xys, int cap) {
    std::vector<cvprim::FastPoint> pts;
    cvprim::fast9_nms(img, w, h, pitch, thr, pts);
    const int n = static_cast<int>(pts.size());
    for (int i = 0; i < std::min(n, cap); ++i) {
        xys[3 * i] = pts[i].x;
        xys[3 * i + 1] = pts[i].y;
        xys[3 * i + 2] = pts[i].score;
    }
    return n;
}
float orc_fast_atan2(float y, float x) { return cvprim::fast_atan2(y, x); }
float orc_cos(float v) { return ref_cos(v); }
float orc_sin(float v) { return ref_sin(v); }

orc_extractor* orc_create(uint32_t max_kpts, float scale_factor, uint32_t num_levels, uint32_t ini_fast_thr,
                          uint32_t min_fast_thr, const float* mask_rects, uint32_t num_mask_rects) {
    if (num_levels == 0) return nullptr;
    // feature/orb_params.cc:16-26 rejects malformed rectangles
    for (uint32_t i = 0; i < num_mask_rects; ++i) {
        const float* r = mask_rects + 4 * i;
        if (r[0] >= r[1] || r[2] >= r[3]) return nullptr;
    }
    auto* e = new orc_extractor;
    e->max_kpts = max_kpts;
    e->scale_factor = scale_factor;
    e->num_levels = num_levels;
    e->ini_thr = ini_fast_thr;
    e->min_thr = min_fast_thr;
    for (uint32_t i = 0; i < num_mask_rects; ++i)
        e->mask_rects.push_back({mask_rects[4 * i], mask_rects[4 * i + 1], mask_rects[4 * i + 2], mask_rects[4 * i + 3]});
    e->initialize();
    return e;
}
void orc_destroy(orc_extractor* e) { delete e; }
void orc_set_max_kpts(orc_extractor* e, uint32_t n) {
    e->max_kpts = n;
    e->initialize();
}
int orc_get_table(const orc_extractor* e, int which, float* out, int cap) {
    const int n = static_cast<int>(e->num_levels);
    if (cap < n) return -2;
    for (int l = 0; l < n; ++l) {
        switch (which) {
            case 0: out[l] = e->scale[l]; break;
            case 1: out[l] = e->inv_scale[l]; break;
            case 2: out[l] = e->sigma_sq[l]; break;
            case 3: out[l] = e->inv_sigma_sq[l]; break;
            case 4: out[l] = static_cast<float>(e->quota[l]); break;
            default: return -1;
        }
    }
    return n;
}

int orc_extract(orc_extractor* e, const uint8_t* img, int w, int h, size_t pitch, const uint8_t* mask, size_t mask_pitch,
                orc_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* n_out) {
    const uint32_t L = e->num_levels;
    *n_out = 0;
    if (!img || w <= 0 || h <= 0) return 0; // orb_extractor.cc:73-75

    // pyramid: orb_extractor.cc:266-275
    e->pyr.assign(L, Image());
    e->blur.assign(L, Image());
    e->pyr[0].alloc(w, h);
    for (int y = 0; y < h; ++y) std::memcpy(e->pyr[0].row(y), img + static_cast<size_t>(y) * pitch, w);
    for (uint32_t l = 1; l < L; ++l) {
        const double s = e->scale[l];
        const int lw = static_cast<int>(std::round(w * 1.0 / s));
        const int lh = static_cast<int>(std::round(h * 1.0 / s));
        e->pyr[l].alloc(lw, lh);
        if (lw > 0 && lh > 0)
            cvprim::resize_linear_u8(e->pyr[l - 1].px.data(), e->pyr[l - 1].w, e->pyr[l - 1].h, e->pyr[l - 1].w,
                                     e->pyr[l].px.data(), lw, lh, lw);
    }

    // rectangle mask, drawn once: orb_extractor.cc:85-88,251-264
    if (!e->rect_mask_ready && !e->mask_rects.empty()) {
        e->rect_mask.alloc(w, h);
        std::fill(e->rect_mask.px.begin(), e->rect_mask.px.end(), 255);
        for (const auto& r : e->mask_rects) {
            const unsigned x0 = static_cast<unsigned>(std::round(static_cast<unsigned>(w) * r[0]));
            const unsigned x1 = static_cast<unsigned>(std::round(static_cast<unsigned>(w) * r[1]));
            const unsigned y0 = static_cast<unsigned>(std::round(static_cast<unsigned>(h) * r[2]));
            const unsigned y1 = static_cast<unsigned>(std::round(static_cast<unsigned>(h) * r[3]));
            cvprim::fill_rect_zero(e->rect_mask.px.data(), w, h, w, static_cast<int>(x0), static_cast<int>(y0),
                                   static_cast<int>(x1), static_cast<int>(y1));
        }
        e->rect_mask_ready = true;
    }
    // mask priority: orb_extractor.cc:93-107
    const uint8_t* m = nullptr;
    size_t mp = 0;
    if (mask) {
        m = mask;
        mp = mask_pitch;
    }
    else if (!e->rect_mask.px.empty()) {
        m = e->rect_mask.px.data();
        mp = e->rect_mask.w;
    }

    // FAST on the cell grid + quadtree: orb_extractor.cc:277-393
    e->cand.assign(L, {});
    e->sel.assign(L, {});
    std::vector<cvprim::FastPoint> cell_pts;
    for (uint32_t l = 0; l < L; ++l) {
        const Image& im = e->pyr[l];
        const float sf = e->scale[l];
        // the reference's unsigned arithmetic wraps (and effectively hangs) when a level is narrower than two borders;
        // such a level cannot hold a keypoint, so it is skipped here.
        if (im.w <= 2 * kBorder || im.h <= 2 * kBorder) continue;
        const unsigned min_bx = kBorder, min_by = kBorder;
        const unsigned max_bx = im.w - kBorder, max_by = im.h - kBorder;
        const unsigned width = max_bx - min_bx, height = max_by - min_by;
        const unsigned ncols = static_cast<unsigned>(std::ceil(width / kCell)) + 1; // integer division, orb_extractor.cc:302
        const unsigned nrows = static_cast<unsigned>(std::ceil(height / kCell)) + 1;
        std::vector<Cand>& cand = e->cand[l];
        for (unsigned i = 0; i < nrows; ++i) {
            const unsigned y0 = min_by + i * kCell;
            if (max_by - kOverlap <= y0) continue;
            unsigned y1 = y0 + kCell + kOverlap;
            if (max_by < y1) y1 = max_by;
            for (unsigned j = 0; j < ncols; ++j) {
                const unsigned x0 = min_bx + j * kCell;
                if (max_bx - kOverlap <= x0) continue;
                unsigned x1 = x0 + kCell + kOverlap;
                if (max_bx < x1) x1 = max_bx;
                if (m && (in_mask(m, mp, y0, x0, sf) || in_mask(m, mp, y1, x0, sf) || in_mask(m, mp, y0, x1, sf)
                          || in_mask(m, mp, y1, x1, sf)))
                    continue;
                cvprim::fast9_nms(im.row(y0) + x0, x1 - x0, y1 - y0, im.w, e->ini_thr, cell_pts);
                if (cell_pts.empty()) cvprim::fast9_nms(im.row(y0) + x0, x1 - x0, y1 - y0, im.w, e->min_thr, cell_pts);
                for (const auto& p : cell_pts) {
                    const float fx = static_cast<float>(p.x) + static_cast<float>(j * kCell);
                    const float fy = static_cast<float>(p.y) + static_cast<float>(i * kCell);
                    if (m && in_mask(m, mp, static_cast<unsigned>(min_by + fy), static_cast<unsigned>(min_bx + fx), sf)) continue;
                    cand.push_back({static_cast<int>(fx), static_cast<int>(fy), p.score});
                }
            }
        }
        e->sel[l] = distribute(cand, min_bx, max_bx, min_by, max_by, e->quota[l]);
    }

    uint32_t total = 0;
    for (uint32_t l = 0; l < L; ++l) total += static_cast<uint32_t>(e->sel[l].size());
    *n_out = total;
    if (total > cap) return -2;

    // orientation (un-blurred level, orb_extractor.cc:396-398), blur + descriptors + scale correction (:126-146)
    uint32_t off = 0;
    for (uint32_t l = 0; l < L; ++l) {
        const auto& sel = e->sel[l];
        if (sel.empty()) continue;
        const Image& im = e->pyr[l];
        Image& bl = e->blur[l];
        bl.alloc(im.w, im.h);
        cvprim::gaussian7_u8(im.px.data(), im.w, im.h, im.w, bl.px.data(), im.w);
        const float size = static_cast<float>(static_cast<unsigned>(kPatch * e->scale[l])); // orb_extractor.cc:383
        for (const Cand& c : sel) {
            orc_keypoint& k = kp_out[off];
            const int px = c.x + kBorder, py = c.y + kBorder;
            k.angle = ic_angle(im, px, py, e->u_max);
            rbrief(bl, px, py, k.angle, desc_out + static_cast<size_t>(off) * 32);
            k.x = static_cast<float>(px);
            k.y = static_cast<float>(py);
            if (l != 0) { // orb_extractor.cc:601-609
                k.x *= e->scale[l];
                k.y *= e->scale[l];
            }
            k.size = size;
            k.response = static_cast<float>(c.score);
            k.octave = static_cast<int32_t>(l);
            k.class_id = -1;
            ++off;
        }
    }
    return 0;
}

int orc_level_size(const orc_extractor* e, int level, int* w, int* h) {
    if (level < 0 || level >= static_cast<int>(e->pyr.size())) return -1;
    *w = e->pyr[level].w;
    *h = e->pyr[level].h;
    return 0;
}
int orc_level_copy(const orc_extractor* e, int level, int blurred, uint8_t* dst, size_t pitch) {
    if (level < 0 || level >= static_cast<int>(e->pyr.size())) return -1;
    const Image& im = blurred ? e->blur[level] : e->pyr[level];
    if (im.px.empty()) return -3;
    for (int y = 0; y < im.h; ++y) std::memcpy(dst + static_cast<size_t>(y) * pitch, im.row(y), im.w);
    return 0;
}
static int dump(const std::vector<Cand>& v, int32_t* xys, int cap) {
    const int n = static_cast<int>(v.size());
    for (int i = 0; i < std::min(n, cap); ++i) {
        xys[3 * i] = v[i].x;
        xys[3 * i + 1] = v[i].y;
        xys[3 * i + 2] = v[i].score;
    }
    return n;
}
int orc_level_candidates(const orc_extractor* e, int level, int32_t* xys, int cap) {
    if (level < 0 || level >= static_cast<int>(e->cand.size())) return -1;
    return dump(e->cand[level], xys, cap);
}
int orc_level_selected(const orc_extractor* e, int level, int32_t* xys, int cap) {
    if (level < 0 || level >= static_cast<int>(e->sel.size())) return -1;
    return dump(e->sel[level], xys, cap);
}

} // extern "C"
