// Host side of the extractor context: parameter tables, level / cell / quadtree geometry, device arena, launch sequence.
#include "ctx.h"

#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: ranges cost nothing unless a profiler injects itself

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace lps {

static thread_local std::string g_error;

void set_error(const std::string& msg) { g_error = msg; }
const std::string& get_error() { return g_error; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    g_error = buf;
    cudaGetLastError();  // clear the sticky-free error state
    return LPSLAM_ERR_CUDA;
}

namespace {

int round_half_even(double v) { return (int)std::lrint(v); }
int floor_int(double v) {
    const int i = (int)v;
    return i - (v < i);
}

struct ArenaPlan {
    size_t off = 0;
    size_t take(size_t bytes) {
        const size_t o = off;
        off = align_up(off + bytes, 256);
        return o;
    }
};

}  // namespace

// OpenCV resizeGeneric_ coefficient tables for INTER_LINEAR, 8-bit. `clamp_reset` is the
// horizontal-axis behaviour (offset clamped AND coefficient reset at both ends); rows are only clamped when read.
void make_axis_tab(int sn, int dn, bool clamp_reset, std::vector<int>& ofs, std::vector<short2>& coef) {
    ofs.resize(dn);
    coef.resize(dn);
    const double scale = (double)sn / dn;
    for (int d = 0; d < dn; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = floor_int(f);
        f -= s;
        if (clamp_reset) {
            if (s < 0) {
                s = 0;
                f = 0.f;
            }
            if (s >= sn - 1) {
                s = sn - 1;
                f = 0.f;
            }
        }
        ofs[d] = s;
        short2 c;
        c.x = (short)round_half_even((1.f - f) * 2048.f);
        c.y = (short)round_half_even(f * 2048.f);
        coef[d] = c;
    }
}

}  // namespace lps

using namespace lps;

// feature/orb_params.cc:38-72 (scale tables: float cumulative products) and feature/orb_extractor.cc:208-226 (quota:
// double geometric series, std::round, last level takes the remainder)
void lpslam_orb_ctx::initialize_tables() {
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
    double desired = max_kpts * (1.0 - 1.0 / scale_factor) / (1.0 - std::pow(1.0 / scale_factor, (double)num_levels));
    unsigned total = 0;
    for (uint32_t l = 0; l + 1 < num_levels; ++l) {
        quota[l] = (unsigned)std::round(desired);
        total += quota[l];
        desired *= 1.0 / scale_factor;
    }
    quota[num_levels - 1] = (unsigned)std::max((int)max_kpts - (int)total, 0);
}

void lpslam_orb_ctx::release_async() {
    for (auto& sl : slot) {
        if (sl.ev_done && sl.busy) cudaEventSynchronize(sl.ev_done);
        cudaFree(sl.landing);
        cudaFree(sl.kp);
        cudaFree(sl.desc);
        cudaFree(sl.xr);
        cudaFree(sl.dep);
        cudaFree(sl.off_dev);
        if (sl.h_off) cudaFreeHost(sl.h_off);
        if (sl.ev_in) cudaEventDestroy(sl.ev_in);
        if (sl.ev_landing_free) cudaEventDestroy(sl.ev_landing_free);
        if (sl.ev_done) cudaEventDestroy(sl.ev_done);
        sl = AsyncSlot{};
    }
    if (copy_in) cudaStreamDestroy(copy_in);
    if (copy_out) cudaStreamDestroy(copy_out);
    copy_in = copy_out = nullptr;
    async_ready = false;
    landing_bytes = 0;
}

void lpslam_orb_ctx::drop_pair_graph() {
    if (pair_graph) cudaGraphExecDestroy(pair_graph);
    pair_graph = nullptr;
    pair_graph_warm = 0;
}

void lpslam_orb_ctx::release_geometry() {
    drop_pair_graph();
    release_async();
    lps::blur_plan_destroy(blur_plan);
    blur_plan = nullptr;
    lps::pyramid_plan_destroy(pyr_plan);
    pyr_plan = nullptr;
    if (arena) cudaFree(arena);
    arena = nullptr;
    arena_bytes = 0;
    w = h = 0;
    has_result = false;
    rect_mask_ready = false;
}

int lpslam_orb_ctx::configure(int width, int height) {
    if (width == w && height == h && arena) return LPSLAM_OK;
    LPS_CUDA(cudaSetDevice(device));
    LPS_CUDA(cudaStreamSynchronize(stream));
    release_geometry();
    const int L = (int)num_levels;
    LPS_ARG(L >= 1 && L <= kMaxLevels, "num_levels out of range (1..16)");
    LPS_ARG(width <= 32767 && height <= 32767, "image larger than 32767 px is not supported");

    std::vector<std::vector<int>> xofs(L), yofs(L);
    std::vector<std::vector<short2>> xc(L), yc(L);
    cells.clear();
    memset(lv, 0, sizeof lv);
    ArenaPlan plan;
    std::vector<size_t> off_img(L), off_blur(L), off_xo(L), off_xc(L), off_yo(L), off_yc(L), off_rt(L), off_wt(L);
    std::vector<std::vector<uint4>> wintab(L);
    std::vector<std::vector<int2>> rowtab(L);
    std::vector<std::vector<uint16_t>> qix(L), qiy(L);
    std::vector<std::vector<int>> qbx(L), qby(L);
    std::vector<std::vector<uint32_t>> qxk(L), qyk(L);
    std::vector<size_t> off_qix(L), off_qiy(L), off_qbx(L), off_qby(L), off_qxk(L), off_qyk(L);
    int cand_total = 0, node_total = 0, sel_total = 0;
    for (int l = 0; l < L; ++l) {
        LevelGeom& g = lv[l];
        if (l == 0) {
            g.w = width;
            g.h = height;
        }
        else {
            // feature/orb_extractor.cc:270-271: size from the ORIGINAL image and the float scale widened to double
            const double sc = scale[l];
            g.w = (int)std::round(width * 1.0 / sc);
            g.h = (int)std::round(height * 1.0 / sc);
        }
        LPS_ARG(g.w > 0 && g.h > 0, "pyramid level collapses to zero size");
        g.pitch = (int)align_up(g.w + 2 * kPadLeft, 64);
        g.frame_stride = align_up((size_t)g.pitch * (g.h + 2 * kApron), 256);
        g.scale = scale[l];
        g.inv_scale = inv_scale[l];
        g.kp_size = (float)(unsigned)(kPatch * scale[l]);  // orb_extractor.cc:383
        g.quota = quota[l];
        off_img[l] = plan.take(g.frame_stride * max_batch + 8 * (size_t)g.pitch);  // slack: the resize prefetches rows past a frame
        off_blur[l] = plan.take(g.frame_stride * max_batch);
        if (l > 0) {
            const LevelGeom& p = lv[l - 1];
            if (p.w == g.w && p.h == g.h) g.mode = 2;
            else if (p.w == 2 * g.w && p.h == 2 * g.h) g.mode = 1;
            else g.mode = 0;
            make_axis_tab(p.w, g.w, true, xofs[l], xc[l]);
            make_axis_tab(p.h, g.h, false, yofs[l], yc[l]);
            off_xo[l] = plan.take(xofs[l].size() * 4);
            off_xc[l] = plan.take(xc[l].size() * 4);
            off_yo[l] = plan.take(yofs[l].size() * 4);
            off_yc[l] = plan.take(yc[l].size() * 4);
            // source-row schedule of the column-walk resize
            rowtab[l].assign(p.h + 2, make_int2(-1, 0));
            g.walk_ok = g.mode == 0 && p.w < 3 * g.w && p.h >= g.h && p.h < 24 * g.h && g.h > 2 * kApron + 1 && g.w > 1;
            for (int y = 0; y < g.h && g.walk_ok; ++y) {
                const int r = yofs[l][y] + 1;
                if (yofs[l][y] < 0 || r > p.h || rowtab[l][r].x >= 0) {
                    g.walk_ok = 0;
                    break;
                }
                rowtab[l][r] = make_int2(y, (int)((uint32_t)(uint16_t)yc[l][y].x | ((uint32_t)(uint16_t)yc[l][y].y << 16)));
            }
            off_rt[l] = plan.take(rowtab[l].size() * sizeof(int2));
            if (g.walk_ok) {
                // loop-invariant window description of the column-walk kernel, one entry per destination column pair
                auto refl = [](int p, int len) {
                    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
                    return p;
                };
                const int n_win = (g.w + 2 * kApron + 2560) / 2;
                wintab[l].resize(2 * (size_t)n_win);
                for (int i = 0; i < n_win; ++i) {
                    const int xg = 2 * i - 2 * kApron;
                    const int xa = refl(xg, g.w), xb = refl(xg + 1, g.w);
                    const int sa = xofs[l][xa], sb = xofs[l][xb];
                    const int base = std::min(sa, sb) & ~3;
                    const uint32_t oa = (uint32_t)std::min(sa - base, 6), ob = (uint32_t)std::min(sb - base, 6);
                    const short2 ca = xc[l][xa], cb = xc[l][xb];
                    wintab[l][2 * i] = make_uint4((uint32_t)base, oa | ((oa + 1) << 4), ob | ((ob + 1) << 4),
                                                  (uint32_t)(uint16_t)ca.x | ((uint32_t)(uint16_t)ca.y << 16));
                    wintab[l][2 * i + 1] = make_uint4((uint32_t)(uint16_t)cb.x | ((uint32_t)(uint16_t)cb.y << 16), xg < g.w + kApron ? 1u : 0u, 0u, 0u);
                }
                off_wt[l] = plan.take(wintab[l].size() * sizeof(uint4));
            }
        }
        // ---- FAST cell grid, feature/orb_extractor.cc:293-332 ----
        g.cell_begin = (int)cells.size();
        g.cell_cols = g.cell_rows = 0;
        if (g.w > 2 * kBorder && g.h > 2 * kBorder) {
            const unsigned max_bx = g.w - kBorder, max_by = g.h - kBorder;
            const unsigned width_b = max_bx - kBorder, height_b = max_by - kBorder;
            const unsigned ncols = width_b / kCell + 1, nrows = height_b / kCell + 1;  // std::ceil of an integer division
            int rows_inc = 0, cols_inc = 0;
            for (unsigned i = 0; i < nrows; ++i) {
                const unsigned y0 = kBorder + i * kCell;
                if (max_by - kOverlap <= y0) continue;
                ++rows_inc;
                const unsigned y1 = std::min(y0 + kCell + kOverlap, max_by);
                cols_inc = 0;
                for (unsigned j = 0; j < ncols; ++j) {
                    const unsigned x0 = kBorder + j * kCell;
                    if (max_bx - kOverlap <= x0) continue;
                    ++cols_inc;
                    const unsigned x1 = std::min(x0 + kCell + kOverlap, max_bx);
                    CellDesc c;
                    c.level = (short)l;
                    c.x0 = (short)x0;
                    c.y0 = (short)y0;
                    c.cw = (short)(x1 - x0);
                    c.ch = (short)(y1 - y0);
                    c.j = (short)j;
                    c.i = (short)i;
                    cells.push_back(c);
                }
            }
            g.cell_cols = cols_inc;
            g.cell_rows = rows_inc;
            // ---- quadtree initial grid, feature/orb_extractor.cc:483-503 ----
            const int min_x = kBorder, max_x = (int)max_bx, min_y = kBorder, max_y = (int)max_by;
            const double ratio = (double)(max_x - min_x) / (max_y - min_y);
            if (ratio > 1) {
                g.nx = (int)std::round(ratio);
                g.ny = 1;
                g.dx = (double)(max_x - min_x) / g.nx;
                g.dy = max_y - min_y;
            }
            else {
                g.nx = 1;
                g.ny = (int)std::round(1 / ratio);
                g.dx = max_x - min_y;  // sic, orb_extractor.cc:500
                g.dy = (double)(max_y - min_y) / g.ny;
            }
            int e = std::max((int)std::ceil(g.dx), (int)std::ceil(g.dy)) + 1;
            int d = 0;
            while (e > 1) {
                e = (e + 1) >> 1;
                ++d;
            }
            g.depth = d + 1;
            g.rbits = 0;
            while ((1 << g.rbits) < g.nx * g.ny) ++g.rbits;
            LPS_ARG(2 * g.depth + g.rbits <= 31 && g.nx * g.ny <= 1024, "image aspect ratio / size outside the quadtree key range");
        }
        else {
            g.nx = g.ny = 1;
            g.dx = g.dy = 1.0;
            g.depth = 1;
        }
        {   // initial-node lookup tables: the double division of the reference, done once on the host
            const int ext_x = std::max(g.w - 2 * kBorder, 1), ext_y = std::max(g.h - 2 * kBorder, 1);
            qix[l].resize(ext_x + 1);
            qiy[l].resize(ext_y + 1);
            for (int x = 0; x <= ext_x; ++x) qix[l][x] = (uint16_t)std::min<unsigned>((unsigned)((double)(float)x / g.dx), 65535u);
            for (int y = 0; y <= ext_y; ++y) qiy[l][y] = (uint16_t)std::min<unsigned>((unsigned)((double)(float)y / g.dy), 65535u);
            const int maxix = *std::max_element(qix[l].begin(), qix[l].end()), maxiy = *std::max_element(qiy[l].begin(), qiy[l].end());
            qbx[l].resize(std::max(g.nx, maxix + 1) + 1);
            qby[l].resize(std::max(g.ny, maxiy + 1) + 1);
            for (size_t i = 0; i < qbx[l].size(); ++i) qbx[l][i] = (int)(g.dx * (unsigned)i);
            for (size_t i = 0; i < qby[l].size(); ++i) qby[l][i] = (int)(g.dy * (unsigned)i);
            // Path-key halves. Whether a keypoint goes right at a depth depends on its x and its initial column only, whether it
            // goes down on its y and its initial row only (orb_extractor_node.cc:6-50: half extents round up), so the key
            // (initial node, then 2 bits per depth: bit 0 = right, bit 1 = bottom) is qxk[x] + qyk[y].
            auto half_key = [&](int v, const std::vector<uint16_t>& qi, const std::vector<int>& qb, int shift) {
                const unsigned i = qi[v];
                int b = qb[i], e = qb[i + 1];
                uint32_t path = 0;
                for (int d = 0; d < g.depth; ++d) {
                    const int hv = (e - b + 1) >> 1;  // cvCeil(extent / 2.0)
                    uint32_t bit = 0;
                    if (b + hv <= v) { bit = 1; b += hv; } else { e = b + hv; }
                    path = (path << 2) | (bit << shift);
                }
                return path;
            };
            qxk[l].resize(ext_x + 1);
            qyk[l].resize(ext_y + 1);
            for (int x = 0; x <= ext_x; ++x) qxk[l][x] = ((uint32_t)qix[l][x] << (2 * g.depth)) + half_key(x, qix[l], qbx[l], 0);
            for (int y = 0; y <= ext_y; ++y) qyk[l][y] = ((uint32_t)(qiy[l][y] * g.nx) << (2 * g.depth)) + half_key(y, qiy[l], qby[l], 1);
            off_qxk[l] = plan.take(qxk[l].size() * 4);
            off_qyk[l] = plan.take(qyk[l].size() * 4);
            off_qix[l] = plan.take(qix[l].size() * 2);
            off_qiy[l] = plan.take(qiy[l].size() * 2);
            off_qbx[l] = plan.take(qbx[l].size() * 4);
            off_qby[l] = plan.take(qby[l].size() * 4);
        }
        const int ncl = g.cell_cols * g.cell_rows;
        g.cand_begin = cand_total;
        g.cand_cap = ncl * kCellSlots;
        LPS_ARG(g.cand_cap < (1 << 24), "level too large: the quadtree packs a candidate index into 24 bits");
        cand_total += g.cand_cap;
        g.node_begin = node_total;
        g.node_cap = 4 * std::max<int>((int)g.quota, g.nx * g.ny) + 8;
        node_total += g.node_cap;
        g.sel_begin = sel_total;
        g.sel_cap = g.node_cap;
        sel_total += g.sel_cap;
    }
    num_cells = (int)cells.size();
    cand_stride = align_up(std::max(cand_total, 1), 64);
    node_stride = align_up(node_total, 64);
    sel_stride = align_up(sel_total, 64);
    out_cap = (uint32_t)std::min<size_t>((size_t)sel_total * max_batch, 0x7FFFFFFFu);

    const size_t B = max_batch;
    const size_t off_lv = plan.take(sizeof(LevelGeom) * kMaxLevels);
    const size_t off_cells = plan.take(sizeof(CellDesc) * std::max(num_cells, 1));
    const size_t off_counts = plan.take(4 * B * std::max(num_cells, 1));
    const size_t off_slots = plan.take(4 * B * std::max(num_cells, 1) * kCellSlots);
    const size_t off_cbase = plan.take(4 * B * std::max(num_cells, 1));
    const size_t off_cxy = plan.take(4 * B * cand_stride);
    const size_t off_csc = plan.take(1 * B * cand_stride);
    size_t off_keys[2], off_vals[2], off_node[2][3], off_tmp[8];
    for (int i = 0; i < 2; ++i) off_keys[i] = plan.take(4 * B * cand_stride);
    for (int i = 0; i < 2; ++i) off_vals[i] = plan.take(4 * B * cand_stride);
    for (int i = 0; i < 2; ++i)
        for (int k = 0; k < 3; ++k) off_node[i][k] = plan.take(4 * B * node_stride);
    for (int i = 0; i < 8; ++i) off_tmp[i] = plan.take(4 * B * node_stride);
    const size_t off_ccount = plan.take(4 * B * L);
    const size_t off_sxy = plan.take(4 * B * sel_stride);
    const size_t off_ssc = plan.take(1 * B * sel_stride);
    const size_t off_scount = plan.take(4 * B * L);
    const size_t off_loff = plan.take(4 * (B * L + 1));
    const size_t off_foff = plan.take(4 * (B + 2));
    const size_t off_kp = plan.take(sizeof(lpslam_keypoint) * (size_t)out_cap);
    const size_t off_desc = plan.take(32 * (size_t)out_cap);
    const size_t off_mask = plan.take((size_t)width * height);
    const size_t off_rmask = plan.take((size_t)width * height);
    const size_t off_srec = plan.take(8 * (size_t)out_cap);
    const size_t off_sxr = plan.take(4 * (size_t)out_cap);
    const size_t off_sdep = plan.take(4 * (size_t)out_cap);
    const size_t off_scorr = plan.take(4 * (size_t)out_cap);
    const size_t off_sperm = plan.take(16 * (size_t)out_cap);
    const size_t off_srows = plan.take(4 * (size_t)(height + 1) * ((max_batch + 1) / 2));

    LPS_CUDA(cudaMalloc(&arena, plan.off));
    arena_bytes = plan.off;
    for (int l = 0; l < L; ++l) {
        lv[l].img = arena + off_img[l] + (size_t)kApron * lv[l].pitch + kPadLeft;
        lv[l].blur = arena + off_blur[l] + (size_t)kApron * lv[l].pitch + kPadLeft;
        lv[l].qix = (const uint16_t*)(arena + off_qix[l]);
        lv[l].qiy = (const uint16_t*)(arena + off_qiy[l]);
        lv[l].qbx = (const int*)(arena + off_qbx[l]);
        lv[l].qby = (const int*)(arena + off_qby[l]);
        lv[l].qxk = (const uint32_t*)(arena + off_qxk[l]);
        lv[l].qyk = (const uint32_t*)(arena + off_qyk[l]);
        LPS_CUDA(cudaMemcpyAsync(arena + off_qxk[l], qxk[l].data(), qxk[l].size() * 4, cudaMemcpyHostToDevice, stream));
        LPS_CUDA(cudaMemcpyAsync(arena + off_qyk[l], qyk[l].data(), qyk[l].size() * 4, cudaMemcpyHostToDevice, stream));
        LPS_CUDA(cudaMemcpyAsync(arena + off_qix[l], qix[l].data(), qix[l].size() * 2, cudaMemcpyHostToDevice, stream));
        LPS_CUDA(cudaMemcpyAsync(arena + off_qiy[l], qiy[l].data(), qiy[l].size() * 2, cudaMemcpyHostToDevice, stream));
        LPS_CUDA(cudaMemcpyAsync(arena + off_qbx[l], qbx[l].data(), qbx[l].size() * 4, cudaMemcpyHostToDevice, stream));
        LPS_CUDA(cudaMemcpyAsync(arena + off_qby[l], qby[l].data(), qby[l].size() * 4, cudaMemcpyHostToDevice, stream));
        if (l > 0) {
            lv[l].xofs = (const int*)(arena + off_xo[l]);
            lv[l].xcoef = (const short2*)(arena + off_xc[l]);
            lv[l].yofs = (const int*)(arena + off_yo[l]);
            lv[l].ycoef = (const short2*)(arena + off_yc[l]);
            lv[l].srctab = (const int2*)(arena + off_rt[l]);
            if (lv[l].walk_ok) {
                lv[l].wintab = (const uint4*)(arena + off_wt[l]);
                LPS_CUDA(cudaMemcpyAsync(arena + off_wt[l], wintab[l].data(), wintab[l].size() * sizeof(uint4), cudaMemcpyHostToDevice, stream));
            }
            LPS_CUDA(cudaMemcpyAsync(arena + off_rt[l], rowtab[l].data(), rowtab[l].size() * sizeof(int2), cudaMemcpyHostToDevice, stream));
            LPS_CUDA(cudaMemcpyAsync(arena + off_xo[l], xofs[l].data(), xofs[l].size() * 4, cudaMemcpyHostToDevice, stream));
            LPS_CUDA(cudaMemcpyAsync(arena + off_xc[l], xc[l].data(), xc[l].size() * 4, cudaMemcpyHostToDevice, stream));
            LPS_CUDA(cudaMemcpyAsync(arena + off_yo[l], yofs[l].data(), yofs[l].size() * 4, cudaMemcpyHostToDevice, stream));
            LPS_CUDA(cudaMemcpyAsync(arena + off_yc[l], yc[l].data(), yc[l].size() * 4, cudaMemcpyHostToDevice, stream));
        }
    }
    lv_dev = (LevelGeom*)(arena + off_lv);
    cells_dev = (CellDesc*)(arena + off_cells);
    LPS_CUDA(cudaMemcpyAsync(lv_dev, lv, sizeof(LevelGeom) * kMaxLevels, cudaMemcpyHostToDevice, stream));
    if (num_cells)
        LPS_CUDA(cudaMemcpyAsync(cells_dev, cells.data(), sizeof(CellDesc) * num_cells, cudaMemcpyHostToDevice, stream));
    cell_counts = (uint32_t*)(arena + off_counts);
    cell_slots = (uint32_t*)(arena + off_slots);

    tree = TreeArgs{};
    tree.lv = lv_dev;
    tree.num_levels = L;
    tree.num_cells = num_cells;
    tree.cells = cells_dev;
    tree.cell_counts = cell_counts;
    tree.cell_slots = cell_slots;
    tree.cand_xy = (uint32_t*)(arena + off_cxy);
    tree.cand_stride = cand_stride;
    tree.cand_sc = arena + off_csc;
    for (int i = 0; i < 2; ++i) {
        tree.keys[i] = (uint32_t*)(arena + off_keys[i]);
        tree.vals[i] = (uint32_t*)(arena + off_vals[i]);
        for (int k = 0; k < 3; ++k) tree.node[i][k] = (int*)(arena + off_node[i][k]);
    }
    for (int i = 0; i < 8; ++i) tree.tmp[i] = (int*)(arena + off_tmp[i]);
    tree.cell_base = (uint32_t*)(arena + off_cbase);
    tree.cell_stride = std::max(num_cells, 1);
    tree.node_stride = node_stride;
    tree.cand_count = (uint32_t*)(arena + off_ccount);
    tree.sel_xys = (uint32_t*)(arena + off_sxy);
    tree.sel_stride = sel_stride;
    tree.sel_sc = arena + off_ssc;
    tree.sel_count = (uint32_t*)(arena + off_scount);

    desc = DescribeArgs{};
    desc.lv = lv_dev;
    desc.num_levels = L;
    desc.sel_xys = tree.sel_xys;
    desc.sel_stride = sel_stride;
    desc.sel_sc = tree.sel_sc;
    desc.sel_count = tree.sel_count;
    desc.level_off = (uint32_t*)(arena + off_loff);
    desc.frame_off = (uint32_t*)(arena + off_foff);
    desc.overflow = desc.frame_off + B + 1;
    desc.kp = (lpslam_keypoint*)(arena + off_kp);
    desc.desc = arena + off_desc;
    desc.out_cap = out_cap;
    mask_dev = arena + off_mask;
    rect_mask_dev = arena + off_rmask;
    stereo = StereoArgs{};
    stereo.lv = lv_dev;
    stereo.num_levels = L;
    stereo.kp = desc.kp;
    stereo.desc = desc.desc;
    stereo.frame_off = desc.frame_off;
    stereo.max_rows = out_cap;
    stereo.rec = (uint2*)(arena + off_srec);
    stereo.x_right = (float*)(arena + off_sxr);
    stereo.depth = (float*)(arena + off_sdep);
    stereo.corr = (int*)(arena + off_scorr);
    stereo.perm = (uint4*)(arena + off_sperm);
    stereo.row_start = (uint32_t*)(arena + off_srows);
    stereo.rows0 = height;
    LPS_CUDA(cudaStreamSynchronize(stream));  // the host-side tables go out of scope
    {
        static const bool blur_v1 = getenv("LPS_BLUR_V1") != nullptr;  // A/B switch: the ld.global version of blur.cu
        if (!blur_v1) {
            // nullptr (a tensor map could not be encoded for this geometry) = the ld.global blur of blur.cu, as pyr_plan
            // does below: same results, and configure() still completes so that later calls do not rebuild the arena
            blur_plan = blur_plan_create(lv, L, (int)max_batch);
            if (!blur_plan) fprintf(stderr, "liblpslam_cuda: TMA blur unavailable for %dx%d (%s), using the ld.global blur\n", width, height,
                                    lpslam_cuda_last_error());
        }
        pyr_plan = pyramid_plan_create(lv, L, (int)max_batch);  // nullptr = geometry outside the kernel's range
    }
    w = width;
    h = height;
    return LPSLAM_OK;
}

int lpslam_orb_ctx::run_pipeline(uint32_t n_frames, const uint8_t* mask_ptr, size_t mask_pitch) {
    const int L = (int)num_levels;
    const int B = (int)n_frames;
    int e = 0;
    // NVTX ranges around the six stages (what nsys / ncu --nvtx group by); the ranges bracket the ENQUEUE of a stage
    static const char* const kStage[] = {"lpslam:pyramid", "lpslam:fast", "lpslam:quadtree", "lpslam:blur", "lpslam:describe"};
    bool range_open = false;
    auto mark = [&](int idx) {  // idx = stage that starts here (5 = end of the pipeline)
        if (timing) cudaEventRecord(ev[idx], stream);
        if (range_open) nvtxRangePop();
        range_open = idx < 5;
        if (range_open) nvtxRangePushA(kStage[idx]);
    };
    (void)e;
    mark(0);  // start of pyramid
    uint32_t pad_mask = 1u;  // level 0 arrives without its apron; so do levels made by the exact-half / copy / wide-ratio paths
    for (int l = 1; l < L && !pyr_plan; ++l)
        if (!resize_writes_apron(lv[l - 1], lv[l])) pad_mask |= 1u << l;
    if (pyr_plan) {
        if (int r = launch_pyramid_tma(pyr_plan, lv_dev, B, stream)) return r;
    }
    else {
        const int chains = std::min(pyr_chains, B);
        auto run_chain = [&](int f0, int f1, cudaStream_t st) {
            for (int l = 1; l < L; ++l) {
                LevelGeom s = lv[l - 1], d = lv[l];
                s.img += (size_t)f0 * s.frame_stride;
                d.img += (size_t)f0 * d.frame_stride;
                launch_resize(s, d, f1 - f0, st);
            }
        };
        if (chains > 1) LPS_CUDA(cudaEventRecord(ev_chain_fork, stream));
        for (int c = 1; c < chains; ++c) {
            LPS_CUDA(cudaStreamWaitEvent(chain[c - 1], ev_chain_fork, 0));
            run_chain(B * c / chains, B * (c + 1) / chains, chain[c - 1]);
            LPS_CUDA(cudaEventRecord(ev_chain_join[c - 1], chain[c - 1]));
        }
        run_chain(0, B / chains, stream);
        for (int c = 1; c < chains; ++c) LPS_CUDA(cudaStreamWaitEvent(stream, ev_chain_join[c - 1], 0));
    }
    // The blur and the apron of level 0 depend on the pyramid only: outside of stage timing they run on the side stream,
    // next to FAST and the quadtree (memory-bound work beside compare-pipe-bound and latency-bound work), and are joined
    // before the descriptors. With stage timing on, everything stays on one stream so that each stage is timed alone.
    const bool fork = !timing && side && pad_mask == 1u;
    if (fork) {
        LPS_CUDA(cudaEventRecord(ev_fork, stream));
        LPS_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        launch_pad_borders(lv_dev, lv, L, pad_mask, B, side);
        if (blur_plan) {
            if (int r = launch_blur_tma(blur_plan, lv_dev, B, side)) return r;
        }
        else launch_blur(lv_dev, lv, L, B, side);
        LPS_CUDA(cudaEventRecord(ev_join, side));
    }
    else {
        launch_pad_borders(lv_dev, lv, L, pad_mask, B, stream);
    }
    mark(1);
    FastArgs fa;
    fa.lv = lv_dev;
    fa.cells = cells_dev;
    fa.num_cells = num_cells;
    fa.batch = B;
    fa.frame0 = 0;
    fa.ini_thr = ini_thr;
    fa.min_thr = min_thr;
    fa.cell_counts = cell_counts;
    fa.cell_slots = cell_slots;
    fa.mask = mask_ptr;
    fa.mask_pitch = mask_pitch;
    tree.batch = B;
    tree.frame0 = 0;
    // Outside of stage timing FAST + quadtree run as sub-batches of frames on their own streams: the quadtree is a
    // latency-bound kernel (block-serial list discipline, ~27 % issue utilisation) and FAST is bound by the compare pipe, so
    // the quadtree of one sub-batch runs beside the FAST of the next instead of after it.
    const int subs = timing ? 1 : std::min(std::min(sub_batches, B), 4);
    if (subs > 1) {
        LPS_CUDA(cudaEventRecord(ev_chain_fork, stream));
        for (int sb = 0; sb < subs; ++sb) {
            cudaStream_t st = sb == 0 ? stream : chain[sb - 1];
            if (sb > 0) LPS_CUDA(cudaStreamWaitEvent(st, ev_chain_fork, 0));
            FastArgs f2 = fa;
            f2.frame0 = B * sb / subs;
            f2.batch = B * (sb + 1) / subs - f2.frame0;
            launch_fast(f2, st);
            TreeArgs t2 = tree;
            t2.frame0 = f2.frame0;
            t2.batch = f2.batch;
            launch_quadtree(t2, lv, st);
            if (sb > 0) LPS_CUDA(cudaEventRecord(ev_chain_join[sb - 1], st));
        }
        for (int sb = 1; sb < subs; ++sb) LPS_CUDA(cudaStreamWaitEvent(stream, ev_chain_join[sb - 1], 0));
    }
    else {
        launch_fast(fa, stream);
        mark(2);
        launch_quadtree(tree, lv, stream);
    }
    mark(3);
    if (fork)
        LPS_CUDA(cudaStreamWaitEvent(stream, ev_join, 0));
    else if (blur_plan) {
        if (int r = launch_blur_tma(blur_plan, lv_dev, B, stream)) return r;
    }
    else
        launch_blur(lv_dev, lv, L, B, stream);
    mark(4);
    desc.batch = B;
    LPS_CUDA(cudaMemsetAsync(desc.overflow, 0, 4, stream));
    launch_offsets(desc, stream);
    launch_describe(desc, stream);
    mark(5);
    LPS_CUDA(cudaGetLastError());
    last_batch = n_frames;
    has_result = true;
    return LPSLAM_OK;
}
