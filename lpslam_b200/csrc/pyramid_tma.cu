// K1 (EXPERIMENT, opt-in with LPS_PYR_TMA=1; pyramid.cu's k_resize_walk chain is the production path) — the image pyramid as ONE launch with the source rows staged through
// shared memory by TMA. Replaces cv::resize(..., INTER_LINEAR) at feature/orb_extractor.cc:266-275; arithmetic as in
// pyramid.cu: 11-bit coefficients, horizontal pass in int32, vertical pass
//   dst = (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2.
//
// Why a second version: the seven dependent k_resize_walk launches ran at 0.35 of the HBM copy ceiling. Each level paid
// its own ramp and tail (the four small levels are launch-latency bound), a thread spent ~17.7 instructions per
// destination pixel (pointer arithmetic and two global loads per 2-pixel window and source row), and occupancy sat at 40 %
// on a long-scoreboard-bound loop. Here
//   * all levels of all frames are one grid of independent warps, ordered level-major; a unit of destination level l waits
//     on a device counter until level l-1 of ITS frame is complete, so frames flow through the levels without a global
//     barrier and without launch gaps (a unit only ever waits for units that were dispatched before it);
//   * a half-warp owns 16 strips of 8 destination columns and streams down a segment of destination rows in stages of 8;
//     the 12 source rows a stage can touch arrive as one cp.async.bulk.tensor box per half-warp (double buffered,
//     completion on an mbarrier per warp and stage, lane 0 is the warp's own producer);
//   * a lane cuts the 12 source bytes its 8 pixels read out of four LDS.32 (3 PRMT re-align them to its first pixel), then
//     spends PRMT + IDP.2A + SHF per pixel and source row; the vertical pass is the one of k_resize_walk
//     (2 IMUL + PRMT + IDP.2A per pixel); output rows leave as one 8-byte store per lane;
//   * the reflected apron of a level (what the blur's halo reads) is written by the same kernel: the left / right apron
//     are ordinary strips computed over the mirrored columns and stored byte-reversed, apron rows are second stores.
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

namespace lps {

namespace {

constexpr int PT_DROWS = 8;                        // destination rows per stage
constexpr int PT_SROWS = 12;                       // source rows per box: ceil(7 * 1.43) + 2
constexpr int PT_BOXW = 192;                       // source bytes per box row: ceil(128 * 1.35) + 2 + 15 (16-byte start) -> 192
constexpr int PT_HALF_BYTES = PT_SROWS * PT_BOXW;  // 2304 = 18 x 128
constexpr int PT_STAGE_BYTES = 2 * PT_HALF_BYTES;
constexpr int PT_STAGES = 2;
constexpr int PT_SCHED_BYTES = PT_STAGES * PT_SROWS * 16;
constexpr int PT_WARP_BYTES = PT_STAGES * PT_STAGE_BYTES + PT_SCHED_BYTES;  // 9216 + 384 = 9600 = 75 x 128

struct PyrMaps {
    CUtensorMap m[kMaxLevels];  // m[l] = level l as a SOURCE (box PT_BOXW x PT_SROWS)
};

struct PyrStrip {     // loop-invariant description of one strip of 8 destination columns (64 bytes)
    int32_t win;      // byte offset of the lane's 12-byte source window inside its half-warp's box row
    int32_t dst_x;    // first destination column the 8 bytes are stored at (apron strips: negative / >= w)
    uint32_t flags;   // bit 0 = stored byte-reversed (apron strip), bit 1 = byte stores, bit 2 = live, bits 8.. = bytes stored (byte-store strips)
    uint32_t pad;
    uint32_t sel[4];  // PRMT selectors of pixels (2k, 2k+1) as two 16-bit halves
    uint32_t coef[8]; // (c0 | c1 << 16) per pixel
};
static_assert(sizeof(PyrStrip) == 64, "PyrStrip layout");

struct PyrUnit {  // one warp: two neighbouring half-bands of one destination level, one segment of rows
    int level;    // destination level
    int y0, stages;
    int strip0[2];  // first strip of each half in the level's strip table (-1 = half unused)
    int box_col[2]; // tensor column of the half's box in the source level
    int pad;
};

struct PyrSched {  // per source row j of a stage (warp-uniform, read as one LDS.128)
    int y;         // destination row completed by box rows (j - 1, j), or -1
    uint32_t b01;  // row coefficients b0 | b1 << 16
    int dup;       // apron row that receives a copy of y, or INT_MIN
    int pad;
};

__device__ __forceinline__ uint32_t ld_acquire(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(32, 24) k_pyramid_tma(const __grid_constant__ PyrMaps maps, const LevelGeom* __restrict__ lvs,
                                                        const PyrUnit* __restrict__ units, const PyrStrip* __restrict__ strips,
                                                        int batch, uint32_t* __restrict__ done, const uint32_t* __restrict__ need,
                                                        uint32_t epoch, int num_levels) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[PT_STAGES];
    uint8_t* const smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const int lane = threadIdx.x;
    const int half = lane >> 4, hl = lane & 15;
    const uint32_t g = blockIdx.x;
    const PyrUnit u = units[g / (uint32_t)batch];
    const int frame = (int)(g % (uint32_t)batch);
    const LevelGeom& dl = lvs[u.level];
    const int h = dl.h, pitch = dl.pitch, src_h = lvs[u.level - 1].h;
    const uint32_t sbase = smem_u32(smem), bar0 = smem_u32(&s_full[0]);
    PyrSched* const sched = reinterpret_cast<PyrSched*>(smem + PT_STAGES * PT_STAGE_BYTES);
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < PT_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
        // the source level of this frame must be complete (level 0 is the input)
        if (u.level > 1) {
            const uint32_t* c = done + (size_t)frame * num_levels + (u.level - 1);
            const uint32_t want = epoch * need[u.level - 1];
            while ((int32_t)(ld_acquire(c) - want) < 0) __nanosleep(200);
        }
        asm volatile("fence.proxy.async;" ::: "memory");  // what other SMs stored with st.global is read by the TMA unit below
    }
    __syncwarp();
    const CUtensorMap* map = &maps.m[u.level - 1];
    const int nh = u.strip0[1] >= 0 ? 2 : 1;
    const int* __restrict__ yofs = dl.yofs;
    const int2* __restrict__ srctab = dl.srctab;
    auto issue = [&](int it) {  // lane 0: the 12 source rows from yofs[first destination row of the stage] on
        const int s = it & (PT_STAGES - 1);
        const uint32_t bar = bar0 + 8 * s, dst = sbase + s * PT_STAGE_BYTES;
        const int r0 = __ldg(yofs + min(u.y0 + PT_DROWS * it, h - 1));
        mbar_expect_tx(bar, nh * PT_HALF_BYTES);
        tma_load_3d(dst, map, u.box_col[0], r0 + kApron, frame, bar);
        if (nh == 2) tma_load_3d(dst + PT_HALF_BYTES, map, u.box_col[1], r0 + kApron, frame, bar);
    };
    // schedule of stage `it`: lane j describes box row j (which destination row the pair (j - 1, j) completes)
    auto make_sched = [&](int it) {
        if (lane < PT_SROWS) {
            const int ys = u.y0 + PT_DROWS * it;
            const int r0 = __ldg(yofs + min(ys, h - 1));
            PyrSched e;
            e.y = -1;
            e.b01 = 0;
            e.dup = INT_MIN;
            e.pad = 0;
            if (lane >= 1) {
                const int2 t = __ldg(srctab + min(r0 + lane, src_h + 1));  // indexed by source row; entry src_h + 1 is always (-1, 0)
                if (t.x >= ys && t.x < ys + PT_DROWS && t.x < h) {
                    e.y = t.x;
                    e.b01 = (uint32_t)t.y;
                    if (t.x >= 1 && t.x <= kApron) e.dup = -t.x;
                    if (t.x >= h - 1 - kApron && t.x <= h - 2) e.dup = 2 * (h - 1) - t.x;
                }
            }
            sched[(it & 1) * PT_SROWS + lane] = e;
        }
    };
    if (lane == 0) {
        issue(0);
        if (u.stages > 1) issue(1);
    }
    make_sched(0);

    // loop-invariant description of this lane's strip
    const int my_strip0 = half ? u.strip0[1] : u.strip0[0];
    PyrStrip sp;
    {
        const uint4* p = reinterpret_cast<const uint4*>(strips + (my_strip0 >= 0 ? my_strip0 + hl : 0));
        const uint4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3);
        sp.win = (int)a.x;
        sp.dst_x = (int)a.y;
        sp.flags = my_strip0 >= 0 ? a.z : 0u;
        sp.sel[0] = b.x, sp.sel[1] = b.y, sp.sel[2] = b.z, sp.sel[3] = b.w;
        sp.coef[0] = c.x, sp.coef[1] = c.y, sp.coef[2] = c.z, sp.coef[3] = c.w;
        sp.coef[4] = d.x, sp.coef[5] = d.y, sp.coef[6] = d.z, sp.coef[7] = d.w;
    }
    const bool live = (sp.flags & 4u) != 0, rev = (sp.flags & 1u) != 0, bytes = (sp.flags & 2u) != 0;
    const uint32_t mis = (uint32_t)sp.win & 3u;
    const uint32_t sel_al = 0x3210u + 0x1111u * mis;  // re-aligns two consecutive words to the window's first byte
    const uint8_t* const lane_smem = smem + half * PT_HALF_BYTES + (sp.win & ~3);
    uint8_t* const dst = dl.img + frame * dl.frame_stride + sp.dst_x;
    // output byte order: as computed, or reversed for the mirrored apron strips
    const uint32_t sel_o0 = rev ? 0x4567u : 0x3210u, sel_o1 = rev ? 0x0123u : 0x7654u;

    auto hrow = [&](const uint8_t* row, uint32_t (&H)[8]) {  // H >> 4 of the lane's 8 pixels on one source row
        const uint32_t* w = reinterpret_cast<const uint32_t*>(row);
        const uint32_t a0 = w[0], a1 = w[1], a2 = w[2], a3 = w[3];
        const uint32_t W0 = __byte_perm(a0, a1, sel_al), W1 = __byte_perm(a1, a2, sel_al), W2 = __byte_perm(a2, a3, sel_al);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t sel = (k & 1) ? sp.sel[k >> 1] >> 16 : sp.sel[k >> 1];  // PRMT reads the low 16 bits only
            const uint32_t px = k < 4 ? __byte_perm(W0, W1, sel) : __byte_perm(W1, W2, sel);
            H[k] = __dp2a_lo(sp.coef[k], px, 0u) >> 4;
        }
    };
    auto emit = [&](const PyrSched e, const uint32_t (&Lo)[8], const uint32_t (&Up)[8]) {
        const uint32_t b0 = e.b01 & 0xFFFFu, b1 = e.b01 >> 16;
        uint32_t sum[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) sum[k] = __dp2a_lo(__byte_perm(b0 * Lo[k], b1 * Up[k], 0x7632), 0x4040u, 128u);  // (t0 + t1 + 2) << 6
        const uint32_t p01 = __byte_perm(sum[0], sum[1], 0x4451), p23 = __byte_perm(sum[2], sum[3], 0x4451);
        const uint32_t p45 = __byte_perm(sum[4], sum[5], 0x4451), p67 = __byte_perm(sum[6], sum[7], 0x4451);
        const uint32_t lo = __byte_perm(p01, p23, 0x5410), hi = __byte_perm(p45, p67, 0x5410);
        const uint32_t o0 = __byte_perm(lo, hi, sel_o0), o1 = __byte_perm(lo, hi, sel_o1);
        if (live) {
            uint8_t* o = dst + (ptrdiff_t)e.y * pitch;
            if (!bytes) {
                *reinterpret_cast<uint2*>(o) = make_uint2(o0, o1);
                if (e.dup != INT_MIN) *reinterpret_cast<uint2*>(dst + (ptrdiff_t)e.dup * pitch) = make_uint2(o0, o1);
            }
            else {
                uint8_t* o2 = e.dup != INT_MIN ? dst + (ptrdiff_t)e.dup * pitch : o;
                const int nb = (int)(sp.flags >> 8);  // right apron: kApron columns; last real strip: up to column w - 1
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    if (k < nb) {
                        const uint8_t v = (uint8_t)((k < 4 ? o0 : o1) >> (8 * (k & 3)));
                        o[k] = v;
                        o2[k] = v;
                    }
            }
        }
    };

    for (int it = 0; it < u.stages; ++it) {
        const int s = it & (PT_STAGES - 1);
        if (it + 1 < u.stages) make_sched(it + 1);
        mbar_wait(bar0 + 8 * s, (it >> 1) & 1);
        __syncwarp();  // the schedule of this stage (written one iteration ago) is visible
        const uint8_t* const st = lane_smem + s * PT_STAGE_BYTES;
        const PyrSched* const sc = sched + s * PT_SROWS;
        uint32_t A[8], B[8];
        hrow(st, A);
#pragma unroll
        for (int j = 1; j < PT_SROWS; j += 2) {
            {
                const PyrSched e = sc[j];
                hrow(st + j * PT_BOXW, B);
                if (e.y >= 0) emit(e, A, B);
            }
            if (j + 1 < PT_SROWS) {
                const PyrSched e = sc[j + 1];
                hrow(st + (j + 1) * PT_BOXW, A);
                if (e.y >= 0) emit(e, B, A);
            }
        }
        __syncwarp();  // every lane is done with stage s
        if (lane == 0 && it + PT_STAGES < u.stages) issue(it + PT_STAGES);
    }
    // this unit's rows are complete: publish (the counter is what units of the next level acquire)
    __syncwarp();
    if (lane == 0) {
        __threadfence();
        atomicAdd(done + (size_t)frame * num_levels + u.level, 1u);
    }
}

}  // namespace

struct PyramidPlan {
    PyrMaps maps;
    int num_levels = 0, max_batch = 0;
    PyrUnit* units_dev = nullptr;
    PyrStrip* strips_dev = nullptr;
    uint32_t* done_dev = nullptr;  // [max_batch][num_levels] units completed, monotonic over launches
    uint32_t* need_dev = nullptr;  // [num_levels] units per level and frame
    int n_units = 0;
    uint32_t epoch = 0;
    int last_batch = -1;
};

void pyramid_plan_destroy(PyramidPlan* p) {
    if (!p) return;
    cudaFree(p->units_dev);
    cudaFree(p->strips_dev);
    cudaFree(p->done_dev);
    cudaFree(p->need_dev);
    delete p;
}

#define PT_REJECT(code)                                                                              \
    do {                                                                                             \
        if (getenv("LPS_PYR_DEBUG")) fprintf(stderr, "pyramid_tma: geometry rejected at check %d (level %d)\n", code, l); \
        return nullptr;                                                                              \
    } while (0)
PyramidPlan* pyramid_plan_create(const LevelGeom* lv, int num_levels, int max_batch) {
    // Opt-in (LPS_PYR_TMA=1): bit-exact on every golden (aprons included), but measured SLOWER than the k_resize_walk chain on
    // B200: 0.22 ms against 0.127 ms per 64 frames of 1280x720 (profiles/r2_pyramid_tma.md). A level of a frame can only
    // start when the level below is complete, so the levels still run as seven phases; units small enough to fill the
    // machine inside one phase pay their set-up (tables, first TMA round trip) per 8 .. 16 rows, and 38 % of all warp
    // instructions of the launch were polls of the completion counters.
    static const bool on = getenv("LPS_PYR_TMA") != nullptr;
    if (!on || num_levels < 2) return nullptr;
    static const int seg_rows = [] {
        const char* e = getenv("LPS_PYR_SEG_ROWS");
        return e ? std::max(PT_DROWS, atoi(e) / PT_DROWS * PT_DROWS) : 64;
    }();
    static const int fill_x10 = [] {
        const char* e = getenv("LPS_PYR_FILL");
        return e ? std::max(1, atoi(e)) : 10;
    }();
    std::vector<PyrStrip> strips;
    std::vector<PyrUnit> units;
    std::vector<uint32_t> need(num_levels, 0);
    auto refl = [](int p, int len) {
        while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
        return p;
    };
    for (int l = 1; l < num_levels; ++l) {
        const LevelGeom &s = lv[l - 1], &d = lv[l];
        if (d.mode != 0 || d.w < 16 || d.h < 2 * kApron + 2 || s.w < d.w || s.h < d.h) PT_REJECT(2);
        std::vector<int> xofs, yofs;
        std::vector<short2> xc, yc;
        make_axis_tab(s.w, d.w, true, xofs, xc);
        make_axis_tab(s.h, d.h, false, yofs, yc);
        // rows: a stage of 8 destination rows must fit the 12-row box, rows strictly increasing (one completion per source row)
        for (int y = 0; y < d.h; ++y) {
            if (yofs[y] < 0 || yofs[y] + 1 > s.h) PT_REJECT(3);
            if (y > 0 && yofs[y] <= yofs[y - 1]) PT_REJECT(4);
            const int ye = std::min(y + PT_DROWS - 1, d.h - 1);
            if (y % PT_DROWS == 0 && yofs[ye] + 1 - yofs[y] >= PT_SROWS) PT_REJECT(5);
        }
        // strips of the level: [left apron (reversed)] [real strips] [right apron (reversed, byte stores)]
        struct Cols {
            int x[8];      // destination columns computed (mirrored for the apron strips), increasing
            int dst_x;
            uint32_t flags;
        };
        std::vector<Cols> cols;
        {
            Cols c{};
            for (int k = 0; k < 8; ++k) c.x[k] = refl(1 + k, d.w);
            c.dst_x = -8;
            c.flags = 1u | 4u;
            cols.push_back(c);
            for (int x0 = 0; x0 < d.w; x0 += 8) {
                for (int k = 0; k < 8; ++k) c.x[k] = std::min(x0 + k, d.w - 1);
                c.dst_x = x0;
                c.flags = 4u;
                if (x0 + 8 > d.w) c.flags |= 2u | ((uint32_t)(d.w - x0) << 8);  // the right apron strip owns columns w ..
                cols.push_back(c);
            }
            for (int k = 0; k < 8; ++k) c.x[k] = std::max(d.w - 9 + k, 0);  // stored reversed: column w + j = pixel w - 2 - j
            c.dst_x = d.w;
            c.flags = 1u | 2u | 4u | ((uint32_t)kApron << 8);
            cols.push_back(c);
        }
        const int level_strip0 = (int)strips.size();
        const int n_strips = (int)cols.size(), halves = (n_strips + 15) / 16;
        std::vector<int> box_col(halves);
        for (int hb = 0; hb < halves; ++hb) {
            int lo = 1 << 30, hi = 0;
            for (int i = 16 * hb; i < std::min(16 * hb + 16, n_strips); ++i)
                for (int k = 0; k < 8; ++k) {
                    lo = std::min(lo, xofs[cols[i].x[k]]);
                    hi = std::max(hi, xofs[cols[i].x[k]] + 1);
                }
            box_col[hb] = (lo + kPadLeft) & ~15;  // tensor column = image column + kPadLeft, boxes start on 16 bytes
            if (hi + kPadLeft - box_col[hb] + 1 > PT_BOXW) PT_REJECT(6);
        }
        for (int i = 0; i < 16 * halves; ++i) {
            PyrStrip ps{};
            if (i < n_strips) {
                const Cols& c = cols[i];
                int base = 1 << 30;
                for (int k = 0; k < 8; ++k) base = std::min(base, xofs[c.x[k]]);
                ps.win = base + kPadLeft - box_col[i / 16];
                if ((ps.win & ~3) + 16 > PT_BOXW) PT_REJECT(7);
                ps.dst_x = c.dst_x;
                ps.flags = c.flags;
                const int nstored = (c.flags & 2u) && !(c.flags & 1u) ? (int)(c.flags >> 8) : 8;  // last real strip: pixels past w - 1 are not stored
                for (int k = 0; k < 8; ++k) {
                    int p = xofs[c.x[k]] - base;  // position inside the re-aligned 12-byte window
                    if (k >= nstored) p = 4;      // any readable bytes will do
                    uint32_t sel;
                    if (k < 4) {
                        if (p > 6) PT_REJECT(8);
                        sel = (uint32_t)p | ((uint32_t)(p + 1) << 4);
                    }
                    else {
                        if (p < 4 || p > 10) PT_REJECT(9);
                        sel = (uint32_t)(p - 4) | ((uint32_t)(p - 3) << 4);
                    }
                    sel |= 0x4400u;  // upper bytes: anything (IDP.2A reads the two low bytes)
                    ps.sel[k >> 1] |= sel << (16 * (k & 1));
                    ps.coef[k] = (uint32_t)(uint16_t)xc[c.x[k]].x | ((uint32_t)(uint16_t)xc[c.x[k]].y << 16);
                }
            }
            strips.push_back(ps);
        }
        const int stages_all = (d.h + PT_DROWS - 1) / PT_DROWS;
        // Levels run as phases (a level of a frame starts when the level below is complete, and all frames advance
        // together), so every level must fill the machine on its own: the segment length is chosen per level so that the
        // level has about `fill` warps per resident warp slot for this batch
        int level_seg = seg_rows;
        {
            const long slots = 148L * 24;
            const long warps_at_8 = (long)((halves + 1) / 2) * stages_all * max_batch;  // warps if every unit were one stage
            const long stages_per_unit = std::max(1L, warps_at_8 / (slots * fill_x10 / 10));
            level_seg = (int)std::min<long>(seg_rows, stages_per_unit * PT_DROWS);
        }
        int segs = std::max(1, (d.h + level_seg / 2) / level_seg);
        const int stages = (stages_all + segs - 1) / segs;
        segs = (stages_all + stages - 1) / stages;
        for (int sg = 0; sg < segs; ++sg)
            for (int hb = 0; hb < halves; hb += 2) {
                PyrUnit u{};
                u.level = l;
                u.y0 = sg * stages * PT_DROWS;
                u.stages = std::min(stages, stages_all - sg * stages);
                u.strip0[0] = level_strip0 + 16 * hb;
                u.box_col[0] = box_col[hb];
                u.strip0[1] = hb + 1 < halves ? level_strip0 + 16 * (hb + 1) : -1;
                u.box_col[1] = hb + 1 < halves ? box_col[hb + 1] : 0;
                units.push_back(u);
                ++need[l];
            }
    }
    PyramidPlan* p = new PyramidPlan;
    p->num_levels = num_levels;
    p->max_batch = max_batch;
    p->n_units = (int)units.size();
    for (int l = 0; l + 1 < num_levels; ++l) {
        const LevelGeom& g = lv[l];
        const uint64_t dims[3] = {(uint64_t)g.pitch, (uint64_t)(g.h + 2 * kApron), (uint64_t)max_batch};
        const uint64_t strides[2] = {(uint64_t)g.pitch, (uint64_t)g.frame_stride};
        const uint32_t box[3] = {(uint32_t)PT_BOXW, (uint32_t)PT_SROWS, 1u};
        if (!tma_encode_u8(&p->maps.m[l], g.img - (size_t)kApron * g.pitch - kPadLeft, 3, dims, strides, box)) {
            pyramid_plan_destroy(p);
            return nullptr;
        }
    }
    bool ok = cudaMalloc(&p->units_dev, sizeof(PyrUnit) * units.size()) == cudaSuccess
              && cudaMalloc(&p->strips_dev, sizeof(PyrStrip) * strips.size()) == cudaSuccess
              && cudaMalloc(&p->done_dev, sizeof(uint32_t) * (size_t)max_batch * num_levels) == cudaSuccess
              && cudaMalloc(&p->need_dev, sizeof(uint32_t) * num_levels) == cudaSuccess;
    ok = ok && cudaMemcpy(p->units_dev, units.data(), sizeof(PyrUnit) * units.size(), cudaMemcpyHostToDevice) == cudaSuccess
         && cudaMemcpy(p->strips_dev, strips.data(), sizeof(PyrStrip) * strips.size(), cudaMemcpyHostToDevice) == cudaSuccess
         && cudaMemcpy(p->need_dev, need.data(), sizeof(uint32_t) * num_levels, cudaMemcpyHostToDevice) == cudaSuccess
         && cudaMemset(p->done_dev, 0, sizeof(uint32_t) * (size_t)max_batch * num_levels) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        pyramid_plan_destroy(p);
        return nullptr;
    }
    return p;
}

int launch_pyramid_tma(PyramidPlan* p, const LevelGeom* lv_dev, int batch, cudaStream_t st) {
    if (p->n_units == 0) return LPSLAM_OK;
    // the completion counters are monotonic: launch number `epoch` waits for epoch x (units per level). Frames beyond a
    // launch's batch keep their old counts, so the counters are re-based whenever the batch size changes
    ++p->epoch;
    if (batch != p->last_batch || p->epoch > 0x00FFFFFFu) {
        LPS_CUDA(cudaMemsetAsync(p->done_dev, 0, sizeof(uint32_t) * (size_t)p->max_batch * p->num_levels, st));
        p->epoch = 1u;
        p->last_batch = batch;
    }
    k_pyramid_tma<<<(unsigned)p->n_units * (unsigned)batch, 32, PT_WARP_BYTES + 128, st>>>(p->maps, lv_dev, p->units_dev, p->strips_dev, batch,
                                                                                        p->done_dev, p->need_dev, p->epoch, p->num_levels);
    return LPSLAM_OK;
}

}  // namespace lps
