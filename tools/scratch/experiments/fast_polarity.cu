// K2 — FAST-9/16 with non-max suppression on the reference's cell grid.
// Replaces the per-cell cv::FAST calls of feature/orb_extractor.cc:277-372 (cells of 64 px + 6 px overlap starting at the
// 19-px border, threshold ini_fast_thr with a per-cell retry at min_fast_thr when the cell is empty AFTER suppression).
// Arithmetic: best = max over the 16 arcs of 9 contiguous ring pixels of min(+-(centre - ring));
// corner iff best > thr; score = best - 1; suppression is strict against the 8 neighbours' scores and never looks outside
// the cell's own window (scores there count as 0).
//
// One block per (cell, frame): the 70x70 window is staged in shared memory, `best` is computed once for all 64x64
// interior pixels (it does not depend on the threshold), then suppression + ordered compaction run from shared memory,
// a second time at the lower threshold if the first pass left the cell empty. Candidates leave in raster order into
// the cell's slot block (<= 1024 entries: strict 8-neighbourhood maxima cannot be denser than one per 2x2).
#include "common.cuh"
#include <cstdlib>

namespace lps {

namespace {

constexpr int WIN = kCell + kOverlap;  // 70
constexpr int WP = 80;                 // staged row pitch (bytes)
constexpr int BP = 68;                 // best-map row pitch: 64 + one zero column each side, padded

__device__ __forceinline__ bool in_mask(const uint8_t* __restrict__ mask, size_t pitch, unsigned y, unsigned x, float sf) {
    // feature/orb_extractor.cc:281-283: mask.at<uchar>(y * scale, x * scale) with float products truncated toward zero
    const int my = (int)__fmul_rn((float)y, sf);
    const int mx = (int)__fmul_rn((float)x, sf);
    return mask[(size_t)my * pitch + mx] == 0;
}

// max over the 16 circular windows of 9 of the minimum of d[] (doubling: 2, 4, 8, then +1)
__device__ __forceinline__ int arc9_maxmin(const int (&d)[16]) {
    int m2[16], m4[16], best = -512;
#pragma unroll
    for (int i = 0; i < 16; ++i) m2[i] = min(d[i], d[(i + 1) & 15]);
#pragma unroll
    for (int i = 0; i < 16; ++i) m4[i] = min(m2[i], m2[(i + 2) & 15]);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int m8 = min(m4[i], m4[(i + 4) & 15]);
        best = max(best, min(m8, d[(i + 8) & 15]));
    }
    return best;
}

#ifndef LPS_FAST_MINBLOCKS
#define LPS_FAST_MINBLOCKS 6
#endif
__global__ void __launch_bounds__(256, LPS_FAST_MINBLOCKS) k_fast_cells(FastArgs a) {
    __shared__ __align__(16) uint8_t s_win[WIN][WP];
    __shared__ __align__(16) uint8_t s_best[kCell + 2][BP];  // [1 + ly][1 + lx], zero apron
    __shared__ uint32_t s_bits[kCell][2];      // kept-pixel bitmap per interior row
    __shared__ uint32_t s_rowbase[kCell + 1];
    __shared__ int s_total;
    __shared__ int s_nlist;
    __shared__ uint32_t s_pb[kCell][2], s_pd[kCell][2];  // pre-test survivors: a bright / a dark arc is possible
    __shared__ int s_wbase[2 * kCell];
    __shared__ uint16_t s_list[2 * kCell * kCell];  // pixel | polarity << 12; a pixel that passed both ways is listed twice

    const int cell_id = blockIdx.x, frame = blockIdx.y + a.frame0, tid = threadIdx.x;
    const CellDesc c = a.cells[cell_id];
    const LevelGeom lv = a.lv[c.level];
    uint32_t* count_out = a.cell_counts + (size_t)frame * a.num_cells + cell_id;
    uint32_t* slots = a.cell_slots + ((size_t)frame * a.num_cells + cell_id) * kCellSlots;

    // cell skipped when one of its four window corners is masked (orb_extractor.cc:334-340)
    if (a.mask) {
        const unsigned x0 = c.x0, y0 = c.y0, x1 = c.x0 + c.cw, y1 = c.y0 + c.ch;
        if (in_mask(a.mask, a.mask_pitch, y0, x0, lv.scale) || in_mask(a.mask, a.mask_pitch, y1, x0, lv.scale)
            || in_mask(a.mask, a.mask_pitch, y0, x1, lv.scale) || in_mask(a.mask, a.mask_pitch, y1, x1, lv.scale)) {
            if (tid == 0) *count_out = 0;
            return;
        }
    }

    const uint8_t* __restrict__ img = lv.img + frame * lv.frame_stride;
    // ---- stage the window: the 16-byte aligned span [x0 - 3, x0 + 77) of every window row (x0 = 19 + 64 j) ----
    for (int idx = tid; idx < WIN * (WP / 16); idx += 256) {
        const int r = idx / (WP / 16), q = idx % (WP / 16);
        uint4 v = make_uint4(0, 0, 0, 0);
        const int gx = c.x0 - 3 + 16 * q;
        if (r < c.ch && gx + 16 <= lv.pitch - kPadLeft) v = __ldg(reinterpret_cast<const uint4*>(img + (size_t)(c.y0 + r) * lv.pitch + gx));
        *reinterpret_cast<uint4*>(&s_win[r][16 * q]) = v;
    }
    for (int idx = tid; idx < (kCell + 2) * BP / 4; idx += 256) reinterpret_cast<uint32_t*>(&s_best[0][0])[idx] = 0;
    if (tid < 2 * kCell) (&s_bits[0][0])[tid] = 0;
    __syncthreads();

    // ---- corner strength, suppression, ordered counts: at ini_fast_thr, then once more at min_fast_thr if the cell is empty ----
    // Per attempt (threshold T):
    //   1. compass pre-test, 4 pixels per thread as packed 16-bit lanes: an arc of 9 always holds two ADJACENT compass
    //      points, so best > T needs max over the 4 adjacent pairs of min(d_i, d_i+1) > T or min of max < -T.
    //      Survivors go into a bitmap, the bitmap into a compact list.
    //   2. exact best for listed pixels only, two pixels per thread in 16x2 SIMD (VIMNMX.U16x2), every lane busy.
    //   3. non-max suppression only around listed pixels with best > T (everything unlisted scores 0).
    const int vw = c.cw - 6, vh = c.ch - 6;  // valid interior extent (<= 64)
    const int warp = tid >> 5, lane = tid & 31;
    int thr = (int)a.ini_thr;
    int total = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        {   // 1a. pre-test -> s_pass
            const int w = tid & 15, lx = 4 * w;
            const uint32_t kT = (uint32_t)(0x8000 - 1 - thr) * 0x00010001u;  // bit15 of (x + kT) set iff half x >= thr + 1
            const int nv = vw - lx;
            const uint32_t vmask = nv >= 4 ? 0xFu : (nv <= 0 ? 0u : ((1u << nv) - 1u));
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                const int ly = (tid >> 4) + 16 * it;
                uint32_t nibb = 0, nibd = 0;
                if (ly < vh && vmask) {
                    const uint32_t* rc = reinterpret_cast<const uint32_t*>(&s_win[ly + 3][0]) + w;
                    const uint32_t* rn = reinterpret_cast<const uint32_t*>(&s_win[ly + 6][0]) + w;
                    const uint32_t* rs = reinterpret_cast<const uint32_t*>(&s_win[ly][0]) + w;
                    const uint32_t w0 = rc[0], w1 = rc[1], w2 = rc[2], w3 = rc[3];
                    // pixel lx sits at staged byte lx + 6
                    const uint32_t C = __funnelshift_r(w1, w2, 16), E = __funnelshift_r(w2, w3, 8), Wt = __funnelshift_r(w0, w1, 24);
                    const uint32_t N = __funnelshift_r(rn[1], rn[2], 16), S = __funnelshift_r(rs[1], rs[2], 16);
                    uint32_t fb[2], fd[2];
#pragma unroll
                    for (int hf = 0; hf < 2; ++hf) {
                        const uint32_t sel = hf ? 0x4342u : 0x4140u;
                        // raw 16-bit lanes: M = brightest "darker one of two adjacent compass pixels", X = darkest
                        // "brighter one"; bright arc possible iff M - c > thr, dark arc iff c - X > thr
                        const uint32_t c16 = __byte_perm(C, 0, sel);
                        const uint32_t pn = __byte_perm(N, 0, sel), pe = __byte_perm(E, 0, sel);
                        const uint32_t ps = __byte_perm(S, 0, sel), pw = __byte_perm(Wt, 0, sel);
                        const uint32_t M = __vmaxu2(__vimax3_u16x2(__vminu2(pn, pe), __vminu2(pe, ps), __vminu2(ps, pw)), __vminu2(pw, pn));
                        const uint32_t X = __vminu2(__vimin3_u16x2(__vmaxu2(pn, pe), __vmaxu2(pe, ps), __vmaxu2(ps, pw)), __vmaxu2(pw, pn));
                        fb[hf] = (M + kT - c16) & 0x80008000u;
                        fd[hf] = (c16 + kT - X) & 0x80008000u;
                    }
                    nibb = (((fb[0] >> 15) & 1u) | ((fb[0] >> 30) & 2u) | ((fb[1] >> 13) & 4u) | ((fb[1] >> 28) & 8u)) & vmask;
                    nibd = (((fd[0] >> 15) & 1u) | ((fd[0] >> 30) & 2u) | ((fd[1] >> 13) & 4u) | ((fd[1] >> 28) & 8u)) & vmask;
                }
                // the eight lanes of a 32-pixel word OR their nibbles together (REDUX)
                const uint32_t group = 0xFFu << (lane & 24);
                const uint32_t wordb = __reduce_or_sync(group, nibb << (4 * (w & 7)));
                const uint32_t wordd = __reduce_or_sync(group, nibd << (4 * (w & 7)));
                if ((w & 7) == 0) {
                    s_pb[ly][w >> 3] = wordb;
                    s_pd[ly][w >> 3] = wordd;
                }
            }
        }
        __syncthreads();
        if (warp == 0) {  // 1b. bitmap -> per-word list offsets
            const uint32_t* pb = &s_pb[0][0];
            const uint32_t* pd = &s_pd[0][0];
            const int c0 = __popc(pb[4 * lane]) + __popc(pd[4 * lane]), c1 = __popc(pb[4 * lane + 1]) + __popc(pd[4 * lane + 1]);
            const int c2 = __popc(pb[4 * lane + 2]) + __popc(pd[4 * lane + 2]), c3 = __popc(pb[4 * lane + 3]) + __popc(pd[4 * lane + 3]);
            int inc = c0 + c1 + c2 + c3;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += t;
            }
            int ex = inc - (c0 + c1 + c2 + c3);
            s_wbase[4 * lane] = ex;
            s_wbase[4 * lane + 1] = ex + c0;
            s_wbase[4 * lane + 2] = ex + c0 + c1;
            s_wbase[4 * lane + 3] = ex + c0 + c1 + c2;
            if (lane == 31) s_nlist = inc;
        }
        __syncthreads();
        if (tid < 128) {
            uint32_t mb = (&s_pb[0][0])[tid], md = (&s_pd[0][0])[tid];
            int o = s_wbase[tid];
            const int idbase = (tid >> 1) * 64 + (tid & 1) * 32;
            while (mb) {  // entry = pixel, bright arc possible
                const int bit = __ffs(mb) - 1;
                mb &= mb - 1;
                s_list[o++] = (uint16_t)(idbase + bit);
            }
            while (md) {  // entry = pixel | 1 << 12, dark arc possible
                const int bit = __ffs(md) - 1;
                md &= md - 1;
                s_list[o++] = (uint16_t)((idbase + bit) | 0x1000);
            }
        }
        __syncthreads();
        const int n = s_nlist;
        // 2. exact best, two listed pixels per thread. A warp takes 64 consecutive list entries per round and a lane the
        //    entries (lane, lane + 32) of them: the 32 pixels one shared-memory byte load touches are then neighbours in
        //    raster order (two or three window rows, few bank conflicts) instead of every other entry of a 64-entry run.
        //    The arc network runs on the raw ring values; the centre is subtracted from its two results only.
        for (int base = 64 * warp; base < n; base += 64 * 8) {
            const int ia = min(base + lane, n - 1), ib = min(base + 32 + lane, n - 1);
            const bool va = base + lane < n, vb = base + 32 + lane < n;
            const int ea = s_list[ia], eb = s_list[ib];
            const int ida = ea & 0xFFF, idb = eb & 0xFFF;
            const uint8_t* pa = &s_win[(ida >> 6) + 3][(ida & 63) + 6];
            const uint8_t* pb = &s_win[(idb >> 6) + 3][(idb & 63) + 6];
            // Only the polarity the pre-test passed is scored: the other one cannot reach the threshold, and a corner's strength is
            // the passing polarity's (an arc of 9 brighter and one of 9 darker pixels do not fit a ring of 16). A dark candidate is
            // scored as a bright one on the inverted pixels (255 - p), so ONE network serves both: the inversion rides in the two
            // IMADs that pack the pixel pair. A pixel that passed both ways is in the list twice.
            const bool dka = (ea >> 12) != 0, dkb = (eb >> 12) != 0;
            const int sa = dka ? -1 : 1, sb = dkb ? -65536 : 65536;
            const uint32_t bias = (dka ? 255u : 0u) | (dkb ? 255u << 16 : 0u);
            int ca = pa[0], cb = pb[0];
            uint32_t d[16];
#define RING(k, off) d[k] = (uint32_t)((int)pb[off] * sb + ((int)pa[off] * sa + (int)bias))
            RING(0, 3 * WP);      RING(1, 3 * WP + 1);   RING(2, 2 * WP + 2);   RING(3, WP + 3);
            RING(4, 3);           RING(5, -WP + 3);      RING(6, -2 * WP + 2);  RING(7, -3 * WP + 1);
            RING(8, -3 * WP);     RING(9, -3 * WP - 1);  RING(10, -2 * WP - 2); RING(11, -WP - 3);
            RING(12, -3);         RING(13, WP - 3);      RING(14, 2 * WP - 2);  RING(15, 3 * WP - 1);
#undef RING
            // max over every window of 9 consecutive ring values of its minimum = 3-input min of three triples, twice
            auto arcs = [](const uint32_t (&v)[16]) {
                uint32_t lo3[16], lo9[16];
#pragma unroll
                for (int k = 0; k < 16; ++k) lo3[k] = __vimin3_u16x2(v[k], v[(k + 1) & 15], v[(k + 2) & 15]);
#pragma unroll
                for (int k = 0; k < 16; ++k) lo9[k] = __vimin3_u16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
                uint32_t A = __vimax3_u16x2(lo9[0], lo9[1], lo9[2]);
#pragma unroll
                for (int k = 3; k < 15; k += 2) A = __vimax3_u16x2(A, lo9[k], lo9[k + 1]);
                return __vmaxu2(A, lo9[15]);
            };
            const uint32_t A = arcs(d);
            // A = brightest "darkest pixel of an arc" of the (possibly inverted) ring: strength = A - centre (inverted alike)
            const int ba = (int)(A & 0xFFFFu) - (dka ? 255 - ca : ca);
            const int bb = (int)(A >> 16) - (dkb ? 255 - cb : cb);
            // s_best starts at 0 and only corners are written (at most one polarity of a pixel is one; a retry at the lower
            // threshold finds the first attempt's corners again, with the same strength)
            if (va && ba > thr) s_best[(ida >> 6) + 1][(ida & 63) + 1] = (uint8_t)ba;
            if (vb && bb > thr) s_best[(idb >> 6) + 1][(idb & 63) + 1] = (uint8_t)bb;
        }
        __syncthreads();
        // 3. suppression around the corners among the listed pixels (every other pixel scores 0)
        for (int i = tid; i < n; i += 256) {
            const int id = s_list[i] & 0xFFF, lx = id & 63, ly = id >> 6;
            const uint8_t* b = &s_best[ly + 1][lx + 1];
            const int own = b[0];
            if (own <= 1) continue;  // not a corner; a strength of 1 (threshold 0) scores 0 and is never kept
            // score = best - 1 for corners, 0 otherwise: strictly greater than all eight neighbours' scores
            const int nb = max(max(max((int)b[-1], (int)b[1]), max((int)b[-BP - 1], (int)b[-BP])),
                               max(max((int)b[-BP + 1], (int)b[BP - 1]), max((int)b[BP], (int)b[BP + 1])));
            if (own > nb) atomicOr(&s_bits[ly][lx >> 5], 1u << (lx & 31));
        }
        __syncthreads();
        if (tid < 32) {
            // exclusive scan of the 64 row counts by one warp
            const int c0 = __popc(s_bits[lane][0]) + __popc(s_bits[lane][1]);
            const int c1 = __popc(s_bits[lane + 32][0]) + __popc(s_bits[lane + 32][1]);
            int i0 = c0, i1 = c1;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t0 = __shfl_up_sync(0xFFFFFFFFu, i0, o), t1 = __shfl_up_sync(0xFFFFFFFFu, i1, o);
                if (lane >= o) {
                    i0 += t0;
                    i1 += t1;
                }
            }
            const int sum0 = __shfl_sync(0xFFFFFFFFu, i0, 31);
            s_rowbase[lane] = i0 - c0;
            s_rowbase[lane + 32] = sum0 + i1 - c1;
            if (lane == 31) {
                s_rowbase[64] = sum0 + i1;
                s_total = sum0 + i1;
            }
        }
        __syncthreads();
        total = s_total;
        if (total > 0 || (int)a.min_thr >= thr) break;
        thr = (int)a.min_thr;  // nothing survived: every s_bits word is still zero, s_best is recomputed as a superset
    }

    // ---- ordered emission (raster order inside the cell), own-pixel mask applied after the emptiness decision ----
    // (orb_extractor.cc:352-369). Masked candidates leave holes, so with a mask the slot list is compacted by a second scan.
    if (!a.mask) {
        for (int idx = tid; idx < kCell * 2; idx += 256) {
            const int ly = idx >> 1, half = idx & 1;
            uint32_t m = s_bits[ly][half];
            int base = s_rowbase[ly] + (half ? __popc(s_bits[ly][0]) : 0);
            while (m) {
                const int bit = __ffs(m) - 1;
                m &= m - 1;
                const int lx = half * 32 + bit;
                slots[base++] = (uint32_t)lx | ((uint32_t)ly << 8) | ((uint32_t)(s_best[ly + 1][lx + 1] - 1) << 16);
            }
        }
        if (tid == 0) *count_out = (uint32_t)total;
    }
    else {
        // rare path: single warp walks the rows in order and drops masked candidates
        if (warp == 0) {
            int n = 0;
            for (int ly = 0; ly < kCell; ++ly) {
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int lx = half * 32 + lane;
                    bool keep = (s_bits[ly][half] >> lane) & 1u;
                    if (keep) {
                        // absolute level coordinates: window origin + 3 + local (orb_extractor.cc:362-367)
                        const unsigned ax = c.x0 + 3 + lx, ay = c.y0 + 3 + ly;
                        if (in_mask(a.mask, a.mask_pitch, ay, ax, lv.scale)) keep = false;
                    }
                    const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
                    if (keep)
                        slots[n + __popc(m & ((1u << lane) - 1))] =
                            (uint32_t)lx | ((uint32_t)ly << 8) | ((uint32_t)(s_best[ly + 1][lx + 1] - 1) << 16);
                    n += __popc(m);
                }
            }
            if (lane == 0) *count_out = (uint32_t)n;
        }
    }
}

}  // namespace

void launch_fast(const FastArgs& a, cudaStream_t st) {
    if (a.num_cells == 0) return;
    // The context's streams are created at the device's greatest priority (api_orb.cu); FAST, the one issue-bound kernel of
    // the step, is launched at the least: when several contexts run side by side, another context's latency-bound tail
    // (quadtree, descriptors, stereo) gets the SM slots FAST blocks free before further FAST blocks do.
    static const int low_prio = [] {
        int least = 0, greatest = 0;
        cudaDeviceGetStreamPriorityRange(&least, &greatest);
        const char* env = getenv("LPSLAM_FAST_LOW_PRIORITY");
        return (env && atoi(env) == 0) ? greatest : least;
    }();
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.num_cells, a.batch);
    cfg.blockDim = dim3(256);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributePriority;
    at[0].val.priority = low_prio;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_fast_cells, a);
}

}  // namespace lps
