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

__global__ void __launch_bounds__(256) k_fast_cells(FastArgs a) {
    __shared__ __align__(16) uint8_t s_win[WIN][WP];
    __shared__ __align__(16) uint8_t s_best[kCell + 2][BP];  // [1 + ly][1 + lx], zero apron
    __shared__ uint32_t s_bits[kCell][2];      // kept-pixel bitmap per interior row
    __shared__ uint32_t s_rowbase[kCell + 1];
    __shared__ int s_total;
    __shared__ int s_nlist;
    __shared__ uint16_t s_list[kCell * kCell];

    const int cell_id = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
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
        if (r < c.ch && gx + 16 <= lv.pitch) v = __ldg(reinterpret_cast<const uint4*>(img + (size_t)(c.y0 + r) * lv.pitch + gx));
        *reinterpret_cast<uint4*>(&s_win[r][16 * q]) = v;
    }
    for (int idx = tid; idx < (kCell + 2) * BP / 4; idx += 256) reinterpret_cast<uint32_t*>(&s_best[0][0])[idx] = 0;
    if (tid == 0) s_nlist = 0;
    __syncthreads();

    // ---- best for every interior pixel with a full ring inside the window ----
    // Pass 1 (cheap, divergent): four compass pixels. An arc of 9 always holds two ADJACENT compass points, so a corner
    // at threshold t needs such a pair both brighter or both darker than the centre by more than t; everything else
    // has best <= t and keeps the zero already in s_best. Survivors are compacted into a list so that pass 2 (the
    // full 16-pixel ring, ~200 integer ops) runs with every lane busy.
    const int vw = c.cw - 6, vh = c.ch - 6;  // valid interior extent (<= 64)
    const int quick_thr = (int)min(a.ini_thr, a.min_thr);
    {
        const int lx = tid & 63;
        const int lane_ = tid & 31;
        for (int ly = tid >> 6; ly < kCell; ly += 4) {
            bool pass = false;
            if (lx < vw && ly < vh) {
                const uint8_t* p = &s_win[ly + 3][lx + 6];
                const int v = p[0];
                const int d0 = v - p[3 * WP], d4 = v - p[3], d8 = v - p[-3 * WP], d12 = v - p[-3];
                const unsigned hi = (d0 > quick_thr) | ((d4 > quick_thr) << 1) | ((d8 > quick_thr) << 2) | ((d12 > quick_thr) << 3);
                const unsigned lo = (d0 < -quick_thr) | ((d4 < -quick_thr) << 1) | ((d8 < -quick_thr) << 2) | ((d12 < -quick_thr) << 3);
                // adjacent pairs (0,4) (4,8) (8,12) (12,0): bit i and bit (i+1) mod 4
                const unsigned hr = ((hi >> 1) | (hi << 3)) & 15u, lr = ((lo >> 1) | (lo << 3)) & 15u;
                pass = ((hi & hr) | (lo & lr)) != 0;
            }
            const unsigned m = __ballot_sync(0xFFFFFFFFu, pass);
            if (m) {
                int base = 0;
                if (lane_ == 0) base = atomicAdd(&s_nlist, __popc(m));
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (pass) s_list[base + __popc(m & ((1u << lane_) - 1))] = (uint16_t)(ly * 64 + lx);
            }
        }
    }
    __syncthreads();
    {
        const int n = s_nlist;
        for (int i = tid; i < n; i += 256) {
            const int id = s_list[i], lx = id & 63, ly = id >> 6;
            const uint8_t* p = &s_win[ly + 3][lx + 6];
            const int v = p[0];
            int d[16];
            d[0] = v - p[3 * WP];
            d[1] = v - p[3 * WP + 1];
            d[2] = v - p[2 * WP + 2];
            d[3] = v - p[WP + 3];
            d[4] = v - p[3];
            d[5] = v - p[-WP + 3];
            d[6] = v - p[-2 * WP + 2];
            d[7] = v - p[-3 * WP + 1];
            d[8] = v - p[-3 * WP];
            d[9] = v - p[-3 * WP - 1];
            d[10] = v - p[-2 * WP - 2];
            d[11] = v - p[-WP - 3];
            d[12] = v - p[-3];
            d[13] = v - p[WP - 3];
            d[14] = v - p[2 * WP - 2];
            d[15] = v - p[3 * WP - 1];
            const int pos = arc9_maxmin(d);
#pragma unroll
            for (int k = 0; k < 16; ++k) d[k] = -d[k];
            const int neg = arc9_maxmin(d);
            s_best[ly + 1][lx + 1] = (uint8_t)max(max(pos, neg), 0);
        }
    }
    __syncthreads();

    // ---- suppression at `thr`, retried once at the lower threshold when nothing survives ----
    const int warp = tid >> 5, lane = tid & 31;
    int thr = (int)a.ini_thr;
    int total = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        // warp w handles rows w, w+8, ...; two ballots (column halves) per row
        for (int ly = warp; ly < kCell; ly += 8) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int lx = half * 32 + lane;
                const uint8_t* b = &s_best[ly + 1][lx + 1];
                const int best = b[0];
                bool keep = false;
                if (best > thr) {
                    const int s = best - 1;
                    auto sc = [thr](int v) { return v > thr ? v - 1 : 0; };
                    keep = s > sc(b[-1]) && s > sc(b[1]) && s > sc(b[-BP - 1]) && s > sc(b[-BP]) && s > sc(b[-BP + 1])
                           && s > sc(b[BP - 1]) && s > sc(b[BP]) && s > sc(b[BP + 1]);
                }
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
                if (lane == 0) s_bits[ly][half] = m;
            }
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
        if (total > 0 || (int)a.min_thr == thr) break;
        thr = (int)a.min_thr;
        __syncthreads();
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
    k_fast_cells<<<dim3(a.num_cells, a.batch), 256, 0, st>>>(a);
}

}  // namespace lps
