// K2, dense variant — FAST-9/16 with non-max suppression on the reference's cell grid (feature/orb_extractor.cc:277-372), same
// contract as k_fast_cells (fast.cu): one block per (cell, frame), candidates in raster order into the cell's slot block,
// per-cell retry at min_fast_thr when nothing survives the suppression.
//
// Why a second version: k_fast_cells scores only the pixels that pass a compass pre-test, through a compacted list. On the bench
// images 25-45 % of all pixels pass, and the list machinery is most of the kernel: 32 byte loads per pixel pair (LSU data pipe at
// 77 %), shared-memory atomics for the corner list and the kept bitmap (15 % of all warp instructions), list building, two extra
// barriers per attempt: 99 thread instructions per pixel (profiles/r1_lines_k_fast_cells.txt). Here the corner strength is computed
// for EVERY pixel, branch-free:
//   * a thread owns 8 consecutive pixels of one row: the 7 window rows its 16 ring positions touch are 14 aligned LDS.64;
//   * ring pixels go into 16-bit x 2 lanes straight from those words (one PRMT per ring position and pixel pair, a funnel shift
//     first only where the two bytes straddle a word: 12 of 64 cases), no byte loads, no address arithmetic;
//   * the exact arc network (VIMNMX3.U16x2, issued at twice the LOP / PRMT rate on B200: tools/pipe_rates.cu) runs on the raw ring
//     values, the centre is subtracted from its two results;
//   * strengths are stored raw; suppression is one more dense pass: keep iff best > max(8 neighbours' best, thr, 1), which equals
//     the reference's "score strictly greater than all 8 neighbours' scores" with non-corners scoring 0;
//   * no list, no atomics on the hot path (kept pixels, ~3 % of all, set bits with atomicOr), three barriers per attempt.
#include "common.cuh"
#include <cstdlib>

namespace lps {

namespace {

constexpr int WIN = kCell + kOverlap;  // 70
constexpr int WP = 80;                 // staged row pitch: window column j at byte j + 3 (the 16-byte aligned global span starts 3 left)
constexpr int BPD = 96;                // strength map pitch: pixel lx at byte lx + 11, row ly at index ly + 1 (zero apron all round)
constexpr int FD_THREADS = 288;        // 32 rows x 9 strips per pass (strip t = pixels 8 t - 3 .. 8 t + 4)

__device__ __forceinline__ bool in_mask_d(const uint8_t* __restrict__ mask, size_t pitch, unsigned y, unsigned x, float sf) {
    const int my = (int)__fmul_rn((float)y, sf);  // feature/orb_extractor.cc:281-283
    const int mx = (int)__fmul_rn((float)x, sf);
    return mask[(size_t)my * pitch + mx] == 0;
}

// bytes (o, o + 1) of the 16-byte register window w[0..3] as two zero-extended 16-bit lanes; o is a compile-time constant
template<int O>
__device__ __forceinline__ uint32_t lanes(const uint32_t (&w)[4]) {
    constexpr int q = O >> 2, r = O & 3;
    if (r == 3) return __byte_perm(__funnelshift_r(w[q], w[q + 1 < 4 ? q + 1 : 3], 24), 0u, 0x4140u);
    return __byte_perm(w[q], 0u, r == 0 ? 0x4140u : (r == 1 ? 0x4241u : 0x4342u));
}

// exact corner strength of the pixel pair whose first pixel sits at window byte C0 of the thread's 16-byte row windows
template<int C0>
__device__ __forceinline__ uint32_t strength_pair(const uint32_t (&rows)[7][4]) {
    uint32_t d[16];
    // ring offsets (dx, dy), k = 0 .. 15; rows[] index = dy + 3
    d[0] = lanes<C0 + 0>(rows[6]);   d[1] = lanes<C0 + 1>(rows[6]);   d[2] = lanes<C0 + 2>(rows[5]);   d[3] = lanes<C0 + 3>(rows[4]);
    d[4] = lanes<C0 + 3>(rows[3]);   d[5] = lanes<C0 + 3>(rows[2]);   d[6] = lanes<C0 + 2>(rows[1]);   d[7] = lanes<C0 + 1>(rows[0]);
    d[8] = lanes<C0 + 0>(rows[0]);   d[9] = lanes<C0 - 1>(rows[0]);   d[10] = lanes<C0 - 2>(rows[1]);  d[11] = lanes<C0 - 3>(rows[2]);
    d[12] = lanes<C0 - 3>(rows[3]);  d[13] = lanes<C0 - 3>(rows[4]);  d[14] = lanes<C0 - 2>(rows[5]);  d[15] = lanes<C0 - 1>(rows[6]);
    const uint32_t c2 = lanes<C0>(rows[3]);
    // min / max over every window of 9 consecutive ring values = 3-input min / max of three triples
    uint32_t lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = __vimin3_u16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
        hi3[k] = __vimax3_u16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
    }
    uint32_t A = 0u, Bm = 0xFFFFFFFFu;
#pragma unroll
    for (int k = 0; k < 16; k += 2) {
        const uint32_t la = __vimin3_u16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
        const uint32_t lb = __vimin3_u16x2(lo3[k + 1], lo3[(k + 4) & 15], lo3[(k + 7) & 15]);
        const uint32_t ha = __vimax3_u16x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
        const uint32_t hb = __vimax3_u16x2(hi3[k + 1], hi3[(k + 4) & 15], hi3[(k + 7) & 15]);
        A = __vimax3_u16x2(A, la, lb);
        Bm = __vimin3_u16x2(Bm, ha, hb);
    }
    // A = brightest "darkest pixel of an arc", Bm = darkest "brightest pixel of an arc": best = max(A - c, c - Bm), floored at 0
    const uint32_t pb = __vmaxu2(A, c2) - c2, pd = c2 - __vminu2(Bm, c2);  // no borrow between the lanes: both differences are >= 0
    return __vmaxu2(pb, pd);
}

__global__ void __launch_bounds__(FD_THREADS, 3) k_fast_dense(FastArgs a) {
    __shared__ __align__(16) uint8_t s_win[WIN][WP];
    __shared__ __align__(16) uint8_t s_best[kCell + 2][BPD];
    __shared__ uint32_t s_bits[kCell][2];
    __shared__ uint32_t s_rowbase[kCell + 1];
    __shared__ int s_total;

    const int cell_id = blockIdx.x, frame = blockIdx.y + a.frame0, tid = threadIdx.x;
    const CellDesc c = a.cells[cell_id];
    const LevelGeom lv = a.lv[c.level];
    uint32_t* count_out = a.cell_counts + (size_t)frame * a.num_cells + cell_id;
    uint32_t* slots = a.cell_slots + ((size_t)frame * a.num_cells + cell_id) * kCellSlots;

    if (a.mask) {  // cell skipped when one of its four window corners is masked (orb_extractor.cc:334-340)
        const unsigned x0 = c.x0, y0 = c.y0, x1 = c.x0 + c.cw, y1 = c.y0 + c.ch;
        if (in_mask_d(a.mask, a.mask_pitch, y0, x0, lv.scale) || in_mask_d(a.mask, a.mask_pitch, y1, x0, lv.scale)
            || in_mask_d(a.mask, a.mask_pitch, y0, x1, lv.scale) || in_mask_d(a.mask, a.mask_pitch, y1, x1, lv.scale)) {
            if (tid == 0) *count_out = 0;
            return;
        }
    }
    const uint8_t* __restrict__ img = lv.img + frame * lv.frame_stride;
    for (int idx = tid; idx < WIN * (WP / 16); idx += FD_THREADS) {  // the 16-byte aligned span [x0 - 3, x0 + 77) of every window row
        const int r = idx / (WP / 16), q = idx % (WP / 16);
        uint4 v = make_uint4(0, 0, 0, 0);
        const int gx = c.x0 - 3 + 16 * q;
        if (r < c.ch && gx + 16 <= lv.pitch - kPadLeft) v = __ldg(reinterpret_cast<const uint4*>(img + (size_t)(c.y0 + r) * lv.pitch + gx));
        *reinterpret_cast<uint4*>(&s_win[r][16 * q]) = v;
    }
    for (int idx = tid; idx < (kCell + 2) * BPD / 16; idx += FD_THREADS) reinterpret_cast<uint4*>(&s_best[0][0])[idx] = make_uint4(0, 0, 0, 0);
    __syncthreads();

    const int vw = c.cw - 6, vh = c.ch - 6;  // valid interior extent (<= 64)
    const int row_in_pass = tid / 9, t = tid - 9 * row_in_pass;
    // byte mask of the strip's valid pixels (lx = 8 t - 3 + j in [0, vw))
    uint32_t vm0 = 0, vm1 = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int lx = 8 * t - 3 + j;
        const uint32_t on = (lx >= 0 && lx < vw) ? 0xFFu : 0u;
        if (j < 4) vm0 |= on << (8 * j);
        else vm1 |= on << (8 * (j - 4));
    }

    // ---- pass A: corner strength of every pixel (threshold independent) ----
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        const int ly = row_in_pass + 32 * pass;
        if (ly >= vh) continue;
        uint32_t rows[7][4];
#pragma unroll
        for (int dy = 0; dy < 7; ++dy) {
            const uint2 lo = *reinterpret_cast<const uint2*>(&s_win[ly + dy][8 * t]);
            const uint2 hi = *reinterpret_cast<const uint2*>(&s_win[ly + dy][8 * t + 8]);
            rows[dy][0] = lo.x, rows[dy][1] = lo.y, rows[dy][2] = hi.x, rows[dy][3] = hi.y;
        }
        // strip pixel j sits at window byte 3 + j of the 16-byte row windows
        const uint32_t b01 = strength_pair<3>(rows), b23 = strength_pair<5>(rows), b45 = strength_pair<7>(rows), b67 = strength_pair<9>(rows);
        const uint32_t o0 = __byte_perm(b01, b23, 0x6420u) & vm0, o1 = __byte_perm(b45, b67, 0x6420u) & vm1;
        *reinterpret_cast<uint2*>(&s_best[ly + 1][8 * t + 8]) = make_uint2(o0, o1);
    }
    __syncthreads();

    // ---- suppression + ordered counts: at ini_fast_thr, then once more at min_fast_thr if the cell came out empty ----
    int thr = (int)a.ini_thr;
    int total = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        if (tid < 2 * kCell) (&s_bits[0][0])[tid] = 0;
        __syncthreads();
        const uint32_t floor2 = (uint32_t)max(thr, 1) * 0x00010001u;
#pragma unroll 1
        for (int pass = 0; pass < 2; ++pass) {
            const int ly = row_in_pass + 32 * pass;
            if (ly >= vh) continue;
            uint32_t w[3][4];  // rows ly - 1, ly, ly + 1 of the strength map, bytes 8 t + 4 .. 8 t + 19: pixel j of the strip at byte 4 + j
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const uint8_t* p = &s_best[ly + r][8 * t + 4];
                w[r][0] = *reinterpret_cast<const uint32_t*>(p);
                const uint2 m = *reinterpret_cast<const uint2*>(p + 4);
                w[r][1] = m.x, w[r][2] = m.y;
                w[r][3] = *reinterpret_cast<const uint32_t*>(p + 12);
            }
            if (((w[1][1] | w[1][2]) == 0u)) continue;  // no strength at all in this strip (flat image regions)
            uint32_t bits = 0;
#define FD_NMS(P)                                                                                               \
    {                                                                                                            \
        const uint32_t up = __vimax3_u16x2(lanes<3 + 2 * P>(w[0]), lanes<4 + 2 * P>(w[0]), lanes<5 + 2 * P>(w[0])); \
        const uint32_t dn = __vimax3_u16x2(lanes<3 + 2 * P>(w[2]), lanes<4 + 2 * P>(w[2]), lanes<5 + 2 * P>(w[2])); \
        const uint32_t sd = __vmaxu2(lanes<3 + 2 * P>(w[1]), lanes<5 + 2 * P>(w[1]));                            \
        const uint32_t lim = __vimax3_u16x2(__vmaxu2(up, dn), sd, floor2);                                       \
        const uint32_t x = __vmaxu2(lanes<4 + 2 * P>(w[1]), lim) ^ lim; /* non-zero lane = strictly greater */   \
        bits |= ((x & 0xFFFFu) ? 1u : 0u) << (2 * P);                                                            \
        bits |= ((x >> 16) ? 2u : 0u) << (2 * P);                                                                \
    }
            FD_NMS(0) FD_NMS(1) FD_NMS(2) FD_NMS(3)
#undef FD_NMS
            if (bits) {  // ~3 % of the pixels survive: the atomics are rare
                // pixel j of the strip is lx = 8 t - 3 + j; invalid pixels hold strength 0 and are never kept
                const unsigned long long v = t < 8 ? (((unsigned long long)bits << (8 * t)) >> 3) : ((unsigned long long)(bits & 7u) << 61);
                const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
                if (lo) atomicOr(&s_bits[ly][0], lo);
                if (hi) atomicOr(&s_bits[ly][1], hi);
            }
        }
        __syncthreads();
        if (tid < 32) {  // exclusive scan of the 64 row counts by one warp
            const int lane = tid;
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
        thr = (int)a.min_thr;
    }

    // ---- ordered emission (raster order inside the cell), own-pixel mask applied after the emptiness decision (orb_extractor.cc:352-369) ----
    if (!a.mask) {
        for (int idx = tid; idx < kCell * 2; idx += FD_THREADS) {
            const int ly = idx >> 1, half = idx & 1;
            uint32_t m = s_bits[ly][half];
            int base = s_rowbase[ly] + (half ? __popc(s_bits[ly][0]) : 0);
            while (m) {
                const int bit = __ffs(m) - 1;
                m &= m - 1;
                const int lx = half * 32 + bit;
                slots[base++] = (uint32_t)lx | ((uint32_t)ly << 8) | ((uint32_t)(s_best[ly + 1][lx + 11] - 1) << 16);
            }
        }
        if (tid == 0) *count_out = (uint32_t)total;
    }
    else if (tid < 32) {  // rare path: one warp walks the rows in order and drops masked candidates
        const int lane = tid;
        int n = 0;
        for (int ly = 0; ly < kCell; ++ly) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int lx = half * 32 + lane;
                bool keep = (s_bits[ly][half] >> lane) & 1u;
                if (keep) {
                    const unsigned ax = c.x0 + 3 + lx, ay = c.y0 + 3 + ly;  // absolute level coordinates (orb_extractor.cc:362-367)
                    if (in_mask_d(a.mask, a.mask_pitch, ay, ax, lv.scale)) keep = false;
                }
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, keep);
                if (keep) slots[n + __popc(m & ((1u << lane) - 1))] = (uint32_t)lx | ((uint32_t)ly << 8) | ((uint32_t)(s_best[ly + 1][lx + 11] - 1) << 16);
                n += __popc(m);
            }
        }
        if (lane == 0) *count_out = (uint32_t)n;
    }
}

}  // namespace

void launch_fast_dense(const FastArgs& a, cudaStream_t st, int priority) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.num_cells, a.batch);
    cfg.blockDim = dim3(FD_THREADS);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributePriority;
    at[0].val.priority = priority;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_fast_dense, a);
}

}  // namespace lps
