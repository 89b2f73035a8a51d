// K1 — image pyramid. Replaces cv::resize(..., INTER_LINEAR) at feature/orb_extractor.cc:266-275.
// Arithmetic: 11-bit fixed-point coefficients, horizontal pass in int32, vertical pass
//   dst = (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2
// Coefficient / offset tables are built on the host (api.cu) with the same float/double steps as OpenCV.
//
// Layout: one launch per level over the whole batch, grid.z = frame. Each thread walks DOWN a strip of 4 output
// columns: the horizontal pass of a source row is kept in registers and reused when consecutive output rows share it
// (vertical scale 1.2 -> 1.2 source rows per output row instead of 2), which cuts the byte-gather count by 40 %.
#include "common.cuh"

namespace lps {

namespace {

constexpr int kRowsPerThread = 8;

__device__ __forceinline__ void hpass4(const uint8_t* __restrict__ row, const int (&sx)[4], const int (&sx1)[4],
                                       const short2 (&c)[4], int (&out)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] = __ldg(row + sx[k]) * c[k].x + __ldg(row + sx1[k]) * c[k].y;
}

__global__ void __launch_bounds__(256) k_resize_linear(LevelGeom src, LevelGeom dst) {
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y0 = (blockIdx.y * blockDim.y + threadIdx.y) * kRowsPerThread;
    if (x0 >= dst.w || y0 >= dst.h) return;
    const uint8_t* __restrict__ s = src.img + blockIdx.z * src.frame_stride;
    uint8_t* __restrict__ d = dst.img + blockIdx.z * dst.frame_stride;

    int sx[4], sx1[4];
    short2 c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, dst.w - 1);
        sx[k] = dst.xofs[x];
        sx1[k] = min(sx[k] + 1, src.w - 1);
        c[k] = dst.xcoef[x];
    }
    int h0[4], h1[4];
    int have0 = -1, have1 = -1;  // source rows currently held in h0 / h1
    const int yend = min(y0 + kRowsPerThread, dst.h);
    for (int y = y0; y < yend; ++y) {
        const int sy = dst.yofs[y];
        const short2 b = dst.ycoef[y];
        const int r0 = min(max(sy, 0), src.h - 1);
        const int r1 = min(max(sy + 1, 0), src.h - 1);
        if (r0 == have1) {
#pragma unroll
            for (int k = 0; k < 4; ++k) h0[k] = h1[k];
        }
        else if (r0 != have0) {
            hpass4(s + (size_t)r0 * src.pitch, sx, sx1, c, h0);
        }
        have0 = r0;
        if (r1 == r0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) h1[k] = h0[k];
        }
        else if (r1 != have1) {
            hpass4(s + (size_t)r1 * src.pitch, sx, sx1, c, h1);
        }
        have1 = r1;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int v = (((b.x * (h0[k] >> 4)) >> 16) + ((b.y * (h1[k] >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            packed |= (uint32_t)v << (8 * k);
        }
        uint8_t* o = d + (size_t)y * dst.pitch + x0;
        if (x0 + 4 <= dst.w) {
            *reinterpret_cast<uint32_t*>(o) = packed;  // pitch and x0 are multiples of 4
        }
        else {
            for (int k = 0; x0 + k < dst.w; ++k) o[k] = (uint8_t)(packed >> (8 * k));
        }
    }
}

// ---- column-walk bilinear kernel (the production path for scale factors below 3) -------------------------------------
// Instruction diet (budget at 60 % of HBM: ~22 issue slots per destination pixel, the first version spent ~46):
//   * a thread owns 4 destination columns (two 2-pixel windows) and walks down a band of destination rows; the row
//     schedule (which source rows a destination row needs, the row coefficients) depends on the block only, so it runs
//     on the uniform datapath and branches never diverge;
//   * the two source bytes of a destination pixel sit in a thread-private 8-byte window (two aligned words, loop
//     invariant offset); one PRMT cuts them out and one IDP.2A applies the coefficient pair (c0, c1);
//   * horizontal sums are kept as H >> 4 and reused by the next destination row when it shares the source row;
//   * vertical: two IMUL, one PRMT that gathers both products' upper halves ((b * Hs) >> 16) and one IDP.2A with weights
//     (64, 64) and addend 128 = ((t0 + t1 + 2) << 6): the output byte is byte 1 of that sum, no clamp is needed;
//   * the words of the source rows the NEXT destination row needs are fetched before the current row is computed.
__device__ __forceinline__ int reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        p = p < 0 ? -p : 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// predicated 16-bit store (keeps the store a single predicated instruction instead of a branch around it)
__device__ __forceinline__ void st_u16_if(uint8_t* p, uint16_t v, bool pred) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %2, 0;\n\t@q st.global.u16 [%0], %1;\n\t}" ::"l"(p), "h"(v), "r"((uint32_t)pred) : "memory");
}

#ifndef LPS_RW_MIN_THREADS
#define LPS_RW_MIN_THREADS (148L * 2048)
#endif
#ifndef LPS_RW_PX
#define LPS_RW_PX 4
#endif
constexpr int RW_PX = LPS_RW_PX;  // destination pixels per thread (4 or 8)
constexpr int RW_NW = RW_PX / 2;  // 2-pixel windows per thread

struct Window {  // loop-invariant description of one 2-pixel window
    const uint8_t* ptr;  // first aligned word of the window in source row 0 of this frame
    uint32_t sel0, sel1; // PRMT selectors cutting (p[sx], p[sx + 1]) out of the two words
    uint32_t c0, c1;     // coefficient pairs (c0 | c1 << 16)
};
struct RowW {  // the words of one source row for all windows of a thread
    uint32_t lo[RW_NW], hi[RW_NW];
};

struct TabE {  // block-local copy of the source-row schedule (one entry per source row of the band)
    int y;            // destination row completed by source rows (r - 1, r); -1 = none inside this band
    uint32_t b0, b1;  // row coefficients
    int dup;          // byte offset (in rows) of the apron row that mirrors y, 0 = none
};
constexpr int kTabMax = 224;  // source rows per band the shared schedule can hold (launcher keeps bands below it)

__global__ void __launch_bounds__(256) k_resize_walk(LevelGeom src, LevelGeom dst, int rows_per_block) {
    // The destination domain is the level PLUS its reflected apron: columns [-RW_PX, w + kApron), rows [-kApron, h + kApron).
    // An apron pixel is computed like the real pixel it mirrors (same tables, same arithmetic), so no separate border
    // pass and no divergent edge code is needed.
    // Lane mapping: a warp owns RW_PX * 32 consecutive destination columns; window g of lane i covers the two columns
    // 64 g + 2 i (+1) of them, so that ONE warp-wide load touches the source bytes of 64 consecutive destination columns
    // (77 contiguous bytes at scale 1.2: a single L1 wavefront instead of three or four).
    __shared__ TabE s_tab[kTabMax];
    // programmatic dependent launch: the next level's blocks may start their set-up (tables only) while this grid drains
    asm volatile("griddepcontrol.launch_dependents;");
    const int warp_x0 = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (RW_PX * 32) - 2 * kApron;
    const int lane2 = 2 * (int)(threadIdx.x & 31);
    const uint8_t* __restrict__ s = src.img + blockIdx.z * src.frame_stride;
    const int spitch = src.pitch, dpitch = dst.pitch;
    // Band of destination rows [y0, y1). The walk is over SOURCE rows: every iteration consumes one source row (its
    // words were fetched two iterations earlier), writes its H >> 4 into the register set that held the row before
    // last, and, when the schedule says that (r - 1, r) completes a destination row, emits it. The two sets swap roles
    // every iteration, so nothing is ever moved, and the control flow is the same for the whole block.
    const int y0 = blockIdx.y * rows_per_block;
    const int y1 = min(y0 + rows_per_block, dst.h);
    const int r_first = dst.yofs[y0], r_last = dst.yofs[y1 - 1] + 1;
    for (int i = threadIdx.x; i <= r_last - r_first; i += blockDim.x) {
        const int2 e = __ldg(dst.srctab + r_first + i);
        TabE t;
        const bool in = e.x >= y0 && e.x < y1;
        t.y = in ? e.x : -1;
        t.b0 = (uint32_t)e.y & 0xFFFFu;
        t.b1 = (uint32_t)e.y >> 16;
        // apron rows are copies of the real rows they mirror (BORDER_REFLECT_101: row -k = row k, row h-1+k = row h-1-k)
        t.dup = 0;
        if (e.x >= 1 && e.x <= kApron) t.dup = -2 * e.x;
        if (e.x >= dst.h - 1 - kApron && e.x <= dst.h - 2) t.dup = 2 * (dst.h - 1 - e.x);
        s_tab[i] = t;
    }

    bool live[RW_NW];
    Window wd[RW_NW];
#pragma unroll
    for (int g = 0; g < RW_NW; ++g) {
        // host-built window table (ctx.cu): what used to be two reflections, four dependent table reads and the selector
        // arithmetic per window is two 128-bit loads
        const int wi = (warp_x0 + 64 * g + lane2 + 2 * kApron) >> 1;
        const uint4 t0 = __ldg(dst.wintab + 2 * wi), t1 = __ldg(dst.wintab + 2 * wi + 1);
        wd[g].ptr = s + t0.x + (ptrdiff_t)r_first * spitch;
        wd[g].sel0 = t0.y;
        wd[g].sel1 = t0.z;
        wd[g].c0 = t0.w;
        wd[g].c1 = t1.x;
        live[g] = t1.y != 0;
    }
    uint8_t* d = dst.img + blockIdx.z * dst.frame_stride + warp_x0 + lane2;  // window g stores at d + 64 g
    asm volatile("" : "+l"(d));  // not to be rematerialised from the kernel parameters at every emitted row
    __syncthreads();
    asm volatile("griddepcontrol.wait;" ::: "memory");  // the source level is complete and visible from here on
    // (the image must be read with ordinary loads: ptxas schedules non-coherent ld.global.nc loads above the wait)

    // fetch the words of the next unread source row (rows below the image are apron rows: readable, and only ever
    // multiplied by a zero coefficient) and advance the window pointers
    auto fetch = [&](RowW& w) {
#pragma unroll
        for (int g = 0; g < RW_NW; ++g) {
            // 4-byte aligned window; ordinary (coherent, L1-cached) loads, kept below the dependency wait
            asm volatile("ld.global.u32 %0, [%2];\n\tld.global.u32 %1, [%2 + 4];" : "=r"(w.lo[g]), "=r"(w.hi[g]) : "l"(wd[g].ptr));
            wd[g].ptr += spitch;
            asm volatile("" : "+l"(wd[g].ptr));  // keep a running 64-bit pointer (no per-row re-derivation from a 32-bit offset)
        }
    };
    auto hsum = [&](const RowW& w, uint32_t (&h)[RW_PX]) {  // H >> 4 of the thread's pixels on one source row
#pragma unroll
        for (int g = 0; g < RW_NW; ++g) {
            h[2 * g] = __dp2a_lo(wd[g].c0, __byte_perm(w.lo[g], w.hi[g], wd[g].sel0), 0u) >> 4;
            h[2 * g + 1] = __dp2a_lo(wd[g].c1, __byte_perm(w.lo[g], w.hi[g], wd[g].sel1), 0u) >> 4;
        }
    };
    // source row r_first + i arrives in `buf`; Lo holds row r - 1; Up receives row r
    auto step = [&](int i, uint32_t (&Lo)[RW_PX], uint32_t (&Up)[RW_PX], RowW& buf) {
        hsum(buf, Up);
        fetch(buf);
        const TabE e = s_tab[i];
        if (e.y >= 0) {  // block-uniform
            uint32_t sum[RW_PX];
            uint16_t out[RW_NW];
#pragma unroll
            for (int k = 0; k < RW_PX; ++k)
                sum[k] = __dp2a_lo(__byte_perm(e.b0 * Lo[k], e.b1 * Up[k], 0x7632), 0x4040u, 128u);  // (t0 + t1 + 2) << 6
#pragma unroll
            for (int g = 0; g < RW_NW; ++g) out[g] = (uint16_t)__byte_perm(sum[2 * g], sum[2 * g + 1], 0x4451);  // byte 1 of each
            uint8_t* o = d + (ptrdiff_t)e.y * dpitch;  // window g stores at o + 64 g (immediate offsets)
#pragma unroll
            for (int g = 0; g < RW_NW; ++g) st_u16_if(o + 64 * g, out[g], live[g]);
            if (e.dup) {  // block-uniform, 8 rows per level
                uint8_t* o2 = o + (ptrdiff_t)e.dup * dpitch;
#pragma unroll
                for (int g = 0; g < RW_NW; ++g) st_u16_if(o2 + 64 * g, out[g], live[g]);
            }
        }
    };

    // Rows are fetched four source rows ahead of their use (a band is a short dependent chain of row loads otherwise: at
    // the small levels a block lives for a dozen rows and the loads' latency was most of its life). The reads run up to
    // four rows past the band's last source row: inside the apron, or inside the slack ctx.cu leaves behind every level.
    uint32_t A[RW_PX], B[RW_PX];
    RowW w0, w1, w2, w3;
    {
        RowW wa;
        fetch(wa);
        fetch(w0);
        fetch(w1);
        fetch(w2);
        fetch(w3);
        hsum(wa, A);
    }
    const int n = r_last - r_first;  // source rows r_first + 1 .. r_last are walked
    int i = 1;
    for (; i + 3 <= n; i += 4) {
        step(i, A, B, w0);
        step(i + 1, B, A, w1);
        step(i + 2, A, B, w2);
        step(i + 3, B, A, w3);
    }
    if (i <= n) step(i, A, B, w0);
    if (i + 1 <= n) step(i + 1, B, A, w1);
    if (i + 2 <= n) step(i + 2, A, B, w2);
}

// exact 2x decimation: OpenCV's INTER_LINEAR takes the INTER_AREA fast path, (a + b + c + d + 2) >> 2
__global__ void __launch_bounds__(256) k_resize_half(LevelGeom src, LevelGeom dst) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= dst.w || y >= dst.h) return;
    const uint8_t* __restrict__ s = src.img + blockIdx.z * src.frame_stride + (size_t)(2 * y) * src.pitch + 2 * x;
    const int v = (s[0] + s[1] + s[src.pitch] + s[src.pitch + 1] + 2) >> 2;
    dst.img[blockIdx.z * dst.frame_stride + (size_t)y * dst.pitch + x] = (uint8_t)v;
}

// pitched 2-D copy (level 0 ingest from a caller-pitched device image, and same-size "resize")
__global__ void __launch_bounds__(256) k_copy2d(const uint8_t* __restrict__ src, size_t spitch, size_t sstride,
                                               uint8_t* __restrict__ dst, int w, int h, int dpitch, size_t dstride,
                                               int vec_ok) {
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (y >= h) return;
    const uint8_t* s = src + blockIdx.z * sstride + (size_t)y * spitch;
    uint8_t* d = dst + blockIdx.z * dstride + (size_t)y * dpitch;
    const int x16 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (x16 >= w) return;
    if (vec_ok && x16 + 16 <= w) {
        *reinterpret_cast<uint4*>(d + x16) = __ldg(reinterpret_cast<const uint4*>(s + x16));
    }
    else {
        for (int k = 0; k < 16 && x16 + k < w; ++k) d[x16 + k] = s[x16 + k];
    }
}

}  // namespace

bool resize_writes_apron(const LevelGeom& src, const LevelGeom& dst) {
    (void)src;
    return dst.walk_ok != 0;
}

void launch_resize(const LevelGeom& src, const LevelGeom& dst, int batch, cudaStream_t st) {
    if (dst.w <= 0 || dst.h <= 0) return;
    if (dst.mode == 2) {
        launch_copy_level0(src.img, src.pitch, src.frame_stride, dst, batch, st);
        return;
    }
    if (dst.mode == 1) {
        dim3 blk(32, 8);
        dim3 grd((dst.w + 31) / 32, (dst.h + 7) / 8, batch);
        k_resize_half<<<grd, blk, 0, st>>>(src, dst);
        return;
    }
    if (dst.walk_ok) {
        // block width: the candidate that wastes the fewest lanes on the last column tile
        // columns -2 kApron .. w + kApron in units of a thread's RW_PX pixels (a warp owns RW_PX * 32 consecutive columns)
        const int strips = ((dst.w + 3 * kApron + RW_PX * 32 - 1) / (RW_PX * 32)) * 32;
        const int vrows = dst.h;
        int best_t = 128, best_waste = 1 << 30;
        for (int t = 64; t <= 256; t += 32) {
            const int waste = (strips + t - 1) / t * t - strips;
            if (waste < best_waste || (waste == best_waste && t == 128)) {
                best_waste = waste;
                best_t = t;
            }
        }
        // band height: long walks amortise the per-thread set-up, short ones keep a small batch spread over the SMs
        const long threads_per_row_of_bands = (long)((strips + best_t - 1) / best_t) * best_t * batch;
        int rows = 64;
        while (rows > 8 && threads_per_row_of_bands * ((vrows + rows - 1) / rows) < LPS_RW_MIN_THREADS) rows >>= 1;
        // the band's source rows must fit the shared schedule (ctx.cu only sets walk_ok when they do at 8 rows)
        while (rows > 8 && (long)rows * src.h / dst.h + 4 > kTabMax) rows >>= 1;
        dim3 grd((strips + best_t - 1) / best_t, (vrows + rows - 1) / rows, batch);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grd;
        cfg.blockDim = dim3(best_t);
        cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, k_resize_walk, src, dst, rows);
        return;
    }
    dim3 blk(32, 8);
    dim3 grd((dst.w + 127) / 128, (dst.h + 8 * kRowsPerThread - 1) / (8 * kRowsPerThread), batch);
    k_resize_linear<<<grd, blk, 0, st>>>(src, dst);
}

void launch_copy_level0(const uint8_t* src, size_t src_pitch, size_t src_frame_stride, const LevelGeom& dst, int batch,
                        cudaStream_t st) {
    dim3 blk(32, 8);
    dim3 grd((dst.w + 16 * 32 - 1) / (16 * 32), (dst.h + 7) / 8, batch);
    const int vec_ok = (reinterpret_cast<uintptr_t>(src) % 16 == 0) && (src_pitch % 16 == 0) && (src_frame_stride % 16 == 0);
    k_copy2d<<<grd, blk, 0, st>>>(src, src_pitch, src_frame_stride, dst.img, dst.w, dst.h, dst.pitch, dst.frame_stride, vec_ok);
}

}  // namespace lps
