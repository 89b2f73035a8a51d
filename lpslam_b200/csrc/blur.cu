// K5 — 7x7 Gaussian blur, sigma 2, BORDER_REFLECT_101. Replaces cv::GaussianBlur at feature/orb_extractor.cc:135-136.
// Arithmetic: separable integer kernel [18 34 48 56 48 34 18] (sum 256); the horizontal pass keeps
// full 16-bit sums, the vertical pass is Q16.16 with round-half-up: dst = (V + 32768) >> 16. Bit-exact with OpenCV 4.x.
//
// One launch blurs every level of every frame of the batch: blockIdx.x walks the concatenated tile lists of the levels,
// blockIdx.y is the frame. A block stages a (128+32) x (32+6) byte window in shared memory with 128-bit loads, runs the
// horizontal pass into a 16-bit shared buffer and the vertical pass out of it; global traffic is one read and one write
// per pixel plus the 6-row / 16-byte halo.
#include "common.cuh"

namespace lps {

namespace {

constexpr int TW = 128, TH = 32, HALO = 3;
constexpr int IN_W = TW + 32;        // bytes per staged row: [tx0 - 16, tx0 + 144), 16-byte aligned in global memory
constexpr int IN_H = TH + 2 * HALO;  // 38 rows

__device__ __forceinline__ int reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        p = p < 0 ? -p : 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

struct BlurTable {
    int num_levels;
    int tile_begin[kMaxLevels + 1];
    int tiles_x[kMaxLevels];
};

__global__ void __launch_bounds__(256) k_gauss7(const LevelGeom* __restrict__ lvs, BlurTable tab) {
    __shared__ __align__(16) uint8_t s_in[IN_H][IN_W];
    __shared__ __align__(16) uint16_t s_h[IN_H][TW];

    int level = 0;
    while (level + 1 < tab.num_levels && (int)blockIdx.x >= tab.tile_begin[level + 1]) ++level;
    const LevelGeom lv = lvs[level];
    const int tile = blockIdx.x - tab.tile_begin[level];
    const int tx0 = (tile % tab.tiles_x[level]) * TW;
    const int ty0 = (tile / tab.tiles_x[level]) * TH;
    const uint8_t* __restrict__ src = lv.img + blockIdx.y * lv.frame_stride;
    uint8_t* __restrict__ dst = lv.blur + blockIdx.y * lv.frame_stride;
    const int tid = threadIdx.x;

    // ---- stage the window (rows reflect at the top/bottom, columns at the left/right) ----
    for (int idx = tid; idx < IN_H * (IN_W / 16); idx += 256) {
        const int r = idx / (IN_W / 16), v = idx % (IN_W / 16);
        const int gy = reflect101(ty0 - HALO + r, lv.h);
        const int gx = tx0 - 16 + 16 * v;
        const uint8_t* row = src + (size_t)gy * lv.pitch;
        uint4 val;
        if (gx >= 0 && gx + 16 <= lv.w) {
            val = __ldg(reinterpret_cast<const uint4*>(row + gx));
        }
        else {
            uint32_t w4[4] = {0, 0, 0, 0};
#pragma unroll
            for (int k = 0; k < 16; ++k) w4[k >> 2] |= (uint32_t)row[reflect101(gx + k, lv.w)] << (8 * (k & 3));
            val = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
        *reinterpret_cast<uint4*>(&s_in[r][16 * v]) = val;
    }
    __syncthreads();

    // ---- horizontal pass: 4 outputs per thread-iteration from 12 staged bytes ----
    for (int idx = tid; idx < IN_H * (TW / 4); idx += 256) {
        const int r = idx / (TW / 4), cg = idx % (TW / 4);
        const uint32_t* w = reinterpret_cast<const uint32_t*>(&s_in[r][4 * cg + 12]);
        const uint32_t w0 = w[0], w1 = w[1], w2 = w[2];
        int p[12];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            p[k] = (w0 >> (8 * k)) & 255;
            p[4 + k] = (w1 >> (8 * k)) & 255;
            p[8 + k] = (w2 >> (8 * k)) & 255;
        }
        // output x (local 4cg + o) is centred on staged byte 16 + 4cg + o = p[4 + o]; taps p[1 + o .. 7 + o]
        uint32_t h[4];
#pragma unroll
        for (int o = 0; o < 4; ++o)
            h[o] = 18 * (p[1 + o] + p[7 + o]) + 34 * (p[2 + o] + p[6 + o]) + 48 * (p[3 + o] + p[5 + o]) + 56 * p[4 + o];
        *reinterpret_cast<uint2*>(&s_h[r][4 * cg]) = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
    }
    __syncthreads();

    // ---- vertical pass: each thread produces a 4-wide x 4-tall patch ----
    const int cg = tid % 32, band = tid / 32;
    const int x = tx0 + 4 * cg;
    if (x >= lv.w) return;
    uint32_t hv[10][4];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint2 q = *reinterpret_cast<const uint2*>(&s_h[band * 4 + i][4 * cg]);
        hv[i][0] = q.x & 0xFFFF;
        hv[i][1] = q.x >> 16;
        hv[i][2] = q.y & 0xFFFF;
        hv[i][3] = q.y >> 16;
    }
#pragma unroll
    for (int rr = 0; rr < 4; ++rr) {
        const int y = ty0 + band * 4 + rr;
        if (y >= lv.h) break;
        uint32_t packed = 0;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            const uint32_t v = 18u * (hv[rr][o] + hv[rr + 6][o]) + 34u * (hv[rr + 1][o] + hv[rr + 5][o])
                               + 48u * (hv[rr + 2][o] + hv[rr + 4][o]) + 56u * hv[rr + 3][o];
            packed |= ((v + 32768u) >> 16) << (8 * o);
        }
        uint8_t* o = dst + (size_t)y * lv.pitch + x;
        if (x + 4 <= lv.w) {
            *reinterpret_cast<uint32_t*>(o) = packed;
        }
        else {
            for (int k = 0; x + k < lv.w; ++k) o[k] = (uint8_t)(packed >> (8 * k));
        }
    }
}

}  // namespace

void launch_blur(const LevelGeom* lv_dev, const LevelGeom* lv_host, int num_levels, int batch, cudaStream_t st) {
    BlurTable tab;
    tab.num_levels = num_levels;
    int total = 0;
    for (int l = 0; l < num_levels; ++l) {
        tab.tile_begin[l] = total;
        const int tx = (lv_host[l].w + TW - 1) / TW, ty = (lv_host[l].h + TH - 1) / TH;
        tab.tiles_x[l] = tx > 0 ? tx : 1;
        total += (lv_host[l].w > 0 && lv_host[l].h > 0) ? tx * ty : 0;
    }
    tab.tile_begin[num_levels] = total;
    if (total == 0) return;
    k_gauss7<<<dim3(total, batch), 256, 0, st>>>(lv_dev, tab);
}

}  // namespace lps
