// K5 — 7x7 Gaussian blur, sigma 2, BORDER_REFLECT_101. Replaces cv::GaussianBlur at feature/orb_extractor.cc:135-136.
// Arithmetic: separable integer kernel k = [18 34 48 56 48 34 18] (sum 256); OpenCV's fixed-point
// path keeps the full 16-bit horizontal sums and rounds once: dst = (sum_ij k_i k_j p_ij + 32768) >> 16. Nothing is
// rounded in between, so this is one exact integer 2-D convolution and it is bit-exact with OpenCV 4.x.
//
// Design (B200): the kernel is bound by the integer pipes, not by HBM, unless the instruction count per pixel is cut to
// ~10, so everything here is about that count.
//   * every level image carries a reflected apron (k_pad_borders, below), so no thread ever mirrors an index;
//   * a thread owns a strip of 8 columns and walks DOWN it; all loads are plain 32/64-bit global loads (the three
//     neighbouring words of a row are shared with the adjacent lanes through L1), there is no shared memory and no
//     barrier, and a warp reads 256 contiguous bytes per row;
//   * horizontal pass: the 7 taps of one pixel are two IDP.4A (4-way byte dot product) on byte quads cut from the row
//     words with PRMT: H(x) = dp4a(p[x-3..x], (18,34,48,56)) + dp4a(p[x+1..x+4], (48,34,18,0));
//   * the 16-bit H values of rows 2m and 2m+1 are packed into one register (a "row pair");
//   * vertical pass: the 7 taps of one output pixel are four IDP.2A (2-way 16-bit x 8-bit dot product) over four row
//     pairs, the rounding constant 32768 rides in as the first accumulator, and the output byte is byte 2 of the sum.
//   per pixel: 2 IDP.4A + 4 IDP.2A + ~2.4 PRMT + 0.5 load/store  (was ~56 instructions per pixel with the staged version)
#include "common.cuh"

namespace lps {

namespace {

constexpr int SW = 8;           // strip width (pixels per thread-row)
constexpr int MACRO_ROWS = 10;  // rows per macro iteration: 5 row pairs = one full rotation of the pair window
constexpr int BLUR_THREADS = 128;
#ifndef LPS_BLUR_MAX_MACROS
#define LPS_BLUR_MAX_MACROS 2  // longest strip a thread walks, in macro iterations of 10 rows (measured on 64 frames of 1280x720: 1 -> 0.1126, 2 -> 0.1065, 3 -> 0.1085, 4 -> 0.1106, 6 -> 0.1126, 8 -> 0.1229, 12 -> 0.127 ms)
#endif

struct BlurTable {
    int num_levels;
    int rows_per_thread;               // multiple of MACRO_ROWS
    int item_begin[kMaxLevels + 1];    // first (strip, row block) item of each level
    int strips[kMaxLevels];            // strips per row
};

constexpr uint32_t W_A = 0x30221200u;  // bytes ( 0, 18, 34, 48)
constexpr uint32_t W_B = 0x12223038u;  // bytes (56, 48, 34, 18)
constexpr uint32_t W_C = 0x38302212u;  // bytes (18, 34, 48, 56)
constexpr uint32_t W_D = 0x00122230u;  // bytes (48, 34, 18,  0)

// the four words of a strip row: rp[-4 .. 12) (the apron makes them readable on every strip)
struct RowWords {
    uint32_t wm, w0, w1, w2;
};
__device__ __forceinline__ RowWords load_row(const uint8_t* __restrict__ rp) {
    RowWords r;
    r.wm = __ldg(reinterpret_cast<const uint32_t*>(rp - 4));
    const uint2 w01 = __ldg(reinterpret_cast<const uint2*>(rp));
    r.w0 = w01.x;
    r.w1 = w01.y;
    r.w2 = __ldg(reinterpret_cast<const uint32_t*>(rp + 8));
    return r;
}

// horizontal sums of the 8 pixels of a strip row
__device__ __forceinline__ void hrow(const RowWords& r, uint32_t (&h)[SW]) {
    uint32_t q[12];  // q[k] = bytes p[k-3 .. k]
    q[3] = r.w0;
    q[7] = r.w1;
    q[11] = r.w2;
    q[0] = __byte_perm(r.wm, r.w0, 0x4321);
    q[1] = __byte_perm(r.wm, r.w0, 0x5432);
    q[2] = __byte_perm(r.wm, r.w0, 0x6543);
    q[4] = __byte_perm(r.w0, r.w1, 0x4321);
    q[5] = __byte_perm(r.w0, r.w1, 0x5432);
    q[6] = __byte_perm(r.w0, r.w1, 0x6543);
    q[8] = __byte_perm(r.w1, r.w2, 0x4321);
    q[9] = __byte_perm(r.w1, r.w2, 0x5432);
    q[10] = __byte_perm(r.w1, r.w2, 0x6543);
#pragma unroll
    for (int k = 0; k < SW; ++k) h[k] = __dp4a(q[k + 4], W_D, __dp4a(q[k], W_C, 0u));
}

__device__ __forceinline__ uint32_t pack4(uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
    // byte 2 of each sum (sum < 2^24) -> one word
    return __byte_perm(__byte_perm(r0, r1, 0x6262), __byte_perm(r2, r3, 0x6262), 0x5410);
}

__global__ void __launch_bounds__(BLUR_THREADS) k_gauss7(const LevelGeom* __restrict__ lvs, BlurTable tab) {
    const int item = blockIdx.x * BLUR_THREADS + threadIdx.x;
    if (item >= tab.item_begin[tab.num_levels]) return;
    int level = 0;
    while (item >= tab.item_begin[level + 1]) ++level;
    const int w = lvs[level].w, h = lvs[level].h, pitch = lvs[level].pitch;
    (void)w;
    const size_t fofs = blockIdx.y * lvs[level].frame_stride;
    const int local = item - tab.item_begin[level];
    const int strips = tab.strips[level];
    const int rb = local / strips, s = local - rb * strips;
    const int y0 = rb * tab.rows_per_thread;
    const uint8_t* __restrict__ src = lvs[level].img + fofs + s * SW;
    uint8_t* __restrict__ dst = lvs[level].blur + fofs + s * SW;
    const int last_row = h + 3;  // the apron holds rows -4 .. h + 3

    // row pair m (virtual rows 2m, 2m + 1) of this strip: the words are fetched one iteration ahead of their use
    auto fetch_pair = [&](int m, RowWords& a, RowWords& b) {
        a = load_row(src + (ptrdiff_t)min(2 * m, last_row) * pitch);
        b = load_row(src + (ptrdiff_t)min(2 * m + 1, last_row) * pitch);
    };
    auto make_pair = [&](const RowWords& a, const RowWords& b, uint32_t (&p)[SW]) {
        uint32_t he[SW], ho[SW];
        hrow(a, he);
        hrow(b, ho);
#pragma unroll
        for (int k = 0; k < SW; ++k) p[k] = __byte_perm(he[k], ho[k], 0x5410);
    };

    uint32_t P[5][SW];  // window of row pairs j-2 .. j+2 (rotating: pair m lives in slot m mod 5, relative to the start)
    const int j0 = y0 >> 1;
    RowWords na, nb;  // words of the pair the next iteration consumes
    {
        RowWords r[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) fetch_pair(j0 - 2 + i, r[2 * i], r[2 * i + 1]);
        fetch_pair(j0 + 2, na, nb);
#pragma unroll
        for (int i = 0; i < 4; ++i) make_pair(r[2 * i], r[2 * i + 1], P[i]);
    }

    const int macros = tab.rows_per_thread / MACRO_ROWS;
    for (int mi = 0; mi < macros; ++mi) {
        const int jb = j0 + mi * (MACRO_ROWS / 2);
        if (2 * jb >= h) break;
#pragma unroll
        for (int u = 0; u < MACRO_ROWS / 2; ++u) {
            const int j = jb + u;
            uint32_t(&A)[SW] = P[u % 5];
            uint32_t(&B)[SW] = P[(u + 1) % 5];
            uint32_t(&C)[SW] = P[(u + 2) % 5];
            uint32_t(&D)[SW] = P[(u + 3) % 5];
            uint32_t(&E)[SW] = P[(u + 4) % 5];
            const RowWords ca = na, cb = nb;
            fetch_pair(j + 3, na, nb);
            make_pair(ca, cb, E);
            uint32_t re[SW], ro[SW];
#pragma unroll
            for (int k = 0; k < SW; ++k) {
                uint32_t a = __dp2a_lo(A[k], W_A, 32768u);  // rows 2j-4 (x0), 2j-3 (x18)
                a = __dp2a_hi(B[k], W_A, a);                // 2j-2 (x34), 2j-1 (x48)
                a = __dp2a_lo(C[k], W_B, a);                // 2j (x56), 2j+1 (x48)
                re[k] = __dp2a_hi(D[k], W_B, a);            // 2j+2 (x34), 2j+3 (x18)
                uint32_t b = __dp2a_lo(B[k], W_C, 32768u);  // 2j-2 (x18), 2j-1 (x34)
                b = __dp2a_hi(C[k], W_C, b);                // 2j (x48), 2j+1 (x56)
                b = __dp2a_lo(D[k], W_D, b);                // 2j+2 (x48), 2j+3 (x34)
                ro[k] = __dp2a_hi(E[k], W_D, b);            // 2j+4 (x18), 2j+5 (x0)
            }
            // rows past the image are parked in apron row h of the blurred level: no branch in the loop body
            *reinterpret_cast<uint2*>(dst + (size_t)min(2 * j, h) * pitch) = make_uint2(pack4(re[0], re[1], re[2], re[3]), pack4(re[4], re[5], re[6], re[7]));
            *reinterpret_cast<uint2*>(dst + (size_t)min(2 * j + 1, h) * pitch) = make_uint2(pack4(ro[0], ro[1], ro[2], ro[3]), pack4(ro[4], ro[5], ro[6], ro[7]));
        }
    }
}

__device__ __forceinline__ int reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        p = p < 0 ? -p : 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// Writes the BORDER_REFLECT_101 apron (kApron rows above / below, kApron columns left / right) of every level of every
// frame, so that the blur never mirrors an index. grid = (row block | apron row, level, frame). Every value is read from a
// real pixel, so blocks are independent: in the first `row_blocks` blocks a thread owns one row (apron rows included) and
// writes its 2 x kApron apron bytes; each of the last 2 x kApron blocks copies one mirrored apron row.
constexpr int PAD_THREADS = 128;
__global__ void __launch_bounds__(PAD_THREADS) k_pad_borders(const LevelGeom* __restrict__ lvs, int row_blocks, uint32_t level_mask) {
    if (!((level_mask >> blockIdx.y) & 1u)) return;
    const int w = lvs[blockIdx.y].w, h = lvs[blockIdx.y].h, pitch = lvs[blockIdx.y].pitch;
    if (w <= 0 || h <= 0) return;
    uint8_t* __restrict__ img = lvs[blockIdx.y].img + blockIdx.z * lvs[blockIdx.y].frame_stride;
    if ((int)blockIdx.x < row_blocks) {
        const int y = (int)blockIdx.x * PAD_THREADS + threadIdx.x - kApron;
        if (y >= h + kApron) return;
        const uint8_t* __restrict__ s = img + (ptrdiff_t)reflect101(y, h) * pitch;
        uint8_t* __restrict__ d = img + (ptrdiff_t)y * pitch;
        uint8_t l[kApron], r[kApron];
        if (w > kApron + 1) {  // the usual case: no repeated mirroring
#pragma unroll
            for (int c = 0; c < kApron; ++c) {
                l[c] = s[1 + c];
                r[c] = s[w - 2 - c];
            }
        }
        else {
#pragma unroll
            for (int c = 0; c < kApron; ++c) {
                l[c] = s[reflect101(-1 - c, w)];
                r[c] = s[reflect101(w + c, w)];
            }
        }
#pragma unroll
        for (int c = 0; c < kApron; ++c) {
            d[-1 - c] = l[c];
            d[w + c] = r[c];
        }
        return;
    }
    const int a = (int)blockIdx.x - row_blocks;  // 0 .. 2 * kApron - 1
    const int y = a < kApron ? a - kApron : h + (a - kApron);
    const uint8_t* __restrict__ s = img + (ptrdiff_t)reflect101(y, h) * pitch;
    uint8_t* __restrict__ d = img + (ptrdiff_t)y * pitch;
    for (int x4 = threadIdx.x; x4 < w / 4; x4 += PAD_THREADS)  // rows start 16-byte aligned
        reinterpret_cast<uint32_t*>(d)[x4] = reinterpret_cast<const uint32_t*>(s)[x4];
    if ((int)threadIdx.x < (w & 3)) d[(w & ~3) + threadIdx.x] = s[(w & ~3) + threadIdx.x];
}

}  // namespace

void launch_pad_borders(const LevelGeom* lv_dev, const LevelGeom* lv_host, int num_levels, uint32_t level_mask, int batch,
                        cudaStream_t st) {
    int hmax = 0, top = 0;
    for (int l = 0; l < num_levels; ++l)
        if ((level_mask >> l) & 1u) {
            hmax = lv_host[l].h > hmax ? lv_host[l].h : hmax;
            top = l + 1;
        }
    num_levels = top;
    if (hmax <= 0) return;
    const int row_blocks = (hmax + 2 * kApron + PAD_THREADS - 1) / PAD_THREADS;
    k_pad_borders<<<dim3(row_blocks + 2 * kApron, num_levels, batch), PAD_THREADS, 0, st>>>(lv_dev, row_blocks, level_mask);
}

void launch_blur(const LevelGeom* lv_dev, const LevelGeom* lv_host, int num_levels, int batch, cudaStream_t st) {
    BlurTable tab;
    tab.num_levels = num_levels;
    // rows per thread: long strips amortise the 8-row prologue; short ones keep a small batch spread over the SMs
    long px = 0;
    for (int l = 0; l < num_levels; ++l) px += (long)lv_host[l].w * lv_host[l].h;
    const long threads_at_10 = px * batch / (SW * MACRO_ROWS);  // threads if every thread took one macro
    int macros = (int)(threads_at_10 / (148L * 1024));
    macros = macros < 1 ? 1 : (macros > LPS_BLUR_MAX_MACROS ? LPS_BLUR_MAX_MACROS : macros);
    tab.rows_per_thread = macros * MACRO_ROWS;
    int total = 0;
    for (int l = 0; l < num_levels; ++l) {
        tab.item_begin[l] = total;
        const int strips = (lv_host[l].w + SW - 1) / SW;
        const int blocks = (lv_host[l].h + tab.rows_per_thread - 1) / tab.rows_per_thread;
        tab.strips[l] = strips > 0 ? strips : 1;
        total += (lv_host[l].w > 0 && lv_host[l].h > 0) ? strips * blocks : 0;
    }
    for (int l = num_levels; l <= kMaxLevels; ++l) tab.item_begin[l] = total;
    if (total == 0) return;
    k_gauss7<<<dim3((total + BLUR_THREADS - 1) / BLUR_THREADS, batch), BLUR_THREADS, 0, st>>>(lv_dev, tab);
}

}  // namespace lps
