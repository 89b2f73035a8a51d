// K5 (production path) — 7x7 Gaussian blur, sigma 2, BORDER_REFLECT_101, with the source rows staged through shared memory
// by TMA. Replaces cv::GaussianBlur at feature/orb_extractor.cc:135-136; arithmetic as in blur.cu:
// one exact integer convolution, dst = (sum_ij k_i k_j p_ij + 32768) >> 16, k = [18 34 48 56 48 34 18].
//
// Why a second version: k_gauss7 (blur.cu) reads its rows with ordinary global loads. ncu (profiles/r1_ncu_summary.md):
// 1.18x DRAM read amplification (26 rows fetched per 20-row strip), 26 % of the stall samples on the first use of a freshly
// loaded row, 45 % occupancy. Here
//   * the unit of work is a HALF-WARP: 16 lanes own a band of 16 x 8 = 128 columns and stream DOWN a long row segment
//     (~240 rows, so the 10-row halo of a segment is ~4 % instead of 30 %);
//   * the rows arrive by cp.async.bulk.tensor (one 160-byte x 10-row box per half-warp and stage: 128 columns + 2 x 16 bytes
//     of halo, a box has to start on a 16-byte boundary; the level's reflected apron is part of the tensor, rows above / below it are zero-filled by the TMA unit
//     and only ever feed outputs that are not stored), double buffered, completion on one mbarrier per warp and stage;
//     lane 0 of every warp is its own producer, there is no block-level synchronisation at all;
//   * a lane reads its 16 bytes of a row as LDS.32 + LDS.64 + LDS.32 with immediate offsets, then runs the same
//     IDP.4A / IDP.2A arithmetic as k_gauss7 (2 + 4 dot-product instructions per pixel) and stores 8 bytes per row;
//   * the two halves of a warp are independent (band, segment) pairs of one level, so ragged level widths cost at most
//     15 idle lanes per level row instead of 31.
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>
#include <cstdlib>
#include <map>
#include <vector>

namespace lps {

namespace {

constexpr int BT_ROWS = 10;                                // rows per stage = one rotation of the 5-slot row-pair window
constexpr int BT_BOXW = 160;                               // bytes per box row: 16 + 16 strips x 8 + 16 (a box must START on a 16-byte boundary of the row: measured, tools/scratch/tma_probe.cu)
constexpr int BT_STAGES = 2;
#ifndef LPS_BT_WARPS
#define LPS_BT_WARPS 1
#endif
constexpr int BT_WARPS = LPS_BT_WARPS;
constexpr int BT_HALF_BYTES = BT_ROWS * BT_BOXW;           // 1600 bytes land per box
constexpr int BT_HALF_STRIDE = 1664;                       // a TMA destination must be 128-byte aligned
constexpr int BT_STAGE_BYTES = 2 * BT_HALF_STRIDE;         // 3328
constexpr int BT_WARP_BYTES = BT_STAGES * BT_STAGE_BYTES;  // 6656
constexpr int SW = 8;
#ifndef LPS_BT_MINB
#define LPS_BT_MINB 6  // resident blocks per SM the register allocation must allow (6 x 4 warps at 80 registers; measured: 72 registers spill and run 10-20 % slower)
#endif

struct BlurMaps {
    CUtensorMap m[kMaxLevels];
};

struct BlurUnit {  // one warp: two independent (band, segment) halves of one level
    int level, macros;
    int xb[2], y0[2], strips[2];  // band origin (pixels), first output row, live strips (0 = half unused)
};

constexpr uint32_t W_A = 0x30221200u;  // bytes ( 0, 18, 34, 48)
constexpr uint32_t W_B = 0x12223038u;  // bytes (56, 48, 34, 18)
constexpr uint32_t W_C = 0x38302212u;  // bytes (18, 34, 48, 56)
constexpr uint32_t W_D = 0x00122230u;  // bytes (48, 34, 18,  0)

// horizontal sums of the 8 pixels of a strip row from its 16 staged bytes p[-4 .. 12)
__device__ __forceinline__ void hrow(const uint32_t wm, const uint2 mid, const uint32_t w2, uint32_t (&h)[SW]) {
    const uint32_t w0 = mid.x, w1 = mid.y;
    uint32_t q[12];  // q[k] = bytes p[k-3 .. k]
    q[3] = w0;
    q[7] = w1;
    q[11] = w2;
    q[0] = __byte_perm(wm, w0, 0x4321);
    q[1] = __byte_perm(wm, w0, 0x5432);
    q[2] = __byte_perm(wm, w0, 0x6543);
    q[4] = __byte_perm(w0, w1, 0x4321);
    q[5] = __byte_perm(w0, w1, 0x5432);
    q[6] = __byte_perm(w0, w1, 0x6543);
    q[8] = __byte_perm(w1, w2, 0x4321);
    q[9] = __byte_perm(w1, w2, 0x5432);
    q[10] = __byte_perm(w1, w2, 0x6543);
#pragma unroll
    for (int k = 0; k < SW; ++k) h[k] = __dp4a(q[k + 4], W_D, __dp4a(q[k], W_C, 0u));
}

__device__ __forceinline__ uint32_t pack4(uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
    return __byte_perm(__byte_perm(r0, r1, 0x6262), __byte_perm(r2, r3, 0x6262), 0x5410);  // byte 2 of each sum
}

// rows 2i, 2i + 1 of a stage -> one packed row pair
__device__ __forceinline__ void make_pair(const uint8_t* st, int i, uint32_t (&p)[SW]) {
    // st points at the lane's pixel 0 of stage row 0 (8-byte aligned); p[-4 .. 0) and p[8 .. 12) are 4-byte aligned only
    const uint8_t* r0 = st + (2 * i) * BT_BOXW;
    const uint32_t a_m = *reinterpret_cast<const uint32_t*>(r0 - 4), a_2 = *reinterpret_cast<const uint32_t*>(r0 + 8);
    const uint2 a_01 = *reinterpret_cast<const uint2*>(r0);
    const uint32_t b_m = *reinterpret_cast<const uint32_t*>(r0 + BT_BOXW - 4), b_2 = *reinterpret_cast<const uint32_t*>(r0 + BT_BOXW + 8);
    const uint2 b_01 = *reinterpret_cast<const uint2*>(r0 + BT_BOXW);
    uint32_t he[SW], ho[SW];
    hrow(a_m, a_01, a_2, he);
    hrow(b_m, b_01, b_2, ho);
#pragma unroll
    for (int k = 0; k < SW; ++k) p[k] = __byte_perm(he[k], ho[k], 0x5410);
}

__global__ void __launch_bounds__(BT_WARPS * 32, LPS_BT_MINB * 4 / BT_WARPS) k_gauss7_tma(const __grid_constant__ BlurMaps maps, const LevelGeom* __restrict__ lvs,
                                                                             const BlurUnit* __restrict__ units, int n_units, int batch) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t s_full[BT_WARPS][BT_STAGES];
    // TMA destinations must be 128-byte aligned: the dynamic segment starts behind the static one, so align by hand
    uint8_t* const smem = smem_raw + ((128u - (smem_u32(smem_raw) & 127u)) & 127u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int half = lane >> 4, hl = lane & 15;
    uint8_t* const warp_smem = smem + warp * BT_WARP_BYTES;
    const uint32_t sbase = smem_u32(warp_smem);
    const uint32_t bar0 = smem_u32(&s_full[warp][0]);
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < BT_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();
    const uint8_t* const lane_smem = warp_smem + half * BT_HALF_STRIDE + 16 + hl * SW;
    const uint32_t total_work = (uint32_t)n_units * (uint32_t)batch;
    uint32_t ph0 = 0, ph1 = 0;  // parity the next completion of stage 0 / 1 will have
    // Warps are independent workers (no block-level barrier anywhere): warp g of the grid owns item g = unit * batch + frame, longest
    // units first, so the tail of the grid is made of short items. A persistent grid drawing items from a device counter measured
    // slower (0.112 against 0.096 ms, profiles/r2_blur_tma.md), so the hardware block scheduler does the balancing.
    const uint32_t g = blockIdx.x * BT_WARPS + warp;
    if (g < total_work) {
        const BlurUnit u = units[g / (uint32_t)batch];
        const int frame = (int)(g % (uint32_t)batch);
        const int h = lvs[u.level].h, pitch = lvs[u.level].pitch;
        const int total = u.macros + 1;  // stage 0 is the 8-row run-in (rows y0 - 4 .. y0 + 3 sit in rows 2 .. 9 of the stage)
        const CUtensorMap* map = &maps.m[u.level];
        const int nh = u.strips[1] > 0 ? 2 : 1;
        auto issue = [&](int it) {  // lane 0: rows y0 - 6 + 10 it .. + 9 of both halves (tensor row = image row + kApron)
            const int s = it & (BT_STAGES - 1);
            const uint32_t bar = bar0 + 8 * s, dst = sbase + s * BT_STAGE_BYTES;
            mbar_expect_tx(bar, nh * BT_HALF_BYTES);
            // the box starts at image column xb - 16 = tensor column xb (kPadLeft = 16)
            tma_load_3d(dst, map, u.xb[0], u.y0[0] - 6 + kApron + BT_ROWS * it, frame, bar);
            if (nh == 2) tma_load_3d(dst + BT_HALF_STRIDE, map, u.xb[1], u.y0[1] - 6 + kApron + BT_ROWS * it, frame, bar);
        };
        if (lane == 0) {
            issue(0);
            if (total > 1) issue(1);
        }
        const int my_xb = half ? u.xb[1] : u.xb[0], my_y0 = half ? u.y0[1] : u.y0[0], my_strips = half ? u.strips[1] : u.strips[0];
        const bool active = hl < my_strips;
        // running output pointer (row 2j); rows past the image are not stored
        uint8_t* o = lvs[u.level].blur + frame * lvs[u.level].frame_stride + my_xb + hl * SW + (size_t)my_y0 * pitch;
        int rows_left = active ? h - my_y0 : 0;  // output rows this lane may still store

        uint32_t P[5][SW];  // row pairs j-2 .. j+2, rotating (pair m of the segment lives in slot m mod 5)
        mbar_wait(bar0, ph0);
        ph0 ^= 1u;
#pragma unroll
        for (int i = 0; i < 4; ++i) make_pair(lane_smem, i + 1, P[i]);
        __syncwarp();
        if (lane == 0 && 2 < total) issue(2);

        for (int it = 1; it < total; ++it) {
            const int s = it & (BT_STAGES - 1);
            if (s) {
                mbar_wait(bar0 + 8, ph1);
                ph1 ^= 1u;
            }
            else {
                mbar_wait(bar0, ph0);
                ph0 ^= 1u;
            }
            const uint8_t* const st = lane_smem + s * BT_STAGE_BYTES;
#pragma unroll
            for (int v = 0; v < BT_ROWS / 2; ++v) {
                uint32_t(&A)[SW] = P[v % 5];
                uint32_t(&B)[SW] = P[(v + 1) % 5];
                uint32_t(&C)[SW] = P[(v + 2) % 5];
                uint32_t(&D)[SW] = P[(v + 3) % 5];
                uint32_t(&E)[SW] = P[(v + 4) % 5];
                make_pair(st, v, E);
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
                if (rows_left > 0) *reinterpret_cast<uint2*>(o) = make_uint2(pack4(re[0], re[1], re[2], re[3]), pack4(re[4], re[5], re[6], re[7]));
                if (rows_left > 1) *reinterpret_cast<uint2*>(o + pitch) = make_uint2(pack4(ro[0], ro[1], ro[2], ro[3]), pack4(ro[4], ro[5], ro[6], ro[7]));
                o += 2 * pitch;
                rows_left -= 2;
            }
            __syncwarp();  // every lane is done with stage s
            if (lane == 0 && it + BT_STAGES < total) issue(it + BT_STAGES);
        }
    }
}

}  // namespace

struct BlurPlan {
    BlurMaps maps;
    int num_levels = 0;
    std::vector<LevelGeom> lv;
    struct Table {
        BlurUnit* dev = nullptr;
        int n = 0;
    };
    std::map<int, Table> tables;  // by batch size
};

BlurPlan* blur_plan_create(const LevelGeom* lv_host, int num_levels, int max_batch) {
    BlurPlan* p = new BlurPlan;
    p->num_levels = num_levels;
    p->lv.assign(lv_host, lv_host + num_levels);
    for (int l = 0; l < num_levels; ++l) {
        const LevelGeom& g = lv_host[l];
        const uint64_t dims[3] = {(uint64_t)g.pitch, (uint64_t)(g.h + 2 * kApron), (uint64_t)max_batch};
        const uint64_t strides[2] = {(uint64_t)g.pitch, (uint64_t)g.frame_stride};
        const uint32_t box[3] = {(uint32_t)BT_BOXW, (uint32_t)BT_ROWS, 1u};
        if (!tma_encode_u8(&p->maps.m[l], g.img - (size_t)kApron * g.pitch - kPadLeft, 3, dims, strides, box)) {
            delete p;
            return nullptr;
        }
    }
    return p;
}

void blur_plan_destroy(BlurPlan* p) {
    if (!p) return;
    for (auto& t : p->tables) cudaFree(t.second.dev);
    delete p;
}

static int build_units(const BlurPlan* p, int batch, std::vector<BlurUnit>& out) {
    static const int seg_rows_env = [] {
        const char* e = getenv("LPS_BLUR_SEG_ROWS");
        return e ? atoi(e) : 50;
    }();
    int target = seg_rows_env > 0 ? seg_rows_env : 50;
    for (;;) {
        out.clear();
        for (int l = 0; l < p->num_levels; ++l) {
            const LevelGeom& g = p->lv[l];
            const int strips = (g.w + SW - 1) / SW, halves = (strips + 15) / 16;
            const int macros_all = (g.h + BT_ROWS - 1) / BT_ROWS;
            int segs = std::max(1, (g.h + target / 2) / target);
            const int macros = (macros_all + segs - 1) / segs;
            segs = (macros_all + macros - 1) / macros;
            BlurUnit u{};
            int k = 0;
            for (int sg = 0; sg < segs; ++sg)
                for (int hb = 0; hb < halves; ++hb) {
                    u.level = l;
                    u.macros = macros;
                    u.xb[k] = 128 * hb;
                    u.y0[k] = sg * macros * BT_ROWS;
                    u.strips[k] = std::min(16, strips - 16 * hb);
                    if (++k == 2) {
                        out.push_back(u);
                        u = BlurUnit{};
                        k = 0;
                    }
                }
            if (k == 1) out.push_back(u);
        }
        if ((long)out.size() * batch >= 148L * 16 || target <= 40) break;
        target /= 2;  // small batches: shorter segments, more warps
    }
    // longest first: what runs in the tail of the grid is short
    std::stable_sort(out.begin(), out.end(), [](const BlurUnit& a, const BlurUnit& b) { return a.macros > b.macros; });
    return (int)out.size();
}

int launch_blur_tma(BlurPlan* p, const LevelGeom* lv_dev, int batch, cudaStream_t st) {
    auto it = p->tables.find(batch);
    if (it == p->tables.end()) {
        std::vector<BlurUnit> units;
        BlurPlan::Table t;
        t.n = build_units(p, batch, units);
        if (t.n > 0) {
            LPS_CUDA(cudaMalloc(&t.dev, sizeof(BlurUnit) * units.size()));
            LPS_CUDA(cudaMemcpy(t.dev, units.data(), sizeof(BlurUnit) * units.size(), cudaMemcpyHostToDevice));
        }
        it = p->tables.emplace(batch, t).first;
    }
    const BlurPlan::Table& t = it->second;
    if (t.n == 0) return LPSLAM_OK;
    const long work = (long)t.n * batch;  // one warp per (unit, frame)
    const int blocks = (int)((work + BT_WARPS - 1) / BT_WARPS);
    k_gauss7_tma<<<blocks, BT_WARPS * 32, BT_WARPS * BT_WARP_BYTES + 128, st>>>(p->maps, lv_dev, t.dev, t.n, batch);
    return LPSLAM_OK;
}

bool tma_encode_u8(CUtensorMap* out, void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                 const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = [] {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) f = nullptr;
        return reinterpret_cast<EncodeFn>(f);
    }();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from this driver");
        return false;
    }
    cuuint64_t gd[5], gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gd[i] = dims[i];
        bx[i] = box[i];
        es[i] = 1;
        if (i + 1 < rank) gs[i] = strides_bytes[i];
    }
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, (cuuint32_t)rank, base, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
        return false;
    }
    return true;
}

}  // namespace lps
