// K4 + K6 — intensity-centroid orientation and rotated 256-bit BRIEF, one warp per keypoint, plus the output assembly
// of feature/orb_extractor.cc:119-146,382-392,601-609 (border shift, octave, size, scale correction, level concatenation).
//   ic_angle            feature/orb_extractor.cc:611-634  (un-blurred level, int32 moments, cv::fastAtan2)
//   compute_orb_descriptor  feature/orb_extractor.cc:644-705 (blurred level, polynomial cos/sin of util/trigonometric.h)
// Float steps use explicit round-to-nearest intrinsics so that no FMA contraction changes a rounding (the reference's
// default build has none): SURVEY Appendix A.4 / A.5.
#include "common.cuh"

namespace lps {

namespace {

// rBRIEF pattern in global memory: every lane reads ITS 8 pairs once (lane-divergent constant-bank reads would serialise)
__device__ const signed char g_pattern[1024] = {
#include "rbrief_pattern.inc"
};
// u_max_ of feature/orb_extractor.cc:228-241 for a half patch of 15
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.1415926535897932384626433832795);
    const float p1 = __fmul_rn(0.9997878412794807f, scale);
    const float p3 = __fmul_rn(-0.3258083974640975f, scale);
    const float p5 = __fmul_rn(0.1555786518463281f, scale);
    const float p7 = __fmul_rn(-0.04432655554792128f, scale);
    const float eps = (float)2.2204460492503131e-16;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    }
    else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        const float c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// util/trigonometric.h:11-46
__device__ __forceinline__ float poly_cos(float v) {
    const float v2 = __fmul_rn(v, v);
    return __fadd_rn(0.99940307f, __fmul_rn(v2, __fadd_rn(-0.49558072f, __fmul_rn(0.03679168f, v2))));
}
__device__ __forceinline__ float ref_cos(float v) {
    const float PI = 3.14159265358979f;
    const float PI_2 = PI / 2.0f;
    const float TWO_PI = 2.0f * PI;
    const float INV_TWO_PI = 1.0f / TWO_PI;
    const float THREE_PI_2 = 3.0f * PI_2;
    const int fl = __float2int_rd(__fmul_rn(v, INV_TWO_PI));
    v = __fsub_rn(v, __fmul_rn((float)fl, TWO_PI));
    v = (0.0f < v) ? v : -v;
    if (v < PI_2) return poly_cos(v);
    if (v < PI) return -poly_cos(__fsub_rn(PI, v));
    if (v < THREE_PI_2) return -poly_cos(__fsub_rn(v, PI));
    return poly_cos(__fsub_rn(TWO_PI, v));
}

// exclusive prefix over (frame, level) keypoint counts -> packed output offsets
__global__ void __launch_bounds__(1024) k_offsets(DescribeArgs a) {
    __shared__ unsigned s_warp[33];
    const int n = a.batch * a.num_levels;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned carry = 0;
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const unsigned v = i < n ? a.sel_count[i] : 0u;
        unsigned inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const unsigned w = s_warp[lane];
            unsigned wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_warp[32] = wi;
        }
        __syncthreads();
        const unsigned ex = carry + s_warp[warp] + inc - v;
        if (i < n) {
            a.level_off[i] = ex;
            if (i % a.num_levels == 0) a.frame_off[i / a.num_levels] = ex;
        }
        carry += s_warp[32];
        __syncthreads();
    }
    if (tid == 0) {
        a.level_off[n] = carry;
        a.frame_off[a.batch] = carry;
        if (carry > a.out_cap) *a.overflow = carry;
    }
}

constexpr int kDescribeBlocksPerSM = 2;
constexpr unsigned kRun = 8;  // consecutive packed rows per warp before it strides on
// The blurred patch a descriptor taps (|row|, |col| <= 18: the pattern's points lie within 13 of the centre, rotated) is
// staged in shared memory by cp.async while the orientation is computed: rows of kTapPitch bytes starting at the 16-byte
// boundary at or below column -18, four lanes per row and eight rows per copy. The 512 taps of a keypoint then are
// shared-memory byte reads instead of 16 fully divergent global loads per lane.
constexpr int kTapReach = 18, kTapRows = 2 * kTapReach + 1, kTapPitch = 80, kTapBuf = 40 * kTapPitch;
constexpr int kDescribeSmem = 8 * 2 * kTapBuf;  // 8 warps, double buffered

// Software pipeline over a warp's keypoints, three deep: the position of keypoint k + 2 is being read while the patch
// rows of k + 1 are in flight and k is computed. Warps take runs of kRun consecutive packed rows, strided by the grid, so
// that the warps resident at one time cover a few frames (their level images stay in L2).
__global__ void __launch_bounds__(256, kDescribeBlocksPerSM) k_describe(DescribeArgs a) {
    const int lane = threadIdx.x & 31;
    const unsigned warps_total = gridDim.x * (blockDim.x >> 5);
    const int n_seg = a.batch * a.num_levels;
    const unsigned total = min(a.level_off[n_seg], a.out_cap);
    // Orientation patch rows are read as 16-byte pieces, three lanes per row and ten rows per load: a load then touches ten
    // lines instead of one line per lane (the L1 wavefront rate is what bounds this kernel). Rows start 16-byte aligned,
    // so a row's 31 pixels sit at bytes sh .. sh + 30 of a 48-byte span, sh = (x - 15) & 15 the same for every row.
    // s_mask[|v|][sh][piece] keeps the bytes inside the disc (|u| <= u_max[|v|], feature/orb_extractor.cc:228-241); with the
    // byte position p as IDP.4A weight, m10 = sum p I - (sh + 15) sum I and m01 = v sum I.
    __shared__ uint4 s_mask[16 * 16 * 3];
    extern __shared__ __align__(16) uint8_t s_tap[];
    uint8_t* my_tap = s_tap + (threadIdx.x >> 5) * 2 * kTapBuf;
    for (int e = threadIdx.x; e < 16 * 16 * 3; e += blockDim.x) {
        const int piece = e % 3, sh = (e / 3) % 16, um = c_umax[e / 48];
        uint32_t m[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            m[w] = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int u = 16 * piece + 4 * w + b - sh - kHalfPatch;
                if ((u < 0 ? -u : u) <= um) m[w] |= 0xFFu << (8 * b);
            }
        }
        s_mask[e] = make_uint4(m[0], m[1], m[2], m[3]);
    }
    __syncthreads();
    const unsigned k_first = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * kRun;
    if (k_first >= total) return;
    float pat[32];  // this lane's 8 point pairs: x0, y0, x1, y1 each
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const char4 q = reinterpret_cast<const char4*>(g_pattern)[lane * 8 + i];
        pat[4 * i] = (float)q.x;
        pat[4 * i + 1] = (float)q.y;
        pat[4 * i + 2] = (float)q.z;
        pat[4 * i + 3] = (float)q.w;
    }
    const int piece = lane % 3, row_in_load = lane / 3;  // lanes 30, 31 idle
    const uint32_t pos0 = 0x03020100u + 0x01010101u * (uint32_t)(16 * piece);

    struct Loc {   // stage 1: where the keypoint is
        unsigned k;
        int frame, level, score;
        uint32_t xy;
        bool ok;
    };
    struct Rows {  // stage 2: its patch rows on the way
        unsigned k;
        int frame, level, score, px, py;
        unsigned sh, tsh;
        uint4 d[4];
        bool ok;
    };
    int lo = 0;
    unsigned seg_begin = 0, seg_end = 0;
    auto locate = [&](unsigned k) {
        Loc L;
        L.k = k;
        L.ok = k < total;
        L.frame = L.level = L.score = 0;
        L.xy = 0;
        if (L.ok) {
            if (k >= seg_end) {  // (frame, level) segment holding packed row k; rows are visited in ascending order
                int hi = n_seg;
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (a.level_off[mid] <= k) lo = mid;
                    else hi = mid;
                }
                seg_begin = a.level_off[lo];
                seg_end = a.level_off[lo + 1];
            }
            // lo / num_levels by a multiply (lo < 2^32 / num_levels): a runtime division here was 3.7 % of the kernel
            L.frame = a.num_levels > 1 ? (int)__umulhi((unsigned)lo, 0xFFFFFFFFu / (unsigned)a.num_levels + 1u) : lo;
            L.level = lo - L.frame * a.num_levels;
            const size_t at = (size_t)L.frame * a.sel_stride + a.lv[L.level].sel_begin + (k - seg_begin);
            L.xy = a.sel_xys[at];
            L.score = a.sel_sc[at];
        }
        return L;
    };
    int tap_turn = 0;  // buffer the next staged patch goes to
    auto fetch = [&](const Loc& L) {
        Rows R;
        R.k = L.k;
        R.ok = L.ok;
        R.frame = L.frame;
        R.level = L.level;
        R.score = L.score;
        R.px = (int)(L.xy & 0xFFFFu) + kBorder;
        R.py = (int)(L.xy >> 16) + kBorder;
        R.sh = R.tsh = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) R.d[i] = make_uint4(0, 0, 0, 0);
        if (L.ok) {
            const LevelGeom& lv = a.lv[L.level];
            const uint8_t* row0 = lv.img + L.frame * lv.frame_stride + (size_t)R.py * lv.pitch + (R.px - kHalfPatch);
            R.sh = (unsigned)(reinterpret_cast<uintptr_t>(row0) & 15);
            const uint8_t* mine = row0 - R.sh + 16 * piece;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = 10 * i + row_in_load;
                R.d[i] = (lane < 30 && r < kPatch) ? __ldg(reinterpret_cast<const uint4*>(mine + (ptrdiff_t)(r - kHalfPatch) * lv.pitch))
                                                   : make_uint4(0, 0, 0, 0);
            }
            // blurred patch -> this warp's buffer `tap_turn`
            const uint8_t* t0 = lv.blur + L.frame * lv.frame_stride + (size_t)R.py * lv.pitch + (R.px - kTapReach);
            R.tsh = (unsigned)(reinterpret_cast<uintptr_t>(t0) & 15);
            const uint8_t* src = t0 - R.tsh + 16 * (lane & 3);
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(my_tap + tap_turn * kTapBuf + 16 * (lane & 3));
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const int r = 8 * i + (lane >> 2);
                if (r < kTapRows)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + r * kTapPitch),
                                 "l"(src + (ptrdiff_t)(r - kTapReach) * lv.pitch));
            }
        }
        asm volatile("cp.async.commit_group;");
        tap_turn ^= 1;
        return R;
    };
    auto next_k = [&](unsigned k) { return ((k + 1) % kRun) ? k + 1 : k + 1 + (warps_total - 1) * kRun; };

    Loc A = locate(k_first);
    Rows B = fetch(A);
    A = locate(next_k(k_first));
    while (B.ok) {
        const Loc A2 = locate(next_k(A.k));
        const LevelGeom& lv = a.lv[B.level];
        // ---- orientation: masked bytes, IDP.4A with the byte positions and with ones ----
        unsigned sum_p = 0;
        int m01 = 0;
        unsigned sum_1 = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = 10 * i + row_in_load;
            const int vv = r - kHalfPatch;
            const int av = min(vv < 0 ? -vv : vv, 15);
            const uint4 mk = s_mask[av * 48 + B.sh * 3 + piece];
            const uint32_t d0 = B.d[i].x & mk.x, d1 = B.d[i].y & mk.y, d2 = B.d[i].z & mk.z, d3 = B.d[i].w & mk.w;
            sum_p = __dp4a(d0, pos0, sum_p);
            sum_p = __dp4a(d1, pos0 + 0x04040404u, sum_p);
            sum_p = __dp4a(d2, pos0 + 0x08080808u, sum_p);
            sum_p = __dp4a(d3, pos0 + 0x0C0C0C0Cu, sum_p);
            unsigned rs = __dp4a(d0, 0x01010101u, 0u);
            rs = __dp4a(d1, 0x01010101u, rs);
            rs = __dp4a(d2, 0x01010101u, rs);
            rs = __dp4a(d3, 0x01010101u, rs);
            sum_1 += rs;
            m01 += vv * (int)rs;
        }
        const int m10 = __reduce_add_sync(0xFFFFFFFFu, (int)sum_p - (int)(B.sh + kHalfPatch) * (int)sum_1);
        m01 = __reduce_add_sync(0xFFFFFFFFu, m01);
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // ---- descriptor: lane = output byte, 8 point pairs each ----
        const float arad = (float)((double)angle * 3.14159265358979323846 / 180.0);
        const float ca = ref_cos(arad);
        const float sa = ref_cos(__fsub_rn(3.14159265358979f / 2.0f, arad));
        // the patch of this keypoint was sent off one iteration ago, into the buffer the next fetch does not use
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncwarp();
        const uint8_t* cb = my_tap + (tap_turn ^ 1) * kTapBuf + kTapReach * kTapPitch + kTapReach + B.tsh;
        int v0[8], v1[8];
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const float x0 = pat[4 * b], y0 = pat[4 * b + 1], x1 = pat[4 * b + 2], y1 = pat[4 * b + 3];
            const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sa), __fmul_rn(y0, ca)));
            const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sa)));
            const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
            const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
            v0[b] = cb[r0 * kTapPitch + q0];
            v1[b] = cb[r1 * kTapPitch + q1];
        }
        // the next keypoint's rows leave now, behind this one's descriptor taps
        const unsigned k = B.k;
        const int level = B.level, px = B.px, py = B.py, score = B.score;
        B = fetch(A);
        A = A2;
        unsigned byte = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) byte |= (unsigned)(v0[b] < v1[b]) << b;
        a.desc[(size_t)k * 32 + lane] = (uint8_t)byte;

        if (lane == 0) {
            lpslam_keypoint kp;
            kp.x = (float)px;
            kp.y = (float)py;
            if (level != 0) {  // correct_keypoint_scale, orb_extractor.cc:601-609
                kp.x = __fmul_rn(kp.x, lv.scale);
                kp.y = __fmul_rn(kp.y, lv.scale);
            }
            kp.size = lv.kp_size;
            kp.angle = angle;
            kp.response = (float)score;
            kp.octave = level;
            kp.class_id = -1;
            a.kp[k] = kp;
        }
    }
}

}  // namespace

void launch_offsets(const DescribeArgs& a, cudaStream_t st) { k_offsets<<<1, 1024, 0, st>>>(a); }

void launch_describe(const DescribeArgs& a, cudaStream_t st) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, kDescribeSmem);
        attr_set = true;
    }
    k_describe<<<sms * kDescribeBlocksPerSM, 256, kDescribeSmem, st>>>(a);
}

}  // namespace lps
