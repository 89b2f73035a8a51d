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

__global__ void __launch_bounds__(256) k_describe(DescribeArgs a) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * (blockDim.x >> 5);
    const int n_seg = a.batch * a.num_levels;
    const unsigned total = min(a.level_off[n_seg], a.out_cap);
    float pat[32];  // this lane's 8 point pairs: x0, y0, x1, y1 each
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const char4 q = reinterpret_cast<const char4*>(g_pattern)[lane * 8 + i];
        pat[4 * i] = (float)q.x;
        pat[4 * i + 1] = (float)q.y;
        pat[4 * i + 2] = (float)q.z;
        pat[4 * i + 3] = (float)q.w;
    }
    for (unsigned k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); k < total; k += warps_total) {
        // segment (frame, level) holding packed row k
        int lo = 0, hi = n_seg;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (a.level_off[mid] <= k) lo = mid;
            else hi = mid;
        }
        const int frame = lo / a.num_levels, level = lo % a.num_levels;
        const unsigned idx = k - a.level_off[lo];
        const LevelGeom& lv = a.lv[level];
        const uint32_t xy = a.sel_xys[(size_t)frame * a.sel_stride + lv.sel_begin + idx];
        const int score = a.sel_sc[(size_t)frame * a.sel_stride + lv.sel_begin + idx];
        const int px = (int)(xy & 0xFFFFu) + kBorder, py = (int)(xy >> 16) + kBorder;
        const int pitch = lv.pitch;

        // ---- orientation: lane = column offset u + 15 (lane 31 idles), rows v = -15..15 ----
        const uint8_t* __restrict__ c0 = lv.img + frame * lv.frame_stride + (size_t)py * pitch + px;
        int m10 = 0, m01 = 0;
        const int u = lane - kHalfPatch;
        const int au = abs(u);
        if (lane < kPatch) {
            // 31 independent predicated byte loads in flight (rows v = -15..15 of this lane's column)
            int val[kPatch];
#pragma unroll
            for (int i = 0; i < kPatch; ++i) {
                const int v = i - kHalfPatch;
                val[i] = (au <= c_umax[v < 0 ? -v : v]) ? (int)__ldg(c0 + v * pitch + u) : 0;
            }
            int colsum = 0;
#pragma unroll
            for (int i = 0; i < kPatch; ++i) {
                colsum += val[i];
                m01 += (i - kHalfPatch) * val[i];
            }
            m10 = u * colsum;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xFFFFFFFFu, m10, o);
            m01 += __shfl_xor_sync(0xFFFFFFFFu, m01, o);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // ---- descriptor: lane = output byte, 8 point pairs each ----
        const float arad = (float)((double)angle * 3.14159265358979323846 / 180.0);
        const float ca = ref_cos(arad);
        const float sa = ref_cos(__fsub_rn(3.14159265358979f / 2.0f, arad));
        const uint8_t* __restrict__ cb = lv.blur + frame * lv.frame_stride + (size_t)py * pitch + px;
        unsigned byte = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const float x0 = pat[4 * b], y0 = pat[4 * b + 1], x1 = pat[4 * b + 2], y1 = pat[4 * b + 3];
            const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sa), __fmul_rn(y0, ca)));
            const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sa)));
            const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sa), __fmul_rn(y1, ca)));
            const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sa)));
            const int v0 = __ldg(cb + r0 * pitch + q0);
            const int v1 = __ldg(cb + r1 * pitch + q1);
            byte |= (unsigned)(v0 < v1) << b;
        }
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
    k_describe<<<sms * 8, 256, 0, st>>>(a);
}

}  // namespace lps
