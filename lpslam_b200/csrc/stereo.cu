// K9 — stereo row matcher with sub-pixel SAD refinement. Replaces match::stereo::compute (match/stereo.cc:18-249).
//   k_stereo_prep    get_right_keypoint_indices_in_each_row (:114-141) folded into a per-keypoint row interval
//   k_stereo_match   find_closest_keypoints_in_stereo (:143-176) + compute_subpixel_disparity (:178-249), warp per left keypoint
//   k_stereo_median  the 2 x median SAD rejection (:94-111), block per stereo pair
// The reference buckets right keypoints by image row; the candidate set of a left keypoint at row y is exactly the right
// keypoints whose interval [floor(y_r - 2 s), ceil(y_r + 2 s)] contains y, scanned in increasing index with a strict <,
// i.e. the minimum of (distance, index). The SAD of centre-subtracted 11x11 patches is integer valued, so it is summed in
// int32 and converted once (the reference's float / double sums are exact for the same reason).
#include "common.cuh"

namespace lps {

namespace {

constexpr unsigned kHammThr = 75;  // (HAMMING_DIST_THR_HIGH + HAMMING_DIST_THR_LOW) / 2, match/stereo.h:99
constexpr int kWin = 5, kSlide = 5;

__device__ __forceinline__ int frame_of_row(const uint32_t* __restrict__ off, int n_frames, uint32_t k) {
    int lo = 0, hi = n_frames;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= k) lo = mid;
        else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(256) k_stereo_prep(StereoArgs a) {
    const uint32_t total = min(a.frame_off[2 * a.n_pairs], a.max_rows);
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < total; k += gridDim.x * blockDim.x) {
        const lpslam_keypoint kp = a.kp[k];
        const float r = __fmul_rn(2.0f, a.lv[kp.octave].scale);
        int max_r = __float2int_ru(__fadd_rn(kp.y, r));
        int min_r = __float2int_rd(__fsub_rn(kp.y, r));
        // rows outside the image make the reference's .at() throw; extractor output never gets there
        min_r = max(min_r, 0);
        max_r = min(max_r, a.rows0 - 1);
        a.rec[k] = make_uint2((uint32_t)min_r | ((uint32_t)max_r << 16), (uint32_t)kp.octave);
        a.x_right[k] = -1.0f;
        a.depth[k] = -1.0f;
        a.corr[k] = -1;
    }
}

// Right-eye keypoints of one pair ordered by image row (counting sort on floor(y)): the candidates of a left keypoint at row
// y are then a contiguous slice of `perm` (rows y - R .. y + R, R = the largest half interval), not the whole right eye.
__global__ void __launch_bounds__(1024) k_stereo_rows(StereoArgs a) {
    extern __shared__ uint32_t s_rows[];  // rows0 + 1 counters, then cursors
    __shared__ uint32_t s_part[33];
    const int pair = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t rbeg = min(a.frame_off[2 * pair + 1], a.max_rows), rend = min(a.frame_off[2 * pair + 2], a.max_rows);
    const int rows = a.rows0;
    for (int r = tid; r <= rows; r += 1024) s_rows[r] = 0;
    __syncthreads();
    for (uint32_t k = rbeg + tid; k < rend; k += 1024) atomicAdd(&s_rows[min(max((int)a.kp[k].y, 0), rows - 1)], 1u);
    __syncthreads();
    // exclusive scan over rows: every thread owns a contiguous chunk
    const int chunk = (rows + 1 + 1023) / 1024;
    const int c0 = min(tid * chunk, rows + 1), c1 = min(c0 + chunk, rows + 1);
    uint32_t sum = 0;
    for (int r = c0; r < c1; ++r) sum += s_rows[r];
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_part[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_part[lane];
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
            if (lane >= o) wi += t;
        }
        s_part[lane] = wi - w;
    }
    __syncthreads();
    uint32_t run = s_part[warp] + inc - sum;
    uint32_t* __restrict__ rs = a.row_start + (size_t)pair * (rows + 1);
    for (int r = c0; r < c1; ++r) {
        const uint32_t c = s_rows[r];
        s_rows[r] = run;
        rs[r] = run;
        run += c;
    }
    __syncthreads();
    for (uint32_t k = rbeg + tid; k < rend; k += 1024) {
        const lpslam_keypoint kp = a.kp[k];
        const uint32_t pos = atomicAdd(&s_rows[min(max((int)kp.y, 0), rows - 1)], 1u);
        const uint2 rec = a.rec[k];
        // everything the candidate scan tests, in scan order: (row interval, octave, x, packed row of the keypoint)
        a.perm[rbeg + pos] = make_uint4(rec.x, rec.y, __float_as_uint(kp.x), k);
    }
}

__device__ __forceinline__ unsigned hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x)
           + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(256) k_stereo_match(StereoArgs a) {
    const int lane = threadIdx.x & 31;
    const int warps_total = gridDim.x * (blockDim.x >> 5);
    const int n_frames = 2 * a.n_pairs;
    const uint32_t total = min(a.frame_off[n_frames], a.max_rows);
    const float max_disp = __fdiv_rn(a.focal_x_baseline, a.true_baseline), min_disp = 0.0f;  // match/stereo.cc:16
    // a right keypoint at y covers rows floor(y - 2 s) .. ceil(y + 2 s): at most ceil(2 s_max) + 1 rows from floor(y)
    const int reach = __float2int_ru(__fmul_rn(2.0f, a.lv[a.num_levels - 1].scale)) + 1;
    for (uint32_t k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); k < total; k += warps_total) {
        const int f = frame_of_row(a.frame_off, n_frames, k);
        if (f & 1) continue;  // right-eye row
        const lpslam_keypoint kl = a.kp[k];
        const int row = (int)kl.y;  // indices_right_in_row.at(y_left): float -> size_t truncation
        const float min_xr = __fsub_rn(kl.x, max_disp), max_xr = __fsub_rn(kl.x, min_disp);
        if (max_xr < 0) continue;
        const uint32_t rbeg = a.frame_off[f + 1], rend = min(a.frame_off[f + 2], a.max_rows);
        const uint4* dl = reinterpret_cast<const uint4*>(a.desc + (size_t)k * 32);
        const uint4 l0 = __ldg(dl), l1 = __ldg(dl + 1);
        uint32_t best = (kHammThr << 16);  // (distance << 16 | right index): strict < in index order == minimum key
        // candidates: right keypoints whose row interval contains `row`; they sit within `reach` rows of it in the row order
        const int pair = f >> 1;
        const uint32_t* __restrict__ rs = a.row_start + (size_t)pair * (a.rows0 + 1);
        const uint32_t s_lo = rs[max(row - reach, 0)], s_hi = rs[min(row + reach + 1, a.rows0)];
        for (uint32_t s = s_lo + lane; s < s_hi; s += 32) {
            const uint4 rec = a.perm[rbeg + s];
            const uint32_t ir = rec.w;
            const int min_r = (int)(rec.x & 0xFFFFu), max_r = (int)(rec.x >> 16), oct = (int)rec.y;
            if (row < min_r || row > max_r) continue;
            if (oct < kl.octave - 1 || oct > kl.octave + 1) continue;
            const float xr = __uint_as_float(rec.z);
            if (xr < min_xr || max_xr < xr) continue;
            const uint4* dr = reinterpret_cast<const uint4*>(a.desc + (size_t)ir * 32);
            const unsigned d = hamming256(l0, l1, __ldg(dr), __ldg(dr + 1));
            const uint32_t key = (d << 16) | (ir - rbeg);
            best = min(best, key);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xFFFFFFFFu, best, o));
        if ((best >> 16) >= kHammThr) continue;
        const uint32_t ir = rbeg + (best & 0xFFFFu);

        // ---- compute_subpixel_disparity on the left keypoint's pyramid level ----
        const int oct = kl.octave;
        const LevelGeom& lv = a.lv[oct];
        const float xr = a.kp[ir].x;
        const int sxl = __float2int_rn(__fmul_rn(kl.x, lv.inv_scale));
        const int syl = __float2int_rn(__fmul_rn(kl.y, lv.inv_scale));
        const int sxr = __float2int_rn(__fmul_rn(xr, lv.inv_scale));
        if (sxr - kSlide - kWin < 0 || lv.w <= sxr + kSlide + kWin) continue;
        // the reference's rowRange / colRange on the left patch would throw outside the image; skip instead
        if (syl - kWin < 0 || syl + kWin >= lv.h || sxl - kWin < 0 || sxl + kWin >= lv.w) continue;
        const uint8_t* __restrict__ L = lv.img + (size_t)f * lv.frame_stride;
        const uint8_t* __restrict__ R = lv.img + (size_t)(f + 1) * lv.frame_stride;
        const int cl = L[(size_t)syl * lv.pitch + sxl];
        int sad[2 * kSlide + 1];
#pragma unroll
        for (int o = 0; o <= 2 * kSlide; ++o) sad[o] = 0;
        for (int p = lane; p < 121; p += 32) {
            const int dy = p / 11 - kWin, dx = p % 11 - kWin;
            const int av = (int)L[(size_t)(syl + dy) * lv.pitch + sxl + dx] - cl;
            const uint8_t* rrow = R + (size_t)(syl + dy) * lv.pitch + sxr + dx;
            const uint8_t* rc = R + (size_t)syl * lv.pitch + sxr;
#pragma unroll
            for (int o = 0; o <= 2 * kSlide; ++o) {
                const int bv = (int)rrow[o - kSlide] - (int)rc[o - kSlide];
                sad[o] += abs(av - bv);
            }
        }
#pragma unroll
        for (int o = 0; o <= 2 * kSlide; ++o) {
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) sad[o] += __shfl_xor_sync(0xFFFFFFFFu, sad[o], s);
        }
        if (lane != 0) continue;
        float best_corr = 4294967296.0f;  // (float)UINT_MAX
        int best_off = 0;
#pragma unroll
        for (int o = 0; o <= 2 * kSlide; ++o) {
            const float c = (float)sad[o];
            if (c < best_corr) {
                best_corr = c;
                best_off = o - kSlide;
            }
        }
        if (best_off == -kSlide || best_off == kSlide) continue;
        float c1 = 0.f, c2 = 0.f, c3 = 0.f;
#pragma unroll
        for (int o = 1; o < 2 * kSlide; ++o) {
            if (o - kSlide == best_off) {
                c1 = (float)sad[o - 1];
                c2 = (float)sad[o];
                c3 = (float)sad[o + 1];
            }
        }
        const float x_delta = (float)__ddiv_rn((double)__fsub_rn(c1, c3),
                                               __dsub_rn(__dmul_rn(2.0, (double)__fadd_rn(c1, c3)), __dmul_rn(4.0, (double)c2)));
        if (x_delta < -1.0f || 1.0f < x_delta) continue;
        float best_xr = __fmul_rn(lv.scale, __fadd_rn((float)(sxr + best_off), x_delta));
        float best_disp = __fsub_rn(kl.x, best_xr);
        if (best_disp < min_disp || max_disp <= best_disp) continue;
        if (best_disp <= 0.0f) {
            best_disp = 0.01f;
            best_xr = __fsub_rn(kl.x, best_disp);
        }
        a.depth[k] = __fdiv_rn(a.focal_x_baseline, best_disp);
        a.x_right[k] = best_xr;
        a.corr[k] = (int)best_corr;
    }
}

// median of the accepted SADs of one pair by a two-pass radix select (values < 2^16), then the 2 x median cut
__global__ void __launch_bounds__(1024) k_stereo_median(StereoArgs a) {
    __shared__ int s_hist[256];
    __shared__ int s_sel[3];
    const int pair = blockIdx.x, tid = threadIdx.x;
    const uint32_t beg = min(a.frame_off[2 * pair], a.max_rows), end = min(a.frame_off[2 * pair + 1], a.max_rows);
    // count accepted
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    for (uint32_t k = beg + tid; k < end; k += 1024) {
        const int c = a.corr[k];
        if (c >= 0) atomicAdd(&s_hist[min(c >> 8, 255)], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int n = 0;
        for (int b = 0; b < 256; ++b) n += s_hist[b];
        s_sel[2] = n;
        int rank = n / 2, b = 0;  // correlation_and_idx_left.at(size / 2) of the sorted list
        if (n > 0) {
            while (rank >= s_hist[b]) {
                rank -= s_hist[b];
                ++b;
            }
        }
        s_sel[0] = b;
        s_sel[1] = rank;
    }
    __syncthreads();
    const int n = s_sel[2];
    if (n == 0) return;
    const int hb = s_sel[0], rank = s_sel[1];
    __syncthreads();
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    for (uint32_t k = beg + tid; k < end; k += 1024) {
        const int c = a.corr[k];
        if (c >= 0 && min(c >> 8, 255) == hb) atomicAdd(&s_hist[c & 255], 1);
    }
    __syncthreads();
    if (tid == 0) {
        int r = rank, b = 0;
        while (r >= s_hist[b]) {
            r -= s_hist[b];
            ++b;
        }
        s_sel[0] = (hb << 8) | b;
    }
    __syncthreads();
    const float median = (float)s_sel[0];
    const float thr = (float)(2.0 * (double)median);
    for (uint32_t k = beg + tid; k < end; k += 1024) {
        const int c = a.corr[k];
        if (c >= 0 && thr < (float)c) {
            a.x_right[k] = -1.0f;
            a.depth[k] = -1.0f;
        }
    }
}

}  // namespace

void launch_stereo(const StereoArgs& a, cudaStream_t st) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    k_stereo_prep<<<sms * 2, 256, 0, st>>>(a);
    k_stereo_rows<<<a.n_pairs, 1024, (size_t)(a.rows0 + 1) * sizeof(uint32_t), st>>>(a);
    k_stereo_match<<<sms * 8, 256, 0, st>>>(a);
    k_stereo_median<<<a.n_pairs, 1024, 0, st>>>(a);
}

}  // namespace lps
