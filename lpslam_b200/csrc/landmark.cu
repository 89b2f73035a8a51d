// Representative descriptor of a landmark: data::landmark::compute_descriptor (data/landmark.cc:127-183), batched over
// landmarks. For a landmark observed n times: all pairwise Hamming distances (match/base.h:18-40), per observation the
// value at position unsigned(0.5 (n - 1)) of its sorted distance row, and the first observation with the smallest such
// median wins (strict '<', initial best 256 / index 0).
//
// One warp per landmark. The k-th smallest of a row comes from a 257-bin histogram in shared memory (distances are
// 0..256) and a warp prefix scan, so n is not limited and nothing is sorted.
#include "common.cuh"

#include <algorithm>

namespace lps {
namespace {

constexpr int LM_WARPS = 8;
constexpr int BINS = 288;  // 257 rounded up to a multiple of 32

__global__ void __launch_bounds__(LM_WARPS * 32) k_landmark_descriptor(const uint8_t* __restrict__ desc, const uint32_t* __restrict__ offsets,
                                                                        uint32_t n_lm, uint32_t* __restrict__ best_idx,
                                                                        uint8_t* __restrict__ desc_out) {
    __shared__ uint32_t s_hist[LM_WARPS][BINS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* hist = s_hist[warp];
    for (uint32_t lm = blockIdx.x * LM_WARPS + warp; lm < n_lm; lm += gridDim.x * LM_WARPS) {
        const uint32_t b = offsets[lm], n = offsets[lm + 1] - b;
        unsigned best_median = 256, best = 0;
        if (n > 0) {
            const unsigned k = (unsigned)__dmul_rn(0.5, (double)(n - 1));
            for (uint32_t i = 0; i < n; ++i) {
                for (int q = lane; q < BINS; q += 32) hist[q] = 0;
                __syncwarp();
                const uint4* di = reinterpret_cast<const uint4*>(desc + (size_t)(b + i) * 32);
                const uint4 a0 = __ldg(di), a1 = __ldg(di + 1);
                for (uint32_t j = lane; j < n; j += 32) {
                    const uint4* dj = reinterpret_cast<const uint4*>(desc + (size_t)(b + j) * 32);
                    const uint4 c0 = __ldg(dj), c1 = __ldg(dj + 1);
                    const unsigned d = __popc(a0.x ^ c0.x) + __popc(a0.y ^ c0.y) + __popc(a0.z ^ c0.z) + __popc(a0.w ^ c0.w)
                                       + __popc(a1.x ^ c1.x) + __popc(a1.y ^ c1.y) + __popc(a1.z ^ c1.z) + __popc(a1.w ^ c1.w);
                    atomicAdd(&hist[d], 1u);
                }
                __syncwarp();
                // smallest d with count(dist <= d) >= k + 1
                unsigned carry = 0, median = 256;
                for (int q0 = 0; q0 < BINS; q0 += 32) {
                    const unsigned v = hist[q0 + lane];
                    unsigned inc = v;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const unsigned t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                        if (lane >= o) inc += t;
                    }
                    const unsigned hit = __ballot_sync(0xFFFFFFFFu, carry + inc >= k + 1);
                    if (hit) {
                        median = q0 + __ffs(hit) - 1;
                        break;
                    }
                    carry += __shfl_sync(0xFFFFFFFFu, inc, 31);
                }
                if (median < best_median) {
                    best_median = median;
                    best = i;
                }
                __syncwarp();
            }
        }
        if (lane == 0) best_idx[lm] = best;
        if (desc_out && n > 0) desc_out[(size_t)lm * 32 + lane] = desc[(size_t)(b + best) * 32 + lane];
    }
}

}  // namespace
}  // namespace lps

using namespace lps;

extern "C" int lpslam_landmark_descriptors(const uint8_t* obs_desc, const uint32_t* offsets, uint32_t n_landmarks, uint32_t* best_idx,
                                           uint8_t* desc_out) {
    LPS_ARG(offsets && best_idx, "null argument");
    if (n_landmarks == 0) return LPSLAM_OK;
    const uint32_t total = offsets[n_landmarks];
    LPS_ARG(total == 0 || obs_desc, "null descriptor array");
    for (uint32_t i = 0; i < n_landmarks; ++i) LPS_ARG(offsets[i] <= offsets[i + 1], "offsets must be non-decreasing");
    ArenaBump ar;
    const size_t o_d = ar.take(32 * (size_t)total), o_o = ar.take(4 * (size_t)(n_landmarks + 1)), o_b = ar.take(4 * (size_t)n_landmarks);
    const size_t o_out = ar.take(32 * (size_t)n_landmarks);
    uint8_t* dev = nullptr;
    LPS_CUDA(scratch_alloc((void**)&dev, ar.off));
    cudaError_t e = cudaSuccess;
    do {
        if (total && (e = cudaMemcpy(dev + o_d, obs_desc, 32 * (size_t)total, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if ((e = cudaMemcpy(dev + o_o, offsets, 4 * (size_t)(n_landmarks + 1), cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if (desc_out && (e = cudaMemset(dev + o_out, 0, 32 * (size_t)n_landmarks)) != cudaSuccess) break;
        const int blocks = (int)std::min<uint32_t>((n_landmarks + LM_WARPS - 1) / LM_WARPS, 148 * 8);
        k_landmark_descriptor<<<blocks, LM_WARPS * 32>>>(dev + o_d, (const uint32_t*)(dev + o_o), n_landmarks, (uint32_t*)(dev + o_b),
                                                         desc_out ? dev + o_out : nullptr);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cudaMemcpy(best_idx, dev + o_b, 4 * (size_t)n_landmarks, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
        if (desc_out && (e = cudaMemcpy(desc_out, dev + o_out, 32 * (size_t)n_landmarks, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
    } while (false);
    scratch_free(dev);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_landmark_descriptors", __FILE__, __LINE__);
    return LPSLAM_OK;
}
