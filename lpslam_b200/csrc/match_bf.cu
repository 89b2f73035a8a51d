// K7 + K8 — brute-force Hamming matcher with ratio test. Replaces match::robust::brute_force_match (match/robust.cc:215-321)
// and carries the Hamming primitive of match/base.h:18-63.
//
// The reference scans keyframe rows in order; each row takes the best and second-best frame descriptor AMONG THOSE NOT YET
// CLAIMED by an earlier row (strict <, so the lowest index wins ties), accepts on best <= 50 and ratio * second >= best,
// and claims its best. The claim makes the loop order-dependent, so the work is split:
//   k_bf_topk     all rows in parallel: the 4 smallest keys (distance << 16 | index) of every row ignoring claims.
//                 In key order, "best / second among unclaimed" is simply the first two unclaimed keys.
//   k_bf_resolve  one warp replays the rows in order against a shared-memory claim bitmap, 32 rows per round (commit up to the
//                 first conflict). A row whose top-4 list runs out of unclaimed entries before the decision is certain is
//                 rescanned exactly by the warp.
#include "common.cuh"
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <numeric>

namespace lps {

namespace {

constexpr int K = 4;
constexpr unsigned HAMMING_DIST_THR_LOW = 50, MAX_HAMMING_DIST = 256;  // match/base.h:13-15
constexpr uint32_t EMPTY = 0xFFFFFFFFu;

__device__ __forceinline__ unsigned hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x)
           + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ void insert4(uint32_t (&t)[K], uint32_t key) {
    // keep t sorted ascending
    if (key < t[3]) {
        t[3] = key;
        if (t[3] < t[2]) { const uint32_t x = t[2]; t[2] = t[3]; t[3] = x; }
        if (t[2] < t[1]) { const uint32_t x = t[1]; t[1] = t[2]; t[2] = x; }
        if (t[1] < t[0]) { const uint32_t x = t[0]; t[0] = t[1]; t[1] = x; }
    }
}

__global__ void __launch_bounds__(256) k_bf_topk(const uint8_t* __restrict__ desc_1, uint32_t n1, const uint8_t* __restrict__ desc_2,
                                                 uint32_t n2, const uint8_t* __restrict__ valid_2, uint4* __restrict__ topk) {
    const int lane = threadIdx.x & 31;
    const uint32_t row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n2) return;
    if (valid_2 && !valid_2[row]) {
        if (lane == 0) topk[row] = make_uint4(EMPTY, EMPTY, EMPTY, EMPTY);
        return;
    }
    const uint4* d2 = reinterpret_cast<const uint4*>(desc_2 + (size_t)row * 32);
    const uint4 a0 = __ldg(d2), a1 = __ldg(d2 + 1);
    uint32_t t[K] = {EMPTY, EMPTY, EMPTY, EMPTY};
    for (uint32_t i = lane; i < n1; i += 32) {
        const uint4* d1 = reinterpret_cast<const uint4*>(desc_1 + (size_t)i * 32);
        const unsigned d = hamming256(a0, a1, __ldg(d1), __ldg(d1 + 1));
        insert4(t, (d << 16) | i);
    }
    // warp merge: four rounds of "smallest head wins and pops"
    uint32_t out[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        uint32_t m = t[0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
        out[r] = m;
        if (t[0] == m && m != EMPTY) {  // keys are unique (index in the low bits)
            t[0] = t[1];
            t[1] = t[2];
            t[2] = t[3];
            t[3] = EMPTY;
        }
    }
    if (lane == 0) topk[row] = make_uint4(out[0], out[1], out[2], out[3]);
}

// One warp, rows in reference order, 32 rows per round: every lane decides its own row from its four kept keys against the
// claim bitmap; rows are committed in order up to the first one whose kept keys contain a keypoint that an earlier row of
// the round just claimed; that row and the ones behind it are decided again. A row whose four keys do not settle the
// decision is rescanned exactly by the whole warp when it is the first pending row.
// s_first (optional, n1 words behind the bitmap): lowest lane claiming a keypoint in the current round.
__global__ void __launch_bounds__(32) k_bf_resolve(const uint8_t* __restrict__ desc_1, uint32_t n1, const uint8_t* __restrict__ desc_2,
                                                   uint32_t n2, const uint4* __restrict__ topk, float lowe_ratio, int first_in_smem,
                                                   int32_t* __restrict__ match_2_in_1, uint32_t* __restrict__ n_matches) {
    extern __shared__ uint32_t s_claim[];  // bitmap over frame keypoints
    const int lane = threadIdx.x;
    const uint32_t words = (n1 + 31) / 32;
    uint32_t* s_first = first_in_smem ? s_claim + words : nullptr;
    for (uint32_t i = lane; i < words; i += 32) s_claim[i] = 0;
    for (uint32_t i = lane; i < n1; i += 32) {
        match_2_in_1[i] = -1;
        if (s_first) s_first[i] = 0xFFFFFFFFu;
    }
    __syncwarp();
    uint32_t num = 0;
    for (uint32_t base = 0; base < n2; base += 32) {
        const uint32_t row = base + lane;
        const uint4 mk = row < n2 ? topk[row] : make_uint4(EMPTY, EMPTY, EMPTY, EMPTY);
        const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
        // rows whose unconditional best is already > 50 can never match and claim nothing
        uint32_t pending = __ballot_sync(0xFFFFFFFFu, mk.x != EMPTY && (mk.x >> 16) <= HAMMING_DIST_THR_LOW);
        while (pending) {
            const bool mine = (pending >> lane) & 1u;
            // first two unclaimed keys of the list
            uint32_t best = EMPTY, second = EMPTY;
            int seen = 0;
            if (mine) {
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    if (k[q] == EMPTY) continue;
                    const uint32_t i1 = k[q] & 0xFFFFu;
                    if ((s_claim[i1 >> 5] >> (i1 & 31)) & 1u) continue;
                    if (seen == 0) best = k[q];
                    else if (seen == 1) second = k[q];
                    ++seen;
                }
            }
            const bool list_complete = (k[K - 1] == EMPTY);  // fewer than K frame keypoints exist: nothing beyond the list
            bool exact = (seen >= 2) || list_complete;
            if (mine && !exact && seen == 1) {
                // second >= the list's last distance; if the ratio test passes with that bound it passes with the truth
                const unsigned lb = k[K - 1] >> 16, bd = best >> 16;
                if (bd > HAMMING_DIST_THR_LOW) exact = true;  // rejected whatever the second is
                else if (!(__fmul_rn(lowe_ratio, (float)lb) < (float)bd)) {
                    exact = true;
                    second = lb << 16;
                }
            }
            const bool rescan = mine && !exact;
            const int first = __ffs(pending) - 1;
            if (__shfl_sync(0xFFFFFFFFu, (int)rescan, first)) {
                // exact rescan of the first pending row over the unclaimed frame keypoints (match/robust.cc:252-274), whole warp
                const uint4* d2 = reinterpret_cast<const uint4*>(desc_2 + (size_t)(base + first) * 32);
                const uint4 a0 = __ldg(d2), a1 = __ldg(d2 + 1);
                uint32_t b = EMPTY, s = EMPTY;
                for (uint32_t i = lane; i < n1; i += 32) {
                    if ((s_claim[i >> 5] >> (i & 31)) & 1u) continue;
                    const uint4* d1 = reinterpret_cast<const uint4*>(desc_1 + (size_t)i * 32);
                    const uint32_t key = (hamming256(a0, a1, __ldg(d1), __ldg(d1 + 1)) << 16) | i;
                    if (key < b) { s = b; b = key; }
                    else if (key < s) s = key;
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, b, o), os = __shfl_xor_sync(0xFFFFFFFFu, s, o);
                    const uint32_t nb = min(b, ob);
                    const uint32_t ns = min(max(b, ob), min(s, os));
                    b = nb;
                    s = ns;
                }
                if (lane == first) {
                    best = b;
                    second = s;
                }
            }
            bool accept = mine && best != EMPTY && (lane == first || !rescan);
            if (accept) {
                const unsigned bd = best >> 16;
                const unsigned sd = second == EMPTY ? MAX_HAMMING_DIST : (second >> 16);
                if (HAMMING_DIST_THR_LOW < bd) accept = false;
                else if (__fmul_rn(lowe_ratio, (float)sd) < (float)bd) accept = false;  // match/robust.cc:285
            }
            const uint32_t i1 = best & 0xFFFFu;
            bool wait = mine && rescan && lane != first;
            if (s_first) {
                if (accept) atomicMin(&s_first[i1], (uint32_t)lane);
                __syncwarp();
                if (mine) {
#pragma unroll
                    for (int q = 0; q < K; ++q)
                        if (k[q] != EMPTY && s_first[k[q] & 0xFFFFu] < (uint32_t)lane) wait = true;
                }
                __syncwarp();
                if (accept) s_first[i1] = 0xFFFFFFFFu;
            }
            else {
                for (uint32_t mm = pending; mm; mm &= mm - 1) {
                    const int M = __ffs(mm) - 1;
                    const int accM = __shfl_sync(0xFFFFFFFFu, (int)accept, M);
                    const uint32_t iM = __shfl_sync(0xFFFFFFFFu, i1, M);
                    if (mine && lane > M && accM) {
#pragma unroll
                        for (int q = 0; q < K; ++q)
                            if (k[q] != EMPTY && (k[q] & 0xFFFFu) == iM) wait = true;
                    }
                }
            }
            const uint32_t waiting = __ballot_sync(0xFFFFFFFFu, wait) & pending;
            const uint32_t commit = waiting ? (pending & ((1u << (__ffs(waiting) - 1)) - 1u)) : pending;
            const bool go = ((commit >> lane) & 1u) && accept;
            if (go) {
                atomicOr(&s_claim[i1 >> 5], 1u << (i1 & 31));
                match_2_in_1[i1] = (int32_t)row;
            }
            num += __popc(__ballot_sync(0xFFFFFFFFu, go));
            pending &= ~commit;
            __syncwarp();
        }
    }
    if (lane == 0) *n_matches = num;
}

__global__ void k_hamming_pairs(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, uint32_t n, uint32_t* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* pa = reinterpret_cast<const uint4*>(a + (size_t)i * 32);
    const uint4* pb = reinterpret_cast<const uint4*>(b + (size_t)i * 32);
    out[i] = hamming256(__ldg(pa), __ldg(pa + 1), __ldg(pb), __ldg(pb + 1));
}

}  // namespace

// match/angle_checker.h:60-134 (histogram_length 30, num_bins_thr 3) — the same std::sort call on the same container as
// the reference, so libstdc++'s treatment of equal bin sizes is reproduced. Host-side: <= n1 scalar entries.
std::vector<int> angle_invalid_matches(const std::vector<std::pair<float, int>>& delta_and_id) {
    constexpr unsigned kLen = 30, kTop = 3;
    const float inv_len = 1.0f / kLen;
    std::vector<std::vector<int>> bins(kLen);
    for (const auto& e : delta_and_id) {
        float delta = e.first;
        if (delta < 0.0) delta += 360.0;
        if (360.0 <= delta) delta -= 360.0;
        const unsigned bin = (unsigned)std::lrint((double)(delta * inv_len));
        bins.at(bin).push_back(e.second);
    }
    std::vector<unsigned int> order(kLen);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&bins](const unsigned int a, const unsigned int b) { return bins.at(a).size() > bins.at(b).size(); });
    std::vector<int> out;
    for (unsigned bin = 0; bin < kLen; ++bin) {
        const bool keep = std::any_of(order.begin(), order.begin() + kTop, [bin](unsigned i) { return bin == i; });
        if (!keep) out.insert(out.end(), bins[bin].begin(), bins[bin].end());
    }
    return out;
}

}  // namespace lps

using namespace lps;

extern "C" {

int lpslam_match_bruteforce(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint8_t* valid_2,
                            const float* angle_1, const float* angle_2, float lowe_ratio, int check_orientation,
                            int32_t* match_2_in_1, uint32_t* n_matches) {
    LPS_ARG(n_matches, "n_matches is null");
    *n_matches = 0;
    LPS_ARG(n1 == 0 || (desc_1 && match_2_in_1), "null frame-side argument");
    LPS_ARG(n2 == 0 || desc_2, "null keyframe descriptors");
    LPS_ARG(n1 <= 65535, "more than 65535 frame keypoints are not supported by this matcher");
    LPS_ARG(!check_orientation || (angle_1 && angle_2), "angles required for the orientation check");
    for (uint32_t i = 0; i < n1; ++i) match_2_in_1[i] = -1;
    if (n1 == 0 || n2 == 0) return LPSLAM_OK;
    uint8_t* dev = nullptr;
    const size_t o_d1 = 0, o_d2 = align_up(32 * (size_t)n1, 256), o_v = o_d2 + align_up(32 * (size_t)n2, 256);
    const size_t o_top = o_v + align_up(n2, 256), o_m = o_top + align_up(16 * (size_t)n2, 256), o_n = o_m + align_up(4 * (size_t)n1, 256);
    LPS_CUDA(scratch_alloc((void**)&dev, o_n + 256));
    int rc = LPSLAM_OK;
    cudaError_t e;
    do {
        if ((e = cudaMemcpy(dev + o_d1, desc_1, 32 * (size_t)n1, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if ((e = cudaMemcpy(dev + o_d2, desc_2, 32 * (size_t)n2, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if (valid_2 && (e = cudaMemcpy(dev + o_v, valid_2, n2, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        k_bf_topk<<<(n2 + 7) / 8, 256>>>(dev + o_d1, n1, dev + o_d2, n2, valid_2 ? dev + o_v : nullptr, (uint4*)(dev + o_top));
        const int first_in_smem = n1 <= 10000;
        k_bf_resolve<<<1, 32, ((n1 + 31) / 32) * 4 + (first_in_smem ? 4 * (size_t)n1 : 0)>>>(
            dev + o_d1, n1, dev + o_d2, n2, (const uint4*)(dev + o_top), lowe_ratio, first_in_smem, (int32_t*)(dev + o_m),
            (uint32_t*)(dev + o_n));
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cudaMemcpy(match_2_in_1, dev + o_m, 4 * (size_t)n1, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
        if ((e = cudaMemcpy(n_matches, dev + o_n, 4, cudaMemcpyDeviceToHost)) != cudaSuccess) break;
    } while (false);
    if (e != cudaSuccess) rc = cuda_fail(e, "lpslam_match_bruteforce", __FILE__, __LINE__);
    scratch_free(dev);
    if (rc == LPSLAM_OK && check_orientation) {
        // appended in keyframe-row order, as the reference does while it scans (match/robust.cc:293-297)
        std::vector<std::pair<int, int>> by_row;
        for (uint32_t i = 0; i < n1; ++i)
            if (match_2_in_1[i] >= 0) by_row.emplace_back(match_2_in_1[i], (int)i);
        std::sort(by_row.begin(), by_row.end());
        std::vector<std::pair<float, int>> deltas;
        for (const auto& p : by_row) deltas.emplace_back(angle_1[p.second] - angle_2[p.first], p.second);
        for (int id : angle_invalid_matches(deltas)) {
            match_2_in_1[id] = -1;
            --*n_matches;
        }
    }
    return rc;
}

int lpslam_match_consecutive_frames(lpslam_orb_ctx* c, float lowe_ratio, int32_t* match_prev_in_cur, uint32_t* n_matches) {
    LPS_ARG(c && match_prev_in_cur && n_matches, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    LPS_CUDA(cudaSetDevice(c->device));
    const uint32_t nf = c->last_batch;
    LPS_CUDA(cudaMemcpyAsync(c->h_frame_off, c->desc.frame_off, sizeof(uint32_t) * (c->max_batch + 2), cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    const uint32_t* off = c->h_frame_off;
    const uint32_t total = off[nf];
    n_matches[0] = 0;
    if (total == 0) return LPSLAM_OK;
    uint32_t max_rows = 0;
    for (uint32_t f = 0; f < nf; ++f) {
        const uint32_t n = off[f + 1] - off[f];
        LPS_ARG(n <= 65535, "more than 65535 keypoints per frame are not supported by this matcher");
        max_rows = std::max(max_rows, n);
    }
    // scratch: kept candidates per pair (one block of max_rows per pair, so that the pairs' kernels are independent), the packed
    // match array and one counter per frame
    uint8_t* dev = nullptr;
    const size_t o_top = 0, top_stride = align_up(16 * (size_t)max_rows, 256);
    const size_t o_m = o_top + top_stride * nf, o_n = o_m + align_up(4 * (size_t)total, 256);
    LPS_CUDA(scratch_alloc((void**)&dev, o_n + align_up(4 * (size_t)nf, 256)));
    cudaError_t e = cudaSuccess;
    do {
        if ((e = cudaMemsetAsync(dev + o_m, 0xFF, 4 * (size_t)total, c->stream)) != cudaSuccess) break;
        if ((e = cudaMemsetAsync(dev + o_n, 0, 4 * (size_t)nf, c->stream)) != cudaSuccess) break;
        // the pairs are independent and each resolve is one serial warp: they are spread over the context's five streams
        cudaStream_t lanes[5] = {c->stream, c->side, c->chain[0], c->chain[1], c->chain[2]};
        cudaEvent_t joins[5] = {nullptr, c->ev_join, c->ev_chain_join[0], c->ev_chain_join[1], c->ev_chain_join[2]};
        if ((e = cudaEventRecord(c->ev_chain_fork, c->stream)) != cudaSuccess) break;
        for (int k = 1; k < 5; ++k) cudaStreamWaitEvent(lanes[k], c->ev_chain_fork, 0);
        for (uint32_t f = 1; f < nf; ++f) {
            cudaStream_t st = lanes[f % 5];
            // side 1 = the current frame f (its keypoints are claimed), side 2 = the previous frame f - 1 (every row valid)
            const uint32_t n1 = off[f + 1] - off[f], n2 = off[f] - off[f - 1];
            if (n1 == 0 || n2 == 0) continue;
            const uint8_t* d1 = c->desc.desc + 32 * (size_t)off[f];
            const uint8_t* d2 = c->desc.desc + 32 * (size_t)off[f - 1];
            uint4* top = (uint4*)(dev + o_top + top_stride * f);
            k_bf_topk<<<(n2 + 7) / 8, 256, 0, st>>>(d1, n1, d2, n2, nullptr, top);
            const int first_in_smem = n1 <= 10000;
            k_bf_resolve<<<1, 32, ((n1 + 31) / 32) * 4 + (first_in_smem ? 4 * (size_t)n1 : 0), st>>>(
                d1, n1, d2, n2, top, lowe_ratio, first_in_smem, (int32_t*)(dev + o_m) + off[f], (uint32_t*)(dev + o_n) + f);
        }
        for (int k = 1; k < 5; ++k) {
            cudaEventRecord(joins[k], lanes[k]);
            cudaStreamWaitEvent(c->stream, joins[k], 0);
        }
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(match_prev_in_cur, dev + o_m, 4 * (size_t)total, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) break;
        if ((e = cudaMemcpyAsync(n_matches, dev + o_n, 4 * (size_t)nf, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess) break;
        e = cudaStreamSynchronize(c->stream);
    } while (false);
    scratch_free(dev);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_consecutive_frames", __FILE__, __LINE__);
    return LPSLAM_OK;
}

int lpslam_hamming_pairs(const uint8_t* a, const uint8_t* b, uint32_t n, uint32_t* dist) {
    LPS_ARG(n == 0 || (a && b && dist), "null argument");
    if (n == 0) return LPSLAM_OK;
    uint8_t* dev = nullptr;
    const size_t sz = align_up(32 * (size_t)n, 256);
    LPS_CUDA(scratch_alloc((void**)&dev, 2 * sz + 4 * (size_t)n));
    cudaError_t e;
    do {
        if ((e = cudaMemcpy(dev, a, 32 * (size_t)n, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        if ((e = cudaMemcpy(dev + sz, b, 32 * (size_t)n, cudaMemcpyHostToDevice)) != cudaSuccess) break;
        k_hamming_pairs<<<(n + 255) / 256, 256>>>(dev, dev + sz, n, (uint32_t*)(dev + 2 * sz));
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        e = cudaMemcpy(dist, dev + 2 * sz, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    } while (false);
    scratch_free(dev);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_hamming_pairs", __FILE__, __LINE__);
    return LPSLAM_OK;
}

}  // extern "C"
