// K10 — frame-to-frame projection matcher. Replaces match::projection::match_current_and_last_frames
// (match/projection.cc:85-208) together with the grid of data/common.cc:97-186 / data/common.h:59-64 and the perspective
// reprojection of camera/perspective.cc:151-169, over flat arrays.
//
//   k_grid_*        assign_keypoints_to_grid: CSR lists per cell, indices ascending inside a cell (= insertion order);
//                   k_grid_build is the four steps in one block for per-frame sizes
//   k_proj_topk     one warp per row (last-frame landmark / local landmark / frame-1 keypoint): reproject (double, no FMA),
//                   the lanes split the cells of get_keypoints_in_cell's window (x-outer / y-inner), apply the octave / window /
//                   x_right gates and keep the 4 best (distance, visit order)
//   k_lm_resolve    replays the rows against the "already holds an observed landmark" set, which changes as matches are made
//                   (projection.cc:163-165,189) - that is what makes the reference loop sequential. Order-exact parallel rounds
//                   in one block (rows whose outcome no earlier unresolved row can change are committed together), a serial
//                   warp loop for contended inputs; a row whose 4 candidates do not settle its decision is rescanned exactly.
// Every host entry point packs its inputs into one page-locked staging buffer (CallStage, scratch.cu): one copy in, kernels, one
// copy out, on a stream of the calling thread; the grid is built from the first part of the copy while the rest is in flight.
// LPS_MATCH_TIMING=1 prints the host / device time split of lpslam_match_reprojected.
#include "common.cuh"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>

namespace lps {

std::vector<int> angle_invalid_matches(const std::vector<std::pair<float, int>>& delta_and_id);  // match_bf.cu

namespace {

constexpr int K = 4;
constexpr unsigned HAMMING_DIST_THR_HIGH = 100, MAX_HAMMING_DIST = 256;  // match/base.h:14-15
constexpr uint32_t EMPTY = 0xFFFFFFFFu;

struct ProjArgs {
    lpslam_camera cam;
    double Tc[16], Tl[16];
    uint32_t n_last, n_curr;
    const uint8_t* last_has_lm;
    const uint8_t* last_outlier;
    const double* last_pos_w;
    const uint8_t* last_desc;
    const int32_t* last_octave;
    const uint8_t* last_obs;  // nullable
    const lpslam_keypoint* kp;
    const uint8_t* desc;
    const float* x_right;
    const uint8_t* curr_taken;
    const float* scale;
    float margin;
    // grid (CSR over cx * rows + cy)
    uint32_t* cell_count;
    uint32_t* cell_start;
    uint32_t* cell_fill;
    uint32_t* cell_items;
    // per-landmark candidates
    uint4* top_key;
    int4* top_idx;
    // results
    int32_t* curr_match_last;
    int2* events;  // (idx_last, idx_curr) in match order
    uint32_t* n_events;
    // mode 1 = match_frame_and_landmarks (projection.cc:11-83): the "last frame" arrays describe local landmarks that come
    // with their cached reprojection (data/landmark.h:112-115); last_has_lm = is_observable_in_tracking_ && !will_be_erased()
    int mode;  // 2 = area::match_in_consistent_area (area.cc:8-112): rows = frame-1 keypoints, window around prev_pts
    const float* prev_pts;     // n_last x 2 (mode 2)
    int32_t* match_2_of_1;     // n_last (mode 2)
    int32_t* match_1_of_2;     // n_curr (mode 2)
    const double* lm_reproj;   // n_last x 2
    const float* lm_x_right;   // n_last
    float lowe_ratio;
    // mode 1 knobs (lpslam_reproj_match_params): octave window relative to the predicted level, acceptance threshold,
    // Lowe ratio on / off, stereo x_right gate on / off
    int level_lo, level_hi, use_ratio, use_xr;
    unsigned hamm_thr;
    // mode 3 = candidate lists instead of the grid (bow_tree.cc:16-110: the frame keypoints of the row's vocabulary node):
    // row il scans cand_items[cand_begin[il] .. cand_end[il]); the ratio test ignores scale levels
    const uint32_t* cand_begin;
    const uint32_t* cand_end;
    const uint32_t* cand_items;
    int claim_in_smem;  // k_lm_resolve: versioned owner table + claim table in shared memory (else a bitmap, serial loop only)
    // fuse::replace_duplication's reprojection-error gate (fuse.cc:199-222)
    int chi_gate;
    const float* inv_sigma_sq;
    int32_t* row_best;   // k_rows_best: keypoint of every row or -1
    uint32_t* row_dist;  // its Hamming distance (256 when the row saw no candidate)
    // robust::match_for_triangulation (robust.cc:18-178) on candidate lists (mode 3): `best_hamm_dist < hamm_dist` skips, so
    // an equal distance REPLACES the best (the last one in visit order wins); candidates pass the epipole gate (:114-123)
    // and check_epipolar_constraint (:126-127, :323-341) before they count
    int last_wins, tri;
    const double* row_bearing;  // n_last x 3
    const double* kp_bearing;   // n_curr x 3
    const uint8_t* row_stereo;  // 0 <= stereo_x_right of the row's keypoint
    const uint8_t* kp_stereo;
    const float* row_scale;     // scale_factors_.at(keypt_1.octave)
    double E[9], epi[3];
};

// robust.cc:323-341 check_epipolar_constraint: double arithmetic in source order, no contraction
__device__ __forceinline__ bool epipolar_inlier(const double* E, const double* b1, const double* b2, float scale_1) {
    double n[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
        n[r] = __dadd_rn(__dadd_rn(__dmul_rn(E[3 * r], b2[0]), __dmul_rn(E[3 * r + 1], b2[1])), __dmul_rn(E[3 * r + 2], b2[2]));
    const double dot = __dadd_rn(__dadd_rn(__dmul_rn(n[0], b1[0]), __dmul_rn(n[1], b1[1])), __dmul_rn(n[2], b1[2]));
    const double nrm = __dsqrt_rn(__dadd_rn(__dadd_rn(__dmul_rn(n[0], n[0]), __dmul_rn(n[1], n[1])), __dmul_rn(n[2], n[2])));
    const double cos_residual = __ddiv_rn(dot, nrm);
    const double residual_rad = __dsub_rn(M_PI / 2.0, fabs(acos(cos_residual)));
    constexpr double residual_rad_thr = 0.2 * M_PI / 180.0;
    return residual_rad < __dmul_rn(residual_rad_thr, (double)scale_1);
}

__device__ __forceinline__ int cv_round(double v) { return __double2int_rn(v); }
__device__ __forceinline__ int cv_floor(double v) { return __double2int_rd(v); }
__device__ __forceinline__ int cv_ceil(double v) { return __double2int_ru(v); }

__device__ __forceinline__ unsigned hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x)
           + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ bool cell_of(const ProjArgs& a, const lpslam_keypoint& k, int& cx, int& cy) {
    cx = cv_round(__dmul_rn((double)__fsub_rn(k.x, a.cam.min_x), a.cam.inv_cell_width));
    cy = cv_round(__dmul_rn((double)__fsub_rn(k.y, a.cam.min_y), a.cam.inv_cell_height));
    return 0 <= cx && cx < (int)a.cam.grid_cols && 0 <= cy && cy < (int)a.cam.grid_rows;
}

__global__ void k_grid_count(ProjArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_curr) return;
    int cx, cy;
    if (cell_of(a, a.kp[i], cx, cy)) atomicAdd(&a.cell_count[cx * a.cam.grid_rows + cy], 1u);
}

__global__ void __launch_bounds__(1024) k_grid_scan(ProjArgs a) {
    __shared__ uint32_t s_w[33];
    const uint32_t n = a.cam.grid_cols * a.cam.grid_rows;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + tid;
        const uint32_t v = i < n ? a.cell_count[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_w[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = s_w[lane];
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (lane >= o) wi += t;
            }
            s_w[lane] = wi - w;
            if (lane == 31) s_w[32] = wi;
        }
        __syncthreads();
        if (i < n) {
            a.cell_start[i] = carry + s_w[warp] + inc - v;
            a.cell_fill[i] = 0;
        }
        carry += s_w[32];
        __syncthreads();
    }
    if (tid == 0) a.cell_start[n] = carry;
}

__global__ void k_grid_fill(ProjArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n_curr) return;
    int cx, cy;
    if (cell_of(a, a.kp[i], cx, cy)) {
        const uint32_t c = cx * a.cam.grid_rows + cy;
        a.cell_items[a.cell_start[c] + atomicAdd(&a.cell_fill[c], 1u)] = i;
    }
}

__global__ void k_grid_sort(ProjArgs a) {
    // push_back order = ascending keypoint index; cells hold a handful of entries
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= a.cam.grid_cols * a.cam.grid_rows) return;
    const uint32_t b = a.cell_start[c], e = a.cell_start[c + 1];
    for (uint32_t i = b + 1; i < e; ++i) {
        const uint32_t v = a.cell_items[i];
        uint32_t j = i;
        while (j > b && a.cell_items[j - 1] > v) {
            a.cell_items[j] = a.cell_items[j - 1];
            --j;
        }
        a.cell_items[j] = v;
    }
}

// The four grid kernels in one block for the per-frame sizes (a few thousand keypoints): three launches and their gaps
// cost more than the work. Same arithmetic and the same CSR result.
__global__ void __launch_bounds__(1024) k_grid_build(ProjArgs a) {
    __shared__ uint32_t s_w[33];
    const uint32_t n = a.cam.grid_cols * a.cam.grid_rows;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (uint32_t i = tid; i < a.n_curr; i += 1024) {
        int cx, cy;
        if (cell_of(a, a.kp[i], cx, cy)) atomicAdd(&a.cell_count[cx * a.cam.grid_rows + cy], 1u);
    }
    __syncthreads();
    uint32_t carry = 0;
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + tid;
        const uint32_t v = i < n ? a.cell_count[i] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_w[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = s_w[lane];
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (lane >= o) wi += t;
            }
            s_w[lane] = wi - w;
            if (lane == 31) s_w[32] = wi;
        }
        __syncthreads();
        if (i < n) {
            a.cell_start[i] = carry + s_w[warp] + inc - v;
            a.cell_fill[i] = 0;
        }
        carry += s_w[32];
        __syncthreads();
    }
    if (tid == 0) a.cell_start[n] = carry;
    __syncthreads();
    for (uint32_t i = tid; i < a.n_curr; i += 1024) {
        int cx, cy;
        if (cell_of(a, a.kp[i], cx, cy)) {
            const uint32_t c = cx * a.cam.grid_rows + cy;
            a.cell_items[a.cell_start[c] + atomicAdd(&a.cell_fill[c], 1u)] = i;
        }
    }
    __syncthreads();
    for (uint32_t c = tid; c < n; c += 1024) {  // push_back order = ascending keypoint index
        const uint32_t b = a.cell_start[c], e = a.cell_start[c + 1];
        for (uint32_t i = b + 1; i < e; ++i) {
            const uint32_t v = a.cell_items[i];
            uint32_t j = i;
            while (j > b && a.cell_items[j - 1] > v) {
                a.cell_items[j] = a.cell_items[j - 1];
                --j;
            }
            a.cell_items[j] = v;
        }
    }
}

void launch_grid(const ProjArgs& a, uint32_t n, uint32_t ncell, cudaStream_t st) {
    if (n <= 16384) {
        k_grid_build<<<1, 1024, 0, st>>>(a);
        return;
    }
    k_grid_count<<<(n + 255) / 256, 256, 0, st>>>(a);
    k_grid_scan<<<1, 1024, 0, st>>>(a);
    k_grid_fill<<<(n + 255) / 256, 256, 0, st>>>(a);
    k_grid_sort<<<(ncell + 255) / 256, 256, 0, st>>>(a);
}

struct Projected {
    bool ok;
    float ref_x, ref_y, x_right, m;
    int min_cx, max_cx, min_cy, max_cy, min_level, max_level;
};

__device__ Projected cell_window(const ProjArgs& a, Projected p);

__device__ Projected project(const ProjArgs& a, uint32_t il, bool forward, bool backward) {
    Projected p;
    p.ok = false;
    if (a.mode == 3) {
        p.ok = a.cand_begin[il] < a.cand_end[il];
        return p;
    }
    if (a.mode == 2) {
        if (!a.last_has_lm[il]) return p;  // area.cc:23-26: only keypoints of scale level 0
        p.min_level = 0;                   // area.cc:29-30: get_keypoints_in_cell(x, y, margin, level_1, level_1)
        p.max_level = 0;
        p.ref_x = a.prev_pts[2 * (size_t)il];
        p.ref_y = a.prev_pts[2 * (size_t)il + 1];
        p.x_right = 0.f;
        p.m = a.margin;
        return cell_window(a, p);
    }
    if (a.mode == 1) {
        if (!a.last_has_lm[il]) return p;
        const int level = a.last_octave[il];
        p.min_level = level + a.level_lo;  // projection.cc:27-29: (level - 1, level)
        p.max_level = level + a.level_hi;
        p.ref_x = (float)a.lm_reproj[2 * (size_t)il];
        p.ref_y = (float)a.lm_reproj[2 * (size_t)il + 1];
        p.x_right = a.lm_x_right[il];
        p.m = __fmul_rn(a.margin, a.scale[level]);
        return cell_window(a, p);
    }
    if (!a.last_has_lm[il] || a.last_outlier[il]) return p;
    const double* pw = a.last_pos_w + 3 * (size_t)il;
    double pc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        pc[i] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.Tc[i * 4], pw[0]), __dmul_rn(a.Tc[i * 4 + 1], pw[1])), __dmul_rn(a.Tc[i * 4 + 2], pw[2])),
                          a.Tc[i * 4 + 3]);
    if (pc[2] <= 0.0) return p;
    const double z_inv = __ddiv_rn(1.0, pc[2]);
    const double u = __dadd_rn(__dmul_rn(__dmul_rn(a.cam.fx, pc[0]), z_inv), a.cam.cx);
    const double v = __dadd_rn(__dmul_rn(__dmul_rn(a.cam.fy, pc[1]), z_inv), a.cam.cy);
    p.x_right = (float)__dsub_rn(u, __dmul_rn(a.cam.focal_x_baseline, z_inv));
    if (!((double)a.cam.min_x < u && u < (double)a.cam.max_x && (double)a.cam.min_y < v && v < (double)a.cam.max_y)) return p;
    const int oct = a.last_octave[il];
    if (forward) {
        p.min_level = oct;
        p.max_level = (int)a.cam.num_scale_levels - 1;
    }
    else if (backward) {
        p.min_level = 0;
        p.max_level = oct;
    }
    else {
        p.min_level = oct - 1;
        p.max_level = oct + 1;
    }
    p.ref_x = (float)u;
    p.ref_y = (float)v;
    p.m = __fmul_rn(a.margin, a.scale[oct]);
    return cell_window(a, p);
}

// data/common.cc:135-153: the cell range of the query window; ok = false when it misses the grid
__device__ Projected cell_window(const ProjArgs& a, Projected p) {
    p.ok = false;
    const int gc = (int)a.cam.grid_cols, gr = (int)a.cam.grid_rows;
    p.min_cx = max(0, cv_floor(__dmul_rn((double)__fsub_rn(__fsub_rn(p.ref_x, a.cam.min_x), p.m), a.cam.inv_cell_width)));
    if (gc <= p.min_cx) return p;
    p.max_cx = min(gc - 1, cv_ceil(__dmul_rn((double)__fadd_rn(__fsub_rn(p.ref_x, a.cam.min_x), p.m), a.cam.inv_cell_width)));
    if (p.max_cx < 0) return p;
    p.min_cy = max(0, cv_floor(__dmul_rn((double)__fsub_rn(__fsub_rn(p.ref_y, a.cam.min_y), p.m), a.cam.inv_cell_height)));
    if (gr <= p.min_cy) return p;
    p.max_cy = min(gr - 1, cv_ceil(__dmul_rn((double)__fadd_rn(__fsub_rn(p.ref_y, a.cam.min_y), p.m), a.cam.inv_cell_height)));
    if (p.max_cy < 0) return p;
    p.ok = true;
    return p;
}

// Visits the candidates of row il that pass the gates: f(ord, idx, dist, octave). `ord` grows with the reference's visit order
// (get_keypoints_in_cell: cells x-outer / y-inner, insertion order inside a cell; candidate lists front to back), so the smallest
// (dist, ord) is the candidate the reference's strict '<' update keeps. part / parts splits the walk between lanes (cell t or list
// entry t goes to lane t % parts); part = 0, parts = 1 is the reference's own sequential order.
constexpr int ORD_BITS = 44;  // key = dist << ORD_BITS | ord

// Which keypoints are no longer available to a row. versioned: p[c] = EMPTY (free), 0 (taken before the call) or row + 1 of the
// row that took it - for row r a keypoint taken by a LATER row (the parallel rounds commit out of order) still counts as free,
// as it was when the sequential loop reached r. Otherwise p is a plain bitmap (tables too large for shared memory: serial only).
struct TakenSet {
    uint32_t* p;
    bool versioned;
    __device__ __forceinline__ bool taken(uint32_t c, uint32_t row) const {
        return versioned ? p[c] <= row : ((p[c >> 5] >> (c & 31)) & 1u) != 0;
    }
    __device__ __forceinline__ void take(uint32_t c, uint32_t row) const {
        if (versioned) p[c] = row + 1;  // one writer per keypoint
        else atomicOr(&p[c >> 5], 1u << (c & 31));
    }
};
template <typename F>
__device__ void visit(const ProjArgs& a, uint32_t il, const Projected& p, const TakenSet* taken, int part, int parts, F f) {
    const uint4* dl = reinterpret_cast<const uint4*>(a.last_desc + (size_t)il * 32);
    const uint4 l0 = __ldg(dl), l1 = __ldg(dl + 1);
    if (a.mode == 3) {
        const uint32_t q0 = a.cand_begin[il], q1 = a.cand_end[il];
        for (uint32_t q = q0 + part; q < q1; q += parts) {
            const uint32_t i = a.cand_items[q];
            if (taken && taken->taken(i, il)) continue;
            const uint4* dc = reinterpret_cast<const uint4*>(a.desc + (size_t)i * 32);
            const unsigned d = hamming256(l0, l1, __ldg(dc), __ldg(dc + 1));
            if (a.tri) {
                if (a.hamm_thr < d) continue;  // robust.cc:108
                const double* b2 = a.kp_bearing + 3 * (size_t)i;
                if (!a.row_stereo[il] && !a.kp_stereo[i]) {  // robust.cc:112-123: nothing within 3 degrees of the epipole
                    const double cos_dist = __dadd_rn(__dadd_rn(__dmul_rn(a.epi[0], b2[0]), __dmul_rn(a.epi[1], b2[1])), __dmul_rn(a.epi[2], b2[2]));
                    if (0.99862953475 < cos_dist) continue;
                }
                if (!epipolar_inlier(a.E, a.row_bearing + 3 * (size_t)il, b2, a.row_scale[il])) continue;
            }
            const uint64_t ord = q - q0;
            f(a.last_wins ? ((1ull << ORD_BITS) - 1 - ord) : ord, i, d, 0);  // last_wins: an equal distance later in the list wins
        }
        return;
    }
    const bool check_level = (0 < p.min_level) || (0 <= p.max_level);
    const int ncy = p.max_cy - p.min_cy + 1, ncell = (p.max_cx - p.min_cx + 1) * ncy;
    for (int t = part; t < ncell; t += parts) {
        const int cx = p.min_cx + t / ncy, cy = p.min_cy + t % ncy;
        const uint32_t c = cx * a.cam.grid_rows + cy;
        const uint32_t q0 = a.cell_start[c], q1 = a.cell_start[c + 1];
        for (uint32_t q = q0; q < q1; ++q) {
            const uint32_t i = a.cell_items[q];
            const lpslam_keypoint k = a.kp[i];
            if (check_level) {
                if (k.octave < p.min_level) continue;
                if (0 <= p.max_level && p.max_level < k.octave) continue;
            }
            if (!(fabsf(__fsub_rn(k.x, p.ref_x)) < p.m && fabsf(__fsub_rn(k.y, p.ref_y)) < p.m)) continue;
            if (taken && taken->taken(i, il)) continue;
            if (a.mode == 0 || (a.mode == 1 && a.use_xr)) {
                const float xr = a.x_right[i];
                if (xr > 0) {
                    if (p.m < fabsf(__fsub_rn(p.x_right, xr))) continue;
                }
            }
            if (a.chi_gate) {
                // double (reprojection) - float (keypoint) differences, float x_right difference, summed in the
                // reference's order; 3 degrees of freedom when the keypoint has a stereo match, 2 otherwise
                const double ex = __dsub_rn(a.lm_reproj[2 * (size_t)il], (double)k.x);
                const double ey = __dsub_rn(a.lm_reproj[2 * (size_t)il + 1], (double)k.y);
                double sq = __dadd_rn(__dmul_rn(ex, ex), __dmul_rn(ey, ey));
                const float kxr = a.x_right[i];
                double chi = (double)5.99146f;
                if (kxr >= 0) {
                    const float exr = __fsub_rn(a.lm_x_right[il], kxr);
                    sq = __dadd_rn(sq, (double)__fmul_rn(exr, exr));
                    chi = (double)7.81473f;
                }
                if (chi < __dmul_rn(sq, (double)a.inv_sigma_sq[k.octave])) continue;
            }
            const uint4* dc = reinterpret_cast<const uint4*>(a.desc + (size_t)i * 32);
            // cells hold fewer than 2^20 keypoints (n_curr < 2^20) and a window fewer than 2^24 cells
            f(((uint64_t)t << 20) | (q - q0), i, hamming256(l0, l1, __ldg(dc), __ldg(dc + 1)), k.octave);
        }
    }
}

__device__ __forceinline__ void motion_flags(const ProjArgs& a, bool& forward, bool& backward) {
    // projection.cc:90-109: trans_wc = -R_cw^T t_cw, trans_lc = R_lw trans_wc + t_lw
    double twc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
        twc[i] = -__dadd_rn(__dadd_rn(__dmul_rn(a.Tc[i], a.Tc[3]), __dmul_rn(a.Tc[4 + i], a.Tc[7])), __dmul_rn(a.Tc[8 + i], a.Tc[11]));
    const double z = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a.Tl[8], twc[0]), __dmul_rn(a.Tl[9], twc[1])), __dmul_rn(a.Tl[10], twc[2])), a.Tl[11]);
    forward = a.cam.is_monocular ? false : z > a.cam.true_baseline;
    backward = a.cam.is_monocular ? false : -z > a.cam.true_baseline;
}

constexpr uint64_t EMPTY64 = ~0ull;

// One warp per row: the lanes split the row's cells (or candidate list), every lane keeps its 4 smallest (dist, ord) keys, four
// rounds of "smallest head wins and pops" merge them. The stored key is dist << 20 | rank: the resolve kernels only compare the
// kept four with each other and look at the distance.
__global__ void __launch_bounds__(128) k_proj_topk(ProjArgs a) {
    const int lane = threadIdx.x & 31;
    const uint32_t il = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (il >= a.n_last) return;
    bool fw, bw;
    motion_flags(a, fw, bw);
    const Projected p = project(a, il, fw, bw);
    uint64_t key[K] = {EMPTY64, EMPTY64, EMPTY64, EMPTY64};
    int idx[K] = {-1, -1, -1, -1};
    if (p.ok) {
        visit(a, il, p, nullptr, lane, 32, [&](uint64_t ord, uint32_t i, unsigned d, int oct) {
            const uint64_t kk = ((uint64_t)d << ORD_BITS) | ord;
            if (kk < key[3]) {
                key[3] = kk;
                idx[3] = (int)(i | ((uint32_t)(oct & 0x7F) << 24));  // octave rides along: the resolve loop never touches a.kp
#pragma unroll
                for (int q = 3; q > 0; --q)
                    if (key[q] < key[q - 1]) {
                        const uint64_t tk = key[q]; key[q] = key[q - 1]; key[q - 1] = tk;
                        const int ti = idx[q]; idx[q] = idx[q - 1]; idx[q - 1] = ti;
                    }
            }
        });
    }
    uint32_t okey[K];
    int oidx[K];
#pragma unroll
    for (int r = 0; r < K; ++r) {
        uint64_t m = key[0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint64_t other = __shfl_xor_sync(0xFFFFFFFFu, m, o);
            m = other < m ? other : m;
        }
        const bool win = m != EMPTY64 && key[0] == m;  // keys are unique: ord is
        const uint32_t src = __ffs(__ballot_sync(0xFFFFFFFFu, win)) - 1;
        oidx[r] = m == EMPTY64 ? -1 : __shfl_sync(0xFFFFFFFFu, idx[0], src & 31);
        okey[r] = m == EMPTY64 ? EMPTY : (((uint32_t)(m >> ORD_BITS) << 20) | (uint32_t)r);
        if (win) {
            key[0] = key[1]; key[1] = key[2]; key[2] = key[3]; key[3] = EMPTY64;
            idx[0] = idx[1]; idx[1] = idx[2]; idx[2] = idx[3]; idx[3] = -1;
        }
    }
    if (lane == 0) {
        a.top_key[il] = make_uint4(okey[0], okey[1], okey[2], okey[3]);
        a.top_idx[il] = make_int4(oidx[0], oidx[1], oidx[2], oidx[3]);
    }
}

// rows that do not compete for keypoints (match/fuse.cc): the first kept candidate is the answer
__global__ void k_rows_best(ProjArgs a) {
    const uint32_t il = blockIdx.x * blockDim.x + threadIdx.x;
    if (il >= a.n_last) return;
    const uint32_t key = a.top_key[il].x;
    const unsigned d = key == EMPTY ? MAX_HAMMING_DIST : (key >> 20);
    a.row_dist[il] = d;
    a.row_best[il] = (key != EMPTY && d <= a.hamm_thr) ? (a.top_idx[il].x & 0x00FFFFFF) : -1;
}

// Ordered resolve for the reprojection / candidate-list matchers (projection.cc:31-79, :145-190, :264-296,
// bow_tree.cc:58-97). The reference loop is sequential because a matched keypoint is no longer available to later rows:
// row r takes the first free entry of its candidate list (in (distance, visit order) = the strict '<' updates of the
// reference) if the threshold / ratio tests pass with the second free entry.
//
// Parallel rounds (one block of 1024 threads, rows dealt round-robin). A row can only ever TAKE a kept candidate whose
// distance passes the threshold, so every unresolved row claims those (atomicMin of the row number on the keypoint). For row r a
// free candidate is "own" when no earlier row claims it (it stays free until r's turn), "contested" otherwise (free or taken by
// then). r is decided as soon as that cannot change the outcome: its first free candidate is own, and the second free one is
// own too, or contested but the ratio test already passes with its distance (whatever replaces it is farther). All decided rows
// of a round are committed together; rows that share a takeable candidate are thereby committed in row order. A row for which
// some outcome of its contested candidates leaves the kept four insufficient (the list was longer) could reach for ANY keypoint
// of its window: it blocks every later row until it is the first unresolved row, then it is rescanned exactly. Inputs where
// everything wants the same keypoints degenerate to a row per round; after a few such rounds the rest goes to the serial loop.
//
// resolved rows: top_key[r].x = EMPTY, top_key[r].y = matched keypoint or EMPTY. The ordered event list is built at the end.
struct RowDecision {
    bool settled, accept;
    int i1;
};

__device__ __forceinline__ RowDecision decide_row(const ProjArgs& a, uint32_t il, const uint32_t (&k)[K], const int (&raw)[K], const TakenSet& tk) {
    uint32_t b1 = EMPTY, b2 = EMPTY;
    int i1 = -1, i2 = -1, o1 = -1, o2 = -1;
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (k[q] == EMPTY) continue;
        const int ixq = raw[q] & 0x00FFFFFF;
        if (tk.taken(ixq, il)) continue;
        if (i1 < 0) {
            b1 = k[q];
            i1 = ixq;
            o1 = raw[q] >> 24;
        }
        else if (i2 < 0) {
            b2 = k[q];
            i2 = ixq;
            o2 = raw[q] >> 24;
        }
    }
    RowDecision r;
    r.settled = (a.use_ratio ? i2 >= 0 : i1 >= 0) || k[K - 1] == EMPTY;
    if (!r.settled) {
        // everything behind the kept four is at least as far as the fourth: that bound often settles the row
        const unsigned lb = k[K - 1] >> 20;
        if (i1 >= 0) {  // the second is missing
            const unsigned bd = b1 >> 20;
            if (a.hamm_thr < bd) r.settled = true;                                       // rejected whatever the second is
            else if (!((float)bd > __fmul_rn(a.lowe_ratio, (float)lb))) r.settled = true;  // passes the ratio test for any second >= lb
        }
        else if (a.hamm_thr < lb) r.settled = true;  // no free candidate can be close enough
    }
    r.i1 = i1;
    r.accept = i1 >= 0;
    if (r.accept) {  // projection.cc:68-79 / :184-190 / :287-296, bow_tree.cc:80-97
        const unsigned best = b1 >> 20, second = i2 >= 0 ? (b2 >> 20) : MAX_HAMMING_DIST;
        if (a.hamm_thr < best) r.accept = false;
        else if (a.use_ratio) {
            const bool same_level = a.mode == 3 || o1 == (i2 >= 0 ? o2 : -1);
            if (same_level && (float)best > __fmul_rn(a.lowe_ratio, (float)second)) r.accept = false;
        }
    }
    return r;
}

// the exact decision of row il from its full candidate list in reference order (one thread)
__device__ __noinline__ RowDecision rescan_row(const ProjArgs& a, uint32_t il, bool fw, bool bw, const TakenSet& tk) {
    uint64_t b1 = EMPTY64, b2 = EMPTY64;
    int i1 = -1, i2 = -1, o1 = -1, o2 = -1;
    const Projected p = project(a, il, fw, bw);
    visit(a, il, p, &tk, 0, 1, [&](uint64_t ord, uint32_t i, unsigned d, int oct) {
        const uint64_t kk = ((uint64_t)d << ORD_BITS) | ord;
        if (kk < b1) {
            b2 = b1;
            i2 = i1;
            o2 = o1;
            b1 = kk;
            i1 = (int)i;
            o1 = oct;
        }
        else if (kk < b2) {
            b2 = kk;
            i2 = (int)i;
            o2 = oct;
        }
    });
    RowDecision r;
    r.settled = true;
    r.i1 = i1;
    r.accept = i1 >= 0;
    if (r.accept) {
        const unsigned best = (unsigned)(b1 >> ORD_BITS), second = i2 >= 0 ? (unsigned)(b2 >> ORD_BITS) : MAX_HAMMING_DIST;
        if (a.hamm_thr < best) r.accept = false;
        else if (a.use_ratio) {
            const bool same_level = a.mode == 3 || o1 == (i2 >= 0 ? o2 : -1);
            if (same_level && (float)best > __fmul_rn(a.lowe_ratio, (float)second)) r.accept = false;
        }
    }
    return r;
}

// records the decision of row il; set_taken = false defers the bitmap update (parallel rounds apply it after a barrier)
__device__ __forceinline__ void commit_row(const ProjArgs& a, uint32_t il, const RowDecision& d, const TakenSet& tk, bool set_taken) {
    uint32_t res = EMPTY;
    if (d.accept) {
        res = (uint32_t)d.i1;
        a.curr_match_last[d.i1] = (int32_t)il;
        const bool blocks = a.last_obs ? a.last_obs[il] != 0 : true;  // has_observation() of the row's landmark
        if (blocks && set_taken) tk.take(d.i1, il);
    }
    uint32_t* key = reinterpret_cast<uint32_t*>(a.top_key) + 4 * (size_t)il;
    key[1] = res;
    key[0] = EMPTY;
}

// Row r in a parallel round. clear: the decision d is what the sequential loop will reach. may_rescan: some outcome of the
// contested candidates leaves the kept four unable to settle the row.
struct RowEval {
    bool clear, may_rescan;
    RowDecision d;
};

__device__ __forceinline__ RowEval eval_row(const ProjArgs& a, uint32_t r, const uint32_t (&k)[K], const int (&raw)[K],
                                            const TakenSet& tk, const uint32_t* s_claim) {
    RowEval e;
    e.clear = false;
    e.may_rescan = false;
    e.d.settled = true;
    e.d.accept = false;
    e.d.i1 = -1;
    const bool need2 = a.use_ratio != 0, is_long = k[K - 1] != EMPTY;
    const unsigned lb = k[K - 1] >> 20;
    auto single_ok = [&](unsigned bd) { return a.hamm_thr < bd || !((float)bd > __fmul_rn(a.lowe_ratio, (float)lb)); };
    int qb = -1, qs = -1, n_own = 0;
    bool b_own = false, s_own = false, contested_single_bad = false;
    unsigned own_bd = 0;
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (k[q] == EMPTY) continue;
        const int c = raw[q] & 0x00FFFFFF;
        if (tk.taken(c, r)) continue;
        const bool own = s_claim[c] >= r;
        if (own) {
            if (n_own == 0) own_bd = k[q] >> 20;
            ++n_own;
        }
        else if (!single_ok(k[q] >> 20)) contested_single_bad = true;
        if (qb < 0) {
            qb = q;
            b_own = own;
        }
        else if (qs < 0) {
            qs = q;
            s_own = own;
        }
    }
    if (is_long) {  // the outcomes with fewer free candidates than the decision needs
        if (n_own == 0 && !(a.hamm_thr < lb)) e.may_rescan = true;  // nothing left free
        if (need2 && n_own == 0 && contested_single_bad) e.may_rescan = true;
        if (need2 && n_own == 1 && !single_ok(own_bd)) e.may_rescan = true;
    }
    if (qb < 0) {  // nothing free now: final when the list was complete or nothing behind it can be close enough
        e.clear = !is_long || a.hamm_thr < lb;
        return e;
    }
    if (!b_own) return e;
    const unsigned bd = k[qb] >> 20;
    e.d.i1 = raw[qb] & 0x00FFFFFF;
    if (a.hamm_thr < bd) {  // rejected whatever the second is
        e.clear = true;
        return e;
    }
    if (!need2) {
        e.clear = true;
        e.d.accept = true;
        return e;
    }
    if (qs >= 0) {
        const unsigned sd = k[qs] >> 20;
        if (s_own) {  // the second is exactly this one
            const bool same_level = a.mode == 3 || (raw[qb] >> 24) == (raw[qs] >> 24);
            e.clear = true;
            e.d.accept = !(same_level && (float)bd > __fmul_rn(a.lowe_ratio, (float)sd));
        }
        else if (!((float)bd > __fmul_rn(a.lowe_ratio, (float)sd))) {  // passes for this second and for anything farther
            e.clear = true;
            e.d.accept = true;
        }
        return e;
    }
    if (!is_long) {  // no second at all: MAX_HAMMING_DIST, and a level of -1 that only candidate lists (mode 3) compare equal
        e.clear = true;
        e.d.accept = !(a.mode == 3 && (float)bd > __fmul_rn(a.lowe_ratio, (float)MAX_HAMMING_DIST));
        return e;
    }
    if (!((float)bd > __fmul_rn(a.lowe_ratio, (float)lb))) {  // passes for any second behind the kept four
        e.clear = true;
        e.d.accept = true;
    }
    return e;  // otherwise: needs the exact list (may_rescan is set)
}

// Serial form, one warp: 32 unresolved rows are decided at a time against the current bitmap and committed in row order up to
// the first lane whose candidates contain a keypoint that an earlier lane of the batch just took; that lane and the ones
// behind it are decided again. A row whose four kept candidates do not settle the decision is rescanned by its lane when it
// is the first pending row. s_claim (optional): lowest lane accepting a keypoint in the current round.
__device__ void resolve_serial(const ProjArgs& a, const TakenSet& tk, uint32_t* s_claim, int lane, bool fw, bool bw, uint32_t first_row) {
    for (uint32_t base = first_row; base < a.n_last; base += 32) {
        const uint32_t il = base + lane;
        const uint4 mk = il < a.n_last ? a.top_key[il] : make_uint4(EMPTY, EMPTY, EMPTY, EMPTY);
        const int4 mi = il < a.n_last ? a.top_idx[il] : make_int4(-1, -1, -1, -1);
        const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
        const int raw[K] = {mi.x, mi.y, mi.z, mi.w};
        uint32_t pending = __ballot_sync(0xFFFFFFFFu, mk.x != EMPTY);
        while (pending) {
            const bool mine = (pending >> lane) & 1u;
            RowDecision d = {true, false, -1};
            if (mine) d = decide_row(a, il, k, raw, tk);
            const bool rescan = mine && !d.settled;
            const int first = __ffs(pending) - 1;
            if (__shfl_sync(0xFFFFFFFFu, (int)rescan, first) && lane == first) d = rescan_row(a, il, fw, bw, tk);
            const bool accept = mine && d.accept && (lane == first || !rescan);
            // a lane waits when it is undecided, or when an earlier pending lane accepts one of its four candidates. Entries
            // behind the ones a lane used cannot change its first / second free entry, so the kept four are all that matters
            // for a lane that was decided from them; an undecided lane is only ever resolved as `first`.
            bool wait = mine && rescan && lane != first;
            if (s_claim) {
                if (accept) atomicMin(&s_claim[d.i1], (uint32_t)lane);
                __syncwarp();
                if (mine) {
#pragma unroll
                    for (int q = 0; q < K; ++q)
                        if (k[q] != EMPTY && s_claim[raw[q] & 0x00FFFFFF] < (uint32_t)lane) wait = true;
                }
                __syncwarp();
                if (accept) s_claim[d.i1] = EMPTY;
            }
            else {
                for (uint32_t mm = pending; mm; mm &= mm - 1) {
                    const int M = __ffs(mm) - 1;
                    const int accM = __shfl_sync(0xFFFFFFFFu, (int)accept, M);
                    const int iM = __shfl_sync(0xFFFFFFFFu, d.i1, M);
                    if (mine && lane > M && accM) {
#pragma unroll
                        for (int q = 0; q < K; ++q)
                            if (k[q] != EMPTY && (raw[q] & 0x00FFFFFF) == iM) wait = true;
                    }
                }
            }
            const uint32_t waiting = __ballot_sync(0xFFFFFFFFu, wait) & pending;
            const uint32_t commit = waiting ? (pending & ((1u << (__ffs(waiting) - 1)) - 1u)) : pending;  // rows before the first waiter
            if ((commit >> lane) & 1u) {
                d.accept = accept;
                commit_row(a, il, d, tk, true);
            }
            pending &= ~commit;
            __syncwarp();
        }
    }
}

constexpr int RESOLVE_THREADS = 1024;

__global__ void __launch_bounds__(RESOLVE_THREADS) k_lm_resolve(const __grid_constant__ ProjArgs a) {
    extern __shared__ uint32_t s_mem[];
    __shared__ uint32_t s_barrier, s_minrow, s_warp_tot[RESOLVE_THREADS / 32], s_done;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // versioned owner table + two claim tables (n_curr words each) while they fit, else a bitmap (serial loop only)
    const TakenSet tk = {s_mem, a.claim_in_smem != 0};
    uint32_t* s_claim = tk.versioned ? s_mem + a.n_curr : nullptr;
    if (tk.versioned) {
        for (uint32_t i = tid; i < a.n_curr; i += RESOLVE_THREADS) {
            s_mem[i] = a.curr_taken[i] ? 0u : EMPTY;
            s_claim[i] = s_claim[a.n_curr + i] = EMPTY;
        }
    }
    else {
        const uint32_t words = (a.n_curr + 31) / 32;
        for (uint32_t w = tid; w < words; w += RESOLVE_THREADS) {
            uint32_t bits = 0;
            for (uint32_t b = 0; b < 32 && w * 32 + b < a.n_curr; ++b) bits |= (a.curr_taken[w * 32 + b] ? 1u : 0u) << b;
            s_mem[w] = bits;
        }
    }
    for (uint32_t i = tid; i < a.n_curr; i += RESOLVE_THREADS) a.curr_match_last[i] = -1;
    bool fw = false, bw = false;
    if (a.mode == 0) motion_flags(a, fw, bw);
    // rows that can never match (no candidate, or the closest one beyond the threshold) are resolved here
    uint32_t todo = 0;  // bit j: row tid + j * RESOLVE_THREADS is unresolved (parallel rounds: at most 32 rows per thread)
    {
        uint32_t j = 0;
        for (uint32_t r = tid; r < a.n_last; r += RESOLVE_THREADS, ++j) {
            const uint32_t k0 = a.top_key[r].x;
            if (k0 == EMPTY || (k0 >> 20) > a.hamm_thr) {
                reinterpret_cast<uint32_t*>(a.top_key)[4 * (size_t)r + 1] = EMPTY;
                reinterpret_cast<uint32_t*>(a.top_key)[4 * (size_t)r] = EMPTY;
            }
            else if (j < 32) todo |= 1u << j;
        }
    }
    if (tid == 0) {
        s_barrier = EMPTY;
        s_minrow = EMPTY;
        s_done = 0;
    }
    __syncthreads();

    const bool parallel = s_claim != nullptr && a.n_last <= 32u * RESOLVE_THREADS;
    uint32_t* clean_claim = s_claim;  // a claim table with every entry EMPTY, for the serial loop
    if (parallel) {
        int idle_rounds = 0, rounds = 0;
        for (;;) {
            const int active = __syncthreads_count(todo != 0);  // threads that still own an unresolved row
            if (active == 0) break;
            // A: claims on the candidates a row could take
            uint32_t* const claim = (rounds & 1) ? s_claim : s_claim + a.n_curr;  // the other table is wiped during C
            uint32_t* const other = (rounds & 1) ? s_claim + a.n_curr : s_claim;
            if (todo) atomicMin(&s_minrow, (uint32_t)tid + (uint32_t)(__ffs(todo) - 1) * RESOLVE_THREADS);
            for (uint32_t m = todo; m; m &= m - 1) {
                const uint32_t r = tid + (uint32_t)(__ffs(m) - 1) * RESOLVE_THREADS;
                const uint4 mk = a.top_key[r];
                const int4 mi = a.top_idx[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
                const int raw[K] = {mi.x, mi.y, mi.z, mi.w};
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    if (k[q] == EMPTY || (k[q] >> 20) > a.hamm_thr) continue;
                    const int c = raw[q] & 0x00FFFFFF;
                    if (!tk.taken(c, r)) atomicMin(&claim[c], r);
                }
            }
            __syncthreads();
            // B: every row's verdict, and the first row that may have to look beyond its kept four
            uint32_t clear = 0, accept = 0;
            bool behind = false;
            for (uint32_t m = todo; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const uint32_t r = tid + (uint32_t)j * RESOLVE_THREADS;
                const uint4 mk = a.top_key[r];
                const int4 mi = a.top_idx[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
                const int raw[K] = {mi.x, mi.y, mi.z, mi.w};
                const RowEval e = eval_row(a, r, k, raw, tk, claim);
                if (e.clear) clear |= 1u << j;
                if (e.d.accept) accept |= 1u << j;
                if (e.may_rescan && !behind) {
                    atomicMin(&s_barrier, r);
                    behind = true;
                }
                if (behind) break;  // later rows of this thread are behind the barrier anyway
            }
            __syncthreads();
            // C: record the decided rows up to the barrier (the owner table is updated after the next barrier)
            const uint32_t barrier = s_barrier, minrow = s_minrow;
            uint32_t done = 0, took = 0;
            for (uint32_t m = todo; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const uint32_t r = tid + (uint32_t)j * RESOLVE_THREADS;
                if (r > barrier) break;
                RowDecision d = {true, false, -1};
                if ((clear >> j) & 1u) {
                    if ((accept >> j) & 1u) {  // the first free candidate
                        const int4 mi = a.top_idx[r];
                        const int raw[K] = {mi.x, mi.y, mi.z, mi.w};
#pragma unroll
                        for (int q = K - 1; q >= 0; --q)
                            if (raw[q] >= 0 && !tk.taken(raw[q] & 0x00FFFFFF, r)) d.i1 = raw[q] & 0x00FFFFFF;
                        d.accept = true;
                    }
                }
                else if (r == minrow) d = rescan_row(a, r, fw, bw, tk);  // every earlier row is resolved: the exact list decides
                else continue;
                commit_row(a, r, d, tk, false);
                if (d.accept) took |= 1u << j;
                done |= 1u << j;
            }
            if (done) atomicAdd(&s_done, (uint32_t)__popc(done));
            for (uint32_t i = tid; i < a.n_curr; i += RESOLVE_THREADS) other[i] = EMPTY;
            __syncthreads();
            // D: owner-table updates of the committed rows, re-arm
            for (uint32_t m = took; m; m &= m - 1) {
                const uint32_t r = tid + (uint32_t)(__ffs(m) - 1) * RESOLVE_THREADS;
                if (a.last_obs ? a.last_obs[r] != 0 : true) tk.take(reinterpret_cast<const uint32_t*>(a.top_key)[4 * (size_t)r + 1], r);
            }
            todo &= ~done;
            const uint32_t n_done = s_done;
            __syncthreads();
            if (tid == 0) {
                s_barrier = EMPTY;
                s_minrow = EMPTY;
                s_done = 0;
            }
            ++rounds;
            clean_claim = (rounds & 1) ? s_claim : s_claim + a.n_curr;
            // contended (hundreds of rows left, a handful per round): the serial loop commits up to 32 rows per step without block barriers
            idle_rounds = (n_done < 8 && active >= 64) ? idle_rounds + 1 : 0;
            if (idle_rounds >= 4) break;
        }
    }
    __syncthreads();
    if (tid == 0) s_minrow = EMPTY;
    __syncthreads();
    for (uint32_t r = tid; r < a.n_last; r += RESOLVE_THREADS)
        if (a.top_key[r].x != EMPTY) {
            atomicMin(&s_minrow, r);
            break;
        }
    __syncthreads();
    // whatever is left (everything when the claim table did not fit), from the first unresolved row on
    if (warp == 0 && s_minrow != EMPTY) resolve_serial(a, tk, clean_claim, lane, fw, bw, s_minrow & ~31u);
    __syncthreads();

    // ordered event list and count: rows in order, 1024 at a time
    uint32_t num = 0;
    for (uint32_t base = 0; base < a.n_last; base += RESOLVE_THREADS) {
        const uint32_t r = base + tid;
        const int res = r < a.n_last ? (int)reinterpret_cast<const uint32_t*>(a.top_key)[4 * (size_t)r + 1] : -1;
        const uint32_t bal = __ballot_sync(0xFFFFFFFFu, res >= 0);
        if (lane == 0) s_warp_tot[warp] = __popc(bal);
        __syncthreads();
        // every warp scans the 32 warp totals itself
        const uint32_t wt = s_warp_tot[lane];
        uint32_t inc = wt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += t;
        }
        const uint32_t before = __shfl_sync(0xFFFFFFFFu, inc - wt, warp), total = __shfl_sync(0xFFFFFFFFu, inc, 31);
        if (res >= 0 && a.events) a.events[num + before + __popc(bal & ((1u << lane) - 1u))] = make_int2((int)r, res);
        num += total;
        __syncthreads();
    }
    if (tid == 0) *a.n_events = num;
}

constexpr uint32_t kResolveTableMax = 16000;  // 3 x 4 B per keypoint of shared memory

cudaError_t launch_lm_resolve(const ProjArgs& a, uint32_t n_kp, cudaStream_t st) {
    const size_t smem = a.claim_in_smem ? 12 * (size_t)n_kp : ((n_kp + 31) / 32) * 4;
    if (smem > 48 * 1024) {
        const cudaError_t e = cudaFuncSetAttribute(k_lm_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_lm_resolve<<<1, RESOLVE_THREADS, smem, st>>>(a);
    return cudaGetLastError();
}

// ---- frame post-processing that follows extraction (data/frame.cc:37-51,75-92), perspective model --------------------
// camera/perspective.cc:98-141. cv::undistortPoints(K, (k1 k2 p1 p2 k3), R = I, P = K, TermCriteria(EPS | MAX_ITER, 20,
// 1e-6)): iterative inversion of the Brown-Conrady model in double, evaluated in OpenCV's operation order with explicit
// round-to-nearest multiplies and adds (no FMA contraction), result rounded to float. The zero rational / thin-prism
// coefficients of the 5-parameter model contribute exact zeros and are dropped.
__device__ __forceinline__ double dm(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double da(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double ds(double a, double b) { return __dsub_rn(a, b); }

// `mu` = the model as cv::undistortPoints sees it: camera/perspective.cc:21-22 stores the camera matrix and the distortion
// vector as cv::Mat_<float>, so every intrinsic goes through float before the double arithmetic; `m` = the double members the
// bearing conversion (:132-141) uses.
__global__ void k_undistort_bearings(lpslam_perspective mu, lpslam_perspective m, uint32_t n, const lpslam_keypoint* __restrict__ in,
                                     lpslam_keypoint* __restrict__ out, double* __restrict__ bearings) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    lpslam_keypoint k = in[i];
    const double k0 = mu.dist[0], k1 = mu.dist[1], p1 = mu.dist[2], p2 = mu.dist[3], k4 = mu.dist[4];
    const double ifx = __ddiv_rn(1.0, mu.fx), ify = __ddiv_rn(1.0, mu.fy);
    const double u = (double)k.x, v = (double)k.y;
    double x = dm(ds(u, mu.cx), ifx), y = dm(ds(v, mu.cy), ify);
    const double x0 = x, y0 = y;
    double error = 1.7976931348623157e308;
    for (int j = 0; j < 20; ++j) {
        if (error < 1e-6) break;
        double r2 = da(dm(x, x), dm(y, y));
        const double icdist = __ddiv_rn(1.0, da(1.0, dm(da(dm(da(dm(k4, r2), k1), r2), k0), r2)));
        if (icdist < 0) {  // calib3d regression 14583: fall back to the un-iterated point
            x = dm(ds(u, mu.cx), ifx);
            y = dm(ds(v, mu.cy), ify);
            break;
        }
        const double dX = da(dm(dm(dm(2.0, p1), x), y), dm(p2, da(r2, dm(dm(2.0, x), x))));
        const double dY = da(dm(p1, da(r2, dm(dm(2.0, y), y))), dm(dm(dm(2.0, p2), x), y));
        x = dm(ds(x0, dX), icdist);
        y = dm(ds(y0, dY), icdist);
        r2 = da(dm(x, x), dm(y, y));
        const double r4 = dm(r2, r2), r6 = dm(r4, r2);
        const double a1 = dm(dm(2.0, x), y), a2 = da(r2, dm(dm(2.0, x), x)), a3 = da(r2, dm(dm(2.0, y), y));
        const double cdist = da(da(da(1.0, dm(k0, r2)), dm(k1, r4)), dm(k4, r6));
        const double xd0 = da(da(dm(dm(x, cdist), 1.0), dm(p1, a1)), dm(p2, a2));
        const double yd0 = da(da(dm(dm(y, cdist), 1.0), dm(p1, a3)), dm(p2, a1));
        const double ex = ds(da(dm(xd0, mu.fx), mu.cx), u), ey = ds(da(dm(yd0, mu.fy), mu.cy), v);
        error = __dsqrt_rn(da(dm(ex, ex), dm(ey, ey)));
    }
    k.x = (float)da(dm(mu.fx, x), mu.cx);
    k.y = (float)da(dm(mu.fy, y), mu.cy);
    k.response = 0.f;   // cv::KeyPoint defaults: undistort_keypoints copies pt, angle, size, octave only
    k.class_id = -1;
    out[i] = k;
    // camera/perspective.cc:132-141
    const double xn = __ddiv_rn(ds((double)k.x, m.cx), m.fx), yn = __ddiv_rn(ds((double)k.y, m.cy), m.fy);
    const double l2 = __dsqrt_rn(da(da(dm(xn, xn), dm(yn, yn)), 1.0));
    bearings[3 * (size_t)i] = __ddiv_rn(xn, l2);
    bearings[3 * (size_t)i + 1] = __ddiv_rn(yn, l2);
    bearings[3 * (size_t)i + 2] = __ddiv_rn(1.0, l2);
}

// area::match_in_consistent_area (area.cc:20-91): rows are replayed in order against the per-keypoint best distance of
// frame 2 ("ignore if the already-matched point is closer in Hamming space", area.cc:47-50); a later row may take a
// keypoint over, which deletes the earlier match. Same batch form as k_lm_resolve: 32 rows are decided per round against
// the current per-keypoint state, rows are committed in order up to the first one whose kept candidates contain a keypoint
// that an earlier row of the round just matched (its distance gate moved), the rest are decided again.
// s_state[i] = matched_dists_in_frm_2[i] << 20 | (matched_indices_1_in_frm_2[i] + 1); s_claim as in k_lm_resolve.
__global__ void __launch_bounds__(32) k_area_resolve(ProjArgs a) {
    extern __shared__ uint32_t s_state[];
    uint32_t* s_claim = s_state + a.n_curr;
    const int lane = threadIdx.x;
    for (uint32_t i = lane; i < a.n_curr; i += 32) {
        s_state[i] = (uint32_t)MAX_HAMMING_DIST << 20;
        s_claim[i] = 0xFFFFFFFFu;
    }
    for (uint32_t i = lane; i < a.n_last; i += 32) a.match_2_of_1[i] = -1;
    __syncwarp();
    int num = 0;
    uint32_t n_ev = 0;
    constexpr unsigned HAMMING_DIST_THR_LOW = 50;  // match/base.h:13
    for (uint32_t base = 0; base < a.n_last; base += 32) {
        const uint32_t il = base + lane;
        const uint4 mk = il < a.n_last ? a.top_key[il] : make_uint4(EMPTY, EMPTY, EMPTY, EMPTY);
        const int4 mi = il < a.n_last ? a.top_idx[il] : make_int4(-1, -1, -1, -1);
        const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
        const int ix[K] = {mi.x & 0x00FFFFFF, mi.y & 0x00FFFFFF, mi.z & 0x00FFFFFF, mi.w & 0x00FFFFFF};
        uint32_t pending = __ballot_sync(0xFFFFFFFFu, mk.x != EMPTY && (mk.x >> 20) <= HAMMING_DIST_THR_LOW);
        while (pending) {
            const bool mine = (pending >> lane) & 1u;
            uint32_t b1 = EMPTY, b2 = EMPTY;
            int i1 = -1;
            bool have2 = false;
            if (mine) {
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    if (k[q] == EMPTY) continue;
                    if ((s_state[ix[q]] >> 20) <= (k[q] >> 20)) continue;  // area.cc:48
                    if (i1 < 0) {
                        b1 = k[q];
                        i1 = ix[q];
                    }
                    else if (!have2) {
                        b2 = k[q];
                        have2 = true;
                    }
                }
            }
            bool settled = have2 || k[K - 1] == EMPTY;
            if (mine && !settled) {
                // everything behind the kept four is at least as far as the fourth: that bound often settles the row
                const unsigned lb = k[K - 1] >> 20;
                if (i1 >= 0) {
                    const unsigned bd = b1 >> 20;
                    if (HAMMING_DIST_THR_LOW < bd) settled = true;
                    else if (!(__fmul_rn((float)lb, a.lowe_ratio) < (float)bd)) settled = true;  // passes for any second >= lb
                }
                else if (HAMMING_DIST_THR_LOW < lb) settled = true;
            }
            const bool rescan = mine && !settled;
            const int first = __ffs(pending) - 1;
            if (__shfl_sync(0xFFFFFFFFu, (int)rescan, first)) {
                // the kept four do not decide best and second: exact rescan in reference order by the row's lane
                if (lane == first) {
                    uint64_t w1 = EMPTY64, w2 = EMPTY64;
                    i1 = -1;
                    const Projected p = project(a, il, false, false);
                    visit(a, il, p, nullptr, 0, 1, [&](uint64_t ord, uint32_t i, unsigned d, int) {
                        if ((s_state[i] >> 20) <= d) return;
                        const uint64_t kk = ((uint64_t)d << ORD_BITS) | ord;
                        if (kk < w1) {
                            w2 = w1;
                            w1 = kk;
                            i1 = (int)i;
                        }
                        else if (kk < w2) {
                            w2 = kk;
                        }
                    });
                    // only the distances are looked at from here on
                    b1 = w1 == EMPTY64 ? EMPTY : (uint32_t)(w1 >> ORD_BITS) << 20;
                    b2 = w2 == EMPTY64 ? EMPTY : (uint32_t)(w2 >> ORD_BITS) << 20;
                }
            }
            bool accept = mine && i1 >= 0 && (lane == first || !rescan);
            const unsigned best = b1 >> 20;
            if (accept) {
                const unsigned second = b2 == EMPTY ? MAX_HAMMING_DIST : (b2 >> 20);
                if (HAMMING_DIST_THR_LOW < best) accept = false;
                else if (__fmul_rn((float)second, a.lowe_ratio) < (float)best) accept = false;  // area.cc:67-69
            }
            bool wait = mine && rescan && lane != first;
            if (mine && accept) atomicMin(&s_claim[i1], (uint32_t)lane);
            __syncwarp();
            if (mine) {
#pragma unroll
                for (int q = 0; q < K; ++q)
                    if (k[q] != EMPTY && s_claim[ix[q]] < (uint32_t)lane) wait = true;
            }
            __syncwarp();
            if (mine && accept) s_claim[i1] = 0xFFFFFFFFu;
            const uint32_t waiting = __ballot_sync(0xFFFFFFFFu, wait) & pending;
            const uint32_t commit = waiting ? (pending & ((1u << (__ffs(waiting) - 1)) - 1u)) : pending;
            const bool go = ((commit >> lane) & 1u) && accept;
            const uint32_t acc_mask = __ballot_sync(0xFFFFFFFFu, go);
            bool took_over = false;
            if (go) {
                // distinct keypoints within a round (a later lane holding the same one waited), earlier owners are earlier rows
                const int prev = (int)(s_state[i1] & 0xFFFFFu) - 1;  // area.cc:76-80: the earlier owner loses its match
                if (0 <= prev) {
                    a.match_2_of_1[prev] = -1;
                    took_over = true;
                }
                a.match_2_of_1[il] = i1;
                s_state[i1] = (best << 20) | (il + 1);
                a.events[n_ev + __popc(acc_mask & ((1u << lane) - 1u))] = make_int2((int)il, i1);
            }
            n_ev += __popc(acc_mask);
            num += __popc(acc_mask) - __popc(__ballot_sync(0xFFFFFFFFu, took_over));
            pending &= ~commit;
            __syncwarp();
        }
    }
    if (lane == 0) {
        a.n_events[0] = n_ev;
        a.n_events[1] = (uint32_t)num;
    }
}

}  // namespace
}  // namespace lps

using namespace lps;

extern "C" int lpslam_match_projection(const lpslam_camera* cam, const double* pose_cw_curr, const double* pose_cw_last,
                                       uint32_t n_last, const uint8_t* last_has_lm, const uint8_t* last_outlier,
                                       const double* last_pos_w, const uint8_t* last_lm_desc, const int32_t* last_octave,
                                       const float* last_angle, const uint8_t* last_lm_has_observation, uint32_t n_curr,
                                       const lpslam_keypoint* curr_undist_kp, const uint8_t* curr_desc, const float* curr_x_right,
                                       const uint8_t* curr_has_observed_lm, const float* scale_factors, float margin,
                                       int check_orientation, int32_t* curr_match_last, uint32_t* n_matches) {
    LPS_ARG(cam && pose_cw_curr && pose_cw_last && n_matches && scale_factors, "null argument");
    *n_matches = 0;
    LPS_ARG(n_curr == 0 || (curr_undist_kp && curr_desc && curr_x_right && curr_has_observed_lm && curr_match_last), "null current-frame argument");
    LPS_ARG(n_last == 0 || (last_has_lm && last_outlier && last_pos_w && last_lm_desc && last_octave), "null last-frame argument");
    LPS_ARG(!check_orientation || last_angle, "last-frame angles required for the orientation check");
    LPS_ARG(cam->grid_cols >= 1 && cam->grid_rows >= 1 && cam->num_scale_levels >= 1 && cam->num_scale_levels <= (uint32_t)kMaxLevels,
            "bad camera grid / level count");
    LPS_ARG(n_curr < (1u << 20), "too many current keypoints");
    for (uint32_t i = 0; i < n_curr; ++i) curr_match_last[i] = -1;
    if (n_curr == 0 || n_last == 0) return LPSLAM_OK;
    for (uint32_t i = 0; i < n_last; ++i)
        LPS_ARG(last_octave[i] >= 0 && (uint32_t)last_octave[i] < cam->num_scale_levels, "last-frame octave out of range");

    const uint32_t ncell = cam->grid_cols * cam->grid_rows;
    ArenaBump ar;  // inputs | zeroed counters | work | outputs
    const size_t o_has = ar.take(n_last), o_out = ar.take(n_last), o_pos = ar.take(24 * (size_t)n_last), o_ld = ar.take(32 * (size_t)n_last);
    const size_t o_oct = ar.take(4 * (size_t)n_last), o_obs = ar.take(n_last), o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)n_curr);
    const size_t o_desc = ar.take(32 * (size_t)n_curr), o_xr = ar.take(4 * (size_t)n_curr), o_tk = ar.take(n_curr);
    const size_t o_sc = ar.take(4 * kMaxLevels), o_cc = ar.take(4 * (size_t)ncell), in_end = ar.off;
    const size_t o_cs = ar.take(4 * (size_t)(ncell + 1));
    const size_t o_cf = ar.take(4 * (size_t)ncell), o_ci = ar.take(4 * (size_t)n_curr), o_tkey = ar.take(16 * (size_t)n_last);
    const size_t o_tidx = ar.take(16 * (size_t)n_last), o_m = ar.take(4 * (size_t)n_curr), o_ne = ar.take(4);
    const size_t o_ev = ar.take(8 * (size_t)n_last);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    int rc = LPSLAM_OK;
    cudaError_t e = cudaSuccess;
    std::vector<int2> events;
    uint32_t n_ev = 0;
    do {
        cs.put(o_has, last_has_lm, n_last);
        cs.put(o_out, last_outlier, n_last);
        cs.put(o_pos, last_pos_w, 24 * (size_t)n_last);
        cs.put(o_ld, last_lm_desc, 32 * (size_t)n_last);
        cs.put(o_oct, last_octave, 4 * (size_t)n_last);
        if (last_lm_has_observation) cs.put(o_obs, last_lm_has_observation, n_last);
        cs.put(o_kp, curr_undist_kp, sizeof(lpslam_keypoint) * (size_t)n_curr);
        cs.put(o_desc, curr_desc, 32 * (size_t)n_curr);
        cs.put(o_xr, curr_x_right, 4 * (size_t)n_curr);
        cs.put(o_tk, curr_has_observed_lm, n_curr);
        cs.put(o_sc, scale_factors, 4 * (size_t)cam->num_scale_levels);
        cs.zero(o_cc, 4 * (size_t)ncell);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.cam = *cam;
        for (int i = 0; i < 16; ++i) {
            a.Tc[i] = pose_cw_curr[i];
            a.Tl[i] = pose_cw_last[i];
        }
        a.n_last = n_last;
        a.n_curr = n_curr;
        a.last_has_lm = dev + o_has;
        a.last_outlier = dev + o_out;
        a.last_pos_w = (const double*)(dev + o_pos);
        a.last_desc = dev + o_ld;
        a.last_octave = (const int32_t*)(dev + o_oct);
        a.last_obs = last_lm_has_observation ? dev + o_obs : nullptr;
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_desc;
        a.x_right = (const float*)(dev + o_xr);
        a.curr_taken = dev + o_tk;
        a.scale = (const float*)(dev + o_sc);
        a.margin = margin;
        a.cell_count = (uint32_t*)(dev + o_cc);
        a.cell_start = (uint32_t*)(dev + o_cs);
        a.cell_fill = (uint32_t*)(dev + o_cf);
        a.cell_items = (uint32_t*)(dev + o_ci);
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.curr_match_last = (int32_t*)(dev + o_m);
        a.events = (int2*)(dev + o_ev);
        a.n_events = (uint32_t*)(dev + o_ne);
        launch_grid(a, n_curr, ncell, cs.st);
        k_proj_topk<<<(n_last + 3) / 4, 128, 0, cs.st>>>(a);
        a.hamm_thr = HAMMING_DIST_THR_HIGH;
        a.use_ratio = 0;
        a.claim_in_smem = n_curr <= kResolveTableMax;
        if ((e = launch_lm_resolve(a, n_curr, cs.st)) != cudaSuccess) break;
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        // matches + event count, and the event list (at most one event per row) only when the orientation check wants it
        if ((e = cs.download(o_m, check_orientation ? ar.off : o_ev)) != cudaSuccess) break;
        memcpy(curr_match_last, cs.out<int32_t>(o_m), 4 * (size_t)n_curr);
        n_ev = *cs.out<uint32_t>(o_ne);
        if (check_orientation) events.assign(cs.out<int2>(o_ev), cs.out<int2>(o_ev) + n_ev);
    } while (false);
    if (e != cudaSuccess) rc = cuda_fail(e, "lpslam_match_projection", __FILE__, __LINE__);
    if (rc) return rc;
    *n_matches = n_ev;
    if (check_orientation) {
        std::vector<std::pair<float, int>> deltas;
        for (const int2& ev : events) deltas.emplace_back(last_angle[ev.x] - curr_undist_kp[ev.y].angle, ev.y);
        for (int id : angle_invalid_matches(deltas)) {
            curr_match_last[id] = -1;
            --*n_matches;
        }
    }
    return LPSLAM_OK;
}

extern "C" int lpslam_match_reprojected(const lpslam_camera* cam, const lpslam_reproj_match_params* prm, uint32_t n_lm,
                                        const uint8_t* lm_active, const double* lm_reproj, const float* lm_x_right,
                                        const int32_t* lm_scale_level, const uint8_t* lm_desc, const uint8_t* lm_has_observation,
                                        const float* lm_angle, uint32_t n_kp, const lpslam_keypoint* kp_undist,
                                        const uint8_t* kp_desc, const float* kp_x_right, const uint8_t* kp_has_observed_lm,
                                        const float* scale_factors, int32_t* kp_match_lm, uint32_t* n_matches) {
    LPS_ARG(cam && prm && n_matches && scale_factors, "null argument");
    const float margin = prm->margin, lowe_ratio = prm->lowe_ratio;
    std::vector<float> zeros_lm, neg_kp;
    if (!prm->use_x_right) {  // the gate is off: the arrays are not looked at
        zeros_lm.assign(n_lm, 0.f);
        neg_kp.assign(n_kp, -1.f);
        lm_x_right = zeros_lm.data();
        kp_x_right = neg_kp.data();
    }
    *n_matches = 0;
    LPS_ARG(n_kp == 0 || (kp_undist && kp_desc && kp_x_right && kp_has_observed_lm && kp_match_lm), "null frame argument");
    LPS_ARG(n_lm == 0 || (lm_active && lm_reproj && lm_x_right && lm_scale_level && lm_desc), "null landmark argument");
    LPS_ARG(prm->hamm_dist_thr <= 256 && prm->level_lo <= prm->level_hi, "bad matcher parameters");
    LPS_ARG(cam->grid_cols >= 1 && cam->grid_rows >= 1 && cam->num_scale_levels >= 1 && cam->num_scale_levels <= (uint32_t)kMaxLevels,
            "bad camera grid / level count");
    LPS_ARG(n_kp < (1u << 20), "too many keypoints");
    for (uint32_t i = 0; i < n_kp; ++i) kp_match_lm[i] = -1;
    if (n_kp == 0 || n_lm == 0) return LPSLAM_OK;
    for (uint32_t i = 0; i < n_lm; ++i)
        LPS_ARG(!lm_active[i] || (lm_scale_level[i] >= 0 && (uint32_t)lm_scale_level[i] < cam->num_scale_levels),
                "landmark scale level out of range");

    const uint32_t ncell = cam->grid_cols * cam->grid_rows;
    ArenaBump ar;  // keypoints + zeroed counters | landmark inputs | work | outputs
    const size_t o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)n_kp), o_cc = ar.take(4 * (size_t)ncell), grid_end = ar.off;
    const size_t o_act = ar.take(n_lm), o_rp = ar.take(16 * (size_t)n_lm), o_lxr = ar.take(4 * (size_t)n_lm), o_lvl = ar.take(4 * (size_t)n_lm);
    const size_t o_ld = ar.take(32 * (size_t)n_lm), o_obs = ar.take(n_lm);
    const size_t o_desc = ar.take(32 * (size_t)n_kp), o_xr = ar.take(4 * (size_t)n_kp), o_tk = ar.take(n_kp);
    const size_t o_sc = ar.take(4 * kMaxLevels), in_end = ar.off;
    const size_t o_cs = ar.take(4 * (size_t)(ncell + 1));
    const size_t o_cf = ar.take(4 * (size_t)ncell), o_ci = ar.take(4 * (size_t)n_kp), o_tkey = ar.take(16 * (size_t)n_lm);
    const size_t o_tidx = ar.take(16 * (size_t)n_lm), o_m = ar.take(4 * (size_t)n_kp), o_ne = ar.take(4), o_ev = ar.take(8 * (size_t)n_lm);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    uint32_t n_ev = 0;
    std::vector<int2> events;
    const bool dbg = getenv("LPS_MATCH_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t0 = now(), t1 = 0, t2 = 0, t3 = 0, t4 = 0, t5 = 0;
    cudaEvent_t evs[4] = {};
    if (dbg) for (auto& ev : evs) cudaEventCreate(&ev);
    do {
        // the grid only needs the keypoints: they go first, and the landmark arrays are packed while that copy and the grid
        // kernel run
        cs.put(o_kp, kp_undist, sizeof(lpslam_keypoint) * (size_t)n_kp);
        cs.zero(o_cc, 4 * (size_t)ncell);
        if ((e = cs.upload(0, grid_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.cam = *cam;
        a.mode = 1;
        a.n_last = n_lm;
        a.n_curr = n_kp;
        a.last_has_lm = dev + o_act;
        a.last_desc = dev + o_ld;
        a.last_octave = (const int32_t*)(dev + o_lvl);
        a.last_obs = lm_has_observation ? dev + o_obs : nullptr;
        a.lm_reproj = (const double*)(dev + o_rp);
        a.lm_x_right = (const float*)(dev + o_lxr);
        a.lowe_ratio = lowe_ratio;
        a.level_lo = prm->level_lo;
        a.level_hi = prm->level_hi;
        a.use_ratio = prm->use_ratio;
        a.use_xr = prm->use_x_right;
        a.hamm_thr = prm->hamm_dist_thr;
        a.events = lm_angle ? (int2*)(dev + o_ev) : nullptr;
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_desc;
        a.x_right = (const float*)(dev + o_xr);
        a.curr_taken = dev + o_tk;
        a.scale = (const float*)(dev + o_sc);
        a.margin = margin;
        a.cell_count = (uint32_t*)(dev + o_cc);
        a.cell_start = (uint32_t*)(dev + o_cs);
        a.cell_fill = (uint32_t*)(dev + o_cf);
        a.cell_items = (uint32_t*)(dev + o_ci);
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.curr_match_last = (int32_t*)(dev + o_m);
        a.n_events = (uint32_t*)(dev + o_ne);
        t1 = now();
        if (dbg) cudaEventRecord(evs[0], cs.st);
        launch_grid(a, n_kp, ncell, cs.st);
        if (dbg) cudaEventRecord(evs[1], cs.st);
        cs.put(o_act, lm_active, n_lm);
        cs.put(o_rp, lm_reproj, 16 * (size_t)n_lm);
        cs.put(o_lxr, lm_x_right, 4 * (size_t)n_lm);
        cs.put(o_lvl, lm_scale_level, 4 * (size_t)n_lm);
        cs.put(o_ld, lm_desc, 32 * (size_t)n_lm);
        if (lm_has_observation) cs.put(o_obs, lm_has_observation, n_lm);
        cs.put(o_desc, kp_desc, 32 * (size_t)n_kp);
        cs.put(o_xr, kp_x_right, 4 * (size_t)n_kp);
        cs.put(o_tk, kp_has_observed_lm, n_kp);
        cs.put(o_sc, scale_factors, 4 * (size_t)cam->num_scale_levels);
        t2 = now();
        if ((e = cs.upload_beside(grid_end, in_end)) != cudaSuccess) break;
        if (dbg) cudaEventRecord(evs[2], cs.st);
        k_proj_topk<<<(n_lm + 3) / 4, 128, 0, cs.st>>>(a);
        a.claim_in_smem = n_kp <= kResolveTableMax;
        if ((e = launch_lm_resolve(a, n_kp, cs.st)) != cudaSuccess) break;
        if (dbg) cudaEventRecord(evs[3], cs.st);
        t3 = now();
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_m, lm_angle ? ar.off : o_ev)) != cudaSuccess) break;
        t4 = now();
        memcpy(kp_match_lm, cs.out<int32_t>(o_m), 4 * (size_t)n_kp);
        n_ev = *cs.out<uint32_t>(o_ne);
        if (lm_angle) events.assign(cs.out<int2>(o_ev), cs.out<int2>(o_ev) + n_ev);
    } while (false);
    t5 = now();
    if (dbg) {
        float g01 = 0, g12 = 0, g23 = 0;
        cudaEventElapsedTime(&g01, evs[0], evs[1]);
        cudaEventElapsedTime(&g12, evs[1], evs[2]);
        cudaEventElapsedTime(&g23, evs[2], evs[3]);
        fprintf(stderr, "reprojected: host pack1+enqueue %.1f, pack2 %.1f, enqueue rest %.1f, wait %.1f, unpack %.1f us | gpu grid %.1f, upload2 %.1f, topk+resolve %.1f us\n", t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, g01 * 1e3, g12 * 1e3, g23 * 1e3);
        for (auto& ev : evs) cudaEventDestroy(ev);
    }
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_reprojected", __FILE__, __LINE__);
    *n_matches = n_ev;
    if (lm_angle) {  // angle_checker over (row angle - keypoint angle, keypoint), projection.cc:297-311
        std::vector<std::pair<float, int>> deltas;
        for (const int2& ev : events) deltas.emplace_back(lm_angle[ev.x] - kp_undist[ev.y].angle, ev.y);
        for (int id : angle_invalid_matches(deltas)) {
            kp_match_lm[id] = -1;
            --*n_matches;
        }
    }
    return LPSLAM_OK;
}

extern "C" int lpslam_match_landmarks(const lpslam_camera* cam, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                                      const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc,
                                      const uint8_t* lm_has_observation, uint32_t n_kp, const lpslam_keypoint* kp_undist,
                                      const uint8_t* kp_desc, const float* kp_x_right, const uint8_t* kp_has_observed_lm,
                                      const float* scale_factors, float margin, float lowe_ratio, int32_t* kp_match_lm,
                                      uint32_t* n_matches) {
    // projection.cc:11-83: window (level - 1, level), threshold HAMMING_DIST_THR_HIGH, ratio test, stereo gate
    const lpslam_reproj_match_params prm = {margin, lowe_ratio, HAMMING_DIST_THR_HIGH, -1, 0, 1, 1};
    return lpslam_match_reprojected(cam, &prm, n_lm, lm_active, lm_reproj, lm_x_right, lm_scale_level, lm_desc, lm_has_observation,
                                    nullptr, n_kp, kp_undist, kp_desc, kp_x_right, kp_has_observed_lm, scale_factors, kp_match_lm,
                                    n_matches);
}

extern "C" int lpslam_frame_postprocess(const lpslam_perspective* model, const lpslam_camera* cam, uint32_t n,
                                        const lpslam_keypoint* keypts, lpslam_keypoint* undist_keypts, double* bearings,
                                        uint32_t* cell_start, uint32_t* cell_items, uint32_t* n_in_grid) {
    LPS_ARG(model && cam && cell_start && n_in_grid, "null argument");
    LPS_ARG(n == 0 || (keypts && undist_keypts && bearings && cell_items), "null keypoint argument");
    LPS_ARG(model->fx != 0.0 && model->fy != 0.0, "focal length must be non-zero");
    LPS_ARG(cam->grid_cols >= 1 && cam->grid_rows >= 1, "bad camera grid");
    const uint32_t ncell = cam->grid_cols * cam->grid_rows;
    *n_in_grid = 0;
    for (uint32_t c = 0; c <= ncell; ++c) cell_start[c] = 0;
    if (n == 0) return LPSLAM_OK;  // perspective.cc:100-103: empty input is passed through
    ArenaBump ar;  // input | zeroed counters | work | outputs
    const size_t o_in = ar.take(sizeof(lpslam_keypoint) * (size_t)n), o_cc = ar.take(4 * (size_t)ncell), in_end = ar.off;
    const size_t o_cf = ar.take(4 * (size_t)ncell);
    const size_t o_out = ar.take(sizeof(lpslam_keypoint) * (size_t)n), o_b = ar.take(24 * (size_t)n);
    const size_t o_cs = ar.take(4 * (size_t)(ncell + 1)), o_ci = ar.take(4 * (size_t)n);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    do {
        cs.put(o_in, keypts, sizeof(lpslam_keypoint) * (size_t)n);
        cs.zero(o_cc, 4 * (size_t)ncell);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        lpslam_perspective mu = *model;  // camera/perspective.cc:21-22: the OpenCV copies of the intrinsics are float matrices
        mu.fx = (double)(float)mu.fx, mu.fy = (double)(float)mu.fy, mu.cx = (double)(float)mu.cx, mu.cy = (double)(float)mu.cy;
        for (int i = 0; i < 5; ++i) mu.dist[i] = (double)(float)mu.dist[i];
        k_undistort_bearings<<<(n + 127) / 128, 128, 0, cs.st>>>(mu, *model, n, (const lpslam_keypoint*)(dev + o_in),
                                                                 (lpslam_keypoint*)(dev + o_out), (double*)(dev + o_b));
        ProjArgs a = {};
        a.cam = *cam;
        a.n_curr = n;
        a.kp = (const lpslam_keypoint*)(dev + o_out);
        a.cell_count = (uint32_t*)(dev + o_cc);
        a.cell_start = (uint32_t*)(dev + o_cs);
        a.cell_fill = (uint32_t*)(dev + o_cf);
        a.cell_items = (uint32_t*)(dev + o_ci);
        launch_grid(a, n, ncell, cs.st);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_out, ar.off)) != cudaSuccess) break;
        memcpy(undist_keypts, cs.out<uint8_t>(o_out), sizeof(lpslam_keypoint) * (size_t)n);
        memcpy(bearings, cs.out<uint8_t>(o_b), 24 * (size_t)n);
        memcpy(cell_start, cs.out<uint8_t>(o_cs), 4 * (size_t)(ncell + 1));
        memcpy(cell_items, cs.out<uint8_t>(o_ci), 4 * (size_t)cell_start[ncell]);
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_frame_postprocess", __FILE__, __LINE__);
    *n_in_grid = cell_start[ncell];
    return LPSLAM_OK;
}

extern "C" int lpslam_match_area(const lpslam_camera* cam, uint32_t n1, const lpslam_keypoint* kp1_undist, const uint8_t* desc1,
                                 float* prev_matched_pts, uint32_t n2, const lpslam_keypoint* kp2_undist, const uint8_t* desc2,
                                 int margin, float lowe_ratio, int check_orientation, int32_t* matched_indices_2_in_frm_1,
                                 uint32_t* n_matches) {
    LPS_ARG(cam && n_matches, "null argument");
    *n_matches = 0;
    LPS_ARG(n1 == 0 || (kp1_undist && desc1 && prev_matched_pts && matched_indices_2_in_frm_1), "null frame-1 argument");
    LPS_ARG(n2 == 0 || (kp2_undist && desc2), "null frame-2 argument");
    LPS_ARG(cam->grid_cols >= 1 && cam->grid_rows >= 1, "bad camera grid");
    LPS_ARG(n2 <= 24000 && n1 < (1u << 20) - 1, "too many keypoints (the per-keypoint state of frame 2 lives in shared memory)");
    for (uint32_t i = 0; i < n1; ++i) matched_indices_2_in_frm_1[i] = -1;
    if (n1 == 0 || n2 == 0) return LPSLAM_OK;
    const uint32_t ncell = cam->grid_cols * cam->grid_rows;
    ArenaBump ar;  // frame-2 keypoints + zeroed counters | other inputs | work | outputs
    const size_t o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)n2), o_cc = ar.take(4 * (size_t)ncell), grid_end = ar.off;
    const size_t o_use = ar.take(n1), o_pp = ar.take(8 * (size_t)n1), o_d1 = ar.take(32 * (size_t)n1), o_lvl = ar.take(4 * (size_t)n1);
    const size_t o_d2 = ar.take(32 * (size_t)n2), in_end = ar.off;
    const size_t o_cs = ar.take(4 * (size_t)(ncell + 1)), o_cf = ar.take(4 * (size_t)ncell);
    const size_t o_ci = ar.take(4 * (size_t)n2), o_tkey = ar.take(16 * (size_t)n1), o_tidx = ar.take(16 * (size_t)n1);
    const size_t o_m12 = ar.take(4 * (size_t)n2), o_m21 = ar.take(4 * (size_t)n1), o_ne = ar.take(8), o_ev = ar.take(8 * (size_t)n1);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    std::vector<int2> events;
    uint32_t ne[2] = {0, 0};
    do {
        // the grid only needs the keypoints of frame 2: they go first, the rest is packed while that copy and the grid kernel run
        cs.put(o_kp, kp2_undist, sizeof(lpslam_keypoint) * (size_t)n2);
        cs.zero(o_cc, 4 * (size_t)ncell);
        if ((e = cs.upload(0, grid_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.cam = *cam;
        a.mode = 2;
        a.n_last = n1;
        a.n_curr = n2;
        a.last_has_lm = dev + o_use;
        a.last_desc = dev + o_d1;
        a.last_octave = (const int32_t*)(dev + o_lvl);
        a.prev_pts = (const float*)(dev + o_pp);
        a.lowe_ratio = lowe_ratio;
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_d2;
        a.margin = (float)margin;
        a.cell_count = (uint32_t*)(dev + o_cc);
        a.cell_start = (uint32_t*)(dev + o_cs);
        a.cell_fill = (uint32_t*)(dev + o_cf);
        a.cell_items = (uint32_t*)(dev + o_ci);
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.match_2_of_1 = (int32_t*)(dev + o_m21);
        a.match_1_of_2 = (int32_t*)(dev + o_m12);
        a.events = (int2*)(dev + o_ev);
        a.n_events = (uint32_t*)(dev + o_ne);
        launch_grid(a, n2, ncell, cs.st);
        uint8_t* use = cs.host + o_use;
        for (uint32_t i = 0; i < n1; ++i) use[i] = kp1_undist[i].octave <= 0;  // area.cc:23-26
        cs.put(o_pp, prev_matched_pts, 8 * (size_t)n1);
        cs.put(o_d1, desc1, 32 * (size_t)n1);
        cs.zero(o_lvl, 4 * (size_t)n1);
        cs.put(o_d2, desc2, 32 * (size_t)n2);
        if ((e = cs.upload_beside(grid_end, in_end)) != cudaSuccess) break;
        k_proj_topk<<<(n1 + 3) / 4, 128, 0, cs.st>>>(a);
        const size_t smem = 8 * (size_t)n2;
        if (smem > 48 * 1024 &&
            (e = cudaFuncSetAttribute(k_area_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
            break;
        k_area_resolve<<<1, 32, smem, cs.st>>>(a);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_m21, check_orientation ? ar.off : o_ev)) != cudaSuccess) break;
        memcpy(matched_indices_2_in_frm_1, cs.out<int32_t>(o_m21), 4 * (size_t)n1);
        ne[0] = cs.out<uint32_t>(o_ne)[0];
        ne[1] = cs.out<uint32_t>(o_ne)[1];
        if (check_orientation) events.assign(cs.out<int2>(o_ev), cs.out<int2>(o_ev) + ne[0]);
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_area", __FILE__, __LINE__);
    int num = (int)ne[1];
    if (check_orientation) {  // area.cc:87-101: every accepted row entered the histogram, also rows that lost their match later
        std::vector<std::pair<float, int>> deltas;
        for (const int2& ev : events) deltas.emplace_back(kp1_undist[ev.x].angle - kp2_undist[ev.y].angle, ev.x);
        for (int id : angle_invalid_matches(deltas)) {
            if (0 <= matched_indices_2_in_frm_1[id]) {
                matched_indices_2_in_frm_1[id] = -1;
                --num;
            }
        }
    }
    for (uint32_t i = 0; i < n1; ++i)  // area.cc:104-108
        if (0 <= matched_indices_2_in_frm_1[i]) {
            prev_matched_pts[2 * i] = kp2_undist[matched_indices_2_in_frm_1[i]].x;
            prev_matched_pts[2 * i + 1] = kp2_undist[matched_indices_2_in_frm_1[i]].y;
        }
    *n_matches = (uint32_t)num;
    return LPSLAM_OK;
}

extern "C" int lpslam_match_candidates(uint32_t n_rows, const uint8_t* row_desc, const uint32_t* row_cand_begin,
                                       const uint32_t* row_cand_end, const uint32_t* cand_items, uint32_t n_items,
                                       const float* row_angle, uint32_t n_kp, const uint8_t* kp_desc, const float* kp_angle,
                                       const uint8_t* kp_taken, uint32_t hamm_dist_thr, float lowe_ratio, int use_ratio,
                                       int32_t* kp_match_row, uint32_t* n_matches) {
    LPS_ARG(n_matches, "null argument");
    *n_matches = 0;
    LPS_ARG(n_kp == 0 || (kp_desc && kp_match_row), "null keypoint argument");
    LPS_ARG(n_rows == 0 || (row_desc && row_cand_begin && row_cand_end), "null row argument");
    LPS_ARG(n_items == 0 || cand_items, "null candidate list");
    LPS_ARG(!row_angle || kp_angle, "keypoint angles required for the orientation check");
    LPS_ARG(n_kp < (1u << 20) && hamm_dist_thr <= 256, "bad sizes / threshold");
    for (uint32_t i = 0; i < n_kp; ++i) kp_match_row[i] = -1;
    if (n_kp == 0 || n_rows == 0) return LPSLAM_OK;
    for (uint32_t r = 0; r < n_rows; ++r)
        LPS_ARG(row_cand_begin[r] <= row_cand_end[r] && row_cand_end[r] <= n_items, "candidate range out of bounds");
    for (uint32_t q = 0; q < n_items; ++q) LPS_ARG(cand_items[q] < n_kp, "candidate index out of range");

    ArenaBump ar;  // inputs | work | outputs
    const size_t o_rd = ar.take(32 * (size_t)n_rows), o_cb = ar.take(4 * (size_t)n_rows), o_ce = ar.take(4 * (size_t)n_rows);
    const size_t o_it = ar.take(4 * (size_t)n_items), o_kd = ar.take(32 * (size_t)n_kp), o_tk = ar.take(n_kp), in_end = ar.off;
    const size_t o_tkey = ar.take(16 * (size_t)n_rows), o_tidx = ar.take(16 * (size_t)n_rows), o_m = ar.take(4 * (size_t)n_kp);
    const size_t o_ne = ar.take(4), o_ev = ar.take(8 * (size_t)n_rows);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    uint32_t n_ev = 0;
    std::vector<int2> events;
    do {
        cs.put(o_rd, row_desc, 32 * (size_t)n_rows);
        cs.put(o_cb, row_cand_begin, 4 * (size_t)n_rows);
        cs.put(o_ce, row_cand_end, 4 * (size_t)n_rows);
        cs.put(o_it, cand_items, 4 * (size_t)n_items);
        cs.put(o_kd, kp_desc, 32 * (size_t)n_kp);
        if (kp_taken) cs.put(o_tk, kp_taken, n_kp);
        else cs.zero(o_tk, n_kp);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.mode = 3;
        a.n_last = n_rows;
        a.n_curr = n_kp;
        a.last_desc = dev + o_rd;
        a.cand_begin = (const uint32_t*)(dev + o_cb);
        a.cand_end = (const uint32_t*)(dev + o_ce);
        a.cand_items = (const uint32_t*)(dev + o_it);
        a.desc = dev + o_kd;
        a.curr_taken = dev + o_tk;
        a.lowe_ratio = lowe_ratio;
        a.use_ratio = use_ratio;
        a.hamm_thr = hamm_dist_thr;
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.curr_match_last = (int32_t*)(dev + o_m);
        a.events = row_angle ? (int2*)(dev + o_ev) : nullptr;
        a.n_events = (uint32_t*)(dev + o_ne);
        k_proj_topk<<<(n_rows + 3) / 4, 128, 0, cs.st>>>(a);
        a.claim_in_smem = n_kp <= kResolveTableMax;
        if ((e = launch_lm_resolve(a, n_kp, cs.st)) != cudaSuccess) break;
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_m, row_angle ? ar.off : o_ev)) != cudaSuccess) break;
        memcpy(kp_match_row, cs.out<int32_t>(o_m), 4 * (size_t)n_kp);
        n_ev = *cs.out<uint32_t>(o_ne);
        if (row_angle) events.assign(cs.out<int2>(o_ev), cs.out<int2>(o_ev) + n_ev);
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_candidates", __FILE__, __LINE__);
    *n_matches = n_ev;
    if (row_angle) {  // bow_tree.cc:91-95,111-117
        std::vector<std::pair<float, int>> deltas;
        for (const int2& ev : events) deltas.emplace_back(row_angle[ev.x] - kp_angle[ev.y], ev.y);
        for (int id : angle_invalid_matches(deltas)) {
            kp_match_row[id] = -1;
            --*n_matches;
        }
    }
    return LPSLAM_OK;
}

// rows that search independently of each other (fuse.cc:69-101 / :176-238, projection.cc:503-533 / :585-615): grid build,
// k_proj_topk, k_rows_best; best_idx[row] = the closest keypoint when its distance is <= hamm_thr
static int rows_best_search(const char* who, uint32_t hamm_thr, const lpslam_camera* cam, float margin, int chi_square_gate,
                            int unsigned_levels, uint32_t n_lm, const uint8_t* lm_active, const double* lm_reproj,
                            const float* lm_x_right, const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp,
                            const lpslam_keypoint* kp_undist, const uint8_t* kp_desc, const float* kp_x_right,
                            const float* scale_factors, const float* inv_level_sigma_sq, int32_t* best_idx,
                            uint32_t* best_dist, uint32_t* n_fused) {
    LPS_ARG(cam && scale_factors && n_fused, "null argument");
    *n_fused = 0;
    LPS_ARG(n_lm == 0 || (lm_active && lm_reproj && lm_scale_level && lm_desc && best_idx), "null landmark argument");
    LPS_ARG(n_kp == 0 || (kp_undist && kp_desc), "null keyframe argument");
    LPS_ARG(!chi_square_gate || (inv_level_sigma_sq && (n_lm == 0 || lm_x_right) && (n_kp == 0 || kp_x_right)),
            "the reprojection-error gate needs x_right of both sides and inv_level_sigma_sq");
    LPS_ARG(cam->grid_cols >= 1 && cam->grid_rows >= 1 && cam->num_scale_levels >= 1 && cam->num_scale_levels <= (uint32_t)kMaxLevels,
            "bad camera grid / level count");
    LPS_ARG(n_kp < (1u << 20), "too many keypoints");
    for (uint32_t i = 0; i < n_lm; ++i) {
        best_idx[i] = -1;
        if (best_dist) best_dist[i] = 256;
    }
    if (n_kp == 0 || n_lm == 0) return LPSLAM_OK;
    std::vector<uint8_t> act(lm_active, lm_active + n_lm);
    for (uint32_t i = 0; i < n_lm; ++i) {
        LPS_ARG(!act[i] || (lm_scale_level[i] >= 0 && (uint32_t)lm_scale_level[i] < cam->num_scale_levels),
                "landmark scale level out of range");
        // fuse.cc:193 compares unsigned levels: `scale_level < pred_scale_level - 1` wraps for a predicted level of 0
        if (unsigned_levels && lm_scale_level[i] == 0) act[i] = 0;
    }
    const uint32_t ncell = cam->grid_cols * cam->grid_rows;
    ArenaBump ar;  // inputs | zeroed counters | work | outputs
    const size_t o_act = ar.take(n_lm), o_rp = ar.take(16 * (size_t)n_lm), o_lxr = ar.take(4 * (size_t)n_lm), o_lvl = ar.take(4 * (size_t)n_lm);
    const size_t o_ld = ar.take(32 * (size_t)n_lm), o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)n_kp);
    const size_t o_desc = ar.take(32 * (size_t)n_kp), o_xr = ar.take(4 * (size_t)n_kp), o_sc = ar.take(4 * kMaxLevels);
    const size_t o_isq = ar.take(4 * kMaxLevels), o_cc = ar.take(4 * (size_t)ncell), in_end = ar.off;
    const size_t o_cs = ar.take(4 * (size_t)(ncell + 1));
    const size_t o_cf = ar.take(4 * (size_t)ncell), o_ci = ar.take(4 * (size_t)n_kp), o_tkey = ar.take(16 * (size_t)n_lm);
    const size_t o_tidx = ar.take(16 * (size_t)n_lm), o_rb = ar.take(4 * (size_t)n_lm), o_rd = ar.take(4 * (size_t)n_lm);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    const uint32_t* dist = nullptr;
    do {
        cs.put(o_act, act.data(), n_lm);
        cs.put(o_rp, lm_reproj, 16 * (size_t)n_lm);
        cs.put(o_lvl, lm_scale_level, 4 * (size_t)n_lm);
        cs.put(o_ld, lm_desc, 32 * (size_t)n_lm);
        cs.put(o_kp, kp_undist, sizeof(lpslam_keypoint) * (size_t)n_kp);
        cs.put(o_desc, kp_desc, 32 * (size_t)n_kp);
        cs.put(o_sc, scale_factors, 4 * (size_t)cam->num_scale_levels);
        if (chi_square_gate) {
            cs.put(o_lxr, lm_x_right, 4 * (size_t)n_lm);
            cs.put(o_xr, kp_x_right, 4 * (size_t)n_kp);
            cs.put(o_isq, inv_level_sigma_sq, 4 * (size_t)cam->num_scale_levels);
        }
        cs.zero(o_cc, 4 * (size_t)ncell);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.cam = *cam;
        a.mode = 1;
        a.n_last = n_lm;
        a.n_curr = n_kp;
        a.last_has_lm = dev + o_act;
        a.last_desc = dev + o_ld;
        a.last_octave = (const int32_t*)(dev + o_lvl);
        a.lm_reproj = (const double*)(dev + o_rp);
        a.lm_x_right = (const float*)(dev + o_lxr);
        a.level_lo = -1;  // fuse.cc:84-87 / :193-195: octave in [pred - 1, pred]
        a.level_hi = 0;
        a.use_xr = 0;
        a.hamm_thr = hamm_thr;
        a.chi_gate = chi_square_gate ? 1 : 0;
        a.inv_sigma_sq = (const float*)(dev + o_isq);
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_desc;
        a.x_right = (const float*)(dev + o_xr);
        a.scale = (const float*)(dev + o_sc);
        a.margin = margin;
        a.cell_count = (uint32_t*)(dev + o_cc);
        a.cell_start = (uint32_t*)(dev + o_cs);
        a.cell_fill = (uint32_t*)(dev + o_cf);
        a.cell_items = (uint32_t*)(dev + o_ci);
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.row_best = (int32_t*)(dev + o_rb);
        a.row_dist = (uint32_t*)(dev + o_rd);
        launch_grid(a, n_kp, ncell, cs.st);
        k_proj_topk<<<(n_lm + 3) / 4, 128, 0, cs.st>>>(a);
        k_rows_best<<<(n_lm + 255) / 256, 256, 0, cs.st>>>(a);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_rb, ar.off)) != cudaSuccess) break;
        memcpy(best_idx, cs.out<int32_t>(o_rb), 4 * (size_t)n_lm);
        dist = cs.out<uint32_t>(o_rd);
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, who, __FILE__, __LINE__);
    for (uint32_t i = 0; i < n_lm; ++i) {
        if (best_dist) best_dist[i] = dist[i];
        *n_fused += best_idx[i] >= 0;
    }
    return LPSLAM_OK;
}

extern "C" int lpslam_match_fuse(const lpslam_camera* cam, float margin, int chi_square_gate, int unsigned_levels, uint32_t n_lm,
                                 const uint8_t* lm_active, const double* lm_reproj, const float* lm_x_right,
                                 const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp,
                                 const lpslam_keypoint* kp_undist, const uint8_t* kp_desc, const float* kp_x_right,
                                 const float* scale_factors, const float* inv_level_sigma_sq, int32_t* best_idx,
                                 uint32_t* best_dist, uint32_t* n_fused) {
    // HAMMING_DIST_THR_LOW, fuse.cc:99 / :236
    return rows_best_search("lpslam_match_fuse", 50, cam, margin, chi_square_gate, unsigned_levels, n_lm, lm_active, lm_reproj,
                            lm_x_right, lm_scale_level, lm_desc, n_kp, kp_undist, kp_desc, kp_x_right, scale_factors,
                            inv_level_sigma_sq, best_idx, best_dist, n_fused);
}

extern "C" int lpslam_match_sim3(const lpslam_camera* cam, float margin, uint32_t n_lm, const uint8_t* lm_active,
                                 const double* lm_reproj, const int32_t* lm_scale_level, const uint8_t* lm_desc, uint32_t n_kp,
                                 const lpslam_keypoint* kp_undist, const uint8_t* kp_desc, const uint8_t* kp_has_matched_lm,
                                 const float* scale_factors, int32_t* kp_match_lm, uint32_t* n_matches) {
    LPS_ARG(n_matches, "null argument");
    *n_matches = 0;
    LPS_ARG(n_lm == 0 || (lm_active && lm_scale_level), "null landmark argument");
    // projection.cc:385-390 compares unsigned levels: `scale_level < pred_scale_level - 1` wraps for a predicted level of 0,
    // every candidate of such a row is skipped
    std::vector<uint8_t> act(lm_active, lm_active + n_lm);
    for (uint32_t i = 0; i < n_lm; ++i)
        if (lm_scale_level[i] == 0) act[i] = 0;
    // octave in [pred - 1, pred], best <= HAMMING_DIST_THR_LOW (:405), no ratio test, no stereo gate; a matched keypoint
    // holds a landmark from then on (:381-383, :409)
    const lpslam_reproj_match_params prm = {margin, 0.f, 50, -1, 0, 0, 0};
    return lpslam_match_reprojected(cam, &prm, n_lm, act.data(), lm_reproj, nullptr, lm_scale_level, lm_desc, nullptr, nullptr, n_kp,
                                    kp_undist, kp_desc, nullptr, kp_has_matched_lm, scale_factors, kp_match_lm, n_matches);
}

extern "C" int lpslam_match_mutual(const lpslam_camera* cam_1, const lpslam_camera* cam_2, float margin, uint32_t n_1,
                                   const uint8_t* lm1_active, const double* lm1_reproj_in_2, const int32_t* lm1_level_in_2,
                                   const uint8_t* lm1_desc, const lpslam_keypoint* kp1_undist, const uint8_t* kp1_desc,
                                   const float* scale_factors_1, uint32_t n_2, const uint8_t* lm2_active,
                                   const double* lm2_reproj_in_1, const int32_t* lm2_level_in_1, const uint8_t* lm2_desc,
                                   const lpslam_keypoint* kp2_undist, const uint8_t* kp2_desc, const float* scale_factors_2,
                                   int32_t* matched_idx_2_in_1, uint32_t* n_matches) {
    LPS_ARG(n_matches, "null argument");
    *n_matches = 0;
    LPS_ARG(n_1 == 0 || matched_idx_2_in_1, "null result");
    std::vector<int32_t> two_of_1(n_1, -1), one_of_2(n_2, -1);
    uint32_t n = 0;
    // landmarks of keyframe 1 into keyframe 2 (projection.cc:453-534), then the other way round (:536-616); both loops
    // compare unsigned levels (:511-513, :593-595) and accept best <= HAMMING_DIST_THR_HIGH (:530, :612)
    int rc = rows_best_search("lpslam_match_mutual", HAMMING_DIST_THR_HIGH, cam_2, margin, 0, 1, n_1, lm1_active, lm1_reproj_in_2,
                              nullptr, lm1_level_in_2, lm1_desc, n_2, kp2_undist, kp2_desc, nullptr, scale_factors_2, nullptr,
                              two_of_1.data(), nullptr, &n);
    if (rc != LPSLAM_OK) return rc;
    rc = rows_best_search("lpslam_match_mutual", HAMMING_DIST_THR_HIGH, cam_1, margin, 0, 1, n_2, lm2_active, lm2_reproj_in_1, nullptr,
                          lm2_level_in_1, lm2_desc, n_1, kp1_undist, kp1_desc, nullptr, scale_factors_1, nullptr, one_of_2.data(),
                          nullptr, &n);
    if (rc != LPSLAM_OK) return rc;
    // projection.cc:618-631: only the cross-matches are recorded
    for (uint32_t i = 0; i < n_1; ++i) {
        const int32_t i2 = two_of_1[i];
        matched_idx_2_in_1[i] = (i2 >= 0 && one_of_2[i2] == (int32_t)i) ? i2 : -1;
        *n_matches += matched_idx_2_in_1[i] >= 0;
    }
    return LPSLAM_OK;
}

extern "C" int lpslam_match_triangulation(uint32_t n_rows, const uint32_t* row_idx_1, const uint32_t* row_cand_begin,
                                          const uint32_t* row_cand_end, const uint32_t* cand_items, uint32_t n_items, uint32_t n_1,
                                          const lpslam_keypoint* kp1_undist, const uint8_t* kp1_desc, const double* bearings_1,
                                          const float* x_right_1, const float* scale_factors_1, uint32_t n_2,
                                          const lpslam_keypoint* kp2_undist, const uint8_t* kp2_desc, const double* bearings_2,
                                          const float* x_right_2, const uint8_t* kp2_has_lm, const double* E_12,
                                          const double* epiplane_in_keyfrm_2, int check_orientation,
                                          int32_t* matched_idx_2_in_1, uint32_t* n_matches) {
    LPS_ARG(n_matches, "null argument");
    *n_matches = 0;
    LPS_ARG(n_1 == 0 || (kp1_undist && kp1_desc && bearings_1 && x_right_1 && matched_idx_2_in_1), "null keyframe-1 argument");
    LPS_ARG(n_2 == 0 || (kp2_undist && kp2_desc && bearings_2 && x_right_2 && kp2_has_lm), "null keyframe-2 argument");
    LPS_ARG(n_rows == 0 || (row_idx_1 && row_cand_begin && row_cand_end), "null row argument");
    LPS_ARG(n_items == 0 || cand_items, "null candidate list");
    LPS_ARG(E_12 && epiplane_in_keyfrm_2 && scale_factors_1, "null geometry argument");
    LPS_ARG(n_2 < (1u << 20) && n_items < (1u << 20), "too many keypoints / candidates");
    for (uint32_t i = 0; i < n_1; ++i) matched_idx_2_in_1[i] = -1;
    if (n_1 == 0 || n_2 == 0 || n_rows == 0) return LPSLAM_OK;
    for (uint32_t r = 0; r < n_rows; ++r) {
        LPS_ARG(row_idx_1[r] < n_1, "row keypoint out of range");
        LPS_ARG(row_cand_begin[r] <= row_cand_end[r] && row_cand_end[r] <= n_items, "candidate range out of bounds");
        LPS_ARG(kp1_undist[row_idx_1[r]].octave >= 0 && kp1_undist[row_idx_1[r]].octave < kMaxLevels, "octave out of range");
    }
    for (uint32_t q = 0; q < n_items; ++q) LPS_ARG(cand_items[q] < n_2, "candidate index out of range");

    ArenaBump ar;  // inputs | work | outputs
    const size_t o_rd = ar.take(32 * (size_t)n_rows), o_cb = ar.take(4 * (size_t)n_rows), o_ce = ar.take(4 * (size_t)n_rows);
    const size_t o_it = ar.take(4 * (size_t)n_items), o_kd = ar.take(32 * (size_t)n_2), o_tk = ar.take(n_2);
    const size_t o_rb = ar.take(24 * (size_t)n_rows), o_kb = ar.take(24 * (size_t)n_2), o_rs = ar.take(n_rows), o_ks = ar.take(n_2);
    const size_t o_rsc = ar.take(4 * (size_t)n_rows), in_end = ar.off;
    const size_t o_tkey = ar.take(16 * (size_t)n_rows), o_tidx = ar.take(16 * (size_t)n_rows), o_m = ar.take(4 * (size_t)n_2);
    const size_t o_ne = ar.take(4), o_ev = ar.take(8 * (size_t)n_rows);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    uint32_t n_ev = 0;
    std::vector<int2> events;
    do {
        // per-row copies of the keyframe-1 fields (rows are keyframe-1 keypoints in feature-vector order), written in place
        uint8_t* rdesc = cs.host + o_rd;
        double* rbear = (double*)(cs.host + o_rb);
        uint8_t* rstereo = cs.host + o_rs;
        float* rscale = (float*)(cs.host + o_rsc);
        for (uint32_t r = 0; r < n_rows; ++r) {
            const uint32_t i1 = row_idx_1[r];
            memcpy(rdesc + 32 * (size_t)r, kp1_desc + 32 * (size_t)i1, 32);
            memcpy(rbear + 3 * (size_t)r, bearings_1 + 3 * (size_t)i1, 24);
            rstereo[r] = 0 <= x_right_1[i1];
            rscale[r] = scale_factors_1[kp1_undist[i1].octave];
        }
        uint8_t* kstereo = cs.host + o_ks;
        for (uint32_t i = 0; i < n_2; ++i) kstereo[i] = 0 <= x_right_2[i];
        cs.put(o_cb, row_cand_begin, 4 * (size_t)n_rows);
        cs.put(o_ce, row_cand_end, 4 * (size_t)n_rows);
        cs.put(o_it, cand_items, 4 * (size_t)n_items);
        cs.put(o_kd, kp2_desc, 32 * (size_t)n_2);
        cs.put(o_tk, kp2_has_lm, n_2);
        cs.put(o_kb, bearings_2, 24 * (size_t)n_2);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        ProjArgs a = {};
        a.mode = 3;
        a.tri = 1;
        a.last_wins = 1;  // robust.cc:108: an equal distance replaces the best
        a.n_last = n_rows;
        a.n_curr = n_2;
        a.last_desc = dev + o_rd;
        a.cand_begin = (const uint32_t*)(dev + o_cb);
        a.cand_end = (const uint32_t*)(dev + o_ce);
        a.cand_items = (const uint32_t*)(dev + o_it);
        a.desc = dev + o_kd;
        a.curr_taken = dev + o_tk;  // robust.cc:87-91: keypoints with a landmark are never candidates
        a.row_bearing = (const double*)(dev + o_rb);
        a.kp_bearing = (const double*)(dev + o_kb);
        a.row_stereo = dev + o_rs;
        a.kp_stereo = dev + o_ks;
        a.row_scale = (const float*)(dev + o_rsc);
        for (int i = 0; i < 9; ++i) a.E[i] = E_12[i];
        for (int i = 0; i < 3; ++i) a.epi[i] = epiplane_in_keyfrm_2[i];
        a.use_ratio = 0;
        a.hamm_thr = 50;  // HAMMING_DIST_THR_LOW, robust.cc:78,108
        a.top_key = (uint4*)(dev + o_tkey);
        a.top_idx = (int4*)(dev + o_tidx);
        a.curr_match_last = (int32_t*)(dev + o_m);
        a.events = (int2*)(dev + o_ev);
        a.n_events = (uint32_t*)(dev + o_ne);
        k_proj_topk<<<(n_rows + 3) / 4, 128, 0, cs.st>>>(a);
        a.claim_in_smem = n_2 <= kResolveTableMax;
        if ((e = launch_lm_resolve(a, n_2, cs.st)) != cudaSuccess) break;
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_ne, ar.off)) != cudaSuccess) break;
        n_ev = *cs.out<uint32_t>(o_ne);
        events.assign(cs.out<int2>(o_ev), cs.out<int2>(o_ev) + n_ev);
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_triangulation", __FILE__, __LINE__);
    *n_matches = n_ev;
    std::vector<std::pair<float, int>> deltas;
    for (const int2& ev : events) {  // events are in row order = the order of robust.cc:133-141
        const uint32_t i1 = row_idx_1[ev.x];
        matched_idx_2_in_1[i1] = ev.y;
        if (check_orientation) deltas.emplace_back(kp1_undist[i1].angle - kp2_undist[ev.y].angle, (int)i1);
    }
    if (check_orientation)
        for (int id : angle_invalid_matches(deltas)) {
            matched_idx_2_in_1[id] = -1;
            --*n_matches;
        }
    return LPSLAM_OK;
}
