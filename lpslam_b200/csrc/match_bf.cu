// K7 + K8 — brute-force Hamming matcher with ratio test. Replaces match::robust::brute_force_match (match/robust.cc:215-321)
// and carries the Hamming primitive of match/base.h:18-63.
//
// The reference scans keyframe rows in order; each row takes the best and second-best frame descriptor AMONG THOSE NOT YET
// CLAIMED by an earlier row (strict <, so the lowest index wins ties), accepts on best <= 50 and ratio * second >= best,
// and claims its best. The claim makes the loop order-dependent, so the work is split:
//   k_bf_topk     all rows in parallel: the 4 smallest keys (distance << 16 | index) of every row ignoring claims.
//                 In key order, "best / second among unclaimed" is simply the first two unclaimed keys.
//   k_bf_resolve  replays the rows in order against the claim set. Order-exact parallel rounds in one block: a row whose first
//                 unclaimed key no earlier unresolved row can take, and whose second is fixed or cannot flip the ratio test, is
//                 decided as the sequential loop would and committed with all other such rows; a row whose top-4 list may run out
//                 of unclaimed entries blocks the rows behind it and is rescanned exactly by the whole block when it is the first
//                 unresolved row. Contended inputs fall back to a serial warp loop (32 rows per round, commit up to the first
//                 conflict).
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

// Which frame keypoints are claimed. versioned: p[c] = EMPTY (free) or row + 1 of the row that claimed it - for row r a keypoint
// claimed by a LATER row (the parallel rounds commit out of order) still counts as free, as it was when the sequential loop
// reached r. Otherwise p is a plain bitmap (table too large for shared memory: serial loop only).
struct ClaimSet {
    uint32_t* p;
    bool versioned;
    __device__ __forceinline__ bool claimed(uint32_t c, uint32_t row) const {
        return versioned ? p[c] <= row : ((p[c >> 5] >> (c & 31)) & 1u) != 0;
    }
    __device__ __forceinline__ void claim(uint32_t c, uint32_t row) const {
        if (versioned) p[c] = row + 1;  // one writer per keypoint
        else atomicOr(&p[c >> 5], 1u << (c & 31));
    }
};

// Decision of one row from its four kept keys against the claim bitmap: the first two unclaimed keys; `exact` = they (or the
// distance bound of the fourth key) settle the row without looking at the rest of the frame.
struct BfDecision {
    uint32_t best, second;
    bool exact, accept;
};

__device__ __forceinline__ bool bf_accept(uint32_t best, uint32_t second, float lowe_ratio) {
    if (best == EMPTY) return false;
    const unsigned bd = best >> 16;
    const unsigned sd = second == EMPTY ? MAX_HAMMING_DIST : (second >> 16);
    if (HAMMING_DIST_THR_LOW < bd) return false;
    return !(__fmul_rn(lowe_ratio, (float)sd) < (float)bd);  // match/robust.cc:285
}

__device__ __forceinline__ BfDecision bf_decide(uint32_t row, const uint32_t (&k)[K], const ClaimSet& cl, float lowe_ratio) {
    BfDecision d;
    d.best = d.second = EMPTY;
    int seen = 0;
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (k[q] == EMPTY) continue;
        const uint32_t i1 = k[q] & 0xFFFFu;
        if (cl.claimed(i1, row)) continue;
        if (seen == 0) d.best = k[q];
        else if (seen == 1) d.second = k[q];
        ++seen;
    }
    const bool list_complete = (k[K - 1] == EMPTY);  // fewer than K frame keypoints exist: nothing beyond the list
    d.exact = (seen >= 2) || list_complete;
    if (!d.exact && seen == 0 && (k[K - 1] >> 16) > HAMMING_DIST_THR_LOW) d.exact = true;  // everything else is farther than the fourth
    if (!d.exact && seen == 1) {
        // second >= the list's last distance; if the ratio test passes with that bound it passes with the truth
        const unsigned lb = k[K - 1] >> 16, bd = d.best >> 16;
        if (bd > HAMMING_DIST_THR_LOW) d.exact = true;  // rejected whatever the second is
        else if (!(__fmul_rn(lowe_ratio, (float)lb) < (float)bd)) {
            d.exact = true;
            d.second = lb << 16;
        }
    }
    d.accept = bf_accept(d.best, d.second, lowe_ratio);
    return d;
}

// Row r in a parallel round (see k_lm_resolve, match_proj.cu): a key is "own" when it is unclaimed and no earlier unresolved row
// could take it (s_first >= r), "contested" when an earlier row might. clear: accept / best are what the sequential loop will
// reach. may_rescan: some outcome of the contested keys leaves fewer unclaimed keys than the decision needs.
struct BfEval {
    bool clear, may_rescan, accept;
    uint32_t best;
};

__device__ __forceinline__ BfEval bf_eval(uint32_t r, const uint32_t (&k)[K], const ClaimSet& cl, const uint32_t* s_first, float lowe_ratio) {
    BfEval e = {false, false, false, EMPTY};
    const bool is_long = k[K - 1] != EMPTY;
    const unsigned lb = k[K - 1] >> 16;
    auto single_ok = [&](unsigned bd) { return bd > HAMMING_DIST_THR_LOW || !(__fmul_rn(lowe_ratio, (float)lb) < (float)bd); };
    int qb = -1, qs = -1, n_own = 0;
    bool b_own = false, s_own = false, contested_single_bad = false;
    unsigned own_bd = 0;
#pragma unroll
    for (int q = 0; q < K; ++q) {
        if (k[q] == EMPTY) continue;
        const uint32_t c = k[q] & 0xFFFFu;
        if (cl.claimed(c, r)) continue;
        const bool own = s_first[c] >= r;
        if (own) {
            if (n_own == 0) own_bd = k[q] >> 16;
            ++n_own;
        }
        else if (!single_ok(k[q] >> 16)) contested_single_bad = true;
        if (qb < 0) {
            qb = q;
            b_own = own;
        }
        else if (qs < 0) {
            qs = q;
            s_own = own;
        }
    }
    if (is_long) {
        if (n_own == 0 && !(lb > HAMMING_DIST_THR_LOW)) e.may_rescan = true;
        if (n_own == 0 && contested_single_bad) e.may_rescan = true;
        if (n_own == 1 && !single_ok(own_bd)) e.may_rescan = true;
    }
    if (qb < 0) {
        e.clear = !is_long || lb > HAMMING_DIST_THR_LOW;
        return e;
    }
    if (!b_own) return e;
    e.best = k[qb];
    const unsigned bd = k[qb] >> 16;
    if (bd > HAMMING_DIST_THR_LOW) {
        e.clear = true;
        return e;
    }
    if (qs >= 0) {
        const unsigned sd = k[qs] >> 16;
        const bool pass = !(__fmul_rn(lowe_ratio, (float)sd) < (float)bd);
        if (s_own || pass) {  // the second is exactly this key, or anything that replaces it is farther
            e.clear = true;
            e.accept = pass;
        }
        return e;
    }
    const unsigned sd = is_long ? lb : MAX_HAMMING_DIST;
    if (!(__fmul_rn(lowe_ratio, (float)sd) < (float)bd)) {
        e.clear = true;
        e.accept = true;
    }
    else if (!is_long) e.clear = true;  // complete list, fails against MAX_HAMMING_DIST
    return e;
}

// exact best / second of `row` over the unclaimed frame keypoints (match/robust.cc:252-274); every participating thread walks
// i = part, part + parts, ... and returns its own partial (b, s)
__device__ __forceinline__ void bf_scan_unclaimed(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t row,
                                                  const ClaimSet& cl, uint32_t part, uint32_t parts, uint32_t& b, uint32_t& s) {
    const uint4* d2 = reinterpret_cast<const uint4*>(desc_2 + (size_t)row * 32);
    const uint4 a0 = __ldg(d2), a1 = __ldg(d2 + 1);
    b = s = EMPTY;
    for (uint32_t i = part; i < n1; i += parts) {
        if (cl.claimed(i, row)) continue;
        const uint4* d1 = reinterpret_cast<const uint4*>(desc_1 + (size_t)i * 32);
        const uint32_t key = (hamming256(a0, a1, __ldg(d1), __ldg(d1 + 1)) << 16) | i;
        if (key < b) { s = b; b = key; }
        else if (key < s) s = key;
    }
}

__device__ __forceinline__ void bf_warp_merge(uint32_t& b, uint32_t& s) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, b, o), os = __shfl_xor_sync(0xFFFFFFFFu, s, o);
        const uint32_t nb = min(b, ob);
        const uint32_t ns = min(max(b, ob), min(s, os));
        b = nb;
        s = ns;
    }
}

// Serial form, one warp, rows in reference order from `first_row` on, 32 rows per round: every lane decides its own row from its
// four kept keys against the claim bitmap; rows are committed in order up to the first one whose kept keys contain a keypoint
// that an earlier row of the round just claimed; that row and the ones behind it are decided again. A row whose four keys do not
// settle the decision is rescanned exactly by the whole warp when it is the first pending row.
// s_first (optional, n1 words behind the bitmap): lowest lane claiming a keypoint in the current round.
__device__ uint32_t bf_resolve_serial(const uint8_t* desc_1, uint32_t n1, const uint8_t* desc_2, uint32_t n2, const uint4* topk,
                                      float lowe_ratio, const ClaimSet& cl, uint32_t* s_first, int32_t* match_2_in_1, int lane,
                                      uint32_t first_row) {
    uint32_t num = 0;
    for (uint32_t base = first_row; base < n2; base += 32) {
        const uint32_t row = base + lane;
        const uint4 mk = row < n2 ? topk[row] : make_uint4(EMPTY, EMPTY, EMPTY, EMPTY);
        const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
        // rows whose unconditional best is already > 50 can never match and claim nothing
        uint32_t pending = __ballot_sync(0xFFFFFFFFu, mk.x != EMPTY && (mk.x >> 16) <= HAMMING_DIST_THR_LOW);
        while (pending) {
            const bool mine = (pending >> lane) & 1u;
            BfDecision d = {EMPTY, EMPTY, true, false};
            if (mine) d = bf_decide(row, k, cl, lowe_ratio);
            const bool rescan = mine && !d.exact;
            const int first = __ffs(pending) - 1;
            if (__shfl_sync(0xFFFFFFFFu, (int)rescan, first)) {
                uint32_t b, s;
                bf_scan_unclaimed(desc_1, n1, desc_2, base + first, cl, lane, 32, b, s);
                bf_warp_merge(b, s);
                if (lane == first) {
                    d.best = b;
                    d.second = s;
                    d.accept = bf_accept(b, s, lowe_ratio);
                }
            }
            const bool accept = mine && d.accept && (lane == first || !rescan);
            const uint32_t i1 = d.best & 0xFFFFu;
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
                if (accept) s_first[i1] = EMPTY;
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
                cl.claim(i1, row);
                match_2_in_1[i1] = (int32_t)row;
            }
            num += __popc(__ballot_sync(0xFFFFFFFFu, go));
            pending &= ~commit;
            __syncwarp();
        }
    }
    return num;
}

constexpr int BF_THREADS = 1024;

// Parallel rounds first (see k_lm_resolve in match_proj.cu for the argument): every unresolved row claims the unclaimed keypoints
// among its four kept keys (atomicMin of the row number); a row whose unclaimed keys all carry its own number is decided exactly as
// the sequential loop would and is committed; a row with fewer than two such keys might have to rescan the whole frame and blocks
// the rows behind it until it is the first unresolved row, then the block rescans it together. Contended inputs fall through to
// the serial loop. topk[row].x = EMPTY marks a resolved row.
__global__ void __launch_bounds__(BF_THREADS) k_bf_resolve(const uint8_t* __restrict__ desc_1, uint32_t n1, const uint8_t* __restrict__ desc_2,
                                                           uint32_t n2, uint4* topk, float lowe_ratio, int first_in_smem,
                                                           int32_t* __restrict__ match_2_in_1, uint32_t* __restrict__ n_matches) {
    extern __shared__ uint32_t s_mem[];
    __shared__ uint32_t s_barrier, s_minrow, s_done, s_count, s_rescan, s_pb[BF_THREADS / 32], s_ps[BF_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // versioned claim table + per-round table (n1 words each) while they fit, else a bitmap (serial loop only)
    const ClaimSet cl = {s_mem, first_in_smem != 0};
    uint32_t* s_first = cl.versioned ? s_mem + n1 : nullptr;
    if (cl.versioned)
        for (uint32_t i = tid; i < n1; i += BF_THREADS) s_mem[i] = s_first[i] = EMPTY;
    else
        for (uint32_t i = tid; i < (n1 + 31) / 32; i += BF_THREADS) s_mem[i] = 0;
    for (uint32_t i = tid; i < n1; i += BF_THREADS) match_2_in_1[i] = -1;
    if (tid == 0) {
        s_barrier = s_minrow = s_rescan = EMPTY;
        s_done = s_count = 0;
    }
    __syncthreads();
    const bool parallel = s_first != nullptr && n2 <= 32u * BF_THREADS;
    if (parallel) {
        uint32_t todo = 0;  // bit j: row tid + j * BF_THREADS is unresolved
        for (uint32_t j = 0; tid + j * BF_THREADS < n2; ++j) {
            const uint32_t k0 = topk[tid + j * BF_THREADS].x;
            if (k0 != EMPTY && (k0 >> 16) <= HAMMING_DIST_THR_LOW) todo |= 1u << j;
            else reinterpret_cast<uint32_t*>(topk)[4 * (size_t)(tid + j * BF_THREADS)] = EMPTY;
        }
        int idle_rounds = 0;
        for (;;) {
            const int active = __syncthreads_count(todo != 0);
            if (active == 0) break;
            // A: claims on the keys a row could take
            if (todo) atomicMin(&s_minrow, (uint32_t)tid + (uint32_t)(__ffs(todo) - 1) * BF_THREADS);
            for (uint32_t m = todo; m; m &= m - 1) {
                const uint32_t r = tid + (uint32_t)(__ffs(m) - 1) * BF_THREADS;
                const uint4 mk = topk[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
                for (int q = 0; q < K; ++q) {
                    if (k[q] == EMPTY || (k[q] >> 16) > HAMMING_DIST_THR_LOW) continue;
                    const uint32_t c = k[q] & 0xFFFFu;
                    if (!cl.claimed(c, r)) atomicMin(&s_first[c], r);
                }
            }
            __syncthreads();
            // B: the first row that may have to rescan the frame
            for (uint32_t m = todo; m; m &= m - 1) {
                const uint32_t r = tid + (uint32_t)(__ffs(m) - 1) * BF_THREADS;
                const uint4 mk = topk[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
                if (bf_eval(r, k, cl, s_first, lowe_ratio).may_rescan) {
                    atomicMin(&s_barrier, r);
                    break;
                }
            }
            __syncthreads();
            // C: decide against the bitmap of the round's start (its updates wait for the barrier)
            const uint32_t barrier = s_barrier, minrow = s_minrow;
            uint32_t done = 0, took = 0;
            for (uint32_t m = todo; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const uint32_t r = tid + (uint32_t)j * BF_THREADS;
                if (r > barrier) break;
                const uint4 mk = topk[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
                const BfEval e = bf_eval(r, k, cl, s_first, lowe_ratio);
                if (!e.clear) {
                    if (r == minrow) s_rescan = r;  // the block rescans it below
                    continue;
                }
                if (e.accept) {
                    match_2_in_1[e.best & 0xFFFFu] = (int32_t)r;
                    took |= 1u << j;
                }
                done |= 1u << j;
            }
            __syncthreads();
            // the committed rows claim their keypoints and retire; every row withdraws this round's claims
            for (uint32_t m = todo; m; m &= m - 1) {
                const int j = __ffs(m) - 1;
                const uint32_t r = tid + (uint32_t)j * BF_THREADS;
                const uint4 mk = topk[r];
                const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
                if ((took >> j) & 1u) {
                    // the first unclaimed key is the one that was accepted (nothing has changed for this row's keys)
#pragma unroll
                    for (int q = 0; q < K; ++q) {
                        const uint32_t c = k[q] & 0xFFFFu;
                        if (k[q] != EMPTY && !cl.claimed(c, r)) {
                            cl.claim(c, r);
                            break;
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < K; ++q)
                    if (k[q] != EMPTY) s_first[k[q] & 0xFFFFu] = EMPTY;
                if ((done >> j) & 1u) reinterpret_cast<uint32_t*>(topk)[4 * (size_t)r] = EMPTY;
            }
            if (took) atomicAdd(&s_count, (uint32_t)__popc(took));
            if (done) atomicAdd(&s_done, (uint32_t)__popc(done));
            __syncthreads();
            const uint32_t rr = s_rescan;
            if (rr != EMPTY) {  // block-wide exact rescan of the first unresolved row (every earlier row is resolved)
                uint32_t b, s;
                bf_scan_unclaimed(desc_1, n1, desc_2, rr, cl, tid, BF_THREADS, b, s);
                bf_warp_merge(b, s);
                if (lane == 0) {
                    s_pb[warp] = b;
                    s_ps[warp] = s;
                }
                __syncthreads();
                if (warp == 0) {
                    b = s_pb[lane];
                    s = s_ps[lane];
                    bf_warp_merge(b, s);
                    if (lane == 0) {
                        if (bf_accept(b, s, lowe_ratio)) {
                            const uint32_t i1 = b & 0xFFFFu;
                            cl.claim(i1, rr);
                            match_2_in_1[i1] = (int32_t)rr;
                            ++s_count;
                        }
                        ++s_done;
                    }
                }
                if (rr % BF_THREADS == (uint32_t)tid) {  // the owner withdraws the row's claims and retires it
                    const uint4 mk = topk[rr];
                    const uint32_t k[K] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
                    for (int q = 0; q < K; ++q)
                        if (k[q] != EMPTY) s_first[k[q] & 0xFFFFu] = EMPTY;
                    reinterpret_cast<uint32_t*>(topk)[4 * (size_t)rr] = EMPTY;
                    done |= 1u << (rr / BF_THREADS);
                }
                __syncthreads();
            }
            todo &= ~done;
            // D: withdraw the remaining claims, re-arm
            for (uint32_t m = todo; m; m &= m - 1) {
                const uint4 mk = topk[tid + (uint32_t)(__ffs(m) - 1) * BF_THREADS];
                if (mk.x != EMPTY) s_first[mk.x & 0xFFFFu] = EMPTY;
                if (mk.y != EMPTY) s_first[mk.y & 0xFFFFu] = EMPTY;
                if (mk.z != EMPTY) s_first[mk.z & 0xFFFFu] = EMPTY;
                if (mk.w != EMPTY) s_first[mk.w & 0xFFFFu] = EMPTY;
            }
            const uint32_t n_done = s_done;
            __syncthreads();
            if (tid == 0) {
                s_barrier = s_minrow = s_rescan = EMPTY;
                s_done = 0;
            }
            idle_rounds = (n_done < 8 && active >= 64) ? idle_rounds + 1 : 0;
            if (idle_rounds >= 4) break;
        }
    }
    __syncthreads();
    if (tid == 0) s_minrow = EMPTY;
    __syncthreads();
    for (uint32_t r = tid; r < n2; r += BF_THREADS) {
        const uint32_t k0 = topk[r].x;
        if (k0 != EMPTY && (k0 >> 16) <= HAMMING_DIST_THR_LOW) {
            atomicMin(&s_minrow, r);
            break;
        }
    }
    __syncthreads();
    if (warp == 0) {
        uint32_t num = s_count;
        if (s_minrow != EMPTY)
            num += bf_resolve_serial(desc_1, n1, desc_2, n2, topk, lowe_ratio, cl, s_first, match_2_in_1, lane, s_minrow & ~31u);
        if (lane == 0) *n_matches = num;
    }
}

cudaError_t launch_bf_resolve(const uint8_t* d1, uint32_t n1, const uint8_t* d2, uint32_t n2, uint4* topk, float lowe_ratio,
                              int32_t* match_2_in_1, uint32_t* n_matches, cudaStream_t st) {
    const int tables = n1 <= 25000;  // 2 x 4 B per frame keypoint of shared memory
    const size_t smem = tables ? 8 * (size_t)n1 : ((n1 + 31) / 32) * 4;
    if (smem > 48 * 1024) {
        const cudaError_t e = cudaFuncSetAttribute(k_bf_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_bf_resolve<<<1, BF_THREADS, smem, st>>>(d1, n1, d2, n2, topk, lowe_ratio, tables, match_2_in_1, n_matches);
    return cudaGetLastError();
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
    ArenaBump ar;  // inputs | work | outputs
    const size_t o_d1 = ar.take(32 * (size_t)n1), o_d2 = ar.take(32 * (size_t)n2), o_v = ar.take(n2), in_end = ar.off;
    const size_t o_top = ar.take(16 * (size_t)n2), o_m = ar.take(4 * (size_t)n1), o_n = ar.take(4);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    int rc = LPSLAM_OK;
    cudaError_t e;
    do {
        cs.put(o_d1, desc_1, 32 * (size_t)n1);
        cs.put(o_d2, desc_2, 32 * (size_t)n2);
        if (valid_2) cs.put(o_v, valid_2, n2);
        if ((e = cs.upload(0, valid_2 ? in_end : o_v)) != cudaSuccess) break;
        k_bf_topk<<<(n2 + 7) / 8, 256, 0, cs.st>>>(dev + o_d1, n1, dev + o_d2, n2, valid_2 ? dev + o_v : nullptr, (uint4*)(dev + o_top));
        if ((e = launch_bf_resolve(dev + o_d1, n1, dev + o_d2, n2, (uint4*)(dev + o_top), lowe_ratio, (int32_t*)(dev + o_m),
                                   (uint32_t*)(dev + o_n), cs.st)) != cudaSuccess)
            break;
        if ((e = cs.download(o_m, ar.off)) != cudaSuccess) break;
        memcpy(match_2_in_1, cs.out<int32_t>(o_m), 4 * (size_t)n1);
        *n_matches = *cs.out<uint32_t>(o_n);
    } while (false);
    if (e != cudaSuccess) rc = cuda_fail(e, "lpslam_match_bruteforce", __FILE__, __LINE__);
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
        cudaError_t launch_err = cudaSuccess;
        for (uint32_t f = 1; f < nf; ++f) {
            cudaStream_t st = lanes[f % 5];
            // side 1 = the current frame f (its keypoints are claimed), side 2 = the previous frame f - 1 (every row valid)
            const uint32_t n1 = off[f + 1] - off[f], n2 = off[f] - off[f - 1];
            if (n1 == 0 || n2 == 0) continue;
            const uint8_t* d1 = c->desc.desc + 32 * (size_t)off[f];
            const uint8_t* d2 = c->desc.desc + 32 * (size_t)off[f - 1];
            uint4* top = (uint4*)(dev + o_top + top_stride * f);
            k_bf_topk<<<(n2 + 7) / 8, 256, 0, st>>>(d1, n1, d2, n2, nullptr, top);
            if ((launch_err = launch_bf_resolve(d1, n1, d2, n2, top, lowe_ratio, (int32_t*)(dev + o_m) + off[f], (uint32_t*)(dev + o_n) + f,
                                                st)) != cudaSuccess)
                break;
        }
        for (int k = 1; k < 5; ++k) {
            cudaEventRecord(joins[k], lanes[k]);
            cudaStreamWaitEvent(c->stream, joins[k], 0);
        }
        if ((e = launch_err) != cudaSuccess || (e = cudaGetLastError()) != cudaSuccess) break;
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
