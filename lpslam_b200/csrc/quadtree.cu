// K3 — quadtree keypoint distribution. Replaces distribute_keypoints_via_tree / initialize_nodes / assign_child_nodes /
// find_keypoints_with_max_response (feature/orb_extractor.cc:401-593) and orb_extractor_node::divide_node
// (feature/orb_extractor_node.cc:6-50). One block per (level, frame).
//
// The reference walks a std::list of nodes that each own a vector of keypoints. Here the geometry is factored out:
//   * a node's four children are fixed by its rectangle alone (half extents round up), so the quadrant a keypoint falls
//     into at every depth is a pure function of its coordinates and its initial node. Each candidate gets a PATH KEY
//     (initial-node index, then 2 bits per depth: bit0 = right, bit1 = bottom). After one radix sort by that key every
//     node of every depth is a contiguous range of the sorted array, and "divide" is three binary searches.
//   * the list discipline (children pushed to the FRONT in order 0..3, parent erased in place, single-keypoint children
//     not flagged leaf and therefore re-divided, the pool of dividable nodes sorted by (count, address) descending) is
//     replayed on index arrays with prefix sums: a breadth pass maps node i / child j to position T-1-(P_i + j).
//   * the address tie-break of the pool sort (orb_extractor.cc:408,455) is pinned as "later created first", which for
//     nodes created in one pass is "nearer the list front first": a stable sort by count over list order.
//   * the winner of a node is its first keypoint with the strictly greatest response in candidate order
//     (orb_extractor.cc:571-593) = arg max of (score, -candidate index).
#include "common.cuh"

namespace lps {

namespace {

#ifndef LPS_QT_THREADS
#define LPS_QT_THREADS 512
#endif
constexpr int QT = LPS_QT_THREADS;  // threads per block (two resident blocks per SM at 512)
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int PART_BITS = 14;  // leading key bits of the one-pass ordering (2^14 shared-memory bins)
constexpr int BIN_WORDS = (1 << PART_BITS) + (1 << (PART_BITS - 5));  // padded: index i lives at i + i / 32
constexpr int SORT_CH = 8;  // sort rounds whose loads are issued together
constexpr int POOL_CACHE = 4 * 1024;  // pool counts cached in shared memory (words), behind the bins
constexpr int F_LEAF = 1 << 8, F_POOL = 1 << 9, DEPTH_MASK = 0xFF;

struct Scan {
    unsigned long long* warp_tot;  // 32 entries of shared memory
    // exclusive block scan of a 64-bit value (three 21-bit fields are packed by the callers); returns the total to all
    __device__ __forceinline__ unsigned long long excl(unsigned long long v, unsigned long long& total) {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        unsigned long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) warp_tot[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const unsigned long long w = lane < (int)(blockDim.x >> 5) ? warp_tot[lane] : 0ull;
            unsigned long long wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, wi, o);
                if (lane >= o) wi += t;
            }
            warp_tot[lane] = wi - w;
            if (lane == 31) warp_tot[32] = wi;
        }
        __syncthreads();
        const unsigned long long res = warp_tot[warp] + inc - v;
        total = warp_tot[32];
        __syncthreads();
        return res;
    }
};

__device__ __forceinline__ int lower_digit(const uint32_t* __restrict__ keys, int lo, int hi, int shift, unsigned q) {
    // first index in [lo, hi) whose 2-bit digit at `shift` is >= q (the range shares all higher bits)
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (((keys[mid] >> shift) & 3u) < q) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ int pad_idx(int lin) { return lin + (lin >> 5); }

// Dsorted = number of leading digits the key array is ordered by (D after the complete sort); a split below that depth
// cannot be answered from a prefix-ordered array: *deeper is raised and the block redoes the level with the complete sort.
// bins != nullptr (prefix ordering): bins[pad_idx(i)] = end of bin i = start of bin i + 1, so the three inner boundaries
// of a node are three shared-memory reads at its children's first bins instead of three binary searches.
__device__ __forceinline__ void split_node(const uint32_t* __restrict__ keys, int lo, int hi, int d, int D, int Dsorted,
                                           const uint32_t* bins, int* deeper, int (&b)[5]) {
    b[0] = lo;
    b[4] = hi;
    if (hi - lo <= 1 || d >= D) {  // a single keypoint always yields exactly one non-empty child
        b[1] = b[2] = b[3] = hi;
        return;
    }
    if (d >= Dsorted) {
        *deeper = 1;
        b[1] = b[2] = b[3] = hi;
        return;
    }
    if (bins) {
        const int sh = 2 * (Dsorted - d - 1);
        const uint32_t child0 = ((keys[lo] >> (2 * (D - d))) << 2) << sh;  // first bin of child 0
        b[1] = (int)bins[pad_idx((int)(child0 + (1u << sh)) - 1)];
        b[2] = (int)bins[pad_idx((int)(child0 + (2u << sh)) - 1)];
        b[3] = (int)bins[pad_idx((int)(child0 + (3u << sh)) - 1)];
        return;
    }
    const int shift = 2 * (D - d - 1);
    b[2] = lower_digit(keys, lo, hi, shift, 2);
    b[1] = lower_digit(keys, lo, b[2], shift, 1);
    b[3] = lower_digit(keys, b[2], hi, shift, 3);
}

// lanes holding the same 9-bit value (8 digit bits + the "inactive" bit), by votes: MATCH.ANY costs over a thousand
// cycles per call here (measured through the phase clocks), nine VOTEs and nine LOP3s do not
__device__ __forceinline__ unsigned match_digit(unsigned d) {
    unsigned peers = 0xFFFFFFFFu;
#pragma unroll
    for (int b = 0; b <= RADIX_BITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const unsigned v = __ballot_sync(0xFFFFFFFFu, bit);
        peers &= bit ? v : ~v;
    }
    return peers;
}

#ifdef LPS_QT_PROFILE  // phase clocks of block (frame 0, level 0): -DLPS_QT_PROFILE, read with lpslam_debug_qt_profile
__device__ unsigned long long g_qt_prof[16];
#define QT_MARK(i) do { if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) g_qt_prof[i] = clock64(); } while (0)
#else
#define QT_MARK(i) do { } while (0)
#endif

__global__ void __launch_bounds__(QT, 2048 / QT > 2 ? 2 : 2048 / QT) k_quadtree(TreeArgs a) {
    extern __shared__ uint32_t s_dyn[];  // RADIX * QT counters (padded) during the sort, pool counts afterwards
    __shared__ unsigned long long s_scan[33];
    __shared__ int s_i[4];

    // grid = (frame, level): the blocks of the fullest level (0) are scheduled first, the light ones fill the tail
    const int level = blockIdx.y, frame = blockIdx.x + a.frame0, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const LevelGeom lv = a.lv[level];
    Scan scan{s_scan};

    const uint32_t* cell_counts = a.cell_counts + (size_t)frame * a.num_cells + lv.cell_begin;
    const uint32_t* cell_slots = a.cell_slots + ((size_t)frame * a.num_cells + lv.cell_begin) * kCellSlots;
    uint32_t* cell_base = a.cell_base + (size_t)frame * a.cell_stride + lv.cell_begin;
    uint32_t* cand_xy = a.cand_xy + (size_t)frame * a.cand_stride + lv.cand_begin;
    uint8_t* cand_sc = a.cand_sc + (size_t)frame * a.cand_stride + lv.cand_begin;
    uint32_t* keys[2] = {a.keys[0] + (size_t)frame * a.cand_stride + lv.cand_begin,
                         a.keys[1] + (size_t)frame * a.cand_stride + lv.cand_begin};
    uint32_t* vals[2] = {a.vals[0] + (size_t)frame * a.cand_stride + lv.cand_begin,
                         a.vals[1] + (size_t)frame * a.cand_stride + lv.cand_begin};
    const size_t nofs = (size_t)frame * a.node_stride + lv.node_begin;
    int* n_lo[2] = {a.node[0][0] + nofs, a.node[1][0] + nofs};
    int* n_hi[2] = {a.node[0][1] + nofs, a.node[1][1] + nofs};
    int* n_meta[2] = {a.node[0][2] + nofs, a.node[1][2] + nofs};
    int* t_b1 = a.tmp[0] + nofs;
    int* t_b2 = a.tmp[1] + nofs;
    int* t_b3 = a.tmp[2] + nofs;
    int* t_p = a.tmp[3] + nofs;   // prefix of non-empty children
    int* t_l = a.tmp[4] + nofs;   // prefix of carried-over nodes
    int* t_pl = a.tmp[5] + nofs;  // pool: node index in list order
    int* t_srt = a.tmp[6] + nofs; // pool sorted: node index
    int* t_proc = a.tmp[7] + nofs;
    uint32_t* sel_xy = a.sel_xys + (size_t)frame * a.sel_stride + lv.sel_begin;
    uint8_t* sel_sc = a.sel_sc + (size_t)frame * a.sel_stride + lv.sel_begin;
    uint32_t* sel_count = a.sel_count + (size_t)frame * a.num_levels + level;
    uint32_t* cand_count = a.cand_count + (size_t)frame * a.num_levels + level;

    QT_MARK(0);
    // ------------------------------------------------------------------ A. candidate list in reference order
    const int ncl = lv.cell_cols * lv.cell_rows;
    int N = 0;
    for (int base = 0; base < ncl; base += QT) {
        const int i = base + tid;
        const unsigned v = i < ncl ? cell_counts[i] : 0u;
        unsigned long long tot;
        const unsigned ex = (unsigned)scan.excl(v, tot);
        if (i < ncl) cell_base[i] = N + ex;
        N += (int)tot;
    }
    if (tid == 0) *cand_count = (uint32_t)N;
    if (N == 0) {
        if (tid == 0) *sel_count = 0;
        return;
    }
    // the one-pass ordering of phase B counts its bins while the keys are made
    const int keybits = 2 * lv.depth + lv.rbits;
    const int D = lv.depth;
    const int Dpart = min(D, max(0, (PART_BITS - lv.rbits) >> 1));  // digits covered by the one-pass ordering
    const int tbits = lv.rbits + 2 * Dpart, drop = keybits - tbits, nbins = 1 << tbits;
    bool full = lv.rbits > PART_BITS;
#ifdef LPS_QT_FORCE_FULL
    full = true;
#endif
    uint32_t* s_bin = s_dyn;  // <= 2^PART_BITS words
    if (!full)
        for (int i = tid; i < nbins + (nbins >> 5) + 1; i += QT) s_bin[i] = 0;
    __syncthreads();
    {
        int ci = warp;
        CellDesc c = a.cells[lv.cell_begin + min(ci, ncl - 1)];
        int cnt = ci < ncl ? (int)cell_counts[ci] : 0, cbase = ci < ncl ? (int)cell_base[ci] : 0;
        while (ci < ncl) {
            // the next cell's header is on its way while this cell's slots are read
            const int cn = ci + QT / 32;
            const CellDesc c_next = a.cells[lv.cell_begin + min(cn, ncl - 1)];
            const int cnt_next = cn < ncl ? (int)cell_counts[cn] : 0, cbase_next = cn < ncl ? (int)cell_base[cn] : 0;
            const uint32_t* slots = cell_slots + (size_t)ci * kCellSlots;
            for (int k0 = lane; k0 < cnt; k0 += 32 * 4) {
                uint32_t sl[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) sl[j] = k0 + 32 * j < cnt ? slots[k0 + 32 * j] : 0u;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int k = k0 + 32 * j;
                    if (k < cnt) {
                        // window-local + 3 (ring margin) + 64 * cell index, relative to the 19-px border (orb_extractor.cc:362-363)
                        const int x = (int)(sl[j] & 255u) + 3 + kCell * c.j;
                        const int y = (int)((sl[j] >> 8) & 255u) + 3 + kCell * c.i;
                        const uint32_t key = __ldg(lv.qxk + x) + __ldg(lv.qyk + y);
                        cand_xy[cbase + k] = (uint32_t)x | ((uint32_t)y << 16);
                        cand_sc[cbase + k] = (uint8_t)(sl[j] >> 16);
                        keys[0][cbase + k] = key;
                        if (!full) atomicAdd(&s_bin[pad_idx((int)(key >> drop))], 1u);
                    }
                }
            }
            ci = cn;
            c = c_next;
            cnt = cnt_next;
            cbase = cbase_next;
        }
    }
    __syncthreads();
    QT_MARK(1);

    // ------------------------------------------------------------------ B. order the candidates by path key
    // First attempt: ONE counting pass over the leading PART_BITS of the key (shared-memory histogram, prefix, scatter by
    // atomics). The tree only ever asks for the digit of a node's depth, the winner of a node does not depend on the order
    // inside it, and with a few hundred nodes wanted out of thousands of candidates the tree stays far above the depth
    // those bits cover. If a split below that depth is asked for after all, the level is redone with the complete sort:
    // LSD radix, 8 bits per pass, stable. There warp w owns the contiguous segment [w * seg, (w + 1) * seg) and walks it 32
    // keys at a time with coalesced loads; inside a round the rank of a key among the equal digits of its warp comes from
    // votes, so order is preserved: earlier segments, earlier rounds, lower lanes first. Counters live in shared memory as
    // [digit][warp] (pitch 33: leaders of one warp hit different banks).
    int m = 0;   // list length, identical in every thread
    int nb = 0;  // node buffer holding the current list
    bool filled = false;
    const int quota = (int)lv.quota;
    const uint32_t* __restrict__ skey;
    const uint32_t* __restrict__ sval;
    int Dsorted;
  while (true) {  // at most twice: prefix ordering, then (rarely) the complete sort
    int cur = 0;
    if (!full) {
        {   // exclusive prefix over the bins: thread t owns a contiguous run
            const int per = (nbins + QT - 1) / QT, b0 = min(nbins, tid * per), b1 = min(nbins, b0 + per);
            unsigned sum = 0;
            for (int i = b0; i < b1; ++i) sum += s_bin[pad_idx(i)];
            unsigned long long tot;
            unsigned run = (unsigned)scan.excl(sum, tot);
            for (int i = b0; i < b1; ++i) {
                const unsigned c = s_bin[pad_idx(i)];
                s_bin[pad_idx(i)] = run;
                run += c;
            }
        }
        __syncthreads();
        for (int k0 = tid; k0 < N; k0 += QT * SORT_CH) {
            uint32_t kk[SORT_CH], sc[SORT_CH];
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j) {
                kk[j] = k0 + j * QT < N ? keys[0][k0 + j * QT] : 0u;
                sc[j] = k0 + j * QT < N ? cand_sc[k0 + j * QT] : 0u;
            }
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j)
                if (k0 + j * QT < N) {
                    const unsigned pos = atomicAdd(&s_bin[pad_idx((int)(kk[j] >> drop))], 1u);
                    keys[1][pos] = kk[j];
                    vals[1][pos] = (sc[j] << 24) | (0xFFFFFFu - (uint32_t)(k0 + j * QT));  // what phase F compares
                }
        }
        __syncthreads();
        cur = 1;
        Dsorted = Dpart;
    }
    else {
    const int seg = ((N + QT - 1) / QT) * 32;
    const int seg_lo = min(N, warp * seg), seg_hi = min(N, (warp + 1) * seg);
    uint32_t* s_cntr = s_dyn;  // RADIX * 33 words
    for (int k = tid; k < N; k += QT) vals[0][k] = (uint32_t)k;
    __syncthreads();
    for (int shift = 0; shift < keybits; shift += RADIX_BITS) {
        const uint32_t* __restrict__ kin = keys[cur];
        const uint32_t* __restrict__ vin = vals[cur];
        uint32_t* __restrict__ kout = keys[cur ^ 1];
        uint32_t* __restrict__ vout = vals[cur ^ 1];
        for (int i = tid; i < RADIX * 33; i += QT) s_cntr[i] = 0;
        __syncthreads();
        for (int base = seg_lo; base < seg_hi; base += 32 * SORT_CH) {
            uint32_t kk[SORT_CH];  // SORT_CH rounds of loads in flight, then ranked from registers
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j) {
                const int k = base + 32 * j + lane;
                kk[j] = k < seg_hi ? kin[k] : 0u;
            }
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j) {
                if (base + 32 * j >= seg_hi) break;  // warp-uniform
                const bool ok = base + 32 * j + lane < seg_hi;
                const unsigned d = ok ? ((kk[j] >> shift) & (RADIX - 1)) : RADIX;  // inactive lanes form their own group
                const unsigned peers = match_digit(d);
                if (ok && (peers & ((1u << lane) - 1)) == 0) s_cntr[d * 33 + warp] += __popc(peers);
                __syncwarp();
            }
        }
        __syncthreads();
        // exclusive scan in (digit major, warp minor) order: thread t owns digit t / 4, warps 8 (t % 4) .. + 8
        {
            const int d = tid / (QT / RADIX), w0 = (tid % (QT / RADIX)) * 8;  // (QT / 32) warps = (QT / RADIX) threads x 8
            unsigned loc[8], sum = 0;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                loc[e] = s_cntr[d * 33 + w0 + e];
                sum += loc[e];
            }
            unsigned long long tot;
            unsigned run = (unsigned)scan.excl(sum, tot);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                s_cntr[d * 33 + w0 + e] = run;
                run += loc[e];
            }
        }
        __syncthreads();
        for (int base = seg_lo; base < seg_hi; base += 32 * SORT_CH) {
            uint32_t kk[SORT_CH], vv[SORT_CH];
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j) {
                const int k = base + 32 * j + lane;
                kk[j] = k < seg_hi ? kin[k] : 0u;
                vv[j] = k < seg_hi ? vin[k] : 0u;
            }
#pragma unroll
            for (int j = 0; j < SORT_CH; ++j) {
                if (base + 32 * j >= seg_hi) break;  // warp-uniform
                const bool ok = base + 32 * j + lane < seg_hi;
                const unsigned d = ok ? ((kk[j] >> shift) & (RADIX - 1)) : RADIX;
                const unsigned peers = match_digit(d);
                const unsigned before = __popc(peers & ((1u << lane) - 1));
                unsigned pos = 0;
                if (ok) pos = s_cntr[d * 33 + warp] + before;
                __syncwarp();
                if (ok) {
                    if (before == 0) s_cntr[d * 33 + warp] += __popc(peers);
                    kout[pos] = kk[j];
                    vout[pos] = vv[j];
                }
                __syncwarp();
            }
        }
        __syncthreads();
        cur ^= 1;
    }
    for (int k = tid; k < N; k += QT) {  // candidate index -> the word phase F compares
        const uint32_t o = vals[cur][k];
        vals[cur][k] = ((uint32_t)cand_sc[o] << 24) | (0xFFFFFFu - o);
    }
    __syncthreads();
    Dsorted = D;
    }
    skey = keys[cur];
    sval = vals[cur];
    if (tid == 0) s_i[2] = 0;
    m = 0;
    nb = 0;
    filled = false;
    QT_MARK(2);

    // ------------------------------------------------------------------ C. initial nodes (orb_extractor.cc:480-553)
    {
        const int ninit = lv.nx * lv.ny;
        for (int base = 0; base < ninit; base += QT) {
            const int r = base + tid;
            int lo = 0, hi = 0;
            if (r < ninit && !full) {  // first and last bin of the initial node
                const int first = r << (2 * Dpart), last = ((r + 1) << (2 * Dpart)) - 1;
                lo = first ? (int)s_bin[pad_idx(first - 1)] : 0;
                hi = (int)s_bin[pad_idx(last)];
            }
            else if (r < ninit) {
                int l = 0, h = N;
                const int sh = 2 * D;
                while (l < h) { const int mid = (l + h) >> 1; if ((skey[mid] >> sh) < (unsigned)r) l = mid + 1; else h = mid; }
                lo = l;
                h = N;
                while (l < h) { const int mid = (l + h) >> 1; if ((skey[mid] >> sh) <= (unsigned)r) l = mid + 1; else h = mid; }
                hi = l;
            }
            unsigned long long tot;
            const int pos = m + (int)scan.excl(hi > lo ? 1u : 0u, tot);
            if (hi > lo) {
                n_lo[nb][pos] = lo;
                n_hi[nb][pos] = hi;
                n_meta[nb][pos] = (hi - lo == 1) ? F_LEAF : 0;
            }
            m += (int)tot;
        }
        __syncthreads();
    }

    QT_MARK(3);

    // ------------------------------------------------------------------ D. breadth passes (orb_extractor.cc:414-445)
    while (true) {
        const int prev = m;
        int T = 0, NL = 0, POOL = 0;
        for (int base = 0; base < m; base += QT) {
            const int i = base + tid;
            unsigned long long v = 0;
            if (i < m) {
                const int meta = n_meta[nb][i];
                if (meta & F_LEAF) {
                    v = 1ull << 21;
                }
                else {
                    int b[5];
                    split_node(skey, n_lo[nb][i], n_hi[nb][i], meta & DEPTH_MASK, D, Dsorted, full ? nullptr : s_bin, &s_i[2], b);
                    t_b1[i] = b[1];
                    t_b2[i] = b[2];
                    t_b3[i] = b[3];
                    unsigned k = 0, pc = 0;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        k += (b[q + 1] - b[q]) > 0;
                        pc += (b[q + 1] - b[q]) > 1;
                    }
                    v = (unsigned long long)k | ((unsigned long long)pc << 42);
                }
            }
            unsigned long long tot;
            const unsigned long long ex = scan.excl(v, tot);
            if (i < m) {
                t_p[i] = T + (int)(ex & 0x1FFFFF);
                t_l[i] = NL + (int)((ex >> 21) & 0x1FFFFF);
            }
            T += (int)(tot & 0x1FFFFF);
            NL += (int)((tot >> 21) & 0x1FFFFF);
            POOL += (int)(tot >> 42);
        }
        const int ob = nb ^ 1;
        for (int i = tid; i < m; i += QT) {
            const int meta = n_meta[nb][i];
            const int lo = n_lo[nb][i], hi = n_hi[nb][i];
            if (meta & F_LEAF) {
                const int dst = T + t_l[i];
                n_lo[ob][dst] = lo;
                n_hi[ob][dst] = hi;
                n_meta[ob][dst] = meta & ~F_POOL;
            }
            else {
                const int b[5] = {lo, t_b1[i], t_b2[i], t_b3[i], hi};
                int f = t_p[i];
                const int dchild = min((meta & DEPTH_MASK) + 1, DEPTH_MASK);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = b[q + 1] - b[q];
                    if (c > 0) {
                        const int dst = T - 1 - f;
                        ++f;
                        n_lo[ob][dst] = b[q];
                        n_hi[ob][dst] = b[q + 1];
                        n_meta[ob][dst] = dchild | (c > 1 ? F_POOL : 0);
                    }
                }
            }
        }
        __syncthreads();
        nb = ob;
        m = T + NL;
        if (quota <= m || m == prev) {
            filled = true;
            break;
        }
        if (quota < m + POOL) break;
    }

    QT_MARK(4);
    // ------------------------------------------------------------------ E. fullest-first splitting (orb_extractor.cc:447-475)
    int* s_cnt = reinterpret_cast<int*>(s_dyn) + BIN_WORDS;  // pool counts cached in shared memory when they fit
    const int s_cnt_cap = POOL_CACHE;
    while (!filled) {
        const int prev = m;
        // pool = nodes created by the previous pass / round holding more than one keypoint, in list order
        int np = 0;
        for (int base = 0; base < m; base += QT) {
            const int i = base + tid;
            const bool in_pool = i < m && (n_meta[nb][i] & F_POOL);
            unsigned long long tot;
            const int pos = np + (int)scan.excl(in_pool ? 1u : 0u, tot);
            if (in_pool) {
                t_pl[pos] = i;
                if (pos < s_cnt_cap) s_cnt[pos] = n_hi[nb][i] - n_lo[nb][i];
            }
            if (i < m) t_proc[i] = -1;
            np += (int)tot;
        }
        __syncthreads();
        if (np == 0) break;  // nothing dividable: the list cannot grow (size == prev_size)
        // stable rank sort by count descending (ties: list order = later-created first)
        for (int x = tid; x < np; x += QT) {
            const int nx_i = t_pl[x];
            const int cx = n_hi[nb][nx_i] - n_lo[nb][nx_i];
            int rank = 0;
            for (int y = 0; y < np; ++y) {
                int cy;
                if (y < s_cnt_cap) cy = s_cnt[y];
                else { const int ny_i = t_pl[y]; cy = n_hi[nb][ny_i] - n_lo[nb][ny_i]; }
                rank += (cy > cx) || (cy == cx && y < x);
            }
            t_srt[rank] = nx_i;
        }
        __syncthreads();
        // split every pool node (in sorted order), find where the list reaches the quota
        if (tid == 0) { s_i[0] = 0x7FFFFFFF; }
        __syncthreads();
        int K = 0;
        for (int base = 0; base < np; base += QT) {
            const int s = base + tid;
            unsigned k = 0;
            if (s < np) {
                const int i = t_srt[s];
                int b[5];
                split_node(skey, n_lo[nb][i], n_hi[nb][i], n_meta[nb][i] & DEPTH_MASK, D, Dsorted, full ? nullptr : s_bin, &s_i[2], b);
                t_b1[s] = b[1];
                t_b2[s] = b[2];
                t_b3[s] = b[3];
#pragma unroll
                for (int q = 0; q < 4; ++q) k += (b[q + 1] - b[q]) > 0;
            }
            unsigned long long tot;
            const int exk = (int)scan.excl(k, tot);
            if (s < np) {
                t_p[s] = K + exk;
                // list size after this node has been replaced by its children
                const int size_after = m + (K + exk + (int)k) - (s + 1);
                if (quota <= size_after) atomicMin(&s_i[0], s);
            }
            K += (int)tot;
        }
        __syncthreads();
        int sstar = s_i[0];
        if (sstar == 0x7FFFFFFF) sstar = np - 1;
        else filled = true;
        __syncthreads();
        if (tid == 0) {
            const int i = t_srt[sstar];
            int b[5] = {n_lo[nb][i], t_b1[sstar], t_b2[sstar], t_b3[sstar], n_hi[nb][i]};
            int k = 0;
            for (int q = 0; q < 4; ++q) k += (b[q + 1] - b[q]) > 0;
            s_i[1] = t_p[sstar] + k;
        }
        for (int s = tid; s <= sstar; s += QT) t_proc[t_srt[s]] = s;
        __syncthreads();
        const int Tp = s_i[1];
        const int ob = nb ^ 1;
        // children of the processed nodes, pushed to the front in processing order
        for (int s = tid; s <= sstar; s += QT) {
            const int i = t_srt[s];
            const int b[5] = {n_lo[nb][i], t_b1[s], t_b2[s], t_b3[s], n_hi[nb][i]};
            const int dchild = min((n_meta[nb][i] & DEPTH_MASK) + 1, DEPTH_MASK);
            int f = t_p[s];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = b[q + 1] - b[q];
                if (c > 0) {
                    const int dst = Tp - 1 - f;
                    ++f;
                    n_lo[ob][dst] = b[q];
                    n_hi[ob][dst] = b[q + 1];
                    n_meta[ob][dst] = dchild | (c > 1 ? F_POOL : 0);
                }
            }
        }
        // every other node keeps its relative order behind them; the pool flag is per round
        int R = 0;
        for (int base = 0; base < m; base += QT) {
            const int i = base + tid;
            const bool keep = i < m && t_proc[i] < 0;
            unsigned long long tot;
            const int pos = Tp + R + (int)scan.excl(keep ? 1u : 0u, tot);
            if (keep) {
                n_lo[ob][pos] = n_lo[nb][i];
                n_hi[ob][pos] = n_hi[nb][i];
                n_meta[ob][pos] = n_meta[nb][i] & ~F_POOL;
            }
            R += (int)tot;
        }
        __syncthreads();
        nb = ob;
        m = Tp + R;
        if (filled || quota <= m || m == prev) break;
    }

    __syncthreads();
    if (full || !s_i[2]) break;
    __syncthreads();
    full = true;  // a split below the ordered prefix was asked for: redo the level with the complete sort
#ifdef LPS_QT_PROFILE
    if (tid == 0) atomicAdd(&g_qt_prof[12], 1ull);
#endif
  }
    QT_MARK(5);
#ifdef LPS_QT_PROFILE
    if (threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) { g_qt_prof[8] = N; g_qt_prof[9] = m; g_qt_prof[10] = keybits; g_qt_prof[11] = quota; }
#endif
    // ------------------------------------------------------------------ F. winner per node, list order
    // sval holds (score << 24) | (2^24 - 1 - candidate index): the greatest word of a node's range is its first keypoint
    // with the strictly greatest response. A warp takes 32 consecutive nodes (bounds read by one load, handed round by
    // shuffles), four nodes' first rounds in flight at a time; the 32 winners' coordinates are fetched together.
    for (int g0 = 0; g0 < m; g0 += QT) {
        const int mine = g0 + warp * 32 + lane;
        const int my_lo = mine < m ? n_lo[nb][mine] : 0, my_hi = mine < m ? n_hi[nb][mine] : 0;
        const int cntn = min(32, m - (g0 + warp * 32));  // warp-uniform, may be <= 0
        uint32_t my_best = 0;
        for (int j = 0; j < cntn; j += 4) {
            uint32_t v[4];
            int lo4[4], hi4[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                lo4[u] = __shfl_sync(0xFFFFFFFFu, my_lo, (j + u) & 31);
                hi4[u] = __shfl_sync(0xFFFFFFFFu, my_hi, (j + u) & 31);
                v[u] = lo4[u] + lane < hi4[u] ? sval[lo4[u] + lane] : 0u;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                for (int k = lo4[u] + 32 + lane; k < hi4[u]; k += 32) v[u] = max(v[u], sval[k]);
                v[u] = __reduce_max_sync(0xFFFFFFFFu, v[u]);
                if (lane == j + u) my_best = v[u];
            }
        }
        if (mine < m) {
            sel_xy[mine] = cand_xy[0xFFFFFFu - (my_best & 0xFFFFFFu)];
            sel_sc[mine] = (uint8_t)(my_best >> 24);
        }
    }
    if (tid == 0) *sel_count = (uint32_t)m;
    QT_MARK(6);
}

}  // namespace

void launch_quadtree(const TreeArgs& a, const LevelGeom* lv_host, cudaStream_t st) {
    (void)lv_host;
    static_assert(BIN_WORDS >= RADIX * 33, "the LSD counters reuse the bin area");
    const size_t smem = (size_t)(BIN_WORDS + POOL_CACHE) * sizeof(uint32_t);
    cudaFuncSetAttribute(k_quadtree, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_quadtree<<<dim3(a.batch, a.num_levels), QT, smem, st>>>(a);
}

}  // namespace lps

#ifdef LPS_QT_PROFILE
extern "C" int lpslam_debug_qt_profile(unsigned long long* out16) {
    return (int)cudaMemcpyFromSymbol(out16, lps::g_qt_prof, sizeof(unsigned long long) * 16);
}
#endif
