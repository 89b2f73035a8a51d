// K11 — brute-force Hamming top-2 against a device-resident descriptor database shard (relocalisation / loop query).
// Generalises the best / second-best scan of match/robust.cc:252-274 (strict <, lowest index wins ties) to a database of
// millions of rows; a shard reports two packed keys (distance << 40 | global row index) per query and shards are merged by
// taking the two smallest keys, which is exactly what one sequential scan over the concatenated shards would keep.
//
//   k_knn_scan    thread = one query (descriptor in registers), block = 128 queries x one database chunk; chunk rows stream
//                 through shared memory in 256-row tiles and are read as warp-wide broadcasts, so there is no cross-thread
//                 reduction and the inner loop is xor / popc / add only.
//   k_knn_reduce  two smallest keys over the chunk (or shard) partials of every query.
#include "common.cuh"

#include <dlfcn.h>

#include <algorithm>

struct lpslam_knn_db {
    int device = 0;
    uint8_t* rows = nullptr;
    uint64_t cap = 0, n = 0, index_base = 0;
    uint64_t* partial = nullptr;  // [chunks][nq_cap][2]
    uint32_t nq_cap = 0, chunks_cap = 0;
    uint8_t* q_dev = nullptr;
    uint64_t* keys_dev = nullptr;
};

namespace lps {
namespace {

constexpr int QB = 128;    // queries per block
constexpr int TILE = 256;  // database rows per shared-memory tile
constexpr uint64_t EMPTY64 = 0xFFFFFFFFFFFFFFFFull;

// 256-bit Hamming distance with HALF the POPC instructions. POPC issues at 16 lanes/clk/SM on B200 (tools/pipe_rates.cu), a
// quarter of the LOP3 rate, so eight popcounts per comparison cap a GPU at 578 G comparisons/s. Seven of the eight xor
// words go through a carry-save adder tree first (4 full adders = 8 LOP3) that leaves ones / twos / fours bit planes:
//   d = popc(ones) + 2 popc(twos) + 4 popc(fours) + popc(x7)
// 4 POPC + 16 LOP3 + 3 IMAD per comparison: the popc pipe and the alu pipe are both busy for ~0.25 SM-clocks.
__device__ __forceinline__ void csa(uint32_t a, uint32_t b, uint32_t c, uint32_t& sum, uint32_t& carry) {
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(sum) : "r"(a), "r"(b), "r"(c));    // a ^ b ^ c
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(carry) : "r"(a), "r"(b), "r"(c));  // majority
}
__device__ __forceinline__ unsigned hamming_regs(const uint32_t (&q)[8], const uint4& b0, const uint4& b1) {
    const uint32_t x0 = q[0] ^ b0.x, x1 = q[1] ^ b0.y, x2 = q[2] ^ b0.z, x3 = q[3] ^ b0.w;
    const uint32_t x4 = q[4] ^ b1.x, x5 = q[5] ^ b1.y, x6 = q[6] ^ b1.z, x7 = q[7] ^ b1.w;
    uint32_t s1, c1, s2, c2, ones, c3, twos, fours;
    csa(x0, x1, x2, s1, c1);
    csa(x3, x4, x5, s2, c2);
    csa(s1, s2, x6, ones, c3);
    csa(c1, c2, c3, twos, fours);
    return __popc(ones) + __popc(x7) + 2u * __popc(twos) + 4u * __popc(fours);
}

__global__ void __launch_bounds__(QB) k_knn_scan(const uint8_t* __restrict__ db, uint64_t n_rows, uint64_t index_base,
                                                 const uint8_t* __restrict__ queries, uint32_t nq, uint32_t rows_per_chunk,
                                                 uint64_t* __restrict__ partial, uint32_t nq_stride) {
    __shared__ __align__(16) uint4 s_tile[2][TILE * 2];
    const int tid = threadIdx.x;
    const uint32_t qi = blockIdx.x * QB + tid;
    const uint64_t row0 = (uint64_t)blockIdx.y * rows_per_chunk;
    const uint64_t row1 = min(n_rows, row0 + rows_per_chunk);
    uint32_t q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (qi < nq) {
        const uint4* qp = reinterpret_cast<const uint4*>(queries + (size_t)qi * 32);
        const uint4 a = __ldg(qp), b = __ldg(qp + 1);
        q[0] = a.x; q[1] = a.y; q[2] = a.z; q[3] = a.w; q[4] = b.x; q[5] = b.y; q[6] = b.z; q[7] = b.w;
    }
    unsigned bd = 0xFFFFu, sd = 0xFFFFu;
    uint64_t bi = 0, si = 0;
    const uint4* dbv = reinterpret_cast<const uint4*>(db);
    auto load_tile = [&](int buf, uint64_t r0) {
        // TILE rows = 2 * TILE uint4, QB threads -> 4 uint4 each, coalesced
#pragma unroll
        for (int k = 0; k < 2 * TILE / QB; ++k) {
            const uint64_t v = 2 * r0 + (uint64_t)k * QB + tid;
            s_tile[buf][k * QB + tid] = (v < 2 * row1) ? __ldg(dbv + v) : make_uint4(0, 0, 0, 0);
        }
    };
    int buf = 0;
    if (row0 < row1) load_tile(0, row0);
    __syncthreads();
    for (uint64_t r0 = row0; r0 < row1; r0 += TILE) {
        if (r0 + TILE < row1) load_tile(buf ^ 1, r0 + TILE);
        const int cnt = (int)min((uint64_t)TILE, row1 - r0);
        const uint4* t = s_tile[buf];
        int r = 0;
        for (; r + 4 <= cnt; r += 4) {
            unsigned d[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) d[u] = hamming_regs(q, t[2 * (r + u)], t[2 * (r + u) + 1]);
            // after the first few tiles almost no row beats the running second best: one test per four rows
            if (min(min(d[0], d[1]), min(d[2], d[3])) < sd) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (d[u] < bd) {
                        sd = bd; si = bi;
                        bd = d[u]; bi = r0 + r + u;
                    }
                    else if (d[u] < sd) {
                        sd = d[u]; si = r0 + r + u;
                    }
                }
            }
        }
        for (; r < cnt; ++r) {
            const unsigned d = hamming_regs(q, t[2 * r], t[2 * r + 1]);
            if (d < bd) {
                sd = bd; si = bi;
                bd = d; bi = r0 + r;
            }
            else if (d < sd) {
                sd = d; si = r0 + r;
            }
        }
        __syncthreads();
        buf ^= 1;
    }
    if (qi < nq) {
        uint64_t* out = partial + ((size_t)blockIdx.y * nq_stride + qi) * 2;
        out[0] = bd == 0xFFFFu ? EMPTY64 : (((uint64_t)bd << 40) | (index_base + bi));
        out[1] = sd == 0xFFFFu ? EMPTY64 : (((uint64_t)sd << 40) | (index_base + si));
    }
}

// two smallest keys over `parts` blocks of [nq_stride][2] keys (parts given either as one array or as separate pointers)
struct PartPtrs {
    const uint64_t* p[16];
};
__global__ void k_knn_reduce(const uint64_t* __restrict__ partial, PartPtrs ptrs, uint32_t parts, uint32_t nq, uint32_t nq_stride,
                             uint64_t* __restrict__ out) {
    const uint32_t qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    uint64_t b = EMPTY64, s = EMPTY64;
    for (uint32_t c = 0; c < parts; ++c) {
        const uint64_t* src = partial ? partial + ((size_t)c * nq_stride + qi) * 2 : ptrs.p[c] + (size_t)qi * 2;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const uint64_t key = src[k];
            if (key < b) {
                s = b;
                b = key;
            }
            else if (key < s) s = key;
        }
    }
    out[(size_t)qi * 2] = b;
    out[(size_t)qi * 2 + 1] = s;
}

__global__ void k_knn_unpack(const uint64_t* __restrict__ keys, uint32_t nq, int64_t* best_idx, uint32_t* best_d, uint32_t* second_d) {
    const uint32_t qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    const uint64_t b = keys[(size_t)qi * 2], s = keys[(size_t)qi * 2 + 1];
    best_idx[qi] = b == EMPTY64 ? -1 : (int64_t)(b & ((1ull << 40) - 1));
    best_d[qi] = b == EMPTY64 ? 256u : (uint32_t)(b >> 40);    // MAX_HAMMING_DIST, match/base.h:15
    second_d[qi] = s == EMPTY64 ? 256u : (uint32_t)(s >> 40);
}

int ensure_scratch(lpslam_knn_db* db, uint32_t nq, uint32_t chunks) {
    if (nq <= db->nq_cap && chunks <= db->chunks_cap) return LPSLAM_OK;
    // allocate into temporaries and commit pointers and capacities together: a failed allocation must not leave enlarged
    // capacities over null buffers (the next query would skip this function and scan into a null pointer)
    const uint32_t nq_cap = std::max(nq, db->nq_cap), chunks_cap = std::max(chunks, db->chunks_cap);
    uint64_t* partial = nullptr;
    uint8_t* q_dev = nullptr;
    uint64_t* keys_dev = nullptr;
    cudaError_t e = cudaMalloc(&partial, (size_t)chunks_cap * nq_cap * 16);
    if (e == cudaSuccess) e = cudaMalloc(&q_dev, (size_t)nq_cap * 32);
    if (e == cudaSuccess) e = cudaMalloc(&keys_dev, (size_t)nq_cap * 16);
    if (e != cudaSuccess) {
        cudaFree(partial);
        cudaFree(q_dev);
        cudaFree(keys_dev);
        return cuda_fail(e, "kNN scratch allocation", __FILE__, __LINE__);
    }
    cudaFree(db->partial);
    cudaFree(db->q_dev);
    cudaFree(db->keys_dev);
    db->partial = partial;
    db->q_dev = q_dev;
    db->keys_dev = keys_dev;
    db->nq_cap = nq_cap;
    db->chunks_cap = chunks_cap;
    return LPSLAM_OK;
}

uint32_t pick_chunks(const lpslam_knn_db* db, uint32_t nq) {
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, db->device);
    const uint32_t qblocks = (nq + QB - 1) / QB;
    // ~8 resident blocks of 128 threads per SM; at least 4 tiles per chunk
    uint32_t chunks = std::max<uint32_t>(1, (uint32_t)(sms * 12) / std::max<uint32_t>(qblocks, 1));
    const uint64_t max_chunks = std::max<uint64_t>(1, db->n / (4 * TILE));
    chunks = (uint32_t)std::min<uint64_t>(chunks, max_chunks);
    return std::max<uint32_t>(chunks, 1);
}

}  // namespace
}  // namespace lps

using namespace lps;

extern "C" {

int lpslam_knn_create(lpslam_knn_db** out, int device, uint64_t capacity_rows, uint64_t index_base) {
    LPS_ARG(out && capacity_rows > 0, "bad kNN database arguments");
    LPS_ARG(index_base + capacity_rows < (1ull << 40), "row index exceeds the 40-bit key field");
    *out = nullptr;
    int ndev = 0;
    LPS_CUDA(cudaGetDeviceCount(&ndev));
    LPS_ARG(device >= 0 && device < ndev, "no such CUDA device");
    LPS_CUDA(cudaSetDevice(device));
    auto* db = new lpslam_knn_db;
    db->device = device;
    db->cap = capacity_rows;
    db->index_base = index_base;
    const cudaError_t e = cudaMalloc(&db->rows, capacity_rows * 32);
    if (e != cudaSuccess) {
        delete db;
        return cuda_fail(e, "cudaMalloc(kNN rows)", __FILE__, __LINE__);
    }
    *out = db;
    return LPSLAM_OK;
}

void lpslam_knn_destroy(lpslam_knn_db* db) {
    if (!db) return;
    cudaSetDevice(db->device);
    cudaDeviceSynchronize();
    cudaFree(db->rows);
    cudaFree(db->partial);
    cudaFree(db->q_dev);
    cudaFree(db->keys_dev);
    delete db;
}

int lpslam_knn_add(lpslam_knn_db* db, const uint8_t* desc_host, uint64_t n_rows) {
    LPS_ARG(db && (n_rows == 0 || desc_host), "null argument");
    if (db->n + n_rows > db->cap) {
        set_error("kNN database capacity exceeded");
        return LPSLAM_ERR_CAPACITY;
    }
    LPS_CUDA(cudaSetDevice(db->device));
    if (n_rows) LPS_CUDA(cudaMemcpy(db->rows + db->n * 32, desc_host, n_rows * 32, cudaMemcpyHostToDevice));
    db->n += n_rows;
    return LPSLAM_OK;
}

int lpslam_knn_add_dev(lpslam_knn_db* db, const uint8_t* desc_dev, uint64_t n_rows) {
    LPS_ARG(db && (n_rows == 0 || desc_dev), "null argument");
    if (db->n + n_rows > db->cap) {
        set_error("kNN database capacity exceeded");
        return LPSLAM_ERR_CAPACITY;
    }
    LPS_CUDA(cudaSetDevice(db->device));
    if (n_rows) LPS_CUDA(cudaMemcpy(db->rows + db->n * 32, desc_dev, n_rows * 32, cudaMemcpyDeviceToDevice));
    db->n += n_rows;
    return LPSLAM_OK;
}

int lpslam_knn_size(const lpslam_knn_db* db, uint64_t* n_rows) {
    LPS_ARG(db && n_rows, "null argument");
    *n_rows = db->n;
    return LPSLAM_OK;
}

int lpslam_knn_query_dev(lpslam_knn_db* db, const uint8_t* q_dev, uint32_t nq, uint64_t* keys_dev, void* stream) {
    LPS_ARG(db && (nq == 0 || (q_dev && keys_dev)), "null argument");
    if (nq == 0) return LPSLAM_OK;
    LPS_CUDA(cudaSetDevice(db->device));
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t chunks = pick_chunks(db, nq);
    const int rc = ensure_scratch(db, nq, chunks);
    if (rc) return rc;
    const uint32_t rows_per_chunk = (uint32_t)((db->n + chunks - 1) / chunks);
    const uint32_t rpc = std::max<uint32_t>((rows_per_chunk + TILE - 1) / TILE * TILE, TILE);
    const uint32_t used = db->n ? (uint32_t)((db->n + rpc - 1) / rpc) : 1;
    k_knn_scan<<<dim3((nq + QB - 1) / QB, used), QB, 0, st>>>(db->rows, db->n, db->index_base, q_dev, nq, rpc, db->partial, db->nq_cap);
    PartPtrs none{};
    k_knn_reduce<<<(nq + 255) / 256, 256, 0, st>>>(db->partial, none, used, nq, db->nq_cap, keys_dev);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

int lpslam_knn_merge_dev(const uint64_t* gathered_dev, uint32_t n_shards, uint32_t nq, uint64_t* merged_dev, void* stream) {
    LPS_ARG(gathered_dev && merged_dev && n_shards >= 1, "null argument");
    if (nq == 0) return LPSLAM_OK;
    PartPtrs none{};
    k_knn_reduce<<<(nq + 255) / 256, 256, 0, (cudaStream_t)stream>>>(gathered_dev, none, n_shards, nq, nq, merged_dev);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

int lpslam_knn_merge_peer_dev(const uint64_t* const* shard_keys_dev, uint32_t n_shards, uint32_t nq, uint64_t* merged_dev, void* stream) {
    LPS_ARG(shard_keys_dev && merged_dev && n_shards >= 1 && n_shards <= 16, "bad shard pointer list (1..16 shards)");
    if (nq == 0) return LPSLAM_OK;
    PartPtrs ptrs{};
    for (uint32_t i = 0; i < n_shards; ++i) {
        LPS_ARG(shard_keys_dev[i], "null shard pointer");
        ptrs.p[i] = shard_keys_dev[i];
    }
    k_knn_reduce<<<(nq + 255) / 256, 256, 0, (cudaStream_t)stream>>>(nullptr, ptrs, n_shards, nq, nq, merged_dev);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

// ---- sharded query with the exchange below the C ABI: per-shard top-2, ncclAllGather of the (nq, 2) keys, merge. NCCL is resolved at
// run time (dlopen / dlsym) from the libnccl the process already has (e.g. the one torch loaded) or the system's, so that the
// library carries no link-time dependency on one NCCL build and still loads on a machine without NCCL.
namespace {
typedef int (*NcclAllGatherFn)(const void*, void*, size_t, int /*ncclDataType_t*/, void* /*ncclComm_t*/, cudaStream_t);
NcclAllGatherFn nccl_all_gather() {
    static NcclAllGatherFn fn = [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        return h ? reinterpret_cast<NcclAllGatherFn>(dlsym(h, "ncclAllGather")) : nullptr;
    }();
    return fn;
}
}  // namespace

int lpslam_knn_sharded_query_dev(lpslam_knn_db* db, void* nccl_comm, uint32_t world, const uint8_t* q_dev, uint32_t nq,
                                 uint64_t* gathered_dev, uint64_t* merged_dev, void* stream) {
    LPS_ARG(db && nccl_comm && world >= 1 && (nq == 0 || (q_dev && gathered_dev && merged_dev)), "null argument");
    if (nq == 0) return LPSLAM_OK;
    NcclAllGatherFn all_gather = nccl_all_gather();
    if (!all_gather) {
        set_error("libnccl.so.2 (ncclAllGather) could not be loaded");
        return LPSLAM_ERR_STATE;
    }
    // the rank's own keys go straight into its slot... ncclAllGather wants a separate send buffer unless it is that slot
    uint64_t* mine = merged_dev;  // used as the send buffer, overwritten by the merge afterwards
    int rc = lpslam_knn_query_dev(db, q_dev, nq, mine, stream);
    if (rc) return rc;
    const int nccl_rc = all_gather(mine, gathered_dev, (size_t)nq * 2, 5 /* ncclUint64 */, nccl_comm, (cudaStream_t)stream);
    if (nccl_rc != 0) {
        set_error("ncclAllGather failed with ncclResult_t " + std::to_string(nccl_rc));
        return LPSLAM_ERR_CUDA;
    }
    return lpslam_knn_merge_dev(gathered_dev, world, nq, merged_dev, stream);
}

int lpslam_knn_unpack(const uint64_t* merged_dev, uint32_t nq, int64_t* best_idx, uint32_t* best_dist, uint32_t* second_dist, void* stream) {
    LPS_ARG(nq == 0 || (merged_dev && best_idx && best_dist && second_dist), "null argument");
    if (nq == 0) return LPSLAM_OK;
    cudaStream_t st = (cudaStream_t)stream;
    // the calling thread's grow-only scratch (scratch.cu): a relocalisation query no longer pays a cudaMalloc + cudaFree
    // (two device synchronisations) per call; the stream is drained below before the buffer is handed back
    uint8_t* dev = nullptr;
    LPS_CUDA(scratch_alloc((void**)&dev, (size_t)nq * 16));
    int64_t* bi = (int64_t*)dev;
    uint32_t* bd = (uint32_t*)(dev + (size_t)nq * 8);
    uint32_t* sd = bd + nq;
    k_knn_unpack<<<(nq + 255) / 256, 256, 0, st>>>(merged_dev, nq, bi, bd, sd);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_idx, bi, (size_t)nq * 8, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_dist, bd, (size_t)nq * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(second_dist, sd, (size_t)nq * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    scratch_free(dev);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_knn_unpack", __FILE__, __LINE__);
    return LPSLAM_OK;
}

int lpslam_knn_query(lpslam_knn_db* db, const uint8_t* q_host, uint32_t nq, int64_t* best_idx, uint32_t* best_dist,
                     uint32_t* second_dist) {
    LPS_ARG(db && (nq == 0 || (q_host && best_idx && best_dist && second_dist)), "null argument");
    if (nq == 0) return LPSLAM_OK;
    LPS_CUDA(cudaSetDevice(db->device));
    int rc = ensure_scratch(db, nq, pick_chunks(db, nq));
    if (rc) return rc;
    LPS_CUDA(cudaMemcpy(db->q_dev, q_host, (size_t)nq * 32, cudaMemcpyHostToDevice));
    rc = lpslam_knn_query_dev(db, db->q_dev, nq, db->keys_dev, nullptr);
    if (rc) return rc;
    return lpslam_knn_unpack(db->keys_dev, nq, best_idx, best_dist, second_dist, nullptr);
}

}  // extern "C"
