// C++ host test of the sharded relocalisation kNN (BASELINE configs[4]) entirely below the C ABI: one host thread per GPU of ONE
// process, an NCCL communicator over all of them (ncclCommInitAll), a database shard per GPU, the same queries on every GPU,
// lpslam_knn_sharded_query_dev on every rank, and the merged result of EVERY rank held against the planted neighbours. Also runs
// lpslam_knn_merge_peer_dev across real peers (cudaDeviceEnablePeerAccess, the other GPUs' key blocks read over NVLink).
//   usage: knn_nccl_test [rows_per_gpu=200000] [queries=2000]
// Built by cpp/Makefile against /usr/include/nccl.h; exits 0 on success.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include <cuda_runtime.h>
#include <nccl.h>

#include "lpslam_cuda.h"

#define CK(x)                                                                            \
    do {                                                                                 \
        if ((x) != 0) {                                                                  \
            std::fprintf(stderr, "%s failed at line %d: %s\n", #x, __LINE__, lpslam_cuda_last_error()); \
            std::exit(1);                                                                \
        }                                                                                \
    } while (0)

int main(int argc, char** argv) {
    const uint64_t rows = argc > 1 ? std::strtoull(argv[1], nullptr, 10) : 200000;
    const uint32_t nq = argc > 2 ? static_cast<uint32_t>(std::atoi(argv[2])) : 2000;
    int world = 0;
    cudaGetDeviceCount(&world);
    if (world < 1) return 2;
    std::vector<int> devs(world);
    for (int i = 0; i < world; ++i) devs[i] = i;
    std::vector<ncclComm_t> comms(world);
    if (ncclCommInitAll(comms.data(), world, devs.data()) != ncclSuccess) return 3;

    // database: random rows; queries: planted rows with ~10 % of the bits flipped
    std::mt19937_64 rng(5);
    std::vector<uint8_t> db(static_cast<size_t>(world) * rows * 32), q(static_cast<size_t>(nq) * 32);
    for (size_t i = 0; i < db.size(); i += 8) {
        const uint64_t v = rng();
        std::memcpy(&db[i], &v, 8);
    }
    std::vector<uint64_t> target(nq);
    std::vector<uint32_t> flipped(nq);
    for (uint32_t i = 0; i < nq; ++i) {
        target[i] = rng() % (world * rows);
        std::memcpy(&q[32 * static_cast<size_t>(i)], &db[32 * target[i]], 32);
        uint32_t n = 0;
        for (int b = 0; b < 256; ++b)
            if (rng() % 10 == 0) {
                q[32 * static_cast<size_t>(i) + b / 8] ^= static_cast<uint8_t>(1u << (b % 8));
                ++n;
            }
        flipped[i] = n;
    }

    std::vector<lpslam_knn_db*> shard(world);
    std::vector<uint8_t*> q_dev(world);
    std::vector<uint64_t*> gathered(world), merged(world), own(world), merged_peer(world);
    std::vector<cudaStream_t> st(world);
    for (int r = 0; r < world; ++r) {
        cudaSetDevice(r);
        for (int p = 0; p < world; ++p)
            if (p != r) cudaDeviceEnablePeerAccess(p, 0);
        cudaGetLastError();
        cudaStreamCreate(&st[r]);
        CK(lpslam_knn_create(&shard[r], r, rows, static_cast<uint64_t>(r) * rows));
        CK(lpslam_knn_add(shard[r], &db[static_cast<size_t>(r) * rows * 32], rows));
        cudaMalloc(&q_dev[r], static_cast<size_t>(nq) * 32);
        cudaMemcpy(q_dev[r], q.data(), static_cast<size_t>(nq) * 32, cudaMemcpyHostToDevice);
        cudaMalloc(&gathered[r], static_cast<size_t>(world) * nq * 16);
        cudaMalloc(&merged[r], static_cast<size_t>(nq) * 16);
        cudaMalloc(&own[r], static_cast<size_t>(nq) * 16);
        cudaMalloc(&merged_peer[r], static_cast<size_t>(nq) * 16);
    }
    int failures = 0;
    // (1) NCCL exchange: one thread per rank
    std::vector<std::thread> th;
    for (int r = 0; r < world; ++r)
        th.emplace_back([&, r] {
            cudaSetDevice(r);
            CK(lpslam_knn_sharded_query_dev(shard[r], comms[r], static_cast<uint32_t>(world), q_dev[r], nq, gathered[r], merged[r], st[r]));
            cudaStreamSynchronize(st[r]);
        });
    for (auto& t : th) t.join();
    // (2) peer-memory exchange: every rank scans into its own block, then reads ALL blocks through peer pointers
    for (int r = 0; r < world; ++r) {
        cudaSetDevice(r);
        CK(lpslam_knn_query_dev(shard[r], q_dev[r], nq, own[r], st[r]));
    }
    for (int r = 0; r < world; ++r) {
        cudaSetDevice(r);
        cudaStreamSynchronize(st[r]);
    }
    for (int r = 0; r < world; ++r) {
        cudaSetDevice(r);
        CK(lpslam_knn_merge_peer_dev(own.data(), static_cast<uint32_t>(world), nq, merged_peer[r], st[r]));
        cudaStreamSynchronize(st[r]);
    }
    for (int r = 0; r < world; ++r) {
        cudaSetDevice(r);
        for (uint64_t* keys : {merged[r], merged_peer[r]}) {
            std::vector<int64_t> idx(nq);
            std::vector<uint32_t> bd(nq), sd(nq);
            CK(lpslam_knn_unpack(keys, nq, idx.data(), bd.data(), sd.data(), st[r]));
            cudaStreamSynchronize(st[r]);
            for (uint32_t i = 0; i < nq; ++i)
                if (idx[i] != static_cast<int64_t>(target[i]) || bd[i] != flipped[i]) {
                    if (failures < 5) std::fprintf(stderr, "rank %d query %u: got row %lld dist %u, planted row %llu dist %u\n", r, i,
                                                   static_cast<long long>(idx[i]), bd[i], static_cast<unsigned long long>(target[i]), flipped[i]);
                    ++failures;
                }
        }
    }
    std::printf("world %d, %llu rows per GPU, %u queries: %s (nccl all-gather + merge, and peer-memory merge, checked on every rank)\n", world,
                static_cast<unsigned long long>(rows), nq, failures ? "FAILED" : "ok");
    for (int r = 0; r < world; ++r) {
        lpslam_knn_destroy(shard[r]);
        ncclCommDestroy(comms[r]);
    }
    return failures ? 1 : 0;
}
