// Per-thread device scratch for the short matcher / pre-processing entry points. They used to cudaMalloc + cudaFree on
// every call (two device synchronisations, ~0.2 ms) for a few hundred microseconds of work; the buffer is now kept per
// host thread and only grows. One buffer per thread: a nested request (none today) falls back to cudaMalloc.
#include "common.cuh"

namespace lps {

namespace {
struct Scratch {
    void* p = nullptr;
    size_t cap = 0;
    int dev = -1;
    bool busy = false;
    ~Scratch() {
        if (p) cudaFree(p);  // at thread exit; an error after the runtime shut down is ignored
    }
};
thread_local Scratch t_scratch;
}  // namespace

cudaError_t scratch_alloc(void** out, size_t bytes) {
    Scratch& s = t_scratch;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (s.busy) return cudaMalloc(out, bytes);
    if (s.dev != dev || s.cap < bytes) {
        if (s.p) cudaFree(s.p);
        s.p = nullptr;
        s.cap = 0;
        const size_t want = align_up(bytes + bytes / 2 + 4096, 1 << 16);
        e = cudaMalloc(&s.p, want);
        if (e != cudaSuccess) return e;
        s.cap = want;
        s.dev = dev;
    }
    s.busy = true;
    *out = s.p;
    return cudaSuccess;
}

void scratch_free(void* p) {
    Scratch& s = t_scratch;
    if (p && p == s.p) {
        s.busy = false;
        return;
    }
    if (p) cudaFree(p);
}

}  // namespace lps
