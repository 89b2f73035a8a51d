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

namespace {
struct StageBuffers {
    uint8_t* dev = nullptr;
    uint8_t* host = nullptr;
    size_t cap = 0;
    cudaStream_t st = nullptr, side = nullptr;
    cudaEvent_t joined = nullptr;
    int device = -1;
    bool busy = false;
    void drop() {
        if (dev) cudaFree(dev);
        if (host) cudaFreeHost(host);
        if (st) cudaStreamDestroy(st);
        if (side) cudaStreamDestroy(side);
        if (joined) cudaEventDestroy(joined);
        dev = host = nullptr;
        st = side = nullptr;
        joined = nullptr;
        cap = 0;
    }
    ~StageBuffers() { drop(); }  // at thread exit; errors after the runtime shut down are ignored
};
thread_local StageBuffers t_stage;

cudaError_t stage_allocate(StageBuffers& b, size_t want) {
    cudaError_t e;
    if (!b.st && (e = cudaStreamCreateWithFlags(&b.st, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if (!b.side && (e = cudaStreamCreateWithFlags(&b.side, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if (!b.joined && (e = cudaEventCreateWithFlags(&b.joined, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaMalloc((void**)&b.dev, want)) != cudaSuccess) return e;
    if ((e = cudaHostAlloc((void**)&b.host, want, cudaHostAllocDefault)) != cudaSuccess) return e;
    b.cap = want;
    return cudaSuccess;
}
}  // namespace

cudaError_t CallStage::open(size_t bytes) {
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    StageBuffers& b = t_stage;
    if (b.busy) {  // nested use on one thread (none today): private buffers for this call
        StageBuffers own;
        if ((e = stage_allocate(own, align_up(bytes + 256, 1 << 12))) != cudaSuccess) return e;
        dev = own.dev, host = own.host, st = own.st, side = own.side, joined = own.joined;
        own.dev = own.host = nullptr;
        own.st = own.side = nullptr;
        own.joined = nullptr;
        owned_ = true;
        return cudaSuccess;
    }
    if (b.device != device || b.cap < bytes) {
        b.drop();
        b.device = device;
        if ((e = stage_allocate(b, align_up(bytes + bytes / 2 + 4096, 1 << 16))) != cudaSuccess) {
            b.drop();
            return e;
        }
    }
    b.busy = true;
    dev = b.dev, host = b.host, st = b.st, side = b.side, joined = b.joined;
    return cudaSuccess;
}

CallStage::~CallStage() {
    if (owned_) {
        if (st) cudaStreamSynchronize(st);
        if (side) cudaStreamSynchronize(side);
        cudaFree(dev);
        cudaFreeHost(host);
        if (st) cudaStreamDestroy(st);
        if (side) cudaStreamDestroy(side);
        if (joined) cudaEventDestroy(joined);
    }
    else if (dev && dev == t_stage.dev) {
        // a call that failed half-way may have left work queued on the stream: the buffers are reused by the next call
        if (cudaStreamQuery(st) == cudaErrorNotReady) cudaStreamSynchronize(st);
        if (cudaStreamQuery(side) == cudaErrorNotReady) cudaStreamSynchronize(side);
        t_stage.busy = false;
    }
}

}  // namespace lps
