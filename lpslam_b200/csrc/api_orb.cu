// C ABI — library + ORB extractor entry points (include/lpslam_cuda.h).
#include "ctx.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace lps {
const std::string& get_error();
}
using namespace lps;

namespace {
struct PackArgs {
    const uint8_t* src[kMaxLevels];
    int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels];
    size_t off[kMaxLevels];
};

// rows of several levels gathered into one tightly packed buffer (row pitch = level width)
__global__ void k_pack_levels(PackArgs a, uint8_t* __restrict__ dst) {
    const int l = blockIdx.y;
    const uint8_t* __restrict__ src = a.src[l];
    uint8_t* out = dst + a.off[l];
    for (int y = blockIdx.x; y < a.h[l]; y += gridDim.x)
        for (int x = threadIdx.x; x < a.w[l]; x += blockDim.x) out[(size_t)y * a.w[l] + x] = src[(size_t)y * a.pitch[l] + x];
}
}  // namespace

extern "C" {

int lpslam_cuda_abi_version(void) { return LPSLAM_CUDA_ABI_VERSION; }

const char* lpslam_cuda_last_error(void) { return get_error().c_str(); }

int lpslam_cuda_device_count(int* count) {
    LPS_ARG(count, "count is null");
    *count = 0;
    LPS_CUDA(cudaGetDeviceCount(count));
    return LPSLAM_OK;
}

int lpslam_cuda_host_alloc(void** ptr, size_t bytes) {
    LPS_ARG(ptr && bytes, "bad host_alloc arguments");
    LPS_CUDA(cudaHostAlloc(ptr, bytes, cudaHostAllocDefault));
    return LPSLAM_OK;
}

int lpslam_cuda_host_free(void* ptr) {
    if (ptr) LPS_CUDA(cudaFreeHost(ptr));
    return LPSLAM_OK;
}

int lpslam_orb_create(lpslam_orb_ctx** out, const lpslam_orb_params* p, int device, uint32_t max_batch) {
    LPS_ARG(out && p, "null argument");
    *out = nullptr;
    LPS_ARG(p->num_levels >= 1 && p->num_levels <= (uint32_t)kMaxLevels, "num_levels out of range (1..16)");
    LPS_ARG(p->scale_factor > 0.f, "scale_factor must be positive");
    LPS_ARG(p->ini_fast_thr <= 255 && p->min_fast_thr <= 255, "FAST thresholds must be <= 255");
    LPS_ARG(max_batch >= 1, "max_batch must be >= 1");
    LPS_ARG(p->num_mask_rects == 0 || p->mask_rects, "mask_rects is null");
    // feature/orb_params.cc:16-26: x_min < x_max and y_min < y_max or std::runtime_error
    for (uint32_t i = 0; i < p->num_mask_rects; ++i) {
        const float* r = p->mask_rects + 4 * i;
        LPS_ARG(r[0] < r[1] && r[2] < r[3], "mask rectangle must satisfy x_min < x_max and y_min < y_max");
    }
    int ndev = 0;
    LPS_CUDA(cudaGetDeviceCount(&ndev));
    LPS_ARG(device >= 0 && device < ndev, "no such CUDA device");
    LPS_CUDA(cudaSetDevice(device));
    auto* c = new lpslam_orb_ctx;
    c->device = device;
    c->max_batch = max_batch;
    c->max_kpts = p->max_num_keypts;
    c->ini_max_kpts = p->ini_max_num_keypts;
    c->scale_factor = p->scale_factor;
    c->num_levels = p->num_levels;
    c->ini_thr = p->ini_fast_thr;
    c->min_thr = p->min_fast_thr;
    for (uint32_t i = 0; i < p->num_mask_rects; ++i)
        c->mask_rects.push_back({p->mask_rects[4 * i], p->mask_rects[4 * i + 1], p->mask_rects[4 * i + 2], p->mask_rects[4 * i + 3]});
    c->initialize_tables();
    // greatest priority for everything but FAST (launch_fast lowers its own launch): see fast.cu
    int prio_least = 0, prio_greatest = 0;
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    if (const char* env = getenv("LPSLAM_FAST_LOW_PRIORITY")) if (atoi(env) == 0) prio_greatest = prio_least;
    cudaError_t e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_greatest);
    if (e == cudaSuccess) e = cudaStreamCreateWithPriority(&c->side, cudaStreamNonBlocking, prio_greatest);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_reader, cudaEventDisableTiming);
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) {
        e = cudaStreamCreateWithPriority(&c->chain[i], cudaStreamNonBlocking, prio_greatest);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_chain_join[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_chain_fork, cudaEventDisableTiming);
    if (const char* env = getenv("LPSLAM_SUBBATCH")) c->sub_batches = std::min(std::max(atoi(env), 1), 4);
    if (const char* env = getenv("LPSLAM_PYR_CHAINS")) c->pyr_chains = std::min(std::max(atoi(env), 1), 4);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&c->h_frame_off, sizeof(uint32_t) * (max_batch + 2), cudaHostAllocDefault);
    for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreate(&c->ev[i]);
    if (e != cudaSuccess) {
        lpslam_orb_destroy(c);
        return cuda_fail(e, "context creation", __FILE__, __LINE__);
    }
    *out = c;
    return LPSLAM_OK;
}

void lpslam_orb_destroy(lpslam_orb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->side) cudaStreamSynchronize(c->side);
    for (int i = 0; i < 3; ++i) {
        if (c->chain[i]) cudaStreamSynchronize(c->chain[i]);
        if (c->chain[i]) cudaStreamDestroy(c->chain[i]);
        if (c->ev_chain_join[i]) cudaEventDestroy(c->ev_chain_join[i]);
    }
    if (c->ev_chain_fork) cudaEventDestroy(c->ev_chain_fork);
    c->release_geometry();
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->ev_reader) cudaEventDestroy(c->ev_reader);
    if (c->side) cudaStreamDestroy(c->side);
    if (c->h_frame_off) cudaFreeHost(c->h_frame_off);
    if (c->pack_dev) cudaFree(c->pack_dev);
    if (c->pack_host) cudaFreeHost(c->pack_host);
    for (auto& ev : c->ev)
        if (ev) cudaEventDestroy(ev);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int lpslam_orb_set_param(lpslam_orb_ctx* c, int which, double value) {
    LPS_ARG(c, "null context");
    c->drop_pair_graph();  // kernel arguments change
    switch (which) {
        case LPSLAM_ORB_MAX_NUM_KEYPTS: c->max_kpts = (uint32_t)value; break;
        case LPSLAM_ORB_SCALE_FACTOR:
            LPS_ARG(value > 0, "scale_factor must be positive");
            c->scale_factor = (float)value;
            break;
        case LPSLAM_ORB_NUM_LEVELS:
            LPS_ARG(value >= 1 && value <= kMaxLevels, "num_levels out of range (1..16)");
            c->num_levels = (uint32_t)value;
            break;
        case LPSLAM_ORB_INI_FAST_THR:
            LPS_ARG(value >= 0 && value <= 255, "FAST threshold must be in 0..255");
            c->ini_thr = (uint32_t)value;
            return LPSLAM_OK;  // no re-initialisation (orb_extractor.cc:188-190)
        case LPSLAM_ORB_MIN_FAST_THR:
            LPS_ARG(value >= 0 && value <= 255, "FAST threshold must be in 0..255");
            c->min_thr = (uint32_t)value;
            return LPSLAM_OK;
        default: LPS_ARG(false, "unknown parameter id");
    }
    // orb_extractor.cc:153-174: these three setters re-run initialize(); geometry is rebuilt on the next extract
    c->initialize_tables();
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->release_geometry();
    return LPSLAM_OK;
}

int lpslam_orb_get_param(const lpslam_orb_ctx* c, int which, double* value) {
    LPS_ARG(c && value, "null argument");
    switch (which) {
        case LPSLAM_ORB_MAX_NUM_KEYPTS: *value = c->max_kpts; break;
        case LPSLAM_ORB_SCALE_FACTOR: *value = c->scale_factor; break;
        case LPSLAM_ORB_NUM_LEVELS: *value = c->num_levels; break;
        case LPSLAM_ORB_INI_FAST_THR: *value = c->ini_thr; break;
        case LPSLAM_ORB_MIN_FAST_THR: *value = c->min_thr; break;
        case LPSLAM_ORB_DEBUG_GRAPH_LAUNCHES: *value = (double)c->pair_graph_launches; break;
        default: LPS_ARG(false, "unknown parameter id");
    }
    return LPSLAM_OK;
}

int lpslam_orb_get_table(const lpslam_orb_ctx* c, int which, float* out, uint32_t cap) {
    LPS_ARG(c && out, "null argument");
    if (cap < c->num_levels) {
        set_error("table buffer too small");
        return LPSLAM_ERR_CAPACITY;
    }
    for (uint32_t l = 0; l < c->num_levels; ++l) {
        switch (which) {
            case 0: out[l] = c->scale[l]; break;
            case 1: out[l] = c->inv_scale[l]; break;
            case 2: out[l] = c->sigma_sq[l]; break;
            case 3: out[l] = c->inv_sigma_sq[l]; break;
            case 4: out[l] = (float)c->quota[l]; break;
            default: LPS_ARG(false, "unknown table id");
        }
    }
    return LPSLAM_OK;
}

static int prepare_rect_mask(lpslam_orb_ctx* c) {
    // feature/orb_extractor.cc:251-264: white image, each rectangle filled with 0. cv::rectangle(..., LINE_AA) leaves a
    // non-zero anti-aliased fringe outside the inclusive rectangle and only `== 0` is ever tested, so the exact
    // inclusive fill is equivalent.
    if (c->rect_mask_ready || c->mask_rects.empty()) return LPSLAM_OK;
    std::vector<uint8_t> m((size_t)c->w * c->h, 255);
    for (const auto& r : c->mask_rects) {
        int x0 = (int)(unsigned)std::round((unsigned)c->w * r[0]), x1 = (int)(unsigned)std::round((unsigned)c->w * r[1]);
        int y0 = (int)(unsigned)std::round((unsigned)c->h * r[2]), y1 = (int)(unsigned)std::round((unsigned)c->h * r[3]);
        if (x0 > x1) std::swap(x0, x1);
        if (y0 > y1) std::swap(y0, y1);
        x0 = std::max(x0, 0);
        y0 = std::max(y0, 0);
        x1 = std::min(x1, c->w - 1);
        y1 = std::min(y1, c->h - 1);
        for (int y = y0; y <= y1; ++y) memset(&m[(size_t)y * c->w + x0], 0, (size_t)std::max(0, x1 - x0 + 1));
    }
    LPS_CUDA(cudaMemcpyAsync(c->rect_mask_dev, m.data(), m.size(), cudaMemcpyHostToDevice, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    c->rect_mask_ready = true;
    return LPSLAM_OK;
}

static int select_mask(lpslam_orb_ctx* c, const uint8_t* mask_host, size_t mask_pitch, const uint8_t** mptr, size_t* mpitch) {
    // mask priority, feature/orb_extractor.cc:93-107
    *mptr = nullptr;
    *mpitch = 0;
    if (mask_host) {
        LPS_CUDA(cudaMemcpy2DAsync(c->mask_dev, c->w, mask_host, mask_pitch, c->w, c->h, cudaMemcpyHostToDevice, c->stream));
        *mptr = c->mask_dev;
        *mpitch = c->w;
    }
    else if (!c->mask_rects.empty()) {
        const int rc = prepare_rect_mask(c);
        if (rc) return rc;
        *mptr = c->rect_mask_dev;
        *mpitch = c->w;
    }
    return LPSLAM_OK;
}

static int fetch_results(lpslam_orb_ctx* c, uint32_t n_frames, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap,
                         uint32_t* offsets) {
    LPS_CUDA(cudaMemcpyAsync(c->h_frame_off, c->desc.frame_off, sizeof(uint32_t) * (c->max_batch + 2), cudaMemcpyDeviceToHost,
                             c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    const uint32_t total = c->h_frame_off[n_frames];
    for (uint32_t f = 0; f <= n_frames; ++f) offsets[f] = c->h_frame_off[f];
    if (c->h_frame_off[c->max_batch + 1] != 0) {
        set_error("internal output buffer overflow");
        return LPSLAM_ERR_CAPACITY;
    }
    if (total > cap) {
        set_error("output capacity too small for the extracted keypoints");
        return LPSLAM_ERR_CAPACITY;
    }
    if (total) {
        LPS_CUDA(cudaMemcpyAsync(kp_out, c->desc.kp, sizeof(lpslam_keypoint) * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
        LPS_CUDA(cudaMemcpyAsync(desc_out, c->desc.desc, 32 * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
    }
    if (c->timing) cudaEventRecord(c->ev[6], c->stream);
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    return LPSLAM_OK;
}

int lpslam_orb_extract_batch(lpslam_orb_ctx* c, const uint8_t* const* imgs, uint32_t n_frames, int width, int height,
                             size_t pitch, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets) {
    LPS_ARG(c && imgs && offsets, "null argument");
    LPS_ARG(n_frames >= 1 && n_frames <= c->max_batch, "n_frames exceeds the context's max_batch");
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width, "bad image geometry");
    LPS_ARG(kp_out && desc_out, "null output buffer");
    LPS_CUDA(cudaSetDevice(c->device));
    int rc = c->configure(width, height);
    if (rc) return rc;
    if (c->timing) cudaEventRecord(c->ev[7], c->stream);
    for (uint32_t f = 0; f < n_frames; ++f) {
        LPS_ARG(imgs[f], "null image pointer");
        LPS_CUDA(cudaMemcpy2DAsync(c->lv[0].img + f * c->lv[0].frame_stride, c->lv[0].pitch, imgs[f], pitch, width, height,
                                   cudaMemcpyHostToDevice, c->stream));
    }
    const uint8_t* m = nullptr;
    size_t mp = 0;
    rc = select_mask(c, nullptr, 0, &m, &mp);
    if (rc) return rc;
    rc = c->run_pipeline(n_frames, m, mp);
    if (rc) return rc;
    return fetch_results(c, n_frames, kp_out, desc_out, cap, offsets);
}

int lpslam_orb_extract(lpslam_orb_ctx* c, const uint8_t* img, int width, int height, size_t pitch, const uint8_t* mask,
                       size_t mask_pitch, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* n_out) {
    LPS_ARG(c && n_out, "null argument");
    *n_out = 0;
    if (!img || width <= 0 || height <= 0) return LPSLAM_OK;  // empty image: outputs untouched (orb_extractor.cc:73-75)
    LPS_ARG(pitch >= (size_t)width, "pitch smaller than width");
    LPS_ARG(!mask || mask_pitch >= (size_t)width, "mask pitch smaller than width");
    LPS_ARG(cap == 0 || (kp_out && desc_out), "null output buffer");
    LPS_CUDA(cudaSetDevice(c->device));
    int rc = c->configure(width, height);
    if (rc) return rc;
    if (c->timing) cudaEventRecord(c->ev[7], c->stream);
    LPS_CUDA(cudaMemcpy2DAsync(c->lv[0].img, c->lv[0].pitch, img, pitch, width, height, cudaMemcpyHostToDevice, c->stream));
    const uint8_t* m = nullptr;
    size_t mp = 0;
    rc = select_mask(c, mask, mask_pitch, &m, &mp);
    if (rc) return rc;
    rc = c->run_pipeline(1, m, mp);
    if (rc) return rc;
    uint32_t offs[2] = {0, 0};
    rc = fetch_results(c, 1, kp_out, desc_out, cap, offs);
    *n_out = offs[1];
    return rc;
}

int lpslam_orb_extract_batch_dev(lpslam_orb_ctx* c, const uint8_t* imgs_dev, uint32_t n_frames, int width, int height,
                                 size_t pitch) {
    LPS_ARG(c && imgs_dev, "null argument");
    LPS_ARG(n_frames >= 1 && n_frames <= c->max_batch, "n_frames exceeds the context's max_batch");
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width, "bad image geometry");
    LPS_CUDA(cudaSetDevice(c->device));
    int rc = c->configure(width, height);
    if (rc) return rc;
    if (imgs_dev != c->lv[0].img)
        launch_copy_level0(imgs_dev, pitch, pitch * (size_t)height, c->lv[0], (int)n_frames, c->stream);
    const uint8_t* m = nullptr;
    size_t mp = 0;
    rc = select_mask(c, nullptr, 0, &m, &mp);
    if (rc) return rc;
    return c->run_pipeline(n_frames, m, mp);
}

int lpslam_orb_fetch_result(lpslam_orb_ctx* c, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets) {
    LPS_ARG(c && kp_out && desc_out && offsets, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    LPS_CUDA(cudaSetDevice(c->device));
    return fetch_results(c, c->last_batch, kp_out, desc_out, cap, offsets);
}

int lpslam_orb_get_result_dev(lpslam_orb_ctx* c, lpslam_orb_result_dev* out) {
    LPS_ARG(c && out, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    out->kp = c->desc.kp;
    out->desc = c->desc.desc;
    out->offsets = c->desc.frame_off;
    out->capacity = c->out_cap;
    return LPSLAM_OK;
}

int lpslam_orb_input_dev(lpslam_orb_ctx* c, int width, int height, uint8_t** ptr, size_t* pitch, size_t* frame_stride) {
    LPS_ARG(c && ptr && pitch && frame_stride, "null argument");
    LPS_ARG(width > 0 && height > 0, "bad image geometry");
    LPS_CUDA(cudaSetDevice(c->device));
    const int rc = c->configure(width, height);
    if (rc) return rc;
    *ptr = c->lv[0].img;
    *pitch = (size_t)c->lv[0].pitch;
    *frame_stride = c->lv[0].frame_stride;
    return LPSLAM_OK;
}

int lpslam_orb_sync(lpslam_orb_ctx* c) {
    LPS_ARG(c, "null context");
    LPS_CUDA(cudaSetDevice(c->device));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    return LPSLAM_OK;
}

void* lpslam_orb_stream(lpslam_orb_ctx* c) { return c ? (void*)c->stream : nullptr; }

int lpslam_orb_level_size(const lpslam_orb_ctx* c, uint32_t level, int* width, int* height) {
    LPS_ARG(c && width && height, "null argument");
    if (!c->arena) {
        set_error("no image geometry configured yet");
        return LPSLAM_ERR_STATE;
    }
    LPS_ARG(level < c->num_levels, "level out of range");
    *width = c->lv[level].w;
    *height = c->lv[level].h;
    return LPSLAM_OK;
}

int lpslam_orb_pyramid(lpslam_orb_ctx* c, uint32_t frame, uint32_t level, int blurred, uint8_t* dst, size_t pitch) {
    LPS_ARG(c && dst, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    LPS_ARG(level < c->num_levels && frame < c->last_batch, "level / frame out of range");
    const LevelGeom& g = c->lv[level];
    LPS_ARG(pitch >= (size_t)g.w, "pitch smaller than the level width");
    LPS_CUDA(cudaSetDevice(c->device));
    const uint8_t* src = (blurred ? g.blur : g.img) + frame * g.frame_stride;
    LPS_CUDA(cudaMemcpy2DAsync(dst, pitch, src, g.pitch, g.w, g.h, cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    return LPSLAM_OK;
}

int lpslam_orb_pyramid_levels(lpslam_orb_ctx* c, uint32_t frame, uint32_t first_level, uint8_t* dst, size_t cap, size_t* offsets,
                              size_t* total) {
    LPS_ARG(c && offsets && total, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    LPS_ARG(first_level < c->num_levels && frame < c->last_batch, "level / frame out of range");
    PackArgs a = {};
    const uint32_t n = c->num_levels - first_level;
    size_t bytes = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const LevelGeom& g = c->lv[first_level + i];
        a.src[i] = g.img + frame * g.frame_stride;
        a.w[i] = g.w;
        a.h[i] = g.h;
        a.pitch[i] = g.pitch;
        a.off[i] = offsets[i] = bytes;
        bytes += (size_t)g.w * g.h;
    }
    *total = bytes;
    if (!dst || cap < bytes) {
        set_error("destination smaller than the packed levels");
        return LPSLAM_ERR_CAPACITY;
    }
    LPS_CUDA(cudaSetDevice(c->device));
    // packed on the device, fetched into the context's page-locked landing buffer in one copy (a copy straight into pageable
    // memory runs at a fraction of the link rate), then handed over
    if (c->pack_cap < bytes) {
        if (c->pack_dev) cudaFree(c->pack_dev);
        if (c->pack_host) cudaFreeHost(c->pack_host);
        c->pack_dev = c->pack_host = nullptr;
        c->pack_cap = 0;
        LPS_CUDA(cudaMalloc((void**)&c->pack_dev, bytes));
        LPS_CUDA(cudaHostAlloc((void**)&c->pack_host, bytes, cudaHostAllocDefault));
        c->pack_cap = bytes;
    }
    k_pack_levels<<<dim3(64, n), 256, 0, c->stream>>>(a, c->pack_dev);
    LPS_CUDA(cudaGetLastError());
    LPS_CUDA(cudaMemcpyAsync(c->pack_host, c->pack_dev, bytes, cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    memcpy(dst, c->pack_host, bytes);
    return LPSLAM_OK;
}

static int debug_points(lpslam_orb_ctx* c, uint32_t frame, uint32_t level, bool selected, int32_t* xys, uint32_t cap,
                        uint32_t* n_out) {
    LPS_ARG(c && n_out, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    LPS_ARG(level < c->num_levels && frame < c->last_batch, "level / frame out of range");
    LPS_CUDA(cudaSetDevice(c->device));
    const LevelGeom& g = c->lv[level];
    const uint32_t* cnt_dev = (selected ? c->tree.sel_count : c->tree.cand_count) + (size_t)frame * c->num_levels + level;
    uint32_t n = 0;
    LPS_CUDA(cudaMemcpyAsync(&n, cnt_dev, 4, cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    *n_out = n;
    if (n > cap) {
        set_error("debug buffer too small");
        return LPSLAM_ERR_CAPACITY;
    }
    if (!n) return LPSLAM_OK;
    LPS_ARG(xys, "null output buffer");
    std::vector<uint32_t> xy(n);
    std::vector<uint8_t> sc(n);
    const size_t stride = selected ? c->sel_stride : c->cand_stride;
    const size_t begin = selected ? g.sel_begin : g.cand_begin;
    const uint32_t* xy_dev = (selected ? c->tree.sel_xys : c->tree.cand_xy) + frame * stride + begin;
    const uint8_t* sc_dev = (selected ? c->tree.sel_sc : c->tree.cand_sc) + frame * stride + begin;
    LPS_CUDA(cudaMemcpyAsync(xy.data(), xy_dev, 4 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaMemcpyAsync(sc.data(), sc_dev, n, cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    for (uint32_t i = 0; i < n; ++i) {
        xys[3 * i] = (int32_t)(xy[i] & 0xFFFFu);
        xys[3 * i + 1] = (int32_t)(xy[i] >> 16);
        xys[3 * i + 2] = sc[i];
    }
    return LPSLAM_OK;
}

int lpslam_orb_debug_candidates(lpslam_orb_ctx* c, uint32_t frame, uint32_t level, int32_t* xys, uint32_t cap, uint32_t* n_out) {
    return debug_points(c, frame, level, false, xys, cap, n_out);
}
int lpslam_orb_debug_selected(lpslam_orb_ctx* c, uint32_t frame, uint32_t level, int32_t* xys, uint32_t cap, uint32_t* n_out) {
    return debug_points(c, frame, level, true, xys, cap, n_out);
}

int lpslam_orb_enable_timing(lpslam_orb_ctx* c, int on) {
    LPS_ARG(c, "null context");
    c->timing = on != 0;
    return LPSLAM_OK;
}

int lpslam_orb_stage_times(lpslam_orb_ctx* c, float* ms, uint32_t cap) {
    LPS_ARG(c && ms, "null argument");
    if (!c->timing || !c->has_result) {
        set_error("timing not enabled or no extraction has run");
        return LPSLAM_ERR_STATE;
    }
    LPS_CUDA(cudaSetDevice(c->device));
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    // events: 7 = before H2D, 0..5 = pipeline stage boundaries, 6 = after D2H
    float t[7] = {0, 0, 0, 0, 0, 0, 0};
    cudaEventElapsedTime(&t[0], c->ev[7], c->ev[0]);
    for (int i = 0; i < 5; ++i) cudaEventElapsedTime(&t[1 + i], c->ev[i], c->ev[i + 1]);
    cudaEventElapsedTime(&t[6], c->ev[5], c->ev[6]);
    cudaGetLastError();
    for (uint32_t i = 0; i < 7 && i < cap; ++i) ms[i] = t[i];
    return 7;
}

}  // extern "C"
