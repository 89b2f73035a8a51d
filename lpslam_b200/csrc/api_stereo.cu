// C ABI — stereo entry points (include/lpslam_cuda.h): fused stereo frames on an extractor context, and the standalone
// match::stereo drop-in that takes host pyramids (match/stereo.h:18-29).
#include "ctx.h"

#include <cstdlib>

#include <algorithm>
#include <cstring>
#include <vector>

using namespace lps;

extern "C" {

int lpslam_stereo_frames_dev(lpslam_orb_ctx* c, const uint8_t* imgs_dev, uint32_t n_pairs, int width, int height,
                             size_t pitch, float focal_x_baseline, float true_baseline) {
    LPS_ARG(c, "null context");
    LPS_ARG(n_pairs >= 1 && 2 * n_pairs <= c->max_batch, "2 * n_pairs exceeds the context's max_batch");
    const int rc = lpslam_orb_extract_batch_dev(c, imgs_dev, 2 * n_pairs, width, height, pitch);
    if (rc) return rc;
    c->stereo.n_pairs = (int)n_pairs;
    c->stereo.focal_x_baseline = focal_x_baseline;
    c->stereo.true_baseline = true_baseline;
    launch_stereo(c->stereo, c->stream);
    if (c->timing) cudaEventRecord(c->ev[5], c->stream);  // the describe stage then includes the matcher
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

int lpslam_stereo_result_dev(lpslam_orb_ctx* c, const float** xr, const float** dep) {
    LPS_ARG(c && xr && dep, "null argument");
    if (!c->has_result) {
        set_error("no extraction has run on this context");
        return LPSLAM_ERR_STATE;
    }
    *xr = c->stereo.x_right;
    *dep = c->stereo.depth;
    return LPSLAM_OK;
}

int lpslam_stereo_frames(lpslam_orb_ctx* c, const uint8_t* const* imgs_l, const uint8_t* const* imgs_r, uint32_t n_pairs,
                         int width, int height, size_t pitch, float focal_x_baseline, float true_baseline,
                         lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets, float* stereo_x_right,
                         float* depths) {
    LPS_ARG(c && imgs_l && imgs_r && offsets && kp_out && desc_out && stereo_x_right && depths, "null argument");
    LPS_ARG(n_pairs >= 1 && 2 * n_pairs <= c->max_batch, "2 * n_pairs exceeds the context's max_batch");
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width, "bad image geometry");
    LPS_CUDA(cudaSetDevice(c->device));
    int rc = c->configure(width, height);
    if (rc) return rc;
    const LevelGeom& g0 = c->lv[0];
    for (uint32_t p = 0; p < n_pairs; ++p) {
        LPS_ARG(imgs_l[p] && imgs_r[p], "null image pointer");
        LPS_CUDA(cudaMemcpy2DAsync(g0.img + (2 * p) * g0.frame_stride, g0.pitch, imgs_l[p], pitch, width, height,
                                   cudaMemcpyHostToDevice, c->stream));
        LPS_CUDA(cudaMemcpy2DAsync(g0.img + (2 * p + 1) * g0.frame_stride, g0.pitch, imgs_r[p], pitch, width, height,
                                   cudaMemcpyHostToDevice, c->stream));
    }
    const uint32_t nf = 2 * n_pairs;
    // the live tracker's call pattern (one pair per call, data/frame.cc:55-93) is launch-bound: ~20 small kernels. They are
    // replayed as one CUDA graph from the third call of a setting on
    static const bool graphs_on = [] {
        const char* e = getenv("LPSLAM_CUDA_GRAPH");
        return !(e && e[0] == '0');
    }();
    const bool want_graph = graphs_on && n_pairs == 1 && !c->timing && !c->pair_graph_off;
    if (c->pair_graph && (!want_graph || c->pair_graph_fxb != focal_x_baseline || c->pair_graph_tb != true_baseline)) c->drop_pair_graph();
    if (want_graph && c->pair_graph) {
        LPS_CUDA(cudaGraphLaunch(c->pair_graph, c->stream));
        ++c->pair_graph_launches;
        c->last_batch = nf;
        c->has_result = true;
    }
    else {
        const bool capture = want_graph && c->pair_graph_warm >= 1;
        if (capture && cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
            cudaGetLastError();
            c->pair_graph_off = true;
        }
        const bool capturing = capture && !c->pair_graph_off;
        rc = lpslam_stereo_frames_dev(c, g0.img, n_pairs, width, height, g0.pitch, focal_x_baseline, true_baseline);
        cudaError_t ce = cudaSuccess;
        if (!rc)
            ce = cudaMemcpyAsync(c->h_frame_off, c->desc.frame_off, sizeof(uint32_t) * (c->max_batch + 2), cudaMemcpyDeviceToHost, c->stream);
        if (capturing) {
            cudaGraph_t g = nullptr;
            const cudaError_t end = cudaStreamEndCapture(c->stream, &g);
            if (!rc && ce == cudaSuccess && end == cudaSuccess && g && cudaGraphInstantiate(&c->pair_graph, g, 0) == cudaSuccess) {
                c->pair_graph_fxb = focal_x_baseline;
                c->pair_graph_tb = true_baseline;
            }
            else {  // not capturable on this driver / configuration: stay with plain launches for this context
                cudaGetLastError();
                c->pair_graph = nullptr;
                c->pair_graph_off = true;
            }
            if (g) cudaGraphDestroy(g);
            if (c->pair_graph_off) {  // nothing ran: do it launch by launch
                rc = lpslam_stereo_frames_dev(c, g0.img, n_pairs, width, height, g0.pitch, focal_x_baseline, true_baseline);
                ce = rc ? cudaSuccess
                        : cudaMemcpyAsync(c->h_frame_off, c->desc.frame_off, sizeof(uint32_t) * (c->max_batch + 2), cudaMemcpyDeviceToHost,
                                          c->stream);
            }
            else {
                LPS_CUDA(cudaGraphLaunch(c->pair_graph, c->stream));
                ++c->pair_graph_launches;
            }
        }
        if (rc) return rc;
        LPS_CUDA(ce);
        if (want_graph) ++c->pair_graph_warm;
    }
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    const uint32_t total = c->h_frame_off[nf];
    for (uint32_t f = 0; f <= nf; ++f) offsets[f] = c->h_frame_off[f];
    if (c->h_frame_off[c->max_batch + 1] != 0 || total > cap) {
        set_error("output capacity too small for the extracted keypoints");
        return LPSLAM_ERR_CAPACITY;
    }
    if (total) {
        LPS_CUDA(cudaMemcpyAsync(kp_out, c->desc.kp, sizeof(lpslam_keypoint) * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
        LPS_CUDA(cudaMemcpyAsync(desc_out, c->desc.desc, 32 * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
        LPS_CUDA(cudaMemcpyAsync(stereo_x_right, c->stereo.x_right, 4 * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
        LPS_CUDA(cudaMemcpyAsync(depths, c->stereo.depth, 4 * (size_t)total, cudaMemcpyDeviceToHost, c->stream));
    }
    LPS_CUDA(cudaStreamSynchronize(c->stream));
    return LPSLAM_OK;
}

namespace {

// packed results of a finished batch -> the slot's own arrays (exact sizes are only known on the device)
__global__ void k_copy_results(const lpslam_keypoint* __restrict__ kp, const uint8_t* __restrict__ desc, const float* __restrict__ xr,
                               const float* __restrict__ dep, const uint32_t* __restrict__ off, uint32_t nf, uint32_t n_off,
                               uint32_t out_cap, lpslam_keypoint* __restrict__ kp_o, uint8_t* __restrict__ desc_o, float* __restrict__ xr_o,
                               float* __restrict__ dep_o, uint32_t* __restrict__ off_o) {
    // off[nf] is the raw carry: beyond out_cap it only says by how much the batch overflowed (collect() then reports
    // LPSLAM_ERR_CAPACITY from the overflow word); the slot arrays hold out_cap rows, as the packed arrays k_describe wrote
    const uint32_t total = min(off[nf], out_cap);
    const uint32_t stride = gridDim.x * blockDim.x, i0 = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t i = i0; i < n_off; i += stride) off_o[i] = off[i];
    const uint32_t* ks = reinterpret_cast<const uint32_t*>(kp);
    uint32_t* kd = reinterpret_cast<uint32_t*>(kp_o);
    for (uint32_t i = i0; i < total * 7u; i += stride) kd[i] = ks[i];
    const uint4* ds = reinterpret_cast<const uint4*>(desc);
    uint4* dd = reinterpret_cast<uint4*>(desc_o);
    for (uint32_t i = i0; i < total * 2u; i += stride) dd[i] = ds[i];
    for (uint32_t i = i0; i < total; i += stride) {
        xr_o[i] = xr[i];
        dep_o[i] = dep[i];
    }
}

int ensure_async(lpslam_orb_ctx* c, int width, int height) {
    const size_t need = (size_t)c->max_batch * width * height;
    if (c->async_ready && c->landing_bytes >= need) return LPSLAM_OK;
    c->release_async();
    int least = 0, greatest = 0;
    cudaDeviceGetStreamPriorityRange(&least, &greatest);
    LPS_CUDA(cudaStreamCreateWithPriority(&c->copy_in, cudaStreamNonBlocking, greatest));
    LPS_CUDA(cudaStreamCreateWithPriority(&c->copy_out, cudaStreamNonBlocking, greatest));
    for (auto& sl : c->slot) {
        LPS_CUDA(cudaMalloc(&sl.landing, need));
        LPS_CUDA(cudaMalloc(&sl.kp, sizeof(lpslam_keypoint) * (size_t)c->out_cap));
        LPS_CUDA(cudaMalloc(&sl.desc, 32 * (size_t)c->out_cap));
        LPS_CUDA(cudaMalloc(&sl.xr, 4 * (size_t)c->out_cap));
        LPS_CUDA(cudaMalloc(&sl.dep, 4 * (size_t)c->out_cap));
        LPS_CUDA(cudaMalloc(&sl.off_dev, 4 * (size_t)(c->max_batch + 2)));
        LPS_CUDA(cudaHostAlloc((void**)&sl.h_off, 4 * (size_t)(c->max_batch + 2), cudaHostAllocDefault));
        LPS_CUDA(cudaEventCreateWithFlags(&sl.ev_in, cudaEventDisableTiming));
        LPS_CUDA(cudaEventCreateWithFlags(&sl.ev_landing_free, cudaEventDisableTiming));
        LPS_CUDA(cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
    }
    c->landing_bytes = need;
    c->async_ready = true;
    return LPSLAM_OK;
}

}  // namespace

namespace {

// frames[k] = host image of frame k (stereo: left / right interleaved). n_pairs == 0: extraction only.
int submit_frames(lpslam_orb_ctx* c, const uint8_t* const* frames, uint32_t nf, uint32_t n_pairs, int width, int height, size_t pitch,
                  float focal_x_baseline, float true_baseline, uint64_t* ticket) {
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width, "bad image geometry");
    LPS_CUDA(cudaSetDevice(c->device));
    int rc = c->configure(width, height);
    if (rc) return rc;
    if ((rc = ensure_async(c, width, height))) return rc;
    lpslam_orb_ctx::AsyncSlot& sl = c->slot[c->seq & 1];
    if (sl.busy) {
        set_error("both in-flight slots are taken: collect the oldest ticket first");
        return LPSLAM_ERR_STATE;
    }
    for (uint32_t k = 0; k < nf; ++k) LPS_ARG(frames[k], "null image pointer");
    // images -> landing buffer (dense rows) on the copy-in stream, once the previous user of the buffer has consumed it
    if (sl.landing_used) LPS_CUDA(cudaStreamWaitEvent(c->copy_in, sl.ev_landing_free, 0));
    const size_t img = (size_t)width * height;
    uint32_t f = 0;
    while (f < nf) {  // images that follow one another in host memory go out as one copy
        uint32_t g = f + 1;
        if (pitch == (size_t)width) {
            while (g < nf && frames[g] == frames[g - 1] + img) ++g;
            LPS_CUDA(cudaMemcpyAsync(sl.landing + f * img, frames[f], (g - f) * img, cudaMemcpyHostToDevice, c->copy_in));
        }
        else
            LPS_CUDA(cudaMemcpy2DAsync(sl.landing + f * img, width, frames[f], pitch, width, height, cudaMemcpyHostToDevice, c->copy_in));
        f = g;
    }
    LPS_CUDA(cudaEventRecord(sl.ev_in, c->copy_in));
    LPS_CUDA(cudaStreamWaitEvent(c->stream, sl.ev_in, 0));
    launch_copy_level0(sl.landing, width, img, c->lv[0], (int)nf, c->stream);
    LPS_CUDA(cudaEventRecord(sl.ev_landing_free, c->stream));
    sl.landing_used = true;
    if (n_pairs)
        rc = lpslam_stereo_frames_dev(c, c->lv[0].img, n_pairs, width, height, c->lv[0].pitch, focal_x_baseline, true_baseline);
    else
        rc = lpslam_orb_extract_batch_dev(c, c->lv[0].img, nf, width, height, c->lv[0].pitch);
    if (rc) return rc;
    // (extraction only: the stereo arrays are copied too, rows of an earlier stereo call or never-read scratch)
    k_copy_results<<<256, 256, 0, c->stream>>>(c->desc.kp, c->desc.desc, c->stereo.x_right, c->stereo.depth, c->desc.frame_off, nf,
                                               c->max_batch + 2, c->out_cap, sl.kp, sl.desc, sl.xr, sl.dep, sl.off_dev);
    LPS_CUDA(cudaMemcpyAsync(sl.h_off, sl.off_dev, 4 * (size_t)(c->max_batch + 2), cudaMemcpyDeviceToHost, c->stream));
    LPS_CUDA(cudaEventRecord(sl.ev_done, c->stream));
    sl.busy = true;
    sl.n_pairs = n_pairs;
    sl.n_frames = nf;
    *ticket = c->seq++;
    return LPSLAM_OK;
}

int collect_frames(lpslam_orb_ctx* c, uint64_t ticket, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap, uint32_t* offsets,
                   float* stereo_x_right, float* depths) {
    LPS_ARG(c->async_ready && ticket < c->seq && ticket + 2 >= c->seq, "unknown ticket");
    lpslam_orb_ctx::AsyncSlot& sl = c->slot[ticket & 1];
    if (!sl.busy) {
        set_error("ticket already collected");
        return LPSLAM_ERR_STATE;
    }
    LPS_CUDA(cudaSetDevice(c->device));
    LPS_CUDA(cudaEventSynchronize(sl.ev_done));
    sl.busy = false;
    const uint32_t nf = sl.n_frames, total = sl.h_off[nf];
    for (uint32_t f = 0; f <= nf; ++f) offsets[f] = sl.h_off[f];
    if (sl.h_off[c->max_batch + 1] != 0 || total > cap) {
        set_error("output capacity too small for the extracted keypoints");
        return LPSLAM_ERR_CAPACITY;
    }
    if (total) {
        LPS_CUDA(cudaMemcpyAsync(kp_out, sl.kp, sizeof(lpslam_keypoint) * (size_t)total, cudaMemcpyDeviceToHost, c->copy_out));
        LPS_CUDA(cudaMemcpyAsync(desc_out, sl.desc, 32 * (size_t)total, cudaMemcpyDeviceToHost, c->copy_out));
        if (stereo_x_right) LPS_CUDA(cudaMemcpyAsync(stereo_x_right, sl.xr, 4 * (size_t)total, cudaMemcpyDeviceToHost, c->copy_out));
        if (depths) LPS_CUDA(cudaMemcpyAsync(depths, sl.dep, 4 * (size_t)total, cudaMemcpyDeviceToHost, c->copy_out));
        LPS_CUDA(cudaStreamSynchronize(c->copy_out));
    }
    return LPSLAM_OK;
}

}  // namespace

int lpslam_stereo_frames_submit(lpslam_orb_ctx* c, const uint8_t* const* imgs_l, const uint8_t* const* imgs_r, uint32_t n_pairs, int width,
                                int height, size_t pitch, float focal_x_baseline, float true_baseline, uint64_t* ticket) {
    LPS_ARG(c && imgs_l && imgs_r && ticket, "null argument");
    LPS_ARG(n_pairs >= 1 && 2 * n_pairs <= c->max_batch, "2 * n_pairs exceeds the context's max_batch");
    std::vector<const uint8_t*> frames(2 * (size_t)n_pairs);
    for (uint32_t p = 0; p < n_pairs; ++p) {
        frames[2 * p] = imgs_l[p];
        frames[2 * p + 1] = imgs_r[p];
    }
    return submit_frames(c, frames.data(), 2 * n_pairs, n_pairs, width, height, pitch, focal_x_baseline, true_baseline, ticket);
}

int lpslam_orb_extract_batch_submit(lpslam_orb_ctx* c, const uint8_t* const* imgs, uint32_t n_frames, int width, int height, size_t pitch,
                                    uint64_t* ticket) {
    LPS_ARG(c && imgs && ticket, "null argument");
    LPS_ARG(n_frames >= 1 && n_frames <= c->max_batch, "n_frames exceeds the context's max_batch");
    return submit_frames(c, imgs, n_frames, 0, width, height, pitch, 0.f, 0.f, ticket);
}

int lpslam_orb_extract_batch_collect(lpslam_orb_ctx* c, uint64_t ticket, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap,
                                     uint32_t* offsets) {
    LPS_ARG(c && offsets && kp_out && desc_out, "null argument");
    return collect_frames(c, ticket, kp_out, desc_out, cap, offsets, nullptr, nullptr);
}

int lpslam_stereo_frames_collect(lpslam_orb_ctx* c, uint64_t ticket, lpslam_keypoint* kp_out, uint8_t* desc_out, uint32_t cap,
                                 uint32_t* offsets, float* stereo_x_right, float* depths) {
    LPS_ARG(c && offsets && kp_out && desc_out && stereo_x_right && depths, "null argument");
    return collect_frames(c, ticket, kp_out, desc_out, cap, offsets, stereo_x_right, depths);
}

int lpslam_match_stereo(const uint8_t* const* pyr_l, const uint8_t* const* pyr_r, const int* level_w, const int* level_h,
                        const size_t* pitch_l, const size_t* pitch_r, uint32_t num_levels, const lpslam_keypoint* kp_l,
                        uint32_t nl, const lpslam_keypoint* kp_r, uint32_t nr, const uint8_t* desc_l, const uint8_t* desc_r,
                        const float* scale_factors, const float* inv_scale_factors, float focal_x_baseline,
                        float true_baseline, float* stereo_x_right, float* depths, uint32_t* n_matches) {
    LPS_ARG(pyr_l && pyr_r && level_w && level_h && pitch_l && pitch_r && scale_factors && inv_scale_factors, "null argument");
    LPS_ARG(num_levels >= 1 && num_levels <= (uint32_t)kMaxLevels, "num_levels out of range (1..16)");
    LPS_ARG(nl == 0 || (kp_l && desc_l && stereo_x_right && depths), "null left-eye argument");
    LPS_ARG(nr == 0 || (kp_r && desc_r), "null right-eye argument");
    if (n_matches) *n_matches = 0;
    if (nl == 0) return LPSLAM_OK;
    for (uint32_t i = 0; i < nl; ++i) LPS_ARG(kp_l[i].octave >= 0 && (uint32_t)kp_l[i].octave < num_levels, "left octave out of range");
    for (uint32_t i = 0; i < nr; ++i) LPS_ARG(kp_r[i].octave >= 0 && (uint32_t)kp_r[i].octave < num_levels, "right octave out of range");

    LevelGeom lv[kMaxLevels];
    memset(lv, 0, sizeof lv);
    size_t img_bytes = 0;
    std::vector<size_t> off(num_levels);
    for (uint32_t l = 0; l < num_levels; ++l) {
        lv[l].w = level_w[l];
        lv[l].h = level_h[l];
        lv[l].pitch = (int)align_up(level_w[l], 64);
        lv[l].frame_stride = align_up((size_t)lv[l].pitch * level_h[l], 256);
        lv[l].scale = scale_factors[l];
        lv[l].inv_scale = inv_scale_factors[l];
        off[l] = img_bytes;
        img_bytes += 2 * lv[l].frame_stride;
    }
    const uint32_t rows = nl + nr;
    // one packed copy in (both pyramids, level table, keypoints, descriptors, offsets), one copy out (x_right, depth)
    ArenaBump ar;
    const size_t o_img = ar.take(img_bytes), o_lv = ar.take(sizeof lv), o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)rows);
    const size_t o_desc = ar.take(32 * (size_t)rows), o_off = ar.take(12), in_end = ar.off;
    const size_t o_rec = ar.take(8 * (size_t)rows), o_corr = ar.take(4 * (size_t)rows), o_perm = ar.take(16 * (size_t)rows);
    const size_t o_rows = ar.take(4 * (size_t)(level_h[0] + 1)), o_xr = ar.take(4 * (size_t)rows), o_dep = ar.take(4 * (size_t)rows);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    do {
        for (uint32_t l = 0; l < num_levels; ++l) {
            lv[l].img = dev + o_img + off[l];
            for (int eye = 0; eye < 2; ++eye) {
                const uint8_t* src = eye ? pyr_r[l] : pyr_l[l];
                const size_t sp = eye ? pitch_r[l] : pitch_l[l];
                uint8_t* dst = cs.host + o_img + off[l] + eye * lv[l].frame_stride;
                if (sp == (size_t)lv[l].pitch) memcpy(dst, src, sp * (size_t)(lv[l].h - 1) + lv[l].w);
                else
                    for (int y = 0; y < lv[l].h; ++y) memcpy(dst + (size_t)y * lv[l].pitch, src + (size_t)y * sp, lv[l].w);
            }
        }
        const uint32_t h_off[3] = {0, nl, rows};
        cs.put(o_lv, lv, sizeof lv);
        cs.put(o_kp, kp_l, sizeof(lpslam_keypoint) * (size_t)nl);
        cs.put(o_kp + sizeof(lpslam_keypoint) * (size_t)nl, kp_r, sizeof(lpslam_keypoint) * (size_t)nr);
        cs.put(o_desc, desc_l, 32 * (size_t)nl);
        cs.put(o_desc + 32 * (size_t)nl, desc_r, 32 * (size_t)nr);
        cs.put(o_off, h_off, sizeof h_off);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        StereoArgs a{};
        a.lv = (const LevelGeom*)(dev + o_lv);
        a.num_levels = (int)num_levels;
        a.n_pairs = 1;
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_desc;
        a.frame_off = (const uint32_t*)(dev + o_off);
        a.max_rows = rows;
        a.rec = (uint2*)(dev + o_rec);
        a.x_right = (float*)(dev + o_xr);
        a.depth = (float*)(dev + o_dep);
        a.corr = (int*)(dev + o_corr);
        a.perm = (uint4*)(dev + o_perm);
        a.row_start = (uint32_t*)(dev + o_rows);
        a.focal_x_baseline = focal_x_baseline;
        a.true_baseline = true_baseline;
        a.rows0 = level_h[0];
        launch_stereo(a, cs.st);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_xr, ar.off)) != cudaSuccess) break;
        memcpy(stereo_x_right, cs.out<float>(o_xr), 4 * (size_t)nl);
        memcpy(depths, cs.out<float>(o_dep), 4 * (size_t)nl);
        if (n_matches) {
            uint32_t n = 0;
            for (uint32_t i = 0; i < nl; ++i) n += !(depths[i] == -1.0f && stereo_x_right[i] == -1.0f);
            *n_matches = n;
        }
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_stereo", __FILE__, __LINE__);
    return LPSLAM_OK;
}

int lpslam_match_stereo_ctx(lpslam_orb_ctx* cl, lpslam_orb_ctx* cr, const lpslam_keypoint* kp_l, uint32_t nl, const lpslam_keypoint* kp_r,
                            uint32_t nr, const uint8_t* desc_l, const uint8_t* desc_r, float focal_x_baseline, float true_baseline,
                            float* stereo_x_right, float* depths, uint32_t* n_matches) {
    LPS_ARG(cl && cr, "null context");
    LPS_ARG(nl == 0 || (kp_l && desc_l && stereo_x_right && depths), "null left-eye argument");
    LPS_ARG(nr == 0 || (kp_r && desc_r), "null right-eye argument");
    if (n_matches) *n_matches = 0;
    if (!cl->has_result || !cr->has_result) {
        set_error("no extraction has run on one of the contexts");
        return LPSLAM_ERR_STATE;
    }
    LPS_ARG(cl->device == cr->device && cl->w == cr->w && cl->h == cr->h && cl->num_levels == cr->num_levels,
            "the two contexts must hold pyramids of the same geometry on the same device");
    if (nl == 0) return LPSLAM_OK;
    const uint32_t num_levels = cl->num_levels;
    for (uint32_t i = 0; i < nl; ++i) LPS_ARG(kp_l[i].octave >= 0 && (uint32_t)kp_l[i].octave < num_levels, "left octave out of range");
    for (uint32_t i = 0; i < nr; ++i) LPS_ARG(kp_r[i].octave >= 0 && (uint32_t)kp_r[i].octave < num_levels, "right octave out of range");
    LPS_CUDA(cudaSetDevice(cl->device));
    // the level table of the left context with "the next frame" = the same level of the right context: the kernels address the
    // right eye as img + frame_stride (pointer difference, modulo 2^64)
    LevelGeom lv[kMaxLevels];
    memset(lv, 0, sizeof lv);
    for (uint32_t l = 0; l < num_levels; ++l) {
        lv[l] = cl->lv[l];
        LPS_ARG(cl->lv[l].pitch == cr->lv[l].pitch, "level pitch differs between the contexts");
        lv[l].frame_stride = (size_t)(cr->lv[l].img - cl->lv[l].img);
    }
    const uint32_t rows = nl + nr;
    ArenaBump ar;
    const size_t o_lv = ar.take(sizeof lv), o_kp = ar.take(sizeof(lpslam_keypoint) * (size_t)rows);
    const size_t o_desc = ar.take(32 * (size_t)rows), o_off = ar.take(12), in_end = ar.off;
    const size_t o_rec = ar.take(8 * (size_t)rows), o_corr = ar.take(4 * (size_t)rows), o_perm = ar.take(16 * (size_t)rows);
    const size_t o_rows = ar.take(4 * (size_t)(cl->h + 1)), o_xr = ar.take(4 * (size_t)rows), o_dep = ar.take(4 * (size_t)rows);
    CallStage cs;
    LPS_CUDA(cs.open(ar.off));
    uint8_t* const dev = cs.dev;
    cudaError_t e = cudaSuccess;
    do {
        const uint32_t h_off[3] = {0, nl, rows};
        cs.put(o_lv, lv, sizeof lv);
        cs.put(o_kp, kp_l, sizeof(lpslam_keypoint) * (size_t)nl);
        cs.put(o_kp + sizeof(lpslam_keypoint) * (size_t)nl, kp_r, sizeof(lpslam_keypoint) * (size_t)nr);
        cs.put(o_desc, desc_l, 32 * (size_t)nl);
        cs.put(o_desc + 32 * (size_t)nl, desc_r, 32 * (size_t)nr);
        cs.put(o_off, h_off, sizeof h_off);
        if ((e = cs.upload(0, in_end)) != cudaSuccess) break;
        // the pyramids may still be in the making on the contexts' own streams (the enqueue-only and submit forms set
        // has_result without waiting): the call's stream runs after whatever they hold now
        if ((e = cudaEventRecord(cl->ev_reader, cl->stream)) != cudaSuccess) break;
        if ((e = cudaStreamWaitEvent(cs.st, cl->ev_reader, 0)) != cudaSuccess) break;
        if (cr != cl) {
            if ((e = cudaEventRecord(cr->ev_reader, cr->stream)) != cudaSuccess) break;
            if ((e = cudaStreamWaitEvent(cs.st, cr->ev_reader, 0)) != cudaSuccess) break;
        }
        StereoArgs a{};
        a.lv = (const LevelGeom*)(dev + o_lv);
        a.num_levels = (int)num_levels;
        a.n_pairs = 1;
        a.kp = (const lpslam_keypoint*)(dev + o_kp);
        a.desc = dev + o_desc;
        a.frame_off = (const uint32_t*)(dev + o_off);
        a.max_rows = rows;
        a.rec = (uint2*)(dev + o_rec);
        a.x_right = (float*)(dev + o_xr);
        a.depth = (float*)(dev + o_dep);
        a.corr = (int*)(dev + o_corr);
        a.perm = (uint4*)(dev + o_perm);
        a.row_start = (uint32_t*)(dev + o_rows);
        a.focal_x_baseline = focal_x_baseline;
        a.true_baseline = true_baseline;
        a.rows0 = cl->h;
        launch_stereo(a, cs.st);
        if ((e = cudaGetLastError()) != cudaSuccess) break;
        if ((e = cs.download(o_xr, ar.off)) != cudaSuccess) break;
        memcpy(stereo_x_right, cs.out<float>(o_xr), 4 * (size_t)nl);
        memcpy(depths, cs.out<float>(o_dep), 4 * (size_t)nl);
        if (n_matches) {
            uint32_t n = 0;
            for (uint32_t i = 0; i < nl; ++i) n += !(depths[i] == -1.0f && stereo_x_right[i] == -1.0f);
            *n_matches = n;
        }
    } while (false);
    if (e != cudaSuccess) return cuda_fail(e, "lpslam_match_stereo_ctx", __FILE__, __LINE__);
    return LPSLAM_OK;
}

}  // extern "C"
