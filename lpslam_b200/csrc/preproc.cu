// Image pre-processing ahead of the extractor, batched and device resident so that the result can be
// written straight into the extractor's level-0 buffer (lpslam_orb_input_dev):
//   k_gray        util::convert_to_grayscale (util/image_converter.cc:8-39) = cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY), 8-bit:
//                 (B 3735 + G 19235 + R 9798 + 2^14) >> 15
//   k_hist/k_lut/k_apply_lut   LpSlam::ImageProcessing::imadjust + stretchlim (src/Utils/ImageProcessing.h:258-370; the
//                 "AdjustIntensity" processor, src/Processor/AdjustIntensityProcessor.h:17-40) for 8UC1: 256-bin histogram,
//                 1 % / 99 % limits with the reference's loops, value map = cvRound(float(v) * float(alpha) + float(beta))
//                 saturated (Mat::convertTo: float multiply then float add, no fused multiply-add)
//   k_remap       LpSlam::ImageProcessing::Undistort (ImageProcessing.h:245-249) = cv::remap(INTER_LINEAR, BORDER_CONSTANT 0)
//                 with cached CV_32FC1 maps: 5-bit sub-pixel positions, weights (32-fy)(32-fx)*32 ..., (sum + 2^14) >> 15
// All three are HBM-bound streaming kernels: 128-bit / 32-bit coalesced loads and stores, no shared-memory staging needed
// (every input byte is used once; the remap gather has 2-D locality that L1/L2 absorb).
#include "common.cuh"

#include <algorithm>
#include <mutex>

namespace lps {
namespace {

__global__ void __launch_bounds__(256) k_gray(const uint8_t* __restrict__ src, int w, int h, size_t pitch, size_t stride, int ch,
                                              int order, uint8_t* __restrict__ dst, size_t dpitch, size_t dstride, int vec) {
    // a thread converts 16 consecutive pixels: 3 (or 4) 128-bit loads, one 128-bit store
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x0 >= w || y >= h) return;
    const uint8_t* s = src + blockIdx.z * stride + (size_t)y * pitch + (size_t)x0 * ch;
    uint8_t* d = dst + blockIdx.z * dstride + (size_t)y * dpitch + x0;
    const int ir = order == 0 ? 0 : 2, ib = 2 - ir;
    if (vec && x0 + 16 <= w) {
        uint32_t px[16];  // one pixel per word, channels in bytes 0..2
        if (ch == 4) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(s) + q);
                px[4 * q] = v.x; px[4 * q + 1] = v.y; px[4 * q + 2] = v.z; px[4 * q + 3] = v.w;
            }
        }
        else {
            uint32_t wd[12];
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(s) + q);
                wd[4 * q] = v.x; wd[4 * q + 1] = v.y; wd[4 * q + 2] = v.z; wd[4 * q + 3] = v.w;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {  // 4 pixels = 3 words
                const uint32_t a = wd[3 * q], b = wd[3 * q + 1], c = wd[3 * q + 2];
                px[4 * q] = a;                              // bytes 0..2
                px[4 * q + 1] = __byte_perm(a, b, 0x6543);  // bytes 3..5
                px[4 * q + 2] = __byte_perm(b, c, 0x5432);  // bytes 6..8
                px[4 * q + 3] = c >> 8;                     // bytes 9..11
            }
        }
        uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const uint32_t r = (px[k] >> (8 * ir)) & 255u, g = (px[k] >> 8) & 255u, b = (px[k] >> (8 * ib)) & 255u;
            o[k >> 2] |= ((b * 3735u + g * 19235u + r * 9798u + (1u << 14)) >> 15) << (8 * (k & 3));
        }
        *reinterpret_cast<uint4*>(d) = make_uint4(o[0], o[1], o[2], o[3]);
    }
    else {
        for (int k = 0; k < 16 && x0 + k < w; ++k) {
            const uint8_t* p = s + k * ch;
            d[k] = (uint8_t)((p[ib] * 3735u + p[1] * 19235u + p[ir] * 9798u + (1u << 14)) >> 15);
        }
    }
}

__global__ void __launch_bounds__(256) k_hist(const uint8_t* __restrict__ src, int w, int h, size_t pitch, size_t stride,
                                              uint32_t* __restrict__ hist, int vec) {
    __shared__ uint32_t s_h[256];
    s_h[threadIdx.x] = 0;
    __syncthreads();
    const uint8_t* img = src + blockIdx.y * stride;
    const int per_row = vec ? (w + 15) / 16 : w;
    const long total = (long)per_row * h;
    for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += 256L * gridDim.x) {
        const int y = (int)(i / per_row), q = (int)(i % per_row);
        const uint8_t* row = img + (size_t)y * pitch;
        if (vec) {
            const int x = 16 * q;
            if (x + 16 <= w) {
                const uint4 v = __ldg(reinterpret_cast<const uint4*>(row + x));
                const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    atomicAdd(&s_h[ws[k] & 255u], 1u);
                    atomicAdd(&s_h[(ws[k] >> 8) & 255u], 1u);
                    atomicAdd(&s_h[(ws[k] >> 16) & 255u], 1u);
                    atomicAdd(&s_h[ws[k] >> 24], 1u);
                }
            }
            else
                for (int k = x; k < w; ++k) atomicAdd(&s_h[row[k]], 1u);
        }
        else
            atomicAdd(&s_h[row[q]], 1u);
    }
    __syncthreads();
    if (s_h[threadIdx.x]) atomicAdd(&hist[blockIdx.y * 256 + threadIdx.x], s_h[threadIdx.x]);
}

struct AdjustArgs {
    int auto_limits;
    double low_in, high_in, low_out, high_out;
    uint32_t pixels;
};

__global__ void __launch_bounds__(256) k_lut(const uint32_t* __restrict__ hist, AdjustArgs a, uint8_t* __restrict__ lut,
                                             double* __restrict__ limits) {
    __shared__ float s_ab[2];
    const uint32_t* hs = hist + blockIdx.x * 256;
    if (threadIdx.x == 0) {
        double lo = 0.0, hi = 1.0;
        if (a.auto_limits) {
            // ImageProcessing.h:258-288: 'unsigned sum += float' is evaluated in float
            const unsigned low_count = (unsigned)__dmul_rn(0.01, (double)a.pixels);
            unsigned sum = 0;
            for (unsigned i = 0; i < 256; ++i) {
                if (sum >= low_count) {
                    lo = (double)i;
                    break;
                }
                sum = (unsigned)__fadd_rn((float)sum, (float)hs[i]);
            }
            const unsigned high_count = (unsigned)__dmul_rn(1 - 0.99, (double)a.pixels);
            sum = 0;
            for (int i = 255; i >= 0; --i) {
                if (sum >= high_count) {
                    hi = (double)i;
                    break;
                }
                sum = (unsigned)__fadd_rn((float)sum, (float)hs[i]);
            }
            lo = __ddiv_rn(lo, 255.0);
            hi = __ddiv_rn(hi, 255.0);
        }
        else {
            lo = a.low_in;
            hi = a.high_in;
        }
        const double alpha = __ddiv_rn(__dsub_rn(a.high_out, a.low_out), __dsub_rn(hi, lo));
        const double beta = __dmul_rn(__dsub_rn(a.high_out, __dmul_rn(hi, alpha)), 255.0);
        s_ab[0] = (float)alpha;
        s_ab[1] = (float)beta;
        if (limits) {
            limits[2 * blockIdx.x] = lo;
            limits[2 * blockIdx.x + 1] = hi;
        }
    }
    __syncthreads();
    const float f = __fadd_rn(__fmul_rn((float)threadIdx.x, s_ab[0]), s_ab[1]);
    int r = 0;
    if (isfinite(f)) r = min(max(__float2int_rn(f), 0), 255);  // cvRound of a non-finite value saturates to 0 on the host
    lut[blockIdx.x * 256 + threadIdx.x] = (uint8_t)r;
}

__global__ void __launch_bounds__(256) k_apply_lut(const uint8_t* __restrict__ src, int w, int h, size_t pitch, size_t stride,
                                                   const uint8_t* __restrict__ lut, uint8_t* __restrict__ dst, size_t dpitch,
                                                   size_t dstride, int vec) {
    __shared__ uint8_t s_l[256];
    s_l[threadIdx.y * blockDim.x + threadIdx.x] = lut[blockIdx.z * 256 + threadIdx.y * blockDim.x + threadIdx.x];
    __syncthreads();
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x0 >= w || y >= h) return;
    const uint8_t* s = src + blockIdx.z * stride + (size_t)y * pitch + x0;
    uint8_t* d = dst + blockIdx.z * dstride + (size_t)y * dpitch + x0;
    if (vec && x0 + 16 <= w) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(s));
        const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
            o[k] = (uint32_t)s_l[ws[k] & 255u] | ((uint32_t)s_l[(ws[k] >> 8) & 255u] << 8) | ((uint32_t)s_l[(ws[k] >> 16) & 255u] << 16)
                   | ((uint32_t)s_l[ws[k] >> 24] << 24);
        *reinterpret_cast<uint4*>(d) = make_uint4(o[0], o[1], o[2], o[3]);
    }
    else
        for (int k = 0; k < 16 && x0 + k < w; ++k) d[k] = s_l[s[k]];
}

__global__ void __launch_bounds__(256) k_remap(const uint8_t* __restrict__ src, int sw, int sh, size_t spitch, size_t sstride,
                                               const float* __restrict__ mx, const float* __restrict__ my, size_t mpitch,
                                               uint8_t* __restrict__ dst, int dw, int dh, size_t dpitch, size_t dstride, int vec) {
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x0 >= dw || y >= dh) return;
    const uint8_t* s = src + blockIdx.z * sstride;
    uint8_t* d = dst + blockIdx.z * dstride + (size_t)y * dpitch + x0;
    float fxs[4], fys[4];
    const bool full = x0 + 4 <= dw;
    if (vec && full) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(mx + (size_t)y * mpitch + x0));
        const float4 b = __ldg(reinterpret_cast<const float4*>(my + (size_t)y * mpitch + x0));
        fxs[0] = a.x; fxs[1] = a.y; fxs[2] = a.z; fxs[3] = a.w;
        fys[0] = b.x; fys[1] = b.y; fys[2] = b.z; fys[3] = b.w;
    }
    else {
        for (int k = 0; k < 4; ++k) {
            const int x = min(x0 + k, dw - 1);
            fxs[k] = __ldg(mx + (size_t)y * mpitch + x);
            fys[k] = __ldg(my + (size_t)y * mpitch + x);
        }
    }
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int sx = __float2int_rn(__fmul_rn(fxs[k], 32.0f)), sy = __float2int_rn(__fmul_rn(fys[k], 32.0f));
        const int ix = min(max(sx >> 5, -32768), 32767), iy = min(max(sy >> 5, -32768), 32767);
        const int fx = sx & 31, fy = sy & 31;
        // taps outside of the image read as 0 (BORDER_CONSTANT): predicated loads from one running address per row
        // (no clamping, no branch between the map fetch and the 16 pixel loads of a thread: all are in flight together)
        const bool x0ok = (unsigned)ix < (unsigned)sw, x1ok = (unsigned)(ix + 1) < (unsigned)sw;
        const bool y0ok = (unsigned)iy < (unsigned)sh, y1ok = (unsigned)(iy + 1) < (unsigned)sh;
        const uint8_t* r0 = s + (ptrdiff_t)(iy * (int)spitch + ix);
        const uint8_t* r1 = r0 + spitch;
        uint32_t p00 = 0, p01 = 0, p10 = 0, p11 = 0;
        if (x0ok && y0ok) p00 = __ldg(r0);
        if (x1ok && y0ok) p01 = __ldg(r0 + 1);
        if (x0ok && y1ok) p10 = __ldg(r1);
        if (x1ok && y1ok) p11 = __ldg(r1 + 1);
        // weights (32-fy)(32-fx)*32 ... and (sum + 2^14) >> 15 == ((top (32-fy) + bottom fy) + 2^9) >> 10 with
        // top = p00 (32-fx) + p01 fx: all integers, the common factor 32 cancels exactly
        const uint32_t ax = 32 - fx, top = p00 * ax + p01 * fx, bot = p10 * ax + p11 * fx;
        out |= ((top * (32 - fy) + bot * fy + (1u << 9)) >> 10) << (8 * k);
    }
    if (vec && full)
        *reinterpret_cast<uint32_t*>(d) = out;
    else
        for (int k = 0; k < 4 && x0 + k < dw; ++k) d[k] = (uint8_t)(out >> (8 * k));
}

inline bool aligned(const void* p, size_t a) { return reinterpret_cast<uintptr_t>(p) % a == 0; }

}  // namespace
}  // namespace lps

using namespace lps;

extern "C" {

int lpslam_gray_dev(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride, int channels, int order,
                    uint32_t n_images, uint8_t* dst_dev, size_t dst_pitch, size_t dst_frame_stride, void* stream) {
    LPS_ARG(src_dev && dst_dev, "null image pointer");
    LPS_ARG(width > 0 && height > 0 && (channels == 3 || channels == 4) && (order == 0 || order == 1), "bad image geometry / channel order");
    LPS_ARG(pitch >= (size_t)width * channels && dst_pitch >= (size_t)width && n_images >= 1, "bad pitch / image count");
    const int vec = aligned(src_dev, 16) && pitch % 16 == 0 && frame_stride % 16 == 0 && aligned(dst_dev, 16) && dst_pitch % 16 == 0
                    && dst_frame_stride % 16 == 0;
    dim3 blk(32, 8), grd((width + 32 * 16 - 1) / (32 * 16), (height + 7) / 8, n_images);
    k_gray<<<grd, blk, 0, (cudaStream_t)stream>>>(src_dev, width, height, pitch, frame_stride, channels, order, dst_dev, dst_pitch,
                                                  dst_frame_stride, vec);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

size_t lpslam_imadjust_scratch_bytes(uint32_t n_images) { return (size_t)n_images * (256 * 4 + 256 + 16); }

int lpslam_imadjust_dev(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride, uint32_t n_images,
                        int auto_limits, double low_in, double high_in, double low_out, double high_out, uint8_t* dst_dev,
                        size_t dst_pitch, size_t dst_frame_stride, void* scratch_dev, void* stream) {
    LPS_ARG(src_dev && dst_dev && scratch_dev, "null pointer");
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width && dst_pitch >= (size_t)width && n_images >= 1, "bad image geometry");
    LPS_ARG(aligned(scratch_dev, 8), "scratch must be 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    // scratch: limits double[2n] | hist u32[256 n] | lut u8[256 n]
    double* limits = (double*)scratch_dev;
    uint32_t* hist = (uint32_t*)((uint8_t*)scratch_dev + 16 * (size_t)n_images);
    uint8_t* lut = (uint8_t*)(hist + 256 * (size_t)n_images);
    const int vec = aligned(src_dev, 16) && pitch % 16 == 0 && frame_stride % 16 == 0;
    const int vec_out = vec && aligned(dst_dev, 16) && dst_pitch % 16 == 0 && dst_frame_stride % 16 == 0;
    if (auto_limits) {
        LPS_CUDA(cudaMemsetAsync(hist, 0, 4 * 256 * (size_t)n_images, st));
        const long units = (long)(vec ? (width + 15) / 16 : width) * height;
        const int blocks = (int)std::min<long>(std::max<long>((units + 256 * 8 - 1) / (256 * 8), 1), 148 * 4);
        k_hist<<<dim3(blocks, n_images), 256, 0, st>>>(src_dev, width, height, pitch, frame_stride, hist, vec);
    }
    AdjustArgs a{auto_limits, low_in, high_in, low_out, high_out, (uint32_t)(width * height)};
    k_lut<<<n_images, 256, 0, st>>>(hist, a, lut, limits);
    dim3 blk(64, 4), grd((width + 64 * 16 - 1) / (64 * 16), (height + 3) / 4, n_images);
    k_apply_lut<<<grd, blk, 0, st>>>(src_dev, width, height, pitch, frame_stride, lut, dst_dev, dst_pitch, dst_frame_stride, vec_out);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

int lpslam_remap_linear_dev(const uint8_t* src_dev, int src_width, int src_height, size_t src_pitch, size_t src_frame_stride,
                            const float* map_x_dev, const float* map_y_dev, size_t map_pitch, uint8_t* dst_dev, int dst_width,
                            int dst_height, size_t dst_pitch, size_t dst_frame_stride, uint32_t n_images, void* stream) {
    LPS_ARG(src_dev && dst_dev && map_x_dev && map_y_dev, "null pointer");
    LPS_ARG(src_width > 0 && src_height > 0 && dst_width > 0 && dst_height > 0 && n_images >= 1, "bad image geometry");
    LPS_ARG(src_pitch >= (size_t)src_width && dst_pitch >= (size_t)dst_width && map_pitch >= (size_t)dst_width, "bad pitch");
    const int vec = aligned(map_x_dev, 16) && aligned(map_y_dev, 16) && map_pitch % 4 == 0 && aligned(dst_dev, 4) && dst_pitch % 4 == 0
                    && dst_frame_stride % 4 == 0;
    dim3 blk(32, 8), grd((dst_width + 127) / 128, (dst_height + 7) / 8, n_images);
    k_remap<<<grd, blk, 0, (cudaStream_t)stream>>>(src_dev, src_width, src_height, src_pitch, src_frame_stride, map_x_dev, map_y_dev,
                                                   map_pitch, dst_dev, dst_width, dst_height, dst_pitch, dst_frame_stride, vec);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}

// ---- host-buffer single-image forms (what the reference call sites hand over: cv::Mat data pointers) ----
namespace {
struct DevBuf {
    uint8_t* p = nullptr;
    ~DevBuf() { scratch_free(p); }
};
}

int lpslam_gray(const uint8_t* src, int width, int height, size_t pitch, int channels, int order, uint8_t* dst, size_t dst_pitch) {
    LPS_ARG(src && dst && width > 0 && height > 0 && (channels == 3 || channels == 4), "bad arguments");
    LPS_ARG(pitch >= (size_t)width * channels && dst_pitch >= (size_t)width, "bad pitch");
    const size_t sp = align_up((size_t)width * channels, 16), dp = align_up((size_t)width, 16);
    DevBuf b;
    LPS_CUDA(scratch_alloc((void**)&b.p, (sp + dp) * height));
    LPS_CUDA(cudaMemcpy2D(b.p, sp, src, pitch, (size_t)width * channels, height, cudaMemcpyHostToDevice));
    int rc = lpslam_gray_dev(b.p, width, height, sp, sp * height, channels, order, 1, b.p + sp * height, dp, dp * height, nullptr);
    if (rc) return rc;
    LPS_CUDA(cudaMemcpy2D(dst, dst_pitch, b.p + sp * height, dp, width, height, cudaMemcpyDeviceToHost));
    return LPSLAM_OK;
}

int lpslam_imadjust(const uint8_t* src, int width, int height, size_t pitch, int auto_limits, double low_in, double high_in,
                    double low_out, double high_out, uint8_t* dst, size_t dst_pitch, double* limits_out) {
    LPS_ARG(src && dst && width > 0 && height > 0 && pitch >= (size_t)width && dst_pitch >= (size_t)width, "bad arguments");
    const size_t sp = align_up((size_t)width, 16);
    const size_t scratch = align_up(lpslam_imadjust_scratch_bytes(1), 16);
    DevBuf b;
    LPS_CUDA(scratch_alloc((void**)&b.p, scratch + 2 * sp * height));
    uint8_t* s = b.p + scratch;
    uint8_t* d = s + sp * height;
    LPS_CUDA(cudaMemcpy2D(s, sp, src, pitch, width, height, cudaMemcpyHostToDevice));
    int rc = lpslam_imadjust_dev(s, width, height, sp, sp * height, 1, auto_limits, low_in, high_in, low_out, high_out, d, sp,
                                 sp * height, b.p, nullptr);
    if (rc) return rc;
    LPS_CUDA(cudaMemcpy2D(dst, dst_pitch, d, sp, width, height, cudaMemcpyDeviceToHost));
    if (limits_out) LPS_CUDA(cudaMemcpy(limits_out, b.p, 16, cudaMemcpyDeviceToHost));
    return LPSLAM_OK;
}

int lpslam_remap_linear(const uint8_t* src, int src_width, int src_height, size_t src_pitch, const float* map_x, const float* map_y,
                        size_t map_pitch, uint8_t* dst, int dst_width, int dst_height, size_t dst_pitch) {
    LPS_ARG(src && dst && map_x && map_y && src_width > 0 && src_height > 0 && dst_width > 0 && dst_height > 0, "bad arguments");
    LPS_ARG(src_pitch >= (size_t)src_width && dst_pitch >= (size_t)dst_width && map_pitch >= (size_t)dst_width, "bad pitch");
    const size_t sp = align_up((size_t)src_width, 16), dp = align_up((size_t)dst_width, 16), mp = align_up((size_t)dst_width, 4);
    DevBuf b;
    const size_t o_d = sp * src_height, o_mx = align_up(o_d + dp * dst_height, 16), o_my = o_mx + 4 * mp * dst_height;
    LPS_CUDA(scratch_alloc((void**)&b.p, o_my + 4 * mp * dst_height));
    LPS_CUDA(cudaMemcpy2D(b.p, sp, src, src_pitch, src_width, src_height, cudaMemcpyHostToDevice));
    LPS_CUDA(cudaMemcpy2D(b.p + o_mx, 4 * mp, map_x, 4 * map_pitch, 4 * (size_t)dst_width, dst_height, cudaMemcpyHostToDevice));
    LPS_CUDA(cudaMemcpy2D(b.p + o_my, 4 * mp, map_y, 4 * map_pitch, 4 * (size_t)dst_width, dst_height, cudaMemcpyHostToDevice));
    int rc = lpslam_remap_linear_dev(b.p, src_width, src_height, sp, sp * src_height, (const float*)(b.p + o_mx), (const float*)(b.p + o_my),
                                     mp, b.p + o_d, dst_width, dst_height, dp, dp * dst_height, 1, nullptr);
    if (rc) return rc;
    LPS_CUDA(cudaMemcpy2D(dst, dst_pitch, b.p + o_d, dp, dst_width, dst_height, cudaMemcpyDeviceToHost));
    return LPSLAM_OK;
}

}  // extern "C"

// ---- imadjust without any allocation (lpslam_imadjust_dev_static) ----
// Some hosts of this library are not allowed to allocate device memory once they run: a camera callback of a real-time tracking
// thread (cudaMalloc takes driver locks and may synchronise the whole device, so its latency has no bound), or a plugin inside a
// process whose own pool reserved the device memory at start-up, where a later cudaMalloc simply fails. lpslam_imadjust_dev asks
// such a caller for a scratch buffer it cannot produce, and lpslam_imadjust allocates on its behalf. This variant keeps the
// histograms and value maps in an array that is part of the library image instead: kAdjustStaticBytes (80 KiB) of every GPU that
// loads liblpslam_cuda.so, from module load until the process ends.
constexpr size_t kAdjustPerImage = 256 * sizeof(uint32_t) + 256;  // hist u32[256] + lut u8[256]
constexpr size_t kAdjustStaticBytes = (size_t)LPSLAM_IMADJUST_STATIC_MAX_IMAGES * kAdjustPerImage;
// Layout: hist u32[256 x MAX] | lut u8[256 x MAX]. Shared by every call on the device; holds whatever the last call left.
static __device__ __align__(256) uint8_t g_adjust_static[kAdjustStaticBytes];

namespace lps {
namespace {

constexpr int kAdjustMaxDevices = 64;
std::mutex g_adjust_mutex;
uint8_t* g_adjust_ptr[kAdjustMaxDevices] = {};  // device address of g_adjust_static, per device ordinal

// In host code `g_adjust_static` is only a handle for the runtime, not an address a kernel can use: the address differs per
// device and has to be asked for with that device current. Looked up once per device, then cached.
int adjust_static_ptr(uint8_t** out) {
    int dev = -1;
    LPS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kAdjustMaxDevices) {
        set_error("lpslam_imadjust_dev_static: device ordinal beyond the " + std::to_string(kAdjustMaxDevices) + " this entry point tracks");
        return LPSLAM_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_adjust_mutex);
    if (!g_adjust_ptr[dev]) {
        void* p = nullptr;
        LPS_CUDA(cudaGetSymbolAddress(&p, g_adjust_static));
        g_adjust_ptr[dev] = (uint8_t*)p;
    }
    *out = g_adjust_ptr[dev];
    return LPSLAM_OK;
}

}  // namespace
}  // namespace lps

extern "C" int lpslam_imadjust_dev_static(const uint8_t* src_dev, int width, int height, size_t pitch, size_t frame_stride,
                                          uint32_t n_images, int auto_limits, double low_in, double high_in, double low_out,
                                          double high_out, uint8_t* dst_dev, size_t dst_pitch, size_t dst_frame_stride,
                                          double* limits_dev, void* stream) {
    LPS_ARG(src_dev && dst_dev, "null pointer");
    LPS_ARG(width > 0 && height > 0 && pitch >= (size_t)width && dst_pitch >= (size_t)width && n_images >= 1, "bad image geometry");
    if (n_images > LPSLAM_IMADJUST_STATIC_MAX_IMAGES) {
        set_error("lpslam_imadjust_dev_static: " + std::to_string(n_images) + " images in one call, the built-in scratch holds "
                  + std::to_string(LPSLAM_IMADJUST_STATIC_MAX_IMAGES) + " (LPSLAM_IMADJUST_STATIC_MAX_IMAGES); split the batch or use "
                  "lpslam_imadjust_dev with a scratch buffer of your own");
        return LPSLAM_ERR_ARG;
    }
    uint8_t* base = nullptr;
    if (int rc = adjust_static_ptr(&base)) return rc;  // first call on this device: one symbol lookup, nothing allocated
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t* hist = (uint32_t*)base;
    uint8_t* lut = base + 256 * sizeof(uint32_t) * (size_t)LPSLAM_IMADJUST_STATIC_MAX_IMAGES;
    const int vec = aligned(src_dev, 16) && pitch % 16 == 0 && frame_stride % 16 == 0;
    const int vec_out = vec && aligned(dst_dev, 16) && dst_pitch % 16 == 0 && dst_frame_stride % 16 == 0;
    if (auto_limits) {
        // the previous call's counts are still in the array: the histograms this call adds into start from zero
        LPS_CUDA(cudaMemsetAsync(hist, 0, 256 * sizeof(uint32_t) * (size_t)n_images, st));
        const long units = (long)(vec ? (width + 15) / 16 : width) * height;
        const int blocks = (int)std::min<long>(std::max<long>((units + 256 * 8 - 1) / (256 * 8), 1), 148 * 4);
        k_hist<<<dim3(blocks, n_images), 256, 0, st>>>(src_dev, width, height, pitch, frame_stride, hist, vec);
    }
    AdjustArgs a{auto_limits, low_in, high_in, low_out, high_out, (uint32_t)(width * height)};
    k_lut<<<n_images, 256, 0, st>>>(hist, a, lut, limits_dev);  // every lut entry of images 0..n-1 is rewritten, nothing to clear
    dim3 blk(64, 4), grd((width + 64 * 16 - 1) / (64 * 16), (height + 3) / 4, n_images);
    k_apply_lut<<<grd, blk, 0, st>>>(src_dev, width, height, pitch, frame_stride, lut, dst_dev, dst_pitch, dst_frame_stride, vec_out);
    LPS_CUDA(cudaGetLastError());
    return LPSLAM_OK;
}
