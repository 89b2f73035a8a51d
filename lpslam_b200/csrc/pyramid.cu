// K1 — image pyramid. Replaces cv::resize(..., INTER_LINEAR) at feature/orb_extractor.cc:266-275.
// Arithmetic: 11-bit fixed-point coefficients, horizontal pass in int32, vertical pass
//   dst = (((b0 * (H0 >> 4)) >> 16) + ((b1 * (H1 >> 4)) >> 16) + 2) >> 2
// Coefficient / offset tables are built on the host (api.cu) with the same float/double steps as OpenCV.
//
// Layout: one launch per level over the whole batch, grid.z = frame. Each thread walks DOWN a strip of 4 output
// columns: the horizontal pass of a source row is kept in registers and reused when consecutive output rows share it
// (vertical scale 1.2 -> 1.2 source rows per output row instead of 2), which cuts the byte-gather count by 40 %.
#include "common.cuh"

namespace lps {

namespace {

constexpr int kRowsPerThread = 8;

__device__ __forceinline__ void hpass4(const uint8_t* __restrict__ row, const int (&sx)[4], const int (&sx1)[4],
                                       const short2 (&c)[4], int (&out)[4]) {
#pragma unroll
    for (int k = 0; k < 4; ++k) out[k] = __ldg(row + sx[k]) * c[k].x + __ldg(row + sx1[k]) * c[k].y;
}

__global__ void __launch_bounds__(256) k_resize_linear(LevelGeom src, LevelGeom dst) {
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y0 = (blockIdx.y * blockDim.y + threadIdx.y) * kRowsPerThread;
    if (x0 >= dst.w || y0 >= dst.h) return;
    const uint8_t* __restrict__ s = src.img + blockIdx.z * src.frame_stride;
    uint8_t* __restrict__ d = dst.img + blockIdx.z * dst.frame_stride;

    int sx[4], sx1[4];
    short2 c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int x = min(x0 + k, dst.w - 1);
        sx[k] = dst.xofs[x];
        sx1[k] = min(sx[k] + 1, src.w - 1);
        c[k] = dst.xcoef[x];
    }
    int h0[4], h1[4];
    int have0 = -1, have1 = -1;  // source rows currently held in h0 / h1
    const int yend = min(y0 + kRowsPerThread, dst.h);
    for (int y = y0; y < yend; ++y) {
        const int sy = dst.yofs[y];
        const short2 b = dst.ycoef[y];
        const int r0 = min(max(sy, 0), src.h - 1);
        const int r1 = min(max(sy + 1, 0), src.h - 1);
        if (r0 == have1) {
#pragma unroll
            for (int k = 0; k < 4; ++k) h0[k] = h1[k];
        }
        else if (r0 != have0) {
            hpass4(s + (size_t)r0 * src.pitch, sx, sx1, c, h0);
        }
        have0 = r0;
        if (r1 == r0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) h1[k] = h0[k];
        }
        else if (r1 != have1) {
            hpass4(s + (size_t)r1 * src.pitch, sx, sx1, c, h1);
        }
        have1 = r1;
        uint32_t packed = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int v = (((b.x * (h0[k] >> 4)) >> 16) + ((b.y * (h1[k] >> 4)) >> 16) + 2) >> 2;
            v = min(max(v, 0), 255);
            packed |= (uint32_t)v << (8 * k);
        }
        uint8_t* o = d + (size_t)y * dst.pitch + x0;
        if (x0 + 4 <= dst.w) {
            *reinterpret_cast<uint32_t*>(o) = packed;  // pitch and x0 are multiples of 4
        }
        else {
            for (int k = 0; x0 + k < dst.w; ++k) o[k] = (uint8_t)(packed >> (8 * k));
        }
    }
}

// exact 2x decimation: OpenCV's INTER_LINEAR takes the INTER_AREA fast path, (a + b + c + d + 2) >> 2
__global__ void __launch_bounds__(256) k_resize_half(LevelGeom src, LevelGeom dst) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= dst.w || y >= dst.h) return;
    const uint8_t* __restrict__ s = src.img + blockIdx.z * src.frame_stride + (size_t)(2 * y) * src.pitch + 2 * x;
    const int v = (s[0] + s[1] + s[src.pitch] + s[src.pitch + 1] + 2) >> 2;
    dst.img[blockIdx.z * dst.frame_stride + (size_t)y * dst.pitch + x] = (uint8_t)v;
}

// pitched 2-D copy (level 0 ingest from a caller-pitched device image, and same-size "resize")
__global__ void __launch_bounds__(256) k_copy2d(const uint8_t* __restrict__ src, size_t spitch, size_t sstride,
                                               uint8_t* __restrict__ dst, int w, int h, int dpitch, size_t dstride,
                                               int vec_ok) {
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (y >= h) return;
    const uint8_t* s = src + blockIdx.z * sstride + (size_t)y * spitch;
    uint8_t* d = dst + blockIdx.z * dstride + (size_t)y * dpitch;
    const int x16 = (blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (x16 >= w) return;
    if (vec_ok && x16 + 16 <= w) {
        *reinterpret_cast<uint4*>(d + x16) = __ldg(reinterpret_cast<const uint4*>(s + x16));
    }
    else {
        for (int k = 0; k < 16 && x16 + k < w; ++k) d[x16 + k] = s[x16 + k];
    }
}

}  // namespace

void launch_resize(const LevelGeom& src, const LevelGeom& dst, int batch, cudaStream_t st) {
    if (dst.w <= 0 || dst.h <= 0) return;
    if (dst.mode == 2) {
        launch_copy_level0(src.img, src.pitch, src.frame_stride, dst, batch, st);
        return;
    }
    if (dst.mode == 1) {
        dim3 blk(32, 8);
        dim3 grd((dst.w + 31) / 32, (dst.h + 7) / 8, batch);
        k_resize_half<<<grd, blk, 0, st>>>(src, dst);
        return;
    }
    dim3 blk(32, 8);
    dim3 grd((dst.w + 127) / 128, (dst.h + 8 * kRowsPerThread - 1) / (8 * kRowsPerThread), batch);
    k_resize_linear<<<grd, blk, 0, st>>>(src, dst);
}

void launch_copy_level0(const uint8_t* src, size_t src_pitch, size_t src_frame_stride, const LevelGeom& dst, int batch,
                        cudaStream_t st) {
    dim3 blk(32, 8);
    dim3 grd((dst.w + 16 * 32 - 1) / (16 * 32), (dst.h + 7) / 8, batch);
    const int vec_ok = (reinterpret_cast<uintptr_t>(src) % 16 == 0) && (src_pitch % 16 == 0) && (src_frame_stride % 16 == 0);
    k_copy2d<<<grd, blk, 0, st>>>(src, src_pitch, src_frame_stride, dst.img, dst.w, dst.h, dst.pitch, dst.frame_stride, vec_ok);
}

}  // namespace lps
