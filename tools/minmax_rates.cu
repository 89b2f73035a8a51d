// Microbenchmark: sustained rate of the 3-input 16x2 integer min/max (VIMNMX3.U16x2, alu pipe) and of the 3-input half2
// min/max (VHMNMX) alone and interleaved 1:1 — small non-negative integers in 16-bit lanes are fp16 subnormals, whose
// order is the order of their bit patterns, so either instruction can evaluate FAST's arc network.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o minmax_rates minmax_rates.cu
#include <cuda_runtime.h>
#include <cstdio>
constexpr int CH = 8, UNROLL = 16, ITERS = 512;
__device__ __forceinline__ unsigned imin3(unsigned a, unsigned b, unsigned c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ unsigned imax3(unsigned a, unsigned b, unsigned c) { return __vimax3_u16x2(a, b, c); }
__device__ __forceinline__ unsigned hmin3(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm volatile("{.reg .b32 t; min.f16x2 t, %1, %2; min.f16x2 %0, t, %3;}" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ unsigned hmax3(unsigned a, unsigned b, unsigned c) {
    unsigned d;
    asm volatile("{.reg .b32 t; max.f16x2 t, %1, %2; max.f16x2 %0, t, %3;}" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ unsigned hmin2(unsigned a, unsigned b) {
    unsigned d;
    asm volatile("min.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ unsigned hmax2(unsigned a, unsigned b) {
    unsigned d;
    asm volatile("max.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
template <int MODE>
__global__ void __launch_bounds__(1024, 2) k(unsigned* out, unsigned b, unsigned c) {
    unsigned a[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) a[i] = (threadIdx.x * 7 + i) & 0x00FF00FF;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < UNROLL; u += 2)
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                // min then max against neighbouring chains: nothing folds
                if (MODE == 0) { a[i] = imin3(a[i], a[(i + 1) % CH], b); a[i] = imax3(a[i], a[(i + 3) % CH], c); }
                if (MODE == 1) { a[i] = hmin3(a[i], a[(i + 1) % CH], b); a[i] = hmax3(a[i], a[(i + 3) % CH], c); }
                if (MODE == 3) { a[i] = hmin2(a[i], a[(i + 1) % CH]); a[i] = hmax2(a[i], a[(i + 3) % CH]); }
                if (MODE == 4) { a[i] = imin3(a[i], a[(i + 1) % CH], b); a[i] = hmax2(a[i], a[(i + 3) % CH]); }
                if (MODE == 2) { a[i] = imin3(a[i], a[(i + 1) % CH], b); a[i] = hmax3(a[i], a[(i + 3) % CH], c); }
            }
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(int sms, const char* name) {
    const int blocks = sms * 2;
    unsigned* out;
    cudaMalloc(&out, blocks * 1024 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, 1024>>>(out, 0x00300010u, 0x00500040u);
    cudaEventRecord(e0);
    k<MODE><<<blocks, 1024>>>(out, 0x00300010u, 0x00500040u);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const double total = 32.0 * ITERS * UNROLL * CH * blocks;
    printf("%-28s %8.1f G warp-instr/s = %.2f warp-instr/clk/SM at %d MHz (%.3f ms)\n", name, total / (ms * 1e-3) / 1e9,
           total / (ms * 1e-3) / (sms * p.clockRate * 1e3), p.clockRate / 1000, ms);
    cudaFree(out);
}
int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    run<0>(p.multiProcessorCount, "VIMNMX3.U16x2");
    run<1>(p.multiProcessorCount, "VHMNMX (half2, 3-input)");
    run<2>(p.multiProcessorCount, "VIMNMX3 + VHMNMX 1:1");
    run<3>(p.multiProcessorCount, "HMNMX2 (half2, 2-input)");
    run<4>(p.multiProcessorCount, "VIMNMX3 + HMNMX2 1:1");
    return 0;
}
