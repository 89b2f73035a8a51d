// Microbenchmark: sustained issue rate (warp instructions / clk / SM) of the integer instructions the kernels lean on,
// measured on the device with clock64 inside fully occupied SMs. Gives the popc / dp4a / prmt denominators DESIGN.md
// quotes. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_rates pipe_rates.cu ; run: ./pipe_rates
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

constexpr int CH = 8;      // independent chains per thread
constexpr int UNROLL = 16; // ops per chain per loop trip
constexpr int ITERS = 512;

enum Op { IMAD, IADD3, LOP3, PRMT, SHF, DP4A, DP2A, POPC, VMNMX, XORPOPCADD, DP4A_PRMT, DP2A_PRMT, IMAD_PRMT, IMAD_IADD, POPC_LOP, IMADHI, IMADWIDE, DP2A_IMAD, DP2A_SHF, SHF_PRMT, VABSDIFF4, IMADHI_PRMT, NOPS };
const char* names[NOPS] = {"IMAD", "IADD3", "LOP3", "PRMT", "SHF", "IDP.4A", "IDP.2A", "POPC", "VIMNMX.U16x2", "LOP3+POPC+IADD (hamming word)",
                           "IDP.4A+PRMT 1:1", "IDP.2A+PRMT 1:1", "IMAD+PRMT 1:1", "IMAD+IADD3 1:1", "POPC+LOP3 1:1",
                           "IMAD.HI", "IMAD.WIDE", "IDP.2A+IMAD 1:1", "IDP.2A+SHF 1:1", "SHF+PRMT 1:1", "VABSDIFF4", "IMAD.HI+PRMT 1:1"};
const int per_step[NOPS] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 3, 2, 2, 2, 2, 2, 1, 1, 2, 2, 2, 1, 2};

template <int OP>
__device__ __forceinline__ void step(unsigned& a, unsigned b, unsigned c) {
    if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == IADD3) asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(b));
    if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == SHF) asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == DP4A) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == DP2A) asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == POPC) asm volatile("popc.b32 %0, %0;" : "+r"(a));
    if (OP == VMNMX) asm volatile("max.u16x2 %0, %0, %1;" : "+r"(a) : "r"(b));
    if (OP == XORPOPCADD) {
        unsigned t;
        asm volatile("xor.b32 %0, %1, %2;" : "=r"(t) : "r"(a), "r"(b));
        asm volatile("popc.b32 %0, %0;" : "+r"(t));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(t));
    }
    if (OP == DP4A_PRMT) {
        asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == DP2A_PRMT) {
        asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == IMAD_PRMT) {
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == IMAD_IADD) {
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(b));
    }
    if (OP == IMADHI) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == IMADWIDE) {
        unsigned long long w;
        asm volatile("mad.wide.u32 %0, %1, %2, %3;" : "=l"(w) : "r"(a), "r"(b), "l"((unsigned long long)c << 20));
        a = (unsigned)(w >> 8);  // compiles to a byte permute at most; counted as one step (upper bound on the cost)
    }
    if (OP == DP2A_IMAD) {
        asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == DP2A_SHF) {
        asm volatile("dp2a.lo.u32.u32 %0, %1, %2, %0;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == SHF_PRMT) {
        asm volatile("shf.r.wrap.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == VABSDIFF4) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    if (OP == IMADHI_PRMT) {
        asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
        asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(a) : "r"(b), "r"(c));
    }
    if (OP == POPC_LOP) {
        asm volatile("popc.b32 %0, %0;" : "+r"(a));
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a) : "r"(b), "r"(c));
    }
}

template <int OP>
__global__ void __launch_bounds__(1024, 2) k(unsigned* out, long long* cyc, unsigned b, unsigned c) {
    unsigned a[CH];
#pragma unroll
    for (int i = 0; i < CH; ++i) a[i] = threadIdx.x * 7 + i;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
#pragma unroll
            for (int i = 0; i < CH; ++i) step<OP>(a[i], b, c);
    }
    __syncthreads();
    const long long t1 = clock64();
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < CH; ++i) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(int sms) {
    const int blocks = sms * 2;
    unsigned* out;
    long long* cyc;
    cudaMalloc(&out, blocks * 1024 * 4);
    cudaMalloc(&cyc, blocks * 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<OP><<<blocks, 1024>>>(out, cyc, 0x3210u, 0x5410u);
    cudaEventRecord(e0);
    k<OP><<<blocks, 1024>>>(out, cyc, 0x3210u, 0x5410u);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, blocks * 8, cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < blocks; ++i) avg += h[i];
    avg /= blocks;
    const double winstr_per_block = 32.0 * ITERS * UNROLL * CH * per_step[OP];
    // two resident blocks per SM run concurrently for ~avg cycles
    const double per_clk_sm = 2.0 * winstr_per_block / avg;
    const double total = winstr_per_block * blocks;
    printf("%-34s %6.2f warp-instr/clk/SM  (%6.1f lanes/clk/SM)   whole-GPU %8.1f G warp-instr/s  (%.3f ms)\n", names[OP], per_clk_sm,
           per_clk_sm * 32, total / (ms * 1e-3) / 1e9, ms);
    cudaFree(out);
    cudaFree(cyc);
    delete[] h;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, %d MHz\n", p.name, p.multiProcessorCount, p.clockRate / 1000);
    const int s = p.multiProcessorCount;
    run<IMAD>(s); run<IADD3>(s); run<LOP3>(s); run<PRMT>(s); run<SHF>(s); run<DP4A>(s); run<DP2A>(s); run<POPC>(s); run<VMNMX>(s);
    run<XORPOPCADD>(s); run<DP4A_PRMT>(s); run<DP2A_PRMT>(s); run<IMAD_PRMT>(s); run<IMAD_IADD>(s); run<POPC_LOP>(s);
    run<IMADHI>(s); run<IMADWIDE>(s); run<DP2A_IMAD>(s); run<DP2A_SHF>(s); run<SHF_PRMT>(s); run<VABSDIFF4>(s); run<IMADHI_PRMT>(s);
    return 0;
}
