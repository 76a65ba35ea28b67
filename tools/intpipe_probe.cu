// intpipe_probe.cu — issue rate of the integer / byte-SIMD instructions the extractor kernels are built from, in
// thread-operations per clock per SM (128 = one per lane per scheduler per clock).  Eight independent chains per
// thread, 8 CTAs x 256 threads per SM.  Prints one JSON line.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/intpipe_probe tools/intpipe_probe.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

enum { IMAD, IADD3, LOP3, DP4A, DP2A, VABSDIFF4, VIMNMX3, VMAXS2, PRMT, FUNNEL, POPC, NOPS };
static const char* kNames[NOPS] = {"imad", "iadd3", "lop3", "dp4a", "dp2a", "vabsdiff4", "vimnmx3_s16x2", "vmax_s16x2", "prmt", "shf", "popc"};

template <int OP>
__device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c) {
    switch (OP) {
        case IMAD: return a * b + c;
        case IADD3: return a + b + c;
        case LOP3: return (a & b) ^ c;
        case DP4A: return __dp4a(a, b, c);
        case DP2A: return __dp2a_lo(a, b, c);
        case VABSDIFF4: return __vabsdiffu4(a, b) + 0u * c;
        case VIMNMX3: return __vimin3_s16x2(a, b, c);
        case VMAXS2: return __vmaxs2(a, b) + 0u * c;
        case PRMT: return __byte_perm(a, b, c);
        case FUNNEL: return __funnelshift_r(a, b, c);
        case POPC: return __popc(a ^ b) + c;
    }
    return 0;
}

template <int OP>
__global__ void __launch_bounds__(256) probe(uint32_t* out, int iters, uint32_t seed) {
    uint32_t x[8], b = seed * (threadIdx.x + 3), c = seed ^ (threadIdx.x * 0x9E3779B9u);
    for (int i = 0; i < 8; i++) x[i] = seed * (threadIdx.x + 1) + i * 0x85EBCA6Bu;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int i = 0; i < 8; i++) x[i] = op<OP>(x[i], b, OP == PRMT ? (c & 0x7777u) : c);
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
static double rate(uint32_t* d, int sms, double mhz) {
    const int blocks = sms * 8, iters = 2048;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<OP><<<blocks, 256>>>(d, iters, 12345u);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); probe<OP><<<blocks, 256>>>(d, iters, 12345u); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double ops = (double)blocks * 256 * iters * 32;
    return ops / (best * 1e-3) / (mhz * 1e6) / sms;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double mhz = khz / 1000.0;
    uint32_t* d; cudaMalloc(&d, (size_t)p.multiProcessorCount * 8 * 256 * 4);
    double r[NOPS];
    r[IMAD] = rate<IMAD>(d, p.multiProcessorCount, mhz); r[IADD3] = rate<IADD3>(d, p.multiProcessorCount, mhz);
    r[LOP3] = rate<LOP3>(d, p.multiProcessorCount, mhz); r[DP4A] = rate<DP4A>(d, p.multiProcessorCount, mhz);
    r[DP2A] = rate<DP2A>(d, p.multiProcessorCount, mhz); r[VABSDIFF4] = rate<VABSDIFF4>(d, p.multiProcessorCount, mhz);
    r[VIMNMX3] = rate<VIMNMX3>(d, p.multiProcessorCount, mhz); r[VMAXS2] = rate<VMAXS2>(d, p.multiProcessorCount, mhz);
    r[PRMT] = rate<PRMT>(d, p.multiProcessorCount, mhz); r[FUNNEL] = rate<FUNNEL>(d, p.multiProcessorCount, mhz);
    r[POPC] = rate<POPC>(d, p.multiProcessorCount, mhz);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz_nominal\": %.0f, \"unit\": \"thread-ops per clock per SM at the nominal clock\"", p.name,
           p.multiProcessorCount, mhz);
    for (int i = 0; i < NOPS; i++) printf(", \"%s\": %.1f", kNames[i], r[i]);
    printf("}\n");
    return 0;
}
