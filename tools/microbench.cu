// microbench.cu — measures the two candidate engines for the Hamming matcher on the box:
//   (a) the integer popc pipe (XOR + POPC over 32-bit words), in G 256-bit comparisons/s
//   (b) mma.sync.m16n8k256 b1 and.popc / xor.popc on the legacy tensor path
// Prints one JSON line; bench.py records the popc figure as the roofline denominator.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

__global__ void popc_kernel(uint32_t* out, int iters, uint32_t seed) {
    uint32_t a[8], acc[4] = {0, 0, 0, 0};
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) acc[r] += __popc(a[i] ^ (acc[(r + 1) & 3] + i));
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc[0] + acc[1] + acc[2] + acc[3];
}

template <int OP>
__global__ void bmma_kernel(int* out, int iters, uint32_t seed) {
    uint32_t a[4], b[2];
    for (int i = 0; i < 4; i++) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
    for (int i = 0; i < 2; i++) b[i] = seed * (threadIdx.x + 7) + i * 0x85EBCA6Bu;
    int c[4][4] = {};
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (OP == 0)
                asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.and.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[r][0]), "+r"(c[r][1]), "+r"(c[r][2]), "+r"(c[r][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
            else
                asm volatile("mma.sync.aligned.m16n8k256.row.col.s32.b1.b1.s32.xor.popc {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+r"(c[r][0]), "+r"(c[r][1]), "+r"(c[r][2]), "+r"(c[r][3])
                             : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
        }
    }
    int s = 0;
    for (int r = 0; r < 4; r++) for (int i = 0; i < 4; i++) s += c[r][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256;
    uint32_t* d; cudaMalloc(&d, (size_t)blocks * threads * 4);
    const int iters = 4096;
    float ms = time_ms([&] { popc_kernel<<<blocks, threads>>>(d, iters, 12345u); });
    const double popc_ops = (double)blocks * threads * iters * 32;
    const double gcmp_popc = popc_ops / 8 / (ms * 1e-3) / 1e9;
    double gcmp_and = 0, gcmp_xor = 0;
    const int biters = 2048;
    ms = time_ms([&] { bmma_kernel<0><<<blocks, threads>>>((int*)d, biters, 7u); });
    if (cudaGetLastError() == cudaSuccess) gcmp_and = (double)blocks * (threads / 32) * biters * 4 * 128 / (ms * 1e-3) / 1e9;
    ms = time_ms([&] { bmma_kernel<1><<<blocks, threads>>>((int*)d, biters, 7u); });
    if (cudaGetLastError() == cudaSuccess) gcmp_xor = (double)blocks * (threads / 32) * biters * 4 * 128 / (ms * 1e-3) / 1e9;
    printf("{\"device\": \"%s\", \"sms\": %d, \"popc_gcmp_s\": %.1f, \"popc_gpopc_s\": %.1f, \"bmma_and_gcmp_s\": %.1f, \"bmma_xor_gcmp_s\": %.1f}\n",
           p.name, sms, gcmp_popc, gcmp_popc * 8, gcmp_and, gcmp_xor);
    return 0;
}
