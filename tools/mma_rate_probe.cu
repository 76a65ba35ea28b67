// mma_rate_probe.cu — raw tcgen05.mma issue rate on this pool's B200 for the shapes a Hamming kNN could use, sustained
// (about half a second per variant, under the power cap), no epilogue: what is the ceiling of kind::i8 at N = 128 (what
// match_knn_mma.cu uses) and N = 256, and would the block-scaled FP4 path (kind::mxf4, +-1 as E2M1, unit UE8M0 scales)
// be faster in practice?  One CTA per SM; operands are +-1 patterns in the K-major SWIZZLE_128B layout; every "tile" is a
// full 256-element contraction of 128 x N rows (8 MMAs of K 32 for s8, 4 MMAs of K 64 for fp4).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/mma_rate_probe tools/mma_rate_probe.cu && tools/mma_rate_probe
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {      // K-major SWIZZLE_128B, 8-row groups 1024 B apart
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_i8(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_mxf4(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t sfa, uint32_t sfb) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc), "r"(sfa), "r"(sfb) : "memory");
}

struct alignas(1024) Smem {
    uint8_t a[2][128 * 128];          // two 128-byte K blocks of 128 rows
    uint8_t b[2][256 * 128];          // two K blocks of up to 256 rows
    unsigned long long bar[2];
    uint32_t tmem_base;
};

// KIND 0: s8 x s8 -> s32 (K 32 per MMA, two K blocks).  KIND 1: mxf4 block32 (K 64 per MMA, one K block).
template <int KIND, int N>
__global__ void __launch_bounds__(160) probe(int tiles, unsigned* sink) {
    extern __shared__ __align__(1024) unsigned char raw[];
    Smem& S = *reinterpret_cast<Smem*>(raw);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) { mbar_init(&S.bar[0], 1); mbar_init(&S.bar[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&S.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // +-1 operands: s8 0x01 / 0xFF, E2M1 0x2 / 0xA, from a hash of the position
    for (int i = tid; i < (int)(sizeof(S.a) + sizeof(S.b)) / 4; i += blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        uint32_t w;
        if (KIND == 0) w = 0x01010101u | ((h & 1) * 0xFEu) | (((h >> 1) & 1) * 0xFE00u) | (((h >> 2) & 1) * 0xFE0000u) | (((h >> 3) & 1) * 0xFE000000u);
        else w = 0x22222222u | ((h & 0x11111111u) << 3);
        reinterpret_cast<uint32_t*>(&S.a[0][0])[i] = w;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = S.tmem_base;
    constexpr int SFCOL = 448;                       // unit scale factors (UE8M0 0x7F) in the last 64 columns, every lane
    if (KIND == 1 && warp < 4) {
        const uint32_t one = 0x7F7F7F7Fu;
        for (int c = 0; c < 64; c++)
            asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tmem + ((uint32_t)(warp * 32) << 16) + SFCOL + c), "r"(one) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (warp == 4 && lane == 0) {
        constexpr int NBUF = (KIND == 1 ? 448 : 512) / N;            // accumulator buffers, used in turn
        const uint32_t idesc = KIND == 0
            ? (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24)
            : (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t aaddr = smem_u32(&S.a[0][0]), baddr = smem_u32(&S.b[0][0]);
        constexpr int GROUP = 8;                                     // tiles per commit
        int g = 0;
        for (int t = 0; t < tiles; t += GROUP, g++) {
            if (g >= 2) mbar_wait(&S.bar[g & 1], ((g >> 1) - 1) & 1);   // the group two back has drained
            for (int u = 0; u < GROUP; u++) {
                const uint32_t d = tmem + (uint32_t)(((t + u) % NBUF) * N);
                if (KIND == 0) {
#pragma unroll
                    for (int kb = 0; kb < 2; kb++) {
                        const uint64_t ad = umma_desc(aaddr + kb * 128 * 128), bd = umma_desc(baddr + kb * 256 * 128);
#pragma unroll
                        for (int k = 0; k < 4; k++) umma_i8(d, ad + 2 * k, bd + 2 * k, idesc, (kb | k) != 0);
                    }
                } else {
                    const uint64_t ad = umma_desc(aaddr), bd = umma_desc(baddr);
#pragma unroll
                    for (int k = 0; k < 4; k++) umma_mxf4(d, ad + 2 * k, bd + 2 * k, idesc, k != 0, tmem + SFCOL, tmem + SFCOL + 16);
                }
            }
            umma_commit(&S.bar[g & 1]);
        }
        // drain: the last two groups
        for (int k = (g >= 2 ? g - 2 : 0); k < g; k++) mbar_wait(&S.bar[k & 1], (k >> 1) & 1);
    }
    __syncwarp();
    if (warp == 4) {                                                 // one accumulator back, whole warp (.sync.aligned)
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t v;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(tmem) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (sink && blockIdx.x == 0 && lane == 0) sink[KIND * 2 + (N == 256)] = v;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

template <int KIND, int N>
int run(const char* name, int sms, unsigned* sink, double* out_tcmps) {
    CK(cudaFuncSetAttribute(probe<KIND, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    probe<KIND, N><<<sms, 160, sizeof(Smem)>>>(1024, sink);          // warm-up
    CK(cudaDeviceSynchronize());
    int tiles = 1 << 16;
    float ms = 0;
    for (int rep = 0; rep < 3; rep++) {                               // grow until a launch takes > 0.4 s
        CK(cudaEventRecord(e0));
        probe<KIND, N><<<sms, 160, sizeof(Smem)>>>(tiles, sink);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms > 400.f) break;
        tiles = (int)(tiles * (500.0 / (ms > 1.f ? ms : 1.f)));
    }
    const double cmps = (double)sms * tiles * 128.0 * N / (ms * 1e-3);
    *out_tcmps = cmps / 1e12;
    std::printf("  \"%s\": {\"tiles_per_cta\": %d, \"ms\": %.2f, \"tcmp_s\": %.3f, \"tera_ops_s\": %.1f},\n", name, tiles, ms, cmps / 1e12, cmps * 512.0 / 1e12);
    return 0;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    unsigned* sink;
    CK(cudaMalloc(&sink, 64));
    double r[4];
    std::printf("{\n  \"device\": \"%s\", \"sms\": %d,\n", p.name, p.multiProcessorCount);
    if (run<0, 128>("i8_m128_n128", p.multiProcessorCount, sink, &r[0])) return 1;
    if (run<0, 256>("i8_m128_n256", p.multiProcessorCount, sink, &r[1])) return 1;
    if (run<1, 128>("mxf4_m128_n128", p.multiProcessorCount, sink, &r[2])) return 1;
    if (run<1, 256>("mxf4_m128_n256", p.multiProcessorCount, sink, &r[3])) return 1;
    unsigned h[4];
    CK(cudaMemcpy(h, sink, 16, cudaMemcpyDeviceToHost));
    std::printf("  \"accumulator_samples\": [%d, %d, %.1f, %.1f],\n", (int)h[0], (int)h[1], *(float*)&h[2], *(float*)&h[3]);
    std::printf("  \"mxf4_over_i8_n128\": %.3f, \"mxf4_over_i8_n256\": %.3f, \"i8_n256_over_n128\": %.3f\n}\n", r[2] / r[0], r[3] / r[1], r[1] / r[0]);
    return 0;
}
