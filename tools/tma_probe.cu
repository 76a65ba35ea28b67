// Probe: 2D TMA box load of u8 windows at arbitrary (unaligned) x into shared memory.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
struct alignas(64) Maps { CUtensorMap m[2]; };
__global__ void probe(const __grid_constant__ Maps maps, int x, int y, uint8_t* out, int boxW, int boxH) {
    extern __shared__ __align__(128) uint8_t sm[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t barA = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(sm);
    if (threadIdx.x == 0) {
        printf("dst %u bar %u\n", dst, barA);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barA) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (threadIdx.x == 0) {
        const uint64_t mp = reinterpret_cast<uint64_t>(&maps.m[0]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barA), "r"(boxW * boxH) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(dst), "l"(mp), "r"(x), "r"(y), "r"(barA) : "memory");
    }
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(barA), "r"(0) : "memory");
    for (int i = threadIdx.x; i < boxW * boxH; i += 32) out[i] = sm[i];
}
int main() {
    const int pitch = 640, rows = 480 * 4, boxW = 32, boxH = 31;
    std::vector<uint8_t> h((size_t)pitch * rows);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)((i % pitch) * 7 + (i / pitch) * 13);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, 4096);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    Maps maps;
    cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)pitch};
    cuuint32_t box[2] = {(cuuint32_t)boxW, (cuuint32_t)boxH}, es[2] = {1, 1};
    CUresult r = ((enc_t)fn)(&maps.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc %d\n", (int)r);
    maps.m[1] = maps.m[0];
    for (int x : {16, 101, 203, 7}) {
        const int y = 777;
        probe<<<1, 32, 4096>>>(maps, x, y, o, boxW, boxH);
        cudaError_t ce = cudaDeviceSynchronize();
        printf("x=%d: %s\n", x, cudaGetErrorString(ce));
        if (ce != cudaSuccess) return 1;
        std::vector<uint8_t> res(boxW * boxH);
        cudaMemcpy(res.data(), o, res.size(), cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int rr = 0; rr < boxH; rr++) for (int c = 0; c < boxW; c++)
            if (res[rr * boxW + c] != h[(size_t)(y + rr) * pitch + x + c]) bad++;
        printf("  mismatches %d\n", bad);
    }
    return 0;
}
