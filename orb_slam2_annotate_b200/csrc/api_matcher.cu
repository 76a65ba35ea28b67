// api_matcher.cu — matcher half of the C ABI (include/orbfe.h): handle, scalar Hamming
// distance and the host/device entry points of the brute-force kNN-2 search.
#include <cstring>
#include <new>
#include <string>

#include "api_common.cuh"
#include "match_internal.cuh"

using namespace orbfe;

struct orbfe_matcher {
    int device = 0;
    cudaStream_t stream = nullptr;
    DevArray<uint8_t> scratch[8];   // grow-only device staging, reused across calls
};

extern "C" {

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1828-1844): 8 x 32-bit XOR + popcount
int orbfe_hamming256(const void* a, const void* b) {
    const unsigned char* pa = (const unsigned char*)a;
    const unsigned char* pb = (const unsigned char*)b;
    int dist = 0;
    for (int i = 0; i < 4; i++) {
        unsigned long long x, y;
        std::memcpy(&x, pa + 8 * i, 8);
        std::memcpy(&y, pb + 8 * i, 8);
        dist += __builtin_popcountll(x ^ y);
    }
    return dist;
}

int orbfe_matcher_create(int device, orbfe_matcher** out) {
    if (!out) return fail(ORBFE_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (device < 0 || orbfe_device_count() <= device)
        return fail(ORBFE_ERR_NO_DEVICE, "no CUDA device: liborbfe has no CPU fallback");
    orbfe_matcher* m = new (std::nothrow) orbfe_matcher();
    if (!m) return fail(ORBFE_ERR_ARG, "out of host memory");
    m->device = device;
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (ce != cudaSuccess) { delete m; return fail(ORBFE_ERR_CUDA, "cudaStreamCreate", ce); }
    *out = m;
    return ORBFE_OK;
}

void orbfe_matcher_destroy(orbfe_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
    for (auto& s : m->scratch) s.release();
    delete m;
}

void* orbfe_matcher_stream(orbfe_matcher* m) { return m ? (void*)m->stream : nullptr; }

int orbfe_matcher_sync(orbfe_matcher* m) {
    if (!m) return fail(ORBFE_ERR_ARG, "matcher is NULL");
    CU(cudaSetDevice(m->device));
    CU(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}

int orbfe_knn2_device(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                      int th, float ratio, int32_t* d_out_idx, int32_t* d_out_d1, int32_t* d_out_d2) {
    if (!m || nq < 0 || nt < 0 || (nq && (!d_query || !d_out_idx || !d_out_d1 || !d_out_d2)) || (nt && !d_train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2_device");
    if (((uintptr_t)d_query | (uintptr_t)d_train) & 15) return fail(ORBFE_ERR_ARG, "descriptor arrays must be 16-byte aligned");
    CU(cudaSetDevice(m->device));
    launch_knn2(d_query, nq, d_train, nt, th, ratio, d_out_idx, d_out_d1, d_out_d2, m->stream);
    CU(cudaGetLastError());
    return ORBFE_OK;
}

int orbfe_knn2(orbfe_matcher* m, const uint8_t* query, int nq, const uint8_t* train, int nt, int th,
               float ratio, int32_t* out_idx, int32_t* out_d1, int32_t* out_d2) {
    if (!m || nq < 0 || nt < 0 || (nq && (!query || !out_idx || !out_d1 || !out_d2)) || (nt && !train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2");
    if (nq == 0) return ORBFE_OK;
    CU(cudaSetDevice(m->device));
    int rc;
    if ((rc = m->scratch[0].ensure((size_t)nq * 32))) return rc;
    if ((rc = m->scratch[1].ensure((size_t)std::max(nt, 1) * 32))) return rc;
    if ((rc = m->scratch[2].ensure((size_t)nq * 12))) return rc;
    int* d_idx = (int*)m->scratch[2].p;
    int* d_d1 = d_idx + nq;
    int* d_d2 = d_d1 + nq;
    CU(cudaMemcpyAsync(m->scratch[0].p, query, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    if (nt) CU(cudaMemcpyAsync(m->scratch[1].p, train, (size_t)nt * 32, cudaMemcpyHostToDevice, m->stream));
    launch_knn2(m->scratch[0].p, nq, m->scratch[1].p, nt, th, ratio, d_idx, d_d1, d_d2, m->stream);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(out_idx, d_idx, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaMemcpyAsync(out_d1, d_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaMemcpyAsync(out_d2, d_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}

}  // extern "C"
