// api_common.cuh — host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>

#include <stdint.h>

#include <algorithm>
#include <string>

#include "../../include/orbfe.h"

namespace orbfe {

int fail(int code, const char* what, cudaError_t ce = cudaSuccess);

#define CU(call)                                                         \
    do {                                                                 \
        cudaError_t _e = (call);                                         \
        if (_e != cudaSuccess) return fail(ORBFE_ERR_CUDA, #call, _e);   \
    } while (0)

template <class T>
struct DevArray {
    T* p = nullptr;
    size_t cap = 0;   // elements
    int ensure(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&p, n * sizeof(T));
        if (e != cudaSuccess) return fail(ORBFE_ERR_CUDA, "cudaMalloc", e);
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// grow-only page-locked host buffer
struct PinnedArray {
    uint8_t* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        n += n / 2;
        cudaError_t e = cudaMallocHost((void**)&p, n);
        if (e != cudaSuccess) return fail(ORBFE_ERR_CUDA, "cudaMallocHost", e);
        cap = n;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

}  // namespace orbfe
