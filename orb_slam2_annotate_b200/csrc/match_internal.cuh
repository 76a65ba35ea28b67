// match_internal.cuh — launchers of the matcher kernels (liborbfe, sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbfe {

// brute-force kNN-2; device pointers, rows 32 bytes, 16-byte aligned. returns launches.
int launch_knn2(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                int* d_idx, int* d_d1, int* d_d2, cudaStream_t s);

}  // namespace orbfe
