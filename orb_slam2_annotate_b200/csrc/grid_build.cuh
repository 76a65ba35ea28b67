// grid_build.cuh — Frame::AssignFeaturesToGrid (reference src/Frame.cc:246-267) as a CTA-wide
// device routine shared by the Frame epilogue (frame.cu) and the matcher's own grid build for
// frames that arrive from the host (match_window.cu).
//
// mGrid[ix][iy] becomes a CSR over cells c = ix*48 + iy: cell_end[c] is one past the last item
// of cell c (cell 0 starts at 0), items hold keypoint indices in push_back order (ascending).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace orbfe {

constexpr int kGridThreads = 256;                 // CTA size block_grid_csr expects
constexpr int kGridCells = 64 * 48;

// cell(i) -> ix*48 + iy of keypoint i, or -1 outside the grid (PosInGrid, Frame.cc:419-430).
// cnt: kGridCells ints and wsum: kGridThreads/32 ints of shared memory.  Ends with the CSR in
// ce / it visible to the whole CTA.
template <class CellFn>
__device__ __forceinline__ void block_grid_csr(int* cnt, int* wsum, int n, CellFn cell, int* __restrict__ ce,
                                               unsigned short* __restrict__ it) {
    constexpr int CPT = kGridCells / kGridThreads;   // cells per thread in the scan (12)
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < kGridCells; i += kGridThreads) cnt[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {
        const int c = cell(i);
        if (c >= 0) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    // exclusive scan over the cells: thread t owns cells [t*CPT, (t+1)*CPT)
    int loc[CPT], s = 0;
#pragma unroll
    for (int j = 0; j < CPT; j++) { loc[j] = cnt[tid * CPT + j]; s += loc[j]; }
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    int base = inc - s;
    for (int w = 0; w < wid; w++) base += wsum[w];
#pragma unroll
    for (int j = 0; j < CPT; j++) { cnt[tid * CPT + j] = base; base += loc[j]; }   // cell starts = fill cursors
    __syncthreads();
    for (int i = tid; i < n; i += kGridThreads) {
        const int c = cell(i);
        if (c >= 0) it[atomicAdd(&cnt[c], 1)] = (unsigned short)i;
    }
    __syncthreads();
    // cursors are now the cell ends; restore push_back (index) order inside each cell
    int start = tid == 0 ? 0 : cnt[tid * CPT - 1];
#pragma unroll 1
    for (int j = 0; j < CPT; j++) {
        const int end = cnt[tid * CPT + j];
        for (int a = start + 1; a < end; a++) {
            const unsigned short v = it[a];
            int b = a - 1;
            while (b >= start && it[b] > v) { it[b + 1] = it[b]; b--; }
            it[b + 1] = v;
        }
        ce[tid * CPT + j] = end;
        start = end;
    }
    __syncthreads();
}

}  // namespace orbfe
