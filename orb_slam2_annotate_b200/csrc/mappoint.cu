// mappoint.cu — K13: MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:266-334,
// SURVEY.md §8(f) rank 4) for a batch of map points: N x N Hamming distances among the
// descriptors that observe a point, the median of every row (element (N-1)/2 of the sorted row,
// the zero on the diagonal included, :313-315), and the first row with the least median (:317).
//
// One warp per map point, one lane per row (rows beyond 32 in further rounds).  A lane keeps its
// own descriptor in registers and finds its row's median without storing or sorting the row: the
// median is the smallest value v with #{j : d(i,j) <= v} > (N-1)/2, located by a nine-step
// bisection over 0..256 that recounts the row each step (the other descriptors are broadcast
// loads from L1).  The warp then takes the minimum of median << 16 | row, which is the
// reference's strict '<' scan.
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int DD_WARPS = 4;

__global__ void __launch_bounds__(DD_WARPS * 32)
distinctive_kernel(const uint8_t* __restrict__ desc, const int* __restrict__ start, int npoints, int* __restrict__ best) {
    const int lane = threadIdx.x & 31;
    const int p = blockIdx.x * DD_WARPS + (threadIdx.x >> 5);
    if (p >= npoints) return;
    const int s = start[p], N = start[p + 1] - s;
    if (N <= 0) {
        if (lane == 0) best[p] = -1;
        return;
    }
    const uint4* d = reinterpret_cast<const uint4*>(desc) + 2 * (long long)s;
    const int need = (N - 1) / 2 + 1;          // rank of the median, counted from 1
    unsigned bestKey = 0xFFFFFFFFu;
    for (int i0 = 0; i0 < N; i0 += 32) {
        const int i = i0 + lane;
        unsigned key = 0xFFFFFFFFu;
        if (i < N) {
            const uint4 a = __ldg(d + 2 * i), b = __ldg(d + 2 * i + 1);
            int lo = 0, hi = 256;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                int cnt = 0;
                for (int j = 0; j < N; j++) {
                    const uint4 u = __ldg(d + 2 * j), v = __ldg(d + 2 * j + 1);
                    const int dist = __popc(a.x ^ u.x) + __popc(a.y ^ u.y) + __popc(a.z ^ u.z) + __popc(a.w ^ u.w) +
                                     __popc(b.x ^ v.x) + __popc(b.y ^ v.y) + __popc(b.z ^ v.z) + __popc(b.w ^ v.w);
                    cnt += dist <= mid;
                }
                if (cnt >= need) hi = mid; else lo = mid + 1;
            }
            key = ((unsigned)lo << 16) | (unsigned)i;      // rows beyond 65535 are refused by the host
        }
        bestKey = min(bestKey, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) bestKey = min(bestKey, __shfl_xor_sync(FULL, bestKey, o));
    if (lane == 0) best[p] = (int)(bestKey & 0xFFFFu);
}

}  // namespace

int launch_distinctive(const uint8_t* d_desc, const int* d_start, int npoints, int* d_best, cudaStream_t st) {
    if (npoints <= 0) return 0;
    distinctive_kernel<<<(npoints + DD_WARPS - 1) / DD_WARPS, DD_WARPS * 32, 0, st>>>(d_desc, d_start, npoints, d_best);
    return 1;
}

}  // namespace orbfe
