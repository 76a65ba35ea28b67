// match_window.cu — K8/K9: the windowed guided searches.
//   * Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea (src/Frame.cc:246-267,
//     :419-430, :358-415): 64x48 grid, keypoint -> cell by round(), query cells by floor/ceil,
//     traversal ix-major, iy, insertion order.  The grid is kept as a CSR whose cell index is
//     ix*48+iy, so one grid column of a query window is ONE contiguous item range.
//   * ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:469-600): the greedy bookkeeping
//     (vMatchedDistance / vnMatches21 written by earlier queries) makes queries order
//     dependent, so one warp owns one frame pair and walks the queries in order, while the
//     candidates of a query are evaluated across the 32 lanes (popc distance, packed
//     dist<<20|rank keys so that the first candidate in traversal order wins ties).
//   * ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) (:51-138): candidates of
//     all map points are enumerated in parallel; the only cross-map-point dependency is
//     "keypoint already taken by an earlier map point with observations" (:95-98), which is
//     resolved by a parallel fixed-point iteration over first-owner indices whose unique
//     fixed point is the sequential result.
#include <climits>

#include "match_internal.cuh"
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int GC = 64, GR = 48, NCELL = GC * GR;   // Frame.h:38-39
constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO = 30;   // ORBmatcher.cc:37-39
constexpr uint32_t KEY_INF = 0xFFFFFFFFu;

__device__ __forceinline__ int hamming256(const uint32_t a[8], const uint8_t* __restrict__ row) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(row));
    const uint4 y = __ldg(reinterpret_cast<const uint4*>(row) + 1);
    return __popc(a[0] ^ x.x) + __popc(a[1] ^ x.y) + __popc(a[2] ^ x.z) + __popc(a[3] ^ x.w) +
           __popc(a[4] ^ y.x) + __popc(a[5] ^ y.y) + __popc(a[6] ^ y.z) + __popc(a[7] ^ y.w);
}
__device__ __forceinline__ void load_row(uint32_t a[8], const uint8_t* __restrict__ row) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(row));
    const uint4 y = __ldg(reinterpret_cast<const uint4*>(row) + 1);
    a[0] = x.x; a[1] = x.y; a[2] = x.z; a[3] = x.w; a[4] = y.x; a[5] = y.y; a[6] = y.z; a[7] = y.w;
}

struct GridGeom { float minX, minY, invW, invH; };

// PosInGrid (Frame.cc:419-430): cell = ix*48+iy or -1
__device__ __forceinline__ int pos_in_grid(const GridGeom& g, float x, float y) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(x, g.minX), g.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(y, g.minY), g.invH));
    if (px < 0 || px >= GC || py < 0 || py >= GR) return -1;
    return px * GR + py;
}

// query window of GetFeaturesInArea (Frame.cc:363-379); false when the window misses the grid
__device__ __forceinline__ bool cell_window(const GridGeom& g, float x, float y, float r, int& cx0,
                                            int& cx1, int& cy0, int& cy1) {
    cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, g.minX), r), g.invW)));
    if (cx0 >= GC) return false;
    cx1 = min(GC - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, g.minX), r), g.invW)));
    if (cx1 < 0) return false;
    cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, g.minY), r), g.invH)));
    if (cy0 >= GR) return false;
    cy1 = min(GR - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, g.minY), r), g.invH)));
    if (cy1 < 0) return false;
    return true;
}

// One warp builds the CSR of a frame: cell_end[c] = end of cell c (start = end of c-1),
// items in keypoint-index order inside a cell (AssignFeaturesToGrid pushes in index order).
__device__ void warp_build_grid(const GridGeom& g, const orbfe_keypoint* __restrict__ kps, int n,
                                int* cell_end, unsigned short* items) {
    const int lane = threadIdx.x & 31;
    for (int c = lane; c < NCELL; c += 32) cell_end[c] = 0;
    __syncwarp();
    for (int i = lane; i < n; i += 32) {
        const int c = pos_in_grid(g, kps[i].x, kps[i].y);
        if (c >= 0) atomicAdd(&cell_end[c], 1);
    }
    __syncwarp();
    int carry = 0;   // exclusive scan: cell_end[c] becomes the start of cell c
    for (int base = 0; base < NCELL; base += 32) {
        const int v = cell_end[base + lane];
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += u;
        }
        cell_end[base + lane] = carry + inc - v;
        carry += __shfl_sync(FULL, inc, 31);
    }
    __syncwarp();
    for (int base = 0; base < n; base += 32) {   // stable fill; cursors end up as cell ends
        const int i = base + lane;
        const int c = i < n ? pos_in_grid(g, kps[i].x, kps[i].y) : -1;
        const unsigned peers = __match_any_sync(FULL, c);
        if (c >= 0) {
            const int rank = __popc(peers & ((1u << lane) - 1u));
            items[cell_end[c] + rank] = (unsigned short)i;
        }
        __syncwarp();
        if (c >= 0 && (peers & ((1u << lane) - 1u)) == 0) cell_end[c] += __popc(peers);
        __syncwarp();
    }
}

__device__ __forceinline__ uint32_t warp_min(uint32_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// ComputeThreeMaxima (ORBmatcher.cc:1777-1821) on bin counts
__device__ void three_maxima(const int* hist, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < HISTO; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}
