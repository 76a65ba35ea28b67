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

#include "orbfe_internal.cuh"
#include "match_internal.cuh"
#include "grid_build.cuh"

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
    cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, g.qminX), r), g.invW)));
    if (cx0 >= GC) return false;
    cx1 = min(GC - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, g.qminX), r), g.invW)));
    if (cx1 < 0) return false;
    cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, g.qminY), r), g.invH)));
    if (cy0 >= GR) return false;
    cy1 = min(GR - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, g.qminY), r), g.invH)));
    if (cy1 < 0) return false;
    return true;
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

// ---------------------------------------------------------------------------------------
// SearchForInitialization — three kernels: grids (CTA per pair), candidate lists (warp per
// query, fully parallel), greedy replay (warp per pair).

constexpr int SI_WARPS = 8;      // warps per frame pair
constexpr int SI_QB = 64;        // queries staged per block by the greedy kernel
constexpr int SI_K = 32;         // candidates kept per query by the parallel phase (the whole sorted warp)
constexpr int SI_CL = 512;       // per-warp candidate list capacity (longer windows take the direct path)

// Candidates of one query in GetFeaturesInArea order, evaluated across the lanes of a warp.
// WITH_FILTER applies the reference's running skip rule vMatchedDistance[i2] <= dist (:515-516);
// without it the caller receives the SI_K smallest (dist, traversal rank) keys instead.
struct QueryCtx {
    const orbfe_keypoint* kp2; const uint8_t* d2; const int* cell_end; const unsigned short* items;
    const unsigned char* mdist; GridGeom grid; float r;
};

__device__ __forceinline__ bool cand_ok(const QueryCtx& c, int i2, float qx, float qy) {
    const orbfe_keypoint k2 = c.kp2[i2];
    if (k2.octave != 0) return false;                     // minLevel = maxLevel = 0 (:494, Frame.cc:381-395)
    return fabsf(__fsub_rn(k2.x, qx)) < c.r && fabsf(__fsub_rn(k2.y, qy)) < c.r;
}

// sequential-rule evaluation of one query by one warp: best / second-best key, best index
__device__ void eval_query_filtered(const QueryCtx& c, const uint32_t q[8], float qx, float qy, uint32_t& g1,
                                    uint32_t& g2, int& best) {
    const int lane = threadIdx.x & 31;
    uint32_t b1 = KEY_INF, b2 = KEY_INF;
    int bi = -1, seq = 0;
    int cx0, cx1, cy0, cy1;
    g1 = g2 = KEY_INF; best = -1;
    if (!cell_window(c.grid, qx, qy, c.r, cx0, cx1, cy0, cy1)) return;
    for (int ix = cx0; ix <= cx1; ix++) {
        const int c0 = ix * GR + cy0;
        const int s = c0 ? c.cell_end[c0 - 1] : 0, e = c.cell_end[ix * GR + cy1];
        for (int j = s + lane; j < e; j += 32) {
            const int i2 = c.items[j];
            if (!cand_ok(c, i2, qx, qy)) continue;
            const int dist = hamming256(q, c.d2 + 32 * (long long)i2);
            if (c.mdist[i2] != 255 && (int)c.mdist[i2] <= dist) continue;           // :515-516
            const uint32_t key = ((uint32_t)dist << 20) | (uint32_t)(seq + j - s);
            if (key < b1) { b2 = b1; b1 = key; bi = i2; }                           // :518-527
            else if (key < b2) b2 = key;
        }
        seq += e - s;
    }
    g1 = warp_min(b1);
    if (g1 == KEY_INF) return;
    const unsigned own = __ballot_sync(FULL, b1 == g1);
    best = __shfl_sync(FULL, bi, __ffs(own) - 1);
    g2 = warp_min(b1 == g1 ? b2 : b1);
}

// ascending sort of one 32-bit key per lane / final merge of a bitonic sequence
__device__ __forceinline__ uint32_t bitonic_merge32(uint32_t v, int lane) {
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint32_t o = __shfl_xor_sync(FULL, v, j);
        v = (lane & j) ? max(v, o) : min(v, o);
    }
    return v;
}
__device__ __forceinline__ uint32_t bitonic_sort32(uint32_t v, int lane) {
#pragma unroll
    for (int k = 2; k < 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const uint32_t o = __shfl_xor_sync(FULL, v, j);
            v = (((lane & k) != 0) != ((lane & j) != 0)) ? max(v, o) : min(v, o);
        }
    }
    return bitonic_merge32(v, lane);
}

// Kernel G — the CSR grid of every pair's second frame, one CTA per pair: shared-memory
// histogram, block scan, atomic fill, then every cell's (few) items are put back into keypoint
// order, which is the order AssignFeaturesToGrid pushes them in.
constexpr int SG_THREADS = 256;
__global__ void __launch_bounds__(SG_THREADS)
search_init_grid_kernel(const __grid_constant__ InitArgs a) {
    extern __shared__ __align__(16) unsigned char sm[];
    int* cur = (int*)sm;                                         // NCELL: count -> start -> end
    unsigned short* its = (unsigned short*)(cur + NCELL);        // cap
    __shared__ int wsum[SG_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, p = blockIdx.x;
    const int fb = a.pairs ? a.pairs[p].y : p + 1;
    const int n2 = a.counts[fb];
    const orbfe_keypoint* kp2 = a.kps + (long long)fb * a.cap;
    for (int c = tid; c < NCELL; c += SG_THREADS) cur[c] = 0;
    __syncthreads();
    // only octave-0 keypoints can pass GetFeaturesInArea(.., 0, 0) (:494, Frame.cc:381-395), and
    // dropping the others keeps the traversal order of those that can
    for (int i = tid; i < n2; i += SG_THREADS) {
        const int c = kp2[i].octave == 0 ? pos_in_grid(a.grid, kp2[i].x, kp2[i].y) : -1;
        if (c >= 0) atomicAdd(&cur[c], 1);
    }
    __syncthreads();
    constexpr int CPT = NCELL / SG_THREADS;                      // 12 consecutive cells per thread
    int cnt[CPT], local = 0;
#pragma unroll
    for (int k = 0; k < CPT; k++) { cnt[k] = cur[tid * CPT + k]; local += cnt[k]; }
    int inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    int off = inc - local;
    for (int w = 0; w < wid; w++) off += wsum[w];
#pragma unroll
    for (int k = 0; k < CPT; k++) { cur[tid * CPT + k] = off; off += cnt[k]; }
    __syncthreads();
    for (int i = tid; i < n2; i += SG_THREADS) {
        const int c = kp2[i].octave == 0 ? pos_in_grid(a.grid, kp2[i].x, kp2[i].y) : -1;
        if (c >= 0) its[atomicAdd(&cur[c], 1)] = (unsigned short)i;
    }
    __syncthreads();
    for (int c = tid; c < NCELL; c += SG_THREADS) {              // insertion sort of the cell's items
        const int s0 = c ? cur[c - 1] : 0, e0 = cur[c];
        for (int i = s0 + 1; i < e0; i++) {
            const unsigned short v = its[i];
            int j = i - 1;
            while (j >= s0 && its[j] > v) { its[j + 1] = its[j]; j--; }
            its[j + 1] = v;
        }
    }
    __syncthreads();
    int* cell_end = a.cell_end + (long long)p * NCELL;
    unsigned short* items = a.items + (long long)p * a.cap;
    for (int c = tid; c < NCELL; c += SG_THREADS) cell_end[c] = cur[c];
    const int nin = cur[NCELL - 1];
    for (int i = tid; i < nin; i += SG_THREADS) items[i] = its[i];
    // queries: keypoints of the first frame on octave 0, in index order (:483-489)
    const int fa = a.pairs ? a.pairs[p].x : p;
    const int n1 = a.counts[fa];
    const orbfe_keypoint* kp1 = a.kps + (long long)fa * a.cap;
    unsigned short* qlist = a.qlist + (long long)p * a.cap;
    int nq = 0;
    for (int base = 0; base < n1; base += SG_THREADS) {
        const int i = base + tid;
        const bool ok = i < n1 && kp1[i].octave <= 0;
        const unsigned b = __ballot_sync(FULL, ok);
        __syncthreads();                                   // wsum of the previous round was consumed
        if (lane == 0) wsum[wid] = __popc(b);
        __syncthreads();
        int off = nq;
        for (int w = 0; w < wid; w++) off += wsum[w];
        if (ok) qlist[off + __popc(b & ((1u << lane) - 1u))] = (unsigned short)i;
        for (int w = 0; w < SG_THREADS / 32; w++) nq += wsum[w];
    }
    if (tid == 0) a.nq[p] = nq;
}

// Kernel A — one warp per query (octave-0 keypoint of the first frame): its SI_K best candidates
// in the second frame without the running filter (which depends on earlier queries).  SA_CTAS
// CTAs share a pair; a warp takes every (SA_CTAS * SI_WARPS)-th query.
constexpr int SA_CTAS = 4;
__global__ void __launch_bounds__(SI_WARPS * 32)
search_init_topk_kernel(const __grid_constant__ InitArgs a) {
    __shared__ unsigned short clist[SI_WARPS][SI_CL];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, p = blockIdx.y;
    const int nq = a.nq[p];
    int q = blockIdx.x * SI_WARPS + wid;
    if (q >= nq) return;
    const int fa = a.pairs ? a.pairs[p].x : p, fb = a.pairs ? a.pairs[p].y : p + 1;
    const orbfe_keypoint* kp1 = a.kps + (long long)fa * a.cap;
    const orbfe_keypoint* kp2 = a.kps + (long long)fb * a.cap;
    const uint8_t* d1 = a.desc + (long long)fa * a.cap * 32;
    const uint8_t* d2 = a.desc + (long long)fb * a.cap * 32;
    const float* prev = a.prev ? a.prev + (long long)p * a.cap * 2 : nullptr;
    const int* cell_end = a.cell_end + (long long)p * NCELL;
    const unsigned short* items = a.items + (long long)p * a.cap;
    const unsigned short* qlist = a.qlist + (long long)p * a.cap;
    QueryCtx c{kp2, d2, cell_end, items, nullptr, a.grid, (float)a.window};

    for (; q < nq; q += SA_CTAS * SI_WARPS) {
        const int i1 = qlist[q];
        const long long slot = (long long)p * a.cap + q;
        const float qx = prev ? prev[2 * i1] : kp1[i1].x, qy = prev ? prev[2 * i1 + 1] : kp1[i1].y;
        int cx0, cx1, cy0, cy1, total = 0;
        bool direct = false;
        __syncwarp();                                  // the previous query's clist readers are done
        if (cell_window(a.grid, qx, qy, c.r, cx0, cx1, cy0, cy1)) {
            // item range of every grid column of the window (one lane per column), then the
            // ranges are concatenated into the warp's candidate list in column order
            for (int cb = cx0; cb <= cx1 && !direct; cb += 32) {
                const int ix = cb + lane;
                int s0 = 0, len = 0;
                if (ix <= cx1) {
                    const int c0 = ix * GR + cy0;
                    s0 = c0 ? cell_end[c0 - 1] : 0;
                    len = cell_end[ix * GR + cy1] - s0;
                }
                int inc = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int u = __shfl_up_sync(FULL, inc, o);
                    if (lane >= o) inc += u;
                }
                const int sum = __shfl_sync(FULL, inc, 31);
                if (total + sum > SI_CL) { direct = true; break; }
                const int off = total + inc - len;
                for (int j = 0; j < len; j++) clist[wid][off + j] = items[s0 + j];
                total += sum;
            }
        }
        __syncwarp();
        if (direct) {
            if (lane == 0) a.tk_total[slot] = 0xFFFF;
            continue;
        }
        uint32_t qd[8];
        load_row(qd, d1 + 32 * (long long)i1);
        uint32_t top = KEY_INF;
        int npass = 0;
        for (int kb = 0; kb < total; kb += 32) {       // 32 candidates per round, one per lane
            const int j = kb + lane;
            uint32_t key = KEY_INF;
            if (j < total) {
                const int i2 = clist[wid][j];
                if (cand_ok(c, i2, qx, qy)) {
                    key = ((uint32_t)hamming256(qd, d2 + 32 * (long long)i2) << 20) | (uint32_t)j;
                    npass++;
                }
            }
            // keep the 32 smallest keys sorted across the lanes: min with the reversed sorted
            // block is a bitonic sequence holding them; five merge steps sort it
            const uint32_t blk = bitonic_sort32(key, lane);
            top = kb == 0 ? blk : bitonic_merge32(min(top, __shfl_sync(FULL, blk, 31 - lane)), lane);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) npass += __shfl_xor_sync(FULL, npass, o);
        if (lane < SI_K) {
            a.tk_key[slot * SI_K + lane] = top;
            a.tk_idx[slot * SI_K + lane] = top == KEY_INF ? 0 : clist[wid][top & 0xFFFFFu];
        }
        if (lane == 0) a.tk_total[slot] = (unsigned short)npass;
    }
}

// Kernel B — one warp per pair replays the reference's greedy bookkeeping over the queries in
// order, from the per-query candidate lists of kernel A (staged through shared memory SI_QB
// queries at a time).  The loop is one long dependent chain, so it is kept short: all state is
// in shared memory, every lane performs the same scalar updates (no divergent single-lane
// sections, no warp syncs: a lane only ever reads what it wrote itself), and the rotation
// histogram is built afterwards from the recorded (query, match) events.  A query whose kept
// list cannot decide falls back to the filtered scan of its window.
__global__ void __launch_bounds__(32)
search_init_greedy_kernel(const __grid_constant__ InitArgs a) {
    extern __shared__ __align__(16) unsigned char sm[];
    __shared__ __align__(16) uint32_t tk_key[SI_QB][SI_K];
    __shared__ __align__(16) unsigned short tk_idx[SI_QB][SI_K];
    __shared__ unsigned short tk_total[SI_QB];
    __shared__ unsigned short qi1[SI_QB];
    __shared__ int hist[32];
    const int lane = threadIdx.x, p = blockIdx.x;
    const int fa = a.pairs ? a.pairs[p].x : p, fb = a.pairs ? a.pairs[p].y : p + 1;
    const int n1 = a.counts[fa], n2 = a.counts[fb], nq = a.nq[p];
    const orbfe_keypoint* kp1 = a.kps + (long long)fa * a.cap;
    const orbfe_keypoint* kp2 = a.kps + (long long)fb * a.cap;
    const uint8_t* d1 = a.desc + (long long)fa * a.cap * 32;
    const uint8_t* d2 = a.desc + (long long)fb * a.cap * 32;
    int* m12 = a.m12 + (long long)p * a.cap;
    float* prev = a.prev ? a.prev + (long long)p * a.cap * 2 : nullptr;
    const unsigned short* qlist = a.qlist + (long long)p * a.cap;
    unsigned short* m12s = (unsigned short*)sm;                  // cap   vnMatches12 (0xFFFF = -1)
    unsigned short* m21 = m12s + a.cap;                          // cap   vnMatches21 (0xFFFF = none)
    unsigned short* ev = m21 + a.cap;                            // cap   match made by query i1 at its turn (rotHist entry)
    unsigned char* mdist = (unsigned char*)(ev + a.cap);         // cap   vMatchedDistance (255 = INT_MAX)
    for (int i = lane; i < n1; i += 32) { m12s[i] = 0xFFFF; ev[i] = 0xFFFF; }
    for (int i = lane; i < n2; i += 32) { m21[i] = 0xFFFF; mdist[i] = 255; }
    hist[lane] = 0;
    __syncwarp();
    QueryCtx c{kp2, d2, a.cell_end + (long long)p * NCELL, a.items + (long long)p * a.cap, mdist, a.grid, (float)a.window};
    const long long slot0 = (long long)p * a.cap;

    for (int qb = 0; qb < nq; qb += SI_QB) {
        const int qn = min(SI_QB, nq - qb);
        {   // stage the block's candidate lists (16-byte pieces; slots are 16-byte multiples)
            const uint4* gk = reinterpret_cast<const uint4*>(a.tk_key + (slot0 + qb) * SI_K);
            const uint4* gi = reinterpret_cast<const uint4*>(a.tk_idx + (slot0 + qb) * SI_K);
            for (int i = lane; i < qn * SI_K / 4; i += 32) reinterpret_cast<uint4*>(&tk_key[0][0])[i] = __ldg(gk + i);
            for (int i = lane; i < qn * SI_K / 8; i += 32) reinterpret_cast<uint4*>(&tk_idx[0][0])[i] = __ldg(gi + i);
            for (int i = lane; i < qn; i += 32) { tk_total[i] = a.tk_total[slot0 + qb + i]; qi1[i] = qlist[qb + i]; }
        }
        __syncwarp();
        for (int qi = 0; qi < qn; qi++) {
            const int total = tk_total[qi];
            uint32_t g1, g2;
            int best;
            bool need_full = total == 0xFFFF;
            if (!need_full) {
                const uint32_t key = tk_key[qi][lane & (SI_K - 1)];
                const int idx = tk_idx[qi][lane & (SI_K - 1)];
                const int md = mdist[idx];
                const bool alive = lane < SI_K && key != KEY_INF && !(md != 255 && md <= (int)(key >> 20));
                unsigned m = __ballot_sync(FULL, alive);
                const int l1 = __ffs(m) - 1;
                m &= m - 1;
                const int l2 = __ffs(m) - 1;
                g1 = __shfl_sync(FULL, key, max(l1, 0));
                best = __shfl_sync(FULL, idx, max(l1, 0));
                g2 = __shfl_sync(FULL, key, max(l2, 0));
                if (l1 < 0) g1 = KEY_INF;
                if (l2 < 0) g2 = KEY_INF;
                // the kept list may hide further candidates: only matters when a match is still possible
                if (total > SI_K && (l1 < 0 || ((int)(g1 >> 20) <= TH_LOW && l2 < 0))) need_full = true;
            }
            if (need_full) {
                const int i1 = qi1[qi];
                const float qx = prev ? prev[2 * i1] : kp1[i1].x, qy = prev ? prev[2 * i1 + 1] : kp1[i1].y;
                uint32_t qd[8];
                load_row(qd, d1 + 32 * (long long)i1);
                __syncwarp();
                eval_query_filtered(c, qd, qx, qy, g1, g2, best);
            }
            if (g1 == KEY_INF) continue;
            const int bestDist = (int)(g1 >> 20);
            const float second = g2 == KEY_INF ? (float)INT_MAX : (float)(int)(g2 >> 20);
            if (bestDist <= TH_LOW && (float)bestDist < __fmul_rn(second, a.ratio)) {   // :532-534
                const int i1 = qi1[qi];
                const int old = m21[best];
                if (old != 0xFFFF) m12s[old] = 0xFFFF;                                  // :536-540
                m12s[i1] = (unsigned short)best;
                m21[best] = (unsigned short)i1;
                mdist[best] = (unsigned char)bestDist;
                ev[i1] = (unsigned short)best;
            }
        }
        __syncwarp();
    }
    // rotation consistency (:549-590): every match event, also one displaced later, is a rotHist entry
    int ind1 = -1, ind2 = -1, ind3 = -1;
    if (a.checkOri) {
        const float factor = 1.0f / HISTO;                                              // :476
        for (int i = lane; i < n1; i += 32) {
            int bin = 255;
            if (ev[i] != 0xFFFF) {
                float rot = __fsub_rn(kp1[i].angle, kp2[ev[i]].angle);
                if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                bin = (int)roundf(__fmul_rn(rot, factor));
                if (bin == HISTO) bin = 0;
                atomicAdd(&hist[bin], 1);
            }
            ev[i] = (unsigned short)bin;
        }
        __syncwarp();
        three_maxima(hist, ind1, ind2, ind3);
    }
    int nm = 0;
    for (int i = lane; i < n1; i += 32) {                                               // :566-597
        int m = m12s[i] == 0xFFFF ? -1 : (int)m12s[i];
        if (a.checkOri) {
            const int b = ev[i];
            if (b != 255 && b != ind1 && b != ind2 && b != ind3) m = -1;
        }
        m12[i] = m;
        if (m >= 0) {
            nm++;
            if (prev) { prev[2 * i] = kp2[m].x; prev[2 * i + 1] = kp2[m].y; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nm += __shfl_xor_sync(FULL, nm, o);
    if (lane == 0) a.nmatches[p] = nm;
}

// ---------------------------------------------------------------------------------------
// SearchByProjection(Frame&, vector<MapPoint*>&, th)

__global__ void __launch_bounds__(kGridThreads) build_grid_kernel(const __grid_constant__ LocalArgs a) {
    __shared__ int cnt[kGridCells];
    __shared__ int wsum[kGridThreads / 32];
    block_grid_csr(cnt, wsum, a.n, [&](int i) { return pos_in_grid(a.grid, a.kps[i].x, a.kps[i].y); }, a.cell_end, a.items);
}

// Enumerates the candidates of one map point in GetFeaturesInArea order; FILL=false counts,
// FILL=true writes the packed list (keypoint index, Hamming distance, octave).  One warp per
// point: a window column is one contiguous run of the CSR (cells ix*48+cy0 .. ix*48+cy1), so
// the lanes take one column each, prefix-sum the run lengths and then walk the concatenated
// runs 32 items at a time — three dependent loads per round instead of one long serial chain
// per point.  Survivors keep their order through a ballot.
constexpr int LC_WARPS = 4;
template <bool FILL>
__global__ void __launch_bounds__(LC_WARPS * 32) local_candidates_kernel(const __grid_constant__ LocalArgs a) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * LC_WARPS + (threadIdx.x >> 5);
    if (i >= a.nmp) return;
    if (FILL && a.cand_off[a.nmp] > a.cand_cap) return;      // list does not fit: the host grows the buffer and retries
    int count = 0;
    uint32_t* out = FILL ? a.cand + a.cand_off[i] : nullptr;
    if (a.in_view[i]) {                                                                 // :58-62
        float rs;
        int minLevel, maxLevel;
        if (a.mode == 0) {
            const int lvl = a.level[i];
            float r = (double)a.view_cos[i] > 0.998 ? 2.5f : 4.0f;                      // :146-152 (double compare)
            if (a.th != 1.0f) r = __fmul_rn(r, a.th);                                   // :56,:69-70
            rs = __fmul_rn(r, a.scale[lvl]);
            minLevel = lvl - 1; maxLevel = lvl;
        } else {                                                                        // :1542-1553
            rs = a.radius[i]; minLevel = a.min_level[i]; maxLevel = a.max_level[i];
        }
        const float x = a.px[i], y = a.py[i];
        const bool check = (minLevel > 0) || (maxLevel >= 0);                           // Frame.cc:381
        int cx0, cx1, cy0, cy1;
        if (cell_window(a.grid, x, y, rs, cx0, cx1, cy0, cy1)) {
            uint32_t q[8];
            if (FILL) load_row(q, a.mpdesc + 32 * (long long)i);
            const float pxr = a.uright ? a.pxr[i] : 0.0f;
            for (int cbase = cx0; cbase <= cx1; cbase += 32) {
                const int ix = cbase + lane;
                int s = 0, len = 0;
                if (ix <= cx1) {
                    const int c0 = ix * GR + cy0;
                    s = c0 ? a.cell_end[c0 - 1] : 0;
                    len = a.cell_end[ix * GR + cy1] - s;
                }
                int incl = len;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int u = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += u;
                }
                const int total = __shfl_sync(FULL, incl, 31);
                for (int t0 = 0; t0 < total; t0 += 32) {
                    const int t = t0 + lane;
                    // column of item t: the first lane whose inclusive prefix exceeds t
                    int c = 0;
#pragma unroll
                    for (int step = 16; step > 0; step >>= 1) {
                        const int v = __shfl_sync(FULL, incl, c + step - 1);
                        if (v <= t) c += step;
                    }
                    c = min(c, 31);
                    const int cs = __shfl_sync(FULL, s, c), cincl = __shfl_sync(FULL, incl, c), clen = __shfl_sync(FULL, len, c);
                    bool pass = t < total;
                    int idx = 0, oct = 0;
                    if (pass) {
                        idx = a.items[cs + (t - (cincl - clen))];
                        const orbfe_keypoint* kp = a.kps + idx;
                        oct = kp->octave;
                        if (check && (oct < minLevel || (maxLevel >= 0 && oct > maxLevel))) pass = false;
                        if (pass && !(fabsf(__fsub_rn(kp->x, x)) < rs && fabsf(__fsub_rn(kp->y, y)) < rs)) pass = false;
                        if (pass && a.uright) {                                         // :100-105
                            const float ur = a.uright[idx];
                            if (ur > 0 && fabsf(__fsub_rn(pxr, ur)) > rs) pass = false;
                        }
                    }
                    const unsigned m = __ballot_sync(FULL, pass);
                    if (FILL && pass) {
                        const int dist = hamming256(q, a.desc + 32 * (long long)idx);
                        out[count + __popc(m & ((1u << lane) - 1u))] =
                            (uint32_t)idx | ((uint32_t)dist << 16) | ((uint32_t)(oct & 0x3F) << 25);
                    }
                    count += __popc(m);
                }
            }
        }
    }
    if (!FILL && lane == 0) a.cand_off[i] = count;
}

// single-CTA exclusive scan of cand_off[0..n) in place; cand_off[n] = total
__global__ void __launch_bounds__(1024) scan_kernel(int* data, int n) {
    __shared__ int wsum[33];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = i < n ? data[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(FULL, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const int w = wsum[lane];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(FULL, wi, o); if (lane >= o) wi += u; }
            wsum[lane] = wi - w;
            if (lane == 31) wsum[32] = wi;
        }
        __syncthreads();
        if (i < n) data[i] = carry + wsum[wid] + inc - v;
        __syncthreads();
        if (tid == 0) carry += wsum[32];
        __syncthreads();
    }
    if (tid == 0) data[n] = carry;
}

// One fixed-point sweep: map point i ignores keypoints owned by an earlier map point.
__global__ void __launch_bounds__(128) local_select_kernel(const __grid_constant__ LocalArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (a.owner_pp && i < a.n && a.owner_pp[i] != a.owner_prev[i]) *a.diff_flag = 1;   // did the previous sweep change anything?
    if (i >= a.nmp) return;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;   // :83-87
    const int s = a.cand_off[i], e = a.cand_off[a.nmp] > a.cand_cap ? s : a.cand_off[i + 1];
    for (int j = s; j < e; j++) {
        const uint32_t c = a.cand[j];
        const int idx = (int)(c & 0xFFFFu), dist = (int)((c >> 16) & 0x1FFu), oct = (int)(c >> 25);
        if (a.occ0[idx] || a.owner_prev[idx] < i) continue;                             // :95-98
        if (dist < bestDist) {                                                          // :109-123
            bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx;
        } else if (dist < bestDist2) {
            bestLevel2 = oct; bestDist2 = dist;
        }
    }
    int ch = -1;
    if (a.mode == 1) {
        if (bestDist <= a.max_dist) ch = bestIdx;                                       // :1585-1589 / :1734 / :440, best only
    } else if (bestDist <= TH_HIGH &&                                                   // :127-135
               !(bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(a.ratio, (float)bestDist2))) {
        ch = bestIdx;
    }
    a.choice[i] = ch;
    if (ch >= 0 && a.obs[i] > 0) atomicMin(&a.owner_next[ch], i);
}

__global__ void __launch_bounds__(256)
guided_init_kernel(int* owners, int nown, int* mp_for_kp, int n, int* zero_block, int nzero, uint8_t* occ_out,
                   const uint8_t* occ0) {
    const int stride = gridDim.x * blockDim.x;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = t; i < nown; i += stride) owners[i] = 0x7F7F7F7F;
    for (int i = t; i < n; i += stride) { mp_for_kp[i] = -1; occ_out[i] = occ0[i]; }
    for (int i = t; i < nzero; i += stride) zero_block[i] = 0;
}

// results: mp_for_kp[idx] = last map point that took idx, occupied |= owned, nmatches
__global__ void __launch_bounds__(128)
local_finalize_kernel(const __grid_constant__ LocalArgs a, int* mp_for_kp, uint8_t* occ_out, int* nmatches) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *a.ncand_out = a.cand_off[a.nmp];
    if (a.owner_pp && i < a.n && a.owner_pp[i] != a.owner_prev[i]) *a.diff_flag = 1;   // last sweep still moving
    if (i < a.n && a.owner_prev[i] != 0x7F7F7F7F) occ_out[i] = 1;   // 0x7F7F7F7F = no owner
    if (i < a.nmp && a.choice[i] >= 0) {
        atomicMax(&mp_for_kp[a.choice[i]], i);
        atomicAdd(nmatches, 1);
    }
}

// SearchByProjection(CurrentFrame, LastFrame) epilogue (:1589-1628): every accepted point is a
// match event (a keypoint can be taken more than once when the earlier taker has no
// observations); events whose rotation bin is not one of the three maxima clear their keypoint
// and are subtracted from the count.
__global__ void __launch_bounds__(128) frame_hist_kernel(const __grid_constant__ LocalArgs a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.nmp || a.choice[i] < 0) return;
    float rot = __fsub_rn(a.angle[i], a.kps[a.choice[i]].angle);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / HISTO));
    if (bin == HISTO) bin = 0;
    atomicAdd(&a.hist[bin], 1);
}

__global__ void __launch_bounds__(128)
frame_finalize_kernel(const __grid_constant__ LocalArgs a, int* mp_for_kp, int* nmatches) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *a.ncand_out = a.cand_off[a.nmp];
    if (a.owner_pp && i < a.n && a.owner_pp[i] != a.owner_prev[i]) *a.diff_flag = 1;   // last sweep still moving
    if (i >= a.nmp || a.choice[i] < 0) return;
    const int k = a.choice[i];
    bool bad = false;
    if (a.checkOri) {
        int ind1, ind2, ind3;
        three_maxima(a.hist, ind1, ind2, ind3);
        float rot = __fsub_rn(a.angle[i], a.kps[k].angle);
        if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
        int bin = (int)roundf(__fmul_rn(rot, 1.0f / HISTO));
        if (bin == HISTO) bin = 0;
        bad = bin != ind1 && bin != ind2 && bin != ind3;
    }
    if (bad) a.kp_bad[k] = 1;
    else atomicAdd(nmatches, 1);
    atomicMax(&mp_for_kp[k], i);
}

// ---------------------------------------------------------------------------------------
// K14: best keypoint inside the search window of every projected map point, each point on its
// own — the candidate loops of Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:1007-1052),
// Fuse(pKF, Scw, ...) (:1180-1204) and both directions of SearchBySim3 (:1333-1365, :1411-1443).
// One warp per point: window columns of KeyFrame::GetFeaturesInArea (KeyFrame.cc:611-650) are
// contiguous CSR runs, the lanes stride a run; key = distance << 20 | traversal rank keeps the
// reference's first-wins rule under the warp minimum.
constexpr int WB_WARPS = 4;
__global__ void __launch_bounds__(WB_WARPS * 32) window_best_kernel(const __grid_constant__ WindowArgs a) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * WB_WARPS + (threadIdx.x >> 5);
    if (i >= a.np) return;
    uint32_t best = KEY_INF;
    int bi = -1;
    if (a.valid[i]) {
        const float x = a.u[i], y = a.v[i], r = a.radius[i];
        const int lvl = a.level[i];
        int cx0, cx1, cy0, cy1;
        if (cell_window(a.grid, x, y, r, cx0, cx1, cy0, cy1)) {
            uint32_t q[8];
            load_row(q, a.pdesc + 32 * (long long)i);
            const float pur = (a.inv_sigma2 && a.uright) ? a.ur[i] : 0.0f;
            int seq = 0;
            for (int ix = cx0; ix <= cx1; ix++) {
                const int c0 = ix * GR + cy0;
                const int s = c0 ? a.cell_end[c0 - 1] : 0, e = a.cell_end[ix * GR + cy1];
                for (int j = s + lane; j < e; j += 32) {
                    const int idx = a.items[j];
                    const orbfe_keypoint kp = a.kps[idx];
                    const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
                    if (!(fabsf(dx) < r && fabsf(dy) < r)) continue;                      // KeyFrame.cc:640-644
                    if (kp.octave < lvl - 1 || kp.octave > lvl) continue;                 // :1016 / :1187 / :1340
                    if (a.inv_sigma2) {                                                   // :1020-1041
                        const float ex = __fsub_rn(x, kp.x), ey = __fsub_rn(y, kp.y);
                        float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                        const float kur = a.uright ? a.uright[idx] : -1.0f;
                        double gate = 5.99;
                        if (kur >= 0) {
                            const float er = __fsub_rn(pur, kur);
                            e2 = __fadd_rn(e2, __fmul_rn(er, er));
                            gate = 7.8;
                        }
                        if ((double)__fmul_rn(e2, a.inv_sigma2[kp.octave]) > gate) continue;
                    }
                    const int dist = hamming256(q, a.desc + 32 * (long long)idx);
                    const uint32_t key = ((uint32_t)dist << 20) | (uint32_t)(seq + j - s);
                    if (key < best) { best = key; bi = idx; }
                }
                seq += e - s;
            }
        }
    }
    const uint32_t g = warp_min(best);
    int idx = -1, dist = 256;
    if (g != KEY_INF) {
        const unsigned own = __ballot_sync(FULL, best == g);
        idx = __shfl_sync(FULL, bi, __ffs(own) - 1);
        dist = (int)(g >> 20);
    }
    if (lane == 0) {
        a.best_dist[i] = dist;
        a.best_idx[i] = dist <= a.max_dist ? idx : -1;
    }
}

// SearchBySim3's agreement test (:1446-1460): keep (i1, idx2) where the reverse search of idx2 returned i1
__global__ void __launch_bounds__(256) sim3_agree_kernel(const int* match1, const int* match2, int n1, int n2, int* m12,
                                                         int* nfound) {
    const int i1 = blockIdx.x * blockDim.x + threadIdx.x;
    if (i1 >= n1) return;
    const int idx2 = match1[i1];
    const bool ok = idx2 >= 0 && idx2 < n2 && match2[idx2] == i1;
    m12[i1] = ok ? idx2 : -1;
    if (ok) atomicAdd(nfound, 1);
}

// ---------------------------------------------------------------------------------------
// K10: Frame::ComputeStereoMatches (src/Frame.cc:512-686).  One warp per left keypoint: the
// row-band candidate search scans the right keypoints directly (a right keypoint belongs to
// the band of row (int)vL iff floor(y - r) <= (int)vL <= ceil(y + r), r = 2*scale[octave],
// which is what the reference's vRowIndices buckets encode, :529-539), keeps the first minimum
// (ascending iR, strict '<'), then refines by sliding an 11x11 window over 11 shifts with the
// L1 norm of the centre-subtracted patches (integer-valued, so exact in int32) and fits the
// parabola in the reference's float op order.  A second single-CTA kernel reproduces the
// median rule (:671-685) with a rank selection.

// the operands of one stereo pair: the arguments themselves, or pair p of an interleaved extractor batch
struct StereoPair {
    int nl, nr;
    const orbfe_keypoint* kl; const orbfe_keypoint* kr;
    const uint8_t* dl; const uint8_t* dr;
    float* uright; float* depth; int* sad;
    long long fl, fr;        // frame indices of the left / right image inside the level stacks
};
__device__ __forceinline__ StereoPair stereo_pair(const StereoArgs& a, int p) {
    StereoPair s;
    if (!a.batched) {
        s.nl = a.nl; s.nr = a.nr; s.kl = a.kl; s.kr = a.kr; s.dl = a.dl; s.dr = a.dr;
        s.uright = a.uright; s.depth = a.depth; s.sad = a.sad; s.fl = s.fr = 0;
        return s;
    }
    const long long fl = 2LL * p, fr = fl + 1;
    s.nl = a.counts[fl]; s.nr = a.counts[fr];
    s.kl = a.kl + fl * a.cap; s.kr = a.kl + fr * a.cap;
    s.dl = a.dl + fl * a.cap * 32; s.dr = a.dl + fr * a.cap * 32;
    s.uright = a.uright + (long long)p * a.cap; s.depth = a.depth + (long long)p * a.cap; s.sad = a.sad + (long long)p * a.cap;
    s.fl = fl; s.fr = fr;
    return s;
}

__global__ void __launch_bounds__(128) stereo_match_kernel(const __grid_constant__ StereoArgs a) {
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * 4 + (threadIdx.x >> 5);
    const StereoPair P = stereo_pair(a, blockIdx.y);
    if (a.batched && iL >= P.nl && iL < a.cap && lane == 0) { P.uright[iL] = -1.0f; P.depth[iL] = -1.0f; P.sad[iL] = -1; }
    if (iL >= P.nl) return;
    if (lane == 0) { P.uright[iL] = -1.0f; P.depth[iL] = -1.0f; P.sad[iL] = -1; }
    const orbfe_keypoint kpL = P.kl[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int row = (int)vL;
    if (row < 0 || row >= a.nrows) return;
    const float minZ = a.mb, minD = 0.0f, maxD = __fdiv_rn(a.mbf, minZ);                 // :542-545
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    uint32_t q[8];
    load_row(q, P.dl + 32 * (long long)iL);
    uint32_t best = KEY_INF;
    int any = 0;
    for (int iR = lane; iR < P.nr; iR += 32) {
        const orbfe_keypoint kpR = P.kr[iR];
        const float r = __fmul_rn(2.0f, a.scale[kpR.octave]);                           // :532
        const int maxr = (int)ceilf(__fadd_rn(kpR.y, r)), minr = (int)floorf(__fsub_rn(kpR.y, r));
        if (row < minr || row > maxr) continue;
        any = 1;
        if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;               // :582-583
        if (!(kpR.x >= minU && kpR.x <= maxU)) continue;                                // :587
        const int dist = hamming256(q, P.dr + 32 * (long long)iR);
        best = min(best, ((uint32_t)dist << 20) | (uint32_t)iR);
    }
    any = __any_sync(FULL, any);
    best = warp_min(best);
    if (!any || maxU < 0) return;                                                       // :565-572
    const int bestDist = best == KEY_INF ? TH_HIGH : (int)(best >> 20);
    if (!(bestDist < TH_HIGH)) return;                   // 'dist<bestDist' with bestDist = TH_HIGH (:574,:591)
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;                                       // :517
    if (!(bestDist < thOrbDist)) return;                                                // :598
    const int bestIdxR = (int)(best & 0xFFFFFu);

    // ---- sub-pixel refinement by correlation (:600-668)
    const float uR0 = P.kr[bestIdxR].x;
    const float sf = a.inv_scale[levelL];
    const float scaleduL = roundf(__fmul_rn(kpL.x, sf)), scaledvL = roundf(__fmul_rn(kpL.y, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, Lh = 5;
    const uint8_t* IL = a.imgl[levelL] + P.fl * a.frameStride[levelL]; const int sl = a.pitchl[levelL];
    const uint8_t* IR = a.imgr[levelL] + P.fr * a.frameStride[levelL]; const int sr = a.pitchr[levelL];
    const int v0 = (int)(scaledvL - w), uL0 = (int)(scaleduL - w);
    const float iniu = scaleduR0 + Lh - w, endu = scaleduR0 + Lh + w + 1;
    if (iniu < 0 || endu >= (float)a.widthr[levelL]) return;                            // :621
    const int cL = IL[(long long)(v0 + w) * sl + uL0 + w];
    // left patch values of this lane: pixels lane, lane+32, ... of the 121
    int pl[4], py[4], px[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int i = lane + 32 * k;
        py[k] = i / 11; px[k] = i - py[k] * 11;
        pl[k] = i < 121 ? (int)IL[(long long)(v0 + py[k]) * sl + uL0 + px[k]] - cL : 0;
    }
    float vDists[11];
    float sadBestF = (float)INT_MAX;
    int sadBest = INT_MAX, bestinc = 0;
#pragma unroll
    for (int inc = -5; inc <= 5; inc++) {
        const int uR0i = (int)(scaleduR0 + (float)inc - w);
        const int cR = IR[(long long)(v0 + w) * sr + uR0i + w];
        int s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = lane + 32 * k;
            if (i < 121) s += abs(pl[k] - ((int)IR[(long long)(v0 + py[k]) * sr + uR0i + px[k]] - cR));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
        const float dist = (float)s;
        if (dist < sadBestF) { sadBest = (int)dist; sadBestF = (float)sadBest; bestinc = inc; }   // :630-634 (int bestDist)
        vDists[inc + 5] = dist;
    }
    if (bestinc == -Lh || bestinc == Lh) return;                                        // :640
    float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
    for (int k = 1; k < 10; k++)
        if (k == bestinc + 5) { dist1 = vDists[k - 1]; dist2 = vDists[k]; dist3 = vDists[k + 1]; }
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                   __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));   // :648
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(a.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));   // :654
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= minD && disparity < maxD) {                                        // :658-668
        if (disparity <= 0) { disparity = 0.01f; bestuR = __fsub_rn(uL, 0.01f); }
        if (lane == 0) {
            P.depth[iL] = __fdiv_rn(a.mbf, disparity);
            P.uright[iL] = bestuR;
            P.sad[iL] = sadBest;
        }
    }
}

// median of the truncated SADs (element size/2 of the sorted list, found by bisection on the value) and the 1.5*1.4*median cut
__global__ void __launch_bounds__(256) stereo_median_kernel(const __grid_constant__ StereoArgs a) {
    constexpr int MCAP = 4096;                       // matched keypoints staged in shared memory (more: global path)
    __shared__ int s_cnt, s_median;
    __shared__ int s_vals[MCAP];
    const int tid = threadIdx.x;
    const StereoPair P = stereo_pair(a, blockIdx.x);
    if (tid == 0) { s_cnt = 0; s_median = -1; }
    __syncthreads();
    for (int i = tid; i < P.nl; i += 256) {
        const int v = P.sad[i];
        if (v >= 0) {
            const int slot = atomicAdd(&s_cnt, 1);   // the order of the list does not matter to a rank selection
            if (slot < MCAP) s_vals[slot] = v;
        }
    }
    __syncthreads();
    const int m = s_cnt;
    if (m == 0) return;
    const int target = m / 2;                                                           // :672
    // element `target` of the sorted list = the smallest value v with #{u <= v} > target: bisection on the value with
    // block-wide counts (the SADs are < 2^17; the first loop only runs if that ever stops being true)
    const bool staged = m <= MCAP;
    auto count_le = [&](int v) {
        int total = 0;
        if (staged) {
            for (int base = 0; base < m; base += 256) total += __syncthreads_count(base + tid < m && s_vals[base + tid] <= v);
        } else {
            for (int base = 0; base < P.nl; base += 256) {
                const int u = base + tid < P.nl ? P.sad[base + tid] : -1;
                total += __syncthreads_count(u >= 0 && u <= v);
            }
        }
        return total;
    };
    int lo = -1, hi = (1 << 17) - 1;
    while (hi < 0x3FFFFFFF && count_le(hi) <= target) hi = 2 * hi + 1;
    while (hi - lo > 1) {
        const int mid = lo + (hi - lo) / 2;
        if (count_le(mid) > target) hi = mid; else lo = mid;
    }
    if (tid == 0) s_median = hi;
    __syncthreads();
    const float thDist = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)s_median);             // :673
    for (int i = tid; i < P.nl; i += 256) {
        const int v = P.sad[i];
        if (v >= 0 && !((float)v < thDist)) { P.uright[i] = -1.0f; P.depth[i] = -1.0f; }   // :675-685
    }
}

}  // namespace

// HBM scratch of the three kernels, per pair: CSR grid, items, and per first-frame keypoint the
// SI_K kept keys / indices and the candidate count
size_t search_init_scratch_bytes(int cap, int npairs) {
    const size_t per = (size_t)NCELL * 4 + (size_t)cap * 2 * 2 + (size_t)cap * SI_K * 4 + (size_t)cap * SI_K * 2 + (size_t)cap * 2 + 4;
    return per * npairs + 7 * 256;
}

void search_init_carve(InitArgs& a, void* scratch, int npairs) {
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    unsigned char* b = (unsigned char*)scratch;
    a.cell_end = (int*)b;            b += up((size_t)npairs * NCELL * 4);
    a.tk_key = (uint32_t*)b;         b += up((size_t)npairs * a.cap * SI_K * 4);
    a.tk_idx = (unsigned short*)b;   b += up((size_t)npairs * a.cap * SI_K * 2);
    a.items = (unsigned short*)b;    b += up((size_t)npairs * a.cap * 2);
    a.qlist = (unsigned short*)b;    b += up((size_t)npairs * a.cap * 2);
    a.tk_total = (unsigned short*)b; b += up((size_t)npairs * a.cap * 2);
    a.nq = (int*)b;
}

int launch_search_init(const InitArgs& a, int npairs, cudaStream_t st) {
    if (npairs <= 0) return 0;
    const size_t gsm = (size_t)NCELL * 4 + (size_t)a.cap * 2;
    // per function and device, shared by every handle and thread: always the one worst case, never a per-call size
    if (gsm > 40 * 1024) cudaFuncSetAttribute(search_init_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    search_init_grid_kernel<<<npairs, SG_THREADS, gsm, st>>>(a);
    search_init_topk_kernel<<<dim3(SA_CTAS, npairs), SI_WARPS * 32, 0, st>>>(a);
    const size_t bsm = (size_t)a.cap * 7 + 16;
    if (bsm > 24 * 1024) cudaFuncSetAttribute(search_init_greedy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    search_init_greedy_kernel<<<npairs, 32, bsm, st>>>(a);
    return 3;
}

int launch_local_grid_and_count(const LocalArgs& a, cudaStream_t st, bool build_grid) {
    if (build_grid) build_grid_kernel<<<1, kGridThreads, 0, st>>>(a);
    if (a.nmp > 0) local_candidates_kernel<false><<<(a.nmp + LC_WARPS - 1) / LC_WARPS, LC_WARPS * 32, 0, st>>>(a);
    scan_kernel<<<1, 1024, 0, st>>>(a.cand_off, a.nmp);
    return 3;
}

int launch_local_fill(const LocalArgs& a, cudaStream_t st) {
    if (a.nmp > 0) local_candidates_kernel<true><<<(a.nmp + LC_WARPS - 1) / LC_WARPS, LC_WARPS * 32, 0, st>>>(a);
    return 1;
}

int launch_local_sweep(const LocalArgs& a, cudaStream_t st) {
    const int m = a.n > a.nmp ? a.n : a.nmp;
    if (m > 0) local_select_kernel<<<(m + 127) / 128, 128, 0, st>>>(a);
    return 2;
}

int launch_local_finalize(const LocalArgs& a, int* mp_for_kp, uint8_t* occ_out, int* nmatches, cudaStream_t st) {
    const int m = a.n > a.nmp ? a.n : a.nmp;
    if (m > 0) local_finalize_kernel<<<(m + 127) / 128, 128, 0, st>>>(a, mp_for_kp, occ_out, nmatches);
    return 1;
}

int launch_frame_finalize(const LocalArgs& a, int* mp_for_kp, int* nmatches, cudaStream_t st) {
    if (a.nmp <= 0) return 0;
    if (a.checkOri) frame_hist_kernel<<<(a.nmp + 127) / 128, 128, 0, st>>>(a);
    const int m = a.n > a.nmp ? a.n : a.nmp;
    frame_finalize_kernel<<<(m + 127) / 128, 128, 0, st>>>(a, mp_for_kp, nmatches);
    return a.checkOri ? 2 : 1;
}

int launch_guided_init(int* owners, int n, int nsweeps, bool reset_first, int* mp_for_kp, int* zero_block, int nzero,
                       uint8_t* occ_out, const uint8_t* occ0, cudaStream_t st) {
    int* first = reset_first ? owners : owners + n;
    const int nown = (nsweeps + (reset_first ? 1 : 0)) * n;
    const int work = nown > nzero ? nown : nzero;
    guided_init_kernel<<<(work + 1023) / 1024 > 0 ? (work + 1023) / 1024 : 1, 256, 0, st>>>(first, nown, mp_for_kp, n, zero_block, nzero,
                                                                                         occ_out, occ0);
    return 1;
}

int launch_build_grid(const orbfe_keypoint* kps, int n, const GridGeom& grid, int* cell_end, unsigned short* items,
                      cudaStream_t st) {
    LocalArgs a{};
    a.kps = kps; a.n = n; a.grid = grid; a.cell_end = cell_end; a.items = items;
    build_grid_kernel<<<1, kGridThreads, 0, st>>>(a);
    return 1;
}

int launch_window_best(const WindowArgs& a, cudaStream_t st) {
    if (a.np <= 0) return 0;
    window_best_kernel<<<(a.np + WB_WARPS - 1) / WB_WARPS, WB_WARPS * 32, 0, st>>>(a);
    return 1;
}

int launch_sim3_agree(const int* match1, const int* match2, int n1, int n2, int* m12, int* nfound, cudaStream_t st) {
    if (n1 <= 0) return 0;
    sim3_agree_kernel<<<(n1 + 255) / 256, 256, 0, st>>>(match1, match2, n1, n2, m12, nfound);
    return 1;
}

int launch_stereo(const StereoArgs& a, cudaStream_t st, int npairs) {
    const int rows = a.batched ? a.cap : a.nl;      // batched: one warp per keypoint slot (counts live on the device)
    if (rows <= 0 || npairs <= 0) return 0;
    stereo_match_kernel<<<dim3((rows + 3) / 4, npairs), 128, 0, st>>>(a);
    stereo_median_kernel<<<npairs, 256, 0, st>>>(a);
    return 2;
}

}  // namespace orbfe
