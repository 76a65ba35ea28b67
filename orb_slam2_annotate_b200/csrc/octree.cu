// octree.cu — K3: DistributeOctTree (reference src/ORBextractor.cc:566-808, DivideNode
// :498-558), one CTA per (frame, level) tree.
//
// The reference walks a std::list with push_front/erase and, in its last phase, sorts
// (count, node pointer) pairs.  Here the same sequence of decisions is produced by a
// level-synchronous formulation that needs no list surgery:
//   * the list is an array of nodes in list order (front first), rebuilt every step;
//   * every keypoint carries the position of the node that owns it; a node's vKeys order is
//     the original input order (DivideNode partitions stably), so "first maximum response"
//     (:789-805) is "maximum response, lowest input index" and the keys never need moving;
//   * a step splits a prefix of the candidates (nodes with more than one key) taken in the
//     reference's processing order: list order in the breadth-first phase (:639-699),
//     (count descending, newest node first) in the budgeted phase (:714-782, with the pointer
//     tie-break pinned to node creation order — SURVEY.md §7 hard part 1).  Children are
//     pushed to the front one parent at a time, so after the step the list reads: children of
//     the last processed parent (n4,n3,n2,n1), ..., children of the first processed parent,
//     then the untouched nodes in their old order.  "Newest first" among equal counts is
//     therefore "smaller list position first";
//   * the early break at size >= N (:774) is the first prefix of the processing order whose
//     running size reaches N, found with a prefix sum over the children counts.
// Prefix sums, the stable rank sort and the per-key updates are block-parallel (256 threads).
#include <cstdio>
#include <cstdlib>

#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int kPtsSmemMax = 4096;  // most keys a tree keeps in shared memory; larger levels use HBM scratch
constexpr unsigned FULL = 0xFFFFFFFFu;


template <int QT>
__device__ __forceinline__ int block_exscan(int* a, int n, int* wsum) {
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (n + QT - 1) / QT;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; i++) s += a[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int v = lane < QT / 32 ? wsum[lane] : 0;
        int vi = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(FULL, vi, o);
            if (lane >= o) vi += u;
        }
        if (lane < QT / 32) wsum[lane] = vi - v;
        if (lane == QT / 32 - 1) wsum[QT / 32] = vi;
    }
    __syncthreads();
    int off = wsum[wid] + inc - s;
    for (int i = b; i < e; i++) { const int v = a[i]; a[i] = off; off += v; }
    const int total = wsum[QT / 32];
    __syncthreads();
    return total;
}

__device__ __forceinline__ int quadrant(uint32_t p, short4 b) {
    const int mx = b.x + ((b.z - b.x + 1) >> 1), my = b.y + ((b.w - b.y + 1) >> 1);   // ceil(/2) :500-501
    const int x = kp_x(p), y = kp_y(p);
    return x < mx ? (y < my ? 0 : 2) : (y < my ? 1 : 3);                               // :532-546
}
__device__ __forceinline__ short4 child_box(short4 b, int q) {
    const short mx = b.x + ((b.z - b.x + 1) >> 1), my = b.y + ((b.w - b.y + 1) >> 1);
    switch (q) {                                                                       // :503-527
        case 0: return make_short4(b.x, b.y, mx, my);
        case 1: return make_short4(mx, b.y, b.z, my);
        case 2: return make_short4(b.x, my, mx, b.w);
        default: return make_short4(mx, my, b.z, b.w);
    }
}

__device__ __forceinline__ int child_count(const int* c4) {
    return (c4[0] > 0) + (c4[1] > 0) + (c4[2] > 0) + (c4[3] > 0);
}

template <int QT>
__global__ void __launch_bounds__(QT)
octree_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ DevBufs d, int NC, int KP, int level0) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int wsum[QT / 32 + 1];
    __shared__ int s_n, s_k0, s_T, s_ns, s_nexp;

    const int l = blockIdx.x + level0, frame = blockIdx.y + g.frame0;
    const LevelGeom& L = g.lv[l];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int N = L.nFeat;

    // the arena as plain locals (a struct with box[cur] indexing ends up in local memory: every access then reloads
    // its pointer from the stack and loses the shared address space)
    short4* const box0 = (short4*)smem_raw;
    short4* const box1 = box0 + NC;
    int* const cnt0 = (int*)(box1 + NC);
    int* const cnt1 = cnt0 + NC;
    int* const s_cnt4 = cnt1 + NC;      // 4 per node: keys per quadrant of a candidate
    int* const s_cand = s_cnt4 + 4 * NC;   // compacted candidate positions (list order)
    int* const s_order = s_cand + NC;   // processing order -> node position
    int* const s_pex = s_order + NC;    // exclusive prefix of children counts, in processing order
    int* const s_prank = s_pex + NC;    // per node: processing rank if split this step, else -1
    int* const s_remap = s_prank + NC;  // per node: new position (survivor) or first child position (split)
    int* const s_tmp = s_remap + NC;    // scan scratch
    uint32_t* const s_pxy = (uint32_t*)(s_tmp + NC);
    uint32_t* const s_pnode = s_pxy + KP;
    int* selc_out = d.selc + (long long)frame * g.nlevels + l;
    uint32_t* sel_out = d.sel + (long long)frame * g.selTotal + L.selBase;

    // ---- gather this level's corners in the reference's vToDistributeKeys order (cell row,
    //      cell column, then row-major inside the cell: ORBextractor.cc:846-894)
    const int ncell = L.nCols * L.nRows;
    const int* cc = d.cellc + (long long)frame * g.ncells + L.cellBase;
    const uint32_t* ck = d.cellk + (long long)frame * g.cellSlots + L.cellSlotOff;
    if (tid == 0) s_n = 0;
    __syncthreads();
    {
        int local = 0;
        for (int c = tid; c < ncell; c += QT) local += cc[c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(FULL, local, o);
        if (lane == 0 && local) atomicAdd(&s_n, local);
    }
    __syncthreads();
    const int n = s_n;
    if (n == 0) {
        if (tid == 0) *selc_out = 0;
        return;
    }
    uint32_t* pxy = s_pxy;
    uint32_t* pnode = s_pnode;
    if (n > KP) {
        const long long off = (long long)frame * g.cellSlots + L.cellSlotOff;
        pxy = d.qt_xy + off;
        pnode = d.qt_node + off;
    }
    {
        int run = 0;
        for (int base = 0; base < ncell; base += NC) {
            const int m = min(NC, ncell - base);
            for (int c = tid; c < m; c += QT) s_tmp[c] = cc[base + c];
            __syncthreads();
            const int tot = block_exscan<QT>(s_tmp, m, wsum);
            // one thread per cell: cells hold a handful of corners, and this way a CTA has
            // hundreds of independent loads in flight instead of one short row per warp
            for (int c = tid; c < m; c += QT) {
                const int k = cc[base + c];
                const uint32_t* src = ck + (long long)(base + c) * L.cellCap;
                uint32_t* dst = pxy + run + s_tmp[c];
                for (int i = 0; i < k; i++) dst[i] = __ldg(src + i);
            }
            run += tot;
            __syncthreads();
        }
    }

    // ---- roots (:570-615): nIni boxes of width hX; empty roots are erased
    const int nIni = L.nIni;
    for (int i = tid; i < nIni; i += QT) {
        box1[i] = make_short4((short)(int)__fmul_rn(L.hX, (float)i), 0,
                                  (short)(int)__fmul_rn(L.hX, (float)(i + 1)),
                                  (short)(L.maxBY - L.minBY));
        cnt1[i] = 0;
    }
    __syncthreads();
    for (int i = tid; i < n; i += QT) {
        const int r = min((int)__fdiv_rn((float)kp_x(pxy[i]), L.hX), nIni - 1);   // :598
        pnode[i] = r;
        atomicAdd(&cnt1[r], 1);
    }
    __syncthreads();
    for (int i = tid; i < nIni; i += QT) s_tmp[i] = cnt1[i] > 0;
    __syncthreads();
    int size = block_exscan<QT>(s_tmp, nIni, wsum);
    for (int i = tid; i < nIni; i += QT)
        if (cnt1[i] > 0) {
            box0[s_tmp[i]] = box1[i];
            cnt0[s_tmp[i]] = cnt1[i];
        }
    __syncthreads();
    for (int i = tid; i < n; i += QT) pnode[i] = s_tmp[pnode[i]];
    __syncthreads();

    // The per-step bookkeeping over the NODES (a few dozen to a few hundred) runs on warp 0 alone with ballots and
    // shuffles — a block-wide version of it is a dozen short loops, three block scans and twenty barriers per step, paid
    // by every warp of the tree; the per-KEY passes (a, f) and the quadratic rank sort of the budgeted phase use all threads.
    int cur = 0, phase = 1;
    bool finish = false;
    while (!finish) {                                                     // :625
        const int prevSize = size;
        const short4* boxc = (cur ? box1 : box0);
        const int* cntc = (cur ? cnt1 : cnt0);
        short4* boxn = (cur ? box0 : box1);
        int* cntn = (cur ? cnt0 : cnt1);
        // a. keys per quadrant of every candidate (a node with more than one key)
        for (int i = tid; i < 4 * size; i += QT) s_cnt4[i] = 0;
        __syncthreads();
        for (int i = tid; i < n; i += QT) {
            const int nd = pnode[i];
            if (cntc[nd] > 1) {
                const int q = quadrant(pxy[i], boxc[nd]);
                atomicAdd(&s_cnt4[4 * nd + q], 1);
                pnode[i] = (uint32_t)nd | ((uint32_t)q << 28);       // pass f takes the quadrant from here (node positions < 2^12)
            }
        }
        __syncthreads();
        // b. (warp 0) the candidates compacted in list order
        auto compact = [&]() {
            int k0 = 0;
            for (int base = 0; base < size; base += 32) {
                const int i = base + lane;
                const bool f = i < size && cntc[i] > 1;
                const unsigned bal = __ballot_sync(FULL, f);
                if (f) s_cand[k0 + __popc(bal & ((1u << lane) - 1u))] = i;
                k0 += __popc(bal);
            }
            __syncwarp();
            return k0;
        };
        // d, e. (warp 0) children per processed node in processing order, running prefix, early break (:774); then the
        //       new list: children of the processed nodes in front (last processed first), survivors behind in their old order
        auto node_step = [&](const int* order, int k0) {
            for (int i = lane; i < size; i += 32) s_prank[i] = -1;
            int carry = 0, k = k0, T = 0;
            bool cut = false;
            for (int base = 0; base < k0; base += 32) {
                const int r = base + lane;
                const int c = r < k0 ? child_count(&s_cnt4[4 * order[r]]) : 0;
                int inc = c;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int u = __shfl_up_sync(FULL, inc, o);
                    if (lane >= o) inc += u;
                }
                const int ex = carry + inc - c;
                if (r < k0) s_pex[r] = ex;
                if (phase == 2) {
                    const unsigned bal = __ballot_sync(FULL, r < k0 && size + ex + c - (r + 1) >= N);
                    if (bal) {
                        const int first = __ffs(bal) - 1;
                        k = base + first + 1;
                        T = __shfl_sync(FULL, ex + c, first);
                        cut = true;
                        break;
                    }
                }
                carry += __shfl_sync(FULL, inc, 31);
            }
            if (!cut) T = carry;
            __syncwarp();
            for (int r = lane; r < k; r += 32) s_prank[order[r]] = r;
            __syncwarp();
            int ns = 0, nexp = 0;
            for (int base = 0; base < size; base += 32) {
                const int i = base + lane;
                const int r = i < size ? s_prank[i] : 0;
                const bool surv = i < size && r < 0;
                const unsigned bal = __ballot_sync(FULL, surv);
                if (surv) {
                    const int np = T + ns + __popc(bal & ((1u << lane) - 1u));
                    boxn[np] = boxc[i];
                    cntn[np] = cntc[i];
                    s_remap[i] = np;
                } else if (i < size) {
                    const int* c4 = &s_cnt4[4 * i];
                    const int base2 = T - (s_pex[r] + child_count(c4));
                    s_remap[i] = base2;
                    int j = 0, cw = 0;
#pragma unroll
                    for (int q = 3; q >= 0; q--)
                        if (c4[q] > 0) {
                            boxn[base2 + j] = child_box(boxc[i], q);
                            cntn[base2 + j] = c4[q];
                            nexp += c4[q] > 1;
                            cw |= j << (8 * q);                          // list slot of quadrant q's child behind base2
                            j++;
                        }
                    s_tmp[i] = cw;
                }
                ns += __popc(bal);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) nexp += __shfl_xor_sync(FULL, nexp, o);
            if (lane == 0) { s_T = T; s_ns = ns; s_nexp = nexp; }
        };
        if (phase == 1) {
            // c. processing order = list order
            if (wid == 0) {
                const int k0 = compact();
                node_step(s_cand, k0);
            }
        } else {
            if (wid == 0) {
                const int k0 = compact();
                if (lane == 0) s_k0 = k0;
            }
            __syncthreads();
            const int k0 = s_k0;
            // c. sort (count, creation order) ascending, walk from the back (:728-731): count descending, then newest
            // (= smaller list position) first.  Rank sort on one packed key per candidate, count << 12 | (4095 - list
            // rank): "count larger, or equal and earlier in the list" is one unsigned compare; the keys are read four at a
            // time (broadcast LDS.128)
            uint32_t* key = reinterpret_cast<uint32_t*>(s_tmp);
            for (int ci = tid; ci < ((k0 + 3) & ~3); ci += QT)
                key[ci] = ci < k0 ? ((uint32_t)cntc[s_cand[ci]] << 12) | (uint32_t)(4095 - ci) : 0u;
            __syncthreads();
            const uint4* key4 = reinterpret_cast<const uint4*>(key);
            for (int ci = tid; ci < k0; ci += QT) {
                const uint32_t me = key[ci];
                int r = 0;
                for (int j = 0; j < (k0 + 3) >> 2; j++) {
                    const uint4 kj = key4[j];
                    r += (kj.x > me) + (kj.y > me) + (kj.z > me) + (kj.w > me);
                }
                s_order[r] = s_cand[ci];
            }
            __syncthreads();
            if (wid == 0) node_step(s_order, k0);
        }
        __syncthreads();
        // f. keys follow their nodes
        for (int i = tid; i < n; i += QT) {
            const uint32_t pq = pnode[i];
            const int nd = (int)(pq & 0x0FFFFFFFu);
            // a split node's key moves to the child of its quadrant (recorded by pass a; s_tmp holds the children's slots)
            pnode[i] = s_remap[nd] + (s_prank[nd] >= 0 ? (s_tmp[nd] >> (8 * (pq >> 28))) & 0xFF : 0);
        }
        const int nToExpand = s_nexp;
        size = s_T + s_ns;
        __syncthreads();
        cur ^= 1;
        if (size >= N || size == prevSize) finish = true;                 // :707, :778
        else if (phase == 1 && size + nToExpand * 3 > N) phase = 2;       // :714
    }

    // ---- one keypoint per node: maximum response, first in vKeys order wins ties (:789-805)
    for (int i = tid; i < size; i += QT) s_tmp[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += QT)
        atomicMax((unsigned*)&s_tmp[pnode[i]],
                  ((unsigned)(kp_score(pxy[i]) + 1) << 20) | (0xFFFFFu - (unsigned)i));
    __syncthreads();
    for (int pos = tid; pos < size; pos += QT) {
        const uint32_t p = pxy[0xFFFFFu - ((unsigned)s_tmp[pos] & 0xFFFFFu)];
        sel_out[pos] = pack_kp(kp_x(p) + L.minBX, kp_y(p) + L.minBY, kp_score(p));   // :912-913
    }
    if (tid == 0) *selc_out = size;
}

// threads per tree by the level's feature budget: the per-step cost of a tree is a fixed number of block-wide
// loops, scans and barriers, paid once per warp, so the small trees of the upper levels run on fewer warps
int tree_threads(const LevelGeom& L) {
    // measured at nfeatures 1000 (budgets 217 181 151 126 105 87 73 60) with the node bookkeeping on warp 0: 256 threads
    // above a budget of 100 and 128 below gives 0.248 ms per 1024 frames; 256 / 128 / 64 split at 170 / 90: 0.254, at
    // 140 / 80: 0.249; all 256: 0.264.  (With block-wide bookkeeping the split mattered more: 0.308 against 0.328 for all
    // 256 and 0.410 for all 64.)  The single-frame latency does not depend on it (the pyramid chain is as long).
    static int t_hi = -1, t_lo = -1;
    if (t_hi < 0) {
        const char* e = std::getenv("ORBFE_OCTREE_SPLIT");        // "hi,lo": budgets above hi -> 256, above lo -> 128, else 64
        int a = 100, b = 0;
        if (e && std::sscanf(e, "%d,%d", &a, &b) != 2) { a = 100; b = 0; }
        t_hi = a; t_lo = b;
    }
    return L.nFeat > t_hi ? 256 : L.nFeat > t_lo ? 128 : 64;
}

template <int T>
void launch_tree(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, dim3 grid, size_t smem, int NC, int KP, int a) {
    octree_kernel<T><<<grid, T, smem, s>>>(g, d, NC, KP, a);
}

// shared-memory arena of a launch over levels [l0, l1): node arrays for the largest selection capacity, keys for twelve
// times the largest budget (the synthetic and the usual camera frames bring 4-6 raw corners per kept one; more go
// through the HBM scratch), at most kPtsSmemMax
void group_capacity(const FrameSetGeom& g, int l0, int l1, int* NC, int* KP) {
    int nc = 0, nmax = 0;
    for (int l = l0; l < l1; l++) { nc = max(nc, g.lv[l].selCap); nmax = max(nmax, g.lv[l].nFeat); }
    *NC = (nc + 3) & ~3;           // the int arrays behind the boxes stay 16-byte aligned (uint4 reads of the sort keys)
    *KP = min(kPtsSmemMax, max(256, (12 * nmax + 63) & ~63));
}

}  // namespace

size_t octree_smem_bytes(const FrameSetGeom& g) {
    int NC, KP;
    group_capacity(g, 0, g.nlevels, &NC, &KP);
    return (size_t)NC * 64 + (size_t)KP * 8;
}

// The attribute is per function and per device, not per handle: it is always set to the one worst case
// ensure_geometry admits (200 KB), so a handle with a small feature budget cannot lower the limit under a
// live handle with a large one (mono Tracking keeps ORBextractor(2 * nFeatures) and ORBextractor(nFeatures)).
int octree_prepare(const FrameSetGeom& g) {
    (void)g;
    cudaError_t e = cudaFuncSetAttribute(octree_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(octree_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(octree_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    return (int)e;
}

int launch_octree(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0, int l1) {
    if (l1 < 0 || l1 > g.nlevels) l1 = g.nlevels;
    if (l0 < 0) l0 = 0;
    int launches = 0;
    for (int a = l0; a < l1;) {                          // runs of levels with the same thread count
        const int T = tree_threads(g.lv[a]);
        int b = a + 1;
        while (b < l1 && tree_threads(g.lv[b]) == T) b++;
        int NC, KP;
        group_capacity(g, a, b, &NC, &KP);
        const size_t smem = (size_t)NC * 64 + (size_t)KP * 8;
        const dim3 grid(b - a, g.batch);
        switch (T) {
            case 256: launch_tree<256>(g, d, s, grid, smem, NC, KP, a); break;
            case 128: launch_tree<128>(g, d, s, grid, smem, NC, KP, a); break;
            default: launch_tree<64>(g, d, s, grid, smem, NC, KP, a); break;
        }
        launches++;
        a = b;
    }
    return launches;
}

}  // namespace orbfe
