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
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int QT = 256;            // threads per tree
constexpr int kPtsSmemMax = 4096;  // most keys a tree keeps in shared memory; larger levels use HBM scratch
constexpr unsigned FULL = 0xFFFFFFFFu;

struct Smem {
    short4* box[2];
    int* cnt[2];
    int* cnt4;      // 4 per node: keys per quadrant of a candidate
    int* cand;      // compacted candidate positions (list order)
    int* order;     // processing order -> node position
    int* pex;       // exclusive prefix of children counts, in processing order
    int* prank;     // per node: processing rank if split this step, else -1
    int* remap;     // per node: new position (survivor) or first child position (split)
    int* tmp;       // scan scratch
    uint32_t* pxy;
    uint32_t* pnode;
};

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

__global__ void __launch_bounds__(QT)
octree_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ DevBufs d, int NC, int KP, int level0) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int wsum[QT / 32 + 1];
    __shared__ int s_n, s_cut, s_nexp;

    const int l = blockIdx.x + level0, frame = blockIdx.y + g.frame0;
    const LevelGeom& L = g.lv[l];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int N = L.nFeat;

    Smem S;
    {
        unsigned char* p = smem_raw;
        S.box[0] = (short4*)p; p += sizeof(short4) * NC;
        S.box[1] = (short4*)p; p += sizeof(short4) * NC;
        S.cnt[0] = (int*)p; p += 4 * NC;
        S.cnt[1] = (int*)p; p += 4 * NC;
        S.cnt4 = (int*)p; p += 16 * NC;
        S.cand = (int*)p; p += 4 * NC;
        S.order = (int*)p; p += 4 * NC;
        S.pex = (int*)p; p += 4 * NC;
        S.prank = (int*)p; p += 4 * NC;
        S.remap = (int*)p; p += 4 * NC;
        S.tmp = (int*)p; p += 4 * NC;
        S.pxy = (uint32_t*)p; p += 4 * KP;
        S.pnode = (uint32_t*)p;
    }
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
    uint32_t* pxy = S.pxy;
    uint32_t* pnode = S.pnode;
    if (n > KP) {
        const long long off = (long long)frame * g.cellSlots + L.cellSlotOff;
        pxy = d.qt_xy + off;
        pnode = d.qt_node + off;
    }
    {
        int run = 0;
        for (int base = 0; base < ncell; base += NC) {
            const int m = min(NC, ncell - base);
            for (int c = tid; c < m; c += QT) S.tmp[c] = cc[base + c];
            __syncthreads();
            const int tot = block_exscan(S.tmp, m, wsum);
            // one thread per cell: cells hold a handful of corners, and this way a CTA has
            // hundreds of independent loads in flight instead of one short row per warp
            for (int c = tid; c < m; c += QT) {
                const int k = cc[base + c];
                const uint32_t* src = ck + (long long)(base + c) * L.cellCap;
                uint32_t* dst = pxy + run + S.tmp[c];
                for (int i = 0; i < k; i++) dst[i] = __ldg(src + i);
            }
            run += tot;
            __syncthreads();
        }
    }

    // ---- roots (:570-615): nIni boxes of width hX; empty roots are erased
    const int nIni = L.nIni;
    for (int i = tid; i < nIni; i += QT) {
        S.box[1][i] = make_short4((short)(int)__fmul_rn(L.hX, (float)i), 0,
                                  (short)(int)__fmul_rn(L.hX, (float)(i + 1)),
                                  (short)(L.maxBY - L.minBY));
        S.cnt[1][i] = 0;
    }
    __syncthreads();
    for (int i = tid; i < n; i += QT) {
        const int r = min((int)__fdiv_rn((float)kp_x(pxy[i]), L.hX), nIni - 1);   // :598
        pnode[i] = r;
        atomicAdd(&S.cnt[1][r], 1);
    }
    __syncthreads();
    for (int i = tid; i < nIni; i += QT) S.tmp[i] = S.cnt[1][i] > 0;
    __syncthreads();
    int size = block_exscan(S.tmp, nIni, wsum);
    for (int i = tid; i < nIni; i += QT)
        if (S.cnt[1][i] > 0) {
            S.box[0][S.tmp[i]] = S.box[1][i];
            S.cnt[0][S.tmp[i]] = S.cnt[1][i];
        }
    __syncthreads();
    for (int i = tid; i < n; i += QT) pnode[i] = S.tmp[pnode[i]];
    __syncthreads();

    int cur = 0, phase = 1;
    bool finish = false;
    while (!finish) {                                                     // :625
        const int prevSize = size;
        const short4* boxc = S.box[cur];
        const int* cntc = S.cnt[cur];
        short4* boxn = S.box[cur ^ 1];
        int* cntn = S.cnt[cur ^ 1];
        // a. keys per quadrant of every candidate (a node with more than one key)
        for (int i = tid; i < 4 * size; i += QT) S.cnt4[i] = 0;
        for (int i = tid; i < size; i += QT) { S.prank[i] = -1; S.tmp[i] = cntc[i] > 1; }
        if (tid == 0) { s_cut = 0x7FFFFFFF; s_nexp = 0; }
        __syncthreads();
        for (int i = tid; i < n; i += QT) {
            const int nd = pnode[i];
            if (cntc[nd] > 1) atomicAdd(&S.cnt4[4 * nd + quadrant(pxy[i], boxc[nd])], 1);
        }
        // b. compact the candidates in list order
        const int k0 = block_exscan(S.tmp, size, wsum);
        for (int i = tid; i < size; i += QT)
            if (cntc[i] > 1) S.cand[S.tmp[i]] = i;
        __syncthreads();
        // c. processing order
        if (phase == 1) {
            for (int ci = tid; ci < k0; ci += QT) S.order[ci] = S.cand[ci];
        } else {
            // sort (count, creation order) ascending, walk from the back (:728-731): count
            // descending, then newest (= smaller list position) first.
            for (int ci = tid; ci < k0; ci += QT) {
                const int me = cntc[S.cand[ci]];
                int r = 0;
                for (int j = 0; j < k0; j++) {
                    const int cj = cntc[S.cand[j]];
                    r += (cj > me) || (cj == me && j < ci);
                }
                S.order[r] = S.cand[ci];
            }
        }
        __syncthreads();
        // d. children per processed node, prefix sums, early break (:774)
        for (int r = tid; r < k0; r += QT) S.pex[r] = child_count(&S.cnt4[4 * S.order[r]]);
        __syncthreads();
        block_exscan(S.pex, k0, wsum);
        int k = k0;
        if (phase == 2) {
            for (int r = tid; r < k0; r += QT) {
                const int c = child_count(&S.cnt4[4 * S.order[r]]);
                if (size + S.pex[r] + c - (r + 1) >= N) atomicMin(&s_cut, r);
            }
            __syncthreads();
            if (s_cut < k0) k = s_cut + 1;
        }
        const int T = k > 0 ? S.pex[k - 1] + child_count(&S.cnt4[4 * S.order[k - 1]]) : 0;
        for (int r = tid; r < k; r += QT) S.prank[S.order[r]] = r;
        __syncthreads();
        // e. survivors keep their relative order behind the new children
        for (int i = tid; i < size; i += QT) S.tmp[i] = S.prank[i] < 0;
        __syncthreads();
        const int ns = block_exscan(S.tmp, size, wsum);
        int nexp = 0;
        for (int i = tid; i < size; i += QT) {
            const int r = S.prank[i];
            if (r < 0) {
                const int np = T + S.tmp[i];
                boxn[np] = boxc[i];
                cntn[np] = cntc[i];
                S.remap[i] = np;
            } else {
                const int* c4 = &S.cnt4[4 * i];
                const int base = T - (S.pex[r] + child_count(c4));
                S.remap[i] = base;
                int j = 0;
#pragma unroll
                for (int q = 3; q >= 0; q--)
                    if (c4[q] > 0) {
                        boxn[base + j] = child_box(boxc[i], q);
                        cntn[base + j] = c4[q];
                        nexp += c4[q] > 1;
                        j++;
                    }
            }
        }
        if (nexp) atomicAdd(&s_nexp, nexp);
        __syncthreads();
        // f. keys follow their nodes
        for (int i = tid; i < n; i += QT) {
            const int nd = pnode[i];
            if (S.prank[nd] >= 0) {
                const int q = quadrant(pxy[i], boxc[nd]);
                const int* c4 = &S.cnt4[4 * nd];
                int j = 0;
                for (int qq = q + 1; qq < 4; qq++) j += c4[qq] > 0;
                pnode[i] = S.remap[nd] + j;
            } else {
                pnode[i] = S.remap[nd];
            }
        }
        const int nToExpand = s_nexp;
        __syncthreads();
        size = T + ns;
        cur ^= 1;
        if (size >= N || size == prevSize) finish = true;                 // :707, :778
        else if (phase == 1 && size + nToExpand * 3 > N) phase = 2;       // :714
    }

    // ---- one keypoint per node: maximum response, first in vKeys order wins ties (:789-805)
    for (int i = tid; i < size; i += QT) S.tmp[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += QT)
        atomicMax((unsigned*)&S.tmp[pnode[i]],
                  ((unsigned)(kp_score(pxy[i]) + 1) << 20) | (0xFFFFFu - (unsigned)i));
    __syncthreads();
    for (int pos = tid; pos < size; pos += QT) {
        const uint32_t p = pxy[0xFFFFFu - ((unsigned)S.tmp[pos] & 0xFFFFFu)];
        sel_out[pos] = pack_kp(kp_x(p) + L.minBX, kp_y(p) + L.minBY, kp_score(p));   // :912-913
    }
    if (tid == 0) *selc_out = size;
}

int node_capacity(const FrameSetGeom& g) {
    int nc = 0;
    for (int l = 0; l < g.nlevels; l++) nc = max(nc, g.lv[l].selCap);
    return nc;
}

}  // namespace

// keys staged in shared memory per tree: twelve times the largest per-level budget (the synthetic and the usual
// camera frames bring 4-6 raw corners per kept one), at most kPtsSmemMax — sized so that six trees share an SM
// at nfeatures 1000 instead of four with the fixed 4096 (the kernel is latency / barrier bound: occupancy pays)
static int key_capacity(const FrameSetGeom& g) {
    int nmax = 0;
    for (int l = 0; l < g.nlevels; l++) nmax = max(nmax, g.lv[l].nFeat);
    return min(kPtsSmemMax, max(1024, (12 * nmax + 63) & ~63));
}

size_t octree_smem_bytes(const FrameSetGeom& g) {
    return (size_t)node_capacity(g) * 64 + (size_t)key_capacity(g) * 8;
}

// The attribute is per function and per device, not per handle: it is always set to the one worst case
// ensure_geometry admits (200 KB), so a handle with a small feature budget cannot lower the limit under a
// live handle with a large one (mono Tracking keeps ORBextractor(2 * nFeatures) and ORBextractor(nFeatures)).
int octree_prepare(const FrameSetGeom& g) {
    (void)g;
    return (int)cudaFuncSetAttribute(octree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
}

int launch_octree(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0, int l1) {
    if (l1 < 0 || l1 > g.nlevels) l1 = g.nlevels;
    if (l0 < 0) l0 = 0;
    if (l1 <= l0) return 0;
    octree_kernel<<<dim3(l1 - l0, g.batch), QT, octree_smem_bytes(g), s>>>(g, d, node_capacity(g), key_capacity(g), l0);
    return 1;
}

}  // namespace orbfe
