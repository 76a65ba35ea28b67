// match_bow.cu — K12: ORBmatcher::SearchByBoW, both forms:
//   SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)      reference src/ORBmatcher.cc:185-325
//   SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)   reference src/ORBmatcher.cc:610-743
// A segmented brute-force 2-nearest-neighbour search: the two DBoW2 feature vectors arrive as CSR
// (ascending node ids, feature indices in vector order); features only meet inside a common
// vocabulary node.  One warp per node of side 1: it finds the node on side 2 by binary search
// (the lower_bound walk of the reference visits exactly the common ids) and then replays the
// reference's order inside the node — side-1 features one after the other, because each accepted
// match removes a side-2 feature from the pool (:232 / :662) — while the lanes share the work of
// one feature: every lane scores its own side-2 candidates and keeps a local best / second-best,
// a warp min over the keys (distance << 16 | position) picks the winner with the reference's tie
// rule (strict '<': the first candidate in vector order), the second-best is the smallest among
// the winner lane's runner-up and the other lanes' bests.  A side-2 feature belongs to one node
// only, so warps never compete for it.  The rotation histogram (30 bins, only 0..12 reachable,
// :271-281) is accumulated with atomics; a second small kernel applies ComputeThreeMaxima and
// clears the matches outside the three leading bins (:298-318).
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int BW_WARPS = 4;
constexpr int HISTO = 30;

__device__ __forceinline__ int rot_bin(float a1, float a2) {
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / HISTO));
    if (bin == HISTO) bin = 0;
    return bin;
}

// ComputeThreeMaxima (src/ORBmatcher.cc:1777-1821) over bin counts
__device__ void three_maxima30(const int* hist, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < HISTO; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
}

__device__ __forceinline__ unsigned warp_min_u32(unsigned v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(FULL, v, o));
    return v;
}

// grid = (node groups, pairs): pair p reads its own argument record
__global__ void __launch_bounds__(BW_WARPS * 32) bow_match_kernel(const BowArgs* __restrict__ pairs) {
    const BowArgs& a = pairs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int node1 = blockIdx.x * BW_WARPS + (threadIdx.x >> 5);
    if (node1 >= a.nnodes1 || a.nnodes2 == 0) return;
    // the same node on side 2 (std::map::lower_bound over ascending ids)
    const uint32_t id = a.node_id1[node1];
    int lo = 0, hi = a.nnodes2;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a.node_id2[mid] < id) lo = mid + 1; else hi = mid;
    }
    if (lo >= a.nnodes2 || a.node_id2[lo] != id) return;
    const int s1 = a.node_start1[node1], e1 = a.node_start1[node1 + 1];
    const int s2 = a.node_start2[lo], n2 = a.node_start2[lo + 1] - s2;
    volatile int* m21 = a.m21;
    for (int i1 = s1; i1 < e1; i1++) {
        const int idx1 = (int)a.indices1[i1];
        if (a.valid1 && !a.valid1[idx1]) continue;                                   // :214-220 / :645-649
        uint32_t q[8];
        {
            const uint4* p = reinterpret_cast<const uint4*>(a.desc1 + 32 * (long long)idx1);
            const uint4 u = __ldg(p), v = __ldg(p + 1);
            q[0] = u.x; q[1] = u.y; q[2] = u.z; q[3] = u.w; q[4] = v.x; q[5] = v.y; q[6] = v.z; q[7] = v.w;
        }
        unsigned k1 = 0xFFFFFFFFu;       // local best key: distance << 16 | position in the node
        int d2 = 256;                    // local second-best distance
        for (int pos = lane; pos < n2; pos += 32) {
            const int idx2 = (int)a.indices2[s2 + pos];
            if (m21[idx2] >= 0) continue;                                            // :232 / :662
            if (a.valid2 && !a.valid2[idx2]) continue;                               // :662-670
            const uint4* p = reinterpret_cast<const uint4*>(a.desc2 + 32 * (long long)idx2);
            const uint4 u = __ldg(p), v = __ldg(p + 1);
            const int dist = __popc(q[0] ^ u.x) + __popc(q[1] ^ u.y) + __popc(q[2] ^ u.z) + __popc(q[3] ^ u.w) +
                             __popc(q[4] ^ v.x) + __popc(q[5] ^ v.y) + __popc(q[6] ^ v.z) + __popc(q[7] ^ v.w);
            const unsigned key = ((unsigned)dist << 16) | (unsigned)pos;
            if (key < k1) { if (k1 != 0xFFFFFFFFu) d2 = (int)(k1 >> 16); k1 = key; }
            else if (dist < d2) d2 = dist;
        }
        const unsigned best = warp_min_u32(k1);
        if (best == 0xFFFFFFFFu) continue;                                           // no free candidate: bestDist1 = 256
        const int mine = k1 == best ? d2 : (k1 == 0xFFFFFFFFu ? 256 : (int)(k1 >> 16));
        const int bestDist2 = (int)warp_min_u32((unsigned)mine);
        const int bestDist1 = (int)(best >> 16);
        if (bestDist1 <= a.th && (float)bestDist1 < __fmul_rn(a.ratio, (float)bestDist2)) {   // :263-265 / :686-688
            const int idx2 = (int)a.indices2[s2 + (int)(best & 0xFFFFu)];
            if (lane == 0) {
                a.m12[idx1] = idx2;
                m21[idx2] = idx1;
                if (a.checkOri) atomicAdd(&a.hist[rot_bin(a.angle1[idx1], a.angle2[idx2])], 1);   // :271-281
                atomicAdd(a.nmatches, 1);
            }
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(256) bow_rotation_kernel(const BowArgs* __restrict__ pairs) {
    const BowArgs& a = pairs[blockIdx.y];
    __shared__ int ind[3];
    if (threadIdx.x == 0) three_maxima30(a.hist, ind[0], ind[1], ind[2]);
    __syncthreads();
    const int idx1 = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx1 >= a.n1) return;
    const int idx2 = a.m12[idx1];
    if (idx2 < 0) return;
    const int bin = rot_bin(a.angle1[idx1], a.angle2[idx2]);
    if (bin == ind[0] || bin == ind[1] || bin == ind[2]) return;
    a.m12[idx1] = -1; a.m21[idx2] = -1;                                              // :311-315
    atomicSub(a.nmatches, 1);
}

// every pair's m12 / m21 to -1, histograms and match counts to 0 (two runs of one block: see api_matcher.cu)
__global__ void __launch_bounds__(256) bow_init_kernel(int* minus1, int nminus1, int* zero, int nzero) {
    const int stride = gridDim.x * blockDim.x, t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = t; i < nminus1; i += stride) minus1[i] = -1;
    for (int i = t; i < nzero; i += stride) zero[i] = 0;
}

// ---------------------------------------------------------------------------------------
// K15: ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:754-928) with CheckDistEpipolarLine
// (:156-175).  The reference declares vbMatched2 but never sets it (:774), so the features of key
// frame 1 do not interact: one warp per CSR position of side 1.  The warp finds its node (binary
// search over node_start1), the same node on side 2, and scores that node's side-2 features across
// the lanes.  The sequential rule "skip when dist > bestDist, replace when the gates pass" (:840-860)
// ends on the LAST candidate, in vector order, among those of minimal distance that pass the
// parallax gate (:845-853) and the epipolar gate: key = distance << 16 | (0xFFFF - position).
__global__ void __launch_bounds__(BW_WARPS * 32) tri_match_kernel(const __grid_constant__ TriArgs a) {
    const int lane = threadIdx.x & 31;
    const int p1 = blockIdx.x * BW_WARPS + (threadIdx.x >> 5);
    if (p1 >= a.nidx1 || a.nnodes2 == 0) return;
    int lo = 0, hi = a.nnodes1;                       // last node with node_start1[node] <= p1
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (a.node_start1[mid] <= p1) lo = mid; else hi = mid;
    }
    const uint32_t id = a.node_id1[lo];
    lo = 0; hi = a.nnodes2;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a.node_id2[mid] < id) lo = mid + 1; else hi = mid;
    }
    if (lo >= a.nnodes2 || a.node_id2[lo] != id) return;
    const int idx1 = (int)a.indices1[p1];
    if (a.has1[idx1]) return;                                                            // :799-803
    const bool bStereo1 = a.ur1 && a.ur1[idx1] >= 0;                                     // :805
    if (a.onlyStereo && !bStereo1) return;
    const orbfe_keypoint kp1 = a.kps1[idx1];
    // epipolar line of kp1 in image 2 (:159-161)
    const float la = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[0]), __fmul_rn(kp1.y, a.F[3])), a.F[6]);
    const float lb = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[1]), __fmul_rn(kp1.y, a.F[4])), a.F[7]);
    const float lc = __fadd_rn(__fadd_rn(__fmul_rn(kp1.x, a.F[2]), __fmul_rn(kp1.y, a.F[5])), a.F[8]);
    const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
    uint32_t q[8];
    {
        const uint4* p = reinterpret_cast<const uint4*>(a.desc1 + 32 * (long long)idx1);
        const uint4 u = __ldg(p), v = __ldg(p + 1);
        q[0] = u.x; q[1] = u.y; q[2] = u.z; q[3] = u.w; q[4] = v.x; q[5] = v.y; q[6] = v.z; q[7] = v.w;
    }
    const int s2 = a.node_start2[lo], n2 = a.node_start2[lo + 1] - s2;
    unsigned best = 0xFFFFFFFFu;
    for (int pos = lane; pos < n2; pos += 32) {
        const int idx2 = (int)a.indices2[s2 + pos];
        if (a.has2[idx2]) continue;                                                      // :826-828
        const bool bStereo2 = a.ur2 && a.ur2[idx2] >= 0;
        if (a.onlyStereo && !bStereo2) continue;
        const uint4* p = reinterpret_cast<const uint4*>(a.desc2 + 32 * (long long)idx2);
        const uint4 u = __ldg(p), v = __ldg(p + 1);
        const int dist = __popc(q[0] ^ u.x) + __popc(q[1] ^ u.y) + __popc(q[2] ^ u.z) + __popc(q[3] ^ u.w) +
                         __popc(q[4] ^ v.x) + __popc(q[5] ^ v.y) + __popc(q[6] ^ v.z) + __popc(q[7] ^ v.w);
        if (dist > a.th) continue;                                                       // :840
        const orbfe_keypoint kp2 = a.kps2[idx2];
        if (!bStereo1 && !bStereo2) {                                                    // :845-853
            const float dx = __fsub_rn(a.ex, kp2.x), dy = __fsub_rn(a.ey, kp2.y);
            if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.0f, a.scale2[kp2.octave])) continue;
        }
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, kp2.x), __fmul_rn(lb, kp2.y)), lc);   // :163
        if (den == 0.0f) continue;                                                       // :167
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < 3.84 * (double)a.sigma2_2[kp2.octave])) continue;           // :172
        const unsigned key = ((unsigned)dist << 16) | (unsigned)(0xFFFF - pos);
        best = min(best, key);
    }
    best = warp_min_u32(best);
    if (best == 0xFFFFFFFFu || lane != 0) return;
    const int idx2 = (int)a.indices2[s2 + (0xFFFF - (int)(best & 0xFFFFu))];
    a.m12[idx1] = idx2;                                                                  // :866
    if (a.checkOri) atomicAdd(&a.hist[rot_bin(kp1.angle, a.kps2[idx2].angle)], 1);       // :869-879
    atomicAdd(a.nmatches, 1);
}

__global__ void __launch_bounds__(256) tri_rotation_kernel(const __grid_constant__ TriArgs a) {
    __shared__ int ind[3];
    if (threadIdx.x == 0) three_maxima30(a.hist, ind[0], ind[1], ind[2]);
    __syncthreads();
    const int idx1 = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx1 >= a.n1) return;
    const int idx2 = a.m12[idx1];
    if (idx2 < 0) return;
    const int bin = rot_bin(a.kps1[idx1].angle, a.kps2[idx2].angle);
    if (bin == ind[0] || bin == ind[1] || bin == ind[2]) return;
    a.m12[idx1] = -1;                                                                    // :905-909
    atomicSub(a.nmatches, 1);
}

}  // namespace

int launch_search_bow(const BowArgs* d_pairs, int npairs, int max_nodes1, int max_n1, bool check_ori, int* minus1,
                      int nminus1, int* zero, int nzero, cudaStream_t st) {
    const int work = nminus1 > nzero ? nminus1 : nzero;
    const int ib = (work + 1023) / 1024;
    bow_init_kernel<<<ib > 0 ? (ib < 1024 ? ib : 1024) : 1, 256, 0, st>>>(minus1, nminus1, zero, nzero);
    int launches = 1;
    if (npairs > 0 && max_nodes1 > 0) {
        bow_match_kernel<<<dim3((max_nodes1 + BW_WARPS - 1) / BW_WARPS, npairs), BW_WARPS * 32, 0, st>>>(d_pairs);
        launches++;
        if (check_ori && max_n1 > 0) { bow_rotation_kernel<<<dim3((max_n1 + 255) / 256, npairs), 256, 0, st>>>(d_pairs); launches++; }
    }
    return launches;
}

}  // namespace orbfe

namespace orbfe {

// minus1: a.m12 (n1 ints); zero: histogram + count (contiguous, nzero ints)
int launch_search_triangulation(const TriArgs& a, int* zero, int nzero, cudaStream_t st) {
    const int work = a.n1 > nzero ? a.n1 : nzero;
    const int ib = (work + 1023) / 1024;
    bow_init_kernel<<<ib > 0 ? (ib < 1024 ? ib : 1024) : 1, 256, 0, st>>>(a.m12, a.n1, zero, nzero);
    int launches = 1;
    if (a.nidx1 > 0) {
        tri_match_kernel<<<(a.nidx1 + BW_WARPS - 1) / BW_WARPS, BW_WARPS * 32, 0, st>>>(a);
        launches++;
        if (a.checkOri && a.n1 > 0) { tri_rotation_kernel<<<(a.n1 + 255) / 256, 256, 0, st>>>(a); launches++; }
    }
    return launches;
}

}  // namespace orbfe
