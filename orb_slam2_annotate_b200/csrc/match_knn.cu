// match_knn.cu — K7: brute-force Hamming 2-nearest-neighbour search with the reference's
// acceptance rule (inner loops of ORBmatcher::SearchByBoW, src/ORBmatcher.cc:233-269;
// distance = ORBmatcher::DescriptorDistance, :1828-1844).
//
// popc formulation: a thread keeps Q query descriptors (8 x 32-bit words each) in registers;
// the CTA streams the train set through shared memory in tiles, every lane reading the same
// train row (a broadcast LDS.128 pair), so one train row costs 8 XOR + 8 POPC + 7 adds per
// query.  Best and second best are tracked as packed keys  dist << 20 | train_index :
//     k1' = min(k1, key)          k2' = min(k2, max(k1, key))
// which reproduces the reference's strict '<' updates exactly: the lowest train index wins a
// tie for the best, and the second-best distance equals the best when two rows tie.  Train
// sets longer than 2^20 rows are processed in segments and merged with the same rule.
#include "orbfe_internal.cuh"
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr int KT = 256;          // threads per CTA
constexpr int TILE = 512;        // train rows per shared-memory tile (16 KB)
constexpr uint32_t KEY_INIT = (256u << 20) | 0xFFFFFu;

template <int Q>
__global__ void __launch_bounds__(KT)
knn2_popc_kernel(const uint4* __restrict__ query, int nq, const uint4* __restrict__ train, int nt,
                 int th, float ratio, int* __restrict__ out_idx, int* __restrict__ out_d1,
                 int* __restrict__ out_d2) {
    __shared__ uint4 tile[TILE * 2];
    const int tid = threadIdx.x;
    const int q0 = (blockIdx.x * KT + tid) * Q;
    uint32_t qw[Q][8];
#pragma unroll
    for (int q = 0; q < Q; q++) {
        const int qi = min(q0 + q, nq - 1);
        const uint4 a = __ldg(&query[2 * (long long)qi]), b = __ldg(&query[2 * (long long)qi + 1]);
        qw[q][0] = a.x; qw[q][1] = a.y; qw[q][2] = a.z; qw[q][3] = a.w;
        qw[q][4] = b.x; qw[q][5] = b.y; qw[q][6] = b.z; qw[q][7] = b.w;
    }
    int D1[Q], D2[Q], I1[Q];
#pragma unroll
    for (int q = 0; q < Q; q++) { D1[q] = 256; D2[q] = 256; I1[q] = -1; }

    for (int seg = 0; seg < nt; seg += (1 << 20)) {
        const int segn = min(nt - seg, 1 << 20);
        uint32_t k1[Q], k2[Q];
#pragma unroll
        for (int q = 0; q < Q; q++) { k1[q] = KEY_INIT; k2[q] = KEY_INIT; }
        for (int base = 0; base < segn; base += TILE) {
            const int m = min(TILE, segn - base);
            __syncthreads();
            for (int i = tid; i < 2 * m; i += KT) tile[i] = __ldg(&train[2 * (long long)(seg + base) + i]);
            __syncthreads();
#pragma unroll 2
            for (int j = 0; j < m; j++) {
                const uint4 a = tile[2 * j], b = tile[2 * j + 1];
#pragma unroll
                for (int q = 0; q < Q; q++) {
                    const int d = __popc(qw[q][0] ^ a.x) + __popc(qw[q][1] ^ a.y) + __popc(qw[q][2] ^ a.z) +
                                  __popc(qw[q][3] ^ a.w) + __popc(qw[q][4] ^ b.x) + __popc(qw[q][5] ^ b.y) +
                                  __popc(qw[q][6] ^ b.z) + __popc(qw[q][7] ^ b.w);
                    const uint32_t key = ((uint32_t)d << 20) | (uint32_t)(base + j);
                    k2[q] = min(k2[q], max(k1[q], key));
                    k1[q] = min(k1[q], key);
                }
            }
        }
        // merge the segment (segments are visited in ascending train order)
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const int s1 = (int)(k1[q] >> 20), s2 = (int)(k2[q] >> 20);
            if (s1 < D1[q]) {
                D2[q] = min(D1[q], s2);
                D1[q] = s1;
                I1[q] = seg + (int)(k1[q] & 0xFFFFFu);
            } else {
                D2[q] = min(D2[q], s1);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < Q; q++) {
        const int qi = q0 + q;
        if (qi < nq) {
            out_d1[qi] = D1[q];
            out_d2[qi] = D2[q];
            const bool ok = D1[q] <= th && (float)D1[q] < __fmul_rn(ratio, (float)D2[q]);   // :263-265
            out_idx[qi] = ok ? I1[q] : -1;
        }
    }
}

}  // namespace

int launch_knn2(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                int* d_idx, int* d_d1, int* d_d2, cudaStream_t s) {
    if (nq <= 0) return 0;
    const uint4* q = reinterpret_cast<const uint4*>(d_query);
    const uint4* t = reinterpret_cast<const uint4*>(d_train);
    // enough CTAs to fill 148 SMs decides how many queries a thread owns
    if (nq >= 148 * KT * 4 * 2) {
        knn2_popc_kernel<4><<<(nq + KT * 4 - 1) / (KT * 4), KT, 0, s>>>(q, nq, t, nt, th, ratio, d_idx, d_d1, d_d2);
    } else if (nq >= 148 * KT * 2) {
        knn2_popc_kernel<2><<<(nq + KT * 2 - 1) / (KT * 2), KT, 0, s>>>(q, nq, t, nt, th, ratio, d_idx, d_d1, d_d2);
    } else {
        knn2_popc_kernel<1><<<(nq + KT - 1) / KT, KT, 0, s>>>(q, nq, t, nt, th, ratio, d_idx, d_d1, d_d2);
    }
    return 1;
}

}  // namespace orbfe
