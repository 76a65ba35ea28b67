// match_knn_mma.cu — K7b: brute-force Hamming 2-nearest-neighbour search on the 5th-generation
// tensor cores (tcgen05.mma kind::i8, accumulators in TMEM), same contract as the popc kernel of
// match_knn.cu (inner loops of ORBmatcher::SearchByBoW, src/ORBmatcher.cc:233-269; distance =
// ORBmatcher::DescriptorDistance, :1828-1844).
//
// A Hamming distance over 256 bits is a contraction: with the bits mapped to signed bytes, 0 -> +1 and 1 -> -1,
//     <a', b'> = 256 - 2 d(a, b)
// so the tensor core returns the distance itself (no |a|, |b| correction terms) and the nearest row is the largest
// accumulator.  HBM carries the 32-byte rows; the bit -> byte expansion happens on the way into shared memory, so
// the tensor core sees s8 operands (K = 256) in the canonical K-major SWIZZLE_128B layout:
//
//   CTA = 256 query rows (two M = 128 tiles: rows 0-127 and 128-255, expanded once) against one
//   segment of the train set, streamed in tiles of 128 rows through a 3-stage shared-memory ring.
//   warps 0-7   epilogue: warp w owns TMEM lanes 32 (w & 3) .. +31 of M tile (w >> 2), one query row per
//               thread; tcgen05.ld brings 32 accumulator columns at a time (the next load in flight while the
//               current one is folded).  The common case is a max tree over the 32 accumulators and ONE compare
//               against thr = 256 - 2 * (current second-best distance): columns arrive in ascending train index,
//               so a later row changes (best, second best) iff its distance is STRICTLY below the second best —
//               the screen is exact.  Only then the packed keys  d << 20 | train index  are formed and the
//               strict-'<' rule  k2' = min(k2, max(k1, key)), k1' = min(k1, key)  runs on the 32 columns
//               (about 2 ln n times per query over n random rows).
//   warps 8-11  producers: one train row per thread and tile: two 16-byte loads, 64 expanded words written as
//               sixteen conflict-free 16-byte stores at the swizzled chunk positions, fence.proxy.async,
//               mbarrier arrive.
//   warp 12     TMEM allocation (all 512 columns: two accumulator stages x two M tiles x 128 columns) and
//               the MMA issue: per tile 2 x 8 tcgen05.mma (M 128, N 128, K 32) from one elected lane,
//               tcgen05.commit onto the stage's "empty" barrier and the accumulator's "full" barrier.
//
// Train sets are cut into segments (grid.y) of at most 2^20 rows — the key's index field — which also
// gives small query sets enough CTAs; a merge kernel folds the per-segment (d1, d2, idx) in ascending
// train order with the popc kernel's rule and applies TH_LOW and the ratio test.
#include <climits>
#include <cstdio>
#include <mutex>

#include "api_common.cuh"
#include "orbfe_internal.cuh"
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr int MQ = 256;                 // query rows per CTA (two M = 128 tiles)
constexpr int NT = 128;                 // train rows per tile
constexpr int STAGES = 3;               // shared-memory ring of expanded train tiles
constexpr int EPI_WARPS = 8, PROD_WARPS = 4;
constexpr int THREADS = (EPI_WARPS + PROD_WARPS + 1) * 32;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;
constexpr int KBLOCK_BYTES_A = MQ * 128;   // one 128-byte K block of the query tile
constexpr int KBLOCK_BYTES_B = NT * 128;   // one 128-byte K block of a train tile

struct alignas(1024) MmaSmem {
    uint8_t a[2][KBLOCK_BYTES_A];                  // [K block][row][128 B], 8-row groups swizzled (64 KB)
    uint8_t b[STAGES][2][KBLOCK_BYTES_B];          // 3 x 32 KB
    unsigned long long full_b[STAGES], empty_b[STAGES], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done = 0;
    while (!done)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
}

// K-major SWIZZLE_128B operand descriptor (rows of 128 bytes, 8-row groups 1024 bytes apart)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// instruction descriptor: s8 x s8 -> s32, A and B K-major, M = 128, N = NT
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(NT >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// four descriptor bits -> four signed bytes, bit i to byte i: 0 -> +1 (0x01), 1 -> -1 (0xFF).  The multiply puts
// bit i at bit 7 of byte i (the shifted copies of the nibble do not overlap), PRMT's sign-replicate mode turns every
// byte into 0x00 / 0xFF from that bit alone, and the OR leaves 0x01 / 0xFF.
__device__ __forceinline__ uint32_t expand4(uint32_t nibble) {
    uint32_t r;
    asm("prmt.b32 %0, %1, 0, 0xBA98;" : "=r"(r) : "r"(nibble * 0x10204080u));
    return r | 0x01010101u;
}

// one 256-bit row -> 256 bytes at row r of a [2][rows][128] swizzled operand block; blockBytes = rows * 128
__device__ __forceinline__ void expand_row(uint8_t* base, int blockBytes, int r, const uint32_t w[8]) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint8_t* blk = base + (i >> 2) * blockBytes + r * 128;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const uint32_t hw = w[i] >> (16 * h);
            uint4 o;
            o.x = expand4(hw & 15u); o.y = expand4((hw >> 4) & 15u);
            o.z = expand4((hw >> 8) & 15u); o.w = expand4((hw >> 12) & 15u);
            const int chunk = ((2 * i + h) & 7) ^ (r & 7);
            *reinterpret_cast<uint4*>(blk + chunk * 16) = o;
        }
    }
}

__device__ __forceinline__ void load_row(const uint4* __restrict__ rows, long long r, bool valid, uint32_t w[8]) {
    uint4 x = make_uint4(0, 0, 0, 0), y = x;
    if (valid) { x = __ldg(&rows[2 * r]); y = __ldg(&rows[2 * r + 1]); }
    w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w; w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
}
#define TMEM_LD32(taddr, v)                                                                                             \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                              \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),      \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),          \
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),         \
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                      \
                 : "r"(taddr) : "memory")

// 32 accumulators (256 - 2 d) of one query against train rows idx0 .. idx0 + 31 of the segment, folded into the
// packed keys (k1, k2) = (best, second best); thr = 256 - 2 * (k2 >> 20).  `valid` < 32 only in a segment's last tile.
__device__ __forceinline__ void fold32(const uint32_t (&v)[32], uint32_t idx0, int valid, uint32_t& k1, uint32_t& k2, int& thr) {
    int m01 = max((int)v[0], (int)v[1]);
#pragma unroll
    for (int i = 2; i < 32; i += 2) m01 = max(m01, max((int)v[i], (int)v[i + 1]));
    if (m01 > thr) {
        constexpr uint32_t NEG = 0u - (1u << 19);                 // (256 - acc) << 19 == d << 20
        const uint32_t c0 = (256u << 19) + idx0;
#pragma unroll
        for (int i = 0; i < 32; i++) {
            const uint32_t key = i < valid ? v[i] * NEG + (c0 + (uint32_t)i) : KEY_NONE;
            k2 = min(k2, max(k1, key));
            k1 = min(k1, key);
        }
        thr = 256 - 2 * (int)(k2 >> 20);
    }
}

__global__ void __launch_bounds__(THREADS, 1)
knn2_mma_kernel(const uint4* __restrict__ query, int nq, const uint4* __restrict__ train, int nt, int segLen,
                int* __restrict__ part_idx, int* __restrict__ part_d1, int* __restrict__ part_d2) {
    // SWIZZLE_128B wants 1024-byte aligned operand blocks; the dynamic shared window of a CTA without static
    // shared memory starts that aligned (checked: a misplaced block traps rather than computes garbage)
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    MmaSmem& S = *reinterpret_cast<MmaSmem*>(smem_raw);
    if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q0 = blockIdx.x * MQ;
    const int seg = blockIdx.y;
    const long long seg0 = (long long)seg * segLen;
    const int segn = (int)min((long long)segLen, (long long)nt - seg0);        // > 0 by construction of the grid
    const int ntiles = (segn + NT - 1) / NT;

    if (tid == 0) {
        for (int i = 0; i < STAGES; i++) { mbar_init(&S.full_b[i], PROD_WARPS); mbar_init(&S.empty_b[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == EPI_WARPS + PROD_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&S.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the query tile, expanded once: thread i < 256 owns row i
    if (tid < MQ) {
        uint32_t w[8];
        load_row(query, q0 + tid, q0 + tid < nq, w);
        expand_row(&S.a[0][0], KBLOCK_BYTES_A, tid, w);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = S.tmem_base;

    if (warp < EPI_WARPS) {
        // ---------------- epilogue: thread <-> query row q0 + tid (TMEM lane tid & 127 of M tile tid >> 7)
        const int mtile = warp >> 2;
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t k1 = KEY_NONE, k2 = KEY_NONE;
        int thr = 256 - 2 * (int)(KEY_NONE >> 20);                 // below every accumulator: the first groups take the exact path
        for (int t = 0; t < ntiles; t++) {
            const int a = t & 1;
            mbar_wait(&S.acc_full[a], (t >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tbase = tmem + lane_base + (uint32_t)(a * 256 + mtile * 128);
            const uint32_t idx0 = (uint32_t)t * NT;
            const int left = segn - (int)idx0;                       // valid columns of this tile (>= NT except in the last one)
            if (left < NT) thr = INT_MIN;                            // padding columns must never win: exact path with masking
            uint32_t v0[32], v1[32];
            TMEM_LD32(tbase, v0);
#pragma unroll
            for (int c = 0; c < NT; c += 64) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                TMEM_LD32(tbase + c + 32, v1);
                fold32(v0, idx0 + c, left - c, k1, k2, thr);
                if (left < NT) thr = INT_MIN;
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c + 64 < NT) TMEM_LD32(tbase + c + 64, v0);
                fold32(v1, idx0 + c + 32, left - c - 32, k1, k2, thr);
                if (left < NT) thr = INT_MIN;
            }
            // this warp's accumulator reads of stage a are complete (wait::ld above)
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.acc_empty[a]);
        }
        const int q = q0 + tid;
        if (q < nq) {
            const long long o = (long long)seg * nq + q;
            const bool h1 = k1 != KEY_NONE, h2 = k2 != KEY_NONE;
            part_d1[o] = h1 ? (int)(k1 >> 20) : 256;
            part_d2[o] = h2 ? (int)(k2 >> 20) : 256;
            part_idx[o] = h1 ? (int)(seg0 + (k1 & 0xFFFFFu)) : -1;
        }
    } else if (warp < EPI_WARPS + PROD_WARPS) {
        // ---------------- producers: thread <-> train row p of every tile
        const int p = tid - EPI_WARPS * 32;
        uint32_t w[8];
        load_row(train, seg0 + p, p < segn, w);
        for (int t = 0; t < ntiles; t++) {
            const int s = t % STAGES;
            const int r = t * NT + p;
            uint32_t wn[8];
            load_row(train, seg0 + r + NT, r + NT < segn, wn);          // next tile's row, in flight during the expansion
            mbar_wait(&S.empty_b[s], ((t / STAGES) & 1) ^ 1);
            expand_row(&S.b[s][0][0], KBLOCK_BYTES_B, p, w);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.full_b[s]);
#pragma unroll
            for (int i = 0; i < 8; i++) w[i] = wn[i];
        }
    } else {
        // ---------------- MMA issue: one elected lane
        if (lane == 0) {
            const uint32_t aaddr = smem_u32(&S.a[0][0]);
            for (int t = 0; t < ntiles; t++) {
                const int s = t % STAGES, a = t & 1;
                mbar_wait(&S.full_b[s], (t / STAGES) & 1);
                mbar_wait(&S.acc_empty[a], ((t >> 1) & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t baddr = smem_u32(&S.b[s][0][0]);
#pragma unroll
                for (int m = 0; m < 2; m++) {
                    const uint32_t d = tmem + (uint32_t)(a * 256 + m * 128);
#pragma unroll
                    for (int kb = 0; kb < 2; kb++) {
                        const uint64_t ad = umma_desc(aaddr + kb * KBLOCK_BYTES_A + m * 128 * 128);
                        const uint64_t bd = umma_desc(baddr + kb * KBLOCK_BYTES_B);
#pragma unroll
                        for (int k = 0; k < 4; k++) umma_i8(d, ad + 2 * k, bd + 2 * k, (kb | k) != 0);   // +32 bytes of K
                    }
                }
                umma_commit(&S.empty_b[s]);
                umma_commit(&S.acc_full[a]);
            }
        }
        __syncwarp();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == EPI_WARPS + PROD_WARPS)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// per-segment partial results -> (idx, d1, d2): segments in ascending train order, the popc kernel's merge rule
// (a later segment's best replaces the running best only when strictly smaller), then TH_LOW / ratio (:263-265)
__global__ void knn2_merge_kernel(int nq, int nseg, int th, float ratio, const int* __restrict__ part_idx,
                                  const int* __restrict__ part_d1, const int* __restrict__ part_d2,
                                  int* __restrict__ out_idx, int* __restrict__ out_d1, int* __restrict__ out_d2) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    int D1 = 256, D2 = 256, I1 = -1;
    for (int s = 0; s < nseg; s++) {
        const long long o = (long long)s * nq + q;
        const int s1 = part_d1[o], s2 = part_d2[o];
        if (s1 < D1) { D2 = min(D1, s2); D1 = s1; I1 = part_idx[o]; }
        else D2 = min(D2, s1);
    }
    out_d1[q] = D1; out_d2[q] = D2;
    const bool ok = D1 <= th && (float)D1 < __fmul_rn(ratio, (float)D2);
    out_idx[q] = ok ? I1 : -1;
}

}  // namespace

// segments of the train set for the tensor-core engine: at most 2^20 rows each, enough of them that small query
// sets still fill the SMs, and — one CTA per SM, every CTA the same work — a count that leaves the last wave of
// (query tiles x segments) CTAs as full as possible: 489 query tiles (1M queries over eight GPUs) are 3.3 waves in
// one segment (82 % of four) and 9.9 waves in three (99 % of ten)
int knn2_tile_segments(int nq, int nt, int mq, int ntile, int* segLen, int minWant) {
    constexpr int kSms = 148;
    const int qtiles = (nq + mq - 1) / mq;
    int want = qtiles >= kSms ? 1 : (kSms + qtiles - 1) / qtiles;
    const int maxSeg = nt / (16 * ntile) > 1 ? nt / (16 * ntile) : 1;  // no segment shorter than 16 tiles
    if (want > maxSeg) want = maxSeg;
    if (want < minWant) want = minWant < maxSeg ? minWant : maxSeg;
    const int minSeg = (nt + (1 << 20) - 1) >> 20;
    if (want < minSeg) want = minSeg;
    int best = want;
    double bestEff = 0.0;
    for (int c = want; c <= want + 7 && c <= (maxSeg > want ? maxSeg : want); c++) {
        const long long ctas = (long long)qtiles * c, waves = (ctas + kSms - 1) / kSms;
        const double eff = (double)ctas / (double)(waves * kSms);
        if (eff > bestEff + 0.01) { bestEff = eff; best = c; }          // more segments only for a fuller tail
    }
    int len = ((nt + best - 1) / best + ntile - 1) / ntile * ntile;
    if (len > (1 << 20)) len = (1 << 20) / ntile * ntile;
    *segLen = len;
    return (nt + len - 1) / len;
}
int knn2_mma_segments(int nq, int nt, int* segLen) { return knn2_tile_segments(nq, nt, MQ, NT, segLen, 1); }

int launch_knn2_merge(int nq, int nseg, int th, float ratio, const int* part_idx, const int* part_d1, const int* part_d2,
                      int* d_idx, int* d_d1, int* d_d2, cudaStream_t s) {
    knn2_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(nq, nseg, th, ratio, part_idx, part_d1, part_d2, d_idx, d_d1, d_d2);
    return 1;
}

size_t knn2_mma_scratch_bytes(int nq, int nt) {
    int len;
    const int nseg = knn2_mma_segments(nq, nt, &len);
    return (size_t)3 * nseg * nq * sizeof(int);
}

int launch_knn2_mma(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                    int* d_idx, int* d_d1, int* d_d2, void* scratch, cudaStream_t s) {
    if (nq <= 0) return 0;
    if (nt <= 0) {            // nothing to compare with: the merge kernel writes (−1, 256, 256)
        knn2_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(nq, 0, th, ratio, nullptr, nullptr, nullptr, d_idx, d_d1, d_d2);
        return 1;
    }
    int segLen;
    const int nseg = knn2_mma_segments(nq, nt, &segLen);
    int* part_idx = (int*)scratch;
    int* part_d1 = part_idx + (size_t)nseg * nq;
    int* part_d2 = part_d1 + (size_t)nseg * nq;
    constexpr int smem = (int)sizeof(MmaSmem);
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &cap);
    if (cap == cudaStreamCaptureStatusNone)
        cudaFuncSetAttribute(knn2_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    knn2_mma_kernel<<<dim3((nq + MQ - 1) / MQ, nseg), THREADS, smem, s>>>(
        reinterpret_cast<const uint4*>(d_query), nq, reinterpret_cast<const uint4*>(d_train), nt, segLen, part_idx, part_d1, part_d2);
    knn2_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(nq, nseg, th, ratio, part_idx, part_d1, part_d2, d_idx, d_d1, d_d2);
    return 2;
}

}  // namespace orbfe

// ---------------------------------------------------------------------------------------------------------------
// orbfe_knn2_device_noalloc — the tensor-core kNN-2 for callers that must not allocate (stream capture, tight
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-segment partial results, the one piece of scratch this engine needs, live in a
// file-scope device array instead of the matcher handle's grow-only buffer: every GPU that loads liborbfe.so
// holds these 8 MiB until the process exits.  All calls on a device share the array and it keeps the last call's
// contents; nothing in it has to start at zero (the kernel writes every (segment, query) slot the merge reads), but
// two calls that overlap on one GPU race on it — serialise them on one stream or with events.
static __device__ __align__(256) unsigned char g_knn2_noalloc_scratch[ORBFE_KNN2_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNoallocMaxDevices = 64;
// device address of the array per device: the name above is a host-side stand-in in host code, so the address a
// kernel may use comes from cudaGetSymbolAddress, asked once per device with that device current
void* g_noalloc_ptr[kNoallocMaxDevices] = {};
std::mutex g_noalloc_mu;
}  // namespace

extern "C" int orbfe_knn2_device_noalloc(int device, void* cuda_stream, const uint8_t* d_query, int nq,
                                         const uint8_t* d_train, int nt, int th, float ratio,
                                         int32_t* d_out_idx, int32_t* d_out_d1, int32_t* d_out_d2) {
    using namespace orbfe;
    if (device < 0 || device >= kNoallocMaxDevices || nq < 0 || nt < 0 ||
        (nq && (!d_query || !d_out_idx || !d_out_d1 || !d_out_d2)) || (nt && !d_train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2_device_noalloc");
    if (((uintptr_t)d_query | (uintptr_t)d_train) & 15) return fail(ORBFE_ERR_ARG, "descriptor arrays must be 16-byte aligned");
    if (nq == 0) return ORBFE_OK;
    const size_t need = knn2_mma_scratch_bytes(nq, nt > 0 ? nt : 1);
    if (need > (size_t)ORBFE_KNN2_NOALLOC_SCRATCH_BYTES) {
        static thread_local char msg[256];
        snprintf(msg, sizeof(msg), "orbfe_knn2_device_noalloc: %d queries x %d train rows need %zu bytes of scratch, the static "
                 "array holds %d (12 bytes x queries x train segments): split the queries or use orbfe_knn2_device",
                 nq, nt, need, (int)ORBFE_KNN2_NOALLOC_SCRATCH_BYTES);
        return fail(ORBFE_ERR_CAPACITY, msg);
    }
    CU(cudaSetDevice(device));
    void* scratch;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_ptr[device]) {
            void* p = nullptr;
            const cudaError_t ce = cudaGetSymbolAddress(&p, g_knn2_noalloc_scratch);
            if (ce != cudaSuccess || !p) return fail(ORBFE_ERR_CUDA, "cudaGetSymbolAddress(g_knn2_noalloc_scratch)", ce);
            g_noalloc_ptr[device] = p;
        }
        scratch = g_noalloc_ptr[device];
    }
    launch_knn2_mma(d_query, nq, d_train, nt, th, ratio, d_out_idx, d_out_d1, d_out_d2, scratch, (cudaStream_t)cuda_stream);
    CU(cudaGetLastError());
    return ORBFE_OK;
}
