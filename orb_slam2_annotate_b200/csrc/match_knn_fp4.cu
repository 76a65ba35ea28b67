// match_knn_fp4.cu — K7c: the brute-force Hamming 2-nearest-neighbour search of match_knn_mma.cu on the block-scaled
// FP4 tensor path (tcgen05.mma kind::mxf4.block_scale, twice the s8 rate: tools/mma_rate_probe.cu measures 16.4 T
// 256-bit comparisons/s raw against 7.8 T for kind::i8 on this pool's B200).  Same contract as the popc kernel
// (inner loops of ORBmatcher::SearchByBoW, src/ORBmatcher.cc:233-269; distance = ORBmatcher::DescriptorDistance,
// :1828-1844) and bit-identical results.
//
// Bits become E2M1 values, 0 -> +1.0 (0x2), 1 -> -1.0 (0xA), every UE8M0 block scale is 2^0 (0x7F), so
//     <a', b'> = 256 - 2 d(a, b)
// exactly (sums of +-1 in fp32).  A 256-bit row is 128 bytes of FP4 — one SWIZZLE_128B row, the whole contraction in
// four K = 64 MMAs.  Unlike the s8 engine nothing is expanded inside the hot kernel: a pre-pass writes both row sets
// once in FP4 form to scratch (4 x the packed size), and the hot kernel's operands arrive by TMA (cp.async.bulk.tensor
// with the 128-byte swizzle, i.e. directly in the canonical K-major layout the MMA descriptors name).  All CTAs walk the
// train set in step, so a tile is fetched from HBM once and from L2 by everyone else.
//
//   CTA = 256 query rows (two M = 128 tiles, one 32 KB TMA box) against one segment of the train set in tiles of 112 rows
//   (TMEM: 2 accumulator stages x 2 M tiles x 112 columns = 448, the unit scale factors in the columns behind them).
//   warps 0-7  epilogue, one query row per thread: tcgen05.ld 16 columns at a time (next load in flight), max tree over
//              the 16 fp32 accumulators and ONE compare against thr = 256 - 2 * (second-best distance) — exact for the
//              same reason as in match_knn_mma.cu (columns arrive in ascending train index) —, the packed-key update
//              d << 20 | index only behind it.
//   warp 8     TMA producer (one lane): query box once, then train boxes through a 6-stage ring.
//   warp 9     TMEM allocation and the MMA issue: per tile 2 x 4 tcgen05.mma (M 128, N 112, K 64), tcgen05.commit onto
//              the stage's "empty" barrier and the accumulator's "full" barrier.
#include <climits>
#include <cstdio>

#include <cuda.h>

#include "api_common.cuh"
#include "orbfe_internal.cuh"
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr int MQ = 256;                 // query rows per CTA
constexpr int NT = 112;                 // train rows per tile (multiple of 16; 4 x 112 accumulator columns + scale factors <= 512)
constexpr int STAGES = 6;
constexpr int EPI_WARPS = 8;
constexpr int THREADS = (EPI_WARPS + 2) * 32;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;
constexpr int A_BYTES = MQ * 128, B_BYTES = NT * 128;
constexpr int ACC_STAGE_COLS = 2 * NT;  // two M tiles
constexpr int SF_COL = 2 * ACC_STAGE_COLS;   // 448: 32 columns of 0x7F in every lane serve as SFA and SFB

struct alignas(1024) Fp4Smem {
    uint8_t a[A_BYTES];                           // 256 rows x 128 B, 8-row groups swizzled (as TMA writes them)
    uint8_t b[STAGES][B_BYTES];                   // 6 x 14 KB
    unsigned long long full_b[STAGES], empty_b[STAGES], acc_full[2], acc_empty[2], a_full;
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
// one TMA box (128 bytes x rows, starting at row `row`) into shared memory, completion counted on `bar`
__device__ __forceinline__ void tma_rows(void* dst, const CUtensorMap* map, int row, unsigned long long* bar, uint32_t bytes) {
    const uint32_t b = smem_u32(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(0), "r"(row), "r"(b) : "memory");
}
// K-major SWIZZLE_128B operand descriptor (rows of 128 bytes, 8-row groups 1024 bytes apart)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)2 << 61);
}
// block-scaled instruction descriptor: A, B = E2M1 (format 1), K-major, N, UE8M0 scales, M = 128, scale-factor ids 0
constexpr uint32_t kIdesc = (1u << 7) | (1u << 10) | ((uint32_t)(NT >> 3) << 17) | (1u << 23) | ((uint32_t)(128 >> 4) << 24);

__device__ __forceinline__ void umma_mxf4(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t accumulate, uint32_t sfa, uint32_t sfb) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%6], p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(kIdesc), "r"(accumulate), "r"(sfa), "r"(sfb) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

#define TMEM_LD16(taddr, v)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];" \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])                 \
                 : "r"(taddr) : "memory")

// 16 accumulators (256 - 2 d, exact in fp32) of one query against train rows idx0 .. idx0 + 15 of the segment, folded
// into the packed keys (k1, k2) = (best, second best); thr = 256 - 2 * (k2 >> 20).  valid < 16 only in a segment's last tile.
__device__ __forceinline__ void fold16(const uint32_t (&v)[16], uint32_t idx0, int valid, uint32_t& k1, uint32_t& k2, float& thr) {
    float m = fmaxf(__uint_as_float(v[0]), __uint_as_float(v[1]));
#pragma unroll
    for (int i = 2; i < 16; i += 2) m = fmaxf(m, fmaxf(__uint_as_float(v[i]), __uint_as_float(v[i + 1])));
    if (m > thr) {
        constexpr uint32_t NEG = 0u - (1u << 19);                 // (256 - acc) << 19 == d << 20
        const uint32_t c0 = (256u << 19) + idx0;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const uint32_t acc = (uint32_t)__float2int_rn(__uint_as_float(v[i]));
            const uint32_t key = i < valid ? acc * NEG + (c0 + (uint32_t)i) : KEY_NONE;
            k2 = min(k2, max(k1, key));
            k1 = min(k1, key);
        }
        thr = (float)(256 - 2 * (int)(k2 >> 20));
    }
}

// 8 bits -> 8 E2M1 nibbles, bit i to nibble i: 0 -> 0x2 (+1.0), 1 -> 0xA (-1.0)
__device__ __forceinline__ uint32_t expand8(uint32_t x) {
    uint32_t t = (x | (x << 12)) & 0x000F000Fu;
    t = (t | (t << 6)) & 0x03030303u;
    t = (t | (t << 3)) & 0x11111111u;
    return (t << 3) | 0x22222222u;
}

// pre-pass: packed 256-bit rows -> 128-byte FP4 rows; one 32-bit word (-> 16 bytes) per thread
__global__ void knn_expand_fp4_kernel(const uint32_t* __restrict__ rows, long long nwords, uint4* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nwords) return;
    const uint32_t w = __ldg(rows + i);
    out[i] = make_uint4(expand8(w & 255u), expand8((w >> 8) & 255u), expand8((w >> 16) & 255u), expand8(w >> 24));
}

__global__ void __launch_bounds__(THREADS, 1)
knn2_fp4_kernel(const __grid_constant__ CUtensorMap qmap, const __grid_constant__ CUtensorMap tmap, int nq, int nt, int segLen,
                int* __restrict__ part_idx, int* __restrict__ part_d1, int* __restrict__ part_d2) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    Fp4Smem& S = *reinterpret_cast<Fp4Smem*>(smem_raw);
    if ((smem_u32(smem_raw) & 1023u) != 0) __trap();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int q0 = blockIdx.x * MQ;
    const int seg = blockIdx.y;
    const long long seg0 = (long long)seg * segLen;
    const int segn = (int)min((long long)segLen, (long long)nt - seg0);        // > 0 by construction of the grid
    const int ntiles = (segn + NT - 1) / NT;

    if (tid == 0) {
        for (int i = 0; i < STAGES; i++) { mbar_init(&S.full_b[i], 1); mbar_init(&S.empty_b[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], EPI_WARPS); }
        mbar_init(&S.a_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == EPI_WARPS + 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&S.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = S.tmem_base;
    // unit scale factors: every byte of columns SF_COL .. SF_COL + 31 in all 128 lanes is UE8M0 2^0, so whichever bytes
    // the MMA reads for a row / column block, the scale is 1 (warp w of the first four owns lanes 32 w .. 32 w + 31)
    if (warp < 4) {
        const uint32_t one = 0x7F7F7F7Fu;
#pragma unroll 1
        for (int c = 0; c < 32; c++)
            asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(tmem + ((uint32_t)(warp * 32) << 16) + SF_COL + c), "r"(one) : "memory");
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    if (warp < EPI_WARPS) {
        // ---------------- epilogue: thread <-> query row q0 + tid (TMEM lane tid & 127 of M tile tid >> 7)
        const int mtile = warp >> 2;
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t k1 = KEY_NONE, k2 = KEY_NONE;
        float thr = (float)(256 - 2 * (int)(KEY_NONE >> 20));      // below every accumulator: the first groups take the exact path
        for (int t = 0; t < ntiles; t++) {
            const int a = t & 1;
            mbar_wait(&S.acc_full[a], (t >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tbase = tmem + lane_base + (uint32_t)(a * ACC_STAGE_COLS + mtile * NT);
            const uint32_t idx0 = (uint32_t)t * NT;
            const int left = segn - (int)idx0;                       // valid columns of this tile (>= NT except in the last one)
            const bool tail = left < NT;                             // padding columns must never win: exact path with masking
            if (tail) thr = -1e30f;
            uint32_t v0[16], v1[16];
            TMEM_LD16(tbase, v0);
#pragma unroll
            for (int c = 0; c < NT; c += 32) {
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (c + 16 < NT) TMEM_LD16(tbase + c + 16, v1);
                fold16(v0, idx0 + c, left - c, k1, k2, thr);
                if (tail) thr = -1e30f;
                if (c + 16 < NT) {
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (c + 32 < NT) TMEM_LD16(tbase + c + 32, v0);
                    fold16(v1, idx0 + c + 16, left - c - 16, k1, k2, thr);
                    if (tail) thr = -1e30f;
                }
            }
            if (tail) thr = (float)(256 - 2 * (int)(k2 >> 20));
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
    } else if (warp == EPI_WARPS) {
        // ---------------- TMA producer: rows past the end of a tensor arrive as zeros (and count as bytes)
        if (lane == 0) {
            tma_rows(S.a, &qmap, q0, &S.a_full, A_BYTES);
            for (int t = 0; t < ntiles; t++) {
                const int s = t % STAGES;
                mbar_wait(&S.empty_b[s], ((t / STAGES) & 1) ^ 1);
                tma_rows(S.b[s], &tmap, (int)(seg0 + (long long)t * NT), &S.full_b[s], B_BYTES);
            }
        }
        __syncwarp();
    } else {
        // ---------------- MMA issue: one elected lane
        if (lane == 0) {
            const uint32_t aaddr = smem_u32(S.a);
            const uint32_t sfa = tmem + SF_COL, sfb = tmem + SF_COL + 16;
            mbar_wait(&S.a_full, 0);
            for (int t = 0; t < ntiles; t++) {
                const int s = t % STAGES, a = t & 1;
                mbar_wait(&S.full_b[s], (t / STAGES) & 1);
                mbar_wait(&S.acc_empty[a], ((t >> 1) & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t bd = umma_desc(smem_u32(S.b[s]));
#pragma unroll
                for (int m = 0; m < 2; m++) {
                    const uint32_t d = tmem + (uint32_t)(a * ACC_STAGE_COLS + m * NT);
                    const uint64_t ad = umma_desc(aaddr + m * 128 * 128);
#pragma unroll
                    for (int k = 0; k < 4; k++) umma_mxf4(d, ad + 2 * k, bd + 2 * k, k != 0, sfa, sfb);   // +32 bytes = 64 FP4 of K
                }
                umma_commit(&S.empty_b[s]);
                umma_commit(&S.acc_full[a]);
            }
        }
        __syncwarp();
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == EPI_WARPS + 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// [rows][128 B] byte tensor, box = 128 bytes x boxRows with the 128-byte swizzle (the MMA's canonical K-major layout)
int encode_rows_map(CUtensorMap* m, const void* base, long long rows, int boxRows) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        const cudaError_t ce = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (ce != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            return fail(ORBFE_ERR_CUDA, "cudaGetDriverEntryPoint(cuTensorMapEncodeTiled)", ce);
        encode = (encode_fn)fn;
    }
    const cuuint64_t dims[2] = {128, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {128};
    const cuuint32_t box[2] = {128, (cuuint32_t)boxRows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ORBFE_ERR_CUDA, "cuTensorMapEncodeTiled failed (FP4 kNN rows)");
    return 0;
}

}  // namespace

// scratch of the FP4 engine: per-segment partial results, then both row sets in FP4 form (128 bytes per row, 1024-byte aligned)
static size_t fp4_partials_bytes(int nq, int nt) {
    int len;
    const int nseg = knn2_tile_segments(nq, nt, MQ, NT, &len);
    return ((size_t)3 * nseg * nq * sizeof(int) + 1023) & ~size_t(1023);
}
size_t knn2_fp4_scratch_bytes(int nq, int nt) {
    return fp4_partials_bytes(nq, nt) + ((size_t)nq + (size_t)nt) * 128 + 1024;
}

int launch_knn2_fp4(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                    int* d_idx, int* d_d1, int* d_d2, void* scratch, cudaStream_t s) {
    if (nq <= 0) return 0;
    if (nt <= 0) return launch_knn2_merge(nq, 0, th, ratio, nullptr, nullptr, nullptr, d_idx, d_d1, d_d2, s);
    int segLen;
    const int nseg = knn2_tile_segments(nq, nt, MQ, NT, &segLen);
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(scratch) + 1023) & ~uintptr_t(1023));
    int* part_idx = (int*)base;
    int* part_d1 = part_idx + (size_t)nseg * nq;
    int* part_d2 = part_d1 + (size_t)nseg * nq;
    uint8_t* q4 = base + fp4_partials_bytes(nq, nt);
    uint8_t* t4 = q4 + (size_t)nq * 128;
    CUtensorMap qmap, tmap;
    int rc = encode_rows_map(&qmap, q4, nq, MQ);
    if (!rc) rc = encode_rows_map(&tmap, t4, nt, NT);
    if (rc) return -1;
    const long long qw = (long long)nq * 8, tw = (long long)nt * 8;
    knn_expand_fp4_kernel<<<(unsigned)((qw + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_query), qw, reinterpret_cast<uint4*>(q4));
    knn_expand_fp4_kernel<<<(unsigned)((tw + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint32_t*>(d_train), tw, reinterpret_cast<uint4*>(t4));
    constexpr int smem = (int)sizeof(Fp4Smem);
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &cap);
    if (cap == cudaStreamCaptureStatusNone)
        cudaFuncSetAttribute(knn2_fp4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    knn2_fp4_kernel<<<dim3((nq + MQ - 1) / MQ, nseg), THREADS, smem, s>>>(qmap, tmap, nq, nt, segLen, part_idx, part_d1, part_d2);
    return 3 + launch_knn2_merge(nq, nseg, th, ratio, part_idx, part_d1, part_d2, d_idx, d_d1, d_d2, s);
}

}  // namespace orbfe
