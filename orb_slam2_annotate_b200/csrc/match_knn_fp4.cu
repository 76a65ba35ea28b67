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
//   CTA = 256 query rows (two M = 128 tiles, one 32 KB TMA box) against one segment of the train set in tiles of 192 rows.
//   One MMA covers a whole tile (M 128, N 192, K 64): an SS-mode MMA reads both operands from shared memory every time,
//   and with narrow N the 128 B/clk of shared memory, not the tensor pipe, sets the pace (measured without an epilogue:
//   N 112 -> 10.4 T comparisons/s, N 160 -> 11.5 T, N 192 -> 14.7 T, N 224 -> 15.5 T).  TMEM then holds two
//   accumulators of 192 columns (+ the unit scale factors in the columns behind them); the (tile, M tile) pairs take them
//   in turn, so the epilogue of one M tile runs under the MMAs of the other.
//   warps 0-15 epilogue, two threads per query row (a warp reads the TMEM lanes of its quadrant: warps w and w + 8 share
//              32 rows and take 96 columns each): tcgen05.ld 32 columns at a time (next load in flight), a balanced tree
//              of three-input integer maxima over the accumulators' bit patterns and ONE compare against the pattern of
//              256 - 2 * (second-best distance) — exact for the same reason as in match_knn_mma.cu (columns arrive in
//              ascending train index).  A group that passes revisits only the triples whose own maximum passed (the
//              packed-key update d << 20 | index, key formed without F2I); the accumulator is handed back as soon as the
//              last columns are in registers.  Measured: hand-off latency, not arithmetic, is what separates the kernel
//              (10.4 T) from its MMA pipeline (14.7 T): three accumulators would hide it but do not fit TMEM at this N.
//   warp 16    TMA producer (one lane): query box once, then train boxes through a 4-stage ring.
//   warp 17    TMEM allocation and the MMA issue: per tile 2 x 4 tcgen05.mma, tcgen05.commit onto the accumulator's
//              "full" barrier and the stage's "empty" barrier.
#include <climits>
#include <cstdio>
#include <cstdlib>

#include <cuda.h>

#include "api_common.cuh"
#include "orbfe_internal.cuh"
#include "match_internal.cuh"

namespace orbfe {

namespace {

constexpr int MQ = 256;                 // query rows per CTA
#ifndef KNN_FP4_NT
#define KNN_FP4_NT 192
#define KNN_FP4_NACC 2
#endif
constexpr int NT = KNN_FP4_NT;          // train rows per tile = N of one MMA; NACC accumulators of NT columns + scale factors <= 512
constexpr int NACC = KNN_FP4_NACC;      // accumulators, taken in turn by the (tile, M tile) pairs
constexpr int STAGES = NT > 170 ? 4 : 5;
static_assert(NACC * NT + 32 <= 512 && NT % 16 == 0, "TMEM budget / MMA shape");
constexpr int EPI_WARPS = 16;             // two per (M tile, lane quadrant): each takes half of a tile's columns
constexpr int THREADS = (EPI_WARPS + 2) * 32;
constexpr uint32_t KEY_NONE = 0xFFFFFFFFu;
constexpr int A_BYTES = MQ * 128, B_BYTES = NT * 128;
constexpr int SF_COL = NACC * NT;       // 480: 32 columns of 0x7F in every lane serve as SFA and SFB

struct alignas(1024) Fp4Smem {
    uint8_t a[A_BYTES];                           // 256 rows x 128 B, 8-row groups swizzled (as TMA writes them)
    uint8_t b[STAGES][B_BYTES];                   // 5 x 20 KB
    unsigned long long full_b[STAGES], empty_b[STAGES], acc_full[NACC], acc_empty[NACC], a_full;
    uint32_t tmem_base;
    uint32_t xk1[MQ], xk2[MQ];                    // the second column half's (best, second best) on its way to the first
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

#define TMEM_LD32(taddr, v)                                                                                              \
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                               \
                 "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];" \
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),       \
                   "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),           \
                   "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),          \
                   "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                       \
                 : "r"(taddr) : "memory")

// The accumulators are fp32 but the screen runs on their bit patterns as signed integers (three-input integer max,
// VIMNMX3: the fp32 three-input max measured four times slower here): non-negative floats order like their patterns,
// and every negative float is a negative integer.  So the integer maximum IS the float maximum whenever that is >= 0;
// otherwise it is merely some negative number — which decides "max > thr" correctly for every thr >= 0.  While the
// threshold is still negative (no second best yet, or one beyond distance 128) the caller passes INT_MIN and every
// group takes the exact path.
__device__ __forceinline__ int imax3(int a, int b, int c) { return max(max(a, b), c); }
// the screen's threshold for a second-best key k2: bit pattern of 256 - 2 d2 when that is >= 0, else "always"
__device__ __forceinline__ int screen_threshold(uint32_t k2) {
    const int t = 256 - 2 * (int)(k2 >> 20);
    return t >= 0 ? __float_as_int((float)t) : INT_MIN;
}
// one accumulator into (best, second best): the key d << 20 | index from the fp32 value without a conversion
// instruction (adding 1.5 * 2^23 leaves the integer in the low mantissa bits; F2I runs on the quarter-rate pipe), then
// the strict-'<' rule  k2' = min(k2, max(k1, key)), k1' = min(k1, key)
__device__ __forceinline__ void take(uint32_t bits, uint32_t cidx, uint32_t& k1, uint32_t& k2) {
    constexpr uint32_t NEG = 0u - (1u << 19);                     // (256 - acc) << 19 == d << 20
    constexpr uint32_t MAGIC = 0x4B400000u;                       // bit pattern of 12582912.0f
    const uint32_t key = __float_as_uint(__fadd_rn(__uint_as_float(bits), 12582912.0f)) * NEG + (cidx - MAGIC * NEG);
    k2 = min(k2, max(k1, key));
    k1 = min(k1, key);
}

// N accumulators (256 - 2 d, exact in fp32) of one query against train rows idx0 .. idx0 + N - 1 of the segment, folded
// into the packed keys (k1, k2) = (best, second best); thr = screen_threshold(k2).  The common case is a balanced tree of
// three-input maxima (depth 4) and one compare.  A thread whose group passes does not walk all N columns — the other
// lanes of its warp, and the seven other warps that share the accumulator, wait for it —: it revisits the tree's first
// level and updates only the triples whose own maximum passed.  MASK: only the first `valid` columns exist (a segment's
// last tile); then, and while there is no usable threshold yet, every column is taken in turn.
template <int N, bool MASK>
__device__ __forceinline__ void fold(const uint32_t (&v)[N], uint32_t idx0, int valid, uint32_t& k1, uint32_t& k2, int& thr) {
    const uint32_t c0 = (256u << 19) + idx0;
    if (MASK || thr == INT_MIN) {
#pragma unroll
        for (int i = 0; i < N; i++)
            if (!MASK || i < valid) take(v[i], c0 + (uint32_t)i, k1, k2);
        thr = screen_threshold(k2);
        return;
    }
    constexpr int T = N / 3;                                      // 10 triples + 2 columns, or 5 triples + 1 column
    int b[T];
#pragma unroll
    for (int j = 0; j < T; j++) b[j] = imax3((int)v[3 * j], (int)v[3 * j + 1], (int)v[3 * j + 2]);
    int m;
    if (N == 32) m = max(imax3(imax3(b[0], b[1], b[2]), imax3(b[3], b[4], b[5]), imax3(b[6], b[7], b[8])), imax3(b[9 % T], (int)v[30 % N], (int)v[31 % N]));
    else m = max(imax3(b[0], b[1], b[2]), imax3(b[3], b[4], (int)v[15]));
    if (m > thr) {
        const int thr0 = thr;                                     // looser than the running one: never skips a column that matters
#pragma unroll
        for (int j = 0; j < T; j++)
            if (b[j] > thr0) {
                take(v[3 * j], c0 + 3u * j, k1, k2);
                take(v[3 * j + 1], c0 + 3u * j + 1u, k1, k2);
                take(v[3 * j + 2], c0 + 3u * j + 2u, k1, k2);
            }
#pragma unroll
        for (int i = 3 * T; i < N; i++)
            if ((int)v[i] > thr0) take(v[i], c0 + (uint32_t)i, k1, k2);
        thr = screen_threshold(k2);
    }
}

// one thread's share of an accumulator of one query row (COLS columns = G groups of 32 and, if COLS is not a multiple
// of 32, one of 16): the next tcgen05.ld is in flight while a group folds
template <bool MASK, int COLS>
__device__ __forceinline__ void fold_half_tile(uint32_t tbase, uint32_t idx0, int left, uint32_t& k1, uint32_t& k2, int& thr,
                                               unsigned long long* release, int lane) {
    constexpr int G = COLS / 32, R = COLS % 32;
    static_assert(R == 0 || R == 16, "column groups are 32 wide plus at most one of 16");
    uint32_t v[2][32], vr[16];
    // the accumulator goes back to the MMA warp as soon as this warp's last columns are in registers, before they fold
    auto give_back = [&]() {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(release);
    };
    TMEM_LD32(tbase, v[0]);
#pragma unroll
    for (int g = 0; g < G; g++) {
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (g + 1 < G) TMEM_LD32(tbase + 32 * (g + 1), v[(g + 1) & 1]);
        else if (R) TMEM_LD16(tbase + 32 * G, vr);
        else give_back();
        fold<32, MASK>(v[g & 1], idx0 + 32 * g, left - 32 * g, k1, k2, thr);
    }
    if (R) {
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        give_back();
        fold<16, MASK>(vr, idx0 + 32 * G, left - 32 * G, k1, k2, thr);
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
    // consecutive CTAs (one wave) take different train segments: every tile is then wanted by 148 / nseg SMs at a time
    // instead of all of them (the few L2 slices that hold a tile's 112 lines could not serve 148 readers in step)
    const int nseg = (int)gridDim.y;
    const int lin = (int)(blockIdx.y * gridDim.x + blockIdx.x);
    const int seg = lin % nseg;
    const int q0 = (lin / nseg) * MQ;
    const long long seg0 = (long long)seg * segLen;
    const int segn = (int)min((long long)segLen, (long long)nt - seg0);        // > 0 by construction of the grid
    const int ntiles = (segn + NT - 1) / NT;

    if (tid == 0) {
        for (int i = 0; i < STAGES; i++) { mbar_init(&S.full_b[i], 1); mbar_init(&S.empty_b[i], 1); }
        for (int i = 0; i < NACC; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], EPI_WARPS / 2); }
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
        // ---------------- epilogue: thread <-> (query row q0 + row, column half): TMEM lane row & 127 of M tile row >> 7
        const int mtile = (warp >> 2) & 1, half = warp >> 3;
        const int row = mtile * 128 + (warp & 3) * 32 + lane;
        const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
        uint32_t k1 = KEY_NONE, k2 = KEY_NONE;
        int thr = screen_threshold(KEY_NONE);                      // INT_MIN: the first groups take the exact path
        for (int t = 0; t < ntiles; t++) {
            const int j = 2 * t + mtile, b = j % NACC;               // this (tile, M tile) pair's accumulator
            mbar_wait(&S.acc_full[b], (j / NACC) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            constexpr int C0 = (NT / 2 + 31) / 32 * 32, C1 = NT - C0;     // the two threads' column shares (96 + 96 of 192)
            const uint32_t tbase = tmem + lane_base + (uint32_t)(b * NT + half * C0);
            const uint32_t idx0 = (uint32_t)t * NT + half * C0;
            const int left = segn - (int)idx0;                       // valid columns from idx0 on (>= the share except in the last tile)
            if (half == 0) {
                if (left >= C0) fold_half_tile<false, C0>(tbase, idx0, C0, k1, k2, thr, &S.acc_empty[b], lane);
                else fold_half_tile<true, C0>(tbase, idx0, left, k1, k2, thr, &S.acc_empty[b], lane);   // padding columns must never win: masked exact path
            } else {
                if (left >= C1) fold_half_tile<false, C1>(tbase, idx0, C1, k1, k2, thr, &S.acc_empty[b], lane);
                else fold_half_tile<true, C1>(tbase, idx0, left, k1, k2, thr, &S.acc_empty[b], lane);
            }
        }
        // the two column halves of a row meet: keys are (distance, index), so the merge is order free
        if (half) { S.xk1[row] = k1; S.xk2[row] = k2; }
        asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
        const int q = q0 + row;
        if (!half && q < nq) {
            const uint32_t o1 = S.xk1[row], o2 = S.xk2[row];
            k2 = min(min(k2, o2), max(k1, o1));
            k1 = min(k1, o1);
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
                const int s = t % STAGES;
                mbar_wait(&S.full_b[s], (t / STAGES) & 1);
                const uint64_t bd = umma_desc(smem_u32(S.b[s]));
                for (int m = 0; m < 2; m++) {
                    // three accumulators in turn: the epilogue of a pair runs under the MMAs of the next two
                    const int j = 2 * t + m, b = j % NACC;
                    mbar_wait(&S.acc_empty[b], ((j / NACC) & 1) ^ 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t d = tmem + (uint32_t)(b * NT);
                    const uint64_t ad = umma_desc(aaddr + m * 128 * 128);
#pragma unroll
                    for (int k = 0; k < 4; k++) umma_mxf4(d, ad + 2 * k, bd + 2 * k, k != 0, sfa, sfb);   // +32 bytes = 64 FP4 of K
                    umma_commit(&S.acc_full[b]);
                }
                umma_commit(&S.empty_b[s]);
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
