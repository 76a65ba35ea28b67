// orient_desc.cu — K4 + K6 + result assembly: IC_Angle (reference src/ORBextractor.cc:78-106),
// computeOrbDescriptor (:111-152) and the keypoint fields operator() finalises (:905-916,
// :1187-1195), one warp per surviving keypoint.
//
//   * windows: the 31x31 patch of the level and the 37x37 patch of the blurred level arrive by
//     TMA (cp.async.bulk.tensor.2d on [frames * rows][pitch] byte tensors, 16-byte aligned box
//     starts), one keypoint ahead: the level window into a single buffer (it is dead as soon as the
//     moments are summed), the blurred window into a double buffer; a warp walks KPW keypoints.  51 KB of
//     shared memory and 64 registers let four CTAs live on an SM.  Measured alone the window traffic
//     (154 L2 sectors per keypoint, 9.3 TB/s) and the arithmetic each take 0.55 ms per 1024 frames; the
//     kernel runs at 0.70, so it is bound by the L2 path as much as by issue (round-1 experiments:
//     a three-phase variant that evaluates atan2 / sin / cos once per eight keypoints, one per lane,
//     removes a quarter of the instructions and changes nothing).
//   * orientation: lanes 0..30 own the patch columns u = -15..15 and walk the row pairs +-v
//     from shared memory; the int32 moments m10 = sum u*I, m01 = sum v*I over the disc
//     |u| <= umax[|v|] are reduced with shuffles (integer sums, so the order is free) and fed
//     to a float32 restatement of cv::fastAtan2 written with explicit round-to-nearest
//     intrinsics so that nvcc cannot contract mul+add into FMA.
//   * descriptor: lane i produces byte i from pattern pairs 8i..8i+7 of the blurred level (its
//     16 pattern points stay in registers as fp16 pairs); x*b + y*a and x*a - y*b keep the
//     reference's float op order (no FMA), cvRound is round-half-even (__float2int_rn).
//   * output slot = keypoints of lower levels + position in this level's list, i.e. the
//     reference's level-by-level concatenation.
#include "orbfe_internal.cuh"
#include <cstring>
#include "sincosf_glibc.h"
#include <cuda_fp16.h>
#include <cuda.h>

namespace orbfe {

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;

// cv::fastAtan2, scalar path: degrees in [0, 360)
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    const float eps = 2.2204460492503131e-16f;   // (float)DBL_EPSILON
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// Pattern as fp16 (x, y) pairs (|coordinate| <= 13 is exact), transposed for the warp: entry
// i * 32 + lane is point 16 * lane + i, i.e. lane's 16 points are 16 coalesced 128-byte rows.
// A warp keeps its 16 words in registers across the keypoints it processes.
struct PatH { uint32_t v[512]; };
constexpr signed char h_pattern[1024] = {
#include "../../include/orbfe_pattern.inc"
};
constexpr uint32_t half_bits(int v) {
    const uint32_t sgn = v < 0 ? 0x8000u : 0u;
    uint32_t n = (uint32_t)(v < 0 ? -v : v);
    if (n == 0) return sgn;
    int e = 0;
    while ((n >> (e + 1)) != 0) e++;
    return sgn | ((uint32_t)(e + 15) << 10) | ((n << (10 - e)) & 0x3FFu);
}
constexpr PatH make_path() {
    PatH p{};
    for (int lane = 0; lane < 32; lane++)
        for (int i = 0; i < 16; i++)
            p.v[i * 32 + lane] = half_bits(h_pattern[32 * lane + 2 * i]) | (half_bits(h_pattern[32 * lane + 2 * i + 1]) << 16);
    return p;
}
__device__ __align__(16) const PatH d_path = make_path();

constexpr int RAW_ROWS = 2 * kHalfPatch + 1, RAW_STRIDE = kTmaRawW;   // 31 rows x 48 bytes
constexpr int BLR_HALF = 18, BLR_ROWS = 2 * BLR_HALF + 1, BLR_STRIDE = kTmaBlurW;   // rotated pattern stays within 18 px: 37 rows x 64 bytes
constexpr int RAW_BYTES = RAW_ROWS * RAW_STRIDE, BLR_BYTES = BLR_ROWS * BLR_STRIDE;
constexpr int RAW_SLOT = (RAW_BYTES + 127) / 128 * 128, BLR_SLOT = (BLR_BYTES + 127) / 128 * 128;   // TMA destinations are 128-byte aligned
constexpr int WARP_BYTES = RAW_SLOT + 2 * BLR_SLOT;   // one level window (dead once the moments are summed) + two blurred windows
static_assert(RAW_ROWS == kTmaRawH && BLR_ROWS == kTmaBlurH, "TMA boxes are built on the host from orbfe_internal.cuh");
constexpr int KPW_BATCH = 4;   // keypoints per warp (the pattern registers and the prologue amortise over them)
constexpr int KPW_SMALL = 1;   // batches <= 4: one keypoint per warp, four times the warps, no serial walk

// One warp per keypoint: the moment rows and the 512 pattern lookups are LDS.U8 on the two windows in shared memory.
// TMA descriptors travel as a kernel parameter: {level, blurred level} per pyramid level
struct alignas(64) LevelMaps { CUtensorMap m[2 * kMaxLevels]; };

template <int KPW>
__global__ void __launch_bounds__(256, 4)
orient_desc_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ DevBufs d,
                   const __grid_constant__ LevelMaps maps, const int* __restrict__ umax) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    __shared__ __align__(8) uint64_t bars[8][3];   // per warp: level window, blurred window 0 / 1
    const int frame = blockIdx.y + g.frame0;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // level and in-level index of a keypoint: inclusive scan of the per-level counts
    int incl = lane < g.nlevels ? d.selc[(long long)frame * g.nlevels + lane] : 0;
#pragma unroll
    for (int o = 1; o < kMaxLevels; o <<= 1) {
        const int t = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += t;
    }
    const int total = __shfl_sync(FULL, incl, kMaxLevels - 1);   // lanes < 16 hold complete prefixes
    if (blockIdx.x == 0 && tid == 0) d.counts[frame] = total;
    if ((blockIdx.x * KPW) * 8 + wid >= total) return;
    uint8_t* wbuf = smem_dyn + wid * WARP_BYTES;
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 3; i++)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bars[wid][i])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t pat[16];
#pragma unroll
    for (int i = 0; i < 16; i++) pat[i] = d_path.v[i * 32 + lane];
    const int um = lane < 31 ? __ldg(&umax[abs(lane - kHalfPatch)]) : -1;

    // level, packed position and window fetch of one keypoint.  The two windows are TMA box
    // loads (cp.async.bulk.tensor.2d) from the level-major [frames * rows][pitch] tensors, issued
    // by lane 0 one keypoint ahead into the other buffer; completion is an mbarrier per buffer.
    auto locate = [&](int kpi, int& lvl, uint32_t& p) {
        lvl = __popc(__ballot_sync(FULL, lane < g.nlevels && kpi >= incl));
        const int before = __shfl_sync(FULL, incl, max(lvl - 1, 0));
        const int k = kpi - (lvl ? before : 0);
        p = d.sel[(long long)frame * g.selTotal + g.lv[lvl].selBase + k];
    };
    // the level window has one buffer (free again as soon as the moments are summed), the blurred window two
    auto fetch = [&](int lvl, uint32_t p, int buf) {
        if (lane == 0) {
            const int x = kp_x(p), y = kp_y(p) + frame * g.lv[lvl].h;
            const uint32_t barR = (uint32_t)__cvta_generic_to_shared(&bars[wid][0]);
            const uint32_t barB = (uint32_t)__cvta_generic_to_shared(&bars[wid][1 + buf]);
            const uint32_t rdst = (uint32_t)__cvta_generic_to_shared(wbuf);
            const uint64_t mraw = reinterpret_cast<uint64_t>(&maps.m[2 * lvl]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barR), "r"(RAW_BYTES) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(rdst), "l"(mraw), "r"((x - kHalfPatch) & ~15), "r"(y - kHalfPatch), "r"(barR) : "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barB), "r"(BLR_BYTES) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(rdst + RAW_SLOT + buf * BLR_SLOT), "l"(mraw + 128), "r"((x - BLR_HALF) & ~15), "r"(y - BLR_HALF), "r"(barB) : "memory");
        }
    };
    auto wait = [&](int which, int phase) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&bars[wid][which]);
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar), "r"(phase) : "memory");
    };

    int lvlN;
    uint32_t pN;
    locate((blockIdx.x * KPW) * 8 + wid, lvlN, pN);
    fetch(lvlN, pN, 0);
    for (int it = 0; it < KPW; it++) {
    const int kpi = (blockIdx.x * KPW + it) * 8 + wid, buf = it & 1;
    const int lvl = lvlN;
    const uint32_t p = pN;
    const bool more = it + 1 < KPW && kpi + 8 < total;
    wait(0, it & 1);
    const LevelGeom& L = g.lv[lvl];
    const int x = kp_x(p), y = kp_y(p), score = kp_score(p);
    // ---- IC_Angle
    int m10 = 0, m01 = 0;
    {
        // the disc is symmetric (checked on the host): column u holds rows |v| <= umax[|u|]
        const uint8_t* c = wbuf + ((x - kHalfPatch) & 15) + min(lane, 30);       // (u, v = -15)
        // rows +v and -v share their predicate; the selects keep the loop branch-free
        int sum = c[kHalfPatch * RAW_STRIDE];         // v = 0 (every column has it; lane 31 is zeroed below)
#pragma unroll
        for (int v = 1; v <= kHalfPatch; v++) {
            const bool in = v <= um;
            const int lo = c[(kHalfPatch - v) * RAW_STRIDE], hi = c[(kHalfPatch + v) * RAW_STRIDE];
            sum += in ? lo + hi : 0;
            m01 += in ? v * (hi - lo) : 0;
        }
        if (lane == 31) sum = 0;
        m10 = (lane - kHalfPatch) * sum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(FULL, m10, o);
        m01 += __shfl_xor_sync(FULL, m01, o);
    }
    // the shuffles above ordered every lane's reads of the level window, and the previous keypoint's readers of the
    // other blurred buffer finished before this iteration: the next keypoint's windows can be fetched now
    if (more) {
        locate(kpi + 8, lvlN, pN);
        fetch(lvlN, pN, buf ^ 1);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // ---- rBRIEF byte `lane`
    const float factorPI = (float)(3.14159265358979323846 / 180.f);   // :108
    const float ang = __fmul_rn(angle, factorPI);
    // cos/sin resolve to glibc's float routines in the reference (:68,:115-116); sincosf_glibc.h
    // restates them bit for bit.  The angle is warp-uniform, so every lane evaluates it without
    // divergence.
    float a, b;
    orbfe_scf::sincosf_glibc(ang, &b, &a);
    wait(1 + buf, (it >> 1) & 1);
    const uint8_t* cb = wbuf + RAW_SLOT + buf * BLR_SLOT + BLR_HALF * BLR_STRIDE + BLR_HALF + ((x - BLR_HALF) & 15);   // patch centre
    int val = 0;
    // cvRound without the conversion pipe (F2I issues at a quarter of the FADD rate): for |v| < 2^22 the float
    // v + 1.5 * 2^23 holds round-half-even(v) in its low mantissa bits, i.e. its bit pattern is kMagicBits + cvRound(v)
    constexpr float kMagic = 12582912.0f;
    constexpr int kMagicBits = 0x4B400000;
    auto cv_round = [](float v) { return __float_as_int(__fadd_rn(v, kMagic)) - kMagicBits; };
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const float2 q0 = __half22float2(*reinterpret_cast<const __half2*>(&pat[2 * j]));
        const float2 q1 = __half22float2(*reinterpret_cast<const __half2*>(&pat[2 * j + 1]));
        const int r0 = cv_round(__fadd_rn(__fmul_rn(q0.x, b), __fmul_rn(q0.y, a)));
        const int c0 = cv_round(__fsub_rn(__fmul_rn(q0.x, a), __fmul_rn(q0.y, b)));
        const int r1 = cv_round(__fadd_rn(__fmul_rn(q1.x, b), __fmul_rn(q1.y, a)));
        const int c1 = cv_round(__fsub_rn(__fmul_rn(q1.x, a), __fmul_rn(q1.y, b)));
        const int t0 = cb[r0 * BLR_STRIDE + c0], t1 = cb[r1 * BLR_STRIDE + c1];
        val |= (t0 < t1) << j;
    }
    const long long slot = (long long)frame * g.outCap + kpi;
    d.desc[slot * 32 + lane] = (uint8_t)val;
    // keypoint record, one float per lane 0..6 — :1187-1193 (coordinates), :905,:915 (size)
    const float fx = lvl ? __fmul_rn((float)x, L.scale) : (float)x;
    const float fy = lvl ? __fmul_rn((float)y, L.scale) : (float)y;
    float f = __int_as_float(-1);
    f = lane == 5 ? __int_as_float(lvl) : f;
    f = lane == 4 ? (float)score : f;
    f = lane == 3 ? angle : f;
    f = lane == 2 ? (float)L.scaledPatch : f;
    f = lane == 1 ? fy : f;
    f = lane == 0 ? fx : f;
    if (lane < 7) reinterpret_cast<float*>(d.kps + slot)[lane] = f;
    if (!more) break;
    }
}

}  // namespace

int launch_orient_desc(const FrameSetGeom& g, const DevBufs& d, const int* d_umax, cudaStream_t s) {
    LevelMaps maps;
    std::memcpy(&maps, d.tmaps, sizeof(CUtensorMap) * 2 * g.nlevels);
    constexpr int smem = 8 * WARP_BYTES;
    // per device and cheap, so set on every eager launch rather than tracked per context (a stream capture replays a
    // schedule whose eager run already applied it)
    const bool small = g.batch <= 4;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(s, &cap);
    if (cap == cudaStreamCaptureStatusNone &&
        cudaFuncSetAttribute(small ? orient_desc_kernel<KPW_SMALL> : orient_desc_kernel<KPW_BATCH>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             smem) != cudaSuccess) return 0;
    if (small) orient_desc_kernel<KPW_SMALL><<<dim3((g.outCap + 8 * KPW_SMALL - 1) / (8 * KPW_SMALL), g.batch), 256, smem, s>>>(g, d, maps, d_umax);
    else orient_desc_kernel<KPW_BATCH><<<dim3((g.outCap + 8 * KPW_BATCH - 1) / (8 * KPW_BATCH), g.batch), 256, smem, s>>>(g, d, maps, d_umax);
    return 1;
}

}  // namespace orbfe
