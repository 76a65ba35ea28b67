// orient_desc.cu — K4 + K6 + result assembly: IC_Angle (reference src/ORBextractor.cc:78-106),
// computeOrbDescriptor (:111-152) and the keypoint fields operator() finalises (:905-916,
// :1187-1195), one warp per surviving keypoint.
//
//   * orientation: lanes 0..30 own the patch columns u = -15..15 and walk the 31 rows, so a
//     warp reads one contiguous 31-byte row segment per step; the int32 moments m10 = sum u*I,
//     m01 = sum v*I over the disc |u| <= umax[|v|] are reduced with shuffles (integer sums, so
//     the order is free) and fed to a float32 restatement of cv::fastAtan2 written with
//     explicit round-to-nearest intrinsics so that nvcc cannot contract mul+add into FMA.
//   * descriptor: lane i produces byte i from pattern pairs 8i..8i+7 of the blurred level;
//     x*b + y*a and x*a - y*b keep the reference's float op order (no FMA), cvRound is
//     round-half-even (__float2int_rn).
//   * output slot = keypoints of lower levels + position in this level's list, i.e. the
//     reference's level-by-level concatenation.
#include "orbfe_internal.cuh"
#include "sincosf_glibc.h"

namespace orbfe {

namespace {

constexpr unsigned FULL = 0xFFFFFFFFu;

__device__ __align__(16) const signed char d_pattern[1024] = {
#include "../../include/orbfe_pattern.inc"
};

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

__global__ void __launch_bounds__(256)
orient_desc_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ DevBufs d,
                   const int* __restrict__ umax) {
    const int frame = blockIdx.y + g.frame0;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int kpi = blockIdx.x * 8 + wid;
    const int* selc = d.selc + (long long)frame * g.nlevels;
    int total = 0, lvl = -1, k = 0;
    for (int l = 0; l < g.nlevels; l++) {
        const int c = selc[l];
        if (lvl < 0 && kpi < total + c) { lvl = l; k = kpi - total; }
        total += c;
    }
    if (blockIdx.x == 0 && tid == 0) d.counts[frame] = total;
    if (lvl < 0) return;
    const LevelGeom& L = g.lv[lvl];
    const uint32_t p = d.sel[(long long)frame * g.selTotal + L.selBase + k];
    const int x = kp_x(p), y = kp_y(p), score = kp_score(p);
    const long long off = L.pyrOff + (long long)frame * L.lvBytes + (long long)y * L.pitch + x;

    // ---- IC_Angle
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        // the disc is symmetric (checked on the host): column u holds rows |v| <= umax[|u|]
        const uint8_t* c = d.pyr + off;
        const int u = lane - kHalfPatch, vm = __ldg(&umax[abs(u)]);
        int sum = 0;
#pragma unroll
        for (int v = -kHalfPatch; v <= kHalfPatch; v++) {     // 31 independent predicated loads
            if (abs(v) <= vm) {
                const int val = c[v * L.pitch + u];
                sum += val;
                m01 += v * val;
            }
        }
        m10 = u * sum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(FULL, m10, o);
        m01 += __shfl_xor_sync(FULL, m01, o);
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
    const uint8_t* cb = d.blur + off;
    const int4* pat4 = reinterpret_cast<const int4*>(d_pattern) + 2 * lane;
    const int4 w0 = __ldg(pat4), w1 = __ldg(pat4 + 1);
    const int words[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    int val = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        // I2F of the sign-extended bytes (PRMT-based sign extension keeps this to two ops each)
        const float x0 = (float)(int)(signed char)(words[j]), y0 = (float)((words[j] << 16) >> 24);
        const float x1 = (float)((words[j] << 8) >> 24), y1 = (float)(words[j] >> 24);
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        const int t0 = cb[r0 * L.pitch + c0], t1 = cb[r1 * L.pitch + c1];
        val |= (t0 < t1) << j;
    }
    const long long slot = (long long)frame * g.outCap + kpi;
    d.desc[slot * 32 + lane] = (uint8_t)val;
    if (lane < 7) {
        float f;
        switch (lane) {
            case 0: f = lvl ? __fmul_rn((float)x, L.scale) : (float)x; break;   // :1187-1193
            case 1: f = lvl ? __fmul_rn((float)y, L.scale) : (float)y; break;
            case 2: f = (float)L.scaledPatch; break;                            // :905,:915
            case 3: f = angle; break;
            case 4: f = (float)score; break;
            case 5: f = __int_as_float(lvl); break;
            default: f = __int_as_float(-1); break;
        }
        reinterpret_cast<float*>(d.kps + slot)[lane] = f;
    }
}

}  // namespace

int launch_orient_desc(const FrameSetGeom& g, const DevBufs& d, const int* d_umax, cudaStream_t s) {
    orient_desc_kernel<<<dim3((g.outCap + 7) / 8, g.batch), 256, 0, s>>>(g, d, d_umax);
    return 1;
}

}  // namespace orbfe
