// sincosf_glibc.h — float sine / cosine with the arithmetic of glibc >= 2.28 sinf/cosf (the
// routines the reference's `cos(angle)` / `sin(angle)` calls resolve to, src/ORBextractor.cc:
// 68,115-116): double-precision range reduction by pi/2 with a 2^24-prescaled 2/pi and a
// degree-7/8 minimax polynomial evaluated in double, rounded once to float.  Algorithm and
// coefficients are those published in ARM's optimized-routines "sincosf" (Szabolcs Nagy /
// Wilco Dijkstra), which glibc adopted.  Only the branch for |x| < 120 is needed here (angles
// are in [0, 2*pi]).  Written without FMA contraction so that host (x86-64 baseline) and device
// agree bit for bit; tests/test_sincosf.py checks it against the live libm.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define SCF_HD __host__ __device__ __forceinline__
#else
#define SCF_HD static inline
#endif

namespace orbfe_scf {

#if defined(__CUDA_ARCH__)
SCF_HD double mul(double a, double b) { return __dmul_rn(a, b); }
SCF_HD double add(double a, double b) { return __dadd_rn(a, b); }
#else
SCF_HD double mul(double a, double b) { volatile double r = a * b; return r; }   // volatile: no contraction
SCF_HD double add(double a, double b) { volatile double r = a + b; return r; }
#endif

SCF_HD uint32_t abstop12(float x) {
    uint32_t u;
#if defined(__CUDA_ARCH__)
    u = __float_as_uint(x);
#else
    memcpy(&u, &x, 4);
#endif
    return (u >> 20) & 0x7ff;
}

struct Poly { double c0, c1, c2, c3, c4, s1, s2, s3; };

// n even: sine polynomial, n odd: cosine polynomial (coefficients negated for table 1)
SCF_HD float poly(double x, double x2, bool neg, int n) {
    const double sg = neg ? -1.0 : 1.0;
    const double c0 = sg * 0x1p0, c1 = sg * -0x1.ffffffd0c621cp-2, c2 = sg * 0x1.55553e1068f19p-5,
                 c3 = sg * -0x1.6c087e89a359dp-10, c4 = sg * 0x1.99343027bf8c3p-16;
    const double s1 = -0x1.555545995a603p-3, s2 = 0x1.1107605230bc4p-7, s3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = mul(x, x2);
        const double t1 = add(s2, mul(x2, s3));
        const double x7 = mul(x3, x2);
        const double s = add(x, mul(x3, s1));
        return (float)add(s, mul(x7, t1));
    } else {
        const double x4 = mul(x2, x2);
        const double t2 = add(c3, mul(x2, c4));
        const double t1 = add(c0, mul(x2, c1));
        const double x6 = mul(x4, x2);
        const double c = add(t1, mul(x4, c2));
        return (float)add(c, mul(x6, t2));
    }
}

SCF_HD double reduce_fast(double x, int* np) {
    const double hpi_inv = 0x1.45F306DC9C883p+23, hpi = 0x1.921FB54442D18p0;
    const double r = mul(x, hpi_inv);
    const int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return add(x, -mul((double)n, hpi));
}

// valid for |y| < 120
SCF_HD float sinf_glibc(float y) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return y;
        return poly(x, mul(x, x), false, 0);
    }
    int n;
    x = reduce_fast(x, &n);
    const double s = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0;   // sign table {1,-1,-1,1}
    return poly(mul(x, s), mul(x, x), (n & 2) != 0, n);
}

SCF_HD float cosf_glibc(float y) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) return 1.0f;
        return poly(x, mul(x, x), false, 1);
    }
    int n;
    x = reduce_fast(x, &n);
    const double s = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0;
    return poly(mul(x, s), mul(x, x), (n & 2) != 0, n ^ 1);
}

// both at once (one range reduction); bit-identical to sinf_glibc / cosf_glibc
SCF_HD void sincosf_glibc(float y, float* sn, float* cs) {
    double x = y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) {
        if (abstop12(y) < abstop12(0x1p-12f)) { *sn = y; *cs = 1.0f; return; }
        const double x2 = mul(x, x);
        *sn = poly(x, x2, false, 0);
        *cs = poly(x, x2, false, 1);
        return;
    }
    int n;
    x = reduce_fast(x, &n);
    const double s = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0;
    const double xs = mul(x, s), x2 = mul(x, x);
    const bool neg = (n & 2) != 0;
    const float ps = poly(xs, x2, neg, 0), pc = poly(xs, x2, neg, 1);
    *sn = (n & 1) ? pc : ps;
    *cs = (n & 1) ? ps : pc;
}

}  // namespace orbfe_scf
