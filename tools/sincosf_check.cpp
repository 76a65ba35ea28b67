// sincosf_check — compares orb_slam2_annotate_b200/csrc/sincosf_glibc.h with the live libm
// sinf/cosf over float inputs in [0, hi] with a bit-pattern stride. Prints mismatch counts.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../orb_slam2_annotate_b200/csrc/sincosf_glibc.h"
int main(int argc, char** argv) {
    const unsigned stride = argc > 1 ? (unsigned)atoi(argv[1]) : 1;
    const float hi = 6.2831855f;
    unsigned uhi; memcpy(&uhi, &hi, 4);
    unsigned long long n = 0, bad_s = 0, bad_c = 0;
    for (unsigned u = 0; u <= uhi; u += stride) {
        float x; memcpy(&x, &u, 4);
        volatile float vx = x;
        const float s = sinf(vx), c = cosf(vx);
        const float s2 = orbfe_scf::sinf_glibc(x), c2 = orbfe_scf::cosf_glibc(x);
        float s3, c3;
        orbfe_scf::sincosf_glibc(x, &s3, &c3);
        if (memcmp(&s3, &s2, 4) || memcmp(&c3, &c2, 4)) { bad_s++; bad_c++; }
        if (memcmp(&s, &s2, 4)) { if (bad_s < 5) printf("sin %a: libm %a mine %a\n", x, s, s2); bad_s++; }
        if (memcmp(&c, &c2, 4)) { if (bad_c < 5) printf("cos %a: libm %a mine %a\n", x, c, c2); bad_c++; }
        n++;
    }
    printf("{\"n\": %llu, \"sin_mismatch\": %llu, \"cos_mismatch\": %llu}\n", n, bad_s, bad_c);
    return (bad_s || bad_c) ? 1 : 0;
}
