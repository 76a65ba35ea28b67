// orc_primitives.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// Restatements of the OpenCV primitives the reference front end calls.  OpenCV itself is an
// un-vendored dependency of the reference (CMakeLists.txt:31-37), so the arithmetic is
// restated from its published algorithms and pinned bit-for-bit against cv2 4.13.0
// (tests/test_oracle_primitives.py, tests/golden/).  Call sites in the reference:
//   resize          src/ORBextractor.cc:1219      copyMakeBorder  :1223,:1229
//   FAST            :874,:880                      GaussianBlur    :1175
//   fastAtan2       :105                           cvRound         :82,:118,:124-125,:1208
#include "orb_oracle.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

extern "C" int orc_cv_round_f(float v) { return (int)lrintf(v); }   // default FE_TONEAREST: half-even
extern "C" int orc_cv_round_d(double v) { return (int)lrint(v); }

// cv::fastAtan2 (scalar path): 7th-order odd polynomial in float32, no FMA contraction
// (this file is compiled with -ffp-contract=off).
extern "C" float orc_fast_atan2(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// cv::resize(..., INTER_LINEAR) for CV_8UC1: 11-bit fixed-point coefficients, int32
// horizontal pass, vertical pass with the >>4 / >>16 / +2 / >>2 rounding ladder.
extern "C" void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, ptrdiff_t ss,
                                     uint8_t* dst, int dw, int dh, ptrdiff_t ds) {
    const double scale_x = 1.0 / ((double)dw / sw);
    const double scale_y = 1.0 / ((double)dh / sh);
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> ia(2 * dw), ib(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        ia[2 * dx] = (short)orc_cv_round_f((1.f - fx) * 2048.f);
        ia[2 * dx + 1] = (short)orc_cv_round_f(fx * 2048.f);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        yofs[dy] = sy;
        ib[2 * dy] = (short)orc_cv_round_f((1.f - fy) * 2048.f);
        ib[2 * dy + 1] = (short)orc_cv_round_f(fy * 2048.f);
    }
    std::vector<int> row0(dw), row1(dw);
    auto hrow = [&](int sy, std::vector<int>& out) {
        sy = sy < 0 ? 0 : (sy >= sh ? sh - 1 : sy);
        const uint8_t* S = src + (ptrdiff_t)sy * ss;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx];
            int s1 = sx + 1 < sw ? S[sx + 1] : S[sx];
            out[dx] = S[sx] * ia[2 * dx] + s1 * ia[2 * dx + 1];
        }
    };
    for (int dy = 0; dy < dh; dy++) {
        hrow(yofs[dy], row0);
        hrow(yofs[dy] + 1, row1);
        const int b0 = ib[2 * dy], b1 = ib[2 * dy + 1];
        uint8_t* D = dst + (ptrdiff_t)dy * ds;
        for (int dx = 0; dx < dw; dx++)
            D[dx] = (uint8_t)((((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2);
    }
}

static inline int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

// cv::copyMakeBorder(src, dst, b,b,b,b, BORDER_REFLECT_101); dst is (w+2b)x(h+2b).
extern "C" void orc_copy_make_border_reflect101(const uint8_t* src, int w, int h, ptrdiff_t ss,
                                                uint8_t* dst, ptrdiff_t ds, int border) {
    for (int y = -border; y < h + border; y++) {
        const uint8_t* S = src + (ptrdiff_t)reflect101(y, h) * ss;
        uint8_t* D = dst + (ptrdiff_t)(y + border) * ds;
        for (int x = -border; x < w + border; x++) D[x + border] = S[reflect101(x, w)];
    }
}

// FAST 9/16 corner score: largest threshold for which the pixel is still a corner, minus 1
// (min/max ladder over the 16 arcs of 9 contiguous ring pixels).
static const int kRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},   {3, 0},  {3, -1},
                                 {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                 {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

static int fast_corner_score(const uint8_t* p, const ptrdiff_t* off, int threshold) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 25; k++) d[k] = v - p[off[k]];
    int a0 = threshold;
    for (int k = 0; k < 16; k += 2) {
        int a = std::min(d[k + 1], d[k + 2]);
        for (int i = 3; i <= 8; i++) a = std::min(a, d[k + i]);
        a0 = std::max(a0, std::min(a, d[k]));
        a0 = std::max(a0, std::min(a, d[k + 9]));
    }
    int b0 = -a0;
    for (int k = 0; k < 16; k += 2) {
        int b = std::max(d[k + 1], d[k + 2]);
        for (int i = 3; i <= 8; i++) b = std::max(b, d[k + i]);
        b0 = std::min(b0, std::max(b, d[k]));
        b0 = std::min(b0, std::max(b, d[k + 9]));
    }
    return -b0 - 1;
}

extern "C" int orc_fast9_16(const uint8_t* img, int w, int h, ptrdiff_t stride, int threshold,
                            int nonmax, int* out, int cap) {
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return 0;
    ptrdiff_t off[25];
    for (int k = 0; k < 25; k++) off[k] = kRing[k % 16][0] + kRing[k % 16][1] * stride;
    // score map for the inset zone; 0 = not a corner
    std::vector<uint8_t> score((size_t)w * h, 0);
    std::vector<uint8_t> is_corner((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++) {
        const uint8_t* row = img + y * stride;
        for (int x = 3; x < w - 3; x++) {
            const uint8_t* p = row + x;
            const int v = p[0];
            // high-speed rejection: a 9-arc contains one pixel of every opposing pair (k, k+8)
            const int lo = v - threshold, hi = v + threshold;
            int m = 3;   // bit0: darker arc still possible, bit1: brighter arc still possible
            for (int k = 0; k < 8 && m; k++) {
                const int q0 = p[off[k]], q1 = p[off[k + 8]];
                m &= ((q0 < lo || q1 < lo) ? 1 : 0) | ((q0 > hi || q1 > hi) ? 2 : 0);
            }
            if (!m) continue;
            int run_d = 0, run_b = 0;
            bool corner = false;
            for (int k = 0; k < 25 && !corner; k++) {
                int q = p[off[k]];
                if (q < lo) { if (++run_d > 8) corner = true; } else run_d = 0;
                if (q > hi) { if (++run_b > 8) corner = true; } else run_b = 0;
            }
            if (corner) {
                is_corner[(size_t)y * w + x] = 1;
                score[(size_t)y * w + x] = (uint8_t)fast_corner_score(p, off, threshold);
            }
        }
    }
    int n = 0;
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            size_t i = (size_t)y * w + x;
            if (!is_corner[i]) continue;
            int s = score[i];
            if (nonmax) {
                // neighbours outside the 3-px inset are zero in OpenCV's row buffers; the
                // score map here is zero there too.
                if (!(s > score[i - 1] && s > score[i + 1] && s > score[i - w - 1] &&
                      s > score[i - w] && s > score[i - w + 1] && s > score[i + w - 1] &&
                      s > score[i + w] && s > score[i + w + 1]))
                    continue;
            }
            if (n < cap) {
                out[3 * n] = x; out[3 * n + 1] = y;
                out[3 * n + 2] = nonmax ? s : 0;
            }
            n++;
        }
    return n;
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 in cv2 4.x:
// 8.8 fixed-point taps, exact 16-bit horizontal pass, 32-bit vertical pass, round at the end.
extern "C" void orc_gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, ptrdiff_t ss,
                                         uint8_t* dst, ptrdiff_t ds) {
    static const int taps[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> hbuf((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (ptrdiff_t)y * ss;
        for (int x = 0; x < w; x++) {
            int acc = 0;
            for (int k = 0; k < 7; k++) acc += taps[k] * S[reflect101(x + k - 3, w)];
            hbuf[(size_t)y * w + x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; y++) {
        const uint16_t* rows[7];
        for (int k = 0; k < 7; k++) rows[k] = &hbuf[(size_t)reflect101(y + k - 3, h) * w];
        uint8_t* D = dst + (ptrdiff_t)y * ds;
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)taps[k] * rows[k][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}
