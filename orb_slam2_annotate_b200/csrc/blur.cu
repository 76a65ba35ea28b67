// blur.cu — K5: cv::GaussianBlur(level, Size(7,7), 2, 2, BORDER_REFLECT_101) on every pyramid
// level (reference src/ORBextractor.cc:1169-1175), cv2 4.x 8-bit path: 8.8 fixed-point taps
// [18,34,48,56,48,34,18], exact 16-bit horizontal sums, 32-bit vertical sums, one rounding
// (v + 32768) >> 16 at the end.
//
// One launch covers all levels of all frames.  A thread owns one 32-bit word column (4 pixels)
// of a 28-row band and slides down it: per input row it reads the three aligned words around
// its column (a warp reads 128 contiguous bytes), forms the eight 4-byte tap windows with
// funnel shifts and evaluates the horizontal filter with two dp4a per pixel; the last seven
// rows of horizontal sums stay in registers (the row loop is unrolled by 7 so the ring buffer
// is statically indexed) and feed the vertical filter.  No shared memory, no barriers.  The
// reflect-101 border costs nothing in the interior: warps that touch the left / right image
// edge fetch each word as a PRMT of two aligned words (selector fixed per thread), bands that
// touch the top / bottom edge reflect the row index.
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int K7 = 4;             // 7-row groups per band
constexpr int RH = 7 * K7;        // output rows per thread (28)
constexpr int BT = 128;           // threads per CTA

struct BlurTiles {
    int base[kMaxLevels + 1];   // first CTA of each level within a frame
};

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// How to fetch the 4 bytes [x, x+4) of any row with reflect-101 columns: two aligned word
// columns and a PRMT selector (the 4 reflected source columns always lie within 8 bytes).
struct WordSrc { int a, b; uint32_t sel; };
__device__ __forceinline__ WordSrc word_source(int x, int w, int last_word) {
    int pos[4], lo = 1 << 30;
#pragma unroll
    for (int k = 0; k < 4; k++) { pos[k] = reflect101(x + k, w); lo = min(lo, pos[k]); }
    WordSrc ws;
    ws.a = lo >> 2;
    ws.b = min(ws.a + 1, last_word);
    ws.sel = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) ws.sel |= (uint32_t)min(pos[k] - 4 * ws.a, 7) << (4 * k);
    if (ws.b == ws.a) ws.sel &= 0x3333u;   // single-word row tail: only the low word is real
    return ws;
}

__device__ __forceinline__ void hfilter(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t* hr) {
    const uint32_t T0 = 18u | (34u << 8) | (48u << 16) | (56u << 24);   // taps for x-3..x
    const uint32_t T1 = 48u | (34u << 8) | (18u << 16);                 // taps for x+1..x+3
    hr[0] = __dp4a(__funnelshift_r(w0, w1, 8), T0, __dp4a(__funnelshift_r(w1, w2, 8), T1, 0u));
    hr[1] = __dp4a(__funnelshift_r(w0, w1, 16), T0, __dp4a(__funnelshift_r(w1, w2, 16), T1, 0u));
    hr[2] = __dp4a(__funnelshift_r(w0, w1, 24), T0, __dp4a(__funnelshift_r(w1, w2, 24), T1, 0u));
    hr[3] = __dp4a(w1, T0, __dp4a(w2, T1, 0u));
}

template <bool EDGE>
__device__ __forceinline__ void load3(const uint8_t* row, int x, const WordSrc* ws, uint32_t& w0,
                                      uint32_t& w1, uint32_t& w2) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(row);
    if (!EDGE) {
        p += x >> 2;
        w0 = __ldg(p - 1); w1 = __ldg(p); w2 = __ldg(p + 1);
    } else {
        w0 = __byte_perm(__ldg(p + ws[0].a), __ldg(p + ws[0].b), ws[0].sel);
        w1 = __byte_perm(__ldg(p + ws[1].a), __ldg(p + ws[1].b), ws[1].sel);
        w2 = __byte_perm(__ldg(p + ws[2].a), __ldg(p + ws[2].b), ws[2].sel);
    }
}

// one thread: word column x (4 pixels), output rows [y0, y0 + RH).  INSIDE: the band and its 3-row
// halo lie inside the image, so rows advance by pointer increments and nothing is reflected.
template <bool EDGE, bool INSIDE>
__device__ __forceinline__ void blur_column(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                            int pitch, int w, int h, int x, int y0, const WordSrc* ws) {
    uint32_t hb[7][4];
    const uint8_t* rp = src + (long long)(y0 - 3) * pitch;     // only dereferenced when INSIDE
#pragma unroll
    for (int i = 0; i < 6; i++) {
        const uint8_t* row = INSIDE ? rp : src + (long long)reflect101(y0 + i - 3, h) * pitch;
        uint32_t w0, w1, w2;
        load3<EDGE>(row, x, ws, w0, w1, w2);
        hfilter(w0, w1, w2, hb[i]);
        rp += pitch;
    }
    uint8_t* op = dst + (long long)y0 * pitch + x;
#pragma unroll 1
    for (int grp = 0; grp < K7; grp++) {
#pragma unroll
        for (int j = 0; j < 7; j++) {
            const int y = y0 + grp * 7 + j;                 // output row; newest input row is y + 3
            const uint8_t* row = INSIDE ? rp : src + (long long)reflect101(y + 3, h) * pitch;
            uint32_t w0, w1, w2;
            load3<EDGE>(row, x, ws, w0, w1, w2);
            hfilter(w0, w1, w2, hb[(6 + j) % 7]);
            rp += pitch;
            // rows y-3 .. y+3 sit in slots j, j+1, ..., j+6 (mod 7)
            uint32_t acc[4];
#pragma unroll
            for (int k = 0; k < 4; k++)
                acc[k] = 18u * (hb[j % 7][k] + hb[(j + 6) % 7][k]) + 34u * (hb[(j + 1) % 7][k] + hb[(j + 5) % 7][k]) +
                         48u * (hb[(j + 2) % 7][k] + hb[(j + 4) % 7][k]) + (56u * hb[(j + 3) % 7][k] + 32768u);
            // (acc + 32768) >> 16 is byte 2 of each sum (acc < 2^24): gather the four bytes with PRMTs
            const uint32_t out = __byte_perm(__byte_perm(acc[0], acc[1], 0x0062), __byte_perm(acc[2], acc[3], 0x0062), 0x5410);
            if (INSIDE || y < h) *reinterpret_cast<uint32_t*>(op) = out;   // pitch % 16 == 0
            op += pitch;
        }
    }
}

// levels narrower than 8 px (never reached by FAST, kept exact for completeness)
__device__ void blur_tiny(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int pitch, int w,
                          int h, int x, int y0) {
    const int taps[7] = {18, 34, 48, 56, 48, 34, 18};
    for (int y = y0; y < min(y0 + RH, h); y++)
        for (int xx = x; xx < min(x + 4, w); xx++) {
            uint32_t acc = 0;
            for (int i = 0; i < 7; i++) {
                const uint8_t* row = src + (long long)reflect101(y + i - 3, h) * pitch;
                uint32_t hs = 0;
                for (int j = 0; j < 7; j++) hs += taps[j] * row[reflect101(xx + j - 3, w)];
                acc += taps[i] * hs;
            }
            dst[(long long)y * pitch + xx] = (uint8_t)((acc + 32768u) >> 16);
        }
}

__global__ void __launch_bounds__(BT, 8)
blur_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ BlurTiles t,
            const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur, int block0) {
    const int bx = (int)blockIdx.x + block0;       // block0: first block of the launched level range
    int l = 0;
    while (l + 1 < g.nlevels && bx >= t.base[l + 1]) l++;
    const LevelGeom& L = g.lv[l];
    // tasks of a level: (band, word column), word column fastest
    const int nwc = (L.w + 3) >> 2, nbands = (L.h + RH - 1) / RH;
    const int task = (bx - t.base[l]) * BT + threadIdx.x;
    const bool active = task < nwc * nbands;
    // interior word columns (all three words inside the row) first, band-major, then the few
    // edge columns: warps are then either all-interior or all-edge (one mixed warp per level)
    const int cimax = L.w >= 12 ? (L.w - 8) >> 2 : 0;           // interior columns are 1..cimax
    const int nint = cimax, nedge = nwc - nint;
    int band = 0, wc = 0;
    if (active) {
        if (task < nint * nbands) {
            band = task / nint;
            wc = 1 + task - band * nint;
        } else {
            const int te = task - nint * nbands;
            band = te / nedge;
            const int e = te - band * nedge;
            wc = e == 0 ? 0 : cimax + e;
        }
    }
    const int x = wc * 4, y0 = band * RH;
    const long long fo = (long long)(blockIdx.y + g.frame0) * L.lvBytes;
    const long long off = L.pyrOff + fo, boff = L.blurOff + fo;
    const bool edge = x < 4 || x + 8 > L.w;
    if (L.w < 8) {
        if (active) blur_tiny(pyr + off, blur + boff, L.pitch, L.w, L.h, x, y0);
        return;
    }
    const bool inside = y0 >= 3 && y0 + RH + 3 <= L.h;
    const bool warp_edge = __any_sync(0xFFFFFFFFu, active && edge);
    const bool warp_inside = __all_sync(0xFFFFFFFFu, !active || inside);
    if (!active) return;
    if (warp_edge) {
        WordSrc ws[3];
        const int last_word = (L.w - 1) >> 2;
#pragma unroll
        for (int j = 0; j < 3; j++) ws[j] = word_source(x + 4 * (j - 1), L.w, last_word);
        if (warp_inside) blur_column<true, true>(pyr + off, blur + boff, L.pitch, L.w, L.h, x, y0, ws);
        else blur_column<true, false>(pyr + off, blur + boff, L.pitch, L.w, L.h, x, y0, ws);
    } else {
        if (warp_inside) blur_column<false, true>(pyr + off, blur + boff, L.pitch, L.w, L.h, x, y0, nullptr);
        else blur_column<false, false>(pyr + off, blur + boff, L.pitch, L.w, L.h, x, y0, nullptr);
    }
}

}  // namespace

int launch_blur(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0, int l1) {
    BlurTiles t;
    int total = 0;
    for (int l = 0; l < g.nlevels; l++) {
        t.base[l] = total;
        const int tasks = ((g.lv[l].w + 3) / 4) * ((g.lv[l].h + RH - 1) / RH);
        total += (tasks + BT - 1) / BT;
    }
    t.base[g.nlevels] = total;
    if (l1 < 0 || l1 > g.nlevels) l1 = g.nlevels;
    if (l0 < 0) l0 = 0;
    const int nblk = t.base[l1] - t.base[l0];
    if (nblk <= 0) return 0;
    blur_kernel<<<dim3(nblk, g.batch), BT, 0, s>>>(g, t, d.pyr, d.blur, t.base[l0]);
    return 1;
}

}  // namespace orbfe
