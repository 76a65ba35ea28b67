// blur.cu — K5: cv::GaussianBlur(level, Size(7,7), 2, 2, BORDER_REFLECT_101) on every pyramid
// level (reference src/ORBextractor.cc:1169-1175), cv2 4.x 8-bit path: 8.8 fixed-point taps
// [18,34,48,56,48,34,18], exact 16-bit horizontal sums, 32-bit vertical sums, one rounding
// (v + 32768) >> 16 at the end.
//
// One launch covers all levels of all frames.  A thread owns one 32-bit word column (4 pixels)
// of a 16-row band and slides down it: per input row it reads the three aligned words around
// its column (a warp reads 128 contiguous bytes), forms the eight 4-byte tap windows with
// funnel shifts and evaluates the horizontal filter with two dp4a per pixel; the last seven
// rows of horizontal sums stay in registers (the loop is fully unrolled so the ring buffer is
// statically indexed) and feed the vertical filter.  No shared memory, no barriers; the
// reflect-101 border is resolved by index on the few edge lanes / rows.
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int RH = 16;            // output rows per warp band
constexpr int BANDS = 4;          // warps per CTA, stacked vertically
constexpr int TWPX = 128;         // tile width in pixels (32 lanes x 4)

struct BlurTiles {
    int base[kMaxLevels + 1];   // first tile of each level within a frame
    int tilesX[kMaxLevels];
};

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

__device__ __forceinline__ uint32_t edge_word(const uint8_t* row, int x, int w) {
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) v |= (uint32_t)row[reflect101(x + k, w)] << (8 * k);
    return v;
}

__global__ void __launch_bounds__(BANDS * 32)
blur_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ BlurTiles t,
            const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur) {
    int l = 0;
    while (l + 1 < g.nlevels && (int)blockIdx.x >= t.base[l + 1]) l++;
    const LevelGeom& L = g.lv[l];
    const int tile = blockIdx.x - t.base[l];
    const int tx = tile % t.tilesX[l], ty = tile / t.tilesX[l];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int x = tx * TWPX + lane * 4;
    const int y0 = (ty * BANDS + wid) * RH;
    if (x >= L.w || y0 >= L.h) return;
    const long long off = L.pyrOff + (long long)blockIdx.y * L.lvBytes;
    const uint8_t* src = pyr + off;
    uint8_t* dst = blur + off;
    const bool interior = x >= 4 && x + 8 <= L.w;
    const uint32_t T0 = 18u | (34u << 8) | (48u << 16) | (56u << 24);   // taps for x-3..x
    const uint32_t T1 = 48u | (34u << 8) | (18u << 16);                 // taps for x+1..x+3

    uint32_t h[7][4];
#pragma unroll
    for (int i = 0; i < RH + 6; i++) {
        const int ry = reflect101(y0 + i - 3, L.h);
        const uint8_t* row = src + (long long)ry * L.pitch;
        uint32_t w0, w1, w2;
        if (interior) {
            const uint32_t* p = reinterpret_cast<const uint32_t*>(row + x);
            w0 = __ldg(p - 1); w1 = __ldg(p); w2 = __ldg(p + 1);
        } else {
            w0 = edge_word(row, x - 4, L.w); w1 = edge_word(row, x, L.w); w2 = edge_word(row, x + 4, L.w);
        }
        uint32_t* hr = h[i % 7];
        hr[0] = __dp4a(__funnelshift_r(w0, w1, 8), T0, __dp4a(__funnelshift_r(w1, w2, 8), T1, 0u));
        hr[1] = __dp4a(__funnelshift_r(w0, w1, 16), T0, __dp4a(__funnelshift_r(w1, w2, 16), T1, 0u));
        hr[2] = __dp4a(__funnelshift_r(w0, w1, 24), T0, __dp4a(__funnelshift_r(w1, w2, 24), T1, 0u));
        hr[3] = __dp4a(w1, T0, __dp4a(w2, T1, 0u));
        if (i >= 6) {
            const int y = y0 + i - 6;
            if (y < L.h) {
                uint32_t out = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t acc = 18u * (h[(i - 6) % 7][k] + h[i % 7][k]) +
                                         34u * (h[(i - 5) % 7][k] + h[(i - 1) % 7][k]) +
                                         48u * (h[(i - 4) % 7][k] + h[(i - 2) % 7][k]) +
                                         56u * h[(i - 3) % 7][k];
                    out |= ((acc + 32768u) >> 16) << (8 * k);
                }
                *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch + x) = out;   // pitch % 16 == 0
            }
        }
    }
}

}  // namespace

int launch_blur(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s) {
    BlurTiles t;
    int total = 0;
    for (int l = 0; l < g.nlevels; l++) {
        t.base[l] = total;
        t.tilesX[l] = (g.lv[l].w + TWPX - 1) / TWPX;
        total += t.tilesX[l] * ((g.lv[l].h + BANDS * RH - 1) / (BANDS * RH));
    }
    t.base[g.nlevels] = total;
    blur_kernel<<<dim3(total, g.batch), BANDS * 32, 0, s>>>(g, t, d.pyr, d.blur);
    return 1;
}

}  // namespace orbfe
