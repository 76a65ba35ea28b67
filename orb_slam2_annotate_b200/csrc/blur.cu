// blur.cu — K5: cv::GaussianBlur(level, Size(7,7), 2, 2, BORDER_REFLECT_101) on every pyramid
// level (reference src/ORBextractor.cc:1169-1175), cv2 4.x 8-bit path: 8.8 fixed-point taps
// [18,34,48,56,48,34,18], exact 16-bit horizontal sums, 32-bit vertical sums, one rounding
// (v + 32768) >> 16 at the end.
//
// One launch covers all levels of all frames: blockIdx.x walks the 128x16 output tiles of a
// frame's levels back to back, blockIdx.y is the frame.  A CTA stages the tile plus its 3-px
// halo in shared memory (reflect-101 resolved while loading), runs the horizontal pass into a
// 16-bit shared buffer and the vertical pass out of it; output rows leave as 32-bit words.
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int TW = 128, TH = 16, R = 3;
constexpr int IN_W = TW + 2 * R + 2;   // 136: keeps rows 8-byte aligned

struct BlurTiles {
    int base[kMaxLevels + 1];   // first tile of each level within a frame
    int tilesX[kMaxLevels];
};

__device__ __forceinline__ int reflect101(int p, int len) {
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

__global__ void __launch_bounds__(256)
blur_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ BlurTiles t,
            const uint8_t* __restrict__ pyr, uint8_t* __restrict__ blur) {
    __shared__ uint8_t in[TH + 2 * R][IN_W];
    __shared__ uint16_t hb[TH + 2 * R][TW];
    int l = 0;
    while (l + 1 < g.nlevels && (int)blockIdx.x >= t.base[l + 1]) l++;
    const LevelGeom& L = g.lv[l];
    const int tile = blockIdx.x - t.base[l];
    const int tx = tile % t.tilesX[l], ty = tile / t.tilesX[l];
    const int x0 = tx * TW, y0 = ty * TH;
    const long long off = L.pyrOff + (long long)blockIdx.y * L.lvBytes;
    const uint8_t* src = pyr + off;
    uint8_t* dst = blur + off;
    const int tid = threadIdx.x;

    for (int i = tid; i < (TH + 2 * R) * (TW + 2 * R); i += 256) {
        const int r = i / (TW + 2 * R), c = i - r * (TW + 2 * R);
        const int sy = reflect101(y0 + r - R, L.h);
        const int sx = reflect101(min(x0 + c - R, L.w + R - 1), L.w);
        in[r][c] = src[(long long)sy * L.pitch + sx];
    }
    __syncthreads();
    for (int i = tid; i < (TH + 2 * R) * TW; i += 256) {
        const int r = i / TW, c = i - r * TW;
        const uint8_t* p = &in[r][c];
        const int acc = 18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3];
        hb[r][c] = (uint16_t)acc;
    }
    __syncthreads();
    // 128x16 outputs, 4 per thread as one 32-bit store: 32 words per row, 16 rows = 512 words
    for (int i = tid; i < TH * (TW / 4); i += 256) {
        const int r = i / (TW / 4), c4 = (i - r * (TW / 4)) * 4;
        const int y = y0 + r, x = x0 + c4;
        if (y >= L.h || x >= L.w) continue;
        uint32_t out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int c = c4 + k;
            const uint32_t acc = 18u * (hb[r][c] + hb[r + 6][c]) + 34u * (hb[r + 1][c] + hb[r + 5][c]) +
                                 48u * (hb[r + 2][c] + hb[r + 4][c]) + 56u * hb[r + 3][c];
            out |= ((acc + 32768u) >> 16) << (8 * k);
        }
        *reinterpret_cast<uint32_t*>(dst + (long long)y * L.pitch + x) = out;   // pitch%16==0: pad is ours
    }
}

}  // namespace

int launch_blur(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s) {
    BlurTiles t;
    int total = 0;
    for (int l = 0; l < g.nlevels; l++) {
        t.base[l] = total;
        t.tilesX[l] = (g.lv[l].w + TW - 1) / TW;
        total += t.tilesX[l] * ((g.lv[l].h + TH - 1) / TH);
    }
    t.base[g.nlevels] = total;
    blur_kernel<<<dim3(total, g.batch), 256, 0, s>>>(g, t, d.pyr, d.blur);
    return 1;
}

}  // namespace orbfe
