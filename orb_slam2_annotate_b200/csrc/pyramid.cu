// pyramid.cu — K1: ComputePyramid (reference src/ORBextractor.cc:1203-1234).
//
// Level l is cv::resize(level l-1, INTER_LINEAR) in OpenCV's 8-bit fixed-point arithmetic:
// 11-bit horizontal/vertical coefficients, int32 horizontal sums, and the vertical
// ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2) >> 2 rounding ladder.  The
// coefficient tables (source offsets and short coefficients) are computed once on the host in
// the same float/double steps OpenCV uses and kept in HBM; a thread produces 4 neighbouring
// output pixels per row and stores them as one 32-bit word, a warp therefore writes 128
// contiguous bytes, and walks down 32 output rows reusing horizontal sums between rows.  The 19-px reflect-101 frame the reference adds around every level is not
// materialised: nothing on the extractor path reads it.
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

constexpr int PR = 32;   // output rows per thread

// A thread owns 4 neighbouring output columns and walks PR output rows.  Its four source
// offsets / coefficient pairs live in registers for the whole walk; the horizontal sums of a
// source row are computed once and reused when the next output row needs the same source row
// (at scale 1.2 the lower row of output y is the upper row of output y+1 five times out of six).
__global__ void __launch_bounds__(128)
resize_level_kernel(const uint8_t* __restrict__ src_base, uint8_t* __restrict__ dst_base,
                    long long sframe_bytes, long long dframe_bytes, int spitch, int dpitch,
                    int dw, int dh, const int2* __restrict__ xtab, const int4* __restrict__ ytab) {
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y0 = blockIdx.y * PR;
    if (x4 >= dw) return;
    const uint8_t* src = src_base + (long long)blockIdx.z * sframe_bytes;
    uint8_t* dst = dst_base + (long long)blockIdx.z * dframe_bytes;
    int sx[4], sx1[4], a0[4], a1[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int2 xt = __ldg(&xtab[min(x4 + i, dw - 1)]);
        sx[i] = xt.x & 0xFFFF; sx1[i] = xt.x >> 16;
        a0[i] = (short)(xt.y & 0xFFFF); a1[i] = (short)(xt.y >> 16);
    }
    auto hrow = [&](int r, int* h) {
        const uint8_t* p = src + (long long)r * spitch;
#pragma unroll
        for (int i = 0; i < 4; i++) h[i] = ((int)p[sx[i]] * a0[i] + (int)p[sx1[i]] * a1[i]) >> 4;
    };
    int prev_row = -1, hprev[4] = {0, 0, 0, 0};
    const int yend = min(y0 + PR, dh);
    for (int y = y0; y < yend; y++) {
        const int4 yt = __ldg(&ytab[y]);          // {row0, row1, b0, b1}
        int h0[4], h1[4];
        if (yt.x == prev_row) {
#pragma unroll
            for (int i = 0; i < 4; i++) h0[i] = hprev[i];
        } else {
            hrow(yt.x, h0);
        }
        if (yt.y == yt.x) {
#pragma unroll
            for (int i = 0; i < 4; i++) h1[i] = h0[i];
        } else {
            hrow(yt.y, h1);
        }
        uint32_t out = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int v = (((yt.z * h0[i]) >> 16) + ((yt.w * h1[i]) >> 16) + 2) >> 2;
            out |= (uint32_t)(v & 0xFF) << (8 * i);
            hprev[i] = h1[i];
        }
        prev_row = yt.y;
        *reinterpret_cast<uint32_t*>(dst + (long long)y * dpitch + x4) = out;   // pitch % 16 == 0
    }
}

}  // namespace

int launch_pyramid(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s) {
    int launches = 0;
    for (int l = 1; l < g.nlevels; l++) {
        const LevelGeom& P = g.lv[l - 1];
        const LevelGeom& L = g.lv[l];
        dim3 block(128);
        dim3 grid((L.w + 511) / 512, (L.h + PR - 1) / PR, g.batch);
        resize_level_kernel<<<grid, block, 0, s>>>(d.pyr + P.pyrOff + (long long)g.frame0 * P.lvBytes,
                                                   d.pyr + L.pyrOff + (long long)g.frame0 * L.lvBytes,
                                                   P.lvBytes, L.lvBytes, P.pitch, L.pitch, L.w, L.h,
                                                   d.xtab[l], d.ytab[l]);
        launches++;
    }
    return launches;
}

}  // namespace orbfe
