// pyramid.cu — K1: ComputePyramid (reference src/ORBextractor.cc:1203-1234).
//
// Level l is cv::resize(level l-1, INTER_LINEAR) in OpenCV's 8-bit fixed-point arithmetic:
// 11-bit horizontal/vertical coefficients, int32 horizontal sums, and the vertical
// ((b0*(h0>>4))>>16) + ((b1*(h1>>4))>>16) + 2) >> 2 rounding ladder.  The
// coefficient tables (source offsets and short coefficients) are computed once on the host in
// the same float/double steps OpenCV uses and kept in HBM; a thread produces 4 neighbouring
// output pixels and stores them as one 32-bit word, a warp therefore writes 128 contiguous
// bytes.  The 19-px reflect-101 frame the reference adds around every level is not
// materialised: nothing on the extractor path reads it.
#include "orbfe_internal.cuh"

namespace orbfe {

namespace {

__global__ void __launch_bounds__(256)
resize_level_kernel(const uint8_t* __restrict__ src_base, uint8_t* __restrict__ dst_base,
                    long long sframe_bytes, long long dframe_bytes, int spitch, int dpitch,
                    int dw, int dh,
                    const int2* __restrict__ xtab, const int4* __restrict__ ytab) {
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x4 >= dw || y >= dh) return;
    const uint8_t* src = src_base + (long long)blockIdx.z * sframe_bytes;
    uint8_t* dst = dst_base + (long long)blockIdx.z * dframe_bytes;
    const int4 yt = __ldg(&ytab[y]);
    const uint8_t* r0 = src + (long long)yt.x * spitch;
    const uint8_t* r1 = src + (long long)yt.y * spitch;
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = x4 + i;
        if (x < dw) {
            const int2 xt = __ldg(&xtab[x]);
            const int sx = xt.x & 0xFFFF, sx1 = xt.x >> 16;
            const int a0 = (short)(xt.y & 0xFFFF), a1 = (short)(xt.y >> 16);
            const int h0 = (int)r0[sx] * a0 + (int)r0[sx1] * a1;
            const int h1 = (int)r1[sx] * a0 + (int)r1[sx1] * a1;
            const int v = (((yt.z * (h0 >> 4)) >> 16) + ((yt.w * (h1 >> 4)) >> 16) + 2) >> 2;
            out |= (uint32_t)(v & 0xFF) << (8 * i);
        }
    }
    *reinterpret_cast<uint32_t*>(dst + (long long)y * dpitch + x4) = out;   // pitch % 16 == 0
}

}  // namespace

int launch_pyramid(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s) {
    int launches = 0;
    for (int l = 1; l < g.nlevels; l++) {
        const LevelGeom& P = g.lv[l - 1];
        const LevelGeom& L = g.lv[l];
        dim3 block(32, 8);
        dim3 grid((L.w + 127) / 128, (L.h + 7) / 8, g.batch);
        resize_level_kernel<<<grid, block, 0, s>>>(d.pyr + P.pyrOff, d.pyr + L.pyrOff,
                                                   P.lvBytes, L.lvBytes, P.pitch, L.pitch, L.w, L.h,
                                                   d.xtab[l], d.ytab[l]);
        launches++;
    }
    return launches;
}

}  // namespace orbfe
