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


// Word-gather variant (every level whose four-column groups span at most 12 source bytes, i.e.
// scale factors up to ~2.6).  Per source row a thread loads the three aligned words that hold
// its (left, right) taps, two PRMTs per output pair arrange them as (L0, R0, L1, R1) and IDP.2A
// multiplies them with the 16-bit coefficient pairs: 3 loads + 8 integer instructions per row
// instead of 8 byte loads + 8 multiplies + their 64-bit addresses.  The vertical pass uses
// IMAD.HI on coefficients pre-shifted by 16.  Selectors and coefficients come from a host-built
// table (xq: three int4 per column group) and stay in registers while the thread walks PRQ rows.
// SIMPLE (every down-scaling level): no output row has row0 == row1 and a source row is the lower
// row of at most one output row, so the emission is a single predicated block per source row.
constexpr int PRQ_BATCH = 16;      // output rows per thread: large batches (throughput: the row tables amortise over 16 rows)
constexpr int PRQ_SMALL = 4;       // batches <= 4: four times the threads and a quarter of the dependent load rounds per thread
                                   // (one 640x480 frame: 7.9 -> about 4 us per level, the levels run one after the other)
constexpr int kBandsPerCta = 5;   // 128 consecutive tasks span at most 1 + ceil(127 / nq) bands; nq >= 32 is required

// (32 registers for full occupancy spill and run slower: 0.70 vs 0.59 ms)
template <bool SIMPLE, int PRQ>
__global__ void __launch_bounds__(128)
resize_level_q_kernel(const uint8_t* __restrict__ src_base, uint8_t* __restrict__ dst_base,
                      long long sframe_bytes, long long dframe_bytes, int spitch, int dpitch,
                      int dw, int dh, const int4* __restrict__ xq, const int4* __restrict__ ytab) {
    // tasks = (band of PRQ output rows, group of 4 output columns), column group fastest, flattened over the
    // CTA's 128 threads: every warp is full whatever the level width (a 2-D mapping leaves 6 .. 29 lanes of
    // the last warp of each row band busy); a CTA then touches up to kBandsPerCta bands
    const int nq = (dw + 3) >> 2, nbands = (dh + PRQ - 1) / PRQ;
    const int task0 = blockIdx.x * blockDim.x, task = task0 + threadIdx.x;
    const int band0 = task0 / nq;
    const int band = task / nq, q = task - band * nq;
    const int x4 = q * 4;
    const int y0 = band * PRQ;
    // the row tables of this block's bands go through shared memory so that a row's pixel loads do not
    // wait behind a dependent global load
    __shared__ int4 syt_all[kBandsPerCta * (PRQ + 1)];
    for (int i = threadIdx.x; i < kBandsPerCta * (PRQ + 1); i += blockDim.x) {
        const int b = i / (PRQ + 1), k = i - b * (PRQ + 1);
        const int y = (band0 + b) * PRQ + k;
        syt_all[i] = k < PRQ && band0 + b < nbands && y < dh ? __ldg(&ytab[y]) : make_int4(-1, -1, 0, 0);
    }
    __syncthreads();
    if (band >= nbands) return;
    const int4* syt = syt_all + (band - band0) * (PRQ + 1);
    const int4 t0 = __ldg(&xq[3 * q]), t1 = __ldg(&xq[3 * q + 1]), t2 = __ldg(&xq[3 * q + 2]);
    const uint32_t selA01 = t0.y, selB01 = t0.z, selA23 = t0.w, selB23 = t1.x;
    const uint32_t c0 = t1.y, c1 = t1.z, c2 = t1.w, c3 = t2.x;
    const uint8_t* src = src_base + (long long)blockIdx.z * sframe_bytes + t0.x;
    uint8_t* dst = dst_base + (long long)blockIdx.z * dframe_bytes + x4;
    struct Raw { uint32_t w0, w1, w2; };
    // the second / third word never leave the row (the last column group's taps are clamped to the row, so what the
    // repeated word holds is not selected): level 0 may alias a caller buffer with nothing mapped behind it
    const int o1 = min(4, spitch - 4 - t0.x) >> 2, o2 = min(8, spitch - 4 - t0.x) >> 2;
    auto load_raw = [&](int r) {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (long long)r * spitch);
        Raw w; w.w0 = __ldg(p); w.w1 = __ldg(p + o1); w.w2 = __ldg(p + o2);
        return w;
    };
    auto conv = [&](const Raw& w, uint32_t* h) {
        const uint32_t p01 = __byte_perm(__byte_perm(w.w0, w.w1, selA01), w.w2, selB01);
        const uint32_t p23 = __byte_perm(__byte_perm(w.w0, w.w1, selA23), w.w2, selB23);
        h[0] = __dp2a_lo(c0, p01, 0u) >> 4;
        h[1] = __dp2a_hi(c1, p01, 0u) >> 4;
        h[2] = __dp2a_lo(c2, p23, 0u) >> 4;
        h[3] = __dp2a_hi(c3, p23, 0u) >> 4;
    };
    // Source-row driven walk: the rows row0(y0) .. row1(yend-1) are visited once each, four at
    // a time with their twelve loads issued together; an output row is emitted when its lower
    // source row has just been converted (its upper row is that one or the one before).
    const int n = min(PRQ, dh - y0);
    const int rl = syt[n - 1].y;
    int yp = 0;
    int4 yt = syt[0];                                 // {row0, row1, b0, b1}
    uint32_t hp[4] = {0, 0, 0, 0};
    for (int r = yt.x; r <= rl; r += 4) {
        Raw w[4];
#pragma unroll
        for (int k = 0; k < 4; k++) w[k] = load_raw(min(r + k, rl));
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (r + k > rl) break;                    // uniform except in a warp that straddles two bands
            uint32_t hc[4];
            conv(w[k], hc);
            for (; yt.y == r + k; ) {                 // sentinel row (-1) ends the table
                const bool same = !SIMPLE && yt.x == yt.y;
                const uint32_t b0 = (uint32_t)yt.z << 16, b1 = (uint32_t)yt.w << 16;
                uint32_t v[4];
#pragma unroll
                for (int i = 0; i < 4; i++)
                    v[i] = (__umulhi(b0, same ? hc[i] : hp[i]) + __umulhi(b1, hc[i]) + 2u) >> 2;      // <= 255
                const uint32_t out = __byte_perm(__byte_perm(v[0], v[1], 0x3340), __byte_perm(v[2], v[3], 0x3340), 0x5410);
                *reinterpret_cast<uint32_t*>(dst + (long long)(y0 + yp) * dpitch) = out;   // pitch % 16 == 0
                yp++;
                yt = syt[yp];
                if (SIMPLE) break;
            }
#pragma unroll
            for (int i = 0; i < 4; i++) hp[i] = hc[i];
        }
    }
}

}  // namespace

int launch_pyramid(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0, int l1) {
    int launches = 0;
    if (l1 < 0 || l1 > g.nlevels) l1 = g.nlevels;
    for (int l = l0 < 1 ? 1 : l0; l < l1; l++) {
        const LevelGeom& P = g.lv[l - 1];
        const LevelGeom& L = g.lv[l];
        dim3 block(128);
        if (d.xq[l] && (L.w + 3) / 4 >= 32) {           // narrower levels: a CTA's tasks would span more than kBandsPerCta bands
            const bool small = g.batch <= 4;
            const int prq = small ? PRQ_SMALL : PRQ_BATCH;
            const int tasks = ((L.w + 3) / 4) * ((L.h + prq - 1) / prq);
            dim3 gridq((tasks + 127) / 128, 1, g.batch);
            auto kern = small ? (d.xqSimple[l] ? resize_level_q_kernel<true, PRQ_SMALL> : resize_level_q_kernel<false, PRQ_SMALL>)
                              : (d.xqSimple[l] ? resize_level_q_kernel<true, PRQ_BATCH> : resize_level_q_kernel<false, PRQ_BATCH>);
            kern<<<gridq, block, 0, s>>>(d.pyr + P.pyrOff + (long long)g.frame0 * P.lvBytes,
                                         d.pyr + L.pyrOff + (long long)g.frame0 * L.lvBytes,
                                         P.lvBytes, L.lvBytes, P.pitch, L.pitch, L.w, L.h, d.xq[l], d.ytab[l]);
            launches++;
            continue;
        }
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
