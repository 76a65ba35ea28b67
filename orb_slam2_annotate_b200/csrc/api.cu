// api.cu — the C ABI of liborbfe (include/orbfe.h): extractor handle, geometry and table
// construction (the reference constructor, src/ORBextractor.cc:415-486), host<->device
// plumbing and the kernel schedule of ORBextractor::operator() (:1119-1197).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <cstdlib>
#include <vector>

#include "api_common.cuh"
#include "frame_internal.cuh"
#include "orbfe_internal.cuh"
#include <cuda.h>

namespace orbfe {

thread_local std::string g_err;

int fail(int code, const char* what, cudaError_t ce) {
    g_err = what;
    if (ce != cudaSuccess) { g_err += ": "; g_err += cudaGetErrorString(ce); }
    return code;
}

static inline int cv_round(float v) { return (int)lrintf(v); }   // round-half-even

}  // namespace orbfe

using namespace orbfe;

struct orbfe_extractor {
    int nfeatures, nlevels, iniTh, minTh, device;
    double scaleFactor;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> per_level;
    int umax[16];
    cudaStream_t stream = nullptr;
    // host-batch pipeline: H2D of chunk c+1 and D2H of chunk c-1 overlap the kernels of chunk c
    cudaStream_t h2dStream = nullptr, d2hStream = nullptr;
    // the Gaussian blur only depends on the pyramid: it runs on a side stream beside FAST + quadtree
    cudaStream_t sideStream = nullptr;
    cudaEvent_t evPyr = nullptr, evBlur = nullptr;
    // second kernel lane of the chunked host pipeline (chunks alternate so one chunk's tail overlaps the next one's head)
    cudaStream_t stream2 = nullptr, sideStream2 = nullptr;
    cudaEvent_t evPyr2 = nullptr, evBlur2 = nullptr;
    std::vector<cudaEvent_t> evUp, evDone;
    // single-frame schedule: per pyramid level one branch for FAST + quadtree and one for the blur, forked off the
    // pyramid chain as soon as the level exists and joined in front of the descriptor kernel
    PinnedArray stageIn, stageOut;               // single-frame path: page-locked staging for pageable caller buffers
    PinnedArray stagePyr;                        // orbfe_pyramid_host
    std::vector<cudaStream_t> lvStream;          // [2 * nlevels]
    std::vector<cudaEvent_t> lvReady, lvDone;    // [nlevels], [2 * nlevels]
    // geometry of the current image size / batch capacity
    int geomW = 0, geomH = 0, geomBatchCap = 0;
    FrameSetGeom g;
    DevBufs d;
    DevArray<uint8_t> pyr, blur, desc;
    DevArray<int> cellc, selc, counts, umax_d;
    DevArray<uint32_t> cellk, sel, qt_xy, qt_node;
    DevArray<orbfe_keypoint> kps;
    DevArray<int2> xtab;
    DevArray<int4> ytab;
    DevArray<int4> xq;
    DevArray<int4> celltab;
    std::vector<CUtensorMap> tmaps;   // {level, blurred level} descriptors, passed to the kernel by value
    std::vector<CUtensorMap> fastMaps; // level descriptors with the FAST window box
    std::vector<CUtensorMap> fastMaps48; // the same with a 48-byte box (windows that fit skip the re-layout)
    int lastBatch = 0, lastLaunches = 0;
    int level0MapFrames = 0;       // frames the level-0 tensor maps cover (the batch capacity, or an alias's frame count)
    // Frame epilogue (orbfe_extractor_set_camera): mvKeysUn + mGrid of every frame of a batch
    bool haveCam = false;
    CamGeom camGeom{};
    DevArray<orbfe_keypoint> kpsUn;
    DevArray<int> gridEnd;
    DevArray<unsigned short> gridItems;
    // single-frame latency path: the kernel schedule of a small batch captured once as a CUDA graph and replayed while
    // every kernel argument (geometry, device pointers, TMA descriptors, camera) stays what it was at capture time
    cudaGraphExec_t graphExec = nullptr;
    std::vector<uint8_t> graphKey;
    int graphLaunches = 0;
    bool graphRefused = false;     // a capture failed once: stay eager for the life of the handle
    // optional per-stage timing (events on the handle's stream)
    bool profiling = false;
    cudaEvent_t ev[ORBFE_NUM_STAGES + 1] = {};
    std::vector<int> h_counts;
    // orbfe_extractor_device_pyramid
    std::vector<const uint8_t*> pv_data;
    std::vector<int32_t> pv_w, pv_h;
    std::vector<ptrdiff_t> pv_pitch, pv_fstride;
};

namespace orbfe {

// Scale pyramid, per-level feature budget and the circular-patch row extents, exactly as the
// reference constructor computes them (ORBextractor.cc:415-486).
static void build_tables(orbfe_extractor* e, int nfeatures, float scaleFactor, int nlevels) {
    e->scaleFactor = scaleFactor;                         // double member set from a float
    e->scale.assign(nlevels, 1.0f); e->sigma2.assign(nlevels, 1.0f);
    e->inv_scale.assign(nlevels, 1.0f); e->inv_sigma2.assign(nlevels, 1.0f);
    for (int i = 1; i < nlevels; i++) {
        e->scale[i] = (float)(e->scale[i - 1] * e->scaleFactor);
        e->sigma2[i] = e->scale[i] * e->scale[i];
    }
    for (int i = 0; i < nlevels; i++) {
        e->inv_scale[i] = 1.0f / e->scale[i];
        e->inv_sigma2[i] = 1.0f / e->sigma2[i];
    }
    e->per_level.assign(nlevels, 0);
    const float factor = (float)(1.0f / e->scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int l = 0; l < nlevels - 1; l++) {
        e->per_level[l] = cv_round(nDesired);
        sum += e->per_level[l];
        nDesired *= factor;
    }
    e->per_level[nlevels - 1] = std::max(nfeatures - sum, 0);
    const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    std::memset(e->umax, 0, sizeof(e->umax));
    for (int v = 0; v <= vmax; ++v) e->umax[v] = (int)lrint(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (e->umax[v0] == e->umax[v0 + 1]) ++v0;
        e->umax[v] = v0;
        ++v0;
    }
}

static int ensure_frame_bufs(orbfe_extractor* e, int batchCap, int outCap) {
    if (outCap > 65535) return fail(ORBFE_ERR_CAPACITY, "frame grid items are 16-bit: more than 65535 keypoint slots per frame");
    int rc;
    if ((rc = e->kpsUn.ensure((size_t)batchCap * outCap))) return rc;
    if ((rc = e->gridEnd.ensure((size_t)batchCap * ORBFE_GRID_CELLS))) return rc;
    if ((rc = e->gridItems.ensure((size_t)batchCap * outCap))) return rc;
    return 0;
}

// the orientation kernel walks patch columns: needs |u| <= umax[|v|]  <=>  |v| <= umax[|u|]
static bool umax_symmetric(const int* umax) {
    for (int u = 0; u <= kHalfPatch; u++)
        for (int v = 0; v <= kHalfPatch; v++)
            if ((u <= umax[v]) != (v <= umax[u])) return false;
    return true;
}

// cv::resize INTER_LINEAR coefficient tables for one level
static void resize_tables(int sw, int sh, int dw, int dh, std::vector<int2>& xt, std::vector<int4>& yt) {
    const double scale_x = 1.0 / ((double)dw / sw), scale_y = 1.0 / ((double)dh / sh);
    xt.resize(dw); yt.resize(dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        const int sx1 = sx + 1 < sw ? sx + 1 : sx;
        const int a0 = (short)cv_round((1.f - fx) * 2048.f), a1 = (short)cv_round(fx * 2048.f);
        xt[dx] = make_int2(sx | (sx1 << 16), (a0 & 0xFFFF) | (a1 << 16));
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        const int b0 = (short)cv_round((1.f - fy) * 2048.f), b1 = (short)cv_round(fy * 2048.f);
        const int r0 = sy < 0 ? 0 : (sy >= sh ? sh - 1 : sy);
        const int r1 = sy + 1 < 0 ? 0 : (sy + 1 >= sh ? sh - 1 : sy + 1);
        yt[dy] = make_int4(r0, r1, b0, b1);
    }
}

// Column-group table of the word-gather resize kernel: for output columns 4q..4q+3 the aligned
// source byte offset of the first tap, PRMT selectors that pull (L, R) tap pairs out of the
// three words loaded there, and the coefficient pairs.  False when a group spans more than 12
// source bytes (then the level uses the byte-gather kernel).
static bool resize_group_table(const std::vector<int2>& xt, int dw, std::vector<int4>& xq) {
    const int nq = (dw + 3) / 4;
    xq.assign((size_t)nq * 3, make_int4(0, 0, 0, 0));
    for (int q = 0; q < nq; q++) {
        int oL[4], oR[4]; unsigned c[4];
        const int base = (xt[4 * q].x & 0xFFFF) & ~3;
        for (int i = 0; i < 4; i++) {
            const int2 t = xt[std::min(4 * q + i, dw - 1)];
            oL[i] = (t.x & 0xFFFF) - base; oR[i] = (t.x >> 16) - base; c[i] = (unsigned)t.y;
            if (oL[i] < 0 || oR[i] < 0 || oL[i] >= 12 || oR[i] >= 12) return false;
            if ((short)(t.y & 0xFFFF) < 0 || (t.y >> 16) < 0) return false;   // IDP.2A operands are unsigned here
        }
        auto sel = [&](int i, int j, unsigned& selA, unsigned& selB) {
            const int o[4] = {oL[i], oR[i], oL[j], oR[j]};
            selA = selB = 0;
            for (int k = 0; k < 4; k++) {
                selA |= (unsigned)(o[k] < 8 ? o[k] : 0) << (4 * k);
                selB |= (unsigned)(o[k] < 8 ? k : 4 + o[k] - 8) << (4 * k);
            }
        };
        unsigned a01, b01, a23, b23;
        sel(0, 1, a01, b01); sel(2, 3, a23, b23);
        xq[3 * q] = make_int4(base, (int)a01, (int)b01, (int)a23);
        xq[3 * q + 1] = make_int4((int)b23, (int)c[0], (int)c[1], (int)c[2]);
        xq[3 * q + 2] = make_int4((int)c[3], 0, 0, 0);
    }
    return true;
}

// TMA descriptor of one level buffer seen as a [frames * h][pitch] byte tensor with a boxW x boxH
// box (the orientation / descriptor kernel fetches its keypoint windows with it).  The encoder is
// a driver entry point; it is resolved through the runtime so that liborbfe.so does not link libcuda.
static int encode_level_map(CUtensorMap* m, void* base, int pitch, long long rows, int boxW, int boxH) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t ce = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (ce != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn)
            return fail(ORBFE_ERR_CUDA, "cudaGetDriverEntryPoint(cuTensorMapEncodeTiled)", ce);
        encode = (encode_fn)fn;
    }
    const cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch};
    const cuuint32_t box[2] = {(cuuint32_t)boxW, (cuuint32_t)boxH};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ORBFE_ERR_CUDA, "cuTensorMapEncodeTiled failed");
    return 0;
}

static inline long long align_up(long long v, long long a) { return (v + a - 1) / a * a; }

// Per-level geometry for a w x h input at batch capacity B: level sizes (:1208), FAST cell
// grid (:823-842), quadtree roots (:570-572), slot capacities and buffer offsets.
static int build_geometry(orbfe_extractor* e, int w, int h, int batchCap, FrameSetGeom& g) {
    std::memset(&g, 0, sizeof(g));
    g.nlevels = e->nlevels; g.iniTh = e->iniTh; g.minTh = e->minTh;
    long long pyrOff = 0, slotOff = 0;
    int cellBase = 0, selBase = 0;
    for (int l = 0; l < e->nlevels; l++) {
        LevelGeom& L = g.lv[l];
        L.w = cv_round((float)w * e->inv_scale[l]);
        L.h = cv_round((float)h * e->inv_scale[l]);
        if (L.w < 1 || L.h < 1) return fail(ORBFE_ERR_ARG, "image too small for the number of pyramid levels");
        L.pitch = (int)align_up(L.w, 16);
        L.lvBytes = (long long)L.pitch * L.h;
        L.pyrOff = pyrOff;
        L.blurOff = pyrOff;
        pyrOff = align_up(pyrOff + L.lvBytes * batchCap, 256);
        L.minBX = kEdgeThreshold - 3; L.minBY = L.minBX;
        L.maxBX = L.w - kEdgeThreshold + 3; L.maxBY = L.h - kEdgeThreshold + 3;
        const float width = (float)(L.maxBX - L.minBX), height = (float)(L.maxBY - L.minBY);
        L.nCols = (int)(width / kCellW); L.nRows = (int)(height / kCellW);
        if (L.nCols > 0 && L.nRows > 0) {
            L.wCell = (int)std::ceil(width / L.nCols);
            L.hCell = (int)std::ceil(height / L.nRows);
        } else {
            L.nCols = L.nRows = 0; L.wCell = L.hCell = 1;
        }
        if (L.wCell + 6 > kFastMaxWin || L.hCell + 6 > kFastMaxWin)
            return fail(ORBFE_ERR_ARG, "FAST cell window exceeds the kernel's shared-memory tile");
        L.cellBase = cellBase;
        cellBase += L.nCols * L.nRows;
        // strict 3x3 maxima cannot be 8-neighbours: at most ceil(zw/2)*ceil(zh/2) per zone
        L.cellCap = ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2);
        L.cellSlotOff = slotOff;
        if ((long long)L.nCols * L.nRows * L.cellCap > (1 << 20))
            return fail(ORBFE_ERR_ARG, "level too large: more than 2^20 candidate corners");
        slotOff += (long long)L.nCols * L.nRows * L.cellCap;
        L.nFeat = e->per_level[l];
        L.nIni = 1; L.hX = width;
        if (L.nCols > 0) {
            L.nIni = (int)std::round(width / (L.maxBY - L.minBY));                     // :570
            if (L.nIni < 1) return fail(ORBFE_ERR_ARG, "image more than twice as tall as wide: the reference divides by zero here");
            L.hX = width / L.nIni;                                                    // :572
        }
        L.selCap = std::max(L.nFeat + 3, 4 * L.nIni) + 5;
        L.selBase = selBase;
        selBase += L.selCap;
        L.scale = e->scale[l];
        L.scaledPatch = (int)(kPatchSize * e->scale[l]);                              // :905
    }
    g.pyrBytes = pyrOff;
    g.ncells = cellBase;
    g.cellSlots = slotOff;
    g.selTotal = selBase;
    g.outCap = selBase;
    return 0;
}

}  // namespace orbfe

namespace orbfe {

// (re)build geometry, device buffers and resize tables for w x h images, `batch` frames
static int ensure_geometry(orbfe_extractor* e, int w, int h, int batch) {
    if (w == e->geomW && h == e->geomH && batch <= e->geomBatchCap) {
        e->g.batch = batch;
        return 0;
    }
    const int cap = std::max(batch, (w == e->geomW && h == e->geomH) ? e->geomBatchCap : 0);
    FrameSetGeom g;
    int rc = build_geometry(e, w, h, cap, g);
    if (rc) return rc;
    CU(cudaSetDevice(e->device));
    CU(cudaStreamSynchronize(e->stream));
    if ((rc = e->pyr.ensure(g.pyrBytes + 256))) return rc;      // slack: the resize kernel loads whole words past a row's last tap
    if ((rc = e->blur.ensure(g.pyrBytes))) return rc;
    if ((rc = e->cellc.ensure((size_t)cap * g.ncells + 1))) return rc;
    if ((rc = e->cellk.ensure((size_t)cap * g.cellSlots + 1))) return rc;
    if ((rc = e->qt_xy.ensure((size_t)cap * g.cellSlots + 1))) return rc;
    if ((rc = e->qt_node.ensure((size_t)cap * g.cellSlots + 1))) return rc;
    if ((rc = e->sel.ensure((size_t)cap * g.selTotal))) return rc;
    if ((rc = e->selc.ensure((size_t)cap * g.nlevels))) return rc;
    if ((rc = e->kps.ensure((size_t)cap * g.outCap))) return rc;
    if ((rc = e->desc.ensure((size_t)cap * g.outCap * 32))) return rc;
    if ((rc = e->counts.ensure(cap))) return rc;
    if (e->haveCam && (rc = ensure_frame_bufs(e, cap, g.outCap))) return rc;
    // resize tables
    std::vector<int2> xall; std::vector<int4> yall;
    std::vector<int4> qall;
    std::vector<size_t> xo(e->nlevels, 0), yo(e->nlevels, 0);
    std::vector<long long> qo(e->nlevels, -1);
    std::vector<char> qsimple(e->nlevels, 0);
    for (int l = 1; l < e->nlevels; l++) {
        std::vector<int2> xt; std::vector<int4> yt;
        resize_tables(g.lv[l - 1].w, g.lv[l - 1].h, g.lv[l].w, g.lv[l].h, xt, yt);
        xo[l] = xall.size(); yo[l] = yall.size();
        std::vector<int4> xq;
        if (resize_group_table(xt, g.lv[l].w, xq)) { qo[l] = (long long)qall.size(); qall.insert(qall.end(), xq.begin(), xq.end()); }
        bool simple = true;
        for (size_t y = 0; y < yt.size(); y++)
            if (yt[y].y != yt[y].x + 1 || (y > 0 && yt[y].y <= yt[y - 1].y)) simple = false;
        qsimple[l] = simple;
        xall.insert(xall.end(), xt.begin(), xt.end());
        yall.insert(yall.end(), yt.begin(), yt.end());
    }
    if ((rc = e->xtab.ensure(xall.size() + 1))) return rc;
    if ((rc = e->ytab.ensure(yall.size() + 1))) return rc;
    if ((rc = e->xq.ensure(qall.size() + 1))) return rc;
    if (!qall.empty()) CU(cudaMemcpy(e->xq.p, qall.data(), qall.size() * sizeof(int4), cudaMemcpyHostToDevice));
    if (!xall.empty()) {
        CU(cudaMemcpy(e->xtab.p, xall.data(), xall.size() * sizeof(int2), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(e->ytab.p, yall.data(), yall.size() * sizeof(int4), cudaMemcpyHostToDevice));
    }
    // FAST cell records (ORBextractor.cc:846-869, integer-valued floats there)
    {
        std::vector<int4> ct(g.ncells > 0 ? g.ncells : 1, make_int4(0, 0, 0, 0));
        for (int l = 0; l < e->nlevels; l++) {
            const LevelGeom& L = g.lv[l];
            for (int ci = 0; ci < L.nRows; ci++)
                for (int cj = 0; cj < L.nCols; cj++) {
                    const int iniY = L.minBY + ci * L.hCell, iniX = L.minBX + cj * L.wCell;
                    int maxY = iniY + L.hCell + 6, maxX = iniX + L.wCell + 6;
                    if (maxY > L.maxBY) maxY = L.maxBY;
                    if (maxX > L.maxBX) maxX = L.maxBX;
                    int ww = maxX - iniX, wh = maxY - iniY;
                    if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6 || ww < 7 || wh < 7) ww = wh = 0;
                    // what the warp kernel's prologue would otherwise derive per cell (fast.cu): whether the window fits
                    // a 48-byte TMA box (40-row layout only), and the multiply-shift constant of its lane -> (row, word
                    // column) mapping, 65536 / nzc + 1 with nzc = word columns of the zone
                    const int rowsK = fast_window_rows(L);
                    const bool direct = rowsK == 40 && ww > 0 && (iniX & 15) + ww <= 48;
                    const int ax0 = rowsK == 40 && !direct ? iniX & ~3 : iniX & ~15;
                    const int zx0 = iniX + 3 - ax0, zx1 = iniX + ww - 3 - ax0;
                    const int nzc = ww > 0 ? ((zx1 - 1) >> 2) - (zx0 >> 2) + 1 : 1;
                    const unsigned mdiv = 65536u / (unsigned)nzc + 1u;
                    for (int a = 0; a <= 32; a++)
                        if ((int)((a * mdiv) >> 16) != a / nzc) return fail(ORBFE_ERR_ARG, "internal: FAST lane-mapping constant is not exact");
                    ct[L.cellBase + ci * L.nCols + cj] =
                        make_int4(iniX | (iniY << 16), ww | (wh << 16), (cj * L.wCell) | ((ci * L.hCell) << 16),
                                  l | (direct ? 16 : 0) | (int)(mdiv << 8));
                }
        }
        if ((rc = e->celltab.ensure(ct.size()))) return rc;
        CU(cudaMemcpy(e->celltab.p, ct.data(), ct.size() * sizeof(int4), cudaMemcpyHostToDevice));
    }
    // TMA descriptors of the level / blurred-level buffers
    {
        static_assert(sizeof(CUtensorMap) == 128, "descriptor stride assumed by the kernel");
        std::vector<CUtensorMap>& maps = e->tmaps;
        maps.resize(2 * e->nlevels);
        e->fastMaps.resize(e->nlevels);
        e->fastMaps48.resize(e->nlevels);
        for (int l = 0; l < e->nlevels; l++) {
            const LevelGeom& L = g.lv[l];
            const int rows = fast_window_rows(L);
            if ((rc = encode_level_map(&e->fastMaps[l], e->pyr.p + L.pyrOff, L.pitch, (long long)L.h * cap, 64, rows ? rows : 40))) return rc;
            if ((rc = encode_level_map(&e->fastMaps48[l], e->pyr.p + L.pyrOff, L.pitch, (long long)L.h * cap, 48, rows ? rows : 40))) return rc;
        }
        for (int l = 0; l < e->nlevels; l++) {
            const LevelGeom& L = g.lv[l];
            if ((rc = encode_level_map(&maps[2 * l], e->pyr.p + L.pyrOff, L.pitch, (long long)L.h * cap, kTmaRawW, kTmaRawH))) return rc;
            if ((rc = encode_level_map(&maps[2 * l + 1], e->blur.p + L.blurOff, L.pitch, (long long)L.h * cap, kTmaBlurW, kTmaBlurH))) return rc;
        }
    }
    DevBufs& d = e->d;
    std::memset(&d, 0, sizeof(d));
    d.tmaps = e->tmaps.data();
    d.fast_tmaps = e->fastMaps.data();
    d.fast_tmaps48 = e->fastMaps48.data();
    d.celltab = e->celltab.p;
    d.pyr = e->pyr.p; d.blur = e->blur.p; d.cellc = e->cellc.p; d.cellk = e->cellk.p;
    d.sel = e->sel.p; d.selc = e->selc.p; d.kps = e->kps.p; d.desc = e->desc.p;
    d.counts = e->counts.p; d.qt_xy = e->qt_xy.p; d.qt_node = e->qt_node.p; d.qtCap = g.cellSlots;
    for (int l = 1; l < e->nlevels; l++) {
        d.xtab[l] = e->xtab.p + xo[l]; d.ytab[l] = e->ytab.p + yo[l];
        d.xq[l] = qo[l] >= 0 ? e->xq.p + qo[l] : nullptr;
        d.xqSimple[l] = qsimple[l] != 0;
    }
    g.batch = batch;
    e->g = g;
    e->geomW = w; e->geomH = h; e->geomBatchCap = cap;
    e->level0MapFrames = cap;
    if (octree_smem_bytes(g) > 200 * 1024) return fail(ORBFE_ERR_ARG, "nfeatures too large for the quadtree kernel's shared memory");
    cudaError_t ce = (cudaError_t)octree_prepare(g);
    if (ce != cudaSuccess) return fail(ORBFE_ERR_CUDA, "cudaFuncSetAttribute(octree)", ce);
    return 0;
}

// the kernel schedule of operator() on the handle's stream; level 0 is already in d.pyr
static int enqueue_extract(orbfe_extractor* e, int frame0 = 0, int count = -1, int lane = 0) {
    int launches = 0;
    const bool prof = e->profiling && frame0 == 0 && lane == 0;
    cudaStream_t st = lane ? e->stream2 : e->stream, side = lane ? e->sideStream2 : e->sideStream;
    cudaEvent_t evPyr = lane ? e->evPyr2 : e->evPyr, evBlur = lane ? e->evBlur2 : e->evBlur;
    const int fullBatch = e->g.batch;
    e->g.frame0 = frame0;
    if (count >= 0) e->g.batch = count;
    if (prof) cudaEventRecord(e->ev[1], st);
    launches += launch_pyramid(e->g, e->d, st);
    if (prof) cudaEventRecord(e->ev[2], st);
    if (prof) {
        // profiling mode serialises the stages so that each one can be timed alone
        launches += launch_fast(e->g, e->d, st);
        cudaEventRecord(e->ev[3], st);
        launches += launch_blur(e->g, e->d, st);
        cudaEventRecord(e->ev[4], st);
        launches += launch_octree(e->g, e->d, st);
        cudaEventRecord(e->ev[5], st);
    } else {
        // FAST saturates the issue slots on its own; the blur runs beside the quadtree, whose
        // barrier-bound CTAs leave slots free
        launches += launch_fast(e->g, e->d, st);
        CU(cudaEventRecord(evPyr, st));
        CU(cudaStreamWaitEvent(side, evPyr, 0));
        launches += launch_blur(e->g, e->d, side);
        CU(cudaEventRecord(evBlur, side));
        launches += launch_octree(e->g, e->d, st);
        CU(cudaStreamWaitEvent(st, evBlur, 0));
    }
    launches += launch_orient_desc(e->g, e->d, e->umax_d.p, st);
    if (prof) cudaEventRecord(e->ev[6], st);
    if (e->haveCam)
        launches += launch_frame_epilogue(e->camGeom, e->d.kps, e->d.counts, e->g.outCap, frame0, e->g.batch, e->kpsUn.p,
                                          e->gridEnd.p, e->gridItems.p, st);
    if (prof) cudaEventRecord(e->ev[7], st);
    e->g.frame0 = 0;
    e->g.batch = fullBatch;
    e->lastLaunches = frame0 == 0 ? launches : e->lastLaunches + launches;
    CU(cudaGetLastError());
    return 0;
}

// Small batches are bound by the latency of their dependent kernels, not by throughput: FAST, the quadtree and the
// blur of level l only need level l, so they fork off the pyramid chain level by level (eight short parallel branches
// instead of pyramid -> FAST -> quadtree in sequence) and join in front of the orientation / descriptor kernel.  Under
// stream capture the branch streams become edges of the graph.
static int enqueue_extract_forked(orbfe_extractor* e) {
    const int L = e->g.nlevels;
    if ((int)e->lvStream.size() < 2 * L) {
        e->lvStream.resize(2 * L, nullptr); e->lvReady.resize(L, nullptr); e->lvDone.resize(2 * L, nullptr);
        for (auto& st : e->lvStream) if (!st) CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        for (auto& ev : e->lvReady) if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        for (auto& ev : e->lvDone) if (!ev) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    }
    cudaStream_t st = e->stream;
    int launches = 0;
    e->g.frame0 = 0;
    for (int l = 0; l < L; l++) {
        if (l > 0) launches += launch_pyramid(e->g, e->d, st, l, l + 1);
        CU(cudaEventRecord(e->lvReady[l], st));
        cudaStream_t a = e->lvStream[2 * l], b = e->lvStream[2 * l + 1];
        CU(cudaStreamWaitEvent(a, e->lvReady[l], 0));
        launches += launch_fast(e->g, e->d, a, l, l + 1);
        launches += launch_octree(e->g, e->d, a, l, l + 1);
        CU(cudaEventRecord(e->lvDone[2 * l], a));
        CU(cudaStreamWaitEvent(b, e->lvReady[l], 0));
        launches += launch_blur(e->g, e->d, b, l, l + 1);
        CU(cudaEventRecord(e->lvDone[2 * l + 1], b));
    }
    for (int i = 0; i < 2 * L; i++) CU(cudaStreamWaitEvent(st, e->lvDone[i], 0));
    launches += launch_orient_desc(e->g, e->d, e->umax_d.p, st);
    if (e->haveCam)
        launches += launch_frame_epilogue(e->camGeom, e->d.kps, e->d.counts, e->g.outCap, 0, e->g.batch, e->kpsUn.p,
                                          e->gridEnd.p, e->gridItems.p, st);
    e->lastLaunches = launches;
    CU(cudaGetLastError());
    return 0;
}

// Everything the launchers pass to the kernels, byte for byte: the replay is valid exactly while this is unchanged.
static void graph_key(const orbfe_extractor* e, std::vector<uint8_t>& key) {
    key.clear();
    auto put = [&](const void* p, size_t n) { const uint8_t* b = (const uint8_t*)p; key.insert(key.end(), b, b + n); };
    put(&e->g, sizeof(e->g)); put(&e->d, sizeof(e->d));
    put(e->tmaps.data(), e->tmaps.size() * sizeof(CUtensorMap)); put(e->fastMaps.data(), e->fastMaps.size() * sizeof(CUtensorMap));
    put(e->fastMaps48.data(), e->fastMaps48.size() * sizeof(CUtensorMap));
    const void* ptrs[4] = {e->umax_d.p, e->kpsUn.p, e->gridEnd.p, e->gridItems.p};
    put(ptrs, sizeof(ptrs));
    const int cam = e->haveCam ? 1 : 0;
    put(&cam, sizeof(cam));
    if (e->haveCam) put(&e->camGeom, sizeof(e->camGeom));
}

// Small batches are launch-latency bound (one 640x480 frame: ~25 dependent launches of a few microseconds each), so their
// schedule is replayed from a graph.  The first call with a given argument set runs eagerly (it also applies the
// per-function attributes), the second captures, later ones replay.
static int enqueue_extract_small(orbfe_extractor* e) {
    static const bool off = std::getenv("ORBFE_NO_GRAPH") != nullptr;
    static const bool nofork = std::getenv("ORBFE_NO_FORK") != nullptr;
    auto enqueue = [&](orbfe_extractor* x) { return nofork ? enqueue_extract(x) : enqueue_extract_forked(x); };
    if (e->profiling) return enqueue_extract(e);
    if (off || e->graphRefused) return enqueue(e);
    std::vector<uint8_t> key;
    graph_key(e, key);
    if (e->graphExec && key == e->graphKey) {
        CU(cudaGraphLaunch(e->graphExec, e->stream));
        e->lastLaunches = e->graphLaunches;
        return 0;
    }
    if (e->graphExec) { cudaGraphExecDestroy(e->graphExec); e->graphExec = nullptr; }
    if (key != e->graphKey) {                        // first sight of this argument set: eager
        e->graphKey.swap(key);
        return enqueue(e);
    }
    cudaGraph_t graph = nullptr;
    CU(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = enqueue(e);
    const cudaError_t ce = cudaStreamEndCapture(e->stream, &graph);
    if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        e->graphKey.clear();
        e->graphRefused = true;
        return rc ? rc : enqueue(e);         // capture refused: stay eager
    }
    e->graphLaunches = e->lastLaunches;
    const cudaError_t ci = cudaGraphInstantiate(&e->graphExec, graph, 0);
    cudaGraphDestroy(graph);
    if (ci != cudaSuccess) { e->graphExec = nullptr; e->graphKey.clear(); e->graphRefused = true; cudaGetLastError(); return enqueue(e); }
    CU(cudaGraphLaunch(e->graphExec, e->stream));
    return 0;
}

// Level 0 either lives in the pyramid buffer (copied there) or aliases the caller's device frames
// when their layout already is the level layout (16-byte aligned base, stride == pitch, frames
// contiguous): every kernel addresses level 0 as d.pyr + lv[0].pyrOff, so the alias is just a
// different offset, plus the level-0 TMA descriptor of the orientation kernel.
static int set_level0_source(orbfe_extractor* e, const uint8_t* alias, int aliasFrames = 0) {
    LevelGeom& L0 = e->g.lv[0];
    const uint8_t* want = alias ? alias : e->pyr.p + L0.blurOff;       // blurOff keeps the in-buffer offset
    // the tensor maps of an alias cover exactly the frames the caller supplied: TMA boxes that reach past the last
    // frame's last row are zero-filled instead of fetched from behind the caller's allocation
    const int frames = alias ? aliasFrames : e->geomBatchCap;
    if (e->d.pyr + L0.pyrOff == want && e->level0MapFrames == frames) return 0;
    L0.pyrOff = want - e->d.pyr;
    e->level0MapFrames = frames;
    int rc = encode_level_map(&e->tmaps[0], (void*)want, L0.pitch, (long long)L0.h * frames, kTmaRawW, kTmaRawH);
    if (rc) return rc;
    const int rows = fast_window_rows(L0);
    if ((rc = encode_level_map(&e->fastMaps[0], (void*)want, L0.pitch, (long long)L0.h * frames, 64, rows ? rows : 40))) return rc;
    return encode_level_map(&e->fastMaps48[0], (void*)want, L0.pitch, (long long)L0.h * frames, 48, rows ? rows : 40);
}

// copy `batch` frames into level 0 of the pyramid (kind: H2D or D2D)
static int upload_level0(orbfe_extractor* e, const uint8_t* imgs, int batch, int w, int h,
                         ptrdiff_t stride, ptrdiff_t frame_stride, cudaMemcpyKind kind) {
    int rc0 = set_level0_source(e, nullptr);
    if (rc0) return rc0;
    const LevelGeom& L0 = e->g.lv[0];
    uint8_t* dst = e->d.pyr + L0.pyrOff;
    if (e->profiling) cudaEventRecord(e->ev[0], e->stream);
    if (batch == 1 || frame_stride == (ptrdiff_t)h * stride) {
        CU(cudaMemcpy2DAsync(dst, L0.pitch, imgs, stride, w, (size_t)h * batch, kind, e->stream));
    } else {
        for (int b = 0; b < batch; b++)
            CU(cudaMemcpy2DAsync(dst + (long long)b * L0.lvBytes, L0.pitch, imgs + b * frame_stride,
                                 stride, w, h, kind, e->stream));
    }
    return 0;
}

}  // namespace orbfe

extern "C" {

const char* orbfe_last_error(void) { return g_err.c_str(); }
const char* orbfe_version(void) { return "orbfe 0.1 (sm_100a)"; }

void* orbfe_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
void orbfe_host_free(void* p) { if (p) cudaFreeHost(p); }

int orbfe_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int orbfe_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th,
                           int max_width, int max_height, int max_batch, int device,
                           orbfe_extractor** out) {
    if (!out) return fail(ORBFE_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (nfeatures < 1 || nlevels < 1 || nlevels > kMaxLevels || !(scale_factor > 1.0f) ||
        ini_th < 0 || min_th < 0 || ini_th > 255 || min_th > 255)
        return fail(ORBFE_ERR_ARG, "bad extractor parameters");
    if (max_width > kMaxDim || max_height > kMaxDim) return fail(ORBFE_ERR_ARG, "image larger than 4096 px");
    if (orbfe_device_count() <= device || device < 0)
        return fail(ORBFE_ERR_NO_DEVICE, "no CUDA device: liborbfe has no CPU fallback");
    orbfe_extractor* e = new (std::nothrow) orbfe_extractor();
    if (!e) return fail(ORBFE_ERR_ARG, "out of host memory");
    e->nfeatures = nfeatures; e->nlevels = nlevels; e->iniTh = ini_th; e->minTh = min_th;
    e->device = device;
    build_tables(e, nfeatures, scale_factor, nlevels);
    if (!umax_symmetric(e->umax)) { delete e; return fail(ORBFE_ERR_ARG, "internal: circular patch table is not symmetric"); }
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->sideStream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->evPyr, cudaEventDisableTiming);
    if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&e->evBlur, cudaEventDisableTiming);
    if (ce != cudaSuccess) { orbfe_extractor_destroy(e); return fail(ORBFE_ERR_CUDA, "cudaStreamCreate", ce); }
    int rc = e->umax_d.ensure(16);
    if (!rc) {
        ce = cudaMemcpy(e->umax_d.p, e->umax, sizeof(e->umax), cudaMemcpyHostToDevice);
        if (ce != cudaSuccess) rc = fail(ORBFE_ERR_CUDA, "cudaMemcpy(umax)", ce);
    }
    if (!rc && max_width > 0 && max_height > 0 && max_batch > 0)
        rc = ensure_geometry(e, max_width, max_height, max_batch);
    if (rc) { orbfe_extractor_destroy(e); return rc; }
    *out = e;
    return ORBFE_OK;
}

void orbfe_extractor_destroy(orbfe_extractor* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream) { cudaStreamSynchronize(e->stream); cudaStreamDestroy(e->stream); }
    if (e->graphExec) cudaGraphExecDestroy(e->graphExec);
    if (e->sideStream) { cudaStreamSynchronize(e->sideStream); cudaStreamDestroy(e->sideStream); }
    if (e->evPyr) cudaEventDestroy(e->evPyr);
    if (e->evBlur) cudaEventDestroy(e->evBlur);
    if (e->h2dStream) cudaStreamDestroy(e->h2dStream);
    if (e->d2hStream) cudaStreamDestroy(e->d2hStream);
    if (e->stream2) cudaStreamDestroy(e->stream2);
    if (e->sideStream2) cudaStreamDestroy(e->sideStream2);
    if (e->evPyr2) cudaEventDestroy(e->evPyr2);
    if (e->evBlur2) cudaEventDestroy(e->evBlur2);
    for (auto st : e->lvStream) if (st) cudaStreamDestroy(st);
    for (auto ev : e->lvReady) if (ev) cudaEventDestroy(ev);
    for (auto ev : e->lvDone) if (ev) cudaEventDestroy(ev);
    for (auto ev : e->evUp) cudaEventDestroy(ev);
    for (auto ev : e->evDone) cudaEventDestroy(ev);
    for (auto& ev : e->ev) if (ev) cudaEventDestroy(ev);
    e->stageIn.release(); e->stageOut.release(); e->stagePyr.release();
    e->pyr.release(); e->blur.release(); e->desc.release(); e->cellc.release(); e->selc.release();
    e->counts.release(); e->umax_d.release(); e->cellk.release(); e->sel.release();
    e->qt_xy.release(); e->qt_node.release(); e->kps.release(); e->kpsUn.release(); e->gridEnd.release(); e->gridItems.release(); e->xtab.release(); e->ytab.release(); e->xq.release(); e->celltab.release();
    delete e;
}

int orbfe_get_scale_tables(const orbfe_extractor* e, float* scale, float* inv_scale, float* sigma2,
                           float* inv_sigma2, int* features_per_level, int* umax) {
    if (!e) return fail(ORBFE_ERR_ARG, "extractor is NULL");
    for (int i = 0; i < e->nlevels; i++) {
        if (scale) scale[i] = e->scale[i];
        if (inv_scale) inv_scale[i] = e->inv_scale[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
        if (features_per_level) features_per_level[i] = e->per_level[i];
    }
    if (umax) for (int i = 0; i < 16; i++) umax[i] = e->umax[i];
    return ORBFE_OK;
}

int orbfe_extractor_levels(const orbfe_extractor* e) { return e ? e->nlevels : 0; }

int orbfe_extractor_max_keypoints(const orbfe_extractor* e) {
    if (!e) return 0;
    if (e->geomW) return e->g.outCap;
    int n = 0;
    for (int l = 0; l < e->nlevels; l++) n += std::max(e->per_level[l] + 3, 64) + 5;
    return n;
}

}  // extern "C"

extern "C" {

int orbfe_extract_batch_device(orbfe_extractor* e, const uint8_t* d_imgs, int batch, int width,
                               int height, ptrdiff_t stride, ptrdiff_t frame_stride) {
    if (!e || !d_imgs || batch < 1 || width < 1 || height < 1 || stride < width)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_extract_batch_device");
    if (width > kMaxDim || height > kMaxDim) return fail(ORBFE_ERR_ARG, "image larger than 4096 px");
    CU(cudaSetDevice(e->device));
    int rc = ensure_geometry(e, width, height, batch);
    if (rc) return rc;
    const LevelGeom& L0 = e->g.lv[0];
    if (stride == L0.pitch && (batch == 1 || frame_stride == (ptrdiff_t)height * stride) &&
        (reinterpret_cast<uintptr_t>(d_imgs) & 15) == 0) {
        // the frames already are level 0: no copy (they must stay valid until the results are fetched)
        if ((rc = set_level0_source(e, d_imgs, batch))) return rc;
        if (e->profiling) cudaEventRecord(e->ev[0], e->stream);
    } else if ((rc = upload_level0(e, d_imgs, batch, width, height, stride, frame_stride, cudaMemcpyDeviceToDevice))) {
        return rc;
    }
    e->lastBatch = batch;
    return batch <= 4 ? enqueue_extract_small(e) : enqueue_extract(e);
}

int orbfe_extractor_sync(orbfe_extractor* e) {
    if (!e) return fail(ORBFE_ERR_ARG, "extractor is NULL");
    CU(cudaSetDevice(e->device));
    CU(cudaStreamSynchronize(e->stream));
    return ORBFE_OK;
}

int orbfe_extractor_device_results(orbfe_extractor* e, const orbfe_keypoint** d_kps,
                                   const uint8_t** d_desc, const int32_t** d_counts, int* cap) {
    if (!e || !e->geomW) return fail(ORBFE_ERR_ARG, "no extraction has run on this handle");
    if (d_kps) *d_kps = e->d.kps;
    if (d_desc) *d_desc = e->d.desc;
    if (d_counts) *d_counts = e->d.counts;
    if (cap) *cap = e->g.outCap;
    return ORBFE_OK;
}

int orbfe_extractor_device_pyramid(orbfe_extractor* e, orbfe_device_pyramid_view* out) {
    if (!e || !out || !e->geomW || !e->lastBatch) return fail(ORBFE_ERR_ARG, "no extraction has run on this handle");
    const int L = e->g.nlevels;
    e->pv_data.resize(L); e->pv_w.resize(L); e->pv_h.resize(L); e->pv_pitch.resize(L); e->pv_fstride.resize(L);
    for (int l = 0; l < L; l++) {
        const LevelGeom& G = e->g.lv[l];
        e->pv_data[l] = e->d.pyr + G.pyrOff;     // level 0 may be the caller's own device frames
        e->pv_w[l] = G.w; e->pv_h[l] = G.h; e->pv_pitch[l] = G.pitch; e->pv_fstride[l] = (ptrdiff_t)G.lvBytes;
    }
    out->nlevels = L; out->d_data = e->pv_data.data(); out->width = e->pv_w.data(); out->height = e->pv_h.data();
    out->pitch = e->pv_pitch.data(); out->frame_stride = e->pv_fstride.data();
    return ORBFE_OK;
}

int orbfe_extractor_fetch(orbfe_extractor* e, int batch, orbfe_keypoint* kps, uint8_t* desc, int cap,
                          int* counts) {
    if (!e || !kps || !desc || !counts || batch < 1 || batch > e->lastBatch)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_extractor_fetch");
    CU(cudaSetDevice(e->device));
    const int oc = e->g.outCap;
    CU(cudaMemcpyAsync(counts, e->d.counts, sizeof(int) * batch, cudaMemcpyDeviceToHost, e->stream));
    if (cap >= oc) {
        // one strided copy each: rows of oc slots into rows of cap slots
        CU(cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbfe_keypoint), e->d.kps, (size_t)oc * sizeof(orbfe_keypoint),
                             (size_t)oc * sizeof(orbfe_keypoint), batch, cudaMemcpyDeviceToHost, e->stream));
        CU(cudaMemcpy2DAsync(desc, (size_t)cap * 32, e->d.desc, (size_t)oc * 32, (size_t)oc * 32, batch,
                             cudaMemcpyDeviceToHost, e->stream));
        CU(cudaStreamSynchronize(e->stream));
        return ORBFE_OK;
    }
    CU(cudaStreamSynchronize(e->stream));
    for (int b = 0; b < batch; b++)
        if (counts[b] > cap) return fail(ORBFE_ERR_CAPACITY, "keypoint capacity too small (see orbfe_extractor_max_keypoints)");
    CU(cudaMemcpy2DAsync(kps, (size_t)cap * sizeof(orbfe_keypoint), e->d.kps, (size_t)oc * sizeof(orbfe_keypoint),
                         (size_t)cap * sizeof(orbfe_keypoint), batch, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaMemcpy2DAsync(desc, (size_t)cap * 32, e->d.desc, (size_t)oc * 32, (size_t)cap * 32, batch,
                         cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    return ORBFE_OK;
}

// Chunked pipeline for large host batches: three streams, chunk c's kernels overlap the upload of
// chunk c+1 and the download of chunk c-1.
static int extract_batch_pipelined(orbfe_extractor* e, const uint8_t* imgs, int batch, int w, int h,
                                   ptrdiff_t stride, ptrdiff_t frame_stride, orbfe_keypoint* kps,
                                   uint8_t* desc, int cap, int* counts, int chunk) {
    // chunk schedule: the transfer link is the bottleneck, so the pipeline's fill (first upload, nothing to
    // overlap with) and drain (last kernels + download) are kept short with small chunks at both ends
    std::vector<int> first, count;
    {
        std::vector<int> sz;
        int rem = batch;
        const int taper = batch >= 6 * chunk ? 2 : 0;
        int tail = 0;
        for (int t = 0; t < taper; t++) { sz.push_back(chunk >> (2 - t)); rem -= sz.back(); tail += chunk >> (2 - t); }
        rem -= tail;
        while (rem > 0) { sz.push_back(std::min(chunk, rem)); rem -= sz.back(); }
        for (int t = taper - 1; t >= 0; t--) sz.push_back(chunk >> (2 - t));
        int f = 0;
        for (int v : sz) { first.push_back(f); count.push_back(v); f += v; }
    }
    const int nchunks = (int)count.size();
    {
        const int rc0 = set_level0_source(e, nullptr);
        if (rc0) return rc0;
    }
    if (!e->h2dStream) {
        CU(cudaStreamCreateWithFlags(&e->h2dStream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&e->d2hStream, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&e->stream2, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&e->sideStream2, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&e->evPyr2, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->evBlur2, cudaEventDisableTiming));
    }
    while ((int)e->evUp.size() < nchunks) {
        cudaEvent_t a, b;
        CU(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        e->evUp.push_back(a); e->evDone.push_back(b);
    }
    const LevelGeom& L0 = e->g.lv[0];
    const int oc = e->g.outCap;
    const size_t kb = sizeof(orbfe_keypoint);
    // everything queued before this call on the main stream must finish before buffers are reused
    CU(cudaEventRecord(e->evDone[0], e->stream));
    CU(cudaStreamWaitEvent(e->h2dStream, e->evDone[0], 0));
    for (int c = 0; c < nchunks; c++) {
        const int f0 = first[c], n = count[c];
        uint8_t* dst = e->d.pyr + L0.pyrOff + (long long)f0 * L0.lvBytes;
        const uint8_t* src = imgs + (long long)f0 * frame_stride;
        if (frame_stride == (ptrdiff_t)h * stride && stride == L0.pitch) {
            CU(cudaMemcpyAsync(dst, src, (size_t)L0.lvBytes * n, cudaMemcpyHostToDevice, e->h2dStream));   // one flat DMA
        } else if (frame_stride == (ptrdiff_t)h * stride) {
            CU(cudaMemcpy2DAsync(dst, L0.pitch, src, stride, w, (size_t)h * n, cudaMemcpyHostToDevice, e->h2dStream));
        } else {
            for (int b = 0; b < n; b++)
                CU(cudaMemcpy2DAsync(dst + (long long)b * L0.lvBytes, L0.pitch, src + b * frame_stride, stride, w, h,
                                     cudaMemcpyHostToDevice, e->h2dStream));
        }
        const int lane = c & 1;
        cudaStream_t st = lane ? e->stream2 : e->stream;
        CU(cudaEventRecord(e->evUp[c], e->h2dStream));
        CU(cudaStreamWaitEvent(st, e->evUp[c], 0));
        int rc = enqueue_extract(e, f0, n, lane);
        if (rc) return rc;
        CU(cudaEventRecord(e->evDone[c], st));
        CU(cudaStreamWaitEvent(e->d2hStream, e->evDone[c], 0));
        const int rowk = std::min(cap, oc);
        CU(cudaMemcpyAsync(counts + f0, e->d.counts + f0, sizeof(int) * n, cudaMemcpyDeviceToHost, e->d2hStream));
        if (cap == oc) {
            CU(cudaMemcpyAsync(kps + (size_t)f0 * cap, e->d.kps + (size_t)f0 * oc, (size_t)oc * kb * n, cudaMemcpyDeviceToHost, e->d2hStream));
            CU(cudaMemcpyAsync(desc + (size_t)f0 * cap * 32, e->d.desc + (size_t)f0 * oc * 32, (size_t)oc * 32 * n,
                               cudaMemcpyDeviceToHost, e->d2hStream));
        } else {
            CU(cudaMemcpy2DAsync(kps + (size_t)f0 * cap, (size_t)cap * kb, e->d.kps + (size_t)f0 * oc, (size_t)oc * kb,
                                 (size_t)rowk * kb, n, cudaMemcpyDeviceToHost, e->d2hStream));
            CU(cudaMemcpy2DAsync(desc + (size_t)f0 * cap * 32, (size_t)cap * 32, e->d.desc + (size_t)f0 * oc * 32,
                                 (size_t)oc * 32, (size_t)rowk * 32, n, cudaMemcpyDeviceToHost, e->d2hStream));
        }
    }
    CU(cudaStreamSynchronize(e->d2hStream));
    CU(cudaStreamSynchronize(e->stream2));
    CU(cudaStreamSynchronize(e->stream));
    if (cap < oc)
        for (int b = 0; b < batch; b++)
            if (counts[b] > cap) return fail(ORBFE_ERR_CAPACITY, "keypoint capacity too small (see orbfe_extractor_max_keypoints)");
    return ORBFE_OK;
}

// Pageable caller memory (a cv::Mat, a std::vector, a numpy array) makes the driver stage every copy itself and
// synchronise around it: measured 64 us of a 170 us single-frame call.  Small batches therefore go through the handle's
// own page-locked buffers: one CPU memcpy in, one DMA each way, and only the valid result rows are copied out.
static bool is_pageable(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

static int extract_small_staged(orbfe_extractor* e, const uint8_t* imgs, int batch, int w, int h, ptrdiff_t stride,
                                ptrdiff_t frame_stride, orbfe_keypoint* kps, uint8_t* desc, int cap, int* counts) {
    int rc;
    const uint8_t* src = imgs;
    ptrdiff_t sstride = stride, sframe = frame_stride;
    if (is_pageable(imgs)) {
        if ((rc = e->stageIn.ensure((size_t)batch * h * w))) return rc;
        CU(cudaStreamSynchronize(e->stream));                      // the previous call's DMA out of the staging buffer
        for (int b = 0; b < batch; b++) {
            uint8_t* dst = e->stageIn.p + (size_t)b * h * w;
            const uint8_t* from = imgs + b * frame_stride;
            if (stride == w) std::memcpy(dst, from, (size_t)h * w);
            else for (int y = 0; y < h; y++) std::memcpy(dst + (size_t)y * w, from + y * stride, w);
        }
        src = e->stageIn.p; sstride = w; sframe = (ptrdiff_t)h * w;
    }
    if ((rc = upload_level0(e, src, batch, w, h, sstride, sframe, cudaMemcpyHostToDevice))) return rc;
    e->lastBatch = batch;
    if ((rc = enqueue_extract_small(e))) return rc;
    if (!is_pageable(kps) && !is_pageable(desc)) return orbfe_extractor_fetch(e, batch, kps, desc, cap, counts);
    const int oc = e->g.outCap;
    const size_t kb = sizeof(orbfe_keypoint);
    const size_t need = 256 + (size_t)batch * oc * (kb + 32);
    if ((rc = e->stageOut.ensure(need))) return rc;
    int* h_counts = reinterpret_cast<int*>(e->stageOut.p);
    orbfe_keypoint* h_kps = reinterpret_cast<orbfe_keypoint*>(e->stageOut.p + 256);
    uint8_t* h_desc = e->stageOut.p + 256 + (size_t)batch * oc * kb;
    CU(cudaMemcpyAsync(h_counts, e->d.counts, sizeof(int) * batch, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaMemcpyAsync(h_kps, e->d.kps, (size_t)batch * oc * kb, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaMemcpyAsync(h_desc, e->d.desc, (size_t)batch * oc * 32, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    for (int b = 0; b < batch; b++) {
        const int n = h_counts[b];
        counts[b] = n;
        if (n > cap) return fail(ORBFE_ERR_CAPACITY, "keypoint capacity too small (see orbfe_extractor_max_keypoints)");
        std::memcpy(kps + (size_t)b * cap, h_kps + (size_t)b * oc, (size_t)n * kb);
        std::memcpy(desc + (size_t)b * cap * 32, h_desc + (size_t)b * oc * 32, (size_t)n * 32);
    }
    return ORBFE_OK;
}

int orbfe_extract_batch(orbfe_extractor* e, const uint8_t* imgs, int batch, int width, int height,
                        ptrdiff_t stride, ptrdiff_t frame_stride, orbfe_keypoint* kps,
                        uint8_t* desc, int cap, int* counts) {
    if (!e || !imgs || !kps || !desc || !counts || batch < 1 || width < 1 || height < 1 || stride < width)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_extract_batch");
    if (width > kMaxDim || height > kMaxDim) return fail(ORBFE_ERR_ARG, "image larger than 4096 px");
    CU(cudaSetDevice(e->device));
    int rc = ensure_geometry(e, width, height, batch);
    if (rc) return rc;
    static const int chunk = [] { const char* v = std::getenv("ORBFE_CHUNK"); const int c = v ? std::atoi(v) : 0; return c >= 8 ? c : 64; }();
    if (batch >= 4 * chunk) {
        e->lastBatch = batch;
        return extract_batch_pipelined(e, imgs, batch, width, height, stride, frame_stride, kps, desc, cap, counts, chunk);
    }
    if (batch <= 4) return extract_small_staged(e, imgs, batch, width, height, stride, frame_stride, kps, desc, cap, counts);
    if ((rc = upload_level0(e, imgs, batch, width, height, stride, frame_stride, cudaMemcpyHostToDevice))) return rc;
    e->lastBatch = batch;
    if ((rc = enqueue_extract(e))) return rc;
    return orbfe_extractor_fetch(e, batch, kps, desc, cap, counts);
}

int orbfe_extract(orbfe_extractor* e, const uint8_t* img, int width, int height, ptrdiff_t stride,
                  orbfe_keypoint* kps, uint8_t* desc, int cap, int* n) {
    if (!e || !n) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_extract");
    *n = 0;
    if (!img || width <= 0 || height <= 0) return ORBFE_OK;   // ORBextractor.cc:1122-1123
    return orbfe_extract_batch(e, img, 1, width, height, stride, (ptrdiff_t)height * stride, kps, desc, cap, n);
}

int orbfe_extractor_set_profiling(orbfe_extractor* e, int on) {
    if (!e) return fail(ORBFE_ERR_ARG, "extractor is NULL");
    CU(cudaSetDevice(e->device));
    if (on && !e->ev[0])
        for (auto& ev : e->ev) CU(cudaEventCreate(&ev));
    e->profiling = on != 0;
    return ORBFE_OK;
}

int orbfe_extractor_stage_times(orbfe_extractor* e, float* ms, int n) {
    if (!e || !ms || n < ORBFE_NUM_STAGES || !e->ev[0]) return fail(ORBFE_ERR_ARG, "profiling was not enabled");
    CU(cudaSetDevice(e->device));
    CU(cudaEventSynchronize(e->ev[ORBFE_NUM_STAGES]));
    for (int i = 0; i < ORBFE_NUM_STAGES; i++) CU(cudaEventElapsedTime(&ms[i], e->ev[i], e->ev[i + 1]));
    return ORBFE_OK;
}

// Frame::ComputeImageBounds once per camera (Frame.cc:97-111), then the epilogue rides on every extract call
int orbfe_extractor_set_camera(orbfe_extractor* e, const orbfe_camera* cam, int cols, int rows, float bounds[4]) {
    if (!e) return fail(ORBFE_ERR_ARG, "extractor is NULL");
    CU(cudaSetDevice(e->device));
    CU(cudaStreamSynchronize(e->stream));
    if (!cam) { e->haveCam = false; return ORBFE_OK; }
    if (cols <= 0 || rows <= 0 || !(cam->fx != 0.0f) || !(cam->fy != 0.0f))
        return fail(ORBFE_ERR_ARG, "bad camera / image size in orbfe_extractor_set_camera");
    int rc;
    if ((rc = e->gridEnd.ensure(ORBFE_GRID_CELLS))) return rc;      // doubles as the corner scratch
    float b[4];
    if ((rc = image_bounds_device(*cam, cols, rows, e->gridEnd.p, e->stream, b))) return rc;
    if (!(b[1] > b[0]) || !(b[3] > b[2])) return fail(ORBFE_ERR_ARG, "camera gives empty image bounds");
    e->camGeom = cam_geom(*cam, b);
    if (e->geomBatchCap && (rc = ensure_frame_bufs(e, e->geomBatchCap, e->g.outCap))) return rc;
    e->haveCam = true;
    if (bounds) std::memcpy(bounds, b, sizeof(b));
    return ORBFE_OK;
}

int orbfe_extractor_device_frames(orbfe_extractor* e, const orbfe_keypoint** d_kps_un, const int32_t** d_cell_end,
                                  const uint16_t** d_items) {
    if (!e || !e->geomW || !e->haveCam || !e->kpsUn.p)
        return fail(ORBFE_ERR_ARG, "no extraction with a camera (orbfe_extractor_set_camera) has run on this handle");
    if (d_kps_un) *d_kps_un = e->kpsUn.p;
    if (d_cell_end) *d_cell_end = e->gridEnd.p;
    if (d_items) *d_items = e->gridItems.p;
    return ORBFE_OK;
}

int orbfe_extractor_fetch_frames(orbfe_extractor* e, int batch, orbfe_keypoint* kps_un, int cap, int32_t* cell_end,
                                 uint16_t* items) {
    if (!e || !e->geomW || !e->haveCam || !e->kpsUn.p || batch < 1 || batch > e->lastBatch)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_extractor_fetch_frames");
    const int oc = e->g.outCap;
    if ((kps_un || items) && cap < oc) {
        // rows are copied whole: the caller's slots must hold what the device rows hold
        std::vector<int> counts(batch);
        CU(cudaSetDevice(e->device));
        CU(cudaMemcpyAsync(counts.data(), e->d.counts, sizeof(int) * batch, cudaMemcpyDeviceToHost, e->stream));
        CU(cudaStreamSynchronize(e->stream));
        for (int b = 0; b < batch; b++)
            if (counts[b] > cap) return fail(ORBFE_ERR_CAPACITY, "keypoint capacity too small (see orbfe_extractor_max_keypoints)");
    }
    CU(cudaSetDevice(e->device));
    const int w = std::min(cap, oc);
    if (kps_un)
        CU(cudaMemcpy2DAsync(kps_un, (size_t)cap * sizeof(orbfe_keypoint), e->kpsUn.p, (size_t)oc * sizeof(orbfe_keypoint),
                             (size_t)w * sizeof(orbfe_keypoint), batch, cudaMemcpyDeviceToHost, e->stream));
    if (items)
        CU(cudaMemcpy2DAsync(items, (size_t)cap * 2, e->gridItems.p, (size_t)oc * 2, (size_t)w * 2, batch,
                             cudaMemcpyDeviceToHost, e->stream));
    if (cell_end)
        CU(cudaMemcpyAsync(cell_end, e->gridEnd.p, sizeof(int) * ORBFE_GRID_CELLS * (size_t)batch, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    return ORBFE_OK;
}

void* orbfe_extractor_stream(orbfe_extractor* e) { return e ? (void*)e->stream : nullptr; }
int orbfe_extractor_last_launches(const orbfe_extractor* e) { return e ? e->lastLaunches : 0; }

int orbfe_pyramid_level(orbfe_extractor* e, int frame, int level, int blurred, uint8_t* dst,
                        ptrdiff_t dst_stride, int* width, int* height) {
    if (!e || !e->geomW || level < 0 || level >= e->nlevels || frame < 0 || frame >= e->lastBatch)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_pyramid_level");
    const LevelGeom& L = e->g.lv[level];
    if (width) *width = L.w;
    if (height) *height = L.h;
    if (!dst) return ORBFE_OK;
    if (dst_stride < L.w) return fail(ORBFE_ERR_ARG, "dst_stride smaller than the level width");
    CU(cudaSetDevice(e->device));
    const uint8_t* src = (blurred ? e->d.blur + L.blurOff : e->d.pyr + L.pyrOff) + (long long)frame * L.lvBytes;
    CU(cudaMemcpy2DAsync(dst, dst_stride, src, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    return ORBFE_OK;
}

int orbfe_pyramid_host(orbfe_extractor* e, int frame, const uint8_t** levels, int32_t* width, int32_t* height,
                       ptrdiff_t* stride) {
    if (!e || !e->geomW || !levels || frame < 0 || frame >= e->lastBatch)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_pyramid_host");
    CU(cudaSetDevice(e->device));
    size_t total = 0;
    for (int l = 0; l < e->nlevels; l++) total += ((size_t)e->g.lv[l].lvBytes + 255) & ~size_t(255);
    int rc = e->stagePyr.ensure(total);
    if (rc) return rc;
    size_t off = 0;
    for (int l = 0; l < e->nlevels; l++) {
        const LevelGeom& L = e->g.lv[l];
        // a level of one frame is contiguous on the device (rows at the level's pitch): one flat copy per level
        CU(cudaMemcpyAsync(e->stagePyr.p + off, e->d.pyr + L.pyrOff + (long long)frame * L.lvBytes, (size_t)L.lvBytes,
                           cudaMemcpyDeviceToHost, e->stream));
        levels[l] = e->stagePyr.p + off;
        if (width) width[l] = L.w;
        if (height) height[l] = L.h;
        if (stride) stride[l] = L.pitch;
        off += ((size_t)L.lvBytes + 255) & ~size_t(255);
    }
    CU(cudaStreamSynchronize(e->stream));
    return ORBFE_OK;
}

int orbfe_level_raw_keypoints(orbfe_extractor* e, int frame, int level, int32_t* xys, int cap, int* n) {
    if (!e || !e->geomW || !n || level < 0 || level >= e->nlevels || frame < 0 || frame >= e->lastBatch)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_level_raw_keypoints");
    const LevelGeom& L = e->g.lv[level];
    const int ncell = L.nCols * L.nRows;
    *n = 0;
    if (!ncell) return ORBFE_OK;
    CU(cudaSetDevice(e->device));
    std::vector<int> cc(ncell);
    std::vector<uint32_t> ck((size_t)ncell * L.cellCap);
    CU(cudaMemcpyAsync(cc.data(), e->d.cellc + (long long)frame * e->g.ncells + L.cellBase, sizeof(int) * ncell,
                       cudaMemcpyDeviceToHost, e->stream));
    CU(cudaMemcpyAsync(ck.data(), e->d.cellk + (long long)frame * e->g.cellSlots + L.cellSlotOff,
                       sizeof(uint32_t) * ck.size(), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    int cnt = 0;
    for (int c = 0; c < ncell; c++)
        for (int k = 0; k < cc[c]; k++, cnt++)
            if (xys && cnt < cap) {
                const uint32_t p = ck[(size_t)c * L.cellCap + k];
                xys[3 * cnt] = kp_x(p); xys[3 * cnt + 1] = kp_y(p); xys[3 * cnt + 2] = kp_score(p);
            }
    *n = cnt;
    return (xys && cnt > cap) ? fail(ORBFE_ERR_CAPACITY, "raw keypoint capacity too small") : ORBFE_OK;
}

}  // extern "C"
