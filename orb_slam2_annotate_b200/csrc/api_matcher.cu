// api_matcher.cu — matcher half of the C ABI (include/orbfe.h): handle, scalar Hamming
// distance and the host/device entry points of the brute-force kNN-2 search.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "api_common.cuh"
#include "frame_internal.cuh"
#include "match_internal.cuh"

using namespace orbfe;

struct orbfe_matcher {
    int device = 0;
    cudaStream_t stream = nullptr;
    DevArray<uint8_t> scratch[8];   // grow-only device staging, reused across calls
    PinnedArray stage, result;      // page-locked host staging of the guided searches
    bool trace = std::getenv("ORBFE_TRACE") != nullptr;
    int knnEngine = ORBFE_KNN_AUTO;   // orbfe_matcher_set_knn_engine
    DevArray<uint8_t> knnScratch;   // per-segment partial results of the tensor-core kNN engine
};

namespace {

// which brute-force engine serves a kNN call: the tensor-core kernel needs enough work to fill its pipeline
// which engine serves a kNN-2 call: the handle's choice, else ORBFE_KNN_ENGINE, else by size (the tensor cores from
// 2^26 comparisons upwards: below that the launch of the pre-pass / merge kernels and the half-empty waves cost more
// than the popc kernel's whole run)
int knn_engine(const orbfe_matcher* m, int nq, int nt) {
    static const char* env = std::getenv("ORBFE_KNN_ENGINE");
    int eng = m->knnEngine;
    if (eng == ORBFE_KNN_AUTO && env)
        eng = !std::strcmp(env, "popc") ? ORBFE_KNN_POPC : !std::strcmp(env, "mma") ? ORBFE_KNN_MMA : !std::strcmp(env, "fp4") ? ORBFE_KNN_FP4 : eng;
    if (eng != ORBFE_KNN_AUTO) return eng;
    return nq >= 256 && (long long)nq * nt >= (1ll << 26) ? ORBFE_KNN_FP4 : ORBFE_KNN_POPC;
}

int knn2_dispatch(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                  int* d_idx, int* d_d1, int* d_d2) {
    const int eng = nq > 0 ? knn_engine(m, nq, nt) : ORBFE_KNN_POPC;
    if (eng == ORBFE_KNN_FP4) {
        int rc = m->knnScratch.ensure(knn2_fp4_scratch_bytes(nq, std::max(nt, 1)) + 256);
        if (rc) return rc;
        if (launch_knn2_fp4(d_query, nq, d_train, nt, th, ratio, d_idx, d_d1, d_d2, m->knnScratch.p, m->stream) < 0) return ORBFE_ERR_CUDA;
    } else if (eng == ORBFE_KNN_MMA) {
        int rc = m->knnScratch.ensure(knn2_mma_scratch_bytes(nq, std::max(nt, 1)) + 256);
        if (rc) return rc;
        launch_knn2_mma(d_query, nq, d_train, nt, th, ratio, d_idx, d_d1, d_d2, m->knnScratch.p, m->stream);
    } else {
        launch_knn2(d_query, nq, d_train, nt, th, ratio, d_idx, d_d1, d_d2, m->stream);
    }
    CU(cudaGetLastError());
    return ORBFE_OK;
}

}  // namespace

extern "C" {

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1828-1844): 8 x 32-bit XOR + popcount
int orbfe_hamming256(const void* a, const void* b) {
    const unsigned char* pa = (const unsigned char*)a;
    const unsigned char* pb = (const unsigned char*)b;
    int dist = 0;
    for (int i = 0; i < 4; i++) {
        unsigned long long x, y;
        std::memcpy(&x, pa + 8 * i, 8);
        std::memcpy(&y, pb + 8 * i, 8);
        dist += __builtin_popcountll(x ^ y);
    }
    return dist;
}

int orbfe_matcher_create(int device, orbfe_matcher** out) {
    if (!out) return fail(ORBFE_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (device < 0 || orbfe_device_count() <= device)
        return fail(ORBFE_ERR_NO_DEVICE, "no CUDA device: liborbfe has no CPU fallback");
    orbfe_matcher* m = new (std::nothrow) orbfe_matcher();
    if (!m) return fail(ORBFE_ERR_ARG, "out of host memory");
    m->device = device;
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (ce != cudaSuccess) { delete m; return fail(ORBFE_ERR_CUDA, "cudaStreamCreate", ce); }
    *out = m;
    return ORBFE_OK;
}

void orbfe_matcher_destroy(orbfe_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) { cudaStreamSynchronize(m->stream); cudaStreamDestroy(m->stream); }
    for (auto& s : m->scratch) s.release();
    m->knnScratch.release();
    m->stage.release(); m->result.release();
    delete m;
}

int orbfe_matcher_fetch(orbfe_matcher* m, void* dst, const void* d_src, size_t bytes) {
    if (!m || (bytes && (!dst || !d_src))) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_matcher_fetch");
    if (!bytes) return ORBFE_OK;
    CU(cudaSetDevice(m->device));
    CU(cudaMemcpyAsync(dst, d_src, bytes, cudaMemcpyDeviceToHost, m->stream));
    return ORBFE_OK;
}

int orbfe_matcher_set_knn_engine(orbfe_matcher* m, int engine) {
    if (!m || engine < ORBFE_KNN_AUTO || engine > ORBFE_KNN_FP4) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_matcher_set_knn_engine");
    m->knnEngine = engine;
    return ORBFE_OK;
}

void* orbfe_matcher_stream(orbfe_matcher* m) { return m ? (void*)m->stream : nullptr; }

int orbfe_matcher_sync(orbfe_matcher* m) {
    if (!m) return fail(ORBFE_ERR_ARG, "matcher is NULL");
    CU(cudaSetDevice(m->device));
    CU(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}

int orbfe_knn2_device(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt,
                      int th, float ratio, int32_t* d_out_idx, int32_t* d_out_d1, int32_t* d_out_d2) {
    if (!m || nq < 0 || nt < 0 || (nq && (!d_query || !d_out_idx || !d_out_d1 || !d_out_d2)) || (nt && !d_train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2_device");
    if (((uintptr_t)d_query | (uintptr_t)d_train) & 15) return fail(ORBFE_ERR_ARG, "descriptor arrays must be 16-byte aligned");
    CU(cudaSetDevice(m->device));
    return knn2_dispatch(m, d_query, nq, d_train, nt, th, ratio, d_out_idx, d_out_d1, d_out_d2);
}

int orbfe_knn2(orbfe_matcher* m, const uint8_t* query, int nq, const uint8_t* train, int nt, int th,
               float ratio, int32_t* out_idx, int32_t* out_d1, int32_t* out_d2) {
    if (!m || nq < 0 || nt < 0 || (nq && (!query || !out_idx || !out_d1 || !out_d2)) || (nt && !train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2");
    if (nq == 0) return ORBFE_OK;
    CU(cudaSetDevice(m->device));
    int rc;
    if ((rc = m->scratch[0].ensure((size_t)nq * 32))) return rc;
    if ((rc = m->scratch[1].ensure((size_t)std::max(nt, 1) * 32))) return rc;
    if ((rc = m->scratch[2].ensure((size_t)nq * 12))) return rc;
    int* d_idx = (int*)m->scratch[2].p;
    int* d_d1 = d_idx + nq;
    int* d_d2 = d_d1 + nq;
    CU(cudaMemcpyAsync(m->scratch[0].p, query, (size_t)nq * 32, cudaMemcpyHostToDevice, m->stream));
    if (nt) CU(cudaMemcpyAsync(m->scratch[1].p, train, (size_t)nt * 32, cudaMemcpyHostToDevice, m->stream));
    if ((rc = knn2_dispatch(m, m->scratch[0].p, nq, m->scratch[1].p, nt, th, ratio, d_idx, d_d1, d_d2))) return rc;
    CU(cudaMemcpyAsync(out_idx, d_idx, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaMemcpyAsync(out_d1, d_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaMemcpyAsync(out_d2, d_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, m->stream));
    CU(cudaStreamSynchronize(m->stream));
    return ORBFE_OK;
}

}  // extern "C"

namespace {

// bump allocator over one grow-only device buffer (256-byte aligned pieces)
struct Arena {
    uint8_t* base = nullptr;
    size_t off = 0, cap = 0;
    template <class T> T* take(size_t n) {
        off = (off + 255) & ~size_t(255);
        T* p = reinterpret_cast<T*>(base + off);
        off += n * sizeof(T);
        return p;
    }
};
inline size_t padded(size_t bytes) { return ((bytes + 255) & ~size_t(255)) + 256; }

int check_frame(const orbfe_frame_view* f) {
    if (!f || f->n < 0 || (f->n && (!f->kps || !f->desc))) return fail(ORBFE_ERR_ARG, "bad frame view");
    if (f->n > 65000) return fail(ORBFE_ERR_ARG, "more than 65000 keypoints in a frame view");
    if (!(f->max_x > f->min_x) || !(f->max_y > f->min_y)) return fail(ORBFE_ERR_ARG, "empty image bounds in frame view");
    return 0;
}

GridGeom grid_of(const orbfe_frame_view* f) {
    // mfGridElementWidthInv / HeightInv (src/Frame.cc:86-87)
    GridGeom g;
    g.minX = f->min_x; g.minY = f->min_y;
    g.qminX = f->keyframe_bounds ? (float)(int)f->min_x : f->min_x;     // KeyFrame.cc:41: int members
    g.qminY = f->keyframe_bounds ? (float)(int)f->min_y : f->min_y;
    g.invW = f->grid_inv_w > 0.0f ? f->grid_inv_w : 64.0f / (f->max_x - f->min_x);
    g.invH = f->grid_inv_h > 0.0f ? f->grid_inv_h : 48.0f / (f->max_y - f->min_y);
    return g;
}

}  // namespace

extern "C" {

int orbfe_search_init(orbfe_matcher* m, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                      float* prev_matched, int32_t* matches12, int window, float nnratio,
                      int check_orientation, int* nmatches) {
    if (!m || !nmatches) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_init");
    int rc;
    if ((rc = check_frame(f1)) || (rc = check_frame(f2))) return rc;
    *nmatches = 0;
    const int n1 = f1->n, n2 = f2->n;
    if (n1 == 0) return ORBFE_OK;
    if (!prev_matched || !matches12) return fail(ORBFE_ERR_ARG, "prev_matched / matches12 is NULL");
    CU(cudaSetDevice(m->device));
    const int cap = std::max(std::max(n1, n2), 1);
    const size_t total = padded((size_t)2 * cap * sizeof(orbfe_keypoint)) + padded((size_t)2 * cap * 32) +
                         padded(8) + padded((size_t)cap * 8) + padded((size_t)cap * 4) + padded(4);
    if ((rc = m->scratch[0].ensure(total))) return rc;
    Arena A; A.base = m->scratch[0].p;
    orbfe_keypoint* d_kps = A.take<orbfe_keypoint>((size_t)2 * cap);
    uint8_t* d_desc = A.take<uint8_t>((size_t)2 * cap * 32);
    int* d_counts = A.take<int>(2);
    float* d_prev = A.take<float>((size_t)cap * 2);
    int* d_m12 = A.take<int>(cap);
    int* d_nm = A.take<int>(1);
    if (A.off > total) return fail(ORBFE_ERR_CAPACITY, "internal: matcher arena overflow");
    cudaStream_t s = m->stream;
    const int counts[2] = {n1, n2};
    CU(cudaMemcpyAsync(d_kps, f1->kps, sizeof(orbfe_keypoint) * n1, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_desc, f1->desc, (size_t)32 * n1, cudaMemcpyHostToDevice, s));
    if (n2) {
        CU(cudaMemcpyAsync(d_kps + cap, f2->kps, sizeof(orbfe_keypoint) * n2, cudaMemcpyHostToDevice, s));
        CU(cudaMemcpyAsync(d_desc + (size_t)cap * 32, f2->desc, (size_t)32 * n2, cudaMemcpyHostToDevice, s));
    }
    CU(cudaMemcpyAsync(d_counts, counts, sizeof(counts), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_prev, prev_matched, sizeof(float) * 2 * n1, cudaMemcpyHostToDevice, s));
    InitArgs a{};
    a.kps = d_kps; a.desc = d_desc; a.counts = d_counts; a.cap = cap; a.pairs = nullptr;
    a.prev = d_prev; a.m12 = d_m12; a.nmatches = d_nm;
    a.grid = grid_of(f2);
    a.window = window; a.ratio = nnratio; a.checkOri = check_orientation;
    if ((rc = m->scratch[3].ensure(search_init_scratch_bytes(cap, 1)))) return rc;
    search_init_carve(a, m->scratch[3].p, 1);
    launch_search_init(a, 1, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(prev_matched, d_prev, sizeof(float) * 2 * n1, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(matches12, d_m12, sizeof(int) * n1, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(nmatches, d_nm, sizeof(int), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return ORBFE_OK;
}

int orbfe_search_init_consecutive_device(orbfe_matcher* m, const orbfe_keypoint* d_kps,
                                         const uint8_t* d_desc, const int32_t* d_counts, int cap,
                                         int nframes, float min_x, float max_x, float min_y, float max_y,
                                         int window, float nnratio, int check_orientation,
                                         int32_t* d_matches12, int32_t* d_nmatches) {
    if (!m || !d_kps || !d_desc || !d_counts || !d_matches12 || !d_nmatches || cap < 1 || cap > 65000 ||
        !(max_x > min_x) || !(max_y > min_y))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_init_consecutive_device");
    if (nframes < 2) return ORBFE_OK;
    CU(cudaSetDevice(m->device));
    InitArgs a{};
    a.kps = d_kps; a.desc = d_desc; a.counts = d_counts; a.cap = cap; a.pairs = nullptr;
    a.prev = nullptr; a.m12 = d_matches12; a.nmatches = d_nmatches;
    a.grid.minX = a.grid.qminX = min_x; a.grid.minY = a.grid.qminY = min_y;
    a.grid.invW = 64.0f / (max_x - min_x); a.grid.invH = 48.0f / (max_y - min_y);
    a.window = window; a.ratio = nnratio; a.checkOri = check_orientation;
    int rc;
    if ((rc = m->scratch[3].ensure(search_init_scratch_bytes(cap, nframes - 1)))) return rc;
    search_init_carve(a, m->scratch[3].p, nframes - 1);
    launch_search_init(a, nframes - 1, m->stream);
    CU(cudaGetLastError());
    return ORBFE_OK;
}

}  // extern "C"

namespace {

// host view standing in for a device-resident frame: sizes, bounds and scale factors are read
// from it, its kps / desc / uright pointers are only tested for NULL
orbfe_frame_view host_side_of(const orbfe_device_frame_view* df) {
    orbfe_frame_view f{};
    f.n = df->n; f.kps = df->d_kps; f.desc = df->d_desc; f.uright = df->d_uright;
    f.min_x = df->min_x; f.max_x = df->max_x; f.min_y = df->min_y; f.max_y = df->max_y;
    f.scale_factors = df->scale_factors; f.nlevels = df->nlevels; f.keyframe_bounds = df->keyframe_bounds;
    f.grid_inv_w = df->grid_inv_w; f.grid_inv_h = df->grid_inv_h;
    return f;
}

int check_device_frame(const orbfe_device_frame_view* df) {
    if (!df) return fail(ORBFE_ERR_ARG, "device frame view is NULL");
    if (df->n > 0 && (!df->d_kps || !df->d_desc || !df->d_cell_end || !df->d_items))
        return fail(ORBFE_ERR_ARG, "NULL device pointer in device frame view");
    if ((uintptr_t)df->d_desc & 15) return fail(ORBFE_ERR_ARG, "device descriptors must be 16-byte aligned");
    return 0;
}

// ---- the two guided searches share one host schedule --------------------------------------
// Everything the call uploads is packed into one pinned staging block and sent with a single
// H2D copy; the kernels of a whole round (grid, candidate count / scan / fill, kSweeps
// fixed-point sweeps, epilogue) are enqueued without looking at intermediate results, and one
// D2H copy brings back the answers together with the two facts that validate them: the
// candidate total (did the list fit the buffer?) and the "changed" flag of the last sweep (did
// the first-owner fixed point converge?).  Extra sweeps past convergence change nothing, so the
// common case costs one synchronisation; otherwise the round is repeated from where it stopped.
constexpr int kSweeps = 2;

struct Upload {
    uint8_t* host; uint8_t* dev; size_t off = 0, cap = 0;
    template <class T> T* put(const void* src, size_t n) {       // copies into the staging block
        off = (off + 255) & ~size_t(255);
        T* d = reinterpret_cast<T*>(dev + off);
        if (src && n) std::memcpy(host + off, src, n * sizeof(T));
        off += n * sizeof(T);
        return d;
    }
};

// df != NULL: the frame side (keypoints, descriptors, uright, grid) is already in device memory.
// Exactly one of mps (mode 0, SearchByProjection(F, MapPoints)) / pts (mode 1, (Current, Last)).
int guided_search(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_device_frame_view* df,
                  const orbfe_mappoints_view* mps, const orbfe_framepoints_view* pts, float th, float nnratio,
                  int check_orientation, const uint8_t* occ_in, uint8_t* occ_out_host, int32_t* out_for_kp,
                  int* nmatches, int max_dist = 100) {
    const int mode = pts ? 1 : 0;
    const int n = f->n, np = mode ? pts->n : mps->n, L = f->nlevels;
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    int rc;
    // ---- uploads
    const size_t frameBytes = df ? 0 : padded((size_t)n * sizeof(orbfe_keypoint)) + padded((size_t)n * 32) + padded((size_t)n * 4);
    const size_t upBytes = frameBytes + padded((size_t)std::max(L, 1) * 4) + padded(n) + padded(np) + 8 * padded((size_t)np * 4) +
                           padded((size_t)np * 32);
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    LocalArgs a{};
    if (df) {
        a.kps = df->d_kps; a.desc = df->d_desc; a.uright = df->d_uright;
    } else {
        a.kps = U.put<orbfe_keypoint>(f->kps, n);
        a.desc = U.put<uint8_t>(f->desc, (size_t)n * 32);
        a.uright = f->uright ? U.put<float>(f->uright, n) : nullptr;
    }
    a.n = n; a.nlevels = L; a.nmp = np; a.grid = grid_of(f); a.mode = mode;
    const uint8_t* d_occ = U.put<uint8_t>(occ_in, n);
    a.occ0 = d_occ;
    if (mode == 0) {
        a.scale = U.put<float>(f->scale_factors, L);
        a.in_view = U.put<uint8_t>(mps->track_in_view, np);
        a.px = U.put<float>(mps->proj_x, np); a.py = U.put<float>(mps->proj_y, np);
        a.pxr = a.uright ? U.put<float>(mps->proj_xr, np) : nullptr;
        a.level = U.put<int>(mps->level, np); a.view_cos = U.put<float>(mps->view_cos, np);
        a.mpdesc = U.put<uint8_t>(mps->desc, (size_t)np * 32);
        a.obs = U.put<int>(mps->observations, np);
        a.th = th; a.ratio = nnratio;
    } else {
        a.in_view = U.put<uint8_t>(pts->valid, np);
        a.px = U.put<float>(pts->u, np); a.py = U.put<float>(pts->v, np);
        a.pxr = a.uright ? U.put<float>(pts->ur, np) : nullptr;
        a.radius = U.put<float>(pts->radius, np);
        a.min_level = U.put<int>(pts->min_level, np); a.max_level = U.put<int>(pts->max_level, np);
        a.angle = U.put<float>(pts->angle, np);
        a.mpdesc = U.put<uint8_t>(pts->desc, (size_t)np * 32);
        a.obs = U.put<int>(pts->observations, np);
        a.th = 1.0f; a.ratio = 1.0f; a.checkOri = check_orientation; a.max_dist = max_dist;
    }
    if (U.off > U.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher staging overflow");
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    // ---- device scratch.  Result block R = [mpkp n | kp_bad n | flags kSweeps | nmatches | ncand | hist 32] ints + occ_out n bytes
    const size_t rInts = (size_t)2 * n + kSweeps + 2 + 32;
    const size_t rBytes = rInts * 4 + (size_t)n;
    const size_t scBytes = padded(3072 * 4) + padded((size_t)n * 2) + padded((size_t)(np + 1) * 4) +
                           padded((size_t)(kSweeps + 1) * n * 4) + padded((size_t)np * 4) + padded(rBytes);
    if ((rc = m->scratch[2].ensure(scBytes))) return rc;
    if ((rc = m->result.ensure(rBytes))) return rc;
    Arena A; A.base = m->scratch[2].p; A.cap = scBytes;
    if (df) {
        a.cell_end = const_cast<int*>(df->d_cell_end); a.items = const_cast<unsigned short*>(df->d_items);
    } else {
        a.cell_end = A.take<int>(3072); a.items = A.take<unsigned short>(n);
    }
    a.cand_off = A.take<int>(np + 1);
    int* d_owner = A.take<int>((size_t)(kSweeps + 1) * n);
    a.choice = A.take<int>(np);
    int* R = A.take<int>(rInts + (n + 3) / 4);
    int* d_mpkp = R; a.kp_bad = R + n;
    int* d_flags = R + 2 * n; int* d_nm = d_flags + kSweeps; int* d_ncand = d_nm + 1; a.hist = d_ncand + 1;
    uint8_t* d_occ_out = reinterpret_cast<uint8_t*>(R + rInts);
    if (A.off > A.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher arena overflow");
    // candidate buffer: start from what earlier calls needed
    if ((rc = m->scratch[1].ensure(std::max<size_t>((size_t)np * 32, 4096)))) return rc;

    a.ncand_out = d_ncand;
    launch_local_grid_and_count(a, s, df == nullptr);
    bool refill = true, fresh = true;
    const int* hR = reinterpret_cast<const int*>(m->result.p);
    for (int round = 0;; round++) {
        if (refill) {
            a.cand = reinterpret_cast<uint32_t*>(m->scratch[1].p);
            a.cand_cap = (int)std::min<size_t>(m->scratch[1].cap / 4, 0x7FFFFFFF);
            launch_local_fill(a, s);
            refill = false;
        }
        // owners to "no owner" (the first one only when starting over), results cleared
        launch_guided_init(d_owner, n, kSweeps, fresh, d_mpkp, a.kp_bad, (int)(rInts - n), d_occ_out, d_occ, s);
        fresh = false;
        for (int k = 0; k < kSweeps; k++) {
            a.owner_prev = d_owner + (size_t)k * n; a.owner_next = d_owner + (size_t)(k + 1) * n;
            a.owner_pp = k ? d_owner + (size_t)(k - 1) * n : nullptr;     // sweep k checks whether sweep k-1 moved anything
            a.diff_flag = k ? d_flags + (k - 1) : nullptr;
            launch_local_sweep(a, s);
        }
        // choice belongs to the last sweep's input, which equals its output at the fixed point
        a.owner_prev = d_owner + (size_t)kSweeps * n;
        a.owner_pp = d_owner + (size_t)(kSweeps - 1) * n; a.diff_flag = d_flags + (kSweeps - 1);
        if (mode == 0) launch_local_finalize(a, d_mpkp, d_occ_out, d_nm, s);
        else launch_frame_finalize(a, d_mpkp, d_nm, s);
        CU(cudaGetLastError());
        CU(cudaMemcpyAsync(m->result.p, R, rBytes, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        const int ncand = hR[2 * n + kSweeps + 1];
        if (ncand > a.cand_cap) {                                           // list did not fit: grow, start over
            if ((rc = m->scratch[1].ensure((size_t)ncand * 4 + 4096))) return rc;
            refill = true; fresh = true;
            continue;
        }
        if (!hR[2 * n + kSweeps - 1]) break;                                // last sweep changed nothing
        if (round > np / kSweeps + 2) return fail(ORBFE_ERR_CUDA, "internal: first-owner sweeps did not converge");
        CU(cudaMemcpyAsync(d_owner, d_owner + (size_t)kSweeps * n, sizeof(int) * n, cudaMemcpyDeviceToDevice, s));
    }
    *nmatches = hR[2 * n + kSweeps];
    if (m->trace) std::fprintf(stderr, "orbfe guided_search mode %d: n %d points %d candidates %d flags %d %d\n", mode, n, np,
                               hR[2 * n + kSweeps + 1], hR[2 * n], hR[2 * n + 1]);
    const int* h_mpkp = hR; const int* h_bad = hR + n;
    if (mode == 0) {
        std::memcpy(occ_out_host, reinterpret_cast<const uint8_t*>(hR + rInts), n);
        for (int i = 0; i < n; i++)
            if (h_mpkp[i] >= 0) out_for_kp[i] = h_mpkp[i];   // F.mvpMapPoints[bestIdx] = pMP (:132); others untouched
    } else {
        for (int i = 0; i < n; i++) {
            if (h_bad[i]) out_for_kp[i] = -2;                 // mvpMapPoints[idx] = NULL (:1621)
            else if (h_mpkp[i] >= 0) out_for_kp[i] = h_mpkp[i];   // mvpMapPoints[bestIdx2] = pMP (:1587)
        }
    }
    return ORBFE_OK;
}

// ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) — src/ORBmatcher.cc:51-138
int search_local_points_impl(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_device_frame_view* df,
                             const orbfe_mappoints_view* mps, float th, float nnratio, uint8_t* kp_occupied,
                             int32_t* mp_for_kp, int* nmatches) {
    if (!m || !mps || !nmatches) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_local_points");
    int rc;
    if ((rc = check_frame(f))) return rc;
    *nmatches = 0;
    const int n = f->n, nmp = mps->n;
    if (n == 0 || nmp <= 0) return ORBFE_OK;
    if (!kp_occupied || !mp_for_kp || !f->scale_factors || !mps->track_in_view || !mps->proj_x || !mps->proj_y ||
        !mps->level || !mps->view_cos || !mps->desc || !mps->observations || (f->uright && !mps->proj_xr))
        return fail(ORBFE_ERR_ARG, "NULL array in frame / map point view");
    // levels must index mvScaleFactors
    for (int i = 0; i < nmp; i++)
        if (mps->track_in_view[i] && (mps->level[i] < 0 || mps->level[i] >= f->nlevels))
            return fail(ORBFE_ERR_ARG, "map point scale level outside the frame's pyramid");
    return guided_search(m, f, df, mps, nullptr, th, nnratio, 0, kp_occupied, kp_occupied, mp_for_kp, nmatches);
}

// ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) — src/ORBmatcher.cc:1484-1630
int search_frame_points_impl(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_device_frame_view* df,
                             const orbfe_framepoints_view* pts, int check_orientation, const uint8_t* kp_occupied,
                             int32_t* match_for_kp, int* nmatches, int max_dist = 100 /* TH_HIGH, :1590 */) {
    if (!m || !pts || !nmatches) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_frame_points");
    int rc;
    if ((rc = check_frame(f))) return rc;
    *nmatches = 0;
    const int n = f->n, np = pts->n;
    if (n == 0 || np <= 0) return ORBFE_OK;
    if (!kp_occupied || !match_for_kp || !pts->valid || !pts->u || !pts->v || !pts->radius || !pts->min_level ||
        !pts->max_level || !pts->angle || !pts->desc || !pts->observations || (f->uright && !pts->ur))
        return fail(ORBFE_ERR_ARG, "NULL array in frame / frame points view");
    if (max_dist < 0 || max_dist > 256) return fail(ORBFE_ERR_ARG, "max_dist outside 0..256");
    return guided_search(m, f, df, nullptr, pts, 1.0f, 1.0f, check_orientation, kp_occupied, nullptr, match_for_kp, nmatches, max_dist);
}

}  // namespace

extern "C" {

int orbfe_search_local_points(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_mappoints_view* mps,
                              float th, float nnratio, uint8_t* kp_occupied, int32_t* mp_for_kp, int* nmatches) {
    return search_local_points_impl(m, f, nullptr, mps, th, nnratio, kp_occupied, mp_for_kp, nmatches);
}

int orbfe_search_local_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* df,
                                       const orbfe_mappoints_view* mps, float th, float nnratio,
                                       uint8_t* kp_occupied, int32_t* mp_for_kp, int* nmatches) {
    int rc;
    if ((rc = check_device_frame(df))) return rc;
    const orbfe_frame_view f = host_side_of(df);
    return search_local_points_impl(m, &f, df, mps, th, nnratio, kp_occupied, mp_for_kp, nmatches);
}

int orbfe_search_frame_points(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_framepoints_view* pts,
                              int check_orientation, const uint8_t* kp_occupied, int32_t* match_for_kp,
                              int* nmatches) {
    return search_frame_points_impl(m, f, nullptr, pts, check_orientation, kp_occupied, match_for_kp, nmatches);
}

int orbfe_search_frame_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* df,
                                       const orbfe_framepoints_view* pts, int check_orientation,
                                       const uint8_t* kp_occupied, int32_t* match_for_kp, int* nmatches) {
    int rc;
    if ((rc = check_device_frame(df))) return rc;
    const orbfe_frame_view f = host_side_of(df);
    return search_frame_points_impl(m, &f, df, pts, check_orientation, kp_occupied, match_for_kp, nmatches);
}

// SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) — src/ORBmatcher.cc:1641-1770 — and
// SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) — :335-451: the first-owner search of
// orbfe_search_frame_points with the accepted distance as an argument
int orbfe_search_projected_points(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_framepoints_view* pts,
                                  int max_dist, int check_orientation, const uint8_t* kp_occupied,
                                  int32_t* match_for_kp, int* nmatches) {
    return search_frame_points_impl(m, f, nullptr, pts, check_orientation, kp_occupied, match_for_kp, nmatches, max_dist);
}

int orbfe_search_projected_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* df,
                                           const orbfe_framepoints_view* pts, int max_dist, int check_orientation,
                                           const uint8_t* kp_occupied, int32_t* match_for_kp, int* nmatches) {
    int rc;
    if ((rc = check_device_frame(df))) return rc;
    const orbfe_frame_view f = host_side_of(df);
    return search_frame_points_impl(m, &f, df, pts, check_orientation, kp_occupied, match_for_kp, nmatches, max_dist);
}

}  // extern "C"

namespace {

int check_window_points(const orbfe_frame_view* f, const orbfe_windowpoints_view* pts, bool gate) {
    if (!pts || pts->n < 0) return fail(ORBFE_ERR_ARG, "bad window points view");
    if (pts->n && (!pts->valid || !pts->u || !pts->v || !pts->radius || !pts->level || !pts->desc))
        return fail(ORBFE_ERR_ARG, "NULL array in window points view");
    if (pts->n && gate && f->uright && !pts->ur) return fail(ORBFE_ERR_ARG, "window points need ur for a frame with uright");
    return 0;
}

// frame side of a window search: uploaded (and its grid built) unless it is device resident
struct FrameSide {
    const orbfe_keypoint* kps; const uint8_t* desc; const float* uright; const int* cell_end; const unsigned short* items;
};
size_t frame_side_bytes(const orbfe_frame_view* f, const orbfe_device_frame_view* df) {
    return df ? 0 : padded((size_t)f->n * sizeof(orbfe_keypoint)) + padded((size_t)f->n * 32) + padded((size_t)f->n * 4);
}
size_t window_points_bytes(int np) { return padded(np) + 4 * padded((size_t)np * 4) + padded((size_t)np * 4) + padded((size_t)np * 32); }

void put_window_points(Upload& U, const orbfe_windowpoints_view* pts, bool with_ur, WindowArgs& a) {
    const int np = pts->n;
    a.np = np;
    a.valid = U.put<uint8_t>(pts->valid, np);
    a.u = U.put<float>(pts->u, np); a.v = U.put<float>(pts->v, np);
    a.ur = with_ur ? U.put<float>(pts->ur, np) : nullptr;
    a.radius = U.put<float>(pts->radius, np);
    a.level = U.put<int>(pts->level, np);
    a.pdesc = U.put<uint8_t>(pts->desc, (size_t)np * 32);
}

// Fuse (both overloads): every point's best keypoint inside its window
int search_window_best_impl(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_device_frame_view* df,
                            const orbfe_windowpoints_view* pts, const float* inv_level_sigma2, int max_dist,
                            int32_t* best_idx, int32_t* best_dist) {
    if (!m) return fail(ORBFE_ERR_ARG, "matcher is NULL");
    int rc;
    if ((rc = check_frame(f))) return rc;
    if ((rc = check_window_points(f, pts, inv_level_sigma2 != nullptr))) return rc;
    const int n = f->n, np = pts->n, L = f->nlevels;
    if (np == 0) return ORBFE_OK;
    if (!best_idx) return fail(ORBFE_ERR_ARG, "best_idx is NULL");
    if (max_dist < 0 || max_dist > 256) return fail(ORBFE_ERR_ARG, "max_dist outside 0..256");
    if (inv_level_sigma2 && L <= 0) return fail(ORBFE_ERR_ARG, "nlevels must be set with inv_level_sigma2");
    if (n == 0) {
        for (int i = 0; i < np; i++) { best_idx[i] = -1; if (best_dist) best_dist[i] = 256; }
        return ORBFE_OK;
    }
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    const size_t upBytes = frame_side_bytes(f, df) + window_points_bytes(np) + padded((size_t)std::max(L, 1) * 4);
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    WindowArgs a{};
    if (df) { a.kps = df->d_kps; a.desc = df->d_desc; a.uright = df->d_uright; }
    else {
        a.kps = U.put<orbfe_keypoint>(f->kps, n);
        a.desc = U.put<uint8_t>(f->desc, (size_t)n * 32);
        a.uright = f->uright ? U.put<float>(f->uright, n) : nullptr;
    }
    a.n = n; a.grid = grid_of(f);
    put_window_points(U, pts, inv_level_sigma2 && a.uright, a);
    a.inv_sigma2 = inv_level_sigma2 ? U.put<float>(inv_level_sigma2, L) : nullptr;
    a.max_dist = max_dist;
    if (U.off > U.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher staging overflow");
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    const size_t scBytes = padded(3072 * 4) + padded((size_t)n * 2) + padded((size_t)np * 8);
    if ((rc = m->scratch[2].ensure(scBytes))) return rc;
    if ((rc = m->result.ensure((size_t)np * 8))) return rc;
    Arena A; A.base = m->scratch[2].p; A.cap = scBytes;
    if (df) { a.cell_end = df->d_cell_end; a.items = df->d_items; }
    else {
        int* ce = A.take<int>(3072); unsigned short* it = A.take<unsigned short>(n);
        launch_build_grid(a.kps, n, a.grid, ce, it, s);
        a.cell_end = ce; a.items = it;
    }
    int* R = A.take<int>((size_t)np * 2);
    a.best_idx = R; a.best_dist = R + np;
    launch_window_best(a, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, R, (size_t)np * 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    std::memcpy(best_idx, m->result.p, (size_t)np * 4);
    if (best_dist) std::memcpy(best_dist, m->result.p + (size_t)np * 4, (size_t)np * 4);
    return ORBFE_OK;
}

}  // namespace

extern "C" {

int orbfe_search_window_best(orbfe_matcher* m, const orbfe_frame_view* f, const orbfe_windowpoints_view* pts,
                             const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist) {
    return search_window_best_impl(m, f, nullptr, pts, inv_level_sigma2, max_dist, best_idx, best_dist);
}

int orbfe_search_window_best_resident(orbfe_matcher* m, const orbfe_device_frame_view* df,
                                      const orbfe_windowpoints_view* pts, const float* inv_level_sigma2, int max_dist,
                                      int32_t* best_idx, int32_t* best_dist) {
    int rc;
    if ((rc = check_device_frame(df))) return rc;
    const orbfe_frame_view f = host_side_of(df);
    return search_window_best_impl(m, &f, df, pts, inv_level_sigma2, max_dist, best_idx, best_dist);
}

// ORBmatcher::SearchBySim3 — src/ORBmatcher.cc:1251-1477: both directions and the agreement test in one go
int orbfe_search_by_sim3(orbfe_matcher* m, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                         const orbfe_windowpoints_view* pts1, const orbfe_windowpoints_view* pts2, int max_dist,
                         int32_t* matches12, int* nfound) {
    if (!m || !nfound) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_by_sim3");
    int rc;
    if ((rc = check_frame(f1)) || (rc = check_frame(f2))) return rc;
    if ((rc = check_window_points(f2, pts1, false)) || (rc = check_window_points(f1, pts2, false))) return rc;
    if (pts1->n != f1->n || pts2->n != f2->n)
        return fail(ORBFE_ERR_ARG, "SearchBySim3 takes one (possibly invalid) point per keypoint of each key frame");
    if (max_dist < 0 || max_dist > 256) return fail(ORBFE_ERR_ARG, "max_dist outside 0..256");
    *nfound = 0;
    const int n1 = f1->n, n2 = f2->n;
    if (n1 == 0) return ORBFE_OK;
    if (!matches12) return fail(ORBFE_ERR_ARG, "matches12 is NULL");
    if (n2 == 0) { for (int i = 0; i < n1; i++) matches12[i] = -1; return ORBFE_OK; }
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    const size_t upBytes = frame_side_bytes(f1, nullptr) + frame_side_bytes(f2, nullptr) + window_points_bytes(n1) + window_points_bytes(n2);
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    WindowArgs a12{}, a21{};        // a12: KF1's points searched in KF2; a21: KF2's points searched in KF1
    a12.kps = U.put<orbfe_keypoint>(f2->kps, n2); a12.desc = U.put<uint8_t>(f2->desc, (size_t)n2 * 32); a12.n = n2; a12.grid = grid_of(f2);
    a21.kps = U.put<orbfe_keypoint>(f1->kps, n1); a21.desc = U.put<uint8_t>(f1->desc, (size_t)n1 * 32); a21.n = n1; a21.grid = grid_of(f1);
    put_window_points(U, pts1, false, a12);
    put_window_points(U, pts2, false, a21);
    a12.max_dist = a21.max_dist = max_dist;
    if (U.off > U.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher staging overflow");
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    const size_t rInts = (size_t)n1 + 1;
    const size_t scBytes = 2 * padded(3072 * 4) + padded((size_t)n1 * 2) + padded((size_t)n2 * 2) + 2 * padded((size_t)n1 * 4) +
                           2 * padded((size_t)n2 * 4) + padded(rInts * 4);
    if ((rc = m->scratch[2].ensure(scBytes))) return rc;
    if ((rc = m->result.ensure(rInts * 4))) return rc;
    Arena A; A.base = m->scratch[2].p; A.cap = scBytes;
    int* ce2 = A.take<int>(3072); unsigned short* it2 = A.take<unsigned short>(n2);
    int* ce1 = A.take<int>(3072); unsigned short* it1 = A.take<unsigned short>(n1);
    a12.best_idx = A.take<int>(n1); a12.best_dist = A.take<int>(n1);
    a21.best_idx = A.take<int>(n2); a21.best_dist = A.take<int>(n2);
    int* R = A.take<int>(rInts);
    launch_build_grid(a12.kps, n2, a12.grid, ce2, it2, s); a12.cell_end = ce2; a12.items = it2;
    launch_build_grid(a21.kps, n1, a21.grid, ce1, it1, s); a21.cell_end = ce1; a21.items = it1;
    CU(cudaMemsetAsync(R + n1, 0, 4, s));
    launch_window_best(a12, s);
    launch_window_best(a21, s);
    launch_sim3_agree(a12.best_idx, a21.best_idx, n1, n2, R, R + n1, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, R, rInts * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    std::memcpy(matches12, m->result.p, (size_t)n1 * 4);
    *nfound = reinterpret_cast<const int*>(m->result.p)[n1];
    return ORBFE_OK;
}

// ORBmatcher::SearchForTriangulation — src/ORBmatcher.cc:754-928
int orbfe_search_triangulation(orbfe_matcher* m, const orbfe_tri_side* s1, const orbfe_tri_side* s2, const float F12[9],
                               float ex, float ey, const float* scale_factors2, const float* level_sigma2_2, int nlevels2,
                               int only_stereo, int check_orientation, int32_t* matches12, int* nmatches) {
    if (!m || !s1 || !s2 || !F12 || !nmatches) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_triangulation");
    *nmatches = 0;
    for (const orbfe_tri_side* sd : {s1, s2}) {
        if (sd->n < 0 || sd->nnodes < 0) return fail(ORBFE_ERR_ARG, "negative size in triangulation side");
        if (sd->n && (!sd->kps || !sd->desc || !sd->has_mappoint)) return fail(ORBFE_ERR_ARG, "NULL array in triangulation side");
        if (sd->nnodes && (!sd->node_id || !sd->node_start || !sd->indices)) return fail(ORBFE_ERR_ARG, "NULL feature vector array");
        if (sd->nnodes && sd->node_start[0] != 0) return fail(ORBFE_ERR_ARG, "node_start[0] must be 0");
        for (int k = 0; k < sd->nnodes; k++) {
            if (sd->node_start[k + 1] < sd->node_start[k]) return fail(ORBFE_ERR_ARG, "node_start must not decrease");
            if (sd->node_start[k + 1] - sd->node_start[k] > 65535) return fail(ORBFE_ERR_CAPACITY, "more than 65535 features in one node");
            if (k && sd->node_id[k] <= sd->node_id[k - 1]) return fail(ORBFE_ERR_ARG, "node ids must ascend");
        }
        const int tot = sd->nnodes ? sd->node_start[sd->nnodes] : 0;
        for (int j = 0; j < tot; j++)
            if (sd->indices[j] >= (uint32_t)sd->n) return fail(ORBFE_ERR_ARG, "feature index outside the key frame");
    }
    const int n1 = s1->n, n2 = s2->n;
    if (n1 == 0) return ORBFE_OK;
    if (!matches12) return fail(ORBFE_ERR_ARG, "matches12 is NULL");
    const int t1 = s1->nnodes ? s1->node_start[s1->nnodes] : 0, t2 = s2->nnodes ? s2->node_start[s2->nnodes] : 0;
    if (n2 == 0 || t1 == 0 || t2 == 0) { for (int i = 0; i < n1; i++) matches12[i] = -1; return ORBFE_OK; }
    if (!scale_factors2 || !level_sigma2_2 || nlevels2 <= 0) return fail(ORBFE_ERR_ARG, "scale tables of key frame 2 missing");
    for (int i = 0; i < n2; i++)
        if (s2->kps[i].octave < 0 || s2->kps[i].octave >= nlevels2) return fail(ORBFE_ERR_ARG, "key frame 2 octave outside its scale tables");
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    int rc;
    auto side_bytes = [](const orbfe_tri_side* sd, int tot) {
        return padded((size_t)sd->n * sizeof(orbfe_keypoint)) + padded((size_t)sd->n * 32) + padded(sd->n) + padded((size_t)sd->n * 4) +
               padded((size_t)sd->nnodes * 4) + padded((size_t)(sd->nnodes + 1) * 4) + padded((size_t)tot * 4);
    };
    const size_t upBytes = side_bytes(s1, t1) + side_bytes(s2, t2) + 2 * padded((size_t)nlevels2 * 4);
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    TriArgs a{};
    a.n1 = n1; a.n2 = n2; a.nidx1 = t1;
    a.kps1 = U.put<orbfe_keypoint>(s1->kps, n1); a.desc1 = U.put<uint8_t>(s1->desc, (size_t)n1 * 32);
    a.has1 = U.put<uint8_t>(s1->has_mappoint, n1); a.ur1 = s1->uright ? U.put<float>(s1->uright, n1) : nullptr;
    a.nnodes1 = s1->nnodes; a.node_id1 = U.put<uint32_t>(s1->node_id, s1->nnodes);
    a.node_start1 = U.put<int>(s1->node_start, (size_t)s1->nnodes + 1); a.indices1 = U.put<uint32_t>(s1->indices, t1);
    a.kps2 = U.put<orbfe_keypoint>(s2->kps, n2); a.desc2 = U.put<uint8_t>(s2->desc, (size_t)n2 * 32);
    a.has2 = U.put<uint8_t>(s2->has_mappoint, n2); a.ur2 = s2->uright ? U.put<float>(s2->uright, n2) : nullptr;
    a.nnodes2 = s2->nnodes; a.node_id2 = U.put<uint32_t>(s2->node_id, s2->nnodes);
    a.node_start2 = U.put<int>(s2->node_start, (size_t)s2->nnodes + 1); a.indices2 = U.put<uint32_t>(s2->indices, t2);
    a.scale2 = U.put<float>(scale_factors2, nlevels2); a.sigma2_2 = U.put<float>(level_sigma2_2, nlevels2);
    std::memcpy(a.F, F12, sizeof(a.F)); a.ex = ex; a.ey = ey;
    a.th = 50; a.onlyStereo = only_stereo; a.checkOri = check_orientation;                // TH_LOW (:816)
    if (U.off > U.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher staging overflow");
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    const size_t rInts = (size_t)n1 + 32;
    if ((rc = m->scratch[2].ensure(rInts * 4))) return rc;
    if ((rc = m->result.ensure(rInts * 4))) return rc;
    int* R = reinterpret_cast<int*>(m->scratch[2].p);
    a.m12 = R; a.hist = R + n1; a.nmatches = R + n1 + 30;
    launch_search_triangulation(a, R + n1, 32, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, R, rInts * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    std::memcpy(matches12, m->result.p, (size_t)n1 * 4);
    *nmatches = reinterpret_cast<const int*>(m->result.p)[n1 + 30];
    return ORBFE_OK;
}

// ORBmatcher::SearchByBoW, both forms — src/ORBmatcher.cc:185-325, :610-743 — for npairs (side 1, side 2)
// pairs in one go.  Sides that are the same object in several pairs (the current frame of a relocalisation
// against all its candidate key frames, Tracking.cc:1486-1510) are uploaded once.
int orbfe_search_bow_batch(orbfe_matcher* m, int npairs, const orbfe_bow_view* s1, const orbfe_bow_view* s2, int th,
                           float nnratio, int check_orientation, int32_t* const* matches12,
                           int32_t* const* matches21, int* nmatches) {
    if (!m || npairs < 0 || (npairs && (!s1 || !s2 || !nmatches))) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_bow_batch");
    if (npairs == 0) return ORBFE_OK;
    // distinct sides, by content of the view (same arrays = same side)
    std::vector<const orbfe_bow_view*> sides;
    std::vector<int> sideOf((size_t)2 * npairs);
    auto same = [](const orbfe_bow_view* a, const orbfe_bow_view* b) {
        return a->n == b->n && a->desc == b->desc && a->angle == b->angle && a->valid == b->valid && a->nnodes == b->nnodes &&
               a->node_id == b->node_id && a->node_start == b->node_start && a->indices == b->indices;
    };
    for (int p = 0; p < npairs; p++)
        for (int k = 0; k < 2; k++) {
            const orbfe_bow_view* v = k ? &s2[p] : &s1[p];
            int found = -1;
            for (int j = (int)sides.size() - 1; j >= 0 && j >= (int)sides.size() - 4; j--)      // neighbours repeat, if anything does
                if (same(sides[j], v)) { found = j; break; }
            if (found < 0) { found = (int)sides.size(); sides.push_back(v); }
            sideOf[2 * p + k] = found;
        }
    size_t upBytes = padded((size_t)npairs * sizeof(BowArgs));
    for (const orbfe_bow_view* v : sides) {
        if (v->n < 0 || v->nnodes < 0 || v->n > 65535) return fail(ORBFE_ERR_ARG, "bad feature / node count in bow view (at most 65535 features)");
        if (v->n && (!v->desc || !v->angle)) return fail(ORBFE_ERR_ARG, "NULL descriptor / angle array in bow view");
        if (v->nnodes && (!v->node_id || !v->node_start || (v->node_start[v->nnodes] && !v->indices)))
            return fail(ORBFE_ERR_ARG, "NULL feature vector array in bow view");
        if (v->nnodes && v->node_start[0] != 0) return fail(ORBFE_ERR_ARG, "bow view: node_start[0] must be 0");
        for (int k = 0; k < v->nnodes; k++) {
            if (v->node_start[k + 1] < v->node_start[k]) return fail(ORBFE_ERR_ARG, "bow view: node_start must not decrease");
            if (k && v->node_id[k] <= v->node_id[k - 1]) return fail(ORBFE_ERR_ARG, "bow view: node ids must ascend");
        }
        const int total = v->nnodes ? v->node_start[v->nnodes] : 0;
        for (int k = 0; k < total; k++)
            if (v->indices[k] >= (uint32_t)v->n) return fail(ORBFE_ERR_ARG, "bow view: feature index out of range");
        upBytes += padded((size_t)v->n * 32) + padded((size_t)v->n * 4) + padded(v->n) + padded((size_t)v->nnodes * 4) +
                   padded((size_t)(v->nnodes + 1) * 4) + padded((size_t)total * 4);
    }
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    int rc;
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    BowArgs* d_pairs = U.put<BowArgs>(nullptr, npairs);
    BowArgs* h_pairs = reinterpret_cast<BowArgs*>(m->stage.p + ((uint8_t*)d_pairs - m->scratch[0].p));
    struct DevSide { const uint8_t* desc; const float* angle; const uint8_t* valid; const uint32_t* node_id; const int* node_start; const uint32_t* indices; };
    std::vector<DevSide> dev(sides.size());
    for (size_t j = 0; j < sides.size(); j++) {
        const orbfe_bow_view* v = sides[j];
        const int total = v->nnodes ? v->node_start[v->nnodes] : 0;
        dev[j].desc = U.put<uint8_t>(v->desc, (size_t)v->n * 32);
        dev[j].angle = U.put<float>(v->angle, v->n);
        dev[j].valid = v->valid ? U.put<uint8_t>(v->valid, v->n) : nullptr;
        dev[j].node_id = U.put<uint32_t>(v->node_id, v->nnodes);
        dev[j].node_start = U.put<int>(v->nnodes ? v->node_start : nullptr, v->nnodes ? v->nnodes + 1 : 0);
        dev[j].indices = U.put<uint32_t>(v->indices, total);
    }
    if (U.off > U.cap) return fail(ORBFE_ERR_CAPACITY, "internal: matcher staging overflow");
    // result block: [m12 | m21 of every pair] [hist 30 + nmatches of every pair]
    size_t nMinus = 0;
    int maxNodes1 = 0, maxN1 = 0;
    for (int p = 0; p < npairs; p++) { nMinus += (size_t)s1[p].n + s2[p].n; maxNodes1 = std::max(maxNodes1, s1[p].nnodes); maxN1 = std::max(maxN1, s1[p].n); }
    const size_t rInts = nMinus + (size_t)31 * npairs;
    if (rInts > 0x7FFFFFFF) return fail(ORBFE_ERR_CAPACITY, "too many features in one SearchByBoW batch");
    if ((rc = m->scratch[2].ensure(rInts * 4))) return rc;
    if ((rc = m->result.ensure(rInts * 4))) return rc;
    int* R = reinterpret_cast<int*>(m->scratch[2].p);
    int* Z = R + nMinus;
    size_t off = 0;
    for (int p = 0; p < npairs; p++) {
        const DevSide& a1 = dev[sideOf[2 * p]]; const DevSide& a2 = dev[sideOf[2 * p + 1]];
        BowArgs a{};
        a.n1 = s1[p].n; a.n2 = s2[p].n; a.nnodes1 = s1[p].nnodes; a.nnodes2 = s2[p].nnodes;
        a.desc1 = a1.desc; a.angle1 = a1.angle; a.valid1 = a1.valid; a.node_id1 = a1.node_id; a.node_start1 = a1.node_start; a.indices1 = a1.indices;
        a.desc2 = a2.desc; a.angle2 = a2.angle; a.valid2 = a2.valid; a.node_id2 = a2.node_id; a.node_start2 = a2.node_start; a.indices2 = a2.indices;
        a.th = th; a.ratio = nnratio; a.checkOri = check_orientation;
        a.m12 = R + off; a.m21 = R + off + a.n1; off += (size_t)a.n1 + a.n2;
        a.hist = Z + 31 * p; a.nmatches = a.hist + 30;
        h_pairs[p] = a;
    }
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    launch_search_bow(d_pairs, npairs, maxNodes1, maxN1, check_orientation != 0, R, (int)nMinus, Z, 31 * npairs, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, R, rInts * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    const int* hR = reinterpret_cast<const int*>(m->result.p);
    off = 0;
    for (int p = 0; p < npairs; p++) {
        if (matches12 && matches12[p]) std::memcpy(matches12[p], hR + off, sizeof(int) * s1[p].n);
        if (matches21 && matches21[p]) std::memcpy(matches21[p], hR + off + s1[p].n, sizeof(int) * s2[p].n);
        off += (size_t)s1[p].n + s2[p].n;
        nmatches[p] = hR[nMinus + 31 * p + 30];
    }
    return ORBFE_OK;
}

int orbfe_search_bow(orbfe_matcher* m, const orbfe_bow_view* s1, const orbfe_bow_view* s2, int th, float nnratio,
                     int check_orientation, int32_t* matches12, int32_t* matches21, int* nmatches) {
    if (!m || !s1 || !s2 || !nmatches) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_search_bow");
    *nmatches = 0;
    return orbfe_search_bow_batch(m, 1, s1, s2, th, nnratio, check_orientation, &matches12, &matches21, nmatches);
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:266-334) for a batch of map points
int orbfe_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* start, int npoints,
                                  int32_t* best) {
    if (!m || npoints < 0 || (npoints && (!start || !best))) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_distinctive_descriptors");
    if (npoints == 0) return ORBFE_OK;
    if (start[0] != 0) return fail(ORBFE_ERR_ARG, "start[0] must be 0");
    for (int p = 0; p < npoints; p++) {
        if (start[p + 1] < start[p]) return fail(ORBFE_ERR_ARG, "start must not decrease");
        if (start[p + 1] - start[p] > 65535) return fail(ORBFE_ERR_CAPACITY, "more than 65535 observations of one map point");
    }
    const size_t total = (size_t)start[npoints];
    if (total && !desc) return fail(ORBFE_ERR_ARG, "desc is NULL");
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    int rc;
    const size_t upBytes = padded(total * 32) + padded((size_t)(npoints + 1) * 4);
    if ((rc = m->scratch[0].ensure(upBytes))) return rc;
    if ((rc = m->stage.ensure(upBytes))) return rc;
    if ((rc = m->scratch[2].ensure((size_t)npoints * 4))) return rc;
    if ((rc = m->result.ensure((size_t)npoints * 4))) return rc;
    Upload U{m->stage.p, m->scratch[0].p, 0, upBytes};
    const uint8_t* d_desc = U.put<uint8_t>(desc, total * 32);
    const int* d_start = U.put<int>(start, (size_t)npoints + 1);
    CU(cudaMemcpyAsync(m->scratch[0].p, m->stage.p, U.off, cudaMemcpyHostToDevice, s));
    int* d_best = reinterpret_cast<int*>(m->scratch[2].p);
    launch_distinctive(d_desc, d_start, npoints, d_best, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, d_best, (size_t)npoints * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    std::memcpy(best, m->result.p, (size_t)npoints * 4);
    return ORBFE_OK;
}

// Frame::ComputeImageBounds (src/Frame.cc:481-512)
int orbfe_image_bounds(orbfe_matcher* m, const orbfe_camera* cam, int cols, int rows, float bounds[4]) {
    if (!m || !cam || !bounds || cols <= 0 || rows <= 0 || !(cam->fx != 0.0f) || !(cam->fy != 0.0f))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_image_bounds");
    CU(cudaSetDevice(m->device));
    int rc;
    if ((rc = m->scratch[4].ensure(256))) return rc;
    return image_bounds_device(*cam, cols, rows, m->scratch[4].p, m->stream, bounds);
}

// Frame::UndistortKeyPoints + Frame::AssignFeaturesToGrid (src/Frame.cc:443-475, :246-267)
int orbfe_frame_assign(orbfe_matcher* m, const orbfe_camera* cam, const float bounds[4], const orbfe_keypoint* kps,
                       int n, orbfe_keypoint* kps_un, int32_t* cell_end, uint16_t* items, int* n_in_grid) {
    if (!m || !cam || !bounds || n < 0 || !cell_end || (n && (!kps || !kps_un || !items)))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_frame_assign");
    if (n > 65535) return fail(ORBFE_ERR_CAPACITY, "frame grid items are 16-bit: more than 65535 keypoints");
    if (!(bounds[1] > bounds[0]) || !(bounds[3] > bounds[2])) return fail(ORBFE_ERR_ARG, "empty image bounds");
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    const int nn = std::max(n, 1);
    const size_t total = 2 * padded((size_t)nn * sizeof(orbfe_keypoint)) + padded(3072 * 4) + padded((size_t)nn * 2) + padded(4);
    int rc;
    if ((rc = m->scratch[4].ensure(total))) return rc;
    Arena A; A.base = m->scratch[4].p; A.cap = total;
    orbfe_keypoint* d_kps = A.take<orbfe_keypoint>(nn);
    orbfe_keypoint* d_un = A.take<orbfe_keypoint>(nn);
    int* d_end = A.take<int>(3072);
    unsigned short* d_items = A.take<unsigned short>(nn);
    int* d_n = A.take<int>(1);
    if (n) CU(cudaMemcpyAsync(d_kps, kps, sizeof(orbfe_keypoint) * n, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_n, &n, sizeof(int), cudaMemcpyHostToDevice, s));
    launch_frame_epilogue(cam_geom(*cam, bounds), d_kps, d_n, nn, 0, 1, d_un, d_end, d_items, s);
    CU(cudaGetLastError());
    if (n) {
        CU(cudaMemcpyAsync(kps_un, d_un, sizeof(orbfe_keypoint) * n, cudaMemcpyDeviceToHost, s));
        CU(cudaMemcpyAsync(items, d_items, sizeof(uint16_t) * n, cudaMemcpyDeviceToHost, s));
    }
    CU(cudaMemcpyAsync(cell_end, d_end, sizeof(int) * 3072, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    if (n_in_grid) *n_in_grid = cell_end[3071];
    return ORBFE_OK;
}

}  // extern "C"

extern "C" int orbfe_stereo_match(orbfe_matcher* m, const orbfe_frame_view* left, const orbfe_frame_view* right,
                                  const orbfe_pyramid_view* pl, const orbfe_pyramid_view* pr,
                                  const float* inv_scale_factors, float mbf, float mb, float* uright,
                                  float* depth) {
    if (!m || !pl || !pr || !inv_scale_factors || !(mb > 0))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_stereo_match");
    int rc;
    if ((rc = check_frame(left)) || (rc = check_frame(right))) return rc;
    const int nl = left->n, nr = right->n, L = left->nlevels;
    if (nl == 0) return ORBFE_OK;
    if (!uright || !depth || !left->scale_factors || L < 1 || L > kMaxLevels || pl->nlevels < L || pr->nlevels < L)
        return fail(ORBFE_ERR_ARG, "bad pyramid / output arguments to orbfe_stereo_match");
    for (int i = 0; i < nl; i++)
        if (left->kps[i].octave < 0 || left->kps[i].octave >= L) return fail(ORBFE_ERR_ARG, "left keypoint octave outside the pyramid");
    for (int i = 0; i < nr; i++)
        if (right->kps[i].octave < 0 || right->kps[i].octave >= L) return fail(ORBFE_ERR_ARG, "right keypoint octave outside the pyramid");
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    size_t total = padded((size_t)nl * sizeof(orbfe_keypoint)) + padded((size_t)nl * 32) +
                   padded((size_t)std::max(nr, 1) * sizeof(orbfe_keypoint)) + padded((size_t)std::max(nr, 1) * 32) +
                   2 * padded((size_t)L * 4) + 3 * padded((size_t)nl * 4);
    for (int l = 0; l < L; l++)
        total += padded((size_t)pl->width[l] * pl->height[l]) + padded((size_t)pr->width[l] * pr->height[l]);
    if ((rc = m->scratch[0].ensure(total))) return rc;
    Arena A; A.base = m->scratch[0].p;
    StereoArgs a{};
    orbfe_keypoint* d_kl = A.take<orbfe_keypoint>(nl);
    uint8_t* d_dl = A.take<uint8_t>((size_t)nl * 32);
    orbfe_keypoint* d_kr = A.take<orbfe_keypoint>(std::max(nr, 1));
    uint8_t* d_dr = A.take<uint8_t>((size_t)std::max(nr, 1) * 32);
    float* d_scale = A.take<float>(L);
    float* d_inv = A.take<float>(L);
    a.uright = A.take<float>(nl); a.depth = A.take<float>(nl); a.sad = A.take<int>(nl);
    CU(cudaMemcpyAsync(d_kl, left->kps, sizeof(orbfe_keypoint) * nl, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_dl, left->desc, (size_t)nl * 32, cudaMemcpyHostToDevice, s));
    if (nr) {
        CU(cudaMemcpyAsync(d_kr, right->kps, sizeof(orbfe_keypoint) * nr, cudaMemcpyHostToDevice, s));
        CU(cudaMemcpyAsync(d_dr, right->desc, (size_t)nr * 32, cudaMemcpyHostToDevice, s));
    }
    CU(cudaMemcpyAsync(d_scale, left->scale_factors, sizeof(float) * L, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_inv, inv_scale_factors, sizeof(float) * L, cudaMemcpyHostToDevice, s));
    for (int l = 0; l < L; l++) {
        uint8_t* dl = A.take<uint8_t>((size_t)pl->width[l] * pl->height[l]);
        uint8_t* dr = A.take<uint8_t>((size_t)pr->width[l] * pr->height[l]);
        CU(cudaMemcpy2DAsync(dl, pl->width[l], pl->data[l], pl->stride[l], pl->width[l], pl->height[l], cudaMemcpyHostToDevice, s));
        CU(cudaMemcpy2DAsync(dr, pr->width[l], pr->data[l], pr->stride[l], pr->width[l], pr->height[l], cudaMemcpyHostToDevice, s));
        a.imgl[l] = dl; a.imgr[l] = dr; a.pitchl[l] = pl->width[l]; a.pitchr[l] = pr->width[l];
        a.widthr[l] = pr->width[l];
    }
    if (A.off > total) return fail(ORBFE_ERR_CAPACITY, "internal: matcher arena overflow");
    a.kl = d_kl; a.dl = d_dl; a.nl = nl; a.kr = d_kr; a.dr = d_dr; a.nr = nr;
    a.scale = d_scale; a.inv_scale = d_inv; a.nlevels = L; a.nrows = pl->height[0]; a.mbf = mbf; a.mb = mb;
    launch_stereo(a, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(uright, a.uright, sizeof(float) * nl, cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(depth, a.depth, sizeof(float) * nl, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return ORBFE_OK;
}

// Frame::ComputeStereoMatches for a batch of interleaved stereo pairs, device resident
extern "C" int orbfe_stereo_match_batch_device(orbfe_matcher* m, const orbfe_device_pyramid_view* pyr,
                                               const orbfe_keypoint* d_kps, const uint8_t* d_desc, const int32_t* d_counts,
                                               int cap, int npairs, const float* scale_factors,
                                               const float* inv_scale_factors, int nlevels, float mbf, float mb,
                                               float* uright, float* depth) {
    if (!m || !pyr || !scale_factors || !inv_scale_factors || !(mb > 0) || npairs < 0 || cap < 0)
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_stereo_match_batch_device");
    if (npairs == 0 || cap == 0) return ORBFE_OK;
    if (!d_kps || !d_desc || !d_counts || !uright || !depth || nlevels < 1 || nlevels > kMaxLevels || pyr->nlevels < nlevels)
        return fail(ORBFE_ERR_ARG, "bad device arrays / pyramid in orbfe_stereo_match_batch_device");
    if ((uintptr_t)d_desc & 15) return fail(ORBFE_ERR_ARG, "device descriptors must be 16-byte aligned");
    CU(cudaSetDevice(m->device));
    cudaStream_t s = m->stream;
    int rc;
    const size_t outN = (size_t)npairs * cap;
    const size_t bytes = 2 * padded((size_t)nlevels * 4) + padded(2 * outN * 4) + padded(outN * 4);
    if ((rc = m->scratch[0].ensure(bytes))) return rc;
    if ((rc = m->stage.ensure(2 * padded((size_t)nlevels * 4)))) return rc;
    if ((rc = m->result.ensure(2 * outN * 4))) return rc;
    Arena A; A.base = m->scratch[0].p;
    float* d_scale = A.take<float>(nlevels);
    float* d_inv = A.take<float>(nlevels);
    StereoArgs a{};
    float* ud = A.take<float>(2 * outN);           // uright | depth, fetched with one copy
    a.uright = ud; a.depth = ud + outN; a.sad = A.take<int>(outN);
    if (A.off > bytes) return fail(ORBFE_ERR_CAPACITY, "internal: matcher arena overflow");
    float* hs = reinterpret_cast<float*>(m->stage.p);
    std::memcpy(hs, scale_factors, sizeof(float) * nlevels);
    std::memcpy(hs + nlevels, inv_scale_factors, sizeof(float) * nlevels);
    CU(cudaMemcpyAsync(d_scale, hs, sizeof(float) * nlevels, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_inv, hs + nlevels, sizeof(float) * nlevels, cudaMemcpyHostToDevice, s));
    for (int l = 0; l < nlevels; l++) {
        a.imgl[l] = a.imgr[l] = pyr->d_data[l];
        a.pitchl[l] = a.pitchr[l] = (int)pyr->pitch[l];
        a.widthr[l] = pyr->width[l];
        a.frameStride[l] = (long long)pyr->frame_stride[l];
    }
    a.batched = 1; a.cap = cap; a.counts = d_counts; a.kl = d_kps; a.dl = d_desc;
    a.scale = d_scale; a.inv_scale = d_inv; a.nlevels = nlevels; a.nrows = pyr->height[0]; a.mbf = mbf; a.mb = mb;
    launch_stereo(a, s, npairs);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m->result.p, a.uright, 2 * outN * 4, cudaMemcpyDeviceToHost, s));   // uright and depth are adjacent
    CU(cudaStreamSynchronize(s));
    std::memcpy(uright, m->result.p, outN * 4);
    std::memcpy(depth, m->result.p + outN * 4, outN * 4);
    return ORBFE_OK;
}
