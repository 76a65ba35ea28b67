// api_comm.cu — BASELINE.json configs[4]: the query-sharded brute-force kNN-2 and its NCCL result gather
// behind the C ABI (include/orbfe.h: orbfe_comm_*, orbfe_knn2_sharded).
//
// The reference has no multi-device path; what is sharded here are the independent best / second-best loops
// of src/ORBmatcher.cc:233-269, one query per loop.  Every rank holds all query rows and the whole train set
// (32 MB for 1M rows: replicated, as the survey chose), searches the contiguous shard of queries it owns and
// contributes 8 bytes per query to one ncclAllGather.  NCCL is resolved with dlopen at the first use, so
// liborbfe.so itself carries no NCCL dependency: inside a torch process "libnccl.so.2" is the copy torch
// loaded, elsewhere the system library.
#include <dlfcn.h>

#include <cstring>
#include <mutex>
#include <new>

#include "api_common.cuh"
#include "match_internal.cuh"

using namespace orbfe;

namespace {

// the slice of nccl.h this file needs (NCCL 2.x ABI: 128-byte id passed by value, ncclUint8 == 1)
struct NcclId { char internal[ORBFE_COMM_ID_BYTES]; };
typedef void* NcclComm;
struct NcclApi {
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);   // already in the process (torch)?
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        api.GetUniqueId = (int (*)(NcclId*))dlsym(h, "ncclGetUniqueId");
        api.CommInitRank = (int (*)(NcclComm*, int, NcclId, int))dlsym(h, "ncclCommInitRank");
        api.CommDestroy = (int (*)(NcclComm))dlsym(h, "ncclCommDestroy");
        api.AllGather = (int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t))dlsym(h, "ncclAllGather");
        api.GetErrorString = (const char* (*)(int))dlsym(h, "ncclGetErrorString");
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather;
    });
    return api;
}

int nccl_fail(const char* what, int r) {
    static thread_local std::string msg;
    msg = what;
    if (nccl().GetErrorString) { msg += ": "; msg += nccl().GetErrorString(r); }
    return fail(ORBFE_ERR_CUDA, msg.c_str());
}

int need_nccl() {
    return nccl().ok ? 0 : fail(ORBFE_ERR_ARG, "NCCL not available: dlopen(libnccl.so.2) failed or lacks the entry points");
}

__global__ void knn2_pack_kernel(const int* __restrict__ idx, const int* __restrict__ d1, const int* __restrict__ d2, int n,
                                 orbfe_knn_result* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    orbfe_knn_result r;
    r.idx = idx[i]; r.d1 = (uint16_t)d1[i]; r.d2 = (uint16_t)d2[i];
    out[i] = r;
}

}  // namespace

namespace orbfe {
int launch_knn2_pack(const int* d_idx, const int* d_d1, const int* d_d2, int n, void* d_packed, cudaStream_t s) {
    if (n <= 0) return 0;
    knn2_pack_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_idx, d_d1, d_d2, n, (orbfe_knn_result*)d_packed);
    return 1;
}
}  // namespace orbfe

struct orbfe_comm {
    NcclComm comm = nullptr;
    int world = 1, rank = 0, device = 0;
    bool owned = false;
    DevArray<int> tmp;      // (idx, d1, d2) of the local shard before packing
};

extern "C" {

int orbfe_comm_unique_id(uint8_t id[ORBFE_COMM_ID_BYTES]) {
    if (!id) return fail(ORBFE_ERR_ARG, "id is NULL");
    int rc = need_nccl();
    if (rc) return rc;
    NcclId nid;
    const int r = nccl().GetUniqueId(&nid);
    if (r) return nccl_fail("ncclGetUniqueId", r);
    std::memcpy(id, nid.internal, ORBFE_COMM_ID_BYTES);
    return ORBFE_OK;
}

int orbfe_comm_create(const uint8_t id[ORBFE_COMM_ID_BYTES], int world, int rank, int device, orbfe_comm** out) {
    if (!out) return fail(ORBFE_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (!id || world < 1 || rank < 0 || rank >= world) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_comm_create");
    if (device < 0 || orbfe_device_count() <= device) return fail(ORBFE_ERR_NO_DEVICE, "no CUDA device: liborbfe has no CPU fallback");
    int rc = need_nccl();
    if (rc) return rc;
    CU(cudaSetDevice(device));
    orbfe_comm* c = new (std::nothrow) orbfe_comm();
    if (!c) return fail(ORBFE_ERR_ARG, "out of host memory");
    NcclId nid;
    std::memcpy(nid.internal, id, ORBFE_COMM_ID_BYTES);
    const int r = nccl().CommInitRank(&c->comm, world, nid, rank);
    if (r) { delete c; return nccl_fail("ncclCommInitRank", r); }
    c->world = world; c->rank = rank; c->device = device; c->owned = true;
    *out = c;
    return ORBFE_OK;
}

int orbfe_comm_wrap(void* nccl_comm, int world, int rank, int device, orbfe_comm** out) {
    if (!out) return fail(ORBFE_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (!nccl_comm || world < 1 || rank < 0 || rank >= world) return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_comm_wrap");
    int rc = need_nccl();
    if (rc) return rc;
    orbfe_comm* c = new (std::nothrow) orbfe_comm();
    if (!c) return fail(ORBFE_ERR_ARG, "out of host memory");
    c->comm = nccl_comm; c->world = world; c->rank = rank; c->device = device; c->owned = false;
    *out = c;
    return ORBFE_OK;
}

void orbfe_comm_destroy(orbfe_comm* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    c->tmp.release();
    if (c->owned && c->comm && nccl().ok) nccl().CommDestroy(c->comm);
    delete c;
}

int orbfe_comm_world(const orbfe_comm* c) { return c ? c->world : 0; }
int orbfe_comm_rank(const orbfe_comm* c) { return c ? c->rank : -1; }

int orbfe_knn2_shard_rows(int nq, int world) { return world > 0 && nq > 0 ? (nq + world - 1) / world : 0; }

int orbfe_knn2_sharded(orbfe_matcher* m, orbfe_comm* c, const uint8_t* d_query, int nq, const uint8_t* d_train,
                       int nt, int th, float ratio, orbfe_knn_result* d_out) {
    if (!m || !c || nq < 0 || nt < 0 || (nq && (!d_query || !d_out)) || (nt && !d_train))
        return fail(ORBFE_ERR_ARG, "bad arguments to orbfe_knn2_sharded");
    if (nq == 0) return ORBFE_OK;
    if ((uintptr_t)d_out & 7) return fail(ORBFE_ERR_ARG, "d_out must be 8-byte aligned");
    CU(cudaSetDevice(c->device));
    const int per = orbfe_knn2_shard_rows(nq, c->world);
    const int b = std::min(nq, c->rank * per), e = std::min(nq, b + per);
    const int n = e - b;
    int rc;
    if ((rc = c->tmp.ensure((size_t)3 * std::max(per, 1)))) return rc;
    int* t_idx = c->tmp.p;
    int* t_d1 = t_idx + per;
    int* t_d2 = t_d1 + per;
    cudaStream_t s = (cudaStream_t)orbfe_matcher_stream(m);
    orbfe_knn_result* mine = d_out + (size_t)c->rank * per;
    if (n > 0) {
        if ((rc = orbfe_knn2_device(m, d_query + (size_t)b * 32, n, d_train, nt, th, ratio, t_idx, t_d1, t_d2))) return rc;
        launch_knn2_pack(t_idx, t_d1, t_d2, n, mine, s);
    }
    if (n < per) CU(cudaMemsetAsync(mine + n, 0xFF, (size_t)(per - n) * sizeof(orbfe_knn_result), s));   // padding rows: idx -1
    CU(cudaGetLastError());
    if (c->world > 1) {
        // in place: this rank's block already sits at recvbuff + rank * count
        const int r = nccl().AllGather(mine, d_out, (size_t)per * sizeof(orbfe_knn_result), /*ncclUint8*/ 1, c->comm, s);
        if (r) return nccl_fail("ncclAllGather", r);
    }
    return ORBFE_OK;
}

}  // extern "C"
