// ref_wrap.cpp — ORACLE support (test infrastructure only).
//
// C-ABI wrapper around the UNMODIFIED reference class ORB_SLAM2::ORBextractor
// (/root/reference/src/ORBextractor.cc compiled in place against oracle/cvshim; see
// oracle/Makefile).  Built into oracle/_ref/liborbref.so (git-ignored; travels to the GPU
// box as a prebuilt file).  With -DORBREF_BUMP the library also replaces operator new/delete
// by a per-extractor monotonic arena so that heap addresses grow in creation order: that pins
// the one allocator-dependent choice of the reference — the (count, pointer) sort in
// DistributeOctTree, ORBextractor.cc:728 — to "later-created node first among equal counts",
// which is the contract this repo implements (SURVEY.md §7 hard part 1).
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "ORBextractor.h"
#include "orb_oracle.h"

#ifdef ORBREF_BUMP
namespace {
struct Arena {
    // plain arrays only: the arena must never call operator new itself
    enum { kMaxChunks = 256 };
    char* chunks[kMaxChunks];
    int nchunks = 0;
    size_t chunk_size = 64u << 20, used = 0;
    int cur = -1;
    void reset() { cur = nchunks ? 0 : -1; used = 0; }
    void* alloc(size_t n) {
        n = (n + 15) & ~size_t(15);
        if (n > chunk_size) return nullptr;
        if (cur < 0 || used + n > chunk_size) {
            if (cur + 1 >= kMaxChunks) return nullptr;
            cur++;
            if (cur >= nchunks) chunks[nchunks++] = (char*)std::malloc(chunk_size);
            used = 0;
        }
        void* p = chunks[cur] + used; used += n; return p;
    }
    bool owns(const void* p) const {
        for (int i = 0; i < nchunks; i++)
            if ((const char*)p >= chunks[i] && (const char*)p < chunks[i] + chunk_size) return true;
        return false;
    }
};
thread_local Arena* g_active = nullptr;
std::vector<Arena*>& all_arenas() { static std::vector<Arena*>* v = new (std::malloc(sizeof(std::vector<Arena*>))) std::vector<Arena*>(); return *v; }
bool any_arena_owns(const void* p) {
    for (Arena* a : all_arenas()) if (a->owns(p)) return true;
    return false;
}
}  // namespace

void* operator new(size_t n) {
    if (g_active) { void* p = g_active->alloc(n); if (p) return p; }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !any_arena_owns(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }
#endif

struct RefHandle {
    ORB_SLAM2::ORBextractor* ext;
#ifdef ORBREF_BUMP
    Arena* arena;
#endif
};

extern "C" {

void* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
    RefHandle* h = (RefHandle*)std::malloc(sizeof(RefHandle));
    h->ext = new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
#ifdef ORBREF_BUMP
    h->arena = new (std::malloc(sizeof(Arena))) Arena();
    all_arenas().push_back(h->arena);   // creation is expected single-threaded (tests/bench)
#endif
    return h;
}

void ref_extractor_destroy(void* hv) {
    RefHandle* h = (RefHandle*)hv;
    delete h->ext;
    // arenas are kept registered (tiny leak) so late deletes stay recognisable
    std::free(h);
}

void ref_tables(void* hv, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2) {
    RefHandle* h = (RefHandle*)hv;
    std::vector<float> a = h->ext->GetScaleFactors(), b = h->ext->GetInverseScaleFactors(),
                       c = h->ext->GetScaleSigmaSquares(), d = h->ext->GetInverseScaleSigmaSquares();
    for (size_t i = 0; i < a.size(); i++) { scale[i] = a[i]; inv_scale[i] = b[i]; sigma2[i] = c[i]; inv_sigma2[i] = d[i]; }
}

// ORBextractor::operator() — returns keypoint count, -1 if cap is too small.
int ref_extract(void* hv, const uint8_t* img, int w, int hgt, ptrdiff_t stride, orc_keypoint* kps,
                uint8_t* desc, int cap) {
    RefHandle* h = (RefHandle*)hv;
    int n = 0;
#ifdef ORBREF_BUMP
    for (cv::Mat& m : h->ext->mvImagePyramid) m = cv::Mat();   // drop last call's arena buffers
    h->arena->reset();
    g_active = h->arena;
#endif
    {
        cv::Mat image(hgt, w, CV_8UC1, (void*)img, (size_t)stride);
        std::vector<cv::KeyPoint> keys;
        cv::Mat descriptors;
        (*h->ext)(image, cv::Mat(), keys, descriptors);
        n = (int)keys.size();
        if (n <= cap) {
            for (int i = 0; i < n; i++) {
                const cv::KeyPoint& k = keys[i];
                kps[i] = orc_keypoint{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
                std::memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32);
            }
        } else n = -1;
    }
#ifdef ORBREF_BUMP
    g_active = nullptr;
#endif
    return n;
}

int ref_level(void* hv, int level, const uint8_t** data, int* w, int* hgt, ptrdiff_t* stride) {
    RefHandle* h = (RefHandle*)hv;
    if (level < 0 || level >= (int)h->ext->mvImagePyramid.size()) return -1;
    const cv::Mat& m = h->ext->mvImagePyramid[level];
    *data = m.data; *w = m.cols; *hgt = m.rows; *stride = (ptrdiff_t)(size_t)m.step;
    return 0;
}

}  // extern "C"
