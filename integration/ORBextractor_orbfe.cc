// ORBextractor_orbfe.cc — link-level drop-in for the reference's src/ORBextractor.cc.
//
// Defines the members of the reference's OWN class ORB_SLAM2::ORBextractor (include/ORBextractor.h:46-114, header
// unchanged) on top of liborbfe's C ABI: take src/ORBextractor.cc out of the library's source list
// (CMakeLists.txt:51-69), add this file and link liborbfe.  Frame::ExtractORB (src/Frame.cc:272-278), the
// `new ORBextractor(...)` calls in Tracking (src/Tracking.cc:122-128) and the scale-table getters compile and behave
// unchanged.  The class has no member that could hold the device handle, so handles live in a side table keyed by the
// object's address; the header's inline destructor cannot release them, they are reclaimed when an extractor is
// constructed at the same address or by orbfe_dropin_release(); at process exit the driver reclaims them (three extractors per process in the reference).
// There is no CPU fallback: a failed call throws std::runtime_error.
#include <cstring>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>

#include "ORBextractor.h"
#include "orbfe.h"

extern "C" void orbfe_dropin_release(void);

namespace {
struct Slot { orbfe_extractor* h = nullptr; bool pyramid = true; std::vector<uint8_t> desc; };
std::mutex g_mutex;
std::map<const void*, Slot>& slots() { static std::map<const void*, Slot> s; return s; }
void check(int rc, const char* what) {
    if (rc != ORBFE_OK) throw std::runtime_error(std::string(what) + ": " + orbfe_last_error());
}
}  // namespace

namespace ORB_SLAM2 {

// src/ORBextractor.cc:415-486
ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST) {
    orbfe_extractor* h = nullptr;
    check(orbfe_extractor_create(_nfeatures, _scaleFactor, _nlevels, _iniThFAST, _minThFAST, 0, 0, 0, 0, &h), "orbfe_extractor_create");
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
    mnFeaturesPerLevel.resize(nlevels); umax.resize(16);
    orbfe_get_scale_tables(h, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                           mnFeaturesPerLevel.data(), umax.data());
    mvImagePyramid.resize(nlevels);
    std::lock_guard<std::mutex> lock(g_mutex);
    Slot& s = slots()[this];
    if (s.h) orbfe_extractor_destroy(s.h);          // an earlier extractor lived at this address
    s.h = h;
}

// src/ORBextractor.cc:1119-1197.  The mask is ignored, as in the reference.
void ORBextractor::operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors) {
    (void)_mask;
    if (_image.empty()) return;                                        // :1122-1123
    Slot* s;
    {
        std::lock_guard<std::mutex> lock(g_mutex);
        s = &slots()[this];                                            // std::map nodes do not move
    }
    cv::Mat image = _image.getMat();
    const int cap = orbfe_extractor_max_keypoints(s->h) + 4096;
    // cv::KeyPoint and orbfe_keypoint are the same 28 bytes (pt.x, pt.y, size, angle, response, octave, class_id):
    // the library writes straight into the caller's vector
    static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
    _keypoints.resize(cap);
    s->desc.resize((size_t)cap * 32);
    int n = 0;
    check(orbfe_extract(s->h, image.data, image.cols, image.rows, (ptrdiff_t)(size_t)image.step,
                        reinterpret_cast<orbfe_keypoint*>(_keypoints.data()), s->desc.data(), cap, &n), "orbfe_extract");
    _keypoints.resize(n);
    if (n == 0) {
        _descriptors.release();                                        // :1145-1146
    } else {
        _descriptors.create(n, 32, CV_8U);                             // :1149
        cv::Mat descriptors = _descriptors.getMat();
        if ((size_t)descriptors.step == 32) std::memcpy(descriptors.data, s->desc.data(), (size_t)n * 32);
        else for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), &s->desc[(size_t)i * 32], 32);
    }
    // Frame::ComputeStereoMatches reads mvImagePyramid (src/Frame.cc:519, :609-626); nothing else does.  One
    // transfer group + one synchronisation into the handle's page-locked buffer; the cv::Mat headers alias it (valid
    // until the next call, like the reference's pyramid, which the next operator() overwrites)
    if (s->pyramid) {
        const uint8_t* lv[16]; int32_t w[16], h[16]; ptrdiff_t st[16];
        check(orbfe_pyramid_host(s->h, 0, lv, w, h, st), "orbfe_pyramid_host");
        for (int l = 0; l < nlevels; l++)
            mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8UC1, const_cast<uint8_t*>(lv[l]), (size_t)st[l]);
    }
}

}  // namespace ORB_SLAM2

// Monocular and RGB-D systems never read mvImagePyramid: Tracking may switch the device-to-host copy of the eight levels
// off for its extractors (one call after each `new ORBextractor(...)`, src/Tracking.cc:122-128).  Declare it where used:
//   extern "C" void orbfe_dropin_set_pyramid_download(const void* extractor, int on);
extern "C" void orbfe_dropin_set_pyramid_download(const void* extractor, int on) {
    std::lock_guard<std::mutex> lock(g_mutex);
    std::map<const void*, Slot>::iterator it = slots().find(extractor);
    if (it != slots().end()) it->second.pyramid = on != 0;
}

// Releases every device handle of the side table (streams, device and page-locked memory): System::Shutdown or the
// host's own teardown calls it while the CUDA context is still alive.  It is deliberately NOT registered with atexit:
// CUDA calls from an exit handler run after the runtime has begun tearing down and can block forever (seen on the
// B200 box).  Extractors must not be used afterwards.
extern "C" void orbfe_dropin_release(void) {
    std::lock_guard<std::mutex> lock(g_mutex);
    for (std::map<const void*, Slot>::iterator it = slots().begin(); it != slots().end(); ++it)
        if (it->second.h) { orbfe_extractor_destroy(it->second.h); it->second.h = nullptr; }
    slots().clear();
}
