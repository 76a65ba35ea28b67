// orbfe_adapter.hpp — header-only C++ host side of liborbfe: drop-in classes with the
// reference's names and signatures, so that Frame.cc / Tracking.cc compile unchanged.
//
//   ORB_SLAM2::ORBextractor   replaces include/ORBextractor.h:46-114 + src/ORBextractor.cc
//   ORB_SLAM2::ORBmatcher     replaces the hot-path members of include/ORBmatcher.h:51-117
//                             (DescriptorDistance, SearchForInitialization,
//                             SearchByProjection(Frame&, vector<MapPoint*>&, th),
//                             SearchByProjection(CurrentFrame, LastFrame, th, bMono)) and offers
//                             both SearchByBoW overloads (:185-325, :610-743), the key-frame guided searches
//                             (SearchByProjection(F, KF) :1641-1770, SearchByProjection(KF, Scw) :335-451, Fuse x2
//                             :940-1237, SearchBySim3 :1251-1477, SearchForTriangulation :754-928),
//                             ComputeStereoMatches(Frame&) for src/Frame.cc:512-686,
//                             ComputeImageBounds(Frame&, im) for :481-512 and
//                             UndistortAndAssign(Frame&) for :443-475 + :246-267
//
// Include AFTER the OpenCV core headers (cv::Mat, cv::KeyPoint, cv::InputArray, cv::OutputArray,
// cv::Point2f must be visible).  The Frame / MapPoint dependent members are templates over the
// reference's own types, so this header needs neither Frame.h nor MapPoint.h; it reads exactly
// the fields the reference routines read.  All arithmetic happens in liborbfe's CUDA kernels;
// this file only flattens objects into the plain views of include/orbfe.h and writes results
// back.  There is no CPU fallback: a failed call throws std::runtime_error.
#ifndef ORBFE_ADAPTER_HPP
#define ORBFE_ADAPTER_HPP

#include <cstring>
#include <cmath>
#include <map>
#include <set>
#include <utility>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "orbfe.h"

// The two classes carry the reference's names by default.  Inside a translation unit that also sees the reference's
// own class of that name (its ORBextractor.h / ORBmatcher.h), give the adapter another name first:
//   #define ORBFE_MATCHER_CLASS ORBmatcherFE      (integration/ORBmatcher_orbfe.cc does this)
#ifndef ORBFE_EXTRACTOR_CLASS
#define ORBFE_EXTRACTOR_CLASS ORBextractor
#endif
#ifndef ORBFE_MATCHER_CLASS
#define ORBFE_MATCHER_CLASS ORBmatcher
#endif

namespace ORB_SLAM2 {

namespace orbfe_detail {
inline void check(int rc, const char* what) {
    if (rc != ORBFE_OK) throw std::runtime_error(std::string(what) + ": " + orbfe_last_error());
}
}  // namespace orbfe_detail

class ORBFE_EXTRACTOR_CLASS {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // include/ORBextractor.h:53-54.  `device` and `downloadPyramid` are additions with defaults:
    // stereo needs mvImagePyramid on the host after every call (src/Frame.cc:519,609-626),
    // monocular / RGB-D callers can switch the copy off.
    ORBFE_EXTRACTOR_CLASS(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST,
                 int device = 0, bool downloadPyramid = true)
        : mpHandle(nullptr), nlevels(nlevels), mbDownloadPyramid(downloadPyramid) {
        orbfe_detail::check(orbfe_extractor_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, 0, 0, 0,
                                                   device, &mpHandle), "orbfe_extractor_create");
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels); mvInvLevelSigma2.resize(nlevels);
        orbfe_get_scale_tables(mpHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data(), nullptr, nullptr);
        this->scaleFactor = scaleFactor;                                   // double member set from a float (:417)
        mvImagePyramid.resize(nlevels);
    }
    ~ORBFE_EXTRACTOR_CLASS() { orbfe_extractor_destroy(mpHandle); }
    ORBFE_EXTRACTOR_CLASS(const ORBFE_EXTRACTOR_CLASS&) = delete;
    ORBFE_EXTRACTOR_CLASS& operator=(const ORBFE_EXTRACTOR_CLASS&) = delete;

    // src/ORBextractor.cc:1119-1197.  Mask is ignored, as in the reference.
    void operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                    cv::OutputArray _descriptors) {
        if (_image.empty()) return;                                        // :1122-1123
        cv::Mat image = _image.getMat();
        const int cap = orbfe_extractor_max_keypoints(mpHandle) + 4096;
        // cv::KeyPoint and orbfe_keypoint are the same 28 bytes: the library writes straight into the caller's vector
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
        _keypoints.resize(cap);
        mDescBuf.resize((size_t)cap * 32);
        int n = 0;
        orbfe_detail::check(orbfe_extract(mpHandle, image.data, image.cols, image.rows, (ptrdiff_t)(size_t)image.step,
                                          reinterpret_cast<orbfe_keypoint*>(_keypoints.data()), mDescBuf.data(), cap, &n),
                            "orbfe_extract");
        _keypoints.resize(n);
        if (n == 0) {
            _descriptors.release();                                        // :1145-1146
        } else {
            _descriptors.create(n, 32, CV_8U);                             // :1149
            cv::Mat descriptors = _descriptors.getMat();
            if ((size_t)descriptors.step == 32) std::memcpy(descriptors.data, mDescBuf.data(), (size_t)n * 32);
            else for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), &mDescBuf[(size_t)i * 32], 32);
        }
        if (mbDownloadPyramid) {
            // one transfer group + one synchronisation into the handle's page-locked buffer; the headers alias it
            // (valid until the next call, like the reference's pyramid, which the next operator() overwrites)
            const uint8_t* lv[16]; int32_t w[16], h[16]; ptrdiff_t st[16];
            orbfe_detail::check(orbfe_pyramid_host(mpHandle, 0, lv, w, h, st), "orbfe_pyramid_host");
            for (int l = 0; l < nlevels; l++)
                mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8UC1, const_cast<uint8_t*>(lv[l]), (size_t)st[l]);
        }
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;   // include/ORBextractor.h:86 (interior pixels of each level)

    orbfe_extractor* Handle() { return mpHandle; }

protected:
    orbfe_extractor* mpHandle;
    double scaleFactor;
    int nlevels;
    bool mbDownloadPyramid;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<uint8_t> mDescBuf;
};

class ORBFE_MATCHER_CLASS {
public:
    static const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:37-39

    ORBFE_MATCHER_CLASS(float nnratio = 0.6f, bool checkOri = true, int device = 0)
        : mfNNratio(nnratio), mbCheckOrientation(checkOri), mpHandle(nullptr) {
        orbfe_detail::check(orbfe_matcher_create(device, &mpHandle), "orbfe_matcher_create");
    }
    ~ORBFE_MATCHER_CLASS() { orbfe_matcher_destroy(mpHandle); }
    ORBFE_MATCHER_CLASS(const ORBFE_MATCHER_CLASS&) = delete;
    ORBFE_MATCHER_CLASS& operator=(const ORBFE_MATCHER_CLASS&) = delete;

    // src/ORBmatcher.cc:1828-1844
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbfe_hamming256(a.data, b.data); }

    // src/ORBmatcher.cc:469-600
    template <class FrameT>
    int SearchForInitialization(FrameT& F1, FrameT& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10) {
        std::vector<orbfe_keypoint> k1, k2;
        std::vector<uint8_t> d1, d2;
        orbfe_frame_view v1 = View(F1, k1, d1), v2 = View(F2, k2, d2);
        vnMatches12.assign(F1.mvKeysUn.size(), -1);
        std::vector<float> prev(2 * vbPrevMatched.size());
        for (size_t i = 0; i < vbPrevMatched.size(); i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_init(mpHandle, &v1, &v2, prev.data(), vnMatches12.data(), windowSize,
                                              mfNNratio, mbCheckOrientation ? 1 : 0, &nmatches), "orbfe_search_init");
        for (size_t i = 0; i < vbPrevMatched.size(); i++) { vbPrevMatched[i].x = prev[2 * i]; vbPrevMatched[i].y = prev[2 * i + 1]; }
        return nmatches;
    }

    // SearchByProjection(Frame&, const vector<MapPoint*>&, th) — src/ORBmatcher.cc:51-138
    template <class FrameT, class MapPointT>
    int SearchByProjection(FrameT& F, const std::vector<MapPointT*>& vpMapPoints, const float th = 3) {
        std::vector<orbfe_keypoint> k;
        std::vector<uint8_t> d;
        orbfe_frame_view fv = View(F, k, d);
        if (!F.mvuRight.empty()) fv.uright = F.mvuRight.data();
        const int n = (int)vpMapPoints.size();
        std::vector<uint8_t> inView(n), desc((size_t)n * 32);
        std::vector<float> px(n), py(n), pxr(n), vcos(n);
        std::vector<int> level(n), obs(n);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpMapPoints[i];
            inView[i] = pMP && pMP->mbTrackInView && !pMP->isBad();        // :60-64
            if (!inView[i]) continue;
            px[i] = pMP->mTrackProjX; py[i] = pMP->mTrackProjY; pxr[i] = pMP->mTrackProjXR;
            level[i] = pMP->mnTrackScaleLevel; vcos[i] = pMP->mTrackViewCos;
            const cv::Mat dmp = pMP->GetDescriptor();
            std::memcpy(&desc[(size_t)i * 32], dmp.data, 32);
            obs[i] = pMP->Observations();
        }
        orbfe_mappoints_view mv = {n, inView.data(), px.data(), py.data(), pxr.data(), level.data(), vcos.data(),
                                   desc.data(), obs.data()};
        std::vector<uint8_t> occupied(F.mvpMapPoints.size(), 0);
        for (size_t i = 0; i < F.mvpMapPoints.size(); i++)                 // :95-98
            occupied[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
        std::vector<int> mpForKp(F.mvpMapPoints.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_local_points(mpHandle, &fv, &mv, th, mfNNratio, occupied.data(),
                                                      mpForKp.data(), &nmatches), "orbfe_search_local_points");
        for (size_t i = 0; i < mpForKp.size(); i++)
            if (mpForKp[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[mpForKp[i]];   // :132
        return nmatches;
    }

    // SearchByProjection(CurrentFrame, LastFrame, th, bMono) — src/ORBmatcher.cc:1484-1630.  The pose
    // arithmetic (:1494-1553) is kept verbatim on the host with the reference's own cv::Mat
    // expressions; candidates, distances, the greedy bookkeeping and the rotation check run on the GPU.
    template <class FrameT>
    int SearchByProjection(FrameT& CurrentFrame, const FrameT& LastFrame, const float th, const bool bMono) {
        // MatT is cv::Mat; spelling it through the frame type keeps these lines dependent, so the
        // header still parses where only a partial cv:: shim is available (tests/cpp)
        typedef typename std::decay<decltype(CurrentFrame.mTcw)>::type MatT;
        const MatT Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
        const MatT tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const MatT twc = -Rcw.t() * tcw;
        const MatT Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
        const MatT tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
        const MatT tlc = Rlw * twc + tlw;
        const bool bForward = tlc.template at<float>(2) > CurrentFrame.mb && !bMono;
        const bool bBackward = -tlc.template at<float>(2) > CurrentFrame.mb && !bMono;

        const int n = LastFrame.N;
        std::vector<uint8_t> valid(n, 0), desc((size_t)n * 32);
        std::vector<float> u(n), v(n), ur(n), radius(n), angle(n);
        std::vector<int> minL(n), maxL(n), obs(n);
        for (int i = 0; i < n; i++) {
            auto* pMP = LastFrame.mvpMapPoints[i];
            if (!pMP || LastFrame.mvbOutlier[i]) continue;                 // :1510-1514
            const MatT x3Dw = pMP->GetWorldPos();
            const MatT x3Dc = Rcw * x3Dw + tcw;
            const float xc = x3Dc.template at<float>(0);
            const float yc = x3Dc.template at<float>(1);
            const float invzc = 1.0 / x3Dc.template at<float>(2);
            if (invzc < 0) continue;                                       // :1522-1524
            u[i] = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;         // :1528-1529
            v[i] = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
            if (u[i] < CurrentFrame.mnMinX || u[i] > CurrentFrame.mnMaxX) continue;
            if (v[i] < CurrentFrame.mnMinY || v[i] > CurrentFrame.mnMaxY) continue;
            const int nLastOctave = LastFrame.mvKeys[i].octave;
            radius[i] = th * CurrentFrame.mvScaleFactors[nLastOctave];     // :1543
            if (bForward) { minL[i] = nLastOctave; maxL[i] = -1; }         // :1547-1552
            else if (bBackward) { minL[i] = 0; maxL[i] = nLastOctave; }
            else { minL[i] = nLastOctave - 1; maxL[i] = nLastOctave + 1; }
            ur[i] = u[i] - CurrentFrame.mbf * invzc;                       // :1575
            angle[i] = LastFrame.mvKeysUn[i].angle;                        // :1594
            const cv::Mat dMP = pMP->GetDescriptor();
            std::memcpy(&desc[(size_t)i * 32], dMP.data, 32);
            obs[i] = pMP->Observations();
            valid[i] = 1;
        }
        std::vector<orbfe_keypoint> k;
        std::vector<uint8_t> d;
        orbfe_frame_view fv = View(CurrentFrame, k, d);
        if (!CurrentFrame.mvuRight.empty()) fv.uright = CurrentFrame.mvuRight.data();
        orbfe_framepoints_view pv = {n, valid.data(), u.data(), v.data(), ur.data(), radius.data(), minL.data(),
                                     maxL.data(), angle.data(), desc.data(), obs.data()};
        std::vector<uint8_t> occupied(CurrentFrame.mvpMapPoints.size(), 0);
        for (size_t i = 0; i < occupied.size(); i++)                       // :1567-1569
            occupied[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
        std::vector<int> matchForKp(occupied.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_frame_points(mpHandle, &fv, &pv, mbCheckOrientation ? 1 : 0, occupied.data(),
                                                      matchForKp.data(), &nmatches), "orbfe_search_frame_points");
        for (size_t i = 0; i < matchForKp.size(); i++) {
            if (matchForKp[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[matchForKp[i]];   // :1587
            else if (matchForKp[i] == -2) CurrentFrame.mvpMapPoints[i] = nullptr;                           // :1621
        }
        return nmatches;
    }

    // SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) — src/ORBmatcher.cc:185-325.  Reads
    // pKF->GetMapPointMatches(), mDescriptors, mvKeysUn, mFeatVec and F.N, mDescriptors, mvKeys, mFeatVec.
    template <class KeyFrameT, class FrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF, FrameT& F, std::vector<MapPointT*>& vpMapPointMatches) {
        const std::vector<MapPointT*> vpMapPointsKF = pKF->GetMapPointMatches();
        vpMapPointMatches = std::vector<MapPointT*>(F.N, static_cast<MapPointT*>(nullptr));
        BowSide s1(pKF->mDescriptors, pKF->mvKeysUn, pKF->mFeatVec, &vpMapPointsKF);
        BowSide s2(F.mDescriptors, F.mvKeys, F.mFeatVec, static_cast<const std::vector<MapPointT*>*>(nullptr));
        std::vector<int32_t> m21(s2.angle.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_bow(mpHandle, &s1.v, &s2.v, TH_LOW, mfNNratio, mbCheckOrientation ? 1 : 0, nullptr,
                                             m21.data(), &nmatches), "orbfe_search_bow");
        for (size_t i = 0; i < m21.size() && i < vpMapPointMatches.size(); i++)
            if (m21[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[m21[i]];                        // :267
        return nmatches;
    }

    // SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) — src/ORBmatcher.cc:610-743
    template <class KeyFrameT, class MapPointT>
    int SearchByBoW(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12) {
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches();
        const std::vector<MapPointT*> vpMapPoints2 = pKF2->GetMapPointMatches();
        vpMatches12 = std::vector<MapPointT*>(vpMapPoints1.size(), static_cast<MapPointT*>(nullptr));
        BowSide s1(pKF1->mDescriptors, pKF1->mvKeysUn, pKF1->mFeatVec, &vpMapPoints1);
        BowSide s2(pKF2->mDescriptors, pKF2->mvKeysUn, pKF2->mFeatVec, &vpMapPoints2);
        std::vector<int32_t> m12(s1.angle.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_bow(mpHandle, &s1.v, &s2.v, TH_LOW - 1, mfNNratio, mbCheckOrientation ? 1 : 0,
                                             m12.data(), nullptr, &nmatches), "orbfe_search_bow");
        for (size_t i = 0; i < m12.size() && i < vpMatches12.size(); i++)
            if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];                               // :690
        return nmatches;
    }

    // ------------------------------------------------------------------------------------------
    // Key-frame guided searches.  As above, the pose arithmetic in front of
    // each candidate loop is the reference's own cv::Mat code, kept on the host; candidates, Hamming
    // distances, the first-owner bookkeeping and the rotation check run on the GPU; map mutation
    // (Replace / AddObservation / AddMapPoint) is replayed on the host in point order.

    // SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>&, th, ORBdist) — src/ORBmatcher.cc:1641-1770
    template <class FrameT, class KeyFrameT, class SetT>
    int SearchByProjection(FrameT& CurrentFrame, KeyFrameT* pKF, const SetT& sAlreadyFound, const float th, const int ORBdist) {
        typedef typename std::decay<decltype(CurrentFrame.mTcw)>::type MatT;
        const MatT Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
        const MatT tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
        const MatT Ow = -Rcw.t() * tcw;
        const auto vpMPs = pKF->GetMapPointMatches();
        const int n = (int)vpMPs.size();
        Points P(n);
        std::vector<float> angle(n, 0.0f);
        for (int i = 0; i < n; i++) {
            auto* pMP = vpMPs[i];
            if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;                    // :1663-1667
            const MatT x3Dw = pMP->GetWorldPos();
            const MatT x3Dc = Rcw * x3Dw + tcw;
            const float xc = x3Dc.template at<float>(0), yc = x3Dc.template at<float>(1);
            const float invzc = 1.0 / x3Dc.template at<float>(2);
            const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
            const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
            if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
            if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
            const MatT PO = x3Dw - Ow;
            const float dist3D = norm(PO);   // cv::norm, found by argument-dependent lookup at instantiation
            if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
            const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
            P.set(i, u, v, 0.0f, th * CurrentFrame.mvScaleFactors[nPredictedLevel], nPredictedLevel, pMP->GetDescriptor());
            angle[i] = pKF->mvKeysUn[i].angle;
        }
        std::vector<int> minL(n), maxL(n), obs(n, 1);
        for (int i = 0; i < n; i++) { minL[i] = P.level[i] - 1; maxL[i] = P.level[i] + 1; }      // :1705
        std::vector<orbfe_keypoint> k;
        std::vector<uint8_t> d;
        orbfe_frame_view fv = View(CurrentFrame, k, d);
        orbfe_framepoints_view pv = {n, P.valid.data(), P.u.data(), P.v.data(), P.ur.data(), P.radius.data(), minL.data(),
                                     maxL.data(), angle.data(), P.desc.data(), obs.data()};
        std::vector<uint8_t> occupied(CurrentFrame.mvpMapPoints.size(), 0);
        for (size_t i = 0; i < occupied.size(); i++) occupied[i] = CurrentFrame.mvpMapPoints[i] != nullptr;   // :1720
        std::vector<int> matchForKp(occupied.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_projected_points(mpHandle, &fv, &pv, ORBdist, mbCheckOrientation ? 1 : 0, occupied.data(),
                                                          matchForKp.data(), &nmatches), "orbfe_search_projected_points");
        for (size_t i = 0; i < matchForKp.size(); i++) {
            if (matchForKp[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[matchForKp[i]];       // :1736
            else if (matchForKp[i] == -2) CurrentFrame.mvpMapPoints[i] = nullptr;              // :1766
        }
        return nmatches;
    }

    // SearchByProjection(KeyFrame*, cv::Mat Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:335-451
    template <class KeyFrameT, class MatT, class MapPointT>
    int SearchByProjection(KeyFrameT* pKF, MatT Scw, const std::vector<MapPointT*>& vpPoints, std::vector<MapPointT*>& vpMatched, int th) {
        Sim3Pose<MatT> S(Scw);
        std::set<MapPointT*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
        spAlreadyFound.erase(static_cast<MapPointT*>(nullptr));
        const int n = (int)vpPoints.size();
        Points P(n);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            ProjectIntoKeyFrame(pKF, pMP, S.Rcw, S.tcw, S.Ow, false, true, (float)th, 0.0f, i, P);   // :361-407
        }
        std::vector<int> minL(n), maxL(n), obs(n, 1);
        std::vector<float> angle(n, 0.0f);
        for (int i = 0; i < n; i++) { minL[i] = P.level[i] - 1; maxL[i] = P.level[i]; }          // :426
        std::vector<orbfe_keypoint> k;
        orbfe_frame_view fv = KeyFrameView(pKF, k);
        orbfe_framepoints_view pv = {n, P.valid.data(), P.u.data(), P.v.data(), P.ur.data(), P.radius.data(), minL.data(),
                                     maxL.data(), angle.data(), P.desc.data(), obs.data()};
        std::vector<uint8_t> occupied(vpMatched.size(), 0);
        for (size_t i = 0; i < occupied.size(); i++) occupied[i] = vpMatched[i] != nullptr;      // :421
        std::vector<int> matchForKp(occupied.size(), -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_projected_points(mpHandle, &fv, &pv, TH_LOW, 0, occupied.data(), matchForKp.data(),
                                                          &nmatches), "orbfe_search_projected_points");
        for (size_t i = 0; i < matchForKp.size(); i++)
            if (matchForKp[i] >= 0) vpMatched[i] = vpPoints[matchForKp[i]];                      // :442
        return nmatches;
    }

    // Fuse(KeyFrame*, const vector<MapPoint*>&, th) — src/ORBmatcher.cc:940-1090.  The search of every
    // point runs first (it reads only keypoints, descriptors and the point's own state on entry), the
    // reference's bookkeeping follows in point order.  A point that an earlier Replace of this very call
    // turned bad is skipped at replay time, as the sequential loop would (:952).
    template <class KeyFrameT, class MapPointT>
    int Fuse(KeyFrameT* pKF, const std::vector<MapPointT*>& vpMapPoints, const float th = 3.0) {
        typedef typename std::decay<decltype(pKF->GetRotation())>::type MatT;
        const MatT Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation(), Ow = pKF->GetCameraCenter();
        const int n = (int)vpMapPoints.size();
        Points P(n);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpMapPoints[i];
            if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;                         // :949-953
            ProjectIntoKeyFrame(pKF, pMP, Rcw, tcw, Ow, false, true, th, pKF->mbf, i, P);          // :955-996
        }
        std::vector<int32_t> best(n, -1);
        std::vector<orbfe_keypoint> k;
        orbfe_frame_view fv = KeyFrameView(pKF, k);
        if (!pKF->mvuRight.empty()) fv.uright = pKF->mvuRight.data();
        orbfe_windowpoints_view wv = {n, P.valid.data(), P.u.data(), P.v.data(), P.ur.data(), P.radius.data(), P.level.data(), P.desc.data()};
        orbfe_detail::check(orbfe_search_window_best(mpHandle, &fv, &wv, pKF->mvInvLevelSigma2.data(), TH_LOW, best.data(), nullptr),
                            "orbfe_search_window_best");
        int nFused = 0;
        for (int i = 0; i < n; i++) {
            if (best[i] < 0) continue;
            MapPointT* pMP = vpMapPoints[i];
            if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
            MapPointT* pMPinKF = pKF->GetMapPoint(best[i]);                                       // :1057-1073
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // Fuse(KeyFrame*, cv::Mat Scw, vpPoints, th, vpReplacePoint) — src/ORBmatcher.cc:1100-1237
    template <class KeyFrameT, class MatT, class MapPointT>
    int Fuse(KeyFrameT* pKF, MatT Scw, const std::vector<MapPointT*>& vpPoints, float th, std::vector<MapPointT*>& vpReplacePoint) {
        Sim3Pose<MatT> S(Scw);
        const std::set<MapPointT*> spAlreadyFound = pKF->GetMapPoints();
        const int n = (int)vpPoints.size();
        Points P(n);
        for (int i = 0; i < n; i++) {
            MapPointT* pMP = vpPoints[i];
            if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
            ProjectIntoKeyFrame(pKF, pMP, S.Rcw, S.tcw, S.Ow, true, true, th, 0.0f, i, P);         // :1127-1168
        }
        std::vector<int32_t> best(n, -1);
        std::vector<orbfe_keypoint> k;
        orbfe_frame_view fv = KeyFrameView(pKF, k);
        orbfe_windowpoints_view wv = {n, P.valid.data(), P.u.data(), P.v.data(), P.ur.data(), P.radius.data(), P.level.data(), P.desc.data()};
        orbfe_detail::check(orbfe_search_window_best(mpHandle, &fv, &wv, nullptr, TH_LOW, best.data(), nullptr), "orbfe_search_window_best");
        int nFused = 0;
        for (int i = 0; i < n; i++) {
            if (best[i] < 0) continue;
            MapPointT* pMP = vpPoints[i];
            MapPointT* pMPinKF = pKF->GetMapPoint(best[i]);                                       // :1208-1222
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, best[i]);
                pKF->AddMapPoint(pMP, best[i]);
            }
            nFused++;
        }
        return nFused;
    }

    // SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th) — src/ORBmatcher.cc:1251-1477
    template <class KeyFrameT, class MatT, class MapPointT>
    int SearchBySim3(KeyFrameT* pKF1, KeyFrameT* pKF2, std::vector<MapPointT*>& vpMatches12, const float& s12, const MatT& R12,
                     const MatT& t12, const float th) {
        const MatT R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
        const MatT sR12 = s12 * R12, sR21 = (1.0 / s12) * R12.t(), t21 = -sR21 * t12;            // :1265-1268
        const std::vector<MapPointT*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
        const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
        std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
        for (int i = 0; i < N1; i++) {                                                           // :1279-1289
            MapPointT* pMP = vpMatches12[i];
            if (!pMP) continue;
            vbAlreadyMatched1[i] = true;
            const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
        Points P1(N1), P2(N2);
        for (int i1 = 0; i1 < N1; i1++) {                                                        // :1294-1331
            MapPointT* pMP = vpMapPoints1[i1];
            if (!pMP || vbAlreadyMatched1[i1] || pMP->isBad()) continue;
            const MatT p3Dc1 = R1w * pMP->GetWorldPos() + t1w;
            ProjectCameraPoint(pKF2, pMP, MatT(sR21 * p3Dc1 + t21), th, i1, P1);
        }
        for (int i2 = 0; i2 < N2; i2++) {                                                        // :1368-1409
            MapPointT* pMP = vpMapPoints2[i2];
            if (!pMP || vbAlreadyMatched2[i2] || pMP->isBad()) continue;
            const MatT p3Dc2 = R2w * pMP->GetWorldPos() + t2w;
            ProjectCameraPoint(pKF1, pMP, MatT(sR12 * p3Dc2 + t12), th, i2, P2);
        }
        std::vector<orbfe_keypoint> k1, k2;
        orbfe_frame_view f1 = KeyFrameView(pKF1, k1), f2 = KeyFrameView(pKF2, k2);
        orbfe_windowpoints_view w1 = {N1, P1.valid.data(), P1.u.data(), P1.v.data(), P1.ur.data(), P1.radius.data(), P1.level.data(), P1.desc.data()};
        orbfe_windowpoints_view w2 = {N2, P2.valid.data(), P2.u.data(), P2.v.data(), P2.ur.data(), P2.radius.data(), P2.level.data(), P2.desc.data()};
        std::vector<int32_t> m12(N1, -1);
        int nFound = 0;
        orbfe_detail::check(orbfe_search_by_sim3(mpHandle, &f1, &f2, &w1, &w2, TH_HIGH, m12.data(), &nFound), "orbfe_search_by_sim3");
        for (int i1 = 0; i1 < N1; i1++)
            if (m12[i1] >= 0) vpMatches12[i1] = vpMapPoints2[m12[i1]];                           // :1455
        return nFound;
    }

    // SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) — src/ORBmatcher.cc:754-928
    template <class KeyFrameT, class MatT>
    int SearchForTriangulation(KeyFrameT* pKF1, KeyFrameT* pKF2, MatT F12, std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                               const bool bOnlyStereo) {
        const MatT Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
        const MatT C2 = R2w * Cw + t2w;                                                          // :762-768
        const float invz = 1.0f / C2.template at<float>(2);
        const float ex = pKF2->fx * C2.template at<float>(0) * invz + pKF2->cx;
        const float ey = pKF2->fy * C2.template at<float>(1) * invz + pKF2->cy;
        TriSide s1(pKF1), s2(pKF2);
        float F[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F[3 * r + c] = F12.template at<float>(r, c);
        std::vector<int32_t> m12(s1.v.n, -1);
        int nmatches = 0;
        orbfe_detail::check(orbfe_search_triangulation(mpHandle, &s1.v, &s2.v, F, ex, ey, pKF2->mvScaleFactors.data(),
                                                       pKF2->mvLevelSigma2.data(), (int)pKF2->mvScaleFactors.size(), bOnlyStereo ? 1 : 0,
                                                       mbCheckOrientation ? 1 : 0, m12.data(), &nmatches), "orbfe_search_triangulation");
        vMatchedPairs.clear();
        vMatchedPairs.reserve(nmatches);
        for (size_t i = 0; i < m12.size(); i++)
            if (m12[i] >= 0) vMatchedPairs.push_back(std::make_pair(i, (size_t)m12[i]));           // :918-925
        return nmatches;
    }

    // Frame::ComputeImageBounds — src/Frame.cc:481-512 (call instead of the member, inside the
    // mbInitialComputations block of the Frame constructors, :97-111)
    template <class FrameT>
    void ComputeImageBounds(FrameT& F, const cv::Mat& imLeft) {
        const orbfe_camera cam = Camera(F);
        float b[4];
        orbfe_detail::check(orbfe_image_bounds(mpHandle, &cam, imLeft.cols, imLeft.rows, b), "orbfe_image_bounds");
        FrameT::mnMinX = b[0]; FrameT::mnMaxX = b[1]; FrameT::mnMinY = b[2]; FrameT::mnMaxY = b[3];
    }

    // Frame::UndistortKeyPoints + Frame::AssignFeaturesToGrid — src/Frame.cc:443-475, :246-267.  One
    // call where the constructor calls AssignFeaturesToGrid() (the image bounds exist by then); the
    // earlier UndistortKeyPoints() call goes away.  Fills mvKeysUn and mGrid.
    template <class FrameT>
    void UndistortAndAssign(FrameT& F) {
        const int N = (int)F.mvKeys.size();
        const orbfe_camera cam = Camera(F);
        const float b[4] = {FrameT::mnMinX, FrameT::mnMaxX, FrameT::mnMinY, FrameT::mnMaxY};
        std::vector<orbfe_keypoint> k(N), un(N);
        for (int i = 0; i < N; i++) k[i] = Kp(F.mvKeys[i]);
        std::vector<int32_t> cellEnd(ORBFE_GRID_CELLS);
        std::vector<uint16_t> items(N);
        int inGrid = 0;
        orbfe_detail::check(orbfe_frame_assign(mpHandle, &cam, b, k.data(), N, un.data(), cellEnd.data(), items.data(),
                                               &inGrid), "orbfe_frame_assign");
        F.mvKeysUn.resize(N);
        for (int i = 0; i < N; i++) {
            cv::KeyPoint kp = F.mvKeys[i];                      // :469-472
            kp.pt.x = un[i].x; kp.pt.y = un[i].y;
            F.mvKeysUn[i] = kp;
        }
        int start = 0;
        for (int ix = 0; ix < ORBFE_GRID_COLS; ix++)
            for (int iy = 0; iy < ORBFE_GRID_ROWS; iy++) {
                const int end = cellEnd[ix * ORBFE_GRID_ROWS + iy];
                F.mGrid[ix][iy].assign(items.begin() + start, items.begin() + end);
                start = end;
            }
    }

    // Frame::ComputeStereoMatches — src/Frame.cc:512-686 (call instead of the member)
    template <class FrameT>
    void ComputeStereoMatches(FrameT& F) {
        const int N = (int)F.mvKeys.size(), L = (int)F.mvScaleFactors.size();
        F.mvuRight = std::vector<float>(N, -1.0f);
        F.mvDepth = std::vector<float>(N, -1.0f);
        if (N == 0) return;
        std::vector<orbfe_keypoint> kl(N), kr(F.mvKeysRight.size());
        for (int i = 0; i < N; i++) kl[i] = Kp(F.mvKeys[i]);
        for (size_t i = 0; i < kr.size(); i++) kr[i] = Kp(F.mvKeysRight[i]);
        const cv::Mat& p0 = F.mpORBextractorLeft->mvImagePyramid[0];
        orbfe_frame_view vl = {N, kl.data(), F.mDescriptors.data, nullptr, 0.f, (float)p0.cols, 0.f, (float)p0.rows,
                               F.mvScaleFactors.data(), L};
        orbfe_frame_view vr = {(int)kr.size(), kr.data(), F.mDescriptorsRight.data, nullptr, 0.f, (float)p0.cols, 0.f,
                               (float)p0.rows, F.mvScaleFactors.data(), L};
        std::vector<const uint8_t*> dl(L), dr(L);
        std::vector<int32_t> wl(L), hl(L), wr(L), hr(L);
        std::vector<ptrdiff_t> sl(L), sr(L);
        for (int l = 0; l < L; l++) {
            const cv::Mat& a = F.mpORBextractorLeft->mvImagePyramid[l];
            const cv::Mat& b = F.mpORBextractorRight->mvImagePyramid[l];
            dl[l] = a.data; wl[l] = a.cols; hl[l] = a.rows; sl[l] = (ptrdiff_t)(size_t)a.step;
            dr[l] = b.data; wr[l] = b.cols; hr[l] = b.rows; sr[l] = (ptrdiff_t)(size_t)b.step;
        }
        orbfe_pyramid_view pl = {L, dl.data(), wl.data(), hl.data(), sl.data()};
        orbfe_pyramid_view pr = {L, dr.data(), wr.data(), hr.data(), sr.data()};
        orbfe_detail::check(orbfe_stereo_match(mpHandle, &vl, &vr, &pl, &pr, F.mvInvScaleFactors.data(), F.mbf, F.mb,
                                               F.mvuRight.data(), F.mvDepth.data()), "orbfe_stereo_match");
    }

    float mfNNratio;
    bool mbCheckOrientation;

protected:
    static orbfe_keypoint Kp(const cv::KeyPoint& k) {
        orbfe_keypoint o = {k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
        return o;
    }
    // per-point arrays of the key-frame searches
    struct Points {
        std::vector<uint8_t> valid, desc;
        std::vector<float> u, v, ur, radius;
        std::vector<int32_t> level;
        explicit Points(int n) : valid(n, 0), desc((size_t)n * 32, 0), u(n, 0.f), v(n, 0.f), ur(n, 0.f), radius(n, 0.f), level(n, 0) {}
        void set(int i, float u_, float v_, float ur_, float r_, int lvl, const cv::Mat& d) {
            valid[i] = 1; u[i] = u_; v[i] = v_; ur[i] = ur_; radius[i] = r_; level[i] = lvl;
            std::memcpy(&desc[(size_t)i * 32], d.data, 32);
        }
    };
    // Scw -> Rcw, tcw, Ow (src/ORBmatcher.cc:343-348 / :1108-1113)
    template <class MatT>
    struct Sim3Pose {
        MatT Rcw, tcw, Ow;
        explicit Sim3Pose(const MatT& Scw) {
            const MatT sRcw = Scw.rowRange(0, 3).colRange(0, 3);
            const float scw = std::sqrt(sRcw.row(0).dot(sRcw.row(0)));
            Rcw = sRcw / scw;
            tcw = Scw.rowRange(0, 3).col(3) / scw;
            Ow = -Rcw.t() * tcw;
        }
    };
    // the gates between GetWorldPos() and GetFeaturesInArea() of the world-point searches
    // (src/ORBmatcher.cc:361-407, :955-996, :1127-1168); invzDouble: "1.0/z" (:1136) versus "1/z" (:371, :966)
    template <class KeyFrameT, class MapPointT, class MatT>
    static void ProjectIntoKeyFrame(KeyFrameT* pKF, MapPointT* pMP, const MatT& Rcw, const MatT& tcw, const MatT& Ow, bool invzDouble,
                                    bool checkNormal, float th, float bf, int i, Points& P) {
        const MatT p3Dw = pMP->GetWorldPos();
        const MatT p3Dc = Rcw * p3Dw + tcw;
        if (p3Dc.template at<float>(2) < 0.0f) return;
        const float invz = invzDouble ? (float)(1.0 / p3Dc.template at<float>(2)) : 1 / p3Dc.template at<float>(2);
        const float x = p3Dc.template at<float>(0) * invz, y = p3Dc.template at<float>(1) * invz;
        const float u = pKF->fx * x + pKF->cx, v = pKF->fy * y + pKF->cy;
        if (!pKF->IsInImage(u, v)) return;
        const float ur = u - bf * invz;
        const MatT PO = p3Dw - Ow;
        const float dist3D = norm(PO);   // cv::norm, found by argument-dependent lookup at instantiation
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) return;
        if (checkNormal) {
            const MatT Pn = pMP->GetNormal();
            if (PO.dot(Pn) < 0.5 * dist3D) return;
        }
        const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
        P.set(i, u, v, ur, th * pKF->mvScaleFactors[nPredictedLevel], nPredictedLevel, pMP->GetDescriptor());
    }
    // SearchBySim3's gates on a point already expressed in the other camera (src/ORBmatcher.cc:1305-1331)
    template <class KeyFrameT, class MapPointT, class MatT>
    static void ProjectCameraPoint(KeyFrameT* pKF, MapPointT* pMP, const MatT& p3Dc, float th, int i, Points& P) {
        if (p3Dc.template at<float>(2) < 0.0) return;
        const float invz = 1.0 / p3Dc.template at<float>(2);
        const float x = p3Dc.template at<float>(0) * invz, y = p3Dc.template at<float>(1) * invz;
        const float u = pKF->fx * x + pKF->cx, v = pKF->fy * y + pKF->cy;
        if (!pKF->IsInImage(u, v)) return;
        const float dist3D = norm(p3Dc);   // cv::norm (ADL)
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) return;
        const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
        P.set(i, u, v, 0.0f, th * pKF->mvScaleFactors[nPredictedLevel], nPredictedLevel, pMP->GetDescriptor());
    }
    // a KeyFrame as a frame view: int bounds (include/KeyFrame.h:191-194) with the Frame's grid
    template <class KeyFrameT>
    static orbfe_frame_view KeyFrameView(KeyFrameT* pKF, std::vector<orbfe_keypoint>& k) {
        k.resize(pKF->mvKeysUn.size());
        for (size_t i = 0; i < k.size(); i++) k[i] = Kp(pKF->mvKeysUn[i]);
        // cell sizes are the Frame's (src/KeyFrame.cc:33): recover its float bounds from them
        const float minX = (float)pKF->mnMinX, minY = (float)pKF->mnMinY;
        orbfe_frame_view v = {(int)k.size(), k.data(), pKF->mDescriptors.data, nullptr, minX,
                              minX + (float)pKF->mnGridCols / pKF->mfGridElementWidthInv, minY,
                              minY + (float)pKF->mnGridRows / pKF->mfGridElementHeightInv, pKF->mvScaleFactors.data(),
                              (int)pKF->mvScaleFactors.size(), 1,
                              pKF->mfGridElementWidthInv, pKF->mfGridElementHeightInv};   // the Frame's own values, bit for bit
        return v;
    }
    // one key frame of SearchForTriangulation flattened
    struct TriSide {
        std::vector<orbfe_keypoint> k;
        std::vector<uint8_t> has;
        std::vector<uint32_t> nodeId, indices;
        std::vector<int32_t> nodeStart;
        orbfe_tri_side v;
        template <class KeyFrameT>
        explicit TriSide(KeyFrameT* pKF) {
            k.resize(pKF->mvKeysUn.size());
            has.assign(k.size(), 0);
            for (size_t i = 0; i < k.size(); i++) { k[i] = Kp(pKF->mvKeysUn[i]); has[i] = pKF->GetMapPoint(i) != nullptr; }
            nodeStart.push_back(0);
            for (auto it = pKF->mFeatVec.begin(); it != pKF->mFeatVec.end(); ++it) {
                nodeId.push_back((uint32_t)it->first);
                indices.insert(indices.end(), it->second.begin(), it->second.end());
                nodeStart.push_back((int32_t)indices.size());
            }
            v.n = (int32_t)k.size(); v.kps = k.data(); v.desc = pKF->mDescriptors.data; v.has_mappoint = has.data();
            v.uright = pKF->mvuRight.empty() ? nullptr : pKF->mvuRight.data();
            v.nnodes = (int32_t)nodeId.size(); v.node_id = nodeId.data(); v.node_start = nodeStart.data(); v.indices = indices.data();
        }
    };
    // one side of SearchByBoW flattened: angles, live-map-point flags, DBoW2::FeatureVector as CSR
    struct BowSide {
        std::vector<float> angle;
        std::vector<uint8_t> valid;
        std::vector<uint32_t> nodeId, indices;
        std::vector<int32_t> nodeStart;
        orbfe_bow_view v;
        template <class FeatVecT, class MapPointT>
        BowSide(const cv::Mat& desc, const std::vector<cv::KeyPoint>& keys, const FeatVecT& fv,
                const std::vector<MapPointT*>* mps) {
            angle.resize(keys.size());
            for (size_t i = 0; i < keys.size(); i++) angle[i] = keys[i].angle;
            if (mps) {
                valid.assign(keys.size(), 0);
                for (size_t i = 0; i < keys.size() && i < mps->size(); i++) valid[i] = (*mps)[i] && !(*mps)[i]->isBad() ? 1 : 0;
            }
            nodeStart.push_back(0);
            for (typename FeatVecT::const_iterator it = fv.begin(); it != fv.end(); ++it) {     // std::map: ascending NodeId
                nodeId.push_back((uint32_t)it->first);
                indices.insert(indices.end(), it->second.begin(), it->second.end());
                nodeStart.push_back((int32_t)indices.size());
            }
            v.n = (int32_t)keys.size(); v.desc = desc.data; v.angle = angle.data(); v.valid = mps ? valid.data() : nullptr;
            v.nnodes = (int32_t)nodeId.size(); v.node_id = nodeId.data(); v.node_start = nodeStart.data(); v.indices = indices.data();
        }
    };
    // mK (3x3 CV_32F) and mDistCoef (4x1 or 5x1 CV_32F) as Tracking builds them (src/Tracking.cc:62-85)
    template <class FrameT>
    static orbfe_camera Camera(const FrameT& F) {
        orbfe_camera c = {F.mK.template at<float>(0, 0), F.mK.template at<float>(1, 1), F.mK.template at<float>(0, 2),
                          F.mK.template at<float>(1, 2), F.mDistCoef.template at<float>(0, 0),
                          F.mDistCoef.template at<float>(1, 0), F.mDistCoef.template at<float>(2, 0),
                          F.mDistCoef.template at<float>(3, 0), F.mDistCoef.rows > 4 ? F.mDistCoef.template at<float>(4, 0) : 0.0f};
        return c;
    }
    // the Frame fields the searches read: mvKeysUn, mDescriptors, mnMinX.., mvScaleFactors
    template <class FrameT>
    static orbfe_frame_view View(FrameT& F, std::vector<orbfe_keypoint>& /*k*/, std::vector<uint8_t>& /*d*/) {
        // cv::KeyPoint and orbfe_keypoint are the same 28 bytes: the view points into the frame's own vector
        static_assert(sizeof(F.mvKeysUn[0]) == sizeof(orbfe_keypoint), "cv::KeyPoint layout");
        orbfe_frame_view v = {(int)F.mvKeysUn.size(), reinterpret_cast<const orbfe_keypoint*>(F.mvKeysUn.data()), F.mDescriptors.data,
                              nullptr, FrameT::mnMinX, FrameT::mnMaxX,
                              FrameT::mnMinY, FrameT::mnMaxY, F.mvScaleFactors.data(), (int)F.mvScaleFactors.size()};
        return v;
    }
    orbfe_matcher* mpHandle;
};

}  // namespace ORB_SLAM2
#endif  // ORBFE_ADAPTER_HPP
