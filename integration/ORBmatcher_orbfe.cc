// ORBmatcher_orbfe.cc — link-level drop-in for the reference's src/ORBmatcher.cc.
//
// Defines every member of the reference's OWN class ORB_SLAM2::ORBmatcher (include/ORBmatcher.h:51-117, header
// unchanged): the twelve searches forward to the adapter templates of include/orbfe_adapter.hpp instantiated on the
// reference's real Frame / KeyFrame / MapPoint types, i.e. candidates, Hamming distances, bookkeeping and rotation
// checks run in liborbfe's CUDA kernels while the pose arithmetic in front of them is the reference's own cv::Mat
// code.  Take src/ORBmatcher.cc out of the source list (CMakeLists.txt:51-69), add this file, link liborbfe.
// One device handle per host thread (Tracking, LocalMapping and LoopClosing each own a thread).
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "ORBmatcher.h"

#define ORBFE_EXTRACTOR_CLASS ORBextractorFE
#define ORBFE_MATCHER_CLASS ORBmatcherFE
#include "orbfe_adapter.hpp"

using namespace std;

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;      // src/ORBmatcher.cc:37-39
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

namespace {
ORBmatcherFE& fe(float nnratio, bool checkOri) {
    static thread_local ORBmatcherFE* m = new ORBmatcherFE();
    m->mfNNratio = nnratio; m->mbCheckOrientation = checkOri;
    return *m;
}
}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbfe_hamming256(a.data, b.data); }   // :1828

int ORBmatcher::SearchByProjection(Frame& F, const vector<MapPoint*>& vpMapPoints, const float th) {                    // :51
    return fe(mfNNratio, mbCheckOrientation).SearchByProjection(F, vpMapPoints, th);
}
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {     // :1484
    return fe(mfNNratio, mbCheckOrientation).SearchByProjection(CurrentFrame, LastFrame, th, bMono);
}
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, const float th,
                                   const int ORBdist) {                                                                  // :1641
    return fe(mfNNratio, mbCheckOrientation).SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist);
}
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, vector<MapPoint*>& vpMatched, int th) {   // :335
    return fe(mfNNratio, mbCheckOrientation).SearchByProjection(pKF, Scw, vpPoints, vpMatched, th);
}
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, vector<MapPoint*>& vpMapPointMatches) {                            // :185
    return fe(mfNNratio, mbCheckOrientation).SearchByBoW(pKF, F, vpMapPointMatches);
}
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12) {                           // :610
    return fe(mfNNratio, mbCheckOrientation).SearchByBoW(pKF1, pKF2, vpMatches12);
}
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vector<cv::Point2f>& vbPrevMatched, vector<int>& vnMatches12,
                                        int windowSize) {                                                                // :469
    return fe(mfNNratio, mbCheckOrientation).SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize);
}
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, vector<pair<size_t, size_t> >& vMatchedPairs,
                                       const bool bOnlyStereo) {                                                         // :754
    return fe(mfNNratio, mbCheckOrientation).SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo);
}
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                             const cv::Mat& t12, const float th) {                                                       // :1251
    return fe(mfNNratio, mbCheckOrientation).SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th);
}
int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th) {                             // :940
    return fe(mfNNratio, mbCheckOrientation).Fuse(pKF, vpMapPoints, th);
}
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint) {   // :1100
    return fe(mfNNratio, mbCheckOrientation).Fuse(pKF, Scw, vpPoints, th, vpReplacePoint);
}

// The three protected helpers of the class (RadiusByViewingCos, CheckDistEpipolarLine, ComputeThreeMaxima) were only ever
// called from inside src/ORBmatcher.cc; nothing refers to them once that file is out of the build, so they get no
// definition here — their arithmetic lives in the device kernels.

}  // namespace ORB_SLAM2
