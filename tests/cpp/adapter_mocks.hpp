// adapter_mocks.hpp — stand-ins with the reference's member names (include/Frame.h, include/KeyFrame.h,
// include/MapPoint.h) for driving the templates of include/orbfe_adapter.hpp without OpenCV / g2o / DBoW2.
// FMat stands in for the CV_32F cv::Mat of the pose arithmetic: value semantics, at most 4x4, float
// accumulation.  The poses used by the tests are identity rotations with a translation, so every product
// is exact and the Python side of the test can redo the projections in float32.
#ifndef ADAPTER_MOCKS_HPP
#define ADAPTER_MOCKS_HPP
#include <cmath>
#include <cstring>
#include <map>
#include <set>
#include <vector>

#include "cvshim.hpp"

struct FMat {
    int rows, cols;
    float a[16];
    FMat() : rows(0), cols(0) { std::memset(a, 0, sizeof a); }
    FMat(int r, int c) : rows(r), cols(c) { std::memset(a, 0, sizeof a); }
    static FMat eye(int n) { FMat m(n, n); for (int i = 0; i < n; i++) m.a[i * n + i] = 1.0f; return m; }
    static FMat vec3(float x, float y, float z) { FMat m(3, 1); m.a[0] = x; m.a[1] = y; m.a[2] = z; return m; }
    float& operator()(int r, int c) { return a[r * cols + c]; }
    float operator()(int r, int c) const { return a[r * cols + c]; }
    template <class T> T at(int i) const { return (T)a[i]; }
    template <class T> T at(int r, int c) const { return (T)a[r * cols + c]; }
    FMat block(int r0, int r1, int c0, int c1) const {
        FMat m(r1 - r0, c1 - c0);
        for (int r = r0; r < r1; r++) for (int c = c0; c < c1; c++) m(r - r0, c - c0) = (*this)(r, c);
        return m;
    }
    FMat rowRange(int r0, int r1) const { return block(r0, r1, 0, cols); }
    FMat colRange(int c0, int c1) const { return block(0, rows, c0, c1); }
    FMat row(int r) const { return block(r, r + 1, 0, cols); }
    FMat col(int c) const { return block(0, rows, c, c + 1); }
    FMat t() const {
        FMat m(cols, rows);
        for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) m(c, r) = (*this)(r, c);
        return m;
    }
    double dot(const FMat& o) const {
        float s = 0.0f;
        for (int i = 0; i < rows * cols; i++) s += a[i] * o.a[i];
        return s;
    }
};
inline FMat operator*(const FMat& x, const FMat& y) {
    FMat m(x.rows, y.cols);
    for (int r = 0; r < x.rows; r++)
        for (int c = 0; c < y.cols; c++) {
            float s = 0.0f;
            for (int k = 0; k < x.cols; k++) s += x(r, k) * y(k, c);
            m(r, c) = s;
        }
    return m;
}
inline FMat operator+(const FMat& x, const FMat& y) { FMat m(x.rows, x.cols); for (int i = 0; i < 16; i++) m.a[i] = x.a[i] + y.a[i]; return m; }
inline FMat operator-(const FMat& x, const FMat& y) { FMat m(x.rows, x.cols); for (int i = 0; i < 16; i++) m.a[i] = x.a[i] - y.a[i]; return m; }
inline FMat operator-(const FMat& x) { FMat m(x.rows, x.cols); for (int i = 0; i < 16; i++) m.a[i] = -x.a[i]; return m; }
inline FMat operator*(float s, const FMat& x) { FMat m(x.rows, x.cols); for (int i = 0; i < 16; i++) m.a[i] = s * x.a[i]; return m; }
inline FMat operator*(double s, const FMat& x) { return (float)s * x; }
inline FMat operator/(const FMat& x, float s) { FMat m(x.rows, x.cols); for (int i = 0; i < 16; i++) m.a[i] = x.a[i] / s; return m; }
inline double norm(const FMat& x) { return std::sqrt(x.dot(x)); }

typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;   // DBoW2::FeatureVector

struct KeyFrame;

struct MapPoint {
    int id;
    bool bad;
    int nobs, level;
    FMat pos, normal;
    unsigned char d[32];
    KeyFrame* kf;          // the key frame that observes it (one is enough here) and the keypoint index there
    int kfIdx;
    // fields Frame::isInFrustum fills for the local-map search (include/MapPoint.h:84-89)
    bool mbTrackInView;
    float mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos;
    int mnTrackScaleLevel;

    MapPoint() : id(-1), bad(false), nobs(1), level(0), kf(nullptr), kfIdx(-1), mbTrackInView(false), mTrackProjX(0), mTrackProjY(0),
                 mTrackProjXR(0), mTrackViewCos(1), mnTrackScaleLevel(0) { std::memset(d, 0, 32); }
    bool isBad() const { return bad; }
    FMat GetWorldPos() const { return pos; }
    FMat GetNormal() const { return normal; }
    cv::Mat GetDescriptor() const { return cv::Mat(1, 32, CV_8UC1, (void*)d, 32); }
    int Observations() const { return nobs; }
    float GetMinDistanceInvariance() const { return 0.0f; }
    float GetMaxDistanceInvariance() const { return 1e9f; }
    template <class T> int PredictScale(const float&, T*) const { return level; }
    bool IsInKeyFrame(KeyFrame* k) const { return kf == k; }
    int GetIndexInKeyFrame(KeyFrame* k) const { return kf == k ? kfIdx : -1; }
    void AddObservation(KeyFrame* k, size_t idx) { kf = k; kfIdx = (int)idx; nobs++; }
    inline void Replace(MapPoint* by);       // src/MapPoint.cc:172-215: this one turns bad, its observations move to `by`
};

struct KeyFrame {          // include/KeyFrame.h
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    FeatureVector mFeatVec;
    std::vector<MapPoint*> mps;
    std::vector<float> mvuRight, mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    float fx, fy, cx, cy, mbf;
    int mnMinX, mnMinY, mnMaxX, mnMaxY, mnGridCols, mnGridRows;
    float mfGridElementWidthInv, mfGridElementHeightInv;
    FMat R, t;
    KeyFrame() : fx(1), fy(1), cx(0), cy(0), mbf(0), mnMinX(0), mnMinY(0), mnMaxX(0), mnMaxY(0), mnGridCols(64), mnGridRows(48),
                 mfGridElementWidthInv(0), mfGridElementHeightInv(0), R(FMat::eye(3)), t(FMat::vec3(0, 0, 0)) {}
    FMat GetRotation() const { return R; }
    FMat GetTranslation() const { return t; }
    FMat GetCameraCenter() const { return -R.t() * t; }
    bool IsInImage(const float& x, const float& y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
    std::vector<MapPoint*> GetMapPointMatches() const { return mps; }
    std::set<MapPoint*> GetMapPoints() const {      // src/KeyFrame.cc:231-244: live points only
        std::set<MapPoint*> s;
        for (size_t i = 0; i < mps.size(); i++) if (mps[i] && !mps[i]->isBad()) s.insert(mps[i]);
        return s;
    }
    MapPoint* GetMapPoint(size_t idx) const { return mps[idx]; }
    void AddMapPoint(MapPoint* p, size_t idx) { mps[idx] = p; }
};

inline void MapPoint::Replace(MapPoint* by) {
    if (by == this) return;
    bad = true;
    if (kf) {
        if (!by->IsInKeyFrame(kf)) { kf->mps[kfIdx] = by; by->AddObservation(kf, kfIdx); }
        else kf->mps[kfIdx] = nullptr;
    }
}

struct Frame {             // include/Frame.h
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    cv::Mat mDescriptors;
    std::vector<float> mvScaleFactors, mvuRight;
    std::vector<MapPoint*> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    FMat mTcw;
    float fx, fy, cx, cy, mb, mbf;
    int N;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    Frame() : mTcw(FMat::eye(4)), fx(1), fy(1), cx(0), cy(0), mb(0.1f), mbf(4.0f), N(0) {}
};
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;

#endif
