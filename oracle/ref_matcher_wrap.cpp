// ref_matcher_wrap.cpp — ORACLE support (test infrastructure only).
//
// Drives the UNMODIFIED reference translation units src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc,
// src/MapPoint.cc, src/Map.cc (+ the vendored Thirdparty/DBoW2 sources) — compiled where they lie under
// /root/reference against oracle/cvshim — through a C interface shaped like the restatement's (orb_oracle.h), so
// that tests/test_oracle_vs_ref_matcher.py can run restatement and reference on the same inputs.  Every search,
// candidate lookup (Frame / KeyFrame::GetFeaturesInArea), grid assignment, stereo match and descriptor median below
// executes reference code; this file only builds the reference's own objects from flat arrays and reads the
// results back.  Objects are built through the reference's constructors where one fits and by filling members
// otherwise (private / protected members are reached with the access-specifier defines below, in this TU only).
// World points are placed at (u, v, 1) in front of axis-aligned cameras with fx = fy = 1, cx = cy = 0, so every
// projection the reference computes is exact and equals the (u, v) the restatement is given.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include "orb_oracle.h"
#include "cvshim.hpp"

#define private public
#define protected public
#include "Frame.h"
#include "KeyFrame.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#include "ORBextractor.h"
#undef private
#undef protected

using namespace ORB_SLAM2;

namespace {

// everything built for a call lives until the next orm_reset()
struct Arena {
    std::vector<Frame*> frames;
    std::vector<KeyFrame*> keyframes;
    std::vector<MapPoint*> points;
    Map* map = nullptr;
    KeyFrame* carrier = nullptr;          // reference key frame handed to the MapPoint constructor
    ~Arena() { clear(); }
    void clear() {
        for (auto* p : points) delete p;
        for (auto* k : keyframes) delete k;
        for (auto* f : frames) delete f;
        points.clear(); keyframes.clear(); frames.clear();
        delete map; map = nullptr; carrier = nullptr;
    }
};
Arena g;

cv::KeyPoint to_cv(const orc_keypoint& k) { return cv::KeyPoint(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id); }

void set_camera(float fx, float fy, float cx, float cy) {
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    Frame::invfx = 1.0f / fx; Frame::invfy = 1.0f / fy;
}

cv::Mat pose(float tx, float ty, float tz) {
    cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
    T.at<float>(0, 3) = tx; T.at<float>(1, 3) = ty; T.at<float>(2, 3) = tz;
    return T;
}

cv::Mat vec3(float x, float y, float z) {
    cv::Mat m(3, 1, CV_32F);
    m.at<float>(0) = x; m.at<float>(1) = y; m.at<float>(2) = z;
    return m;
}

// A Frame holding the view's keypoints / descriptors; the grid is filled by the reference's AssignFeaturesToGrid
// (src/Frame.cc:246-267) with the static bounds of the view.
Frame* make_frame(const orc_frame_view* v, float tx = 0.f, float ty = 0.f, float tz = 0.f) {
    Frame* F = new Frame();
    g.frames.push_back(F);
    F->mpORBvocabulary = nullptr; F->mpORBextractorLeft = F->mpORBextractorRight = nullptr;
    F->mTimeStamp = 0.0; F->mbf = 0.f; F->mb = 0.f; F->mThDepth = 0.f;
    F->mnId = Frame::nNextId++;
    F->mpReferenceKF = nullptr;
    F->N = v->n;
    F->mvKeys.resize(v->n);
    for (int i = 0; i < v->n; i++) F->mvKeys[i] = to_cv(v->kps[i]);
    F->mvKeysUn = F->mvKeys;
    F->mDescriptors.create(std::max(v->n, 1), 32, CV_8UC1);
    if (v->n) std::memcpy(F->mDescriptors.data, v->desc, (size_t)v->n * 32);
    if (!v->n) F->mDescriptors = F->mDescriptors.rowRange(0, 0);
    F->mvuRight.assign(v->n, -1.0f);
    F->mvDepth.assign(v->n, -1.0f);
    if (v->uright) F->mvuRight.assign(v->uright, v->uright + v->n);
    F->mvpMapPoints.assign(v->n, static_cast<MapPoint*>(nullptr));
    F->mvbOutlier.assign(v->n, false);
    F->mnScaleLevels = v->nlevels;
    F->mvScaleFactors.assign(v->scale_factors, v->scale_factors + v->nlevels);
    F->mfScaleFactor = v->nlevels > 1 ? v->scale_factors[1] : 1.2f;
    F->mfLogScaleFactor = std::log(F->mfScaleFactor);
    F->mvInvScaleFactors.resize(v->nlevels); F->mvLevelSigma2.resize(v->nlevels); F->mvInvLevelSigma2.resize(v->nlevels);
    for (int l = 0; l < v->nlevels; l++) {      // as ORBextractor::ORBextractor fills them (src/ORBextractor.cc:424-436)
        F->mvLevelSigma2[l] = F->mvScaleFactors[l] * F->mvScaleFactors[l];
        F->mvInvScaleFactors[l] = 1.0f / F->mvScaleFactors[l];
        F->mvInvLevelSigma2[l] = 1.0f / F->mvLevelSigma2[l];
    }
    Frame::mnMinX = v->min_x; Frame::mnMaxX = v->max_x; Frame::mnMinY = v->min_y; Frame::mnMaxY = v->max_y;
    // src/Frame.cc:104-105
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
    Frame::mbInitialComputations = false;
    F->mK = cv::Mat::eye(3, 3, CV_32F);
    F->mDistCoef = cv::Mat::zeros(4, 1, CV_32F);
    F->AssignFeaturesToGrid();
    F->SetPose(pose(tx, ty, tz));
    return F;
}

Map* the_map() {
    if (!g.map) g.map = new Map();
    return g.map;
}

KeyFrame* make_keyframe(Frame* F) {
    KeyFrame* K = new KeyFrame(*F, the_map(), nullptr);       // src/KeyFrame.cc:30-61: truncates the bounds to int
    g.keyframes.push_back(K);
    return K;
}

// A map point at `pos` whose descriptor, observation count, bad flag and predicted level are given.  The level
// comes out of the reference's own PredictScale (src/MapPoint.cc:425-458): mfMaxDistance is placed half a level
// above dist * 1.2^level, which also keeps the distance-invariance gates open.
MapPoint* make_point(const cv::Mat& pos, const uint8_t* desc, int nobs, bool bad, int level, float dist, float scaleFactor,
                     float normal_z = 1e6f) {
    if (!g.carrier) {
        static const float one = 1.0f;
        static const orc_keypoint k0 = {0.f, 0.f, 31.f, 0.f, 0.f, 0, -1};
        static const uint8_t d0[32] = {0};
        orc_frame_view v = {1, &k0, d0, nullptr, Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY, &one, 1, 0};
        const float keep[6] = {Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY, Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv};
        g.carrier = make_keyframe(make_frame(&v));
        Frame::mnMinX = keep[0]; Frame::mnMaxX = keep[1]; Frame::mnMinY = keep[2]; Frame::mnMaxY = keep[3];
        Frame::mfGridElementWidthInv = keep[4]; Frame::mfGridElementHeightInv = keep[5];
    }
    MapPoint* p = new MapPoint(pos, g.carrier, the_map());
    g.points.push_back(p);
    p->mDescriptor.create(1, 32, CV_8UC1);
    std::memcpy(p->mDescriptor.data, desc, 32);
    p->nObs = nobs;
    p->mbBad = bad;
    p->mfMinDistance = 0.0f;
    p->mfMaxDistance = dist * std::pow(scaleFactor, (float)level - 0.5f);
    p->mNormalVector = vec3(0.f, 0.f, normal_z);
    return p;
}

int index_of(const std::vector<MapPoint*>& v, MapPoint* p) {
    for (size_t i = 0; i < v.size(); i++) if (v[i] == p) return (int)i;
    return -1;
}

DBoW2::FeatureVector feat_vec(int nnodes, const uint32_t* node_id, const int* node_start, const uint32_t* indices) {
    DBoW2::FeatureVector fv;
    for (int k = 0; k < nnodes; k++)
        for (int j = node_start[k]; j < node_start[k + 1]; j++) fv.addFeature(node_id[k], indices[j]);
    return fv;
}

}  // namespace

extern "C" {

void orm_reset(void) { g.clear(); }

/* Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea of the reference on a view */
void orm_assign_grid(const orc_frame_view* v, int* cell_start, int* cell_items) {
    Frame* F = make_frame(v);
    int n = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = n;
            for (size_t j = 0; j < F->mGrid[ix][iy].size(); j++) cell_items[n++] = (int)F->mGrid[ix][iy][j];
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
}
int orm_features_in_area(const orc_frame_view* v, float x, float y, float r, int minLevel, int maxLevel, int* out, int cap) {
    std::vector<size_t> idx;
    if (v->keyframe_bounds) {
        if (minLevel != -1 || maxLevel != -1) return -1;          // KeyFrame::GetFeaturesInArea has no level arguments
        idx = make_keyframe(make_frame(v))->GetFeaturesInArea(x, y, r);
    } else {
        idx = make_frame(v)->GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    }
    for (size_t i = 0; i < idx.size() && (int)i < cap; i++) out[i] = (int)idx[i];
    return (int)idx.size();
}

/* ORBmatcher::SearchForInitialization — src/ORBmatcher.cc:469-600 */
int orm_search_for_initialization(const orc_frame_view* f1, const orc_frame_view* f2, float* prev_matched, int* matches12,
                                  int window, float nnratio, int check_ori) {
    Frame* F1 = make_frame(f1);
    Frame* F2 = make_frame(f2);
    std::vector<cv::Point2f> prev(f1->n);
    for (int i = 0; i < f1->n; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_ori != 0);
    const int n = matcher.SearchForInitialization(*F1, *F2, prev, m12, window);
    for (int i = 0; i < f1->n; i++) { matches12[i] = m12[i]; prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y; }
    return n;
}

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) — src/ORBmatcher.cc:51-138 */
int orm_search_by_projection(const orc_frame_view* f, const orc_mappoints_view* mps, float th, float nnratio,
                             const uint8_t* kp_occupied, int* out_mp_for_kp) {
    Frame* F = make_frame(f);
    static const uint8_t zero[32] = {0};
    MapPoint* owner = make_point(vec3(0, 0, 1), zero, 1, false, 0, 1.f, 1.2f);     // a point with observations on the occupied keypoints
    for (int i = 0; i < f->n; i++) if (kp_occupied[i]) F->mvpMapPoints[i] = owner;
    std::vector<MapPoint*> pts(mps->n);
    for (int i = 0; i < mps->n; i++) {
        MapPoint* p = make_point(vec3(0, 0, 1), mps->desc + 32 * (size_t)i, mps->observations[i], false, 0, 1.f, 1.2f);
        p->mbTrackInView = mps->track_in_view[i] != 0;
        p->mTrackProjX = mps->proj_x[i]; p->mTrackProjY = mps->proj_y[i]; p->mTrackProjXR = mps->proj_xr[i];
        p->mnTrackScaleLevel = mps->level[i]; p->mTrackViewCos = mps->view_cos[i];
        pts[i] = p;
    }
    ORBmatcher matcher(nnratio, true);
    const int n = matcher.SearchByProjection(*F, pts, th);
    for (int i = 0; i < f->n; i++) {
        MapPoint* p = F->mvpMapPoints[i];
        out_mp_for_kp[i] = (p && p != owner) ? index_of(pts, p) : -1;
    }
    return n;
}

/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) — src/ORBmatcher.cc:1484-1630.  The last
 * frame's point i sits at (u[i], v[i], 1) seen by cameras at the origin; last_octave / angle are its keypoint's.
 * bf > 0 with uright in the view runs the stereo branch (ur = u - bf). */
int orm_search_by_projection_frame(const orc_frame_view* f, int npts, const uint8_t* has_point, const uint8_t* outlier,
                                   const float* u, const float* v, const int* last_octave, const float* angle, const uint8_t* desc,
                                   const int* observations, float th, int mono, float bf, int check_ori, const uint8_t* kp_occupied,
                                   int* out_match_for_kp) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    Frame* C = make_frame(f);
    C->mbf = bf; C->mb = bf;                                   // fx = 1
    std::vector<orc_keypoint> lk(npts);
    std::vector<uint8_t> ld((size_t)std::max(npts, 1) * 32, 0);
    for (int i = 0; i < npts; i++) { lk[i] = orc_keypoint{u[i], v[i], 31.f, angle[i], 0.f, last_octave[i], -1}; }
    orc_frame_view lv = *f;
    lv.n = npts; lv.kps = lk.data(); lv.desc = ld.data(); lv.uright = nullptr;
    Frame* L = make_frame(&lv);
    L->mbf = bf; L->mb = bf;
    static const uint8_t zero[32] = {0};
    MapPoint* owner = make_point(vec3(0, 0, 1), zero, 1, false, 0, 1.f, 1.2f);
    for (int i = 0; i < f->n; i++) if (kp_occupied[i]) C->mvpMapPoints[i] = owner;
    for (int i = 0; i < npts; i++) {
        if (has_point[i]) L->mvpMapPoints[i] = make_point(vec3(u[i], v[i], 1.f), desc + 32 * (size_t)i, observations[i], false, 0, 1.f, 1.2f);
        L->mvbOutlier[i] = outlier[i] != 0;
    }
    ORBmatcher matcher(0.9f, check_ori != 0);
    const int n = matcher.SearchByProjection(*C, *L, th, mono != 0);
    for (int i = 0; i < f->n; i++) {
        MapPoint* p = C->mvpMapPoints[i];
        if (kp_occupied[i]) out_match_for_kp[i] = p == owner ? -1 : (p ? index_of(L->mvpMapPoints, p) : -2);
        else out_match_for_kp[i] = p ? index_of(L->mvpMapPoints, p) : -1;
    }
    return n;
}

}  // extern "C"
