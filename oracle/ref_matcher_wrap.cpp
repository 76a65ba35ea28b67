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

#ifdef ORM_DROPIN            // the drop-in build also drives the adapter's Frame-level replacements on the reference's Frame type
#define ORBFE_EXTRACTOR_CLASS ORBextractorFE
#define ORBFE_MATCHER_CLASS ORBmatcherFE
#include "orbfe_adapter.hpp"
#endif
#include <chrono>
static double g_last_call_ms = 0.0;

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
thread_local Arena g;      // one arena per calling thread (bench.py times the reference with one thread per core)

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
    const auto t0 = std::chrono::steady_clock::now();
    const int n = matcher.SearchForInitialization(*F1, *F2, prev, m12, window);
    g_last_call_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    for (int i = 0; i < f1->n; i++) { matches12[i] = m12[i]; prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y; }
    return n;
}

/* duration of the reference call itself inside the last wrapper above (without building the Frame objects around it) */
double orm_last_call_ms(void) { return g_last_call_ms; }

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


/* ---- SearchByBoW, both overloads — src/ORBmatcher.cc:185-325, :610-743.  A side is the restatement's orc_bow_view;
 * keypoints only carry their angle.  valid == NULL on side 1 means every feature has a live map point. */
static Frame* bow_frame(const orc_bow_view* s, std::vector<orc_keypoint>& k) {
    static const float scale[8] = {1.f, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985984f, 3.5831808f};
    k.resize(s->n);
    for (int i = 0; i < s->n; i++) k[i] = orc_keypoint{(float)(i % 600) + 10.f, (float)(i / 600) * 7.f + 10.f, 31.f, s->angle[i], 0.f, 0, -1};
    orc_frame_view v = {s->n, k.data(), s->desc, nullptr, 0.f, 640.f, 0.f, 480.f, scale, 8, 0};
    Frame* F = make_frame(&v);
    F->mFeatVec = feat_vec(s->nnodes, s->node_id, s->node_start, s->indices);
    return F;
}
static void bow_points(KeyFrame* K, const orc_bow_view* s) {
    // valid[i]: a live map point; the others alternate between "no point" and "bad point" (both are skipped)
    for (int i = 0; i < s->n; i++) {
        const bool live = !s->valid || s->valid[i];
        if (live || (i & 1)) K->mvpMapPoints[i] = make_point(vec3(0, 0, 1), s->desc + 32 * (size_t)i, 1, !live, 0, 1.f, 1.2f);
    }
}
int orm_search_by_bow_frame(const orc_bow_view* s1, const orc_bow_view* s2, float nnratio, int check_ori, int* m21) {
    std::vector<orc_keypoint> k1, k2;
    KeyFrame* K = make_keyframe(bow_frame(s1, k1));
    bow_points(K, s1);
    Frame* F = bow_frame(s2, k2);
    std::vector<MapPoint*> matches;
    ORBmatcher matcher(nnratio, check_ori != 0);
    const int n = matcher.SearchByBoW(K, *F, matches);
    const std::vector<MapPoint*> kfPoints = K->GetMapPointMatches();
    for (int i = 0; i < s2->n; i++) m21[i] = matches[i] ? index_of(kfPoints, matches[i]) : -1;
    return n;
}
int orm_search_by_bow_keyframes(const orc_bow_view* s1, const orc_bow_view* s2, float nnratio, int check_ori, int* m12) {
    std::vector<orc_keypoint> k1, k2;
    KeyFrame* K1 = make_keyframe(bow_frame(s1, k1));
    KeyFrame* K2 = make_keyframe(bow_frame(s2, k2));
    bow_points(K1, s1); bow_points(K2, s2);
    std::vector<MapPoint*> matches;
    ORBmatcher matcher(nnratio, check_ori != 0);
    const int n = matcher.SearchByBoW(K1, K2, matches);
    const std::vector<MapPoint*> p2 = K2->GetMapPointMatches();
    for (int i = 0; i < s1->n; i++) m12[i] = matches[i] ? index_of(p2, matches[i]) : -1;
    return n;
}

/* ---- key-frame guided searches.  Points come as the restatement's orc_projpoints_view plus the gates the reference
 * evaluates itself: in_view is NOT consulted here (the reference decides from the projection), `found` marks points of
 * sAlreadyFound / bad points, normal_ok the viewing-angle gate. */
static std::vector<MapPoint*> world_points(const orc_projpoints_view* p, const uint8_t* bad, const uint8_t* normal_ok, float scaleFactor) {
    std::vector<MapPoint*> v(p->n);
    for (int i = 0; i < p->n; i++) {
        const float dist = (float)std::sqrt((double)p->u[i] * p->u[i] + (double)p->v[i] * p->v[i] + 1.0);
        v[i] = make_point(vec3(p->u[i], p->v[i], 1.f), p->desc + 32 * (size_t)i, 1, bad && bad[i], p->level[i], dist, scaleFactor,
                          normal_ok && !normal_ok[i] ? -1.0f : 1e6f);
    }
    return v;
}
/* SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) — src/ORBmatcher.cc:1641-1770.  Point i is the map point
 * of key-frame keypoint i (angle = p->angle[i]); has_point 0 = no map point there. */
int orm_sbp_frame_keyframe(const orc_frame_view* f, const orc_projpoints_view* p, const uint8_t* has_point, const uint8_t* bad,
                           const uint8_t* found, float th, int orb_dist, int check_ori, const uint8_t* kp_occupied, int* match_for_kp) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    Frame* C = make_frame(f);
    std::vector<orc_keypoint> kk(p->n);
    std::vector<uint8_t> kd((size_t)std::max(p->n, 1) * 32, 0);
    for (int i = 0; i < p->n; i++) kk[i] = orc_keypoint{p->u[i], p->v[i], 31.f, p->angle[i], 0.f, 0, -1};
    orc_frame_view kv = *f;
    kv.n = p->n; kv.kps = kk.data(); kv.desc = kd.data(); kv.uright = nullptr;
    KeyFrame* K = make_keyframe(make_frame(&kv));
    const std::vector<MapPoint*> pts = world_points(p, bad, nullptr, C->mfScaleFactor);
    std::set<MapPoint*> already;
    for (int i = 0; i < p->n; i++) {
        if (has_point[i]) K->mvpMapPoints[i] = pts[i];
        if (found[i]) already.insert(pts[i]);
    }
    static const uint8_t zero[32] = {0};
    MapPoint* owner = make_point(vec3(0, 0, 1), zero, 1, false, 0, 1.f, 1.2f);
    for (int i = 0; i < f->n; i++) if (kp_occupied[i]) C->mvpMapPoints[i] = owner;
    ORBmatcher matcher(0.9f, check_ori != 0);
    const int n = matcher.SearchByProjection(*C, K, already, th, orb_dist);
    for (int i = 0; i < f->n; i++) {
        MapPoint* q = C->mvpMapPoints[i];
        match_for_kp[i] = (q && q != owner) ? index_of(pts, q) : -1;
    }
    return n;
}
/* SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) — src/ORBmatcher.cc:335-451, Scw = identity */
int orm_sbp_keyframe_sim3(const orc_frame_view* f, const orc_projpoints_view* p, const uint8_t* bad, const uint8_t* normal_ok, int th,
                          const uint8_t* kp_matched, int* match_for_kp) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    KeyFrame* K = make_keyframe(make_frame(f));
    const std::vector<MapPoint*> pts = world_points(p, bad, normal_ok, K->mfScaleFactor);
    static const uint8_t zero[32] = {0};
    MapPoint* owner = make_point(vec3(0, 0, 1), zero, 1, false, 0, 1.f, 1.2f);
    std::vector<MapPoint*> matched(f->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < f->n; i++) if (kp_matched[i]) matched[i] = owner;
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchByProjection(K, cv::Mat::eye(4, 4, CV_32F), pts, matched, th);
    for (int i = 0; i < f->n; i++) match_for_kp[i] = (matched[i] && matched[i] != owner) ? index_of(pts, matched[i]) : -1;
    return n;
}
/* Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint) — src/ORBmatcher.cc:1100-1237 — and Fuse(KeyFrame*, vpMapPoints, th) —
 * :940-1090 (with_pose = 1: the reprojection gates run, bf / mvuRight as in the view).  kp_has_point: key-frame keypoints that
 * hold a map point on entry.  best_idx[i] = the keypoint point i was fused with (through AddObservation, a Replace in either
 * direction, or vpReplacePoint), -1 = not fused. */
int orm_fuse(const orc_frame_view* f, const orc_projpoints_view* p, const uint8_t* bad, const uint8_t* normal_ok, float th, int with_pose,
             float bf, const uint8_t* kp_has_point, const int* kp_point_obs, const int* point_obs, int* best_idx) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    Frame* F = make_frame(f);
    F->mbf = bf; F->mb = bf;
    KeyFrame* K = make_keyframe(F);
    std::vector<MapPoint*> pts = world_points(p, bad, normal_ok, K->mfScaleFactor);
    for (int i = 0; i < p->n; i++) pts[i]->nObs = point_obs ? point_obs[i] : 1;
    std::vector<MapPoint*> own(f->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < f->n; i++)
        if (kp_has_point && kp_has_point[i]) {
            own[i] = make_point(vec3(f->kps[i].x, f->kps[i].y, 1.f), f->desc + 32 * (size_t)i, 0, false, 0, 1.f, 1.2f);
            own[i]->AddObservation(K, i);
            own[i]->nObs = kp_point_obs ? kp_point_obs[i] : 1;
            K->mvpMapPoints[i] = own[i];
            the_map()->AddMapPoint(own[i]);
        }
    for (int i = 0; i < p->n; i++) the_map()->AddMapPoint(pts[i]);
    ORBmatcher matcher(0.6f, true);
    int n;
    std::vector<MapPoint*> replace(p->n, static_cast<MapPoint*>(nullptr));
    if (with_pose) n = matcher.Fuse(K, pts, th);
    else n = matcher.Fuse(K, cv::Mat::eye(4, 4, CV_32F), pts, th, replace);
    for (int i = 0; i < p->n; i++) {
        MapPoint* q = pts[i];
        int idx = -1;
        if (replace[i]) idx = replace[i]->GetIndexInKeyFrame(K);
        else if (q->IsInKeyFrame(K)) idx = q->GetIndexInKeyFrame(K);
        else if (q->mbBad && q->mpReplaced) {                       // pMP->Replace(pMPinKF): follow to the survivor
            MapPoint* r = q->mpReplaced;
            while (r && !r->IsInKeyFrame(K) && r->mpReplaced) r = r->mpReplaced;
            if (r && r->IsInKeyFrame(K)) idx = r->GetIndexInKeyFrame(K);
        }
        best_idx[i] = idx;
    }
    return n;
}
/* SearchBySim3(pKF1, pKF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th) — src/ORBmatcher.cc:1251-1477.  Point i of p1 is the map
 * point of KF1's keypoint i (projected into KF2 at p1->(u, v)), likewise p2; has1 / has2 = keypoint holds a point;
 * pre12[i] >= 0: vpMatches12[i] preset to KF2's point pre12[i]. */
int orm_search_by_sim3(const orc_frame_view* f1, const orc_frame_view* f2, const orc_projpoints_view* p1, const orc_projpoints_view* p2,
                       const uint8_t* has1, const uint8_t* has2, const uint8_t* bad1, const uint8_t* bad2, const int* pre12, float th,
                       int* matches12) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    KeyFrame* K1 = make_keyframe(make_frame(f1));
    KeyFrame* K2 = make_keyframe(make_frame(f2));
    const std::vector<MapPoint*> a = world_points(p1, bad1, nullptr, K1->mfScaleFactor), b = world_points(p2, bad2, nullptr, K2->mfScaleFactor);
    for (int i = 0; i < f1->n; i++) if (has1[i]) { K1->mvpMapPoints[i] = a[i]; a[i]->mObservations[K1] = i; }
    for (int i = 0; i < f2->n; i++) if (has2[i]) { K2->mvpMapPoints[i] = b[i]; b[i]->mObservations[K2] = i; }
    std::vector<MapPoint*> m12(f1->n, static_cast<MapPoint*>(nullptr));
    for (int i = 0; i < f1->n; i++) if (pre12[i] >= 0) m12[i] = b[pre12[i]];
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchBySim3(K1, K2, m12, 1.0f, cv::Mat::eye(3, 3, CV_32F), cv::Mat::zeros(3, 1, CV_32F), th);
    for (int i = 0; i < f1->n; i++) matches12[i] = m12[i] ? index_of(b, m12[i]) : -1;
    return n;
}
/* SearchForTriangulation — src/ORBmatcher.cc:754-928.  The epipole (ex, ey) is produced by camera 2 sitting at t2w = (ex, ey, 1). */
int orm_search_for_triangulation(const orc_tri_side* s1, const orc_tri_side* s2, const float* F12, float ex, float ey,
                                 const float* scale_factors, int nlevels, float w, float h, int only_stereo, int check_ori, int* matches12) {
    set_camera(1.f, 1.f, 0.f, 0.f);
    KeyFrame* K[2];
    const orc_tri_side* s[2] = {s1, s2};
    static const uint8_t zero[32] = {0};
    for (int c = 0; c < 2; c++) {
        orc_frame_view v = {s[c]->n, s[c]->kps, s[c]->desc, s[c]->uright, 0.f, w, 0.f, h, scale_factors, nlevels, 0};
        Frame* F = make_frame(&v, c ? ex : 0.f, c ? ey : 0.f, c ? 1.f : 0.f);
        F->mFeatVec = feat_vec(s[c]->nnodes, s[c]->node_id, s[c]->node_start, s[c]->indices);
        K[c] = make_keyframe(F);
        for (int i = 0; i < s[c]->n; i++) if (s[c]->has_mappoint[i]) K[c]->mvpMapPoints[i] = make_point(vec3(0, 0, 1), zero, 1, false, 0, 1.f, 1.2f);
    }
    cv::Mat Fm(3, 3, CV_32F);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Fm.at<float>(r, c) = F12[3 * r + c];
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher matcher(0.6f, check_ori != 0);
    const int n = matcher.SearchForTriangulation(K[0], K[1], Fm, pairs, only_stereo != 0);
    for (int i = 0; i < s1->n; i++) matches12[i] = -1;
    for (size_t i = 0; i < pairs.size(); i++) matches12[pairs[i].first] = (int)pairs[i].second;
    return n;
}

/* MapPoint::ComputeDistinctiveDescriptors — src/MapPoint.cc:266-334: point p is observed by one key frame per row of its set */
void orm_distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* best) {
    static const float one = 1.0f;
    static const uint8_t zero[32] = {0};
    for (int p = 0; p < npoints; p++) {
        const int n = start[p + 1] - start[p];
        std::vector<orc_keypoint> k(std::max(n, 1), orc_keypoint{5.f, 5.f, 31.f, 0.f, 0.f, 0, -1});
        orc_frame_view v = {n, k.data(), desc + 32 * (size_t)start[p], nullptr, 0.f, 640.f, 0.f, 480.f, &one, 1, 0};
        Frame* F = make_frame(&v);
        MapPoint* mp = make_point(vec3(0, 0, 1), zero, 0, false, 0, 1.f, 1.2f);
        // one key frame per observation, each seeing the point at its own row
        std::vector<KeyFrame*> ks(n);
        for (int i = 0; i < n; i++) ks[i] = make_keyframe(F);
        std::sort(ks.begin(), ks.end());                       // std::map order = pointer order: row i goes to the i-th key frame
        for (int i = 0; i < n; i++) mp->mObservations[ks[i]] = i;
        mp->mDescriptor.release();
        mp->ComputeDistinctiveDescriptors();
        best[p] = -1;
        if (n) {
            const cv::Mat d = mp->GetDescriptor();
            // the reference keeps the first row (in its iteration order) with the least median; rows may repeat, so report
            // the winner by content: the lowest row index holding the chosen bytes
            for (int i = 0; i < n && best[p] < 0; i++) if (!std::memcmp(d.data, desc + 32 * (size_t)(start[p] + i), 32)) best[p] = i;
        }
    }
}

/* Frame::ComputeStereoMatches (src/Frame.cc:512-686) on given keypoint lists: the two extractors only supply their
 * mvImagePyramid (built here, single-threaded, by running them on the images; their own keypoints are discarded). */
void orm_compute_stereo_matches(const orc_frame_view* fl, const orc_frame_view* fr, const uint8_t* left, const uint8_t* right, int w, int h,
                                float fx, float bf, float* uright, float* depth) {
    ORBextractor exL(1000, 1.2f, fl->nlevels, 20, 7), exR(1000, 1.2f, fl->nlevels, 20, 7);
    cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
    {
        std::vector<cv::KeyPoint> k; cv::Mat d;
        exL(imL, cv::Mat(), k, d);
        exR(imR, cv::Mat(), k, d);
    }
    Frame* F = make_frame(fl);
    F->mvKeysRight.resize(fr->n);
    for (int i = 0; i < fr->n; i++) F->mvKeysRight[i] = to_cv(fr->kps[i]);
    F->mDescriptorsRight.create(std::max(fr->n, 1), 32, CV_8UC1);
    if (fr->n) std::memcpy(F->mDescriptorsRight.data, fr->desc, (size_t)fr->n * 32);
    F->mpORBextractorLeft = &exL; F->mpORBextractorRight = &exR;
    F->mbf = bf; F->mb = bf / fx;                                  // src/Frame.cc:109
    F->ComputeStereoMatches();
    for (int i = 0; i < fl->n; i++) { uright[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i]; }
    F->mpORBextractorLeft = F->mpORBextractorRight = nullptr;
}


/* Frame::Frame(imGray, timeStamp, extractor, voc, K, distCoef, bf, thDepth) — src/Frame.cc:173-243: extraction, then the
 * epilogue UndistortKeyPoints (:443-475), ComputeImageBounds (:477-510) and AssignFeaturesToGrid (:246-267).  Returns N;
 * kps = mvKeys, kps_un = mvKeysUn, bounds = mnMinX, mnMaxX, mnMinY, mnMaxY, grid in CSR form. */
int orm_mono_frame(const uint8_t* img, int w, int h, int nfeatures, const float* cam, const float* dist, int ndist, orc_keypoint* kps,
                   orc_keypoint* kps_un, float* bounds, int* cell_start, int* cell_items, int cap) {
    // the extractor outlives the call, as Tracking's does (src/Tracking.cc:122-128)
    static thread_local ORBextractor* pex = nullptr;
    static thread_local int pexFeatures = 0;
    if (!pex || pexFeatures != nfeatures) { delete pex; pex = new ORBextractor(nfeatures, 1.2f, 8, 20, 7); pexFeatures = nfeatures; }
    ORBextractor& ex = *pex;
    cv::Mat im(h, w, CV_8UC1, (void*)img, (size_t)w);
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3];
    cv::Mat D(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; i++) D.at<float>(i) = dist[i];
    Frame::mbInitialComputations = true;
    Frame* F = new Frame(im, 0.0, &ex, nullptr, K, D, 0.0f, 0.0f);
    g.frames.push_back(F);
    F->mpORBextractorLeft = nullptr;
    if (F->N > cap) return -1;
    for (int i = 0; i < F->N; i++) {
        const cv::KeyPoint &a = F->mvKeys[i], &b = F->mvKeysUn[i];
        kps[i] = orc_keypoint{a.pt.x, a.pt.y, a.size, a.angle, a.response, a.octave, a.class_id};
        kps_un[i] = orc_keypoint{b.pt.x, b.pt.y, b.size, b.angle, b.response, b.octave, b.class_id};
    }
    bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY;
    int n = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = n;
            for (size_t j = 0; j < F->mGrid[ix][iy].size(); j++) cell_items[n++] = (int)F->mGrid[ix][iy][j];
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
    return F->N;
}


/* Frame::Frame(imLeft, imRight, ...) — src/Frame.cc:58-117: both extractions on two threads (:78-81), UndistortKeyPoints,
 * ComputeStereoMatches, AssignFeaturesToGrid.  Used with the drop-in build, whose extraction does not depend on the heap
 * layout (with the unmodified extractor on glibc malloc the quadtree ties of two concurrent threads vary from run to run).
 * Returns N; the right image's keypoints come back too. */
int orm_stereo_frame(const uint8_t* left, const uint8_t* right, int w, int h, int nfeatures, float fx, float fy, float cx, float cy,
                     float bf, orc_keypoint* kps, uint8_t* desc, float* uright, float* depth, orc_keypoint* kps_right, uint8_t* desc_right,
                     int* n_right, int cap) {
    ORBextractor exL(nfeatures, 1.2f, 8, 20, 7), exR(nfeatures, 1.2f, 8, 20, 7);
    cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = fx; K.at<float>(1, 1) = fy; K.at<float>(0, 2) = cx; K.at<float>(1, 2) = cy;
    cv::Mat D = cv::Mat::zeros(4, 1, CV_32F);
    Frame::mbInitialComputations = true;
    Frame* F = new Frame(imL, imR, 0.0, &exL, &exR, nullptr, K, D, bf, 40.0f);
    g.frames.push_back(F);
    F->mpORBextractorLeft = F->mpORBextractorRight = nullptr;       // the extractors die with this call
    const int nr = (int)F->mvKeysRight.size();
    if (F->N > cap || nr > cap) return -1;
    for (int i = 0; i < F->N; i++) {
        const cv::KeyPoint& k = F->mvKeys[i];
        kps[i] = orc_keypoint{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
        uright[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i];
    }
    if (F->N) std::memcpy(desc, F->mDescriptors.data, (size_t)F->N * 32);
    for (int i = 0; i < nr; i++) {
        const cv::KeyPoint& k = F->mvKeysRight[i];
        kps_right[i] = orc_keypoint{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
    }
    if (nr) std::memcpy(desc_right, F->mDescriptorsRight.data, (size_t)nr * 32);
    *n_right = nr;
    return F->N;
}


#ifdef ORM_DROPIN
/* The adapter's replacement for Frame::ComputeStereoMatches (include/orbfe_adapter.hpp) on a reference
 * Frame: same setup as orm_compute_stereo_matches, but the device routine runs instead of the member. */
void orm_dropin_compute_stereo_matches(const orc_frame_view* fl, const orc_frame_view* fr, const uint8_t* left, const uint8_t* right, int w,
                                       int h, float fx, float bf, float* uright, float* depth) {
    ORBextractor exL(1000, 1.2f, fl->nlevels, 20, 7), exR(1000, 1.2f, fl->nlevels, 20, 7);
    cv::Mat imL(h, w, CV_8UC1, (void*)left, (size_t)w), imR(h, w, CV_8UC1, (void*)right, (size_t)w);
    {
        std::vector<cv::KeyPoint> k; cv::Mat d;
        exL(imL, cv::Mat(), k, d);
        exR(imR, cv::Mat(), k, d);
    }
    Frame* F = make_frame(fl);
    F->mvKeysRight.resize(fr->n);
    for (int i = 0; i < fr->n; i++) F->mvKeysRight[i] = to_cv(fr->kps[i]);
    F->mDescriptorsRight.create(std::max(fr->n, 1), 32, CV_8UC1);
    if (fr->n) std::memcpy(F->mDescriptorsRight.data, fr->desc, (size_t)fr->n * 32);
    F->mpORBextractorLeft = &exL; F->mpORBextractorRight = &exR;
    F->mbf = bf; F->mb = bf / fx;
    ORBmatcherFE fe;
    fe.ComputeStereoMatches(*F);
    for (int i = 0; i < fl->n; i++) { uright[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i]; }
    F->mpORBextractorLeft = F->mpORBextractorRight = nullptr;
}

/* The adapter's replacements for ComputeImageBounds + UndistortKeyPoints + AssignFeaturesToGrid applied to a
 * Frame the reference's monocular constructor built: its own epilogue results are wiped first. */
int orm_dropin_epilogue(const uint8_t* img, int w, int h, int nfeatures, const float* cam, const float* dist, int ndist, orc_keypoint* kps,
                        orc_keypoint* kps_un, float* bounds, int* cell_start, int* cell_items, int cap) {
    static thread_local ORBextractor* pex = nullptr;
    static thread_local int pexFeatures = 0;
    if (!pex || pexFeatures != nfeatures) { delete pex; pex = new ORBextractor(nfeatures, 1.2f, 8, 20, 7); pexFeatures = nfeatures; }
    cv::Mat im(h, w, CV_8UC1, (void*)img, (size_t)w);
    cv::Mat K = cv::Mat::eye(3, 3, CV_32F);
    K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3];
    cv::Mat D(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; i++) D.at<float>(i) = dist[i];
    Frame::mbInitialComputations = true;
    Frame* F = new Frame(im, 0.0, pex, nullptr, K, D, 0.0f, 0.0f);
    g.frames.push_back(F);
    F->mpORBextractorLeft = nullptr;
    if (F->N > cap) return -1;
    F->mvKeysUn.clear();
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++) for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) F->mGrid[ix][iy].clear();
    Frame::mnMinX = Frame::mnMaxX = Frame::mnMinY = Frame::mnMaxY = 0.0f;
    ORBmatcherFE fe;
    fe.ComputeImageBounds(*F, im);
    fe.UndistortAndAssign(*F);
    for (int i = 0; i < F->N; i++) {
        const cv::KeyPoint &a = F->mvKeys[i], &b = F->mvKeysUn[i];
        kps[i] = orc_keypoint{a.pt.x, a.pt.y, a.size, a.angle, a.response, a.octave, a.class_id};
        kps_un[i] = orc_keypoint{b.pt.x, b.pt.y, b.size, b.angle, b.response, b.octave, b.class_id};
    }
    bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY;
    int n = 0;
    for (int ix = 0; ix < FRAME_GRID_COLS; ix++)
        for (int iy = 0; iy < FRAME_GRID_ROWS; iy++) {
            cell_start[ix * FRAME_GRID_ROWS + iy] = n;
            for (size_t j = 0; j < F->mGrid[ix][iy].size(); j++) cell_items[n++] = (int)F->mGrid[ix][iy][j];
        }
    cell_start[FRAME_GRID_COLS * FRAME_GRID_ROWS] = n;
    return F->N;
}
#endif

}  // extern "C"
