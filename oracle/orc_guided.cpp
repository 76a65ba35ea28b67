// orc_guided.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// Plain-array restatements of the remaining guided searches of ORBmatcher (SURVEY.md §8(f) ranks 3
// and 4), one function per reference routine:
//   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)      src/ORBmatcher.cc:1641-1770
//   SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)            src/ORBmatcher.cc:335-451
//   Fuse(KeyFrame*, vpMapPoints, th)                                       src/ORBmatcher.cc:940-1090
//   Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)                     src/ORBmatcher.cc:1100-1237
//   SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th)               src/ORBmatcher.cc:1251-1477
//   SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo)    src/ORBmatcher.cc:754-928
//   CheckDistEpipolarLine                                                  src/ORBmatcher.cc:156-175
//   KeyFrame::GetFeaturesInArea (= Frame's without a level filter)         src/KeyFrame.cc:611-650
// As with the other matcher routines the pose arithmetic in front of each loop body (cv::Mat
// products, PredictScale, the distance / viewing-angle gates) is the caller's and arrives as the
// per-point `valid`, (u, v), `ur`, `radius` and predicted `level`; what is restated here is the
// part that reads keypoints and descriptors.  Map mutation (Replace, AddObservation, AddMapPoint)
// stays with the caller as well: the Fuse routines return the chosen keypoint per map point.
// Parity: pinned against the unmodified src/ORBmatcher.cc (oracle/_ref/liborbref_matcher.so,
// tests/test_oracle_vs_ref_matcher.py::guided) besides the hand-derived known-answer tests
// (tests/test_oracle_guided.py).
#include "orb_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <vector>

namespace {
const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // ORBmatcher.cc:37-39
const int GRID_COLS = 64, GRID_ROWS = 48;

struct Grid {
    std::vector<int> cs, ci, cand;
    explicit Grid(const orc_frame_view* f) : cs(GRID_COLS * GRID_ROWS + 1), ci(std::max(f->n, 1)), cand(std::max(f->n, 1)) {
        orc_assign_grid(f, cs.data(), ci.data());
    }
    // KeyFrame::GetFeaturesInArea(x, y, r) — KeyFrame.cc:611-650 (no level filter)
    int area(const orc_frame_view* f, float x, float y, float r, int minLevel = -1, int maxLevel = -1) {
        return orc_features_in_area(f, cs.data(), ci.data(), x, y, r, minLevel, maxLevel, cand.data(), f->n);
    }
};

int rot_bin(float a1, float a2) {
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}
}  // namespace

// SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist) — :1641-1770, loop
// body from :1705 (GetFeaturesInArea with nPredictedLevel-1 .. nPredictedLevel+1).  kp_occupied:
// CurrentFrame.mvpMapPoints[i2] != NULL (:1720).  match_for_kp as in orc_search_by_projection_frame.
extern "C" int orc_sbp_frame_keyframe(const orc_frame_view* f, const orc_projpoints_view* pts, int orb_dist,
                                      int check_orientation, const uint8_t* kp_occupied, int* match_for_kp) {
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    Grid G(f);
    std::vector<uint8_t> occ(kp_occupied, kp_occupied + f->n);
    for (int i = 0; i < pts->n; i++) {
        if (!pts->valid[i]) continue;                                                  // :1663-1698
        const int lvl = pts->level[i];
        const int nc = G.area(f, pts->u[i], pts->v[i], pts->radius[i], lvl - 1, lvl + 1);   // :1705
        if (nc == 0) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int c = 0; c < nc; c++) {
            const int i2 = G.cand[c];
            if (occ[i2]) continue;                                                     // :1720-1721
            const int dist = orc_descriptor_distance(dMP, f->desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= orb_dist) {                                                    // :1734
            match_for_kp[bestIdx2] = i; occ[bestIdx2] = 1;
            nmatches++;
            if (check_orientation) rotHist[rot_bin(pts->angle[i], f->kps[bestIdx2].angle)].push_back(bestIdx2);   // :1739-1748
        }
    }
    if (check_orientation) {                                                           // :1754-1772
        int hist[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) hist[i] = (int)rotHist[i].size();
        orc_three_maxima(hist, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int k : rotHist[i]) { match_for_kp[k] = -2; nmatches--; }
    }
    return nmatches;
}

// SearchByProjection(KeyFrame* pKF, Scw, vpPoints, vpMatched, th) — :335-451, loop body from :404.
// kp_matched: vpMatched[idx] != NULL on entry.  match_for_kp[idx] = iMP written where vpMatched[idx] = pMP.
extern "C" int orc_sbp_keyframe_sim3(const orc_frame_view* f, const orc_projpoints_view* pts,
                                     const uint8_t* kp_matched, int* match_for_kp) {
    int nmatches = 0;
    Grid G(f);
    std::vector<uint8_t> occ(kp_matched, kp_matched + f->n);
    for (int i = 0; i < pts->n; i++) {
        if (!pts->valid[i]) continue;                                                  // :361-402
        const int lvl = pts->level[i];
        const int nc = G.area(f, pts->u[i], pts->v[i], pts->radius[i]);                // :409
        if (nc == 0) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx = -1;
        for (int c = 0; c < nc; c++) {
            const int idx = G.cand[c];
            if (occ[idx]) continue;                                                    // :421-422
            const int kpLevel = f->kps[idx].octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;                          // :426-427
            const int dist = orc_descriptor_distance(dMP, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_LOW) { match_for_kp[bestIdx] = i; occ[bestIdx] = 1; nmatches++; }   // :440-444
    }
    return nmatches;
}

// The search of both Fuse overloads: best_idx[i] = bestIdx where bestDist <= TH_LOW, else -1.
// inv_level_sigma2 != NULL: Fuse(pKF, vpMapPoints, th) with its reprojection gate (:1009-1037, 7.8 for
// keypoints with a right coordinate, 5.99 otherwise); NULL: Fuse(pKF, Scw, ...) (:1180-1200).
extern "C" void orc_fuse_search(const orc_frame_view* f, const orc_projpoints_view* pts,
                                const float* inv_level_sigma2, int* best_idx, int* best_dist) {
    Grid G(f);
    for (int i = 0; i < pts->n; i++) {
        best_idx[i] = -1; best_dist[i] = 256;
        if (!pts->valid[i]) continue;                                                  // :955-996 / :1127-1168
        const float u = pts->u[i], v = pts->v[i];
        const int lvl = pts->level[i];
        const int nc = G.area(f, u, v, pts->radius[i]);                                // :998 / :1170
        if (nc == 0) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx = -1;
        for (int c = 0; c < nc; c++) {
            const int idx = G.cand[c];
            const orc_keypoint& kp = f->kps[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < lvl - 1 || kpLevel > lvl) continue;                          // :1016 / :1187
            if (inv_level_sigma2) {
                if (f->uright && f->uright[idx] >= 0) {                                // :1020-1031
                    const float ex = u - kp.x, ey = v - kp.y, er = pts->ur[i] - f->uright[idx];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_level_sigma2[kpLevel] > 7.8) continue;
                } else {                                                               // :1033-1041
                    const float ex = u - kp.x, ey = v - kp.y;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_level_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = orc_descriptor_distance(dMP, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        best_dist[i] = bestDist;
        if (bestDist <= TH_LOW) best_idx[i] = bestIdx;                                 // :1055 / :1206
    }
}

// One direction of SearchBySim3 (:1294-1365 / :1368-1443): vnMatch[i] = bestIdx where bestDist <= TH_HIGH
static void sim3_direction(const orc_frame_view* f, const orc_projpoints_view* pts, std::vector<int>& vnMatch) {
    Grid G(f);
    vnMatch.assign(std::max(pts->n, 1), -1);
    for (int i = 0; i < pts->n; i++) {
        if (!pts->valid[i]) continue;
        const int lvl = pts->level[i];
        const int nc = G.area(f, pts->u[i], pts->v[i], pts->radius[i]);
        if (nc == 0) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = INT_MAX, bestIdx = -1;
        for (int c = 0; c < nc; c++) {
            const int idx = G.cand[c];
            const int oct = f->kps[idx].octave;
            if (oct < lvl - 1 || oct > lvl) continue;
            const int dist = orc_descriptor_distance(dMP, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;
    }
}

// SearchBySim3 — :1251-1477.  pts1: KF1's map points (one per KF1 keypoint, pts1->n == f1->n) projected
// into KF2; pts2 the other way round; valid[] already folds !pMP, vbAlreadyMatched and isBad (:1296-1300).
// matches12[i1] = idx2 for the pairs that agree in both directions (:1446-1460); returns nFound.
extern "C" int orc_search_by_sim3(const orc_frame_view* f1, const orc_frame_view* f2, const orc_projpoints_view* pts1,
                                  const orc_projpoints_view* pts2, int* matches12) {
    std::vector<int> vnMatch1, vnMatch2;
    sim3_direction(f2, pts1, vnMatch1);
    sim3_direction(f1, pts2, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < pts1->n; i1++) {
        matches12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && idx2 < pts2->n) {
            if (vnMatch2[idx2] == i1) { matches12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

// CheckDistEpipolarLine — :156-175.  F12 row-major 3x3, sigma2 = pKF2->mvLevelSigma2[kp2.octave]
static bool check_dist_epipolar_line(const orc_keypoint& kp1, const orc_keypoint& kp2, const float* F12, float sigma2) {
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * sigma2;
}

static int lower_bound_node(const orc_tri_side* s, uint32_t id) {
    int lo = 0, hi = s->nnodes;
    while (lo < hi) {
        const int mid = (lo + hi) / 2;
        if (s->node_id[mid] < id) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// SearchForTriangulation — :754-928.  (ex, ey): epipole of KF1's centre in KF2 (:762-768); scale_factors2 /
// level_sigma2_2: pKF2->mvScaleFactors / mvLevelSigma2.  vbMatched2 is declared but never set in the
// reference (:774), so every KF1 feature searches the whole node.  matches12[s1->n] out; returns nmatches.
extern "C" int orc_search_for_triangulation(const orc_tri_side* s1, const orc_tri_side* s2, const float* F12, float ex,
                                            float ey, const float* scale_factors2, const float* level_sigma2_2,
                                            int only_stereo, int check_orientation, int* matches12) {
    int nmatches = 0;
    for (int i = 0; i < s1->n; i++) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int it1 = 0, it2 = 0;
    while (it1 < s1->nnodes && it2 < s2->nnodes) {                                     // :791
        if (s1->node_id[it1] == s2->node_id[it2]) {
            for (int i1 = s1->node_start[it1]; i1 < s1->node_start[it1 + 1]; i1++) {
                const int idx1 = (int)s1->indices[i1];
                if (s1->has_mappoint[idx1]) continue;                                  // :799-803
                const bool bStereo1 = s1->uright && s1->uright[idx1] >= 0;             // :805
                if (only_stereo && !bStereo1) continue;
                const orc_keypoint& kp1 = s1->kps[idx1];
                const uint8_t* d1 = s1->desc + 32 * (size_t)idx1;
                int bestDist = TH_LOW, bestIdx2 = -1;                                  // :816-817
                for (int i2 = s2->node_start[it2]; i2 < s2->node_start[it2 + 1]; i2++) {
                    const int idx2 = (int)s2->indices[i2];
                    if (s2->has_mappoint[idx2]) continue;                              // :826-828 (vbMatched2 stays false)
                    const bool bStereo2 = s2->uright && s2->uright[idx2] >= 0;
                    if (only_stereo && !bStereo2) continue;
                    const int dist = orc_descriptor_distance(d1, s2->desc + 32 * (size_t)idx2);
                    if (dist > TH_LOW || dist > bestDist) continue;                    // :840
                    const orc_keypoint& kp2 = s2->kps[idx2];
                    if (!bStereo1 && !bStereo2) {                                      // :845-853
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * scale_factors2[kp2.octave]) continue;
                    }
                    if (check_dist_epipolar_line(kp1, kp2, F12, level_sigma2_2[kp2.octave])) {   // :856-860
                        bestIdx2 = idx2; bestDist = dist;
                    }
                }
                if (bestIdx2 >= 0) {                                                   // :863-880
                    matches12[idx1] = bestIdx2;
                    nmatches++;
                    if (check_orientation) rotHist[rot_bin(kp1.angle, s2->kps[bestIdx2].angle)].push_back(idx1);
                }
            }
            it1++; it2++;
        } else if (s1->node_id[it1] < s2->node_id[it2]) {
            it1 = lower_bound_node(s1, s2->node_id[it2]);                              // :887
        } else {
            it2 = lower_bound_node(s2, s1->node_id[it1]);
        }
    }
    if (check_orientation) {                                                           // :895-913
        int hist[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) hist[i] = (int)rotHist[i].size();
        orc_three_maxima(hist, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { matches12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}
