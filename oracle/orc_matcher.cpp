// orc_matcher.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// Plain-array restatements of the reference matcher routines, line by line:
//   ORBmatcher::DescriptorDistance        src/ORBmatcher.cc:1828-1844
//   ORBmatcher::ComputeThreeMaxima        src/ORBmatcher.cc:1777-1821
//   ORBmatcher::SearchForInitialization   src/ORBmatcher.cc:469-600
//   ORBmatcher::SearchByProjection(F,MPs) src/ORBmatcher.cc:51-138, RadiusByViewingCos :146-152
//   Frame::AssignFeaturesToGrid/PosInGrid src/Frame.cc:246-267, :419-430
//   Frame::GetFeaturesInArea              src/Frame.cc:358-415
//   Frame::ComputeStereoMatches           src/Frame.cc:512-686
//   brute-force kNN-2 + ratio             inner loops of SearchByBoW, src/ORBmatcher.cc:233-269
// Parity: every routine here is pinned against the unmodified reference TUs — src/ORBmatcher.cc,
// Frame.cc, KeyFrame.cc compiled in place against oracle/cvshim (oracle/_ref/liborbref_matcher.so,
// tests/test_oracle_vs_ref_matcher.py, live and against tests/golden/ref_matcher.npz) — besides the
// hand-derived known-answer tests (tests/test_oracle_matcher.py).
#include "orb_oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

namespace {
const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // ORBmatcher.cc:37-39
const int GRID_COLS = 64, GRID_ROWS = 48;                  // Frame.h:38-39
}

// the reference's bit-hack popcount, kept verbatim in spirit (8 x 32-bit words)
extern "C" int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t wa, wb;
        std::memcpy(&wa, a + 4 * i, 4); std::memcpy(&wb, b + 4 * i, 4);
        uint32_t v = wa ^ wb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

static inline float grid_inv_w(const orc_frame_view* f) { return (float)GRID_COLS / (f->max_x - f->min_x); }
static inline float grid_inv_h(const orc_frame_view* f) { return (float)GRID_ROWS / (f->max_y - f->min_y); }

// AssignFeaturesToGrid + PosInGrid (round(), Frame.cc:422-423).  CSR index: cell = ix*48+iy,
// items keep insertion (keypoint index) order.
extern "C" void orc_assign_grid(const orc_frame_view* f, int* cell_start, int* cell_items) {
    const float iw = grid_inv_w(f), ih = grid_inv_h(f);
    std::vector<int> cell(f->n, -1);
    std::memset(cell_start, 0, sizeof(int) * (GRID_COLS * GRID_ROWS + 1));
    for (int i = 0; i < f->n; i++) {
        int px = (int)std::round((f->kps[i].x - f->min_x) * iw);
        int py = (int)std::round((f->kps[i].y - f->min_y) * ih);
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
        cell[i] = px * GRID_ROWS + py;
        cell_start[cell[i] + 1]++;
    }
    for (int c = 0; c < GRID_COLS * GRID_ROWS; c++) cell_start[c + 1] += cell_start[c];
    std::vector<int> fill(cell_start, cell_start + GRID_COLS * GRID_ROWS);
    for (int i = 0; i < f->n; i++) if (cell[i] >= 0) cell_items[fill[cell[i]]++] = i;
}

// GetFeaturesInArea — Frame.cc:358-415
extern "C" int orc_features_in_area(const orc_frame_view* f, const int* cell_start,
                                    const int* cell_items, float x, float y, float r,
                                    int minLevel, int maxLevel, int* out, int cap) {
    const float iw = grid_inv_w(f), ih = grid_inv_h(f);
    // KeyFrame holds its bounds as int (KeyFrame.h:191-194); Frame as float
    const float qmin_x = f->keyframe_bounds ? (float)(int)f->min_x : f->min_x;
    const float qmin_y = f->keyframe_bounds ? (float)(int)f->min_y : f->min_y;
    int n = 0;
    const int nMinCellX = std::max(0, (int)std::floor((x - qmin_x - r) * iw));
    if (nMinCellX >= GRID_COLS) return 0;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - qmin_x + r) * iw));
    if (nMaxCellX < 0) return 0;
    const int nMinCellY = std::max(0, (int)std::floor((y - qmin_y - r) * ih));
    if (nMinCellY >= GRID_ROWS) return 0;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - qmin_y + r) * ih));
    if (nMaxCellY < 0) return 0;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const int c = ix * GRID_ROWS + iy;
            for (int j = cell_start[c]; j < cell_start[c + 1]; j++) {
                const orc_keypoint& kp = f->kps[cell_items[j]];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float distx = kp.x - x, disty = kp.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) {
                    if (n < cap) out[n] = cell_items[j];
                    n++;
                }
            }
        }
    return n;
}

// ComputeThreeMaxima — ORBmatcher.cc:1777-1821 (on bin counts)
extern "C" void orc_three_maxima(const int* counts, int L, int* ind1, int* ind2, int* ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = counts[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

// SearchForInitialization — ORBmatcher.cc:469-600
extern "C" int orc_search_for_initialization(const orc_frame_view* f1, const orc_frame_view* f2,
                                             float* prev, int* m12, int window, float nnratio,
                                             int check_ori) {
    int nmatches = 0;
    for (int i = 0; i < f1->n; i++) m12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> matchedDist(f2->n, INT_MAX), m21(f2->n, -1);
    std::vector<int> cs(GRID_COLS * GRID_ROWS + 1), ci(std::max(f2->n, 1)), cand(std::max(f2->n, 1));
    orc_assign_grid(f2, cs.data(), ci.data());
    for (int i1 = 0; i1 < f1->n; i1++) {
        const orc_keypoint& kp1 = f1->kps[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;                                                  // :487-489
        int nc = orc_features_in_area(f2, cs.data(), ci.data(), prev[2 * i1], prev[2 * i1 + 1],
                                      (float)window, level1, level1, cand.data(), f2->n);
        if (nc == 0) continue;
        const uint8_t* d1 = f1->desc + 32 * (size_t)i1;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int c = 0; c < nc; c++) {                                             // :506-529
            const int i2 = cand[c];
            const int dist = orc_descriptor_distance(d1, f2->desc + 32 * (size_t)i2);
            if (matchedDist[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {                                                  // :532
            if (bestDist < (float)bestDist2 * nnratio) {                           // :534
                if (m21[bestIdx2] >= 0) { m12[m21[bestIdx2]] = -1; nmatches--; }
                m12[i1] = bestIdx2; m21[bestIdx2] = i1; matchedDist[bestIdx2] = bestDist;
                nmatches++;
                if (check_ori) {                                                   // :549-559
                    float rot = f1->kps[i1].angle - f2->kps[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (check_ori) {                                                               // :566-590
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (m12[idx1] >= 0) { m12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < f1->n; i1++)                                             // :595-597
        if (m12[i1] >= 0) { prev[2 * i1] = f2->kps[m12[i1]].x; prev[2 * i1 + 1] = f2->kps[m12[i1]].y; }
    return nmatches;
}

// SearchByProjection(Frame&, vector<MapPoint*>&, th) — ORBmatcher.cc:51-138
extern "C" int orc_search_by_projection(const orc_frame_view* f, const orc_mappoints_view* mps,
                                        float th, float nnratio, uint8_t* occupied,
                                        int* mp_for_kp) {
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> cs(GRID_COLS * GRID_ROWS + 1), ci(std::max(f->n, 1)), cand(std::max(f->n, 1));
    orc_assign_grid(f, cs.data(), ci.data());
    for (int iMP = 0; iMP < mps->n; iMP++) {
        if (!mps->track_in_view[iMP]) continue;
        const int lvl = mps->level[iMP];
        float r = mps->view_cos[iMP] > 0.998 ? 2.5f : 4.0f;                        // :146-152
        if (bFactor) r *= th;
        const float rs = r * f->scale_factors[lvl];
        int nc = orc_features_in_area(f, cs.data(), ci.data(), mps->proj_x[iMP], mps->proj_y[iMP],
                                      rs, lvl - 1, lvl, cand.data(), f->n);
        if (nc == 0) continue;
        const uint8_t* dmp = mps->desc + 32 * (size_t)iMP;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int c = 0; c < nc; c++) {
            const int idx = cand[c];
            if (occupied[idx]) continue;                                           // :95-98
            if (f->uright && f->uright[idx] > 0) {                                 // :100-105
                const float er = std::fabs(mps->proj_xr[iMP] - f->uright[idx]);
                if (er > rs) continue;
            }
            const int dist = orc_descriptor_distance(dmp, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel;
                bestLevel = f->kps[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) { bestLevel2 = f->kps[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {                                                 // :127-135
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            mp_for_kp[bestIdx] = iMP;
            if (mps->observations[iMP] > 0) occupied[bestIdx] = 1;
            nmatches++;
        }
    }
    return nmatches;
}

// SearchByProjection(CurrentFrame, LastFrame, th, bMono) — ORBmatcher.cc:1484-1630 (the loop body
// from :1540 on; projection and the forward / backward decision are inputs)
extern "C" int orc_search_by_projection_frame(const orc_frame_view* f, const orc_framepoints_view* pts,
                                              int check_orientation, const uint8_t* kp_occupied,
                                              int* match_for_kp) {
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];                                          // :1489-1492
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> cs(GRID_COLS * GRID_ROWS + 1), ci(std::max(f->n, 1)), cand(std::max(f->n, 1));
    orc_assign_grid(f, cs.data(), ci.data());
    // CurrentFrame.mvpMapPoints as far as the routine can tell: 0 = NULL or no observations,
    // 1 = holds a point with observations
    std::vector<uint8_t> occ(kp_occupied, kp_occupied + f->n);
    for (int i = 0; i < pts->n; i++) {
        if (!pts->valid[i]) continue;                                                // :1510-1538
        const float u = pts->u[i], v = pts->v[i], radius = pts->radius[i];
        const int nc = orc_features_in_area(f, cs.data(), ci.data(), u, v, radius, pts->min_level[i],
                                            pts->max_level[i], cand.data(), f->n);      // :1547-1552
        if (nc == 0) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int c = 0; c < nc; c++) {
            const int i2 = cand[c];
            if (occ[i2]) continue;                                                   // :1567-1569
            if (f->uright && f->uright[i2] > 0) {                                    // :1571-1577
                const float er = std::fabs(pts->ur[i] - f->uright[i2]);
                if (er > radius) continue;
            }
            const int dist = orc_descriptor_distance(dMP, f->desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }                 // :1583-1587
        }
        if (bestDist <= TH_HIGH) {                                                   // :1590-1608
            match_for_kp[bestIdx2] = i;
            occ[bestIdx2] = pts->observations[i] > 0;
            nmatches++;
            if (check_orientation) {
                float rot = pts->angle[i] - f->kps[bestIdx2].angle;
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (check_orientation) {                                                         // :1611-1628
        int hist[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) hist[i] = (int)rotHist[i].size();
        orc_three_maxima(hist, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (size_t j = 0; j < rotHist[i].size(); j++) {
                    match_for_kp[rotHist[i][j]] = -2;
                    nmatches--;
                }
    }
    return nmatches;
}

// ComputeStereoMatches — Frame.cc:512-686
extern "C" void orc_compute_stereo_matches(const orc_frame_view* L, const orc_frame_view* R,
                                           const orc_pyramid_view* pl, const orc_pyramid_view* pr,
                                           const float* inv_scale, float mbf, float mb,
                                           float* uright, float* depth) {
    const int N = L->n;
    for (int i = 0; i < N; i++) { uright[i] = -1.0f; depth[i] = -1.0f; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = pl->h[0];
    std::vector<std::vector<int>> rowIdx(nRows);
    for (int iR = 0; iR < R->n; iR++) {                                            // :529-539
        const float kpY = R->kps[iR].y;
        const float r = 2.0f * L->scale_factors[R->kps[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);   // reference indexes unchecked
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> distIdx;
    for (int iL = 0; iL < N; iL++) {
        const orc_keypoint& kpL = L->kps[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<int>& cands = rowIdx[(int)vL];
        if (cands.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        int bestIdxR = 0;
        const uint8_t* dL = L->desc + 32 * (size_t)iL;
        for (int iR : cands) {                                                     // :575-596
            const orc_keypoint& kpR = R->kps[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = orc_descriptor_distance(dL, R->desc + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {                                                // :598
            const float uR0 = R->kps[bestIdxR].x;
            const float sf = inv_scale[kpL.octave];
            const float scaleduL = std::round(kpL.x * sf), scaledvL = std::round(kpL.y * sf);
            const float scaleduR0 = std::round(uR0 * sf);
            const int w = 5, Lh = 5;
            const int lvl = kpL.octave;
            const uint8_t* IL = pl->data[lvl]; const ptrdiff_t sl = pl->stride[lvl];
            const uint8_t* IR = pr->data[lvl]; const ptrdiff_t sr = pr->stride[lvl];
            const int v0 = (int)(scaledvL - w), uL0 = (int)(scaleduL - w);
            const float cL = (float)IL[(v0 + w) * sl + uL0 + w];
            int sadBest = INT_MAX, bestinc = 0;                                    // int! :613
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + Lh - w, endu = scaleduR0 + Lh + w + 1;
            if (iniu < 0 || endu >= pr->w[lvl]) continue;                          // :621
            for (int inc = -Lh; inc <= Lh; inc++) {
                const int uR0i = (int)(scaleduR0 + inc - w);
                const float cR = (float)IR[(v0 + w) * sr + uR0i + w];
                float dist = 0;                                                    // L1 norm (:629)
                for (int yy = 0; yy < 2 * w + 1; yy++)
                    for (int xx = 0; xx < 2 * w + 1; xx++)
                        dist += std::fabs(((float)IL[(v0 + yy) * sl + uL0 + xx] - cL) -
                                          ((float)IR[(v0 + yy) * sr + uR0i + xx] - cR));
                if (dist < sadBest) { sadBest = (int)dist; bestinc = inc; }        // :630-634
                vDists[Lh + inc] = dist;
            }
            if (bestinc == -Lh || bestinc == Lh) continue;
            const float dist1 = vDists[Lh + bestinc - 1], dist2 = vDists[Lh + bestinc],
                        dist3 = vDists[Lh + bestinc + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = L->scale_factors[kpL.octave] * ((float)scaleduR0 + (float)bestinc + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01f; bestuR = uL - 0.01f; }
                depth[iL] = mbf / disparity;
                uright[iL] = bestuR;
                distIdx.push_back({sadBest, iL});
            }
        }
    }
    if (distIdx.empty()) return;   // reference would index an empty vector (UB); defined here
    std::sort(distIdx.begin(), distIdx.end());                                     // :671
    const float median = (float)distIdx[distIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)distIdx.size() - 1; i >= 0; i--) {
        if (distIdx[i].first < thDist) break;
        uright[distIdx[i].second] = -1; depth[distIdx[i].second] = -1;
    }
}

// brute-force kNN-2 with TH_LOW + ratio, as in SearchByBoW's inner loop (ORBmatcher.cc:233-269):
// strict '<' updates => first (lowest index) best wins ties; accept d1<=th && d1 < ratio*d2.
extern "C" void orc_knn2(const uint8_t* q, int m, const uint8_t* t, int n, int th, float ratio,
                         int* out_idx, int* out_d1, int* out_d2) {
    for (int i = 0; i < m; i++) {
        int b1 = 256, b2 = 256, bi = -1;
        const uint64_t* a = (const uint64_t*)(q + 32 * (size_t)i);
        for (int j = 0; j < n; j++) {
            const uint64_t* b = (const uint64_t*)(t + 32 * (size_t)j);
            int d = __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
                    __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
            if (d < b1) { b2 = b1; b1 = d; bi = j; }
            else if (d < b2) b2 = d;
        }
        out_d1[i] = b1; out_d2[i] = b2;
        out_idx[i] = (b1 <= th && (float)b1 < ratio * (float)b2) ? bi : -1;
    }
}
