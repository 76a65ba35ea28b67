// orc_bow.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// Plain-array restatement of the two bag-of-words guided searches:
//   ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)      src/ORBmatcher.cc:185-325
//   ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&)   src/ORBmatcher.cc:610-743
// Both walk the two DBoW2::FeatureVector maps (std::map<NodeId, vector<unsigned>>, ascending node
// ids) in lockstep and, inside every common node, go through side 1's features in vector order,
// looking for the best and second-best Hamming distance among side 2's features of the same node
// that are still free.  They differ in three places, all expressed through the arguments:
//   * who must own a live map point: side 1 only (:214-220; the Frame side has no filter) or both
//     sides (:645-649, :662-670)            -> the `valid` arrays (NULL = every feature counts);
//   * the acceptance threshold: bestDist1 <= TH_LOW (:263) or bestDist1 < TH_LOW (:686)
//                                            -> `th` with <= semantics (50 or 49);
//   * which index the result vector uses: side 2's (:267) or side 1's (:690)
//                                            -> both m12 and m21 are returned.
// DBoW2 itself (vocabulary, transform) is a third-party dependency outside the path: the feature
// vectors arrive as data.  Parity: pinned against the unmodified src/ORBmatcher.cc with real
// DBoW2::FeatureVector objects (oracle/_ref/liborbref_matcher.so, tests/test_oracle_vs_ref_matcher.py::search_by_bow)
// besides the hand-derived known-answer tests (tests/test_oracle_matcher.py).
#include "orb_oracle.h"

#include <cmath>
#include <vector>

namespace {
const int HISTO_LENGTH = 30;   // ORBmatcher.cc:39

// std::map::lower_bound over the ascending node id array
int lower_bound_node(const orc_bow_view* s, uint32_t id) {
    int lo = 0, hi = s->nnodes;
    while (lo < hi) {
        const int mid = (lo + hi) / 2;
        if (s->node_id[mid] < id) lo = mid + 1; else hi = mid;
    }
    return lo;
}
}  // namespace

extern "C" int orc_search_by_bow(const orc_bow_view* s1, const orc_bow_view* s2, int th, float nnratio,
                                 int check_orientation, int* m12, int* m21) {
    for (int i = 0; i < s1->n; i++) m12[i] = -1;
    for (int i = 0; i < s2->n; i++) m21[i] = -1;            // doubles as vpMapPointMatches[realIdxF] / vbMatched2
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int it1 = 0, it2 = 0;
    while (it1 < s1->nnodes && it2 < s2->nnodes) {                                       // :204 / :637
        if (s1->node_id[it1] == s2->node_id[it2]) {
            for (int i1 = s1->node_start[it1]; i1 < s1->node_start[it1 + 1]; i1++) {
                const int idx1 = (int)s1->indices[i1];
                if (s1->valid && !s1->valid[idx1]) continue;                             // :214-220 / :645-649
                const uint8_t* d1 = s1->desc + 32 * (size_t)idx1;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;                     // :224-226
                for (int i2 = s2->node_start[it2]; i2 < s2->node_start[it2 + 1]; i2++) {
                    const int idx2 = (int)s2->indices[i2];
                    if (m21[idx2] >= 0) continue;                                        // :232 / :662
                    if (s2->valid && !s2->valid[idx2]) continue;                         // :662-670
                    const int dist = orc_descriptor_distance(d1, s2->desc + 32 * (size_t)idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= th) {                                                   // :263 / :686
                    if ((float)bestDist1 < nnratio * (float)bestDist2) {                 // :265 / :688
                        m12[idx1] = bestIdx2; m21[bestIdx2] = idx1;
                        if (check_orientation) {                                         // :271-281 / :693-703
                            float rot = s1->angle[idx1] - s2->angle[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(idx1);
                        }
                        nmatches++;
                    }
                }
            }
            it1++; it2++;
        } else if (s1->node_id[it1] < s2->node_id[it2]) {
            it1 = lower_bound_node(s1, s2->node_id[it2]);                                // :289 / :712
        } else {
            it2 = lower_bound_node(s2, s1->node_id[it1]);
        }
    }
    if (check_orientation) {                                                             // :298-318 / :721-740
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) {
                m21[m12[idx1]] = -1; m12[idx1] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}
