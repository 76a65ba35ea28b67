// orc_mappoint.cpp — ORACLE (test infrastructure only; see orb_oracle.h).
//
// MapPoint::ComputeDistinctiveDescriptors (reference src/MapPoint.cc:266-334), the N x N Hamming
// part: among the descriptors that observe a map point pick the one whose median distance to
// the set is least.  Restated on plain arrays for a batch of points; the gathering of the rows
// from the observing key frames (:283-289) is the caller's.
#include "orb_oracle.h"

#include <algorithm>
#include <climits>
#include <vector>

extern "C" void orc_distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* best) {
    for (int p = 0; p < npoints; p++) {
        const int N = start[p + 1] - start[p];
        if (N <= 0) { best[p] = -1; continue; }                       // :291-292: nothing to choose from
        const uint8_t* d = desc + 32 * (size_t)start[p];
        std::vector<float> D((size_t)N * N);                          // float Distances[N][N] (:296)
        for (int i = 0; i < N; i++) {
            D[(size_t)i * N + i] = 0;
            for (int j = i + 1; j < N; j++) {
                const int dij = orc_descriptor_distance(d + 32 * (size_t)i, d + 32 * (size_t)j);
                D[(size_t)i * N + j] = (float)dij; D[(size_t)j * N + i] = (float)dij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;                        // :309-310
        for (int i = 0; i < N; i++) {
            std::vector<int> v(D.begin() + (size_t)i * N, D.begin() + (size_t)(i + 1) * N);
            std::sort(v.begin(), v.end());
            const int median = v[(size_t)(0.5 * (N - 1))];            // :315
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best[p] = BestIdx;
    }
}
