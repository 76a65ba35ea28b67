// adapter_guided.cpp — instantiates the map-point dependent templates of include/orbfe_adapter.hpp on the
// stand-ins of adapter_mocks.hpp and runs them on two synthetic frames the way Tracking.cc, LocalMapping.cc and
// LoopClosing.cc call the reference: the four SearchByProjection overloads, both Fuse overloads, SearchBySim3
// and SearchForTriangulation.  Every input the calls saw and every result is written as a blob list
// ([int32 nbytes][bytes] ...) that tests/test_adapter.py replays against the CPU oracle.
// usage: adapter_guided in.bin out.bin    in.bin = int32 w, h, sx, sy then 2*w*h bytes (frame 1 = frame 0 moved by (sx, sy))
#include <cstdio>
#include <cstdlib>
#include <deque>

#include "adapter_mocks.hpp"
#include "orbfe_adapter.hpp"

static FILE* g_out;
static void put(const void* p, size_t nbytes) {
    int n = (int)nbytes;
    fwrite(&n, 4, 1, g_out);
    if (nbytes) fwrite(p, 1, nbytes, g_out);
}
template <class T> static void put(const std::vector<T>& v) { put(v.data(), v.size() * sizeof(T)); }
static void put_int(int v) { put(&v, 4); }

struct Rng {
    unsigned s;
    explicit Rng(unsigned seed) : s(seed) {}
    unsigned next() { s = s * 1664525u + 1013904223u; return s >> 8; }
    int below(int n) { return (int)(next() % (unsigned)n); }
    float jitter(int halfRange64) { return (float)(below(2 * halfRange64 + 1) - halfRange64) / 64.0f; }   // multiples of 1/64
};

static std::vector<int> ids_of(const std::vector<MapPoint*>& v) {
    std::vector<int> o(v.size());
    for (size_t i = 0; i < v.size(); i++) o[i] = v[i] ? v[i]->id : -1;
    return o;
}

static FeatureVector fake_feat_vec(const std::vector<cv::KeyPoint>& keys, const cv::Mat& desc) {
    FeatureVector fv;
    for (size_t i = 0; i < keys.size(); i++) fv[(unsigned)((desc.data[32 * i] & 7) | (keys[i].octave << 3))].push_back((unsigned)i);
    return fv;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    FILE* f = fopen(argv[1], "rb");
    int hdr[4];
    if (!f || fread(hdr, 4, 4, f) != 4) return 2;
    const int w = hdr[0], h = hdr[1];
    const float sx = (float)hdr[2], sy = (float)hdr[3];
    std::vector<unsigned char> buf((size_t)2 * w * h);
    if (fread(buf.data(), 1, buf.size(), f) != buf.size()) return 2;
    fclose(f);
    g_out = fopen(argv[2], "wb");
    if (!g_out) return 2;

    ORB_SLAM2::ORBextractor extractor(1000, 1.2f, 8, 20, 7);
    Frame F[2];
    KeyFrame KF[2];
    Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
    const FMat t1 = FMat::vec3(sx, sy, 0.0f);              // world = camera 0; camera 1 = world + t1
    for (int i = 0; i < 2; i++) {
        cv::Mat im(h, w, CV_8UC1, buf.data() + (size_t)i * w * h, (size_t)w);
        extractor(im, cv::Mat(), F[i].mvKeys, F[i].mDescriptors);
        F[i].mvKeysUn = F[i].mvKeys;
        F[i].N = (int)F[i].mvKeys.size();
        F[i].mvScaleFactors = extractor.GetScaleFactors();
        F[i].mvpMapPoints.assign(F[i].N, nullptr);
        F[i].mvbOutlier.assign(F[i].N, false);
        KeyFrame& K = KF[i];
        K.mvKeysUn = F[i].mvKeys; K.mDescriptors = F[i].mDescriptors;
        K.mFeatVec = fake_feat_vec(F[i].mvKeys, F[i].mDescriptors);
        K.mvScaleFactors = extractor.GetScaleFactors(); K.mvLevelSigma2 = extractor.GetScaleSigmaSquares();
        K.mvInvLevelSigma2 = extractor.GetInverseScaleSigmaSquares();
        K.mnMaxX = w; K.mnMaxY = h;
        K.mfGridElementWidthInv = 64.0f / (float)w; K.mfGridElementHeightInv = 48.0f / (float)h;
        K.mps.assign(F[i].N, nullptr);
        std::vector<float> kp(7 * (size_t)F[i].N);
        for (int k = 0; k < F[i].N; k++) {
            const cv::KeyPoint& p = F[i].mvKeys[k];
            float* o = &kp[7 * (size_t)k];
            o[0] = p.pt.x; o[1] = p.pt.y; o[2] = p.size; o[3] = p.angle; o[4] = p.response;
            std::memcpy(&o[5], &p.octave, 4); std::memcpy(&o[6], &p.class_id, 4);
        }
        put(kp);                                                             // blobs 0/2: keypoints, 1/3: descriptors
        put(F[i].mDescriptors.data, (size_t)F[i].N * 32);
    }
    F[1].mTcw(0, 3) = sx; F[1].mTcw(1, 3) = sy;
    KF[1].t = t1;
    const int N0 = F[0].N, N1 = F[1].N;

    // map points: A[i] seen at keypoint i of frame 0, B[j] at keypoint j of frame 1 (world = camera 0)
    std::deque<MapPoint> pool;
    Rng rng(12345);
    std::vector<MapPoint*> A(N0, nullptr), B(N1, nullptr);
    for (int side = 0; side < 2; side++) {
        const int n = side ? N1 : N0;
        for (int i = 0; i < n; i++) {
            if (side == 0 ? i % 4 == 3 : i % 3 == 0) continue;
            pool.push_back(MapPoint());
            MapPoint& p = pool.back();
            p.id = (int)pool.size() - 1;
            const cv::KeyPoint& k = F[side].mvKeys[i];
            float x = k.pt.x + rng.jitter(160), y = k.pt.y + rng.jitter(160);
            if (side) { x = x - sx; y = y - sy; }
            p.pos = FMat::vec3(x, y, 1.0f);
            p.normal = FMat::vec3(0.0f, 0.0f, i % 13 == 0 ? -1.0f : 1e6f);
            p.level = std::min(7, k.octave + rng.below(2));
            p.nobs = 1 + rng.below(6);
            p.bad = side == 0 ? i % 17 == 0 : i % 10 == 1;
            std::memcpy(p.d, F[side].mDescriptors.data + 32 * (size_t)i, 32);
            const int flips = rng.below(41);
            for (int b = 0; b < flips; b++) { const int bit = rng.below(256); p.d[bit >> 3] ^= (unsigned char)(1 << (bit & 7)); }
            if (side) { p.kf = &KF[1]; p.kfIdx = i; B[i] = &p; } else A[i] = &p;
        }
    }
    {   // blob 4: the pool as rows of 8 floats (pos, normal z, level, nobs, bad), blob 5: descriptors, 6/7: ids of A / B
        std::vector<float> rows(8 * pool.size());
        std::vector<unsigned char> d(32 * pool.size());
        for (size_t i = 0; i < pool.size(); i++) {
            const MapPoint& p = pool[i];
            float* o = &rows[8 * i];
            o[0] = p.pos.a[0]; o[1] = p.pos.a[1]; o[2] = p.pos.a[2]; o[3] = p.normal.a[2]; o[4] = (float)p.level; o[5] = (float)p.nobs;
            o[6] = p.bad ? 1.0f : 0.0f; o[7] = 0.0f;
            std::memcpy(&d[32 * i], p.d, 32);
        }
        put(rows); put(d); put(ids_of(A)); put(ids_of(B));
    }
    KF[0].mps = A; KF[1].mps = B;

    // ---- case A: SearchByProjection(Frame&, vector<MapPoint*>, th) — Tracking.cc:1072 (SearchLocalPoints)
    {
        std::vector<MapPoint*> local;
        for (int i = 0; i < N0; i++) {
            if (!A[i]) continue;
            MapPoint& p = *A[i];
            p.mbTrackInView = i % 5 != 0;
            p.mTrackProjX = p.pos.a[0] + sx; p.mTrackProjY = p.pos.a[1] + sy; p.mTrackProjXR = p.mTrackProjX - 4.0f;
            p.mnTrackScaleLevel = p.level; p.mTrackViewCos = i % 2 ? 0.9995f : 0.9f;
            local.push_back(&p);
        }
        for (int j = 0; j < N1; j++) F[1].mvpMapPoints[j] = j % 6 == 0 ? B[j] : nullptr;
        put(ids_of(local)); put(ids_of(F[1].mvpMapPoints));                  // blobs 8, 9
        ORB_SLAM2::ORBmatcher m(0.8f, true);
        put_int(m.SearchByProjection(F[1], local, 3.0f));                    // 10
        put(ids_of(F[1].mvpMapPoints));                                      // 11
    }
    // ---- case B: SearchByProjection(CurrentFrame, LastFrame, th, bMono) — Tracking.cc:880 (TrackWithMotionModel)
    {
        F[0].mvpMapPoints = A;
        for (int i = 0; i < N0; i++) F[0].mvbOutlier[i] = i % 11 == 0;
        for (int j = 0; j < N1; j++) F[1].mvpMapPoints[j] = j % 6 == 0 ? B[j] : nullptr;
        put(ids_of(F[1].mvpMapPoints));                                      // 12
        ORB_SLAM2::ORBmatcher m(0.9f, true);
        put_int(m.SearchByProjection(F[1], F[0], 7.0f, true));               // 13
        put(ids_of(F[1].mvpMapPoints));                                      // 14
    }
    // ---- case C: SearchByProjection(Frame&, KeyFrame*, set<MapPoint*>, th, ORBdist) — Tracking.cc:1452 (Relocalization)
    {
        std::set<MapPoint*> found;
        for (int i = 0; i < N0; i += 7) if (A[i]) found.insert(A[i]);
        for (int j = 0; j < N1; j++) F[1].mvpMapPoints[j] = j % 8 == 0 ? B[j] : nullptr;
        put(ids_of(F[1].mvpMapPoints));                                      // 15
        ORB_SLAM2::ORBmatcher m(0.9f, true);
        put_int(m.SearchByProjection(F[1], &KF[0], found, 10.0f, 100));      // 16
        put(ids_of(F[1].mvpMapPoints));                                      // 17
    }
    FMat Scw = FMat::eye(4);                                                 // 2 * [I | t1]
    for (int r = 0; r < 3; r++) Scw(r, r) = 2.0f;
    Scw(0, 3) = 2.0f * sx; Scw(1, 3) = 2.0f * sy;
    std::vector<MapPoint*> loopPoints;                                       // LoopClosing's mvpLoopMapPoints: bad ones may be inside
    for (int i = 0; i < N0; i++) if (A[i]) loopPoints.push_back(A[i]);
    put(ids_of(loopPoints));                                                 // 18
    // ---- case D: SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) — LoopClosing.cc:382 (ComputeSim3)
    {
        std::vector<MapPoint*> matched(N1, nullptr);
        for (int j = 0; j < N1; j += 5) matched[j] = B[j] ? B[j] : loopPoints[j % loopPoints.size()];
        put(ids_of(matched));                                                // 19
        ORB_SLAM2::ORBmatcher m(0.75f, true);
        put_int(m.SearchByProjection(&KF[1], Scw, loopPoints, matched, 10)); // 20
        put(ids_of(matched));                                                // 21
    }
    // ---- case G: SearchBySim3 — LoopClosing.cc:349 (ComputeSim3)
    {
        std::vector<MapPoint*> m12(N0, nullptr);
        for (int i = 0; i < N0; i += 9) { const int j = (i * 7) % N1; if (A[i] && B[j]) m12[i] = B[j]; }
        put(ids_of(m12));                                                    // 22
        const float s12 = 1.0f;
        const FMat R12 = FMat::eye(3), t12 = FMat::vec3(-sx, -sy, 0.0f);    // camera 2 -> camera 1
        ORB_SLAM2::ORBmatcher m(0.75f, true);
        put_int(m.SearchBySim3(&KF[0], &KF[1], m12, s12, R12, t12, 7.5f));   // 23
        put(ids_of(m12));                                                    // 24
    }
    // ---- case H: SearchForTriangulation — LocalMapping.cc:281 (CreateNewMapPoints)
    {
        KeyFrame K2 = KF[1];
        K2.t = FMat::vec3(300.0f, 200.0f, 1.0f);                             // epipole of camera 0 in K2: (300, 200)
        FMat F12(3, 3);                                                      // [t]x for the image-plane shift (sx, sy)
        F12(0, 2) = sy; F12(1, 2) = -sx; F12(2, 0) = -sy; F12(2, 1) = sx;
        std::vector<std::pair<size_t, size_t> > pairs;
        ORB_SLAM2::ORBmatcher m(0.6f, false);
        put_int(m.SearchForTriangulation(&KF[0], &K2, F12, pairs, false));   // 25
        std::vector<int> flat;
        for (size_t i = 0; i < pairs.size(); i++) { flat.push_back((int)pairs[i].first); flat.push_back((int)pairs[i].second); }
        put(flat);                                                           // 26
    }
    // ---- case F: Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint) — LoopClosing.cc:649 (SearchAndFuse)
    {
        std::vector<MapPoint*> replace(loopPoints.size(), nullptr);
        ORB_SLAM2::ORBmatcher m(0.8f);
        put_int(m.Fuse(&KF[1], Scw, loopPoints, 4.0f, replace));             // 27
        put(ids_of(replace)); put(ids_of(KF[1].mps));                        // 28, 29
        for (int i = 0; i < N0; i++) if (A[i] && A[i]->kf) { A[i]->kf = nullptr; A[i]->kfIdx = -1; A[i]->nobs--; }   // undo
        KF[1].mps = B;
    }
    // ---- case E: Fuse(KeyFrame*, vector<MapPoint*>, th) — LocalMapping.cc:599 (SearchInNeighbors)
    {
        std::vector<MapPoint*> cand;
        for (int i = 0; i < N0; i++) {
            cand.push_back(A[i]);                                            // null entries stay in (:949)
            if (i % 23 == 0 && A[i]) cand.push_back(A[i]);                   // a duplicate: skipped at its second turn
        }
        put(ids_of(cand));                                                   // 30
        ORB_SLAM2::ORBmatcher m;
        put_int(m.Fuse(&KF[1], cand, 3.0f));                                 // 31
        put(ids_of(KF[1].mps));                                              // 32
        std::vector<int> st(3 * pool.size());
        for (size_t i = 0; i < pool.size(); i++) { st[3 * i] = pool[i].bad; st[3 * i + 1] = pool[i].nobs; st[3 * i + 2] = pool[i].kf ? pool[i].kfIdx : -1; }
        put(st);                                                             // 33
    }
    fclose(g_out);
    printf("adapter_guided: %d + %d keypoints, %zu map points\n", N0, N1, pool.size());
    return 0;
}
