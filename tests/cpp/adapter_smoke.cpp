// adapter_smoke.cpp — drives include/orbfe_adapter.hpp the way Frame.cc / Tracking.cc drive the
// reference classes: ORB_SLAM2::ORBextractor::operator()(image, mask, keypoints, descriptors) on
// two frames, then ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, 100)
// on a minimal Frame stand-in with the reference's field names.  cv:: types come from the test
// shim (oracle/cvshim) because OpenCV headers are absent here.
// usage: adapter_smoke in.bin out.bin     in.bin = int32 w, h, then 2*w*h bytes (two frames)
#include <cstdio>
#include <cstdlib>
#include <map>
#include <vector>

#include "cvshim.hpp"
#include "orbfe_adapter.hpp"

struct MapPoint {
    bool bad;
    bool isBad() const { return bad; }
};
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;   // DBoW2::FeatureVector

struct Frame {   // the members of ORB_SLAM2::Frame the matcher reads (include/Frame.h)
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    cv::Mat mDescriptors;
    std::vector<float> mvScaleFactors;
    cv::Mat mK, mDistCoef;
    int N;
    FeatureVector mFeatVec;
    std::vector<std::size_t> mGrid[64][48];
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
};
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;

struct KeyFrame {   // the members SearchByBoW reads (include/KeyFrame.h)
    std::vector<cv::KeyPoint> mvKeysUn;
    cv::Mat mDescriptors;
    FeatureVector mFeatVec;
    std::vector<MapPoint*> mps;
    std::vector<MapPoint*> GetMapPointMatches() const { return mps; }
};

// stand-in for the vocabulary: node id from the octave and the low four bits of the first descriptor byte
static FeatureVector fake_feat_vec(const std::vector<cv::KeyPoint>& keys, const cv::Mat& desc) {
    FeatureVector fv;
    for (size_t i = 0; i < keys.size(); i++)
        fv[(unsigned)((desc.data[32 * i] & 15) | (keys[i].octave << 4))].push_back((unsigned)i);
    return fv;
}

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    FILE* f = fopen(argv[1], "rb");
    int wh[2];
    if (!f || fread(wh, 4, 2, f) != 2) return 2;
    const int w = wh[0], h = wh[1];
    std::vector<unsigned char> buf((size_t)2 * w * h);
    if (fread(buf.data(), 1, buf.size(), f) != buf.size()) return 2;
    fclose(f);
    ORB_SLAM2::ORBextractor extractor(2000, 1.2f, 8, 20, 7);       // mpIniORBextractor, Tracking.cc:128
    Frame F[2];
    Frame::mnMinX = 0; Frame::mnMaxX = (float)w; Frame::mnMinY = 0; Frame::mnMaxY = (float)h;
    for (int i = 0; i < 2; i++) {
        cv::Mat im(h, w, CV_8UC1, buf.data() + (size_t)i * w * h, (size_t)w);
        extractor(im, cv::Mat(), F[i].mvKeys, F[i].mDescriptors);   // Frame::ExtractORB, Frame.cc:275
        F[i].mvKeysUn = F[i].mvKeys;
        F[i].mvScaleFactors = extractor.GetScaleFactors();
    }
    if ((int)extractor.mvImagePyramid.size() != 8 || extractor.mvImagePyramid[7].empty()) return 3;
    std::vector<cv::Point2f> prev(F[0].mvKeysUn.size());
    for (size_t i = 0; i < prev.size(); i++) prev[i] = F[0].mvKeysUn[i].pt;   // Tracking.cc:609-616
    std::vector<int> m12;
    ORB_SLAM2::ORBmatcher matcher(0.9f, true);                      // Tracking.cc:641
    const int nm = matcher.SearchForInitialization(F[0], F[1], prev, m12, 100);
    FILE* o = fopen(argv[2], "wb");
    for (int i = 0; i < 2; i++) {
        int n = (int)F[i].mvKeys.size();
        fwrite(&n, 4, 1, o);
        for (int k = 0; k < n; k++) {
            const cv::KeyPoint& kp = F[i].mvKeys[k];
            float v[5] = {kp.pt.x, kp.pt.y, kp.size, kp.angle, kp.response};
            int q[2] = {kp.octave, kp.class_id};
            fwrite(v, 4, 5, o); fwrite(q, 4, 2, o);
        }
        fwrite(F[i].mDescriptors.data, 1, (size_t)n * 32, o);
    }
    int n1 = (int)m12.size();
    fwrite(&nm, 4, 1, o); fwrite(&n1, 4, 1, o); fwrite(m12.data(), 4, n1, o);
    // Frame constructor epilogue with the TUM1 camera (Examples/Monocular/TUM1.yaml): bounds, mvKeysUn, mGrid
    const float Kv[4] = {517.306408f, 516.469215f, 318.643040f, 255.313989f};
    const float Dv[5] = {0.262383f, -0.953104f, -0.005358f, 0.002628f, 1.163314f};
    float Kbuf[9] = {Kv[0], 0.f, Kv[2], 0.f, Kv[1], Kv[3], 0.f, 0.f, 1.f}, Dbuf[5];
    for (int r = 0; r < 5; r++) Dbuf[r] = Dv[r];
    F[0].mK = cv::Mat(3, 3, CV_32F, Kbuf, 3 * sizeof(float));          // the shim only allocates 8-bit matrices
    F[0].mDistCoef = cv::Mat(5, 1, CV_32F, Dbuf, sizeof(float));
    cv::Mat im0(h, w, CV_8UC1, buf.data(), (size_t)w);
    matcher.ComputeImageBounds(F[0], im0);
    matcher.UndistortAndAssign(F[0]);
    const float bnd[4] = {Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY};
    fwrite(bnd, 4, 4, o);
    for (size_t k = 0; k < F[0].mvKeysUn.size(); k++) { float v[2] = {F[0].mvKeysUn[k].pt.x, F[0].mvKeysUn[k].pt.y}; fwrite(v, 4, 2, o); }
    for (int ix = 0; ix < 64; ix++)
        for (int iy = 0; iy < 48; iy++) {
            int cnt = (int)F[0].mGrid[ix][iy].size();
            fwrite(&cnt, 4, 1, o);
            for (int j = 0; j < cnt; j++) { int v = (int)F[0].mGrid[ix][iy][j]; fwrite(&v, 4, 1, o); }
        }
    // SearchByBoW(KeyFrame*, Frame&, ..) and SearchByBoW(KeyFrame*, KeyFrame*, ..): frame 0 as key frame (every
    // third feature without a map point, every seventh with a bad one) against frame 1
    MapPoint good = {false}, bad = {true};
    KeyFrame KF[2];
    for (int i = 0; i < 2; i++) {
        KF[i].mvKeysUn = F[i].mvKeys; KF[i].mDescriptors = F[i].mDescriptors;
        KF[i].mFeatVec = fake_feat_vec(F[i].mvKeys, F[i].mDescriptors);
        KF[i].mps.resize(F[i].mvKeys.size());
        for (size_t k = 0; k < KF[i].mps.size(); k++) KF[i].mps[k] = k % 3 == 0 ? nullptr : (k % 7 == 0 ? &bad : &good);
        F[i].N = (int)F[i].mvKeys.size(); F[i].mFeatVec = KF[i].mFeatVec;
    }
    ORB_SLAM2::ORBmatcher bow(0.75f, true);                          // Tracking.cc:841 / :1484
    std::vector<MapPoint*> vpMatches, vpMatches12;
    const int nb1 = bow.SearchByBoW(&KF[0], F[1], vpMatches);
    const int nb2 = bow.SearchByBoW(&KF[0], &KF[1], vpMatches12);
    int cnt1 = 0, cnt2 = 0;
    for (size_t k = 0; k < vpMatches.size(); k++) cnt1 += vpMatches[k] != nullptr;
    for (size_t k = 0; k < vpMatches12.size(); k++) cnt2 += vpMatches12[k] != nullptr;
    const int bowOut[6] = {nb1, cnt1, (int)vpMatches.size(), nb2, cnt2, (int)vpMatches12.size()};
    fwrite(bowOut, 4, 6, o);
    fclose(o);
    printf("adapter_smoke: %zu + %zu keypoints, %d init matches, DescriptorDistance(0,1)=%d\n", F[0].mvKeys.size(),
           F[1].mvKeys.size(), nm,
           ORB_SLAM2::ORBmatcher::DescriptorDistance(F[0].mDescriptors.rowRange(0, 1), F[0].mDescriptors.rowRange(1, 2)));
    return 0;
}
