/* orb_oracle.h — CPU ORACLE for the ORB front end.  TEST INFRASTRUCTURE ONLY.
 *
 * A plain C/C++ restatement of the reference algorithm (saber/ORB_SLAM2_Annotate:
 * src/ORBextractor.cc, src/ORBmatcher.cc, src/Frame.cc) and of the OpenCV primitives it
 * calls (semantics pinned to cv2 4.13.0, see tests/test_oracle_primitives.py and
 * tests/golden/).  Nothing under orb_slam2_annotate_b200/ may include, link or call this
 * library: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs use it, and only as the checker / the timed CPU baseline.
 *
 * Parity pinning: the reference ships no tests or golden vectors.  The
 * oracle is pinned against (a) cv2 4.13.0 run live in the build container plus committed
 * fixtures made from it (tests/golden/make_golden.py) for every OpenCV primitive, and
 * (b) the UNMODIFIED reference translation unit src/ORBextractor.cc compiled against
 * oracle/cvshim (oracle/_ref/liborbref.so, built by oracle/Makefile) for the extractor.
 * The matcher restatements follow the reference line by line and are pinned the same way: the
 * UNMODIFIED src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc and the vendored DBoW2
 * sources compile in place against oracle/cvshim (plus a stand-in for include/Converter.h, the
 * one header that needs Eigen / g2o) into oracle/_ref/liborbref_matcher.so; ref_matcher_wrap.cpp
 * builds the reference's own Frame / KeyFrame / MapPoint objects from the same flat arrays and
 * tests/test_oracle_vs_ref_matcher.py compares every routine, live and against the committed
 * fixture tests/golden/ref_matcher.npz.  Hand-derived known-answer tests stay beside that.
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* == cv::KeyPoint field-for-field (28 B) == */
typedef struct {
    float x, y, size, angle, response;
    int octave, class_id;
} orc_keypoint;

/* ---- OpenCV primitive restatements (cv2 4.13.0 semantics) ---- */
int   orc_cv_round_f(float v);                       /* cvRound(float): round-half-even */
int   orc_cv_round_d(double v);
float orc_fast_atan2(float y, float x);              /* cv::fastAtan2, degrees [0,360) */
void  orc_resize_linear_u8(const uint8_t* src, int sw, int sh, ptrdiff_t sstride,
                           uint8_t* dst, int dw, int dh, ptrdiff_t dstride);
void  orc_copy_make_border_reflect101(const uint8_t* src, int w, int h, ptrdiff_t sstride,
                                      uint8_t* dst, ptrdiff_t dstride, int border);
/* FAST-9/16 (cv::FAST(img, kps, threshold, nonmax)), row-major output (x,y,score). */
int   orc_fast9_16(const uint8_t* img, int w, int h, ptrdiff_t stride, int threshold,
                   int nonmax, int* out_xys, int cap);
void  orc_gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, ptrdiff_t sstride,
                               uint8_t* dst, ptrdiff_t dstride);

/* ---- extractor (ORBextractor.cc) ---- */
typedef struct orc_extractor orc_extractor;
orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels,
                                    int iniThFAST, int minThFAST);
void  orc_extractor_destroy(orc_extractor* e);
/* tables: out arrays sized nlevels */
void  orc_extractor_tables(const orc_extractor* e, float* scale, float* inv_scale,
                           float* sigma2, float* inv_sigma2, int* features_per_level,
                           int* umax16);
/* operator(): returns number of keypoints (<= cap) or -1 if cap too small. */
int   orc_extract(orc_extractor* e, const uint8_t* img, int w, int h, ptrdiff_t stride,
                  orc_keypoint* kps, uint8_t* desc, int cap);
/* taps into the last call, for stage-by-stage parity tests */
int   orc_level_size(const orc_extractor* e, int level, int* w, int* h);
const uint8_t* orc_level_image(const orc_extractor* e, int level, ptrdiff_t* stride);
const uint8_t* orc_level_blurred(const orc_extractor* e, int level, ptrdiff_t* stride);
/* raw per-level FAST list (x,y,score triplets, coordinates relative to minBorder) in
 * vToDistributeKeys order; returns count */
int   orc_level_raw(const orc_extractor* e, int level, const int** xys);
/* quadtree alone (DistributeOctTree): in = (x,y,score) triplets relative to minBorder,
 * out = indices into the input in result order. returns count. */
int   orc_distribute_octtree(const int* xys, int n, int minX, int maxX, int minY, int maxY,
                             int N, int* out_idx, int cap);

/* ---- matcher (ORBmatcher.cc / Frame.cc) ---- */
int   orc_descriptor_distance(const uint8_t* a, const uint8_t* b);

/* Frame view: what the matcher routines read from a Frame. */
typedef struct {
    int n;
    const orc_keypoint* kps;      /* mvKeysUn (== mvKeys when undistorted input)        */
    const uint8_t* desc;          /* n x 32                                              */
    const float* uright;          /* mvuRight or NULL (treated as all -1)                */
    float min_x, max_x, min_y, max_y;   /* mnMinX.. image bounds                         */
    const float* scale_factors;   /* mvScaleFactors                                      */
    int nlevels;
    int keyframe_bounds;          /* 1: KeyFrame — GetFeaturesInArea subtracts the int-truncated
                                     mnMinX / mnMinY (KeyFrame.h:191-194, KeyFrame.cc:616-628) */
} orc_frame_view;

/* 64x48 grid as in Frame::AssignFeaturesToGrid; CSR arrays sized 64*48+1 / n */
void  orc_assign_grid(const orc_frame_view* f, int* cell_start /*3073*/, int* cell_items /*n*/);
int   orc_features_in_area(const orc_frame_view* f, const int* cell_start,
                           const int* cell_items, float x, float y, float r, int minLevel,
                           int maxLevel, int* out, int cap);
void  orc_three_maxima(const int* hist_counts, int L, int* ind1, int* ind2, int* ind3);
/* SearchForInitialization: prev_matched in/out (2*n1 floats), matches12 out (n1). */
int   orc_search_for_initialization(const orc_frame_view* f1, const orc_frame_view* f2,
                                    float* prev_matched, int* matches12, int window,
                                    float nnratio, int check_ori);
/* SearchByProjection(Frame, MapPoints, th). map point arrays of length nmp.
 * kp_occupied[n]: 1 if F.mvpMapPoints[idx] exists with Observations()>0 (in/out: set when
 * a map point with observations>0 is assigned). out_mp_for_kp[n]: map point index assigned
 * to keypoint or -1 (caller initialises to -1). */
typedef struct {
    int n;
    const uint8_t* track_in_view;  /* mbTrackInView && !isBad                            */
    const float* proj_x; const float* proj_y; const float* proj_xr;
    const int* level;              /* mnTrackScaleLevel                                  */
    const float* view_cos;         /* mTrackViewCos                                      */
    const uint8_t* desc;           /* n x 32                                             */
    const int* observations;       /* Observations()                                     */
} orc_mappoints_view;
int   orc_search_by_projection(const orc_frame_view* f, const orc_mappoints_view* mps,
                               float th, float nnratio, uint8_t* kp_occupied,
                               int* out_mp_for_kp);
/* SearchByProjection(CurrentFrame, LastFrame, th, bMono) — ORBmatcher.cc:1484-1630.  The pose
 * arithmetic (:1494-1553) is the caller's: per last-frame point valid flag (map point present,
 * not outlier, invzc >= 0, projection inside the image bounds), projection (u, v), right
 * coordinate ur = u - mbf*invzc, radius = th*scale[octave], GetFeaturesInArea level arguments,
 * keypoint angle, map point descriptor and Observations().  kp_occupied: 1 where
 * CurrentFrame.mvpMapPoints[idx] holds a point with Observations()>0.  out_match_for_kp:
 * >= 0 index of the last-frame point the keypoint ends up with, -1 untouched (caller
 * initialises), -2 reset to NULL by the rotation check.  Returns nmatches. */
typedef struct {
    int n;
    const uint8_t* valid;
    const float* u; const float* v; const float* ur; const float* radius;
    const int* min_level; const int* max_level;
    const float* angle;
    const uint8_t* desc;
    const int* observations;
} orc_framepoints_view;
int   orc_search_by_projection_frame(const orc_frame_view* f, const orc_framepoints_view* pts,
                                     int check_orientation, const uint8_t* kp_occupied,
                                     int* out_match_for_kp);
/* ComputeStereoMatches. pyramids: arrays of nlevels pointers + dims for left and right. */
typedef struct {
    int nlevels;
    const uint8_t* const* data; const int* w; const int* h; const ptrdiff_t* stride;
} orc_pyramid_view;
void  orc_compute_stereo_matches(const orc_frame_view* left, const orc_frame_view* right,
                                 const orc_pyramid_view* pl, const orc_pyramid_view* pr,
                                 const float* inv_scale_factors, float mbf, float mb,
                                 float* uright_out, float* depth_out);
/* SearchByBoW(KeyFrame*, Frame&, ..) / SearchByBoW(KeyFrame*, KeyFrame*, ..) — ORBmatcher.cc:185-325,
 * :610-743 (orc_bow.cpp).  A side = descriptors, keypoint angles, "has a live map point" flags
 * (NULL = all) and its DBoW2::FeatureVector flattened in map order.  th has <= semantics (TH_LOW for
 * the Frame variant, TH_LOW-1 for the KeyFrame pair).  m12[s1->n], m21[s2->n] out; returns nmatches. */
typedef struct {
    int n;
    const uint8_t* desc;
    const float* angle;
    const uint8_t* valid;
    int nnodes;
    const uint32_t* node_id;      /* ascending */
    const int* node_start;        /* nnodes + 1 */
    const uint32_t* indices;
} orc_bow_view;
int   orc_search_by_bow(const orc_bow_view* s1, const orc_bow_view* s2, int th, float nnratio,
                        int check_orientation, int* m12, int* m21);
/* ---- remaining guided searches (orc_guided.cpp) ----
 * Map points projected into a frame / key frame by the caller: valid = every gate in front of the
 * candidate loop passed, (u, v) projection, ur = u - bf*invz (only read by Fuse's reprojection gate),
 * radius = th * mvScaleFactors[level], level = nPredictedLevel, angle = key frame keypoint angle (only
 * read by the Frame/KeyFrame SearchByProjection), desc = pMP->GetDescriptor(). */
typedef struct {
    int n;
    const uint8_t* valid;
    const float* u; const float* v; const float* ur; const float* radius;
    const int* level;
    const float* angle;
    const uint8_t* desc;
} orc_projpoints_view;
int   orc_sbp_frame_keyframe(const orc_frame_view* f, const orc_projpoints_view* pts, int orb_dist,
                             int check_orientation, const uint8_t* kp_occupied, int* match_for_kp);
int   orc_sbp_keyframe_sim3(const orc_frame_view* f, const orc_projpoints_view* pts,
                            const uint8_t* kp_matched, int* match_for_kp);
void  orc_fuse_search(const orc_frame_view* f, const orc_projpoints_view* pts,
                      const float* inv_level_sigma2, int* best_idx, int* best_dist);
int   orc_search_by_sim3(const orc_frame_view* f1, const orc_frame_view* f2, const orc_projpoints_view* pts1,
                         const orc_projpoints_view* pts2, int* matches12);
/* One key frame of SearchForTriangulation: mvKeysUn, descriptors, "GetMapPoint(idx) != NULL" flags,
 * mvuRight (NULL = monocular: all -1) and mFeatVec flattened in map order. */
typedef struct {
    int n;
    const orc_keypoint* kps;
    const uint8_t* desc;
    const uint8_t* has_mappoint;
    const float* uright;
    int nnodes;
    const uint32_t* node_id;
    const int* node_start;
    const uint32_t* indices;
} orc_tri_side;
int   orc_search_for_triangulation(const orc_tri_side* s1, const orc_tri_side* s2, const float* F12, float ex,
                                   float ey, const float* scale_factors2, const float* level_sigma2_2,
                                   int only_stereo, int check_orientation, int* matches12);
/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:266-334) for npoints points: point p owns
 * descriptor rows [start[p], start[p+1]); best[p] = index inside its set, -1 for an empty set. */
void  orc_distinctive_descriptors(const uint8_t* desc, const int* start, int npoints, int* best);
/* Frame constructor epilogue (orc_frame.cpp).  cam = {fx, fy, cx, cy}; dist = {k1, k2, p1, p2, k3}.
 * Frame::UndistortKeyPoints (Frame.cc:443-475): xy / out are n (x, y) pairs. */
void  orc_undistort_points(const float* xy, int n, const float cam[4], const float dist[5], float* out);
/* Frame::ComputeImageBounds (Frame.cc:481-512): bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY} */
void  orc_image_bounds(const float cam[4], const float dist[5], int cols, int rows, float bounds[4]);
/* brute-force kNN-2 (+ratio): out_idx[m] = best train index or -1 if rejected,
 * out_d1/out_d2 best & second distances (always written). */
void  orc_knn2(const uint8_t* q, int m, const uint8_t* t, int n, int th, float ratio,
               int* out_idx, int* out_d1, int* out_d2);

#ifdef __cplusplus
}
#endif
#endif
