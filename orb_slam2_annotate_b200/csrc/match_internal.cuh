// match_internal.cuh — launchers of the matcher kernels (liborbfe, sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "orbfe_internal.cuh"

namespace orbfe {

// brute-force kNN-2; device pointers, rows 32 bytes, 16-byte aligned. returns launches.
int launch_knn2(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                int* d_idx, int* d_d1, int* d_d2, cudaStream_t s);

// tensor-core engine (match_knn_mma.cu: tcgen05.mma kind::i8 on bit -> byte expanded rows, accumulators in TMEM);
// scratch = knn2_mma_scratch_bytes(nq, nt) bytes of device memory for the per-segment partial results
size_t knn2_mma_scratch_bytes(int nq, int nt);
int launch_knn2_mma(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                    int* d_idx, int* d_d1, int* d_d2, void* scratch, cudaStream_t s);
// block-scaled FP4 engine (match_knn_fp4.cu: tcgen05.mma kind::mxf4 on +-1 E2M1 rows written by a pre-pass, operands by
// TMA); scratch = knn2_fp4_scratch_bytes(nq, nt): partial results + both row sets at 128 bytes per row
size_t knn2_fp4_scratch_bytes(int nq, int nt);
int launch_knn2_fp4(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int th, float ratio,
                    int* d_idx, int* d_d1, int* d_d2, void* scratch, cudaStream_t s);
// shared by the two tensor-core engines: train segments for CTAs of mq queries and tiles of ntile rows; the merge of
// the per-segment partial results (TH_LOW / ratio test included)
int knn2_tile_segments(int nq, int nt, int mq, int ntile, int* segLen, int minWant = 1);
int launch_knn2_merge(int nq, int nseg, int th, float ratio, const int* part_idx, const int* part_d1, const int* part_d2,
                      int* d_idx, int* d_d1, int* d_d2, cudaStream_t s);
// (idx, d1, d2) of rows [first, first + n) -> packed orbfe_knn_result rows (the sharded search's gather payload)
int launch_knn2_pack(const int* d_idx, const int* d_d1, const int* d_d2, int n, void* d_packed, cudaStream_t s);

// minX / minY place keypoints (PosInGrid); qminX / qminY are what GetFeaturesInArea subtracts: the same for a Frame,
// the int-truncated bounds for a KeyFrame (include/KeyFrame.h:191-194)
struct GridGeom { float minX, minY, invW, invH, qminX, qminY; };

struct InitArgs {
    const orbfe_keypoint* kps;   // [frames][cap]
    const uint8_t* desc;         // [frames][cap][32]
    const int* counts;           // [frames]
    int cap;
    const int2* pairs;           // frame indices (F1, F2) per pair, or NULL for (p, p+1)
    float* prev;                 // [pairs][cap][2] vbPrevMatched in/out, or NULL: F1 keypoint positions
    int* m12;                    // [pairs][cap] vnMatches12
    int* nmatches;               // [pairs]
    // scratch (search_init_carve): CSR grid of the second frame and per-query candidate lists
    int* cell_end;               // [pairs][64*48]
    unsigned short* items;       // [pairs][cap]
    uint32_t* tk_key;            // [pairs][cap][32]
    unsigned short* tk_idx;      // [pairs][cap][32]
    unsigned short* tk_total;    // [pairs][cap]
    unsigned short* qlist;       // [pairs][cap] octave-0 keypoints of the first frame, in order
    int* nq;                     // [pairs]
    GridGeom grid;
    int window;
    float ratio;
    int checkOri;
};

struct LocalArgs {
    const orbfe_keypoint* kps;   // frame keypoints [n]
    const uint8_t* desc;         // [n][32]
    const float* uright;         // [n] or NULL
    int n;
    const float* scale;          // mvScaleFactors
    int nlevels;
    // map points
    int nmp;
    const uint8_t* in_view;
    const float* px; const float* py; const float* pxr;
    const int* level; const float* view_cos;
    const uint8_t* mpdesc;
    const int* obs;
    float th, ratio;
    // mode 1 = SearchByProjection(CurrentFrame, LastFrame) (ORBmatcher.cc:1484-1630): per point
    // search radius and octave range come precomputed, best candidate only, rotation histogram
    int mode;
    const float* radius; const int* min_level; const int* max_level; const float* angle;
    int checkOri; int* hist;     // [32] rotation histogram (mode 1)
    int max_dist;                // mode 1: accepted best distance (TH_HIGH, ORBdist or TH_LOW depending on the overload)
    int* kp_bad;                 // [n] keypoint lost its match to the rotation check (mode 1)
    GridGeom grid;
    // grid CSR of the frame
    int* cell_end; unsigned short* items;
    // candidate lists: cand_off[nmp+1] (exclusive scan of counts), cand[] packed idx | dist<<16 | octave<<25
    int* cand_off; uint32_t* cand;
    int cand_cap;                // entries cand can hold; a longer list is not written (the host retries)
    // fixed-point state
    const uint8_t* occ0;         // keypoint already holds a map point with observations
    int* owner_prev; int* owner_next; int* choice;
    // convergence check folded into the next kernel: *diff_flag = 1 where owner_pp != owner_prev
    // (owner_pp = the sweep before owner_prev; NULL = nothing to compare)
    const int* owner_pp; int* diff_flag;
    int* ncand_out;              // the epilogue copies the candidate total here (read back with the results)
};

// K14: independent best-in-window search of projected map points (Fuse x2, SearchBySim3)
struct WindowArgs {
    const orbfe_keypoint* kps; const uint8_t* desc; const float* uright; int n;   // key frame side (mvKeysUn, mDescriptors, mvuRight)
    GridGeom grid; const int* cell_end; const unsigned short* items;               // mGrid as CSR
    int np;
    const uint8_t* valid; const float* u; const float* v; const float* ur; const float* radius;
    const int* level;            // nPredictedLevel: candidates of octave level-1 .. level
    const uint8_t* pdesc;
    const float* inv_sigma2;     // mvInvLevelSigma2 (device) or NULL: no reprojection gate
    int max_dist;                // best_idx = -1 above it
    int* best_idx; int* best_dist;
};
int launch_build_grid(const orbfe_keypoint* kps, int n, const GridGeom& grid, int* cell_end, unsigned short* items,
                      cudaStream_t st);
int launch_window_best(const WindowArgs& a, cudaStream_t st);
int launch_sim3_agree(const int* match1, const int* match2, int n1, int n2, int* m12, int* nfound, cudaStream_t st);

struct StereoArgs {
    const orbfe_keypoint* kl; const uint8_t* dl; int nl;
    const orbfe_keypoint* kr; const uint8_t* dr; int nr;
    const float* scale; const float* inv_scale; int nlevels;
    // per level: image pointers, pitch, width (left and right pyramids)
    const uint8_t* imgl[kMaxLevels]; const uint8_t* imgr[kMaxLevels];
    int pitchl[kMaxLevels], pitchr[kMaxLevels], widthr[kMaxLevels];
    int nrows;          // rows of level 0 (mvImagePyramid[0].rows)
    float mbf, mb;
    float* uright; float* depth;
    int* sad;           // per left keypoint: truncated best SAD, or -1 when no match
    // batched, device-resident form (orbfe_stereo_match_batch_device): pair p = frames 2p (left) and 2p+1 (right) of
    // one extractor batch; kl / dl are the batch's [frames][cap] arrays, imgl[l] the level stacks, outputs [pairs][cap]
    int batched, cap;
    const int* counts;
    long long frameStride[kMaxLevels];
};

// SearchByBoW: two sides = descriptors, keypoint angles, live-map-point flags (NULL = all) and the
// DBoW2::FeatureVector as CSR (ascending node ids, feature indices in vector order)
struct BowArgs {
    int n1, n2;
    const uint8_t* desc1; const uint8_t* desc2;
    const float* angle1; const float* angle2;
    const uint8_t* valid1; const uint8_t* valid2;
    int nnodes1, nnodes2;
    const uint32_t* node_id1; const uint32_t* node_id2;
    const int* node_start1; const int* node_start2;
    const uint32_t* indices1; const uint32_t* indices2;
    int th; float ratio; int checkOri;
    int* m12; int* m21;          // [n1], [n2]: partner index or -1
    int* hist;                   // [30]
    int* nmatches;
};
// d_pairs: npairs argument records in device memory; minus1 / zero: the result block's "-1" part
// (every pair's m12, m21) and "0" part (every pair's histogram + count)
int launch_search_bow(const BowArgs* d_pairs, int npairs, int max_nodes1, int max_n1, bool check_ori, int* minus1,
                      int nminus1, int* zero, int nzero, cudaStream_t st);

// K15: SearchForTriangulation — both key frames' mvKeysUn, descriptors, "has a map point" flags,
// mvuRight (NULL = monocular) and feature vectors as CSR; F12 row-major, epipole (ex, ey)
struct TriArgs {
    int n1, n2, nidx1;           // nidx1 = node_start1[nnodes1]
    const orbfe_keypoint* kps1; const orbfe_keypoint* kps2;
    const uint8_t* desc1; const uint8_t* desc2;
    const uint8_t* has1; const uint8_t* has2;
    const float* ur1; const float* ur2;
    int nnodes1, nnodes2;
    const uint32_t* node_id1; const uint32_t* node_id2;
    const int* node_start1; const int* node_start2;
    const uint32_t* indices1; const uint32_t* indices2;
    float F[9]; float ex, ey;
    const float* scale2; const float* sigma2_2;   // pKF2->mvScaleFactors / mvLevelSigma2 (device)
    int th, onlyStereo, checkOri;
    int* m12; int* hist; int* nmatches;
};
int launch_search_triangulation(const TriArgs& a, int* zero, int nzero, cudaStream_t st);

// MapPoint::ComputeDistinctiveDescriptors for a batch: point p owns rows [start[p], start[p+1]) of desc
int launch_distinctive(const uint8_t* d_desc, const int* d_start, int npoints, int* d_best, cudaStream_t st);

size_t search_init_scratch_bytes(int cap, int npairs);
void search_init_carve(InitArgs& a, void* scratch, int npairs);
int launch_search_init(const InitArgs& a, int npairs, cudaStream_t st);
// build_grid = false: a.cell_end / a.items already hold the frame's grid (device-resident frame)
int launch_local_grid_and_count(const LocalArgs& a, cudaStream_t st, bool build_grid = true);
int launch_local_fill(const LocalArgs& a, cudaStream_t st);
int launch_local_sweep(const LocalArgs& a, cudaStream_t st);
// one launch instead of five memsets / copies: owners (first = first sweep input, included when
// reset_first) to "no owner", mp_for_kp to -1, `nzero` ints at zero_block to 0, occ_out = occ0
int launch_guided_init(int* owners, int n, int nsweeps, bool reset_first, int* mp_for_kp, int* zero_block, int nzero,
                       uint8_t* occ_out, const uint8_t* occ0, cudaStream_t st);
int launch_local_finalize(const LocalArgs& a, int* mp_for_kp, uint8_t* occ_out, int* nmatches, cudaStream_t st);
int launch_frame_finalize(const LocalArgs& a, int* mp_for_kp, int* nmatches, cudaStream_t st);
int launch_stereo(const StereoArgs& a, cudaStream_t st, int npairs = 1);

}  // namespace orbfe
