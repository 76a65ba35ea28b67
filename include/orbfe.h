/* orbfe.h — C ABI of the B200-native ORB front end (liborbfe.so).
 *
 * This is the drop-in boundary for the reference's feature front end: every entry point
 * below replaces one call the reference makes on its CPU path, and is what a maintainer of
 * saber/ORB_SLAM2_Annotate binds from Frame.cc / ORBmatcher.cc.  Plain
 * pointers and sizes only; no torch / OpenCV types.  All functions return 0 on success or a
 * negative orbfe_status; nothing here ever falls back to a CPU implementation.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   orbfe_extractor_create        ORBextractor::ORBextractor          src/ORBextractor.cc:415-486
 *   orbfe_extract / _batch        ORBextractor::operator()            src/ORBextractor.cc:1119-1197
 *                                 (called from Frame::ExtractORB      src/Frame.cc:272-278)
 *   orbfe_pyramid_level / _host   public member mvImagePyramid        include/ORBextractor.h:86
 *   orbfe_get_scale_tables        GetScaleFactors() & friends         include/ORBextractor.h:64-84
 *   orbfe_hamming256              ORBmatcher::DescriptorDistance      src/ORBmatcher.cc:1828-1844
 *   orbfe_knn2                    best/second-best loop + ratio test  src/ORBmatcher.cc:233-269
 *   orbfe_search_init             ORBmatcher::SearchForInitialization src/ORBmatcher.cc:469-600
 *   orbfe_search_local_points     ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th)
 *                                                                     src/ORBmatcher.cc:51-138
 *   orbfe_search_frame_points     ORBmatcher::SearchByProjection(Frame& Current, const Frame& Last, th, bMono)
 *                                                                     src/ORBmatcher.cc:1484-1630
 *   orbfe_stereo_match            Frame::ComputeStereoMatches         src/Frame.cc:512-686
 */
#ifndef ORBFE_H
#define ORBFE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    ORBFE_OK = 0,
    ORBFE_ERR_ARG = -1,       /* null pointer, bad size, unsupported parameter              */
    ORBFE_ERR_CUDA = -2,      /* a CUDA runtime call failed; see orbfe_last_error()         */
    ORBFE_ERR_CAPACITY = -3,  /* caller buffer / handle capacity too small                  */
    ORBFE_ERR_NO_DEVICE = -4  /* no CUDA device: there is no CPU fallback                   */
} orbfe_status;

/* cv::KeyPoint field-for-field (28 bytes): pt.x pt.y size angle response octave class_id */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbfe_keypoint;

/* Page-locked host memory for image batches and result buffers: transfers from/to memory
 * obtained here run at full PCIe rate and asynchronously; ordinary memory also works. */
void* orbfe_host_alloc(size_t bytes);
void orbfe_host_free(void* p);

const char* orbfe_last_error(void);          /* thread-local, never NULL                    */
const char* orbfe_version(void);
int orbfe_device_count(void);                 /* 0 when no usable CUDA device                */

/* ------------------------------------------------------------------ extractor ---------- */
typedef struct orbfe_extractor orbfe_extractor;

/* One handle == one CUDA stream + private device scratch sized for max_batch frames of at
 * most max_width x max_height pixels.  Handles are independent: calls on different handles
 * may run concurrently from different host threads (the reference runs the left and right
 * extractor of a stereo frame on two threads, src/Frame.cc:78-81). */
int orbfe_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast,
                           int min_th_fast, int max_width, int max_height, int max_batch,
                           int device, orbfe_extractor** out);
void orbfe_extractor_destroy(orbfe_extractor* e);

/* Scale tables exactly as the reference constructor computes them; any pointer may be NULL.
 * scale/inv_scale/sigma2/inv_sigma2: nlevels floats; features_per_level: nlevels ints;
 * umax: 16 ints. */
int orbfe_get_scale_tables(const orbfe_extractor* e, float* scale, float* inv_scale,
                           float* sigma2, float* inv_sigma2, int* features_per_level,
                           int* umax);
int orbfe_extractor_levels(const orbfe_extractor* e);
/* upper bound on keypoints one frame can return (sum over levels of budget + 3) */
int orbfe_extractor_max_keypoints(const orbfe_extractor* e);

/* operator(): one 8-bit grey image in HOST memory -> keypoints + 32-byte descriptors in
 * caller-owned HOST buffers of capacity `cap` keypoints.  Synchronous.  *n receives the
 * keypoint count.  Empty image (w<=0 || h<=0 || img==NULL) returns ORBFE_OK with *n = 0,
 * matching the reference's silent return (src/ORBextractor.cc:1122-1123). */
int orbfe_extract(orbfe_extractor* e, const uint8_t* img, int width, int height,
                  ptrdiff_t stride, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n);

/* Batched operator(): `batch` images of identical size, HOST memory, frame b at
 * imgs + b*frame_stride (rows `stride` bytes apart).  Frame b's results are written to
 * kps + b*cap and desc + b*cap*32; counts[b] = number of keypoints. */
int orbfe_extract_batch(orbfe_extractor* e, const uint8_t* imgs, int batch, int width,
                        int height, ptrdiff_t stride, ptrdiff_t frame_stride,
                        orbfe_keypoint* kps, uint8_t* desc, int cap, int* counts);

/* Device-resident variant for pipelines that already hold frames in HBM: enqueues the
 * kernels on the handle's stream and returns without synchronising.  d_imgs is a device
 * pointer (same layout as above).  Results stay on the device; fetch pointers with
 * orbfe_extractor_device_results after orbfe_extractor_sync.  When d_imgs is 16-byte aligned,
 * stride is a multiple of 16 equal to the padded width and the frames are contiguous, the frames
 * are used in place as pyramid level 0 (no copy): they must stay valid and unmodified until the
 * handle has been synchronised, and orbfe_pyramid_level(.., 0, ..) reads them.  No byte outside
 * [d_imgs, d_imgs + batch * height * stride) is read: the kernels' word loads stay inside a row's pitch
 * and the TMA descriptors of the alias cover exactly `batch` frames (boxes past the end are zero-filled). */
int orbfe_extract_batch_device(orbfe_extractor* e, const uint8_t* d_imgs, int batch, int width,
                               int height, ptrdiff_t stride, ptrdiff_t frame_stride);
int orbfe_extractor_sync(orbfe_extractor* e);
/* device pointers of the last batch: kps [batch][cap] orbfe_keypoint, desc [batch][cap][32],
 * counts [batch] int32; *cap = per-frame slot capacity */
int orbfe_extractor_device_results(orbfe_extractor* e, const orbfe_keypoint** d_kps,
                                   const uint8_t** d_desc, const int32_t** d_counts, int* cap);
/* copy the last batch's results to host buffers (same layout as orbfe_extract_batch) */
int orbfe_extractor_fetch(orbfe_extractor* e, int batch, orbfe_keypoint* kps, uint8_t* desc,
                          int cap, int* counts);
void* orbfe_extractor_stream(orbfe_extractor* e);   /* cudaStream_t of the handle          */
/* number of kernel launches the last extract call enqueued (for bench accounting) */
int orbfe_extractor_last_launches(const orbfe_extractor* e);

/* Per-stage device time of the last extract call, measured with CUDA events on the handle's
 * stream (enable before the call).  Stages, in order: level-0 upload/copy, pyramid resize
 * (all levels), FAST cells, Gaussian blur, quadtree, orientation + descriptors, Frame epilogue
 * (undistortion + grid; zero unless orbfe_extractor_set_camera is active). */
#define ORBFE_NUM_STAGES 7
int orbfe_extractor_set_profiling(orbfe_extractor* e, int on);
int orbfe_extractor_stage_times(orbfe_extractor* e, float* ms, int n);

/* mvImagePyramid[level] of frame `frame` of the last call, copied to HOST (lazy D2H; only
 * the stereo matcher and tests need it).  dst has room for width*height bytes, rows packed
 * at `dst_stride`.  Pass dst == NULL to query the level size only.  blurred != 0 returns the
 * 7x7 Gaussian-blurred working image instead (src/ORBextractor.cc:1169-1175). */
int orbfe_pyramid_level(orbfe_extractor* e, int frame, int level, int blurred, uint8_t* dst,
                        ptrdiff_t dst_stride, int* width, int* height);

/* All levels of frame `frame` of the last call in ONE device-to-host transfer group with a single synchronisation,
 * into page-locked memory owned by the handle (include/ORBextractor.h:86: what Frame::ComputeStereoMatches reads,
 * src/Frame.cc:519, :609-626).  levels[l] points at the first pixel of level l (rows stride[l] bytes apart, width[l] x
 * height[l] pixels); the memory stays valid until the next call of this function on the handle — the same lifetime the
 * reference's mvImagePyramid has, which the next operator() overwrites.  A cv::Mat header over it costs nothing. */
int orbfe_pyramid_host(orbfe_extractor* e, int frame, const uint8_t** levels, int32_t* width, int32_t* height,
                       ptrdiff_t* stride);
/* raw per-level FAST corners of frame `frame` (before the quadtree), reference order
 * (cell row, cell col, y, x); triplets (x, y, score) relative to minBorder. *n = count.
 * Test tap for stage-by-stage parity. */
int orbfe_level_raw_keypoints(orbfe_extractor* e, int frame, int level, int32_t* xys, int cap,
                              int* n);

/* ------------------------------------------------------------------ frame epilogue ----- */
/* What the Frame constructor does right after the extractor returns (SURVEY.md section 8(f)
 * rank 2): Frame::UndistortKeyPoints (src/Frame.cc:443-475, cv::undistortPoints with R = I,
 * P = K), Frame::ComputeImageBounds (:481-512) and Frame::AssignFeaturesToGrid (:246-267).
 * Camera as Frame holds it: mK entries and mDistCoef = (k1 k2 p1 p2 [k3]); k3 = 0 when the
 * settings file has none.  As in the reference only k1 decides whether anything is undistorted. */
typedef struct {
    float fx, fy, cx, cy;
    float k1, k2, p1, p2, k3;
} orbfe_camera;

#define ORBFE_GRID_COLS 64   /* FRAME_GRID_COLS, include/Frame.h:39 */
#define ORBFE_GRID_ROWS 48   /* FRAME_GRID_ROWS, include/Frame.h:38 */
#define ORBFE_GRID_CELLS (ORBFE_GRID_COLS * ORBFE_GRID_ROWS)

/* Fuses the epilogue into the extractor: bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY} of a
 * cols x rows image are computed once (the reference does it for the first frame only,
 * mbInitialComputations) and every later extract call on the handle also leaves mvKeysUn and
 * mGrid of each frame in device memory.  cam == NULL switches the epilogue off again. */
int orbfe_extractor_set_camera(orbfe_extractor* e, const orbfe_camera* cam, int cols, int rows,
                               float bounds[4]);
/* device pointers of the last batch: kps_un [batch][cap] (mvKeysUn; cap as reported by
 * orbfe_extractor_device_results), cell_end [batch][ORBFE_GRID_CELLS] and items [batch][cap]:
 * mGrid[ix][iy] of a frame is items[cell_end[c-1] .. cell_end[c]) with c = ix*48 + iy
 * (cell 0 starts at 0), keypoint indices in push_back order. */
int orbfe_extractor_device_frames(orbfe_extractor* e, const orbfe_keypoint** d_kps_un,
                                  const int32_t** d_cell_end, const uint16_t** d_items);
/* the same copied to HOST: kps_un [batch][cap], cell_end [batch][ORBFE_GRID_CELLS],
 * items [batch][cap]; any pointer may be NULL */
int orbfe_extractor_fetch_frames(orbfe_extractor* e, int batch, orbfe_keypoint* kps_un, int cap,
                                 int32_t* cell_end, uint16_t* items);

/* ------------------------------------------------------------------ matcher ------------ */
/* Scalar Hamming distance between two 256-bit descriptors, host side, for the reference's
 * scalar callers (e.g. MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:303-311). */
int orbfe_hamming256(const void* a, const void* b);

typedef struct orbfe_matcher orbfe_matcher;
/* One handle == one CUDA stream + scratch; stateless between calls, so one handle per host
 * thread reproduces the reference's "ORBmatcher is a stack temporary" threading model. */
int orbfe_matcher_create(int device, orbfe_matcher** out);
void orbfe_matcher_destroy(orbfe_matcher* m);
void* orbfe_matcher_stream(orbfe_matcher* m);
int orbfe_matcher_sync(orbfe_matcher* m);

/* What the matcher routines read from a Frame (all HOST pointers). */
typedef struct {
    int32_t n;                        /* N                                                  */
    const orbfe_keypoint* kps;        /* mvKeysUn                                           */
    const uint8_t* desc;              /* mDescriptors, n x 32                               */
    const float* uright;              /* mvuRight or NULL (= all -1)                        */
    float min_x, max_x, min_y, max_y; /* mnMinX, mnMaxX, mnMinY, mnMaxY                     */
    const float* scale_factors;       /* mvScaleFactors                                     */
    int32_t nlevels;
    int32_t keyframe_bounds;          /* 1: the view is a KeyFrame.  KeyFrame holds mnMinX.. as int
                                       * (include/KeyFrame.h:191-194, truncated copies of the
                                       * Frame's floats, src/KeyFrame.cc:41) while mGrid and
                                       * mfGridElement*Inv are the Frame's (:33, :48-53): the cell
                                       * window of KeyFrame::GetFeaturesInArea (:616-628) is then
                                       * computed from (float)(int)min_x / min_y.  0: Frame.     */
    float grid_inv_w, grid_inv_h;     /* mfGridElementWidthInv / HeightInv when the caller holds them
                                       * (a KeyFrame copies its Frame's, src/KeyFrame.cc:33, and has
                                       * no float bounds to derive them from bit for bit); 0 = derive
                                       * 64 / (max_x - min_x), 48 / (max_y - min_y) as Frame.cc:86-87 */
} orbfe_frame_view;

/* Frame::ComputeImageBounds for callers without an extractor handle (HOST output). */
int orbfe_image_bounds(orbfe_matcher* m, const orbfe_camera* cam, int cols, int rows,
                       float bounds[4]);
/* Frame::UndistortKeyPoints + Frame::AssignFeaturesToGrid of one frame, HOST pointers:
 * kps[n] (mvKeys) -> kps_un[n] (mvKeysUn), cell_end[ORBFE_GRID_CELLS], items[n] (layout as in
 * orbfe_extractor_device_frames); *n_in_grid = keypoints that landed inside the grid. */
int orbfe_frame_assign(orbfe_matcher* m, const orbfe_camera* cam, const float bounds[4],
                       const orbfe_keypoint* kps, int n, orbfe_keypoint* kps_un,
                       int32_t* cell_end, uint16_t* items, int* n_in_grid);

/* A Frame whose keypoints, descriptors and grid already sit in device memory (the extractor's
 * results + epilogue): d_kps = mvKeysUn, d_cell_end / d_items = mGrid as above; d_uright may be
 * NULL.  scale_factors stays a HOST pointer. */
typedef struct {
    int32_t n;
    const orbfe_keypoint* d_kps;
    const uint8_t* d_desc;
    const float* d_uright;
    const int32_t* d_cell_end;
    const uint16_t* d_items;
    float min_x, max_x, min_y, max_y;
    const float* scale_factors;
    int32_t nlevels;
    int32_t keyframe_bounds;          /* as in orbfe_frame_view */
    float grid_inv_w, grid_inv_h;     /* as in orbfe_frame_view (0 = derive) */
} orbfe_device_frame_view;

/* Brute-force 2-nearest-neighbour Hamming search with the reference's acceptance rule
 * (strict '<' updates so the lowest train index wins ties; accept iff d1 <= th and
 * (float)d1 < ratio*(float)d2): out_idx[i] = train index or -1, out_d1/out_d2 = best and
 * second-best distances (256 when absent).  HOST pointers. */
int orbfe_knn2(orbfe_matcher* m, const uint8_t* query, int nq, const uint8_t* train, int nt,
               int th, float ratio, int32_t* out_idx, int32_t* out_d1, int32_t* out_d2);
/* Same with DEVICE pointers, asynchronous on the matcher's stream. */
int orbfe_knn2_device(orbfe_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_train,
                      int nt, int th, float ratio, int32_t* d_out_idx, int32_t* d_out_d1,
                      int32_t* d_out_d2);

/* Results of the device-resident entry points (orbfe_knn2_device, orbfe_search_init_consecutive_device, ...) to HOST
 * memory: an asynchronous copy of `bytes` bytes on the matcher's stream, ordered behind the searches enqueued before it;
 * orbfe_matcher_sync makes it visible.  Page-locked `dst` (orbfe_host_alloc) keeps it asynchronous. */
int orbfe_matcher_fetch(orbfe_matcher* m, void* dst, const void* d_src, size_t bytes);

/* Which engine serves orbfe_knn2 / _device / _sharded.  ORBFE_KNN_POPC: 8 x (XOR + POPC) per comparison on
 * the integer pipe, the reference's own form (src/ORBmatcher.cc:1835-1841).  ORBFE_KNN_MMA and ORBFE_KNN_FP4: the
 * distance as a contraction on the tensor cores, <a', b'> = 256 - 2 d over bits mapped to +-1 — as signed bytes expanded
 * in shared memory (tcgen05.mma kind::i8) or as E2M1 values with unit block scales written once by a pre-pass
 * (tcgen05.mma kind::mxf4.block_scale, twice the rate; needs 128 bytes of scratch per row), accumulators in TMEM.  All
 * three give identical results; ORBFE_KNN_AUTO (default) takes the FP4 engine from 2^26 comparisons per call upwards.
 * The environment variable ORBFE_KNN_ENGINE = popc | mma | fp4 overrides AUTO for a whole process. */
enum { ORBFE_KNN_AUTO = 0, ORBFE_KNN_POPC = 1, ORBFE_KNN_MMA = 2, ORBFE_KNN_FP4 = 3 };
int orbfe_matcher_set_knn_engine(orbfe_matcher* m, int engine);

/* Allocation-free variant of orbfe_knn2_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  No handle: `device` is the CUDA device, `cuda_stream` a cudaStream_t of that device (NULL = the
 * default stream); inputs and outputs are the caller's DEVICE buffers as in orbfe_knn2_device; the search always runs
 * on the tensor-core engine.  The call performs no cudaMalloc / cudaFree or any other allocation: the engine's scratch
 * (12 bytes x nq x train segments) is a static device array of ORBFE_KNN2_NOALLOC_SCRATCH_BYTES that
 * every GPU loading the library holds until the process exits.  Sizes that need more are refused with
 * ORBFE_ERR_CAPACITY and a message (one segment per 2^20 train rows once nq >= 37 633, so up to 699 050 queries
 * against up to 1 048 576 train rows; smaller query sets use up to 148 / ceil(nq / 256) segments).  All calls on one
 * GPU share the array, which keeps the previous call's contents (nothing in it needs clearing: every slot read is
 * written first by the same call); calls that OVERLAP on one GPU — different streams without an event between
 * them, or different threads — race on it and return garbage.  Asynchronous on `cuda_stream`. */
#define ORBFE_KNN2_NOALLOC_SCRATCH_BYTES (8 << 20)
int orbfe_knn2_device_noalloc(int device, void* cuda_stream, const uint8_t* d_query, int nq, const uint8_t* d_train,
                              int nt, int th, float ratio, int32_t* d_out_idx, int32_t* d_out_d1, int32_t* d_out_d2);

/* ---- BASELINE.json configs[4]: brute-force kNN-2 with the QUERY rows sharded over the GPUs of one node and an
 * NCCL all-gather of the per-query results (the reference has no counterpart: its
 * loops are src/ORBmatcher.cc:233-269).  One process per GPU.  NCCL is bound at run time (dlopen of
 * libnccl.so.2: the copy the host framework already loaded, else the system one), so liborbfe.so does not
 * link it and every other entry point works without it. */
typedef struct orbfe_comm orbfe_comm;
#define ORBFE_COMM_ID_BYTES 128
/* rank 0 creates the id (ncclGetUniqueId) and hands the 128 bytes to the other ranks by any means */
int orbfe_comm_unique_id(uint8_t id[ORBFE_COMM_ID_BYTES]);
/* collective over all ranks (ncclCommInitRank); `device` is this rank's CUDA device */
int orbfe_comm_create(const uint8_t id[ORBFE_COMM_ID_BYTES], int world, int rank, int device, orbfe_comm** out);
/* a communicator the caller already owns (ncclComm_t); orbfe_comm_destroy leaves it alive */
int orbfe_comm_wrap(void* nccl_comm, int world, int rank, int device, orbfe_comm** out);
void orbfe_comm_destroy(orbfe_comm* c);
int orbfe_comm_world(const orbfe_comm* c);
int orbfe_comm_rank(const orbfe_comm* c);

/* one gathered result row: 8 bytes per query */
typedef struct orbfe_knn_result {
    int32_t idx;        /* accepted train row, or -1 */
    uint16_t d1, d2;    /* best and second-best distance (256 when absent) */
} orbfe_knn_result;

/* Rows per shard: rank r owns query rows [r * per, min((r + 1) * per, nq)). */
int orbfe_knn2_shard_rows(int nq, int world);
/* Every rank passes the same nq query rows and the same train set (DEVICE pointers, replicated); rank r
 * searches its shard, packs (idx, d1, d2) into 8 bytes per query and one in-place ncclAllGather on the matcher's
 * stream leaves all nq results on every rank.  d_out must hold world * orbfe_knn2_shard_rows(nq, world)
 * rows (the tail past nq is padding).  Asynchronous: orbfe_matcher_sync before reading. */
int orbfe_knn2_sharded(orbfe_matcher* m, orbfe_comm* c, const uint8_t* d_query, int nq, const uint8_t* d_train,
                       int nt, int th, float ratio, orbfe_knn_result* d_out);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize).
 * prev_matched: 2*f1->n floats (x,y), in/out; matches12: f1->n ints out; *nmatches out. */
int orbfe_search_init(orbfe_matcher* m, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                      float* prev_matched, int32_t* matches12, int window, float nnratio,
                      int check_orientation, int* nmatches);

/* Batched, device-resident SearchForInitialization over a stream of frames laid out as the
 * extractor's device results ([nframes][cap] keypoints / descriptors, counts[nframes]): pair p
 * matches frame p (F1) against frame p+1 (F2) with vbPrevMatched = F1's keypoint positions (the
 * state right after Tracking::MonocularInitialization stores the initial frame,
 * src/Tracking.cc:609-616).  d_matches12 [nframes-1][cap], d_nmatches [nframes-1].
 * Asynchronous on the matcher's stream. */
int orbfe_search_init_consecutive_device(orbfe_matcher* m, const orbfe_keypoint* d_kps,
                                         const uint8_t* d_desc, const int32_t* d_counts, int cap,
                                         int nframes, float min_x, float max_x, float min_y,
                                         float max_y, int window, float nnratio,
                                         int check_orientation, int32_t* d_matches12,
                                         int32_t* d_nmatches);

/* Map-point side of SearchByProjection(Frame&, vector<MapPoint*>&, th) (HOST pointers). */
typedef struct {
    int32_t n;
    const uint8_t* track_in_view;     /* mbTrackInView && !isBad()                          */
    const float* proj_x;              /* mTrackProjX                                        */
    const float* proj_y;              /* mTrackProjY                                        */
    const float* proj_xr;             /* mTrackProjXR                                       */
    const int32_t* level;             /* mnTrackScaleLevel                                  */
    const float* view_cos;            /* mTrackViewCos                                      */
    const uint8_t* desc;              /* GetDescriptor(), n x 32                            */
    const int32_t* observations;      /* Observations()                                     */
} orbfe_mappoints_view;
/* kp_occupied[f->n] in/out: 1 where F.mvpMapPoints[idx] holds a point with Observations()>0.
 * mp_for_kp[f->n] out: map point index assigned to each keypoint, -1 if none (the caller
 * initialises it to -1 and writes F.mvpMapPoints[idx] = vpMapPoints[mp_for_kp[idx]]). */
int orbfe_search_local_points(orbfe_matcher* m, const orbfe_frame_view* f,
                              const orbfe_mappoints_view* mps, float th, float nnratio,
                              uint8_t* kp_occupied, int32_t* mp_for_kp, int* nmatches);

/* Same search against a device-resident frame: nothing of the frame is uploaded and the grid
 * is not rebuilt.  The map-point side and the results stay HOST arrays. */
int orbfe_search_local_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* f,
                                       const orbfe_mappoints_view* mps, float th, float nnratio,
                                       uint8_t* kp_occupied, int32_t* mp_for_kp, int* nmatches);

/* Points of the last frame projected into the current one by the motion model, for
 * SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1484-1630).  The
 * pose arithmetic (x3Dc = Rcw*x3Dw + tcw, u, v, invzc, bForward / bBackward, :1494-1553) stays
 * with the caller, which passes per point i of the last frame (all HOST arrays of length n):
 *   valid        LastFrame.mvpMapPoints[i] != NULL && !mvbOutlier[i] && invzc >= 0 && (u, v)
 *                inside [mnMinX, mnMaxX] x [mnMinY, mnMaxY]                          (:1510-1538)
 *   u, v         projection                                                           (:1528-1529)
 *   ur           u - mbf * invzc (only read when the frame has uright)                (:1575)
 *   radius       th * CurrentFrame.mvScaleFactors[nLastOctave]                        (:1543)
 *   min_level, max_level   arguments of GetFeaturesInArea: (o, -1) forward, (0, o) backward,
 *                (o - 1, o + 1) otherwise, o = LastFrame.mvKeys[i].octave             (:1547-1552)
 *   angle        LastFrame.mvKeysUn[i].angle                                          (:1594)
 *   desc         pMP->GetDescriptor(), n x 32; observations  pMP->Observations()      (:1557,:1568) */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* u;
    const float* v;
    const float* ur;
    const float* radius;
    const int32_t* min_level;
    const int32_t* max_level;
    const float* angle;
    const uint8_t* desc;
    const int32_t* observations;
} orbfe_framepoints_view;
/* kp_occupied[f->n] in: 1 where CurrentFrame.mvpMapPoints[idx] holds a point with
 * Observations()>0.  match_for_kp[f->n] out: >= 0 index i of the last-frame point whose map
 * point the keypoint ends up holding, -1 untouched, -2 reset to NULL by the rotation check.
 * *nmatches = the reference's return value. */
int orbfe_search_frame_points(orbfe_matcher* m, const orbfe_frame_view* f,
                              const orbfe_framepoints_view* pts, int check_orientation,
                              const uint8_t* kp_occupied, int32_t* match_for_kp, int* nmatches);

int orbfe_search_frame_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* f,
                                       const orbfe_framepoints_view* pts, int check_orientation,
                                       const uint8_t* kp_occupied, int32_t* match_for_kp,
                                       int* nmatches);

/* The same first-owner search with the accepted best distance as an argument; it serves the two other
 * SearchByProjection overloads that assign map points to keypoints as they go:
 *   SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist)
 *     (src/ORBmatcher.cc:1641-1770, relocalisation): point i = pKF's i-th map point; valid folds
 *     pMP != NULL, !isBad, !sAlreadyFound.count, the image-bounds and distance gates (:1663-1698);
 *     radius = th * mvScaleFactors[nPredictedLevel], min_level / max_level = nPredictedLevel -/+ 1
 *     (:1703-1705); angle = pKF->mvKeysUn[i].angle; observations = 1 for every point and
 *     kp_occupied = (CurrentFrame.mvpMapPoints[i2] != NULL), because any held point blocks (:1720);
 *     max_dist = ORBdist.
 *   SearchByProjection(KeyFrame* pKF, Scw, vpPoints, vpMatched, th) (:335-451, loop closing): point
 *     i = vpPoints[i]; valid folds !isBad, !spAlreadyFound.count and the gates of :361-402;
 *     min_level = nPredictedLevel - 1, max_level = nPredictedLevel (:426); kp_occupied =
 *     (vpMatched[idx] != NULL) (:421); observations = 1; max_dist = TH_LOW (:440); no orientation
 *     check, f->uright = NULL.  match_for_kp[idx] = i where vpMatched[idx] = vpPoints[i]. */
int orbfe_search_projected_points(orbfe_matcher* m, const orbfe_frame_view* f,
                                  const orbfe_framepoints_view* pts, int max_dist,
                                  int check_orientation, const uint8_t* kp_occupied,
                                  int32_t* match_for_kp, int* nmatches);
int orbfe_search_projected_points_resident(orbfe_matcher* m, const orbfe_device_frame_view* f,
                                           const orbfe_framepoints_view* pts, int max_dist,
                                           int check_orientation, const uint8_t* kp_occupied,
                                           int32_t* match_for_kp, int* nmatches);

/* Map points projected into a key frame by the caller, for the searches in which every point looks
 * for its best keypoint on its own (HOST arrays of length n):
 *   valid    every gate in front of the candidate loop passed (isBad / IsInKeyFrame / depth > 0 /
 *            IsInImage / distance invariance / viewing angle)
 *   u, v     projection;  ur = u - bf*invz (read only with inv_level_sigma2 and a frame with uright)
 *   radius   th * pKF->mvScaleFactors[nPredictedLevel];  level = nPredictedLevel: candidates are
 *            the keypoints of KeyFrame::GetFeaturesInArea(u, v, radius) (src/KeyFrame.cc:611-650)
 *            with octave in [level-1, level]
 *   desc     pMP->GetDescriptor(), n x 32 */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* u;
    const float* v;
    const float* ur;
    const float* radius;
    const int32_t* level;
    const uint8_t* desc;
} orbfe_windowpoints_view;
/* The search of ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) (src/ORBmatcher.cc:940-1090; pass
 * inv_level_sigma2 = pKF->mvInvLevelSigma2, f->nlevels floats, to enable its reprojection gate: e2 *
 * invSigma2 > 7.8 with a right coordinate, > 5.99 without, :1020-1041) and of Fuse(KeyFrame*, Scw,
 * vpPoints, th, vpReplacePoint) (:1100-1237; inv_level_sigma2 = NULL).  best_dist[i] = smallest
 * distance among the candidates (256: none; may be NULL), best_idx[i] = its keypoint when best_dist
 * <= max_dist (TH_LOW in both overloads), else -1.  What happens to the pair — Replace, AddObservation
 * / AddMapPoint, vpReplacePoint — is map mutation and stays with the caller, in point order. */
int orbfe_search_window_best(orbfe_matcher* m, const orbfe_frame_view* f,
                             const orbfe_windowpoints_view* pts, const float* inv_level_sigma2,
                             int max_dist, int32_t* best_idx, int32_t* best_dist);
int orbfe_search_window_best_resident(orbfe_matcher* m, const orbfe_device_frame_view* f,
                                      const orbfe_windowpoints_view* pts,
                                      const float* inv_level_sigma2, int max_dist,
                                      int32_t* best_idx, int32_t* best_dist);
/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1251-1477).  pts1[i1] = map point of key frame 1's
 * keypoint i1 projected into key frame 2 (pts1->n == f1->n; valid folds pMP != NULL,
 * !vbAlreadyMatched1, !isBad and the gates of :1300-1326), pts2 the other way round.  Both searches
 * accept best distance <= max_dist (TH_HIGH); matches12[i1] = idx2 where the two agree (:1446-1460),
 * else -1; *nfound = the reference's return value. */
int orbfe_search_by_sim3(orbfe_matcher* m, const orbfe_frame_view* f1, const orbfe_frame_view* f2,
                         const orbfe_windowpoints_view* pts1, const orbfe_windowpoints_view* pts2,
                         int max_dist, int32_t* matches12, int* nfound);

/* One key frame of ORBmatcher::SearchForTriangulation (HOST pointers): mvKeysUn, mDescriptors,
 * has_mappoint[i] = (GetMapPoint(i) != NULL), mvuRight (NULL = monocular, all -1) and mFeatVec
 * flattened in map order as in orbfe_bow_view. */
typedef struct {
    int32_t n;
    const orbfe_keypoint* kps;
    const uint8_t* desc;
    const uint8_t* has_mappoint;
    const float* uright;
    int32_t nnodes;
    const uint32_t* node_id;
    const int32_t* node_start;
    const uint32_t* indices;
} orbfe_tri_side;
/* SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:754-928).
 * F12 row-major 3x3; (ex, ey) = epipole of key frame 1's centre in key frame 2 (:762-768, computed
 * by the caller); scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2 (nlevels2
 * floats).  matches12[s1->n]: index in key frame 2 or -1 — vMatchedPairs is its list of (i, >= 0)
 * entries in ascending i (:918-925); *nmatches = the reference's return value. */
int orbfe_search_triangulation(orbfe_matcher* m, const orbfe_tri_side* s1, const orbfe_tri_side* s2,
                               const float F12[9], float ex, float ey, const float* scale_factors2,
                               const float* level_sigma2_2, int nlevels2, int only_stereo,
                               int check_orientation, int32_t* matches12, int* nmatches);

/* One side of ORBmatcher::SearchByBoW (HOST pointers): descriptors, the keypoint angles the
 * rotation check reads, "holds a live map point" flags and the side's DBoW2::FeatureVector
 * (std::map<NodeId, std::vector<unsigned int>>) flattened in map order. */
typedef struct {
    int32_t n;                  /* features                                              */
    const uint8_t* desc;        /* mDescriptors, n x 32                                  */
    const float* angle;         /* KeyFrame: mvKeysUn[i].angle; Frame: mvKeys[i].angle   */
    const uint8_t* valid;       /* pMP != NULL && !pMP->isBad(); NULL = every feature    */
    int32_t nnodes;
    const uint32_t* node_id;    /* ascending NodeIds                                     */
    const int32_t* node_start;  /* nnodes + 1 offsets into indices                       */
    const uint32_t* indices;    /* feature indices, vector order inside a node           */
} orbfe_bow_view;
/* SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) (src/ORBmatcher.cc:185-325): s1 = pKF,
 * s2 = F with valid = NULL, th = TH_LOW (accept bestDist1 <= th, :263); the caller sets
 * vpMapPointMatches[i2] = vpMapPointsKF[matches21[i2]] where matches21[i2] >= 0.
 * SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12) (:610-743): both sides carry valid
 * flags, th = TH_LOW - 1 (bestDist1 < TH_LOW, :686); vpMatches12[i1] = vpMapPoints2[matches12[i1]].
 * matches12[s1->n] and matches21[s2->n] are both written (partner index or -1; either may be
 * NULL); *nmatches = the reference's return value. */
int orbfe_search_bow(orbfe_matcher* m, const orbfe_bow_view* s1, const orbfe_bow_view* s2, int th,
                     float nnratio, int check_orientation, int32_t* matches12,
                     int32_t* matches21, int* nmatches);

/* npairs searches in one call (s1[p] against s2[p]; matches12[p] / matches21[p] as above, the
 * pointer arrays or single entries may be NULL; nmatches[npairs]).  A side that appears in several
 * pairs with the same arrays — the current frame of Tracking::Relocalization against every
 * candidate key frame, src/Tracking.cc:1486-1510 — is uploaded once. */
int orbfe_search_bow_batch(orbfe_matcher* m, int npairs, const orbfe_bow_view* s1,
                           const orbfe_bow_view* s2, int th, float nnratio, int check_orientation,
                           int32_t* const* matches12, int32_t* const* matches21, int* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:266-334) for npoints map points in
 * one call (HOST pointers): point p owns the descriptor rows desc[32*start[p] .. 32*start[p+1])
 * — the rows the reference gathers from the point's observations, in map order, bad key frames
 * left out (:283-289).  best[p] = index inside the point's own set of the descriptor with the
 * least median Hamming distance to the set (median = element (N-1)/2 of the sorted row, the
 * first row wins ties), -1 for an empty set; the caller copies that row into mDescriptor. */
int orbfe_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* start,
                                  int npoints, int32_t* best);

/* Pyramid of one extractor call, HOST pointers (interior pixels of mvImagePyramid[l]). */
typedef struct {
    int32_t nlevels;
    const uint8_t* const* data;
    const int32_t* width;
    const int32_t* height;
    const ptrdiff_t* stride;
} orbfe_pyramid_view;
/* Frame::ComputeStereoMatches: fills uright[left->n] (mvuRight) and depth[left->n]
 * (mvDepth).  mbf, mb as in Frame (maxD = mbf/mb). */
int orbfe_stereo_match(orbfe_matcher* m, const orbfe_frame_view* left,
                       const orbfe_frame_view* right, const orbfe_pyramid_view* pyr_left,
                       const orbfe_pyramid_view* pyr_right, const float* inv_scale_factors,
                       float mbf, float mb, float* uright, float* depth);

/* mvImagePyramid of every frame of the extractor's last batch where it lies, in DEVICE memory: level l
 * of frame b starts at d_data[l] + b * frame_stride[l], rows pitch[l] bytes apart, interior pixels only.
 * The arrays belong to the handle and stay valid until its next extract call or size change. */
typedef struct {
    int32_t nlevels;
    const uint8_t* const* d_data;
    const int32_t* width;
    const int32_t* height;
    const ptrdiff_t* pitch;
    const ptrdiff_t* frame_stride;
} orbfe_device_pyramid_view;
int orbfe_extractor_device_pyramid(orbfe_extractor* e, orbfe_device_pyramid_view* out);
/* Frame::ComputeStereoMatches (src/Frame.cc:512-686) for npairs stereo pairs that went through the
 * extractor as ONE batch with left and right images interleaved (frame 2p = left, 2p+1 = right of pair
 * p): keypoints, descriptors, counts (orbfe_extractor_device_results) and both pyramids
 * (orbfe_extractor_device_pyramid) are read in place, nothing is uploaded.  The extractor's work must
 * have completed (any host-buffer extract call, or orbfe_extractor_sync).  uright / depth: HOST arrays
 * [npairs][cap], -1 where there is no match and past the pair's left keypoint count. */
int orbfe_stereo_match_batch_device(orbfe_matcher* m, const orbfe_device_pyramid_view* pyr,
                                    const orbfe_keypoint* d_kps, const uint8_t* d_desc,
                                    const int32_t* d_counts, int cap, int npairs,
                                    const float* scale_factors, const float* inv_scale_factors,
                                    int nlevels, float mbf, float mb, float* uright, float* depth);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_H */
