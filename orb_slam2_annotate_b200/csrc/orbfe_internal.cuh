// orbfe_internal.cuh — shared device-side definitions of liborbfe (B200 / sm_100a).
//
// Data layout in HBM (one extractor handle, capacity B frames):
//   pyr   [L][B][h_l][pitch_l]   level-major: every level is one stack of B images, interior
//                                pixels only
//                                (the reference's 19-px replicated frame around each level is
//                                never read on this path: SURVEY.md §8a), row pitch = width
//                                rounded up to 16 B so that rows can be moved as uint4
//   blur  same geometry          7x7 sigma-2 Gaussian of every level
//   cellc [B][ncells]            FAST corners found per 30-px cell (after per-cell NMS)
//   cellk [B][cell_slots]        the corners of each cell, row-major inside the cell, packed
//                                x | y<<12 | score<<24 (coordinates relative to minBorder)
//   sel   [B][L][sel_cap]        quadtree survivors per level (same packing) + selc [B][L]
//   kps   [B][cap] orbfe_keypoint, desc [B][cap][32], counts [B]      -> the results
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbfe.h"

namespace orbfe {

constexpr int kMaxLevels = 16;
// TMA boxes of the orientation / descriptor kernel.  A box must start on a 16-byte boundary, so
// the 31 (37) wanted columns sit at offset 0..15 inside 48 (64) fetched ones.
constexpr int kTmaRawW = 48, kTmaRawH = 31, kTmaBlurW = 64, kTmaBlurH = 37;
constexpr int kEdgeThreshold = 19;   // ORBextractor.cc:75
constexpr int kHalfPatch = 15;       // ORBextractor.cc:74
constexpr int kPatchSize = 31;       // ORBextractor.cc:73
constexpr int kCellW = 30;           // ORBextractor.cc:819
constexpr int kMaxDim = 4096;        // 12-bit packed coordinates
constexpr int kFastMaxWin = 72;      // largest FAST cell window side handled in shared memory

struct LevelGeom {
    int w, h, pitch;            // interior size, row pitch in bytes
    int minBX, minBY, maxBX, maxBY;
    int nCols, nRows, wCell, hCell;
    int cellBase;               // first cell of this level inside a frame's cell arrays
    int cellCap;                // slot capacity (keypoints) per cell of this level
    int nFeat;                  // mnFeaturesPerLevel[l]
    int nIni;                   // quadtree roots
    int selCap;                 // survivor slot capacity of this level
    int selBase;                // offset of this level inside a frame's sel array
    int scaledPatch;            // (int)(PATCH_SIZE * scale)
    float hX;                   // root width
    float scale;                // mvScaleFactor[l]
    long long pyrOff;           // byte offset of the level's image stack from d.pyr (level 0 may alias the caller's device frames)
    long long blurOff;          // byte offset of the blurred stack inside d.blur
    long long lvBytes;          // bytes per frame of this level (pitch * h)
    long long cellSlotOff;      // u32 offset of the level's cell slots inside a frame
};

struct FrameSetGeom {
    int nlevels, batch;
    int frame0;                 // first frame this launch processes (chunked pipelining)
    int ncells;                 // cells per frame (all levels)
    int selTotal;               // survivor slots per frame (all levels)
    int outCap;                 // result slots per frame
    int iniTh, minTh;
    long long pyrBytes;         // total bytes of pyr (and of blur) at batch capacity
    long long cellSlots;        // u32 slots per frame in cellk
    LevelGeom lv[kMaxLevels];
};

// device pointers of one extractor handle
struct DevBufs {
    uint8_t* pyr;
    uint8_t* blur;
    int* cellc;
    uint32_t* cellk;
    uint32_t* sel;
    int* selc;
    orbfe_keypoint* kps;
    uint8_t* desc;
    int* counts;
    // resize tables, per level (level 0 entries unused)
    const int2* xtab[kMaxLevels];   // per dst x: {sx, a0 | a1<<16}
    const int4* ytab[kMaxLevels];   // per dst y: {row0, row1, b0, b1}
    const int4* xq[kMaxLevels];     // per 4 dst columns: {base, selA01, selB01, selA23}{selB23, c0..c2}{c3,-,-,-}; null = use xtab
    const void* tmaps;              // HOST array CUtensorMap[2 * nlevels]: {level, blurred level} as [frames * h][pitch] u8 tensors
    const int4* celltab;            // per FAST cell: {iniX | iniY<<16, ww | wh<<16 (0 = degenerate), offX | offY<<16, level | direct<<4 | mdiv<<8}
    const void* fast_tmaps;         // HOST array CUtensorMap[nlevels]: level as [frames * h][pitch] u8, box 64 x window rows
    const void* fast_tmaps48;       // the same with a 48-byte box (cells whose window fits it skip the re-layout)
    bool xqSimple[kMaxLevels];      // row table strictly down-scaling: row1 == row0 + 1 and strictly increasing
    // quadtree point scratch for levels whose corner count exceeds the shared-memory arrays
    uint32_t* qt_xy;                // [B][L][qtCap]
    uint32_t* qt_node;              // [B][L][qtCap]
    long long qtCap;
};

__host__ __device__ inline uint32_t pack_kp(int x, int y, int score) {
    return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)score << 24);
}
__host__ __device__ inline int kp_x(uint32_t p) { return (int)(p & 0xFFFu); }
__host__ __device__ inline int kp_y(uint32_t p) { return (int)((p >> 12) & 0xFFFu); }
__host__ __device__ inline int kp_score(uint32_t p) { return (int)(p >> 24); }

// kernel launchers (each returns the number of kernel launches it enqueued)
// [l0, l1) restricts a launcher to a range of pyramid levels (l1 < 0: up to the last level): the single-frame
// schedule forks per level, so that FAST / quadtree / blur of level l start as soon as level l exists
int launch_pyramid(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0 = 1, int l1 = -1);
int launch_blur(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0 = 0, int l1 = -1);
int fast_window_rows(const LevelGeom& L);
int launch_fast(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0 = 0, int l1 = -1);
int launch_octree(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0 = 0, int l1 = -1);
int launch_orient_desc(const FrameSetGeom& g, const DevBufs& d, const int* d_umax, cudaStream_t s);
size_t octree_smem_bytes(const FrameSetGeom& g);
int octree_prepare(const FrameSetGeom& g);   // cudaFuncSetAttribute; returns 0 / cudaError

}  // namespace orbfe
