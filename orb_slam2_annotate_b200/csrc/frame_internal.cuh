// frame_internal.cuh — launchers of the Frame epilogue kernels (frame.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/orbfe.h"
#include "orbfe_internal.cuh"

namespace orbfe {

// camera intrinsics widened to double the way cv::undistortPoints does (cvConvert of the float
// matrices), plus the grid geometry of Frame (mnMinX, mnMinY, mfGridElementWidthInv/HeightInv)
struct CamGeom {
    double fx, fy, cx, cy, ifx, ify, k[5];
    int undistort;              // mDistCoef.at<float>(0) != 0 (Frame.cc:445)
    float minX, minY, invW, invH;
};

// one CTA per frame: kps [frames][cap] -> kps_un [frames][cap], cell_end [frames][3072],
// items [frames][cap]
int launch_frame_epilogue(const CamGeom& c, const orbfe_keypoint* kps, const int* counts, int cap, int frame0,
                          int nframes, orbfe_keypoint* kps_un, int* cell_end, unsigned short* items, cudaStream_t s);
int launch_undistort_points(const CamGeom& c, const float2* in, int n, float2* out, cudaStream_t s);


CamGeom cam_geom(const orbfe_camera& cam, const float bounds[4]);   // bounds may be NULL (no grid)
int image_bounds_device(const orbfe_camera& cam, int cols, int rows, void* d_scratch, cudaStream_t s, float bounds[4]);

}  // namespace orbfe
