// frame.cu — K11: the Frame constructor's epilogue on the device:
// Frame::UndistortKeyPoints (reference src/Frame.cc:443-475) and Frame::AssignFeaturesToGrid /
// PosInGrid (:246-267, :419-430), one CTA per frame, so that mvKeysUn and mGrid are born in HBM
// next to the descriptors and the guided searches never upload a frame again.
//
// cv::undistortPoints (OpenCV calib3d, cvUndistortPointsInternal) is evaluated the way the
// prebuilt library does for the reference's call (float K / distCoef widened to double, R = I,
// P = K, five fixed iterations): every double operation is an explicit round-to-nearest
// intrinsic so that the compiler cannot contract a*b+c into an FMA, which the x86-64 baseline
// build of OpenCV does not do either; the result is narrowed to float once.
//
// mGrid[ix][iy] becomes a CSR over cells c = ix*48 + iy (GetFeaturesInArea walks ix outer, iy
// inner, so a query window is a run of consecutive cells per column): cell_end[c] is one past
// the last item of cell c, items hold keypoint indices in push_back order (ascending).
#include "api_common.cuh"
#include "frame_internal.cuh"
#include "grid_build.cuh"

namespace orbfe {

namespace {

constexpr int FT = kGridThreads;        // threads per frame
constexpr int NCELLS = kGridCells;

__device__ __forceinline__ void undistort_one(const CamGeom& c, float px, float py, float& ox, float& oy) {
    const double u = (double)px, v = (double)py;
    double x = __dmul_rn(__dsub_rn(u, c.cx), c.ifx);
    double y = __dmul_rn(__dsub_rn(v, c.cy), c.ify);
    const double x0 = x, y0 = y;
    const double k0 = c.k[0], k1 = c.k[1], k2 = c.k[2], k3 = c.k[3], k4 = c.k[4];
#pragma unroll 1
    for (int j = 0; j < 5; j++) {
        const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k4, r2), k1), r2), k0), r2));
        const double icdist = __ddiv_rn(1.0, den);
        if (icdist < 0) {                 // give the normalised input back
            x = x0; y = y0;
            break;
        }
        const double dX = __dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, k2), x), y),
                                    __dmul_rn(k3, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x))));
        const double dY = __dadd_rn(__dmul_rn(k2, __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                    __dmul_rn(__dmul_rn(__dmul_rn(2.0, k3), x), y));
        x = __dmul_rn(__dsub_rn(x0, dX), icdist);
        y = __dmul_rn(__dsub_rn(y0, dY), icdist);
    }
    // RR = K * I: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1 / (0*x + 0*y + 1)
    const double zx = __dmul_rn(0.0, x), zy = __dmul_rn(0.0, y);
    const double xx = __dadd_rn(__dadd_rn(__dmul_rn(c.fx, x), zy), c.cx);
    const double yy = __dadd_rn(__dadd_rn(zx, __dmul_rn(c.fy, y)), c.cy);
    const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(zx, zy), 1.0));
    ox = __double2float_rn(__dmul_rn(xx, ww));
    oy = __double2float_rn(__dmul_rn(yy, ww));
}

// PosInGrid (Frame.cc:419-430): round() of a float argument, half away from zero
__device__ __forceinline__ int cell_of(const CamGeom& c, float x, float y) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(x, c.minX), c.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(y, c.minY), c.invH));
    if (px < 0 || px >= 64 || py < 0 || py >= 48) return -1;
    return px * 48 + py;
}

__global__ void __launch_bounds__(FT)
frame_epilogue_kernel(const __grid_constant__ CamGeom c, const orbfe_keypoint* __restrict__ kps,
                      const int* __restrict__ counts, int cap, int frame0, orbfe_keypoint* __restrict__ kps_un,
                      int* __restrict__ cell_end, unsigned short* __restrict__ items) {
    __shared__ int cnt[NCELLS];
    __shared__ int wsum[FT / 32];
    const int f = frame0 + blockIdx.x, tid = threadIdx.x;
    const int n = min(counts[f], cap);
    const orbfe_keypoint* src = kps + (long long)f * cap;
    orbfe_keypoint* dst = kps_un + (long long)f * cap;
    int* ce = cell_end + (long long)f * NCELLS;
    unsigned short* it = items + (long long)f * cap;

    // mvKeysUn[i] = mvKeys[i] with the undistorted position (:466-473)
    for (int i = tid; i < n; i += FT) {
        orbfe_keypoint kp = src[i];
        if (c.undistort) undistort_one(c, kp.x, kp.y, kp.x, kp.y);
        dst[i] = kp;
    }
    __syncthreads();
    block_grid_csr(cnt, wsum, n, [&](int i) { return cell_of(c, dst[i].x, dst[i].y); }, ce, it);
}

__global__ void undistort_points_kernel(const __grid_constant__ CamGeom c, const float2* __restrict__ in, int n,
                                        float2* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2 p = in[i];
    if (c.undistort) undistort_one(c, p.x, p.y, p.x, p.y);
    out[i] = p;
}

}  // namespace

int launch_frame_epilogue(const CamGeom& c, const orbfe_keypoint* kps, const int* counts, int cap, int frame0,
                          int nframes, orbfe_keypoint* kps_un, int* cell_end, unsigned short* items, cudaStream_t s) {
    if (nframes <= 0) return 0;
    frame_epilogue_kernel<<<nframes, FT, 0, s>>>(c, kps, counts, cap, frame0, kps_un, cell_end, items);
    return 1;
}

int launch_undistort_points(const CamGeom& c, const float2* in, int n, float2* out, cudaStream_t s) {
    if (n <= 0) return 0;
    undistort_points_kernel<<<(n + 127) / 128, 128, 0, s>>>(c, in, n, out);
    return 1;
}

CamGeom cam_geom(const orbfe_camera& cam, const float bounds[4]) {
    CamGeom c{};
    c.fx = cam.fx; c.fy = cam.fy; c.cx = cam.cx; c.cy = cam.cy;       // cvConvert(float K -> double)
    c.ifx = 1. / c.fx; c.ify = 1. / c.fy;
    c.k[0] = cam.k1; c.k[1] = cam.k2; c.k[2] = cam.p1; c.k[3] = cam.p2; c.k[4] = cam.k3;
    c.undistort = cam.k1 != 0.0f;                                     // Frame.cc:445
    if (bounds) {
        c.minX = bounds[0]; c.minY = bounds[2];
        c.invW = 64.0f / (bounds[1] - bounds[0]);                     // Frame.cc:101-102
        c.invH = 48.0f / (bounds[3] - bounds[2]);
    }
    return c;
}

// Frame::ComputeImageBounds (Frame.cc:481-512); the corners take the device path of the keypoints.
// d_scratch: at least 64 bytes of device memory.  Synchronises the stream.
int image_bounds_device(const orbfe_camera& cam, int cols, int rows, void* d_scratch, cudaStream_t s, float bounds[4]) {
    if (cam.k1 == 0.0f) {
        bounds[0] = 0.0f; bounds[1] = (float)cols; bounds[2] = 0.0f; bounds[3] = (float)rows;
        return 0;
    }
    const CamGeom c = cam_geom(cam, nullptr);
    const float corners[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
    float m[8];
    float2* d_in = reinterpret_cast<float2*>(d_scratch);
    float2* d_out = d_in + 4;
    CU(cudaMemcpyAsync(d_in, corners, sizeof(corners), cudaMemcpyHostToDevice, s));
    launch_undistort_points(c, d_in, 4, d_out, s);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(m, d_out, sizeof(m), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    bounds[0] = std::min(m[0], m[4]);            // :496-499
    bounds[1] = std::max(m[2], m[6]);
    bounds[2] = std::min(m[1], m[3]);
    bounds[3] = std::max(m[5], m[7]);
    return 0;
}

}  // namespace orbfe
