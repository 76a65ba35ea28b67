// orc_frame.cpp — CPU oracle (TEST INFRASTRUCTURE ONLY) for the Frame constructor's epilogue:
// Frame::UndistortKeyPoints (reference src/Frame.cc:443-475), Frame::ComputeImageBounds
// (:481-512).  (AssignFeaturesToGrid / PosInGrid, :246-267 and :419-430, are in orc_matcher.cpp:
// orc_assign_grid.)
//
// cv::undistortPoints lives in OpenCV (a prebuilt dependency, absent from /root/reference): its
// published algorithm (modules/calib3d/src/undistort.dispatch.cpp, cvUndistortPointsInternal,
// 4.x) is restated here for the reference's call shape — CV_32FC2 points, float K and distCoef
// (k1 k2 p1 p2 [k3]) widened to double, R = identity, P = K, the fixed five iterations of the
// default TermCriteria(MAX_ITER, 5, 0.01) — in double precision without FMA contraction, result
// narrowed to float.  Pinned bit-for-bit against cv2 4.13 vectors (tests/golden/undistort_cv2.npz,
// made by tests/golden/make_golden.py).
#include <algorithm>
#include <cmath>

#include "orb_oracle.h"

namespace {

struct Cam {
    double fx, fy, cx, cy, ifx, ify, k[5];
};

Cam widen(const float cam[4], const float dist[5]) {
    Cam c;
    c.fx = cam[0]; c.fy = cam[1]; c.cx = cam[2]; c.cy = cam[3];
    c.ifx = 1. / c.fx; c.ify = 1. / c.fy;
    for (int i = 0; i < 5; i++) c.k[i] = dist[i];
    return c;
}

void undistort_one(const Cam& c, float px, float py, float* ox, float* oy) {
    const double* k = c.k;
    double x = px, y = py;
    const double u = x, v = y;
    x = (x - c.cx) * c.ifx;
    y = (y - c.cy) * c.ify;
    const double x0 = x, y0 = y;                 // untilt is the identity (no tauX, tauY)
    for (int j = 0; j < 5; j++) {
        const double r2 = x * x + y * y;
        // k4..k6 are zero: the numerator 1 + ((k6*r2 + k5)*r2 + k4)*r2 is exactly 1
        const double icdist = 1. / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
        if (icdist < 0) {                        // regression_14583: give the normalised input back
            x = (u - c.cx) * c.ifx;
            y = (v - c.cy) * c.ify;
            break;
        }
        const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x);   // s1..s4 are zero
        const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    // RR = P * I = K: xx = fx*x + 0*y + cx, ww = 1/(0*x + 0*y + 1) = 1
    const double xx = c.fx * x + 0. * y + c.cx;
    const double yy = 0. * x + c.fy * y + c.cy;
    const double ww = 1. / (0. * x + 0. * y + 1.);
    *ox = (float)(xx * ww);
    *oy = (float)(yy * ww);
}

}  // namespace

extern "C" {

// Frame::UndistortKeyPoints: xy = n (x, y) pairs of mvKeys[i].pt -> mvKeysUn[i].pt
void orc_undistort_points(const float* xy, int n, const float cam[4], const float dist[5], float* out) {
    if (dist[0] == 0.0f) {                       // :445-450 — only k1 is looked at
        std::copy(xy, xy + 2 * n, out);
        return;
    }
    const Cam c = widen(cam, dist);
    for (int i = 0; i < n; i++) undistort_one(c, xy[2 * i], xy[2 * i + 1], out + 2 * i, out + 2 * i + 1);
}

// Frame::ComputeImageBounds: bounds = {mnMinX, mnMaxX, mnMinY, mnMaxY}
void orc_image_bounds(const float cam[4], const float dist[5], int cols, int rows, float bounds[4]) {
    if (dist[0] != 0.0f) {
        const float corners[8] = {0.f, 0.f, (float)cols, 0.f, 0.f, (float)rows, (float)cols, (float)rows};
        float m[8];
        orc_undistort_points(corners, 4, cam, dist, m);
        bounds[0] = std::min(m[0], m[4]);        // :496-499
        bounds[1] = std::max(m[2], m[6]);
        bounds[2] = std::min(m[1], m[3]);
        bounds[3] = std::max(m[5], m[7]);
    } else {
        bounds[0] = 0.0f; bounds[1] = (float)cols; bounds[2] = 0.0f; bounds[3] = (float)rows;
    }
}

}  // extern "C"
