// cvshim.hpp — ORACLE support (test infrastructure only).
//
// A minimal stand-in for the slice of the OpenCV C++ API that the reference translation unit
// src/ORBextractor.cc (+ include/ORBextractor.h) touches, so that the UNMODIFIED reference TU
// can be compiled here where OpenCV C++ headers do not exist.  The image
// primitives forward to the oracle restatements in orc_primitives.cpp (pinned to cv2 4.13.0).
// The same shim carries the CV_32F matrix algebra (row/col/t/dot/norm, +, -, *, /, Mat_<float> << ...,
// undistortPoints forwarded to orc_frame.cpp) that src/ORBmatcher.cc, src/Frame.cc, src/KeyFrame.cc and
// src/MapPoint.cc use, so those TUs compile unmodified as well (oracle/ref_matcher_wrap.cpp).  Matrix
// products accumulate in double like cv::gemm on CV_32F, but each operator rounds to float on its own
// (OpenCV fuses A*B+C into one gemm): identical whenever the products are exact, which the pinning
// scenarios (axis-aligned poses) guarantee; general pose arithmetic is outside the oracle's scope.
// Semantics that matter and are reproduced:
//   * Mat is a ref-counted header over a buffer; operator()(Rect)/rowRange/colRange are views;
//   * `m = Mat::zeros(r,c,t)` writes INTO an existing same-shape matrix (MatExpr assignment),
//     which computeDescriptors (ORBextractor.cc:1106) relies on to fill the caller's rows;
//   * Point2i(float,int) truncates (implicit conversion), Point2f *= float is a float multiply;
//   * cvRound is round-half-even, cvFloor/cvCeil as in OpenCV.
#ifndef CVSHIM_HPP
#define CVSHIM_HPP

#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "../orb_oracle.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_64F 6
#define CV_PI 3.1415926535897932384626433832795

typedef unsigned char uchar;

static inline int cvRound(double v) { return orc_cv_round_d(v); }
static inline int cvRound(float v) { return orc_cv_round_f(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
             int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct MatStep {
    size_t v;
    MatStep(size_t s = 0) : v(s) {}
    operator size_t() const { return v; }
};

struct MatZerosExpr { int rows, cols, type; };

class Mat {
public:
    int rows, cols;
    uchar* data;
    MatStep step;
    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(CV_8UC1), esz_(1) {}
    Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    Mat(const MatZerosExpr& z) : Mat() { *this = z; }
    // external (non-owning) buffer
    Mat(int r, int c, int type, void* ext, size_t stp) : rows(r), cols(c), data((uchar*)ext), step(stp), type_(type), esz_(elem(type)) {}

    static int elem(int type) { return type == CV_32F ? 4 : type == CV_64F ? 8 : 1; }
    void create(int r, int c, int type) {
        assert(type == CV_8UC1 || type == CV_32F || type == CV_64F);
        if (data && r == rows && c == cols && type == type_) return;
        type_ = type; esz_ = elem(type);
        buf_ = std::shared_ptr<uchar>(new uchar[(size_t)r * c * esz_ + 8], std::default_delete<uchar[]>());
        data = buf_.get(); rows = r; cols = c; step = MatStep((size_t)c * esz_);
    }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = MatStep(0); }
    Mat& operator=(const MatZerosExpr& z) {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) std::memset(data + y * (size_t)step, 0, (size_t)cols * esz_);
        return *this;
    }
    static MatZerosExpr zeros(int r, int c, int type) { return MatZerosExpr{r, c, type}; }
    static Mat eye(int r, int c, int type) { Mat m = zeros(r, c, type); for (int i = 0; i < r && i < c; i++) m.at<float>(i, i) = 1.0f; return m; }
    static Mat ones(int r, int c, int type) { Mat m(r, c, type); for (int y = 0; y < r; y++) for (int x = 0; x < c; x++) m.at<float>(y, x) = 1.0f; return m; }

    Mat operator()(const Rect& r) const {
        Mat m(*this);
        m.data = data + (size_t)r.y * step + (size_t)r.x * esz_; m.rows = r.height; m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * esz_; m.cols = b - a; return m; }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const {
        Mat m; if (!data) return m;
        m.create(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * esz_);
        return m;
    }
    void copyTo(Mat& dst) const {     // writes through an existing same-shape header (a view stays a view), as cv::Mat::copyTo
        if (!data) { dst.release(); return; }
        dst.create(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * esz_);
    }
    void copyTo(Mat&& view) const { copyTo(view); }      // a temporary view: Rwc.copyTo(Twc.rowRange(0,3).colRange(0,3))
    Mat reshape(int) const { return *this; }            // channel count is not modelled: an N x 2 float block either way
    void convertTo(Mat& dst, int type) const {          // 8-bit -> float only (Frame.cc:627-628)
        assert(type == CV_32F && type_ == CV_8UC1); (void)type;
        Mat m(rows, cols, CV_32F);
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) m.at<float>(y, x) = (float)at<uchar>(y, x);
        dst = m;
    }
    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
    // single index: element i of a row or column vector
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
    size_t step1() const { return step; }
    int type() const { return type_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Mat t() const {
        Mat m(cols, rows, CV_32F);
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    double dot(const Mat& o) const {                    // cv::Mat::dot on CV_32F accumulates in double
        double s = 0.0;
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) s += (double)at<float>(y, x) * (double)o.at<float>(y, x);
        return s;
    }

private:
    std::shared_ptr<uchar> buf_;
    int type_, esz_;
};

// CV_32F algebra (each operator materialises its result; see the header comment)
inline Mat operator*(const Mat& a, const Mat& b) {
    assert(a.cols == b.rows);
    Mat m(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            double s = 0.0;
            for (int k = 0; k < a.cols; k++) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
            m.at<float>(y, x) = (float)s;
        }
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b) {
    assert(a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    return m;
}
inline Mat operator-(const Mat& a, const Mat& b) {
    assert(a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    return m;
}
inline Mat operator*(const Mat& a, double s) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)((double)a.at<float>(y, x) * s);
    return m;
}
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)((double)a.at<float>(y, x) / s);
    return m;
}
inline Mat operator-(const Mat& a) { return a * -1.0; }
enum { NORM_L1 = 2, NORM_L2 = 4 };
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }
inline double norm(const Mat& a, const Mat& b, int normType) {
    assert(normType == NORM_L1); (void)normType;
    double s = 0.0;
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) s += std::fabs((double)a.at<float>(y, x) - (double)b.at<float>(y, x));
    return s;
}
// (cv::Mat_<float>(r, c) << a, b, c)
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
};
template <typename T> struct MatCommaInit_ {
    Mat m; int i;
    template <typename V> MatCommaInit_& operator,(V v) { m.at<T>(i / m.cols, i % m.cols) = (T)v; i++; return *this; }
    operator Mat() const { return m; }
};
template <typename T, typename V> inline MatCommaInit_<T> operator<<(const Mat_<T>& m, V v) {
    MatCommaInit_<T> ci{m, 0};
    ci, v;
    return ci;
}
// cv::undistortPoints(src, dst, K, D, R = empty, P = K) in the call shape of Frame::UndistortKeyPoints / ComputeImageBounds
// (src/Frame.cc:463, :493): N x 2 float block in place; forwards to the restatement pinned against cv2 (orc_frame.cpp)
inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P) {
    assert(R.empty() && P.data == K.data && src.cols == 2 && src.type() == CV_32F); (void)R; (void)P;
    const float cam[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    float dist[5] = {D.at<float>(0), D.at<float>(1), D.at<float>(2), D.at<float>(3), D.rows * D.cols > 4 ? D.at<float>(4) : 0.0f};
    std::vector<float> in((size_t)2 * src.rows), out((size_t)2 * src.rows);
    for (int i = 0; i < src.rows; i++) { in[2 * i] = src.at<float>(i, 0); in[2 * i + 1] = src.at<float>(i, 1); }
    orc_undistort_points(in.data(), src.rows, cam, dist, out.data());
    if (dst.data != src.data) dst.create(src.rows, 2, CV_32F);
    for (int i = 0; i < src.rows; i++) { dst.at<float>(i, 0) = out[2 * i]; dst.at<float>(i, 1) = out[2 * i + 1]; }
}
// parse-only: the vocabulary's load / save members are never instantiated
struct FileNode {
    enum { SEQ = 5 };
    int type() const { return 0; }
    size_t size() const { return 0; }
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    template <typename T> operator T() const { return T(); }
};
struct FileStorage {
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode getFirstTopLevelNode() const { return FileNode(); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }


// InputArray / OutputArray: thin proxies over a Mat
class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray(Mat& m) { m_ = &m; }
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16, INTER_LINEAR = 1 };

inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nonmax = true) {
    kps.clear();
    if (img.cols < 7 || img.rows < 7) return;
    std::vector<int> buf((size_t)3 * img.cols * img.rows);
    int n = orc_fast9_16(img.data, img.cols, img.rows, (ptrdiff_t)(size_t)img.step, threshold, nonmax,
                         buf.data(), img.cols * img.rows);
    kps.reserve(n);
    for (int i = 0; i < n; i++)
        kps.push_back(KeyPoint((float)buf[3 * i], (float)buf[3 * i + 1], 7.f, -1, (float)buf[3 * i + 2]));
}

inline void resize(const Mat& src, Mat& dst, Size sz, double, double, int) {
    // dst already has size sz in every reference call (ORBextractor.cc:1219): written in place
    if (dst.rows != sz.height || dst.cols != sz.width) dst.create(sz.height, sz.width, CV_8UC1);
    orc_resize_linear_u8(src.data, src.cols, src.rows, (ptrdiff_t)(size_t)src.step, dst.data, dst.cols,
                         dst.rows, (ptrdiff_t)(size_t)dst.step);
}

inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int) {
    assert(top == bottom && left == right && top == left); (void)bottom; (void)right; (void)left;
    if (dst.rows != src.rows + 2 * top || dst.cols != src.cols + 2 * top)
        dst.create(src.rows + 2 * top, src.cols + 2 * top, CV_8UC1);
    // src may be the interior ROI of dst (ORBextractor.cc:1223): interior is copied onto itself
    std::vector<uchar> tmp((size_t)src.rows * src.cols);
    for (int y = 0; y < src.rows; y++) std::memcpy(&tmp[(size_t)y * src.cols], src.ptr(y), src.cols);
    orc_copy_make_border_reflect101(tmp.data(), src.cols, src.rows, src.cols, dst.data,
                                    (ptrdiff_t)(size_t)dst.step, top);
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sx, double sy, int) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2); (void)ksize; (void)sx; (void)sy;
    std::vector<uchar> tmp((size_t)src.rows * src.cols);
    orc_gaussian_blur_7x7_s2(src.data, src.cols, src.rows, (ptrdiff_t)(size_t)src.step, tmp.data(), src.cols);
    if (dst.rows != src.rows || dst.cols != src.cols) dst.create(src.rows, src.cols, CV_8UC1);
    for (int y = 0; y < src.rows; y++) std::memcpy(dst.ptr(y), &tmp[(size_t)y * src.cols], src.cols);
}

struct KeyPointsFilter {   // only referenced by the dead ComputeKeyPointsOld; must link, never runs
    static void retainBest(std::vector<KeyPoint>& kps, int n) {
        if (n >= 0 && (int)kps.size() > n) kps.resize(n);
    }
};

}  // namespace cv
#endif
