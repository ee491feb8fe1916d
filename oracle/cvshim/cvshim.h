// oracle/cvshim/cvshim.h -- TEST INFRASTRUCTURE ONLY.
//
// A minimal `cv::` facade that provides exactly the OpenCV surface used by the reference files
// this repo compiles VERBATIM from /root/reference (see oracle/Makefile) without OpenCV C++
// being installed:
//   src/ORBextractor.cc                                   (liborbref.so and liborbmref.so)
//   src/ORBmatcher.cc, src/Frame.cc, src/MapPoint.cc,
//   src/CameraModels/Pinhole.cpp, Thirdparty/DBoW2/DBoW2/*  (liborbmref.so)
// All pixel arithmetic is delegated to oracle/cvprims.h, which is pinned bit-for-bit against
// cv2 4.13 by tests/test_oracle_prims.py.  The small float matrix algebra (gemm, transpose,
// 3x3 inverse, norm, dot, undistortPoints, knnMatch) restates OpenCV's arithmetic (evaluation
// order and accumulator width) and is pinned against cv2 by tests/test_oracle_cvshim.py through
// the C hooks in oracle/ref_match_capi.cpp.  This is not a general OpenCV replacement.
#ifndef ORACLE_CVSHIM_H
#define ORACLE_CVSHIM_H

#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <iostream>
#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "../cvprims.h"

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> 3) & 511) + 1)
#define CV_MAKETYPE(depth, cn) (CV_MAT_DEPTH(depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_16SC1 CV_MAKETYPE(CV_16S, 1)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

typedef unsigned char uchar;

static inline int cvRound(double v) { return cvprims::cv_round(v); }
static inline int cvRound(float v) { return cvprims::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return cvprims::cv_floor(v); }
static inline int cvCeil(double v) { return cvprims::cv_ceil(v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };
enum { GEMM_1_T = 1, GEMM_2_T = 2 };

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) {
        x = (T)(x * s);
        y = (T)(y * s);
        return *this;
    }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
template <typename T>
static inline Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) {
    return Point_<T>(a.x - b.x, a.y - b.y);
}
template <typename T>
static inline Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) {
    return Point_<T>(a.x + b.x, a.y + b.y);
}

template <typename T>
struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};
struct Range {
    int start, end;
    Range() : start(0), end(0) {}
    Range(int s, int e) : start(s), end(e) {}
};
struct Scalar {
    double val[4];
    Scalar(double v0 = 0, double v1 = 0, double v2 = 0, double v3 = 0) {
        val[0] = v0; val[1] = v1; val[2] = v2; val[3] = v3;
    }
};

struct KeyPoint {
    Point2f pt;
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0,
             int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave),
          class_id(_class_id) {}
};

struct DMatch {
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(std::numeric_limits<float>::max()) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
};

static inline size_t cvshim_depth_size(int depth) {
    static const size_t s[8] = {1, 1, 2, 2, 4, 4, 8, 2};
    return s[depth & 7];
}

class MatExpr;

// Dense 2-D matrix with shared ownership and ROI views.  `step` is in bytes.
class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;
    int mtype;
    std::shared_ptr<uchar> owner;

    Mat() : rows(0), cols(0), data(nullptr), step(0), mtype(CV_8UC1) {}
    Mat(int r, int c, int type) : mtype(type) { alloc(r, c, type); }
    Mat(Size sz, int type) : mtype(type) { alloc(sz.height, sz.width, type); }
    // external data, not owned
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0)
        : rows(r), cols(c), data((uchar*)ext), step(ext_step), mtype(type) {
        if (!step) step = (size_t)c * elemSize();
    }
    Mat(const Mat&) = default;
    Mat(const MatExpr& e);
    Mat& operator=(const Mat&) = default;
    // OpenCV semantics: an expression assigned to a matrix of the same size and type is
    // evaluated INTO its buffer (this is what makes `M.rowRange(..).colRange(..) = A*B` work)
    Mat& operator=(const MatExpr& e);

    size_t elemSize() const { return cvshim_depth_size(CV_MAT_DEPTH(mtype)) * CV_MAT_CN(mtype); }
    size_t elemSize1() const { return cvshim_depth_size(CV_MAT_DEPTH(mtype)); }
    int type() const { return mtype; }
    int depth() const { return CV_MAT_DEPTH(mtype); }
    int channels() const { return CV_MAT_CN(mtype); }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }

    void alloc(int r, int c, int type) {
        mtype = type;
        rows = r;
        cols = c;
        step = (size_t)c * elemSize();
        size_t n = (size_t)r * step;
        owner.reset((uchar*)malloc(n ? n : 1), free);
        data = owner.get();
    }
    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == mtype) return;
        alloc(r, c, type);
    }
    void release() {
        owner.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step / elemSize1(); }

    Mat operator()(const Rect& r) const {
        Mat m(*this);
        m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
        m.rows = r.height;
        m.cols = r.width;
        return m;
    }
    Mat operator()(const Range& rr, const Range& cr) const {
        return rowRange(rr.start, rr.end).colRange(cr.start, cr.end);
    }
    Mat rowRange(int a, int b) const {
        Mat m(*this);
        m.data = data + (ptrdiff_t)a * (ptrdiff_t)step;
        m.rows = b - a;
        return m;
    }
    Mat colRange(int a, int b) const {
        Mat m(*this);
        m.data = data + (ptrdiff_t)a * (ptrdiff_t)elemSize();
        m.cols = b - a;
        return m;
    }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const {
        Mat m;
        if (empty()) { m.mtype = mtype; return m; }
        m.alloc(rows, cols, mtype);
        size_t rb = (size_t)cols * elemSize();
        for (int y = 0; y < rows; ++y) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, rb);
        return m;
    }
    // dst a view of the right shape: copied into; otherwise (re)allocated, as OpenCV does
    void copyTo(Mat& dst) const {
        if (!(dst.data && dst.rows == rows && dst.cols == cols && dst.mtype == mtype)) dst.alloc(rows, cols, mtype);
        size_t rb = (size_t)cols * elemSize();
        for (int y = 0; y < rows; ++y) memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, rb);
    }
    void copyTo(Mat&& dst) const {
        assert(dst.rows == rows && dst.cols == cols && dst.mtype == mtype);
        copyTo(dst);
    }
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        memset(m.data, 0, (size_t)r * m.step);
        return m;
    }
    static Mat zeros(Size s, int type) { return zeros(s.height, s.width, type); }
    static Mat eye(int r, int c, int type);
    static Mat ones(int r, int c, int type);

    template <typename T>
    T& at(int y, int x) { return *(T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T>
    const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T>
    T& at(int i) { return *(T*)elem1(i, sizeof(T)); }
    template <typename T>
    const T& at(int i) const { return *(const T*)elem1(i, sizeof(T)); }
    uchar* elem1(int i, size_t esz) const {
        if (rows == 1 || isContinuous()) return data + (size_t)i * esz;
        if (cols == 1) return data + (size_t)i * step;
        return data + (size_t)(i / cols) * step + (size_t)(i % cols) * esz;
    }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T>
    T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T>
    const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }

    // generic element access as double (only used by the small-matrix algebra)
    double getd(int y, int x) const {
        const uchar* p = data + (size_t)y * step + (size_t)x * elemSize1();
        switch (depth()) {
            case CV_8U: return *p;
            case CV_8S: return *(const signed char*)p;
            case CV_16U: return *(const unsigned short*)p;
            case CV_16S: return *(const short*)p;
            case CV_32S: return *(const int*)p;
            case CV_32F: return *(const float*)p;
            default: return *(const double*)p;
        }
    }

    Mat reshape(int cn, int new_rows = 0) const {
        assert(new_rows == 0 && isContinuous());
        (void)new_rows;
        Mat m(*this);
        int total_ch = cols * channels();
        assert(total_ch % cn == 0);
        m.cols = total_ch / cn;
        m.mtype = CV_MAKETYPE(depth(), cn);
        return m;
    }
    void convertTo(Mat& dst, int rtype) const;
    MatExpr t() const;
    MatExpr inv() const;
    double dot(const Mat& o) const;
};

// Lazy expression node, restating the folding rules of OpenCV's MatExpr that decide WHICH gemm
// call (and so which rounding) an expression becomes (modules/core/src/matrix_expressions.cpp):
//   PLAIN  alpha * op(src), op = identity or transpose      (MatOp_AddEx scale / MatOp_T)
//   GEMM   alpha * op(A) * op(B) + beta * C                  (MatOp_GEMM)
// `-R.t()*t` is ONE gemm(R, t, -1, GEMM_1_T); `R*p + t` is ONE gemm(R, p, 1, t, 1).  Everything
// else is evaluated to a matrix first.  Assignment into an existing matrix of the same shape
// writes through (see Mat::operator=).
class MatExpr {
public:
    enum Kind { PLAIN, GEMM };
    Kind kind;
    Mat a, b, c;  // PLAIN: a;  GEMM: a, b, optional c
    bool a_t, b_t;
    double alpha, beta;
    MatExpr() : kind(PLAIN), a_t(false), b_t(false), alpha(1), beta(0) {}
    explicit MatExpr(const Mat& mm) : kind(PLAIN), a(mm), a_t(false), b_t(false), alpha(1), beta(0) {}
    static MatExpr plain(const Mat& m, bool t, double alpha) {
        MatExpr e(m);
        e.a_t = t;
        e.alpha = alpha;
        return e;
    }
    bool is_mat_prod() const { return kind == GEMM && (c.empty() || beta == 0); }
    Mat eval() const;
    MatExpr t() const {
        if (kind == PLAIN) return plain(a, !a_t, alpha);
        return plain(eval(), true, 1);
    }
    MatExpr inv() const { return eval().inv(); }
    double dot(const Mat& o) const { return eval().dot(o); }
    Mat row(int y) const { return eval().row(y); }
    Mat col(int x) const { return eval().col(x); }
    Mat rowRange(int r0, int r1) const { return eval().rowRange(r0, r1); }
    Mat colRange(int c0, int c1) const { return eval().colRange(c0, c1); }
    Mat clone() const { return eval().clone(); }
};
inline Mat::Mat(const MatExpr& e) : Mat(e.eval()) {}
inline Mat& Mat::operator=(const MatExpr& e) {
    Mat r = e.eval();
    if (data && rows == r.rows && cols == r.cols && mtype == r.mtype && data != r.data) {
        Mat src = r.clone();  // the expression may alias this matrix
        src.copyTo(*this);
    } else {
        *this = r;
    }
    return *this;
}

inline Mat Mat::eye(int r, int c, int type) {
    Mat m = zeros(r, c, type);
    for (int i = 0; i < (r < c ? r : c); ++i) {
        if (CV_MAT_DEPTH(type) == CV_32F) m.at<float>(i, i) = 1.f;
        else if (CV_MAT_DEPTH(type) == CV_64F) m.at<double>(i, i) = 1.0;
        else if (CV_MAT_DEPTH(type) == CV_8U) m.at<uchar>(i, i) = 1;
        else abort();
    }
    return m;
}
inline Mat Mat::ones(int r, int c, int type) {
    Mat m(r, c, type);
    for (int y = 0; y < r; ++y)
        for (int x = 0; x < c; ++x) {
            if (CV_MAT_DEPTH(type) == CV_32F) m.at<float>(y, x) = 1.f;
            else if (CV_MAT_DEPTH(type) == CV_64F) m.at<double>(y, x) = 1.0;
            else if (CV_MAT_DEPTH(type) == CV_8U) m.at<uchar>(y, x) = 1;
            else abort();
        }
    return m;
}

// Mat::convertTo without scaling: saturate_cast<rtype>(src) per element.  A different type always
// yields a new buffer (so `roi.convertTo(roi, CV_16S)` detaches the ROI from its parent, as in
// OpenCV: src/Frame.cc:897).
inline void Mat::convertTo(Mat& dst, int rtype) const {
    rtype = CV_MAKETYPE(CV_MAT_DEPTH(rtype), channels());
    if (rtype == mtype) { Mat c = clone(); dst = c; return; }
    Mat out(rows, cols, rtype);
    const int n = cols * channels();
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < n; ++x) {
            double v = getd(y, x);
            uchar* p = out.data + (size_t)y * out.step + (size_t)x * out.elemSize1();
            switch (CV_MAT_DEPTH(rtype)) {
                case CV_8U: { int iv = (depth() >= CV_32F) ? cvRound(v) : (int)v; *p = (uchar)(iv < 0 ? 0 : iv > 255 ? 255 : iv); break; }
                case CV_16S: { int iv = (depth() >= CV_32F) ? cvRound(v) : (int)v; *(short*)p = (short)(iv < -32768 ? -32768 : iv > 32767 ? 32767 : iv); break; }
                case CV_32S: *(int*)p = (depth() >= CV_32F) ? cvRound(v) : (int)v; break;
                case CV_32F: *(float*)p = (float)v; break;
                case CV_64F: *(double*)p = v; break;
                default: abort();
            }
        }
    dst = out;
}

template <typename T>
struct DataDepth;
template <> struct DataDepth<float> { enum { value = CV_32F }; };
template <> struct DataDepth<double> { enum { value = CV_64F }; };
template <> struct DataDepth<uchar> { enum { value = CV_8U }; };
template <> struct DataDepth<int> { enum { value = CV_32S }; };
template <> struct DataDepth<short> { enum { value = CV_16S }; };

template <typename T>
class Mat_;
// (cv::Mat_<float>(3,1) << x, y, z)
template <typename T>
class MatCommaInitializer_ {
public:
    Mat m;
    int idx;
    explicit MatCommaInitializer_(const Mat& mm) : m(mm), idx(0) {}
    template <typename V>
    MatCommaInitializer_& operator,(V v) {
        assert(idx < (int)m.total());
        m.at<T>(idx / m.cols, idx % m.cols) = (T)v;
        ++idx;
        return *this;
    }
    operator Mat() const { return m; }
    operator Mat_<T>() const;
};
template <typename T>
class Mat_ : public Mat {
public:
    Mat_() : Mat() { mtype = CV_MAKETYPE(DataDepth<T>::value, 1); }
    Mat_(int r, int c) : Mat(r, c, CV_MAKETYPE(DataDepth<T>::value, 1)) {}
    Mat_(const Mat& m) : Mat(m) { assert(m.empty() || m.depth() == DataDepth<T>::value); }
    T& operator()(int y, int x) { return this->template at<T>(y, x); }
    const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
    T& operator()(int i) { return this->template at<T>(i); }
};
template <typename T>
MatCommaInitializer_<T>::operator Mat_<T>() const { return Mat_<T>(m); }
template <typename T, typename V>
static inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, V v) {
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}

// ---- small-matrix float algebra, restating OpenCV's arithmetic --------------------------------
namespace shim {

// cv::gemm(A, B, alpha, noArray(), 0, D, flags) for CV_32F / CV_64F (modules/core/src/matmul.simd.hpp).
//  * flags == 0 and 2 <= inner <= 4 and (inner == D.cols or inner == D.rows): the unrolled
//    small-matrix path -- products and sums in the ELEMENT type, left to right, then
//    (T)(t*alpha) with alpha a double;
//  * otherwise GEMMSingleMul: accumulation in double for both float and double,
//    d = T(s*alpha).
inline Mat gemm(const Mat& A0, const Mat& B0, double alpha, int flags, const Mat* C = nullptr, double beta = 0) {
    assert(A0.mtype == B0.mtype && (A0.mtype == CV_32FC1 || A0.mtype == CV_64FC1));
    const bool at = flags & GEMM_1_T, bt = flags & GEMM_2_T;
    const int M = at ? A0.cols : A0.rows, K = at ? A0.rows : A0.cols;
    const int Kb = bt ? B0.cols : B0.rows, N = bt ? B0.rows : B0.cols;
    assert(K == Kb);
    (void)Kb;
    if (C && beta == 0) C = nullptr;
    // Pinned against cv2 (tests/test_oracle_cvshim.py) for the scale factors the compiled reference files
    // produce: alpha = +-1 (or a power of two), beta in {0, +-1}.  cv2's rounding for other factors is not
    // reproduced here, so they are refused instead of silently computed differently.
    {
        int e;
        const double fa = std::frexp(std::fabs(alpha), &e);
        if (fa != 0.5 || !(beta == 0 || beta == 1 || beta == -1 || !C)) {
            fprintf(stderr, "cvshim: gemm with alpha = %g, beta = %g is not pinned against OpenCV\n", alpha, beta);
            abort();
        }
    }
    assert(!C || (C->rows == M && C->cols == N && C->mtype == A0.mtype));
    Mat D(M, N, A0.mtype);
    const bool small = flags == 0 && K >= 2 && K <= 4 && (K == N || K == M);
    if (A0.mtype == CV_32FC1) {
        for (int i = 0; i < M; ++i)
            for (int j = 0; j < N; ++j) {
                const float cv = C ? C->at<float>(i, j) : 0.f;
                if (small) {
                    float t = A0.at<float>(i, 0) * B0.at<float>(0, j);
                    for (int k = 1; k < K; ++k) t = t + A0.at<float>(i, k) * B0.at<float>(k, j);
                    D.at<float>(i, j) = (float)((double)t * alpha + (double)cv * beta);
                } else {
                    double s = 0;
                    for (int k = 0; k < K; ++k) {
                        float a = at ? A0.at<float>(k, i) : A0.at<float>(i, k);
                        float b = bt ? B0.at<float>(j, k) : B0.at<float>(k, j);
                        s += (double)a * (double)b;
                    }
                    D.at<float>(i, j) = C ? (float)(s * alpha + (double)cv * beta) : (float)(s * alpha);
                }
            }
    } else {
        for (int i = 0; i < M; ++i)
            for (int j = 0; j < N; ++j) {
                const double cv = C ? C->at<double>(i, j) : 0.0;
                double s = 0;
                if (small) {
                    s = A0.at<double>(i, 0) * B0.at<double>(0, j);
                    for (int k = 1; k < K; ++k) s = s + A0.at<double>(i, k) * B0.at<double>(k, j);
                } else {
                    for (int k = 0; k < K; ++k) {
                        double a = at ? A0.at<double>(k, i) : A0.at<double>(i, k);
                        double b = bt ? B0.at<double>(j, k) : B0.at<double>(k, j);
                        s += a * b;
                    }
                }
                D.at<double>(i, j) = s * alpha + cv * beta;
            }
    }
    return D;
}

inline Mat transpose(const Mat& a) {
    Mat d(a.cols, a.rows, a.mtype);
    const size_t esz = a.elemSize();
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x)
            memcpy(d.data + (size_t)x * d.step + (size_t)y * esz, a.data + (size_t)y * a.step + (size_t)x * esz, esz);
    return d;
}

// a*alpha + b*beta + gamma elementwise.  OpenCV evaluates a float MatExpr of this family as
// addWeighted / convertTo / add / subtract; for CV_32F: pure a+b and a-b are float adds,
// scaling alone is src*(float)alpha (cvtScale with float coefficients).
inline Mat lincomb(const Mat& a, double alpha, const Mat* b, double beta, double gamma) {
    Mat d(a.rows, a.cols, a.mtype);
    assert(!b || (b->rows == a.rows && b->cols == a.cols && b->mtype == a.mtype));
    const int n = a.cols * a.channels();
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < n; ++x) {
            switch (a.depth()) {
                case CV_32F: {
                    float va = ((const float*)(a.data + (size_t)y * a.step))[x];
                    float r;
                    if (b) {
                        float vb = ((const float*)(b->data + (size_t)y * b->step))[x];
                        if (alpha == 1 && beta == 1) r = va + vb;
                        else if (alpha == 1 && beta == -1) r = va - vb;
                        else r = va * (float)alpha + vb * (float)beta;
                        if (gamma != 0) r = r + (float)gamma;
                    } else {
                        if (gamma == 0) r = (alpha == 1) ? va : (alpha == -1) ? -va : va * (float)alpha;
                        else if (alpha == 1) r = (float)((double)va + gamma);
                        else r = va * (float)alpha + (float)gamma;
                    }
                    ((float*)(d.data + (size_t)y * d.step))[x] = r;
                    break;
                }
                case CV_64F: {
                    double va = ((const double*)(a.data + (size_t)y * a.step))[x];
                    double vb = b ? ((const double*)(b->data + (size_t)y * b->step))[x] : 0;
                    ((double*)(d.data + (size_t)y * d.step))[x] = va * alpha + vb * beta + gamma;
                    break;
                }
                case CV_16S: {  // saturating short arithmetic (src/Frame.cc:898: IL - IL.at<short>(w,w))
                    int va = ((const short*)(a.data + (size_t)y * a.step))[x];
                    int vb = b ? ((const short*)(b->data + (size_t)y * b->step))[x] : 0;
                    assert((alpha == 1 || alpha == -1) && (beta == 1 || beta == -1 || !b));
                    int r = va * (int)alpha + vb * (int)beta + cvRound(gamma);
                    ((short*)(d.data + (size_t)y * d.step))[x] = (short)(r < -32768 ? -32768 : r > 32767 ? 32767 : r);
                    break;
                }
                default: abort();
            }
        }
    return d;
}

// cv::invert(DECOMP_LU) for 2x2 / 3x3 (modules/core/src/lapack.cpp): closed form with the
// determinant and the cofactors in double.
inline Mat invert_small(const Mat& a) {
    assert(a.rows == a.cols && (a.rows == 2 || a.rows == 3) && (a.mtype == CV_32FC1 || a.mtype == CV_64FC1));
    Mat d = Mat::zeros(a.rows, a.cols, a.mtype);
    auto S = [&](int y, int x) { return a.getd(y, x); };
    auto put = [&](int y, int x, double v) {
        if (a.mtype == CV_32FC1) d.at<float>(y, x) = (float)v;
        else d.at<double>(y, x) = v;
    };
    if (a.rows == 2) {
        double det = S(0, 0) * S(1, 1) - S(0, 1) * S(1, 0);
        if (det != 0) {
            det = 1. / det;
            double t0 = S(0, 0) * det, t1 = S(1, 1) * det;
            put(1, 1, t0);
            put(0, 0, t1);
            t0 = -S(0, 1) * det;
            t1 = -S(1, 0) * det;
            put(0, 1, t0);
            put(1, 0, t1);
        }
    } else {
        double det = S(0, 0) * (S(1, 1) * S(2, 2) - S(1, 2) * S(2, 1)) -
                     S(0, 1) * (S(1, 0) * S(2, 2) - S(1, 2) * S(2, 0)) +
                     S(0, 2) * (S(1, 0) * S(2, 1) - S(1, 1) * S(2, 0));
        if (det != 0) {
            det = 1. / det;
            double t[9];
            t[0] = (S(1, 1) * S(2, 2) - S(1, 2) * S(2, 1)) * det;
            t[1] = (S(0, 2) * S(2, 1) - S(0, 1) * S(2, 2)) * det;
            t[2] = (S(0, 1) * S(1, 2) - S(0, 2) * S(1, 1)) * det;
            t[3] = (S(1, 2) * S(2, 0) - S(1, 0) * S(2, 2)) * det;
            t[4] = (S(0, 0) * S(2, 2) - S(0, 2) * S(2, 0)) * det;
            t[5] = (S(0, 2) * S(1, 0) - S(0, 0) * S(1, 2)) * det;
            t[6] = (S(1, 0) * S(2, 1) - S(1, 1) * S(2, 0)) * det;
            t[7] = (S(0, 1) * S(2, 0) - S(0, 0) * S(2, 1)) * det;
            t[8] = (S(0, 0) * S(1, 1) - S(0, 1) * S(1, 0)) * det;
            for (int i = 0; i < 9; ++i) put(i / 3, i % 3, t[i]);
        }
    }
    return d;
}

}  // namespace shim

inline MatExpr Mat::t() const { return MatExpr::plain(*this, true, 1); }
inline MatExpr Mat::inv() const { return MatExpr(shim::invert_small(*this)); }
// Mat::dot: accumulation in double for float data (dotProd_ generic path, lengths < one SIMD block)
inline double Mat::dot(const Mat& o) const {
    assert(rows == o.rows && cols == o.cols && mtype == o.mtype);
    double r = 0;
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols * channels(); ++x) {
            if (depth() == CV_32F) r += (double)ptr<float>(y)[x] * o.ptr<float>(y)[x];
            else if (depth() == CV_64F) r += ptr<double>(y)[x] * o.ptr<double>(y)[x];
            else abort();
        }
    return r;
}

inline Mat MatExpr::eval() const {
    if (kind == PLAIN) {
        if (!a_t) return alpha == 1 ? a : shim::lincomb(a, alpha, nullptr, 0, 0);
        Mat t = shim::transpose(a);  // MatOp_T::assign: transpose, then convertTo(alpha)
        return alpha == 1 ? t : shim::lincomb(t, alpha, nullptr, 0, 0);
    }
    return shim::gemm(a, b, alpha, (a_t ? GEMM_1_T : 0) | (b_t ? GEMM_2_T : 0), c.empty() ? nullptr : &c, beta);
}

// MatOp::multiply / MatOp_AddEx::multiply / MatOp_T::multiply: a product of two scaled and/or
// transposed matrices is one gemm; any other operand is evaluated first.
static inline MatExpr operator*(const MatExpr& x, const MatExpr& y) {
    MatExpr e;
    e.kind = MatExpr::GEMM;
    if (x.kind == MatExpr::PLAIN) { e.a = x.a; e.a_t = x.a_t; e.alpha = x.alpha; }
    else { e.a = x.eval(); e.a_t = false; e.alpha = 1; }
    if (y.kind == MatExpr::PLAIN) { e.b = y.a; e.b_t = y.a_t; e.alpha *= y.alpha; }
    else { e.b = y.eval(); e.b_t = false; }
    return e;
}
static inline MatExpr operator*(const Mat& a, const Mat& b) { return MatExpr(a) * MatExpr(b); }
static inline MatExpr operator*(const MatExpr& a, const Mat& b) { return a * MatExpr(b); }
static inline MatExpr operator*(const Mat& a, const MatExpr& b) { return MatExpr(a) * b; }
// scaling: PLAIN and GEMM expressions fold the factor into alpha (and beta)
static inline MatExpr operator*(const MatExpr& x, double s) {
    MatExpr e = x;
    e.alpha *= s;
    e.beta *= s;
    return e;
}
static inline MatExpr operator*(const Mat& a, double s) { return MatExpr::plain(a, false, s); }
static inline MatExpr operator*(double s, const Mat& a) { return a * s; }
static inline MatExpr operator*(double s, const MatExpr& a) { return a * s; }
static inline MatExpr operator/(const Mat& a, double s) { return a * (1. / s); }
static inline MatExpr operator/(const MatExpr& a, double s) { return a * (1. / s); }
static inline MatExpr operator-(const Mat& a) { return a * -1.0; }
static inline MatExpr operator-(const MatExpr& a) { return a * -1.0; }

// MatOp_GEMM::add / subtract: product (without C) +- a scaled, untransposed matrix becomes the
// product's C operand; otherwise both sides are evaluated and combined elementwise
// (MatOp_AddEx: add / subtract for unit factors, addWeighted otherwise).
static inline MatExpr cvshim_addsub(const MatExpr& x, const MatExpr& y, double sign) {
    if (x.is_mat_prod() && y.kind == MatExpr::PLAIN && !y.a_t) {
        MatExpr e = x;
        e.c = y.a;
        e.beta = sign * y.alpha;
        return e;
    }
    if (y.is_mat_prod() && x.kind == MatExpr::PLAIN && !x.a_t) {
        MatExpr e = y;
        e.alpha *= sign;
        e.c = x.a;
        e.beta = x.alpha;
        return e;
    }
    Mat xa, ya;
    double al = 1, be = sign;
    if (x.kind == MatExpr::PLAIN && !x.a_t) { xa = x.a; al = x.alpha; } else xa = x.eval();
    if (y.kind == MatExpr::PLAIN && !y.a_t) { ya = y.a; be = sign * y.alpha; } else ya = y.eval();
    return MatExpr(shim::lincomb(xa, al, &ya, be, 0));
}
static inline MatExpr operator+(const MatExpr& a, const MatExpr& b) { return cvshim_addsub(a, b, 1); }
static inline MatExpr operator+(const Mat& a, const Mat& b) { return MatExpr(a) + MatExpr(b); }
static inline MatExpr operator+(const MatExpr& a, const Mat& b) { return a + MatExpr(b); }
static inline MatExpr operator+(const Mat& a, const MatExpr& b) { return MatExpr(a) + b; }
static inline MatExpr operator-(const MatExpr& a, const MatExpr& b) { return cvshim_addsub(a, b, -1); }
static inline MatExpr operator-(const Mat& a, const Mat& b) { return MatExpr(a) - MatExpr(b); }
static inline MatExpr operator-(const MatExpr& a, const Mat& b) { return a - MatExpr(b); }
static inline MatExpr operator-(const Mat& a, const MatExpr& b) { return MatExpr(a) - b; }
// Mat - scalar (src/Frame.cc:898, :920)
static inline MatExpr operator-(const Mat& a, const Scalar& s) { return MatExpr(shim::lincomb(a, 1, nullptr, 0, -s.val[0])); }
static inline MatExpr operator-(const Mat& a, short s) { return a - Scalar((double)s); }

// cv::norm(src): NORM_L2, float data, sum of squares accumulated in double
static inline double norm(const Mat& a, int normType = NORM_L2) {
    assert(normType == NORM_L2 && (a.depth() == CV_32F || a.depth() == CV_64F));
    (void)normType;
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols * a.channels(); ++x) {
            double v = a.depth() == CV_32F ? (double)a.ptr<float>(y)[x] : a.ptr<double>(y)[x];
            s += v * v;
        }
    return std::sqrt(s);
}
static inline double norm(const MatExpr& a, int normType = NORM_L2) { return norm(a.eval(), normType); }
// cv::norm(a, b, NORM_L1) on CV_16S (src/Frame.cc:924): exact integer sum
static inline double norm(const Mat& a, const Mat& b, int normType) {
    assert(normType == NORM_L1 && a.mtype == CV_16SC1 && b.mtype == CV_16SC1 && a.rows == b.rows && a.cols == b.cols);
    (void)normType;
    long long s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::abs((int)a.ptr<short>(y)[x] - (int)b.ptr<short>(y)[x]);
    return (double)s;
}

static inline void hconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.rows == b.rows && a.mtype == b.mtype);
    Mat d(a.rows, a.cols + b.cols, a.mtype);
    a.copyTo(d.colRange(0, a.cols));
    b.copyTo(d.colRange(a.cols, a.cols + b.cols));
    dst = d;
}
static inline void vconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.empty() || b.empty() || (a.cols == b.cols && a.mtype == b.mtype));
    if (a.empty()) { dst = b.clone(); return; }
    if (b.empty()) { dst = a.clone(); return; }
    Mat d(a.rows + b.rows, a.cols, a.mtype);
    a.copyTo(d.rowRange(0, a.rows));
    b.copyTo(d.rowRange(a.rows, a.rows + b.rows));
    dst = d;
}

class _InputArray {
public:
    Mat* m;
    Mat tmp;
    _InputArray(const Mat& mm) : m(const_cast<Mat*>(&mm)) {}
    _InputArray(const MatExpr& e) : m(&tmp), tmp(e.eval()) {}
    Mat getMat() const { return *m; }
    bool empty() const { return m->empty(); }
};
class _OutputArray : public _InputArray {
public:
    _OutputArray(Mat& mm) : _InputArray(mm) {}
    void create(int r, int c, int type) const { m->create(r, c, type); }
    void release() const { m->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

static inline float fastAtan2(float y, float x) { return cvprims::fast_atan2_deg(y, x); }

// cv::undistortPoints(src, dst, cameraMatrix, distCoeffs, R = empty, P) with the default
// TermCriteria(MAX_ITER, 5, 0.01) (modules/calib3d/src/undistort.dispatch.cpp,
// cvUndistortPointsInternal): all arithmetic in double, 5 fixed-point iterations, k1 k2 p1 p2 [k3]
// (the reference passes 4 or 5 coefficients, src/Tracking.cc), then P applied.  src/dst: N x 1
// CV_32FC2 (src/Frame.cc:756-758).
static inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& dist, const Mat& R, const Mat& P) {
    assert(src.mtype == CV_32FC2 && R.empty());
    (void)R;
    double k[14] = {0};
    const int nd = (int)dist.total();
    for (int i = 0; i < nd && i < 14; ++i) k[i] = dist.depth() == CV_32F ? (double)dist.at<float>(i) : dist.at<double>(i);
    const double fx = K.getd(0, 0), fy = K.getd(1, 1), ifx = 1. / fx, ify = 1. / fy, cx = K.getd(0, 2), cy = K.getd(1, 2);
    double RR[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (!P.empty())
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) RR[i][j] = P.getd(i, j);  // PP * I
    Mat out(src.rows, src.cols, CV_32FC2);
    for (int r = 0; r < src.rows; ++r)
        for (int c = 0; c < src.cols; ++c) {
            const float* s = src.ptr<float>(r) + 2 * c;
            double x = s[0], y = s[1];
            const double u = x, v = y;
            x = (x - cx) * ifx;
            y = (y - cy) * ify;
            if (nd) {
                const double x0 = x, y0 = y;
                for (int j = 0; j < 5; ++j) {
                    double r2 = x * x + y * y;
                    double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
                    if (icdist < 0) {
                        x = (u - cx) * ifx;
                        y = (v - cy) * ify;
                        break;
                    }
                    double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
                    double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
                    x = (x0 - deltaX) * icdist;
                    y = (y0 - deltaY) * icdist;
                }
            }
            double xx = RR[0][0] * x + RR[0][1] * y + RR[0][2];
            double yy = RR[1][0] * x + RR[1][1] * y + RR[1][2];
            double ww = 1. / (RR[2][0] * x + RR[2][1] * y + RR[2][2]);
            float* d = out.ptr<float>(r) + 2 * c;
            d[0] = (float)(xx * ww);
            d[1] = (float)(yy * ww);
        }
    dst = out;
}

// cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, matches, k): k smallest distances per query
// row, lowest train index first on ties (pinned against cv2 4.13 by the tests).
class BFMatcher {
public:
    explicit BFMatcher(int normType = NORM_L2, bool crossCheck = false) {
        assert(normType == NORM_HAMMING && !crossCheck);
        (void)normType; (void)crossCheck;
    }
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch> >& matches, int k) const {
        matches.clear();
        if (q.empty() || t.empty()) return;
        assert(q.cols == 32 && t.cols == 32 && q.depth() == CV_8U);
        matches.resize(q.rows);
        std::vector<std::pair<int, int> > best;
        for (int i = 0; i < q.rows; ++i) {
            best.clear();
            const uint64_t* a = (const uint64_t*)q.ptr(i);
            uint64_t av[4];
            memcpy(av, a, 32);
            for (int j = 0; j < t.rows; ++j) {
                uint64_t bv[4];
                memcpy(bv, t.ptr(j), 32);
                int d = __builtin_popcountll(av[0] ^ bv[0]) + __builtin_popcountll(av[1] ^ bv[1]) +
                        __builtin_popcountll(av[2] ^ bv[2]) + __builtin_popcountll(av[3] ^ bv[3]);
                // insertion into the sorted top-k, strict < keeps the lower index first on ties
                int pos = (int)best.size();
                while (pos > 0 && d < best[pos - 1].first) --pos;
                if (pos < k) {
                    best.insert(best.begin() + pos, std::make_pair(d, j));
                    if ((int)best.size() > k) best.pop_back();
                }
            }
            for (size_t n = 0; n < best.size(); ++n) matches[i].push_back(DMatch(i, best[n].second, (float)best[n].first));
        }
    }
};

// Declared only: named by Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h's yml save/load members,
// which are never instantiated here.
class FileNode {
public:
    FileNode operator[](const char*) const { abort(); }
    FileNode operator[](const std::string&) const { abort(); }
    FileNode operator[](int) const { abort(); }
    size_t size() const { abort(); }
    int type() const { abort(); }
    enum { SEQ = 5 };
    operator int() const { abort(); }
    operator double() const { abort(); }
    operator std::string() const { abort(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) { abort(); }
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { abort(); }
    FileNode operator[](const char*) const { abort(); }
    void release() {}
};
template <typename T>
static inline FileStorage& operator<<(FileStorage& fs, const T&) { abort(); return fs; }

// Optional per-thread stage clock (off by default): seconds spent inside the three OpenCV primitives that
// dominate the extractor, measured in situ so bench.py's strawman guard can say what real OpenCV would change.
struct StageClock {
    bool on = false;
    double resize = 0, fast = 0, blur = 0;
};
inline StageClock& stage_clock() {
    static thread_local StageClock c;
    return c;
}
struct StageTimer {
    double* acc;
    timespec t0;
    explicit StageTimer(double* a) : acc(stage_clock().on ? a : nullptr) {
        if (acc) clock_gettime(CLOCK_MONOTONIC, &t0);
    }
    ~StageTimer() {
        if (!acc) return;
        timespec t1;
        clock_gettime(CLOCK_MONOTONIC, &t1);
        *acc += (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    }
};

// Optional per-thread observer of every cv::FAST call (used by oracle/ref_capi.cpp to dump the
// per-cell candidate lists, which are locals of ComputeKeyPointsOctTree in the reference).
struct FastObserver {
    virtual void on_fast(const Mat& img, int threshold, const std::vector<KeyPoint>& kps) = 0;
    virtual ~FastObserver() {}
};
inline FastObserver*& fast_observer() {
    static thread_local FastObserver* o = nullptr;
    return o;
}

static inline void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold,
                        bool nonmaxSuppression) {
    assert(nonmaxSuppression);
    (void)nonmaxSuppression;
    StageTimer _st(&stage_clock().fast);
    Mat img = image.getMat();
    std::vector<cvprims::FastKp> v;
    cvprims::fast9_nms(img.data, img.rows, img.cols, img.step, threshold, v);
    keypoints.clear();
    keypoints.reserve(v.size());
    for (const auto& k : v)
        keypoints.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1, (float)k.score));
    if (fast_observer()) fast_observer()->on_fast(img, threshold, keypoints);
}

static inline void resize(InputArray src, OutputArray dst, Size dsize, double fx, double fy,
                          int interpolation) {
    assert(fx == 0 && fy == 0 && interpolation == INTER_LINEAR);
    (void)fx; (void)fy; (void)interpolation;
    StageTimer _st(&stage_clock().resize);
    Mat s = src.getMat();
    dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat d = dst.getMat();
    cvprims::resize_linear_u8(s.data, s.rows, s.cols, s.step, d.data, d.rows, d.cols, d.step);
}

static inline void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left,
                                  int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)borderType;
    Mat s = src.getMat();
    dst.create(s.rows + top + bottom, s.cols + left + right, CV_8UC1);
    Mat d = dst.getMat();
    cvprims::copy_make_border_reflect101(s.data, s.rows, s.cols, s.step, d.data, d.step, top,
                                         bottom, left, right);
}

static inline void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sx, double sy,
                                int borderType) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 &&
           borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    StageTimer _st(&stage_clock().blur);
    Mat s = src.getMat();
    dst.create(s.rows, s.cols, CV_8UC1);
    Mat d = dst.getMat();
    cvprims::gaussian7_sigma2_u8(s.data, s.rows, s.cols, s.step, d.data, d.step);
}

struct KeyPointsFilter {
    // Only reached from ORBextractor::ComputeKeyPointsOld, which is dead code
    // (call commented out at src/ORBextractor.cc:1083).
    static void retainBest(std::vector<KeyPoint>&, int) {
        fprintf(stderr, "cvshim: KeyPointsFilter::retainBest is not implemented (dead code)\n");
        abort();
    }
};

}  // namespace cv
#endif  // ORACLE_CVSHIM_H
