// oracle/cvshim/cvshim.h -- TEST INFRASTRUCTURE ONLY.
//
// A minimal `cv::` facade that provides exactly the OpenCV surface used by the reference's
// src/ORBextractor.cc + include/ORBextractor.h, so that the reference file can be compiled
// VERBATIM from /root/reference (see oracle/Makefile) without OpenCV C++ being installed.
// All pixel arithmetic is delegated to oracle/cvprims.h, which is pinned bit-for-bit against
// cv2 4.13 by tests/test_oracle_prims.py.  This is not a general OpenCV replacement.
#ifndef ORACLE_CVSHIM_H
#define ORACLE_CVSHIM_H

#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "../cvprims.h"

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

typedef unsigned char uchar;

static inline int cvRound(double v) { return cvprims::cv_round(v); }
static inline int cvRound(float v) { return cvprims::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return cvprims::cv_floor(v); }
static inline int cvCeil(double v) { return cvprims::cv_ceil(v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

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

// 8-bit single channel matrix with shared ownership and ROI views.
class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;
    std::shared_ptr<uchar> owner;

    Mat() : rows(0), cols(0), data(nullptr), step(0) {}
    Mat(int r, int c, int type) { alloc(r, c); (void)type; }
    Mat(Size sz, int type) { alloc(sz.height, sz.width); (void)type; }
    // external data, not owned
    Mat(int r, int c, int type, void* ext, size_t ext_step)
        : rows(r), cols(c), data((uchar*)ext), step(ext_step) { (void)type; }

    void alloc(int r, int c) {
        rows = r;
        cols = c;
        step = (size_t)c;
        size_t n = (size_t)r * (size_t)c;
        owner.reset((uchar*)malloc(n ? n : 1), free);
        data = owner.get();
    }
    void create(int r, int c, int type) {
        (void)type;
        if (data && r == rows && c == cols) return;
        alloc(r, c);
    }
    void release() {
        owner.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step; }

    Mat operator()(const Rect& r) const {
        Mat m(*this);
        m.data = data + (size_t)r.y * step + r.x;
        m.rows = r.height;
        m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const {
        Mat m(*this);
        m.data = data + (size_t)a * step;
        m.rows = b - a;
        return m;
    }
    Mat colRange(int a, int b) const {
        Mat m(*this);
        m.data = data + a;
        m.cols = b - a;
        return m;
    }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat clone() const {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; ++y) memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }
    void copyTo(Mat dst) const {  // dst is a view of the right shape (row copy)
        assert(dst.rows == rows && dst.cols == cols);
        for (int y = 0; y < rows; ++y) memcpy(dst.data + (size_t)y * dst.step, data + (size_t)y * step, cols);
    }
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        memset(m.data, 0, (size_t)r * c);
        return m;
    }
    template <typename T>
    T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x); }
    template <typename T>
    const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T>
    T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
};

class _InputArray {
public:
    Mat* m;
    Mat tmp;
    _InputArray(const Mat& mm) : m(const_cast<Mat*>(&mm)) {}
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
