// oracle/cvprims.h -- TEST INFRASTRUCTURE ONLY (CPU oracle). Never linked into the product.
//
// Plain C++ restatement of the un-vendored OpenCV primitives that the reference's
// ORB front-end calls.  The reference call sites are
//   cv::resize(INTER_LINEAR)          src/ORBextractor.cc:1165
//   cv::copyMakeBorder(REFLECT_101)   src/ORBextractor.cc:1167,1172
//   cv::FAST(img, kps, th, true)      src/ORBextractor.cc:808,827
//   cv::GaussianBlur(7x7, s=2)        src/ORBextractor.cc:1115
//   cv::fastAtan2                     src/ORBextractor.cc:101
//   cvRound / cvFloor / cvCeil        src/ORBextractor.cc:79,113,117-118,440,458,1157
// OpenCV itself is NOT in /root/reference (CMakeLists.txt:32-38 asks for OpenCV 3, the
// fork was built with 3.4.1).  The arithmetic below is pinned to OpenCV 4.13 (the only
// OpenCV present in this image, Python module `cv2`); tests/test_oracle_prims.py checks
// every function here bit-for-bit against cv2 4.13.
//
// Everything works on raw (ptr, rows, cols, step) views so that the same code backs
//   * oracle/cvshim/  (a tiny cv:: facade used to compile the reference file verbatim), and
//   * oracle/orb_oracle.cpp (our own restatement of the extractor algorithm).
#ifndef ORACLE_CVPRIMS_H
#define ORACLE_CVPRIMS_H

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

namespace cvprims {

// ---- rounding helpers (OpenCV: cvRound = round-half-even via SSE cvtsd2si / lrint) --------
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_floor(double v) {
    int i = (int)v;
    return i - (i > v);
}
static inline int cv_ceil(double v) {
    int i = (int)v;
    return i + (i < v);
}

static inline int reflect101(int p, int len) {
    // BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

// ---- cv::copyMakeBorder(..., BORDER_REFLECT_101 [| BORDER_ISOLATED]) ----------------------
// dst is (rows+top+bottom) x (cols+left+right).
static inline void copy_make_border_reflect101(const uint8_t* src, int rows, int cols, size_t sstep,
                                               uint8_t* dst, size_t dstep, int top, int bottom,
                                               int left, int right) {
    const int drows = rows + top + bottom, dcols = cols + left + right;
    for (int y = 0; y < drows; ++y) {
        const uint8_t* s = src + (size_t)reflect101(y - top, rows) * sstep;
        uint8_t* d = dst + (size_t)y * dstep;
        if (d + left != s) memmove(d + left, s, (size_t)cols);
        for (int x = 0; x < left; ++x) d[x] = s[reflect101(x - left, cols)];
        for (int x = 0; x < right; ++x) d[left + cols + x] = s[reflect101(cols + x, cols)];
        (void)dcols;
    }
}

// ---- cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1 --------------------------
// OpenCV imgproc/resize.cpp: resizeGeneric_ + HResizeLinear<uchar,int,short,2048> +
// VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>> (the 8U specialisation).
struct LinearTab {
    std::vector<int> ofs;       // source index of the first tap
    std::vector<short> a0, a1;  // 11-bit fixed point coefficients
};

static inline void linear_tab(int ssize, int dsize, bool clamp_coeffs, LinearTab& t) {
    t.ofs.resize(dsize);
    t.a0.resize(dsize);
    t.a1.resize(dsize);
    const double inv_scale = (double)dsize / ssize;
    const double scale = 1. / inv_scale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (clamp_coeffs) {  // x axis: coefficients are zeroed at the clamps
            if (s < 0) { f = 0; s = 0; }
            if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        }
        t.ofs[d] = s;
        float c0 = 1.f - f, c1 = f;
        int i0 = cv_round(c0 * 2048.f), i1 = cv_round(c1 * 2048.f);
        t.a0[d] = (short)std::max(-32768, std::min(32767, i0));
        t.a1[d] = (short)std::max(-32768, std::min(32767, i1));
    }
}

static inline void resize_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep,
                                    uint8_t* dst, int drows, int dcols, size_t dstep) {
    LinearTab tx, ty;
    linear_tab(scols, dcols, true, tx);
    linear_tab(srows, drows, false, ty);  // y axis: rows are clipped, coefficients are not
    std::vector<int> h0(dcols), h1(dcols);
    int cached0 = -1, cached1 = -1;  // source rows currently held in h0 / h1
    auto hrow = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src + (size_t)sy * sstep;
        for (int x = 0; x < dcols; ++x) {
            int sx = tx.ofs[x];
            int sx1 = sx + 1 < scols ? sx + 1 : sx;  // a1 == 0 there
            out[x] = S[sx] * tx.a0[x] + S[sx1] * tx.a1[x];
        }
    };
    for (int y = 0; y < drows; ++y) {
        int sy0 = ty.ofs[y], sy1 = sy0 + 1;
        sy0 = sy0 < 0 ? 0 : (sy0 >= srows ? srows - 1 : sy0);
        sy1 = sy1 < 0 ? 0 : (sy1 >= srows ? srows - 1 : sy1);
        if (sy0 == cached1) { h0.swap(h1); std::swap(cached0, cached1); }
        if (sy0 != cached0) { hrow(sy0, h0); cached0 = sy0; }
        if (sy1 == cached0) { h1 = h0; cached1 = sy1; }
        else if (sy1 != cached1) { hrow(sy1, h1); cached1 = sy1; }
        const int b0 = ty.a0[y], b1 = ty.a1[y];
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < dcols; ++x) {
            int v = (((b0 * (h0[x] >> 4)) >> 16) + ((b1 * (h1[x] >> 4)) >> 16) + 2) >> 2;
            D[x] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
}

// ---- cv::FAST(img, keypoints, threshold, nonmaxSuppression=true), TYPE_9_16 ---------------
// OpenCV features2d/fast.cpp FAST_t<16> + fast_score.cpp cornerScore<16>.
struct FastKp {
    int x, y;   // sub-image coordinates
    int score;  // KeyPoint::response (size = 7, angle = -1, octave = 0, class_id = -1)
};

static const int kFastRing[16][2] = {{0, 3},  {1, 3},   {2, 2},   {3, 1},  {3, 0},  {3, -1},
                                     {2, -2}, {1, -3},  {0, -3},  {-1, -3}, {-2, -2}, {-3, -1},
                                     {-3, 0}, {-3, 1},  {-2, 2},  {-1, 3}};

static inline int fast_corner_score16(const uint8_t* ptr, const int pixel[25], int threshold) {
    const int K = 8, N = K * 3 + 1;
    int k, v = ptr[0];
    short d[N];
    for (k = 0; k < N; k++) d[k] = (short)(v - ptr[pixel[k]]);
    int a0 = threshold;
    for (k = 0; k < 16; k += 2) {
        int a = std::min((int)d[k + 1], (int)d[k + 2]);
        a = std::min(a, (int)d[k + 3]);
        if (a <= a0) continue;
        a = std::min(a, (int)d[k + 4]);
        a = std::min(a, (int)d[k + 5]);
        a = std::min(a, (int)d[k + 6]);
        a = std::min(a, (int)d[k + 7]);
        a = std::min(a, (int)d[k + 8]);
        a0 = std::max(a0, std::min(a, (int)d[k]));
        a0 = std::max(a0, std::min(a, (int)d[k + 9]));
    }
    int b0 = -a0;
    for (k = 0; k < 16; k += 2) {
        int b = std::max((int)d[k + 1], (int)d[k + 2]);
        b = std::max(b, (int)d[k + 3]);
        b = std::max(b, (int)d[k + 4]);
        b = std::max(b, (int)d[k + 5]);
        if (b >= b0) continue;
        b = std::max(b, (int)d[k + 6]);
        b = std::max(b, (int)d[k + 7]);
        b = std::max(b, (int)d[k + 8]);
        b0 = std::min(b0, std::max(b, (int)d[k]));
        b0 = std::min(b0, std::max(b, (int)d[k + 9]));
    }
    return -b0 - 1;
}

static inline void fast9_nms(const uint8_t* img, int rows, int cols, size_t step, int threshold,
                             std::vector<FastKp>& out) {
    out.clear();
    const int K = 8, N = 25;
    if (rows < 7 || cols < 7) return;
    int pixel[25];
    for (int k = 0; k < 16; ++k) pixel[k] = kFastRing[k][0] + kFastRing[k][1] * (int)step;
    for (int k = 16; k < 25; ++k) pixel[k] = pixel[k - 16];
    threshold = std::min(std::max(threshold, 0), 255);
    uint8_t tab[512];
    for (int i = -255; i <= 255; i++)
        tab[i + 255] = (uint8_t)(i < -threshold ? 1 : i > threshold ? 2 : 0);

    std::vector<uint8_t> bufv((size_t)cols * 3, 0);
    uint8_t* buf[3] = {bufv.data(), bufv.data() + cols, bufv.data() + 2 * cols};
    std::vector<int> cpv((size_t)(cols + 1) * 3, 0);
    int* cpbuf[3] = {cpv.data() + 1, cpv.data() + (cols + 1) + 1, cpv.data() + 2 * (cols + 1) + 1};

    for (int i = 3; i < rows - 2; i++) {
        const uint8_t* ptr = img + (size_t)i * step + 3;
        uint8_t* curr = buf[(i - 3) % 3];
        int* cornerpos = cpbuf[(i - 3) % 3];
        memset(curr, 0, (size_t)cols);
        int ncorners = 0;
        if (i < rows - 3) {
            for (int j = 3; j < cols - 3; j++, ptr++) {
                int v = ptr[0];
                const uint8_t* t = &tab[0] - v + 255;
                int d = t[ptr[pixel[0]]] | t[ptr[pixel[8]]];
                if (d == 0) continue;
                d &= t[ptr[pixel[2]]] | t[ptr[pixel[10]]];
                d &= t[ptr[pixel[4]]] | t[ptr[pixel[12]]];
                d &= t[ptr[pixel[6]]] | t[ptr[pixel[14]]];
                if (d == 0) continue;
                d &= t[ptr[pixel[1]]] | t[ptr[pixel[9]]];
                d &= t[ptr[pixel[3]]] | t[ptr[pixel[11]]];
                d &= t[ptr[pixel[5]]] | t[ptr[pixel[13]]];
                d &= t[ptr[pixel[7]]] | t[ptr[pixel[15]]];
                bool corner = false;
                if (d & 1) {
                    int vt = v - threshold, count = 0;
                    for (int k = 0; k < N; k++) {
                        int x = ptr[pixel[k]];
                        if (x < vt) {
                            if (++count > K) { corner = true; break; }
                        } else
                            count = 0;
                    }
                }
                if (!corner && (d & 2)) {
                    int vt = v + threshold, count = 0;
                    for (int k = 0; k < N; k++) {
                        int x = ptr[pixel[k]];
                        if (x > vt) {
                            if (++count > K) { corner = true; break; }
                        } else
                            count = 0;
                    }
                }
                if (corner) {
                    cornerpos[ncorners++] = j;
                    curr[j] = (uint8_t)fast_corner_score16(ptr, pixel, threshold);
                }
            }
        }
        cornerpos[-1] = ncorners;
        if (i == 3) continue;
        const uint8_t* prev = buf[(i - 4 + 3) % 3];
        const uint8_t* pprev = buf[(i - 5 + 3) % 3];
        cornerpos = cpbuf[(i - 4 + 3) % 3];
        ncorners = cornerpos[-1];
        for (int k = 0; k < ncorners; k++) {
            int j = cornerpos[k];
            int score = prev[j];
            if (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] &&
                score > pprev[j] && score > pprev[j + 1] && score > curr[j - 1] &&
                score > curr[j] && score > curr[j + 1]) {
                out.push_back(FastKp{j, i - 1, score});
            }
        }
    }
}

// ---- cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 ----------
// OpenCV >= 4.1.1 (here 4.13) fixed-point path: getGaussianKernelBitExact -> 8.8 kernel
// {18,34,48,56,48,34,18} (sums to 256), horizontal pass exact in ufixedpoint16, vertical pass
// ufixedpoint32, result = (sum + 32768) >> 16.  In-place safe (uses a temporary).
static inline void gaussian7_sigma2_u8(const uint8_t* src, int rows, int cols, size_t sstep,
                                       uint8_t* dst, size_t dstep) {
    static const int kk[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> h((size_t)rows * cols);
    for (int y = 0; y < rows; ++y) {
        const uint8_t* S = src + (size_t)y * sstep;
        uint16_t* H = h.data() + (size_t)y * cols;
        for (int x = 0; x < cols; ++x) {
            int acc = 0;
            if (x >= 3 && x + 3 < cols) {
                for (int k = 0; k < 7; ++k) acc += kk[k] * S[x + k - 3];
            } else {
                for (int k = 0; k < 7; ++k) acc += kk[k] * S[reflect101(x + k - 3, cols)];
            }
            H[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < rows; ++y) {
        const uint16_t* R[7];
        for (int k = 0; k < 7; ++k) R[k] = h.data() + (size_t)reflect101(y + k - 3, rows) * cols;
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < cols; ++x) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; ++k) acc += (uint32_t)kk[k] * R[k][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// ---- cv::cvtColor(src, dst, CV_{BGR,RGB,BGRA,RGBA}2GRAY) for CV_8U --------------------------
// Reference call sites: src/Tracking.cc:354-379, 425-438, 479-492.  Integer luma:
//   bits = 15 (OpenCV 4.x, the one pinned here against cv2 4.13): (R*9798 + G*19235 + B*3735 + 2^14) >> 15
//   bits = 14 (OpenCV 2.4/3.x, the fork's 3.4.1 build; published constants R2Y 4899, G2Y 9617,
//              B2Y 1868, yuv_shift 14 -- not pinnable here): (R*4899 + G*9617 + B*1868 + 2^13) >> 14
// channels = 3 or 4 (alpha ignored); rgb_order != 0 means the first channel is R.
static inline void cvt_gray_u8(const uint8_t* src, int rows, int cols, size_t sstep, int channels,
                               int rgb_order, int bits, uint8_t* dst, size_t dstep) {
    const int cr = bits == 15 ? 9798 : 4899, cg = bits == 15 ? 19235 : 9617, cb = bits == 15 ? 3735 : 1868;
    for (int y = 0; y < rows; ++y) {
        const uint8_t* S = src + (size_t)y * sstep;
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < cols; ++x) {
            const int c0 = S[x * channels], c1 = S[x * channels + 1], c2 = S[x * channels + 2];
            const int r = rgb_order ? c0 : c2, b = rgb_order ? c2 : c0;
            D[x] = (uint8_t)((r * cr + c1 * cg + b * cb + (1 << (bits - 1))) >> bits);
        }
    }
}

// ---- cv::remap(src, dst, map1, map2, INTER_LINEAR) for CV_8UC{1,3,4}, CV_32FC1 maps ----------
// Reference call site: Examples/Stereo/stereo_euroc.cc:170-171 (maps from initUndistortRectifyMap,
// :117-118), BORDER_CONSTANT with value 0 (the default).  OpenCV's fixed-point path
// (imgwarp.cpp remap -> remapBilinear): sx = cvRound(mapx * 32) with the product in float32,
// integer part sx >> 5 (saturated to short), fraction sx & 31; the 32x32 weight table holds
// round(wy*wx*32768) -- exact multiples of 32 -- so the pixel is
//   (a00*v00 + a01*v01 + a10*v10 + a11*v11 + 512) >> 10,  a = (32-fx)(32-fy), fx(32-fy), (32-fx)fy, fx*fy
// (the saturated entry of the zero-fraction case, {32767,0,0,1}, gives v00 as well); source
// pixels outside the image read as 0.  Pinned bit-exact against cv2 4.13 in tests/test_oracle_prims.py.
static inline void remap_linear_u8(const uint8_t* src, int srows, int scols, size_t sstep, int channels,
                                   const float* mapx, const float* mapy, size_t mstep_floats, uint8_t* dst,
                                   int drows, int dcols, size_t dstep) {
    for (int y = 0; y < drows; ++y) {
        uint8_t* D = dst + (size_t)y * dstep;
        for (int x = 0; x < dcols; ++x) {
            const int sx = cv_round(mapx[(size_t)y * mstep_floats + x] * 32.0f);
            const int sy = cv_round(mapy[(size_t)y * mstep_floats + x] * 32.0f);
            const int fx = sx & 31, fy = sy & 31;
            int ix = sx >> 5, iy = sy >> 5;
            ix = ix < -32768 ? -32768 : ix > 32767 ? 32767 : ix;
            iy = iy < -32768 ? -32768 : iy > 32767 ? 32767 : iy;
            const int a00 = (32 - fx) * (32 - fy), a01 = fx * (32 - fy), a10 = (32 - fx) * fy, a11 = fx * fy;
            for (int c = 0; c < channels; ++c) {
                int v[4];
                for (int k = 0; k < 4; ++k) {
                    const int xx = ix + (k & 1), yy = iy + (k >> 1);
                    v[k] = (xx >= 0 && xx < scols && yy >= 0 && yy < srows)
                               ? src[(size_t)yy * sstep + (size_t)xx * channels + c] : 0;
                }
                D[x * channels + c] = (uint8_t)((a00 * v[0] + a01 * v[1] + a10 * v[2] + a11 * v[3] + 512) >> 10);
            }
        }
    }
}

// ---- cv::fastAtan2(y, x) -> degrees in [0, 360) --------------------------------------------
// OpenCV core/mathfuncs_core: 7th-order odd polynomial, float32, no FMA contraction
// (compile the including TU with -ffp-contract=off).
static inline float fast_atan2_deg(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

}  // namespace cvprims
#endif  // ORACLE_CVPRIMS_H
