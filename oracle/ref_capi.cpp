// oracle/ref_capi.cpp -- TEST INFRASTRUCTURE ONLY.
//
// C wrapper around the reference's own ORB_SLAM3::ORBextractor (include/ORBextractor.h:43-109,
// src/ORBextractor.cc compiled VERBATIM from /root/reference by oracle/Makefile) so that the
// Python tests and bench.py's cpu_baseline / --impl reference legs can drive it through ctypes.
// Built into oracle/_ref/liborbref.so (git-ignored; travels to the GPU box with gpurun).
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"  // the reference header, found via -I/root/reference/include

namespace {

struct FastCall {
    const uchar* data;
    int rows, cols, th;
    std::vector<cv::KeyPoint> kps;
};

struct Recorder : cv::FastObserver {
    std::vector<FastCall> calls;
    void on_fast(const cv::Mat& img, int th, const std::vector<cv::KeyPoint>& kps) override {
        calls.push_back(FastCall{img.data, img.rows, img.cols, th, kps});
    }
};

// Derived only to reach the protected stage functions; no behaviour is changed.
struct Probe : ORB_SLAM3::ORBextractor {
    Probe(int n, float s, int l, int a, int b) : ORB_SLAM3::ORBextractor(n, s, l, a, b) {}
    using ORB_SLAM3::ORBextractor::ComputePyramid;
    using ORB_SLAM3::ORBextractor::ComputeKeyPointsOctTree;
    const std::vector<int>& perLevel() const { return mnFeaturesPerLevel; }
    const std::vector<int>& uMax() const { return umax; }
    const std::vector<cv::Point>& pat() const { return pattern; }

    // stage dumps of the last orbref_stages() call
    std::vector<std::vector<int> > cand;                  // per level: x,y,response triplets
    std::vector<std::vector<cv::KeyPoint> > survivors;    // per level, level coordinates, with angle
};

}  // namespace

extern "C" {

void* orbref_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    return new Probe(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
}

void orbref_destroy(void* h) { delete (Probe*)h; }

// 0 = canonical creation-order tie-break (bump arena, default); 1 = plain malloc, i.e. the
// allocator-dependent behaviour of a stock build (used only to REPORT the difference rate).
void orbref_set_malloc_mode(int use_malloc) { orb_arena::tls().enabled = !use_malloc; }

// In-situ stage clock of the calling thread (cvshim.h StageClock): on != 0 starts it from zero; the call returns the
// seconds accumulated so far inside cv::resize / cv::FAST / cv::GaussianBlur through out3 (may be NULL).
void orbref_stage_clock(int on, double* out3) {
    cv::StageClock& c = cv::stage_clock();
    if (out3) {
        out3[0] = c.resize;
        out3[1] = c.fast;
        out3[2] = c.blur;
    }
    c.on = on != 0;
    c.resize = c.fast = c.blur = 0;
}

int orbref_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                  int* n_per_level) {
    Probe* p = (Probe*)h;
    int n = p->GetLevels();
    std::vector<float> a = p->GetScaleFactors(), b = p->GetInverseScaleFactors(),
                       c = p->GetScaleSigmaSquares(), d = p->GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; ++i) {
        scale[i] = a[i];
        inv_scale[i] = b[i];
        sigma2[i] = c[i];
        inv_sigma2[i] = d[i];
        n_per_level[i] = p->perLevel()[i];
    }
    return n;
}

int orbref_umax(void* h, int* out16) {
    Probe* p = (Probe*)h;
    for (int i = 0; i < 16; ++i) out16[i] = p->uMax()[i];
    return 16;
}

int orbref_pattern(void* h, int* out1024) {
    Probe* p = (Probe*)h;
    for (int i = 0; i < 512; ++i) {
        out1024[2 * i] = p->pat()[i].x;
        out1024[2 * i + 1] = p->pat()[i].y;
    }
    return 512;
}

// ORBextractor::operator() (src/ORBextractor.cc:1068).  kps: cap records of 7 x 4 bytes
// {x, y, size, angle, response (float), octave, class_id (int32)}.  Returns monoIndex
// (or -1 for an empty image, as the reference does); *n_out = number of keypoints, or
// -2 if cap is too small (nothing written).
int orbref_extract(void* h, const uint8_t* img, int w, int h_, size_t stride, int lap0, int lap1,
                   void* kps, uint8_t* desc, int cap, int* n_out) {
    Probe* p = (Probe*)h;
    orb_arena::reset();
    cv::Mat image = (img && w > 0 && h_ > 0) ? cv::Mat(h_, w, CV_8UC1, (void*)img, stride) : cv::Mat();
    std::vector<cv::KeyPoint> keys;
    cv::Mat descriptors;
    std::vector<int> lap(2);
    lap[0] = lap0;
    lap[1] = lap1;
    int mono = (*p)(image, cv::Mat(), keys, descriptors, lap);
    int n = (int)keys.size();
    if (n > cap) {
        *n_out = -2;
        return mono;
    }
    *n_out = n;
    static_assert(sizeof(cv::KeyPoint) == 28, "KeyPoint layout");
    if (n) {
        memcpy(kps, keys.data(), (size_t)n * sizeof(cv::KeyPoint));
        for (int i = 0; i < n; ++i) memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
    }
    return mono;
}

// mvImagePyramid[level] of the last extract / stages call (include/ORBextractor.h:83).
int orbref_level_size(void* h, int level, int* w, int* h_) {
    Probe* p = (Probe*)h;
    *w = p->mvImagePyramid[level].cols;
    *h_ = p->mvImagePyramid[level].rows;
    return 0;
}

// pad = 0: the level ROI; pad = 19: the whole padded buffer (ROI + EDGE_THRESHOLD border).
int orbref_get_level(void* h, int level, int pad, uint8_t* dst, size_t dst_stride) {
    Probe* p = (Probe*)h;
    const cv::Mat& m = p->mvImagePyramid[level];
    for (int y = -pad; y < m.rows + pad; ++y)
        memcpy(dst + (size_t)(y + pad) * dst_stride, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - pad,
               (size_t)m.cols + 2 * pad);
    return 0;
}

// Runs ComputePyramid + ComputeKeyPointsOctTree (src/ORBextractor.cc:1079-1082) and keeps
//   * the per-level FAST candidate lists in the reference's order (vToDistributeKeys,
//     :776-854; coordinates relative to (16,16) as in the reference), and
//   * the per-level octree survivors with orientation (allKeypoints, :856-877).
int orbref_stages(void* h, const uint8_t* img, int w, int h_, size_t stride) {
    Probe* p = (Probe*)h;
    orb_arena::reset();
    cv::Mat image(h_, w, CV_8UC1, (void*)img, stride);
    p->ComputePyramid(image);
    Recorder rec;
    cv::fast_observer() = &rec;
    std::vector<std::vector<cv::KeyPoint> > all;
    p->ComputeKeyPointsOctTree(all);
    cv::fast_observer() = nullptr;
    const int nl = p->GetLevels();
    p->cand.assign(nl, std::vector<int>());
    for (size_t c = 0; c < rec.calls.size(); ++c) {
        const FastCall& fc = rec.calls[c];
        // a cell that came back empty at iniThFAST is re-run at minThFAST on the same view
        if (c + 1 < rec.calls.size() && rec.calls[c + 1].data == fc.data && fc.kps.empty()) continue;
        int lvl = -1;
        for (int l = 0; l < nl; ++l) {
            const cv::Mat& m = p->mvImagePyramid[l];
            const uchar* lo = m.data;
            const uchar* hi = m.data + (size_t)(m.rows - 1) * m.step + m.cols;
            if (fc.data >= lo && fc.data < hi) { lvl = l; break; }
        }
        if (lvl < 0) return -1;
        const cv::Mat& m = p->mvImagePyramid[lvl];
        size_t off = (size_t)(fc.data - m.data);
        int iniY = (int)(off / m.step), iniX = (int)(off % m.step);
        for (const cv::KeyPoint& k : fc.kps) {
            p->cand[lvl].push_back((int)k.pt.x + iniX - 16);
            p->cand[lvl].push_back((int)k.pt.y + iniY - 16);
            p->cand[lvl].push_back((int)k.response);
        }
    }
    p->survivors = all;
    return 0;
}

int orbref_stage_candidates(void* h, int level, int* xyr, int cap) {
    Probe* p = (Probe*)h;
    int n = (int)p->cand[level].size() / 3;
    if (xyr && n <= cap) memcpy(xyr, p->cand[level].data(), (size_t)n * 3 * sizeof(int));
    return n;
}

int orbref_stage_survivors(void* h, int level, void* kps, int cap) {
    Probe* p = (Probe*)h;
    int n = (int)p->survivors[level].size();
    if (kps && n <= cap) memcpy(kps, p->survivors[level].data(), (size_t)n * sizeof(cv::KeyPoint));
    return n;
}

}  // extern "C"
