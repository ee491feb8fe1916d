// orb-slam3-mac_b200/host/ORBextractor.h -- source-level drop-in for the reference's
// include/ORBextractor.h:43-109: the same class name, namespace, constructor, operator(), getters
// and public mvImagePyramid member, implemented purely by marshalling to the C ABI of
// include/orbx.h (liborbx_b200.so).  Replace the reference's header + src/ORBextractor.cc with this
// header and link -lorbx_b200; Frame.cc / Tracking.cc compile unchanged.
//
// Builds against real OpenCV (cv::Mat, cv::KeyPoint, cv::InputArray) where it is installed, and
// against oracle/cvshim in this repository's tests (tests/test_cpp_dropin.py).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv/cv.h>

#include "orbx.h"

namespace ORB_SLAM3 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // Same five arguments as the reference (src/ORBextractor.cc:408).  The device buffers are sized
    // lazily for the first image and grow if a larger one arrives.
    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_),
          minThFAST(minThFAST_), handle(nullptr), maxW(0), maxH(0) {
        mvImagePyramid.resize(nlevels);
        ensure(752, 480);  // builds the tables; any later size re-creates the handle if larger
    }
    ~ORBextractor() { orbx_destroy(handle); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Compute the ORB features and descriptors on an image (mask is ignored, as in the reference).
    int operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        (void)_mask;
        if (_image.empty()) return -1;  // src/ORBextractor.cc:1072-1073
        cv::Mat image = _image.getMat();
        if (image.type() != CV_8UC1) throw std::invalid_argument("ORBextractor: image must be CV_8UC1");
        ensure(image.cols, image.rows);
        const int cap = orbx_max_keypoints(handle);
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");
        std::vector<orbx_keypoint> kps(cap);
        std::vector<uint8_t> desc((size_t)cap * 32);
        int n = 0, mono = 0;
        int rc = orbx_extract(handle, image.data, image.cols, image.rows, (size_t)image.step, vLappingArea[0],
                              vLappingArea[1], kps.data(), desc.data(), cap, &n, &mono);
        if (rc != ORBX_OK) throw std::runtime_error(std::string("orbx_extract: ") + orbx_last_error_string());
        _keypoints.resize(n);
        if (n == 0) {
            _descriptors.release();  // :1090-1091
        } else {
            std::memcpy(static_cast<void*>(_keypoints.data()), kps.data(), (size_t)n * sizeof(orbx_keypoint));
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), &desc[(size_t)i * 32], 32);
        }
        // mvImagePyramid (read by Frame::ComputeStereoMatches, src/Frame.cc:809-918): ROI views into
        // buffers with the reference's 19-px BORDER_REFLECT_101 frame
        for (int l = 0; l < nlevels && mbCopyPyramid; ++l) {
            int lw = 0, lh = 0;
            orbx_level_size(handle, image.cols, image.rows, l, &lw, &lh);
            cv::Mat temp(lh + 38, lw + 38, CV_8UC1);
            orbx_get_pyramid_level(handle, 0, 0, l, 19, temp.data, (size_t)temp.step);
            mvImagePyramid[l] = temp(cv::Rect(19, 19, lw, lh));
        }
        return mono;
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;

    // Additions of the B200 build (not in the reference class).  Handle(): the C-ABI handle, for
    // same-GPU consumers of the device-resident pyramid (orbx_device_pyramid ->
    // orbm_compute_stereo_matches).  mbCopyPyramid = false skips the per-frame
    // host copy of mvImagePyramid once Frame::ComputeStereoMatches runs on the device.
    orbx_handle* Handle() const { return handle; }
    bool mbCopyPyramid = true;

protected:
    void ensure(int w, int h) {
        if (handle && w <= maxW && h <= maxH) return;
        if (handle) orbx_destroy(handle);
        handle = nullptr;
        maxW = w > maxW ? w : maxW;
        maxH = h > maxH ? h : maxH;
        orbx_config cfg;
        std::memset(&cfg, 0, sizeof cfg);
        cfg.nfeatures = nfeatures;
        cfg.scale_factor = (float)scaleFactor;
        cfg.nlevels = nlevels;
        cfg.ini_th_fast = iniThFAST;
        cfg.min_th_fast = minThFAST;
        cfg.max_width = maxW;
        cfg.max_height = maxH;
        cfg.max_batch = 1;
        cfg.n_slots = 1;
        if (orbx_create(&cfg, &handle) != ORBX_OK)
            throw std::runtime_error(std::string("orbx_create: ") + orbx_last_error_string());
        mvScaleFactor.resize(nlevels);
        mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels);
        mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels);
        orbx_get_tables(handle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                        mvInvLevelSigma2.data(), mnFeaturesPerLevel.data());
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbx_handle* handle;
    int maxW, maxH;
};

}  // namespace ORB_SLAM3

#endif
