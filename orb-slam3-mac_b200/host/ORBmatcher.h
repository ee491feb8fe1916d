// orb-slam3-mac_b200/host/ORBmatcher.h -- source-level drop-in for the tracking half of the reference's
// include/ORBmatcher.h:34-104: the same class name, namespace, constructor, constants and the searches the tracking
// thread runs on every frame, implemented purely by marshalling the reference's own Frame / MapPoint objects to the
// C ABI of include/orbm.h (liborbx_b200.so) and writing the matches back where the reference writes them
// (F.mvpMapPoints, vnMatches12, vbPrevMatched).  Replace the reference's header with this one and src/ORBmatcher.cc
// with host/ORBmatcher.cc, link -lorbx_b200; Frame.cc / MapPoint.cc / Tracking.cc compile unchanged.
//
//   DescriptorDistance                                           src/ORBmatcher.cc:2353-2369  (host, as the reference)
//   SearchByProjection(Frame&, vector<MapPoint*>&, th, far...)   :48-218   -> orbm_search_by_projection_points[_stereo]
//   SearchByProjection(Frame& Current, const Frame& Last, ...)   :1965-2181 -> orbm_search_by_projection_frames[_stereo]
//   SearchForInitialization                                      :710-825  -> orbm_search_for_initialization
//   SearchByBoW(KeyFrame*, Frame&, vpMapPointMatches)            :273-475  -> orbm_search_by_bow
//   SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)               :827-967  -> orbm_search_by_bow_kf
//
// The other LocalMapping / LoopClosing searches (the Sim3 / relocalisation SearchByProjection, Fuse x2,
// SearchBySim3, SearchForTriangulation x2) have their C-ABI entry points too (include/orbm.h); INTEGRATION.md shows
// their bindings, this class does not wrap them yet.
//
// Builds against real OpenCV where it is installed and against oracle/cvshim + oracle/matchshim in this repository's
// tests (tests/test_cpp_dropin.py: the reference's src/ORBmatcher.cc and this class are driven through the same
// wrapper on the same Frame / MapPoint objects and must leave them in the same state).
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <vector>
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "MapPoint.h"
#include "KeyFrame.h"
#include "Frame.h"

struct orbm_context;

namespace ORB_SLAM3 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);

    // Computes the Hamming distance between two ORB descriptors
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);

    // Search matches between Frame keypoints and projected MapPoints. Returns number of matches
    // Used to track the local map (Tracking)
    int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3,
                           const bool bFarPoints = false, const float thFarPoints = 50.0f);

    // Project MapPoints tracked in last frame into the current frame and search matches.
    // Used to track from previous frame (Tracking)
    int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);

    // Search matches between MapPoints in a KeyFrame and ORB in a Frame.
    // Brute force constrained to ORB that belong to the same vocabulary node (at a certain level)
    // Used in Relocalisation and Loop Detection
    int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
    int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);

    // Matching for the Map Initialization (only used in the monocular case)
    int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                std::vector<int>& vnMatches12, int windowSize = 10);

public:
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

    // Addition of the B200 build: the calling thread's matcher context (created on first use on the device named by
    // the environment variable ORBM_DEVICE, default 0; destroyed with the thread).  Tracking, LocalMapping and
    // LoopClosing each construct ORBmatcher objects on their own thread, so each gets its own stream and buffers.
    static orbm_context* Context();

protected:
    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM3

#endif  // ORBMATCHER_H
