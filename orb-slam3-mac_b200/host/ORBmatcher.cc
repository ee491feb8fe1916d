// orb-slam3-mac_b200/host/ORBmatcher.cc -- takes the place of the reference's src/ORBmatcher.cc for the methods
// host/ORBmatcher.h declares.  Nothing here compares descriptors or walks the grid: every search snapshots the
// reference's objects into the POD views of include/orbm.h, calls the CUDA path, and writes the matches back.
#include "ORBmatcher.h"

#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>

#include "orbm.h"

namespace ORB_SLAM3 {

const int ORBmatcher::TH_HIGH = 100;      // src/ORBmatcher.cc:38-40
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

namespace {

struct ThreadContext {
    orbm_context* ctx = nullptr;
    ~ThreadContext() {
        if (ctx) orbm_destroy(ctx);
    }
};

[[noreturn]] void fail(const char* what) {
    throw std::runtime_error(std::string(what) + ": " + orbx_last_error_string());
}

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint layout");

// The left (or only) camera of a frame as the searches see it.  A fisheye stereo frame (Nleft != -1) searches
// mvKeys, a rectified / monocular one mvKeysUn (Frame::GetFeaturesInArea, src/Frame.cc:694-696).
orbm_frame_view view_of(const Frame& F) {
    orbm_frame_view v;
    std::memset(&v, 0, sizeof v);
    const bool fisheye = F.Nleft != -1;
    v.keys = reinterpret_cast<const orbx_keypoint*>(fisheye ? F.mvKeys.data() : F.mvKeysUn.data());
    v.desc = F.mDescriptors.data;
    v.n = fisheye ? F.Nleft : F.N;
    v.min_x = Frame::mnMinX;
    v.min_y = Frame::mnMinY;
    v.max_x = Frame::mnMaxX;
    v.max_y = Frame::mnMaxY;
    v.grid_w_inv = Frame::mfGridElementWidthInv;
    v.grid_h_inv = Frame::mfGridElementHeightInv;
    v.scale_factors = F.mvScaleFactors.data();
    v.nlevels = (int)F.mvScaleFactors.size();
    v.u_right = (!fisheye && (int)F.mvuRight.size() == F.N) ? F.mvuRight.data() : nullptr;
    return v;
}

orbm_frame_right right_of(const Frame& F) {
    orbm_frame_right r;
    std::memset(&r, 0, sizeof r);
    r.keys = reinterpret_cast<const orbx_keypoint*>(F.mvKeysRight.data());
    r.desc = F.mDescriptors.data + (size_t)F.Nleft * 32;   // rows Nleft.. of the concatenated matrix (src/Frame.cc:1101)
    r.n = F.Nright;
    r.left_to_right = (int)F.mvLeftToRightMatch.size() == F.Nleft ? F.mvLeftToRightMatch.data() : nullptr;
    r.right_to_left = (int)F.mvRightToLeftMatch.size() == F.Nright ? F.mvRightToLeftMatch.data() : nullptr;
    return r;
}

// Observations() of what every keypoint holds now (-1: nothing), the state the accept loops start from.
std::vector<int32_t> attached(const Frame& F) {
    std::vector<int32_t> obs(F.N > 0 ? F.N : 1, -1);
    for (int i = 0; i < F.N; ++i)
        if (F.mvpMapPoints[i]) obs[i] = F.mvpMapPoints[i]->Observations();
    return obs;
}

void copy_descriptor(uint8_t* dst, const cv::Mat& d) { std::memcpy(dst, d.data, 32); }

}  // namespace

orbm_context* ORBmatcher::Context() {
    static thread_local ThreadContext tc;
    if (!tc.ctx) {
        const char* e = std::getenv("ORBM_DEVICE");
        if (orbm_create(e ? std::atoi(e) : 0, &tc.ctx) != ORBX_OK) fail("orbm_create");
    }
    return tc.ctx;
}

// Bit set count operation from
// http://graphics.stanford.edu/~seander/bithacks.html#CountBitsSetParallel (as the reference says at :2351); any
// exact popcount gives the same integer.
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
    uint32_t wa[8], wb[8];
    std::memcpy(wa, a.data, 32);
    std::memcpy(wb, b.data, 32);
    int dist = 0;
    for (int i = 0; i < 8; ++i) dist += __builtin_popcount(wa[i] ^ wb[i]);
    return dist;
}

// src/ORBmatcher.cc:48-218.  The map points arrive with the tracking fields Frame::isInFrustum left in them.
int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th,
                                   const bool bFarPoints, const float thFarPoints) {
    const int m = (int)vpMapPoints.size();
    const bool fisheye = F.Nleft != -1;
    std::vector<orbm_map_point> mps(m > 0 ? m : 1);
    std::vector<orbm_map_point_right> mpr(fisheye ? (m > 0 ? m : 1) : 0);
    std::vector<uint8_t> desc((size_t)(m > 0 ? m : 1) * 32);
    for (int i = 0; i < m; ++i) {
        MapPoint* p = vpMapPoints[i];
        orbm_map_point& q = mps[i];
        std::memset(&q, 0, sizeof q);
        q.in_view = p->mbTrackInView ? 1 : 0;
        q.proj_x = p->mTrackProjX;
        q.proj_y = p->mTrackProjY;
        q.proj_xr = p->mTrackProjXR;
        q.view_cos = p->mTrackViewCos;
        q.depth = p->mTrackDepth;
        q.level = p->mnTrackScaleLevel;
        q.observations = p->Observations();
        q.bad = p->isBad() ? 1 : 0;
        if (fisheye) {
            orbm_map_point_right& r = mpr[i];
            std::memset(&r, 0, sizeof r);
            r.in_view = p->mbTrackInViewR ? 1 : 0;
            r.proj_x = p->mTrackProjXR;
            r.proj_y = p->mTrackProjYR;
            r.view_cos = p->mTrackViewCosR;
            r.level = p->mnTrackScaleLevelR;
        }
        copy_descriptor(&desc[(size_t)i * 32], p->GetDescriptor());
    }
    std::vector<int32_t> obs = attached(F), match(F.N > 0 ? F.N : 1, -1);
    const orbm_frame_view v = view_of(F);
    int n;
    if (fisheye) {
        const orbm_frame_right r = right_of(F);
        n = orbm_search_by_projection_points_stereo(Context(), &v, &r, mps.data(), mpr.data(), desc.data(), m, th,
                                                    bFarPoints, thFarPoints, mfNNratio, obs.data(), match.data());
    } else {
        n = orbm_search_by_projection_points(Context(), &v, mps.data(), desc.data(), m, th, bFarPoints, thFarPoints,
                                             mfNNratio, obs.data(), match.data());
    }
    if (n < 0) fail("orbm_search_by_projection_points");
    for (int i = 0; i < F.N; ++i)
        if (match[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[match[i]];   // :138, :204
    return n;
}

// src/ORBmatcher.cc:1965-2181.  The projection of the last frame's map points with the current pose (:1975-2013,
// :2087-2092) stays on the host, in the reference's own cv::Mat expressions; the window search, the distances, the
// accept loop and the rotation histogram are the library's.
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat tlc = Rlw * twc + tlw;
    const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;

    const bool fisheye = CurrentFrame.Nleft != -1;
    const int n_last = LastFrame.N;
    std::vector<orbm_last_obs> last(n_last > 0 ? n_last : 1);
    std::vector<float> uv_right(fisheye ? 2 * (size_t)(n_last > 0 ? n_last : 1) : 0);
    std::vector<uint8_t> desc((size_t)(n_last > 0 ? n_last : 1) * 32);
    std::vector<MapPoint*> point(n_last > 0 ? n_last : 1, static_cast<MapPoint*>(NULL));
    for (int i = 0; i < n_last; ++i) {
        orbm_last_obs& o = last[i];
        std::memset(&o, 0, sizeof o);
        MapPoint* pMP = LastFrame.mvpMapPoints[i];
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        cv::Mat x3Dw = pMP->GetWorldPos();
        cv::Mat x3Dc = Rcw * x3Dw + tcw;
        const float invzc = 1.0 / x3Dc.at<float>(2);
        const cv::Point2f uv = CurrentFrame.mpCamera->project(x3Dc);
        const bool left_last = LastFrame.Nleft == -1 || i < LastFrame.Nleft;
        o.valid = 1;
        o.u = uv.x;
        o.v = uv.y;
        o.inv_z = invzc;
        o.octave = left_last ? LastFrame.mvKeys[i].octave : LastFrame.mvKeysRight[i - LastFrame.Nleft].octave;
        o.angle = LastFrame.Nleft == -1 ? LastFrame.mvKeysUn[i].angle
                                        : (left_last ? LastFrame.mvKeys[i].angle : LastFrame.mvKeysRight[i - LastFrame.Nleft].angle);
        o.observations = pMP->Observations();
        if (fisheye) {
            cv::Mat x3Dr = CurrentFrame.mTrl.colRange(0, 3).rowRange(0, 3) * x3Dc + CurrentFrame.mTrl.col(3);
            const cv::Point2f uvr = CurrentFrame.mpCamera->project(x3Dr);
            uv_right[2 * (size_t)i] = uvr.x;
            uv_right[2 * (size_t)i + 1] = uvr.y;
        }
        copy_descriptor(&desc[(size_t)i * 32], pMP->GetDescriptor());
        point[i] = pMP;
    }
    const std::vector<int32_t> before = attached(CurrentFrame);
    std::vector<int32_t> obs = before, match(CurrentFrame.N > 0 ? CurrentFrame.N : 1, -1);
    const orbm_frame_view v = view_of(CurrentFrame);
    int n;
    if (fisheye) {
        const orbm_frame_right r = right_of(CurrentFrame);
        n = orbm_search_by_projection_frames_stereo(Context(), &v, &r, last.data(), uv_right.data(), desc.data(), n_last, th,
                                                    bForward, bBackward, mbCheckOrientation, obs.data(), match.data());
    } else {
        n = orbm_search_by_projection_frames(Context(), &v, last.data(), desc.data(), n_last, th, bForward, bBackward,
                                             CurrentFrame.mbf, mbCheckOrientation, obs.data(), match.data());
    }
    if (n < 0) fail("orbm_search_by_projection_frames");
    for (int i = 0; i < CurrentFrame.N; ++i) {
        if (match[i] >= 0) CurrentFrame.mvpMapPoints[i] = point[match[i]];                      // :2063
        else if (obs[i] < 0 && before[i] >= 0) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint*>(NULL);   // :2172
    }
    return n;
}

namespace {

// One side of SearchByBoW: keypoints (only .angle is read: the reference takes mvKeysUn for a rectified / monocular
// side, mvKeys ++ mvKeysRight for a fisheye one, :394-398, :422-427), descriptors, and the DBoW2::FeatureVector
// (std::map: ascending node ids) as CSR.
struct BowSide {
    std::vector<cv::KeyPoint> keys;
    std::vector<uint32_t> ids, items;
    std::vector<int32_t> ptr;
    orbm_bow_view view;
    BowSide(const std::vector<cv::KeyPoint>& un, const std::vector<cv::KeyPoint>& left, const std::vector<cv::KeyPoint>& right,
            bool fisheye, int nleft, const cv::Mat& desc, int n, const DBoW2::FeatureVector& fv) {
        if (!fisheye) {
            keys = un;
        } else {
            keys.assign(left.begin(), left.begin() + nleft);
            keys.insert(keys.end(), right.begin(), right.end());
        }
        keys.resize(n > 0 ? n : 1);
        ptr.push_back(0);
        for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
            ids.push_back(it->first);
            items.insert(items.end(), it->second.begin(), it->second.end());
            ptr.push_back((int32_t)items.size());
        }
        if (ids.empty()) ids.push_back(0);
        if (items.empty()) items.push_back(0);
        std::memset(&view, 0, sizeof view);
        view.keys = reinterpret_cast<const orbx_keypoint*>(keys.data());
        view.desc = desc.data;
        view.n = n;
        view.nnodes = (int)ptr.size() - 1;
        view.node_ids = ids.data();
        view.node_ptr = ptr.data();
        view.node_items = items.data();
    }
};

}  // namespace

// src/ORBmatcher.cc:273-475
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
    const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    const int nkf = (int)vpMapPointsKF.size();
    std::vector<uint8_t> ok(nkf > 0 ? nkf : 1, 0);
    for (int i = 0; i < nkf; ++i) ok[i] = (vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad()) ? 1 : 0;   // :302-308
    const BowSide K(pKF->mvKeysUn, pKF->mvKeys, pKF->mvKeysRight, pKF->mpCamera2 != NULL, pKF->NLeft, pKF->mDescriptors, nkf,
                    pKF->mFeatVec);
    const BowSide S(F.mvKeysUn, F.mvKeys, F.mvKeysRight, F.Nleft != -1, F.Nleft, F.mDescriptors, F.N,
                    F.mFeatVec);
    std::vector<int32_t> matches(F.N > 0 ? F.N : 1, -1);
    const int n = orbm_search_by_bow(Context(), &K.view, ok.data(), &S.view, F.Nleft, mfNNratio, mbCheckOrientation,
                                     matches.data());
    if (n < 0) fail("orbm_search_by_bow");
    for (int j = 0; j < F.N; ++j)
        if (matches[j] >= 0) vpMapPointMatches[j] = vpMapPointsKF[matches[j]];   // :361, :384
    return n;
}

// src/ORBmatcher.cc:827-967
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
    const std::vector<MapPoint*> vp1 = pKF1->GetMapPointMatches(), vp2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPoint*>(vp1.size(), static_cast<MapPoint*>(NULL));
    const int n1 = (int)vp1.size(), n2 = (int)vp2.size();
    std::vector<uint8_t> ok1(n1 > 0 ? n1 : 1, 0), ok2(n2 > 0 ? n2 : 1, 0);
    for (int i = 0; i < n1; ++i) ok1[i] = (vp1[i] && !vp1[i]->isBad()) ? 1 : 0;   // :868-872
    for (int i = 0; i < n2; ++i) ok2[i] = (vp2[i] && !vp2[i]->isBad()) ? 1 : 0;   // :892-898
    const std::vector<cv::KeyPoint> none;
    const BowSide A(pKF1->mvKeysUn, none, none, false, 0, pKF1->mDescriptors, n1, pKF1->mFeatVec);
    const BowSide B(pKF2->mvKeysUn, none, none, false, 0, pKF2->mDescriptors, n2, pKF2->mFeatVec);
    const int n_un1 = pKF1->NLeft != -1 ? (int)pKF1->mvKeysUn.size() : -1;   // :863-865
    const int n_un2 = pKF2->NLeft != -1 ? (int)pKF2->mvKeysUn.size() : -1;   // :887-889
    std::vector<int32_t> m12(n1 > 0 ? n1 : 1, -1);
    const int n = orbm_search_by_bow_kf(Context(), &A.view, ok1.data(), n_un1, &B.view, ok2.data(), n_un2, mfNNratio,
                                        mbCheckOrientation, m12.data());
    if (n < 0) fail("orbm_search_by_bow_kf");
    for (int i = 0; i < n1; ++i)
        if (m12[i] >= 0) vpMatches12[i] = vp2[m12[i]];   // :925
    return n;
}

// src/ORBmatcher.cc:710-825
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                        std::vector<int>& vnMatches12, int windowSize) {
    std::vector<float> prev(2 * (size_t)(F1.N > 0 ? F1.N : 1));
    for (int i = 0; i < F1.N; ++i) {
        prev[2 * (size_t)i] = vbPrevMatched[i].x;
        prev[2 * (size_t)i + 1] = vbPrevMatched[i].y;
    }
    std::vector<int32_t> m12(F1.N > 0 ? F1.N : 1, -1);
    const orbm_frame_view v1 = view_of(F1), v2 = view_of(F2);
    const int n = orbm_search_for_initialization(Context(), &v1, &v2, prev.data(), m12.data(), windowSize, mfNNratio,
                                                 mbCheckOrientation);
    if (n < 0) fail("orbm_search_for_initialization");
    vnMatches12 = std::vector<int>(m12.begin(), m12.begin() + F1.N);   // :713
    for (int i = 0; i < F1.N; ++i) vbPrevMatched[i] = cv::Point2f(prev[2 * (size_t)i], prev[2 * (size_t)i + 1]);   // :819-821
    return n;
}

}  // namespace ORB_SLAM3
