// oracle/ref_match_capi.cpp -- TEST INFRASTRUCTURE ONLY.
//
// C wrapper around the reference's OWN matcher code, compiled VERBATIM from /root/reference by
// oracle/Makefile (target `mref` -> oracle/_ref/liborbmref.so; never copied into this repo):
//     src/ORBmatcher.cc  src/Frame.cc  src/MapPoint.cc  src/ORBextractor.cc  src/CameraModels/Pinhole.cpp
//     Thirdparty/DBoW2/DBoW2/{FORB,BowVector,FeatureVector,ScoringObject}.cpp (+ TemplatedVocabulary.h)
// against oracle/cvshim (cv::) and oracle/matchshim/facade.h (KeyFrame / Map / IMU stand-ins).
//
// Every orbr_* entry point takes the same POD snapshot as the matching orbo_* function of
// oracle/match_oracle.cpp (and as the product's C ABI, include/orbm.h), builds REAL ORB_SLAM3::Frame /
// MapPoint / KeyFrame objects from it, calls the reference method, and reads the result back out of the
// objects.  The searches that do geometry before they search (SearchByProjection(Cur, Last), the Sim3 /
// reloc / Fuse / SearchBySim3 projections, SearchForTriangulation's epipole) are driven with identity
// poses and a camera whose project() is a table lookup keyed by the point's x coordinate, so the
// reference's own arithmetic reproduces the snapshot's (u, v, level) EXACTLY and its gates all pass;
// what is compared is everything after that: candidate windows, order, distances, accept logic.
//
// This file is compiled with -fno-access-control (private members such as Frame::AssignFeaturesToGrid
// and MapPoint::mDescriptor are set directly); the reference files are compiled with their own access
// rules.
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>

#include "Frame.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#include "ORBextractor.h"
#include "CameraModels/Pinhole.h"

using namespace ORB_SLAM3;

// ---- bodies of the facade's stand-ins ---------------------------------------------------------------
namespace ORB_SLAM3 {

long unsigned int KeyFrame::nNextId = 0;

// src/KeyFrame.cc:49-100: member-wise copy of the frame, grid included
KeyFrame::KeyFrame(Frame& F, Map* pMap)
    : mnFrameId(F.mnId), mnGridCols(FRAME_GRID_COLS), mnGridRows(FRAME_GRID_ROWS),
      mfGridElementWidthInv(F.mfGridElementWidthInv), mfGridElementHeightInv(F.mfGridElementHeightInv),
      fx(F.fx), fy(F.fy), cx(F.cx), cy(F.cy), invfx(F.invfx), invfy(F.invfy), mbf(F.mbf), mb(F.mb),
      mThDepth(F.mThDepth), N(F.N), mvKeys(F.mvKeys), mvKeysUn(F.mvKeysUn), mvuRight(F.mvuRight), mvDepth(F.mvDepth),
      mDescriptors(F.mDescriptors.clone()), mBowVec(F.mBowVec), mFeatVec(F.mFeatVec), mnScaleLevels(F.mnScaleLevels),
      mfScaleFactor(F.mfScaleFactor), mfLogScaleFactor(F.mfLogScaleFactor), mvScaleFactors(F.mvScaleFactors),
      mvLevelSigma2(F.mvLevelSigma2), mvInvLevelSigma2(F.mvInvLevelSigma2), mnMinX(F.mnMinX), mnMinY(F.mnMinY),
      mnMaxX(F.mnMaxX), mnMaxY(F.mnMaxY), mvpMapPoints(F.mvpMapPoints), mbBad(false), mpMap(pMap), mpCamera(F.mpCamera),
      mpCamera2(F.mpCamera2), mvLeftToRightMatch(F.mvLeftToRightMatch), mvRightToLeftMatch(F.mvRightToLeftMatch),
      mTlr(F.mTlr.clone()), mTrl(F.mTrl), mvKeysRight(F.mvKeysRight), NLeft(F.Nleft), NRight(F.Nright) {
    mnId = nNextId++;
    mGrid.resize(mnGridCols);
    if (F.Nleft != -1) mGridRight.resize(mnGridCols);
    for (int i = 0; i < mnGridCols; i++) {
        mGrid[i].resize(mnGridRows);
        if (F.Nleft != -1) mGridRight[i].resize(mnGridRows);
        for (int j = 0; j < mnGridRows; j++) {
            mGrid[i][j] = F.mGrid[i][j];
            if (F.Nleft != -1) mGridRight[i][j] = F.mGridRight[i][j];
        }
    }
    SetPose(F.mTcw.empty() ? cv::Mat::eye(4, 4, CV_32F) : F.mTcw);
}

// src/KeyFrame.cc:118-135 (camera part)
void KeyFrame::SetPose(const cv::Mat& Tcw_) {
    Tcw_.copyTo(Tcw);
    cv::Mat Rcw = Tcw.rowRange(0, 3).colRange(0, 3);
    cv::Mat tcw = Tcw.rowRange(0, 3).col(3);
    cv::Mat Rwc = Rcw.t();
    Ow = -Rwc * tcw;
    Twc = cv::Mat::eye(4, 4, Tcw.type());
    Rwc.copyTo(Twc.rowRange(0, 3).colRange(0, 3));
    Ow.copyTo(Twc.rowRange(0, 3).col(3));
}

// src/KeyFrame.cc:1176-1260
cv::Mat KeyFrame::GetRightPose() {
    cv::Mat Rrl = mTlr.rowRange(0, 3).colRange(0, 3).t();
    cv::Mat Rlw = Tcw.rowRange(0, 3).colRange(0, 3).clone();
    cv::Mat Rrw = Rrl * Rlw;
    cv::Mat tlw = Tcw.rowRange(0, 3).col(3).clone();
    cv::Mat trl = -Rrl * mTlr.rowRange(0, 3).col(3);
    cv::Mat trw = Rrl * tlw + trl;
    cv::Mat Trw;
    cv::hconcat(Rrw, trw, Trw);
    return Trw.clone();
}
cv::Mat KeyFrame::GetRightCameraCenter() {
    cv::Mat Rwl = Tcw.rowRange(0, 3).colRange(0, 3).t();
    cv::Mat tlr = mTlr.rowRange(0, 3).col(3);
    cv::Mat twl = Ow.clone();
    cv::Mat twr = Rwl * tlr + twl;
    return twr.clone();
}
cv::Mat KeyFrame::GetRightRotation() {
    cv::Mat Rrl = mTlr.rowRange(0, 3).colRange(0, 3).t();
    cv::Mat Rlw = Tcw.rowRange(0, 3).colRange(0, 3).clone();
    cv::Mat Rrw = Rrl * Rlw;
    return Rrw.clone();
}
cv::Mat KeyFrame::GetRightTranslation() {
    cv::Mat Rrl = mTlr.rowRange(0, 3).colRange(0, 3).t();
    cv::Mat tlw = Tcw.rowRange(0, 3).col(3).clone();
    cv::Mat trl = -Rrl * mTlr.rowRange(0, 3).col(3);
    cv::Mat trw = Rrl * tlw + trl;
    return trw.clone();
}

// src/KeyFrame.cc:700-712
void KeyFrame::EraseMapPointMatch(MapPoint* pMP) {
    tuple<size_t, size_t> indexes = pMP->GetIndexInKeyFrame(this);
    size_t leftIndex = get<0>(indexes), rightIndex = get<1>(indexes);
    if (leftIndex != (size_t)-1) mvpMapPoints[leftIndex] = static_cast<MapPoint*>(NULL);
    if (rightIndex != (size_t)-1) mvpMapPoints[rightIndex] = static_cast<MapPoint*>(NULL);
}
std::set<MapPoint*> KeyFrame::GetMapPoints() {
    std::set<MapPoint*> s;
    for (size_t i = 0, iend = mvpMapPoints.size(); i < iend; i++) {
        if (!mvpMapPoints[i]) continue;
        MapPoint* pMP = mvpMapPoints[i];
        if (!pMP->isBad()) s.insert(pMP);
    }
    return s;
}

// src/KeyFrame.cc:770-814: the window query of a keyframe (no octave gates; bounds are the ints above)
std::vector<size_t> KeyFrame::GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight) const {
    std::vector<size_t> vIndices;
    vIndices.reserve(N);
    float factorX = r;
    float factorY = r;
    const int nMinCellX = max(0, (int)floor((x - mnMinX - factorX) * mfGridElementWidthInv));
    if (nMinCellX >= mnGridCols) return vIndices;
    const int nMaxCellX = min((int)mnGridCols - 1, (int)ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = max(0, (int)floor((y - mnMinY - factorY) * mfGridElementHeightInv));
    if (nMinCellY >= mnGridRows) return vIndices;
    const int nMaxCellY = min((int)mnGridRows - 1, (int)ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
    if (nMaxCellY < 0) return vIndices;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<size_t> vCell = (!bRight) ? mGrid[ix][iy] : mGridRight[ix][iy];
            for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                const cv::KeyPoint& kpUn = (NLeft == -1) ? mvKeysUn[vCell[j]] : (!bRight) ? mvKeys[vCell[j]] : mvKeysRight[vCell[j]];
                const float distx = kpUn.pt.x - x;
                const float disty = kpUn.pt.y - y;
                if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
            }
        }
    }
    return vIndices;
}
bool KeyFrame::IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }

// The fisheye camera model is out of scope (cv::fisheye, SVD); nothing in the oracle reaches it.
#define KB8_ABSENT(sig) sig { fprintf(stderr, "matchshim: KannalaBrandt8 is not part of the oracle\n"); abort(); }
KB8_ABSENT(cv::Point2f KannalaBrandt8::project(const cv::Point3f&))
KB8_ABSENT(cv::Point2f KannalaBrandt8::project(const cv::Mat&))
KB8_ABSENT(Eigen::Vector2d KannalaBrandt8::project(const Eigen::Vector3d&))
KB8_ABSENT(cv::Mat KannalaBrandt8::projectMat(const cv::Point3f&))
KB8_ABSENT(float KannalaBrandt8::uncertainty2(const Eigen::Matrix<double, 2, 1>&))
KB8_ABSENT(cv::Point3f KannalaBrandt8::unproject(const cv::Point2f&))
KB8_ABSENT(cv::Mat KannalaBrandt8::unprojectMat(const cv::Point2f&))
KB8_ABSENT(cv::Mat KannalaBrandt8::projectJac(const cv::Point3f&))
typedef Eigen::Matrix<double, 2, 3> EigenMat23;
KB8_ABSENT(EigenMat23 KannalaBrandt8::projectJac(const Eigen::Vector3d&))
KB8_ABSENT(cv::Mat KannalaBrandt8::unprojectJac(const cv::Point2f&))
KB8_ABSENT(bool KannalaBrandt8::ReconstructWithTwoViews(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&,
                                                        const std::vector<int>&, cv::Mat&, cv::Mat&, std::vector<cv::Point3f>&,
                                                        std::vector<bool>&))
KB8_ABSENT(cv::Mat KannalaBrandt8::toK())
KB8_ABSENT(bool KannalaBrandt8::epipolarConstrain(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&, const cv::Mat&,
                                                  const cv::Mat&, const float, const float))
KB8_ABSENT(bool KannalaBrandt8::matchAndtriangulate(const cv::KeyPoint&, const cv::KeyPoint&, GeometricCamera*, cv::Mat&, cv::Mat&,
                                                    const float, const float, cv::Mat&))
KB8_ABSENT(float KannalaBrandt8::TriangulateMatches(GeometricCamera*, const cv::KeyPoint&, const cv::KeyPoint&, const cv::Mat&,
                                                    const cv::Mat&, const float, const float, cv::Mat&))

}  // namespace ORB_SLAM3

namespace {

// ---- cameras that turn the reference's projection calls into table lookups ---------------------------
// A point handed to project() carries its own index in its x coordinate (exact in float32 below 2^24);
// indices >= kRightBase address the second table (the right camera of a fisheye rig, reached by the
// reference through mTrl's translation).  The pair tests (epipolarConstrain / matchAndtriangulate) look
// the pair up by the keypoints' class_id, which the wrapper sets to the feature index.
const int kRightBase = 1 << 20;

struct TableCamera : GeometricCamera {
    const float* uv = nullptr;    // 2 floats per point
    const float* uv_r = nullptr;  // right table
    const uint8_t* pair_ok = nullptr;
    int n2 = 0;
    bool pair_default = true;
    cv::Point2f lookup(float x) const {
        const int i = (int)x;
        if (i >= kRightBase) return cv::Point2f(uv_r[2 * (i - kRightBase)], uv_r[2 * (i - kRightBase) + 1]);
        return cv::Point2f(uv[2 * i], uv[2 * i + 1]);
    }
    cv::Point2f project(const cv::Point3f& p3D) override { return lookup(p3D.x); }
    cv::Point2f project(const cv::Mat& m3D) override { return lookup(m3D.at<float>(0)); }
    bool pair(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2) const {
        return pair_ok ? pair_ok[(size_t)kp1.class_id * n2 + kp2.class_id] != 0 : pair_default;
    }
    bool epipolarConstrain(GeometricCamera*, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat&, const cv::Mat&,
                           const float, const float) override { return pair(kp1, kp2); }
    bool matchAndtriangulate(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, GeometricCamera*, cv::Mat&, cv::Mat&, const float,
                             const float, cv::Mat& x3D) override {
        x3D = cv::Mat::zeros(3, 1, CV_32F);
        return pair(kp1, kp2);
    }
    static void absent() { fprintf(stderr, "ref_match_capi: TableCamera method not used by the matcher\n"); abort(); }
    Eigen::Vector2d project(const Eigen::Vector3d&) override { absent(); return Eigen::Vector2d(); }
    cv::Mat projectMat(const cv::Point3f&) override { absent(); return cv::Mat(); }
    float uncertainty2(const Eigen::Matrix<double, 2, 1>&) override { return 1.f; }
    cv::Point3f unproject(const cv::Point2f&) override { absent(); return cv::Point3f(); }
    cv::Mat unprojectMat(const cv::Point2f&) override { absent(); return cv::Mat(); }
    cv::Mat projectJac(const cv::Point3f&) override { absent(); return cv::Mat(); }
    Eigen::Matrix<double, 2, 3> projectJac(const Eigen::Vector3d&) override { absent(); return Eigen::Matrix<double, 2, 3>(); }
    cv::Mat unprojectJac(const cv::Point2f&) override { absent(); return cv::Mat(); }
    bool ReconstructWithTwoViews(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&, const std::vector<int>&, cv::Mat&,
                                 cv::Mat&, std::vector<cv::Point3f>&, std::vector<bool>&) override { absent(); return false; }
    cv::Mat toK() override { absent(); return cv::Mat(); }
};

// The reference's own Pinhole (src/CameraModels/Pinhole.cpp compiled verbatim: epipolarConstrain :129-150),
// with only the epipole projection replaced by a fixed point.
struct PinholeEp : Pinhole {
    cv::Point2f ep;
    PinholeEp(const std::vector<float>& p, cv::Point2f e) : Pinhole(p), ep(e) {}
    cv::Point2f project(const cv::Mat&) override { return ep; }
};

}  // namespace

// ---- POD snapshots: the layouts of include/orbm.h (and of oracle/match_oracle.cpp) ---------------------
namespace {

struct Kp {  // cv::KeyPoint layout
    float x, y, size, angle, response;
    int32_t octave, class_id;
};
static_assert(sizeof(cv::KeyPoint) == sizeof(Kp), "KeyPoint layout");
struct FrameView {  // orbm_frame_view
    const Kp* keys;
    const uint8_t* desc;
    int32_t n;
    float min_x, min_y, max_x, max_y;
    float grid_w_inv, grid_h_inv;
    const float* scale_factors;
    int32_t nlevels;
    const float* u_right;
};
struct MapPointPod {  // orbm_map_point
    float proj_x, proj_y, proj_xr, view_cos, depth;
    int32_t level, observations;
    uint8_t in_view, bad, pad[2];
};
struct FrameRight {  // orbm_frame_right
    const Kp* keys;
    const uint8_t* desc;
    int32_t n;
    const int32_t* left_to_right;
    const int32_t* right_to_left;
};
struct MapPointRightPod {  // orbm_map_point_right
    float proj_x, proj_y, view_cos;
    int32_t level;
    uint8_t in_view, pad[3];
};
struct BowView {  // orbm_bow_view
    const Kp* keys;
    const uint8_t* desc;
    int32_t n, nnodes;
    const uint32_t* node_ids;
    const int32_t* node_ptr;
    const uint32_t* node_items;
};
struct LastObs {  // orbm_last_obs
    float u, v, inv_z, angle;
    int32_t octave, observations;
    uint8_t valid, pad[3];
};
struct ProjPoint {  // orbm_proj_point
    float u, v, ur, angle;
    int32_t level;
    uint8_t valid, pad[3];
};

// Frame's image bounds and grid pitch are STATIC members in the reference (include/Frame.h:196-197, 245-248):
// one wrapper call at a time.
std::mutex g_mu;

std::vector<cv::KeyPoint> to_keypoints(const Kp* k, int n) {
    std::vector<cv::KeyPoint> v(n);
    if (n) memcpy((void*)v.data(), k, (size_t)n * sizeof(Kp));
    for (int i = 0; i < n; ++i) v[i].class_id = i;  // the pair tests of TableCamera identify features by it
    return v;
}
cv::Mat to_desc(const uint8_t* d, int n) {
    cv::Mat m(n, 32, CV_8UC1);
    if (n) memcpy(m.data, d, (size_t)n * 32);
    return m;
}

void set_scales(Frame& F, const float* sf, int nlevels) {
    F.mnScaleLevels = nlevels;
    F.mvScaleFactors.assign(sf, sf + nlevels);
    F.mvInvScaleFactors.resize(nlevels);
    F.mvLevelSigma2.resize(nlevels);
    F.mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; ++i) {  // src/ORBextractor.cc:413-429
        F.mvLevelSigma2[i] = sf[i] * sf[i];
        F.mvInvScaleFactors[i] = 1.0f / sf[i];
        F.mvInvLevelSigma2[i] = 1.0f / F.mvLevelSigma2[i];
    }
    F.mfScaleFactor = nlevels > 1 ? sf[1] : 1.2f;
    F.mfLogScaleFactor = log(F.mfScaleFactor);  // src/Frame.cc:96
}

// A Frame as its constructors leave it (src/Frame.cc:88-190, 1034-1124), from a snapshot: keypoints,
// descriptors, bounds, scale tables, then the reference's own AssignFeaturesToGrid.  R != NULL: a fisheye
// stereo frame (Nleft = v.n, Nright = R->n, descriptors concatenated as :1101 does).
void fill_frame(Frame& F, const FrameView& v, const FrameRight* R = nullptr) {
    Frame::mnMinX = v.min_x;
    Frame::mnMinY = v.min_y;
    Frame::mnMaxX = v.max_x;
    Frame::mnMaxY = v.max_y;
    Frame::mfGridElementWidthInv = v.grid_w_inv;
    Frame::mfGridElementHeightInv = v.grid_h_inv;
    Frame::fx = Frame::fy = 1.f;
    Frame::cx = Frame::cy = 0.f;
    Frame::invfx = Frame::invfy = 1.f;
    F.mnId = Frame::nNextId++;
    F.mvKeys = to_keypoints(v.keys, v.n);
    F.mvKeysUn = F.mvKeys;
    F.mDescriptors = to_desc(v.desc, v.n);
    set_scales(F, v.scale_factors, v.nlevels);
    F.mbf = 0.f;
    F.mb = 1.f;
    F.mThDepth = 0.f;
    F.mpCamera = F.mpCamera2 = nullptr;
    F.mpReferenceKF = nullptr;
    if (!R) {
        F.N = v.n;
        F.Nleft = F.Nright = -1;
        F.mvuRight.assign(v.n, -1.0f);
        if (v.u_right) F.mvuRight.assign(v.u_right, v.u_right + v.n);
    } else {
        F.Nleft = v.n;
        F.Nright = R->n;
        F.N = v.n + R->n;
        F.mvKeysRight = to_keypoints(R->keys, R->n);
        F.mDescriptorsRight = to_desc(R->desc, R->n);
        cv::vconcat(F.mDescriptors, F.mDescriptorsRight, F.mDescriptors);  // src/Frame.cc:1101
        F.mvLeftToRightMatch.assign(v.n, -1);
        F.mvRightToLeftMatch.assign(R->n, -1);
        if (R->left_to_right) F.mvLeftToRightMatch.assign(R->left_to_right, R->left_to_right + v.n);
        if (R->right_to_left) F.mvRightToLeftMatch.assign(R->right_to_left, R->right_to_left + R->n);
        F.mvuRight.assign(v.n, -1.0f);
        F.mTrl = cv::Mat::zeros(3, 4, CV_32F);
        F.mTlr = cv::Mat::zeros(3, 4, CV_32F);
        for (int i = 0; i < 3; ++i) F.mTrl.at<float>(i, i) = F.mTlr.at<float>(i, i) = 1.f;
    }
    F.mvDepth.assign(F.mvuRight.size(), -1.0f);
    F.mvpMapPoints.assign(F.N, static_cast<MapPoint*>(NULL));
    F.mvbOutlier.assign(F.N, false);
    F.AssignFeaturesToGrid();  // src/Frame.cc:377-408 (private there)
}

void set_desc(MapPoint& mp, const uint8_t* d) {
    mp.mDescriptor = cv::Mat(1, 32, CV_8UC1);
    memcpy(mp.mDescriptor.data, d, 32);
}

// kp_obs (in): Observations() of the map point attached to keypoint i, -1 = none
void attach_existing(Frame& F, MapPoint* att, const int32_t* kp_obs) {
    for (int i = 0; i < F.N; ++i) {
        att[i].nObs = kp_obs[i];
        F.mvpMapPoints[i] = kp_obs[i] >= 0 ? &att[i] : static_cast<MapPoint*>(NULL);
    }
}
// kp_match (out): index into `mp` of the point this call attached; kp_obs (out): Observations() of whatever is attached
void read_attachments(Frame& F, MapPoint* mp, int m, int32_t* kp_obs, int32_t* kp_match) {
    for (int i = 0; i < F.N; ++i) {
        MapPoint* p = F.mvpMapPoints[i];
        kp_match[i] = (p >= mp && p < mp + m) ? (int32_t)(p - mp) : -1;
        kp_obs[i] = p ? p->Observations() : -1;
    }
}

cv::Mat eye4() { return cv::Mat::eye(4, 4, CV_32F); }

}  // namespace

extern "C" {

// ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2353-2369
int orbr_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    return ORBmatcher::DescriptorDistance(to_desc(a, 1), to_desc(b, 1));
}

// Frame::AssignFeaturesToGrid + Frame::GetFeaturesInArea, src/Frame.cc:377-408, 645-726
int orbr_features_in_area(const FrameView* Fp, float x, float y, float r, int minLevel, int maxLevel, int32_t* out, int cap) {
    std::lock_guard<std::mutex> lk(g_mu);
    Frame F;
    fill_frame(F, *Fp);
    std::vector<size_t> v = F.GetFeaturesInArea(x, y, r, minLevel, maxLevel);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int32_t)v[i];
    return (int)v.size();
}

// SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints), src/ORBmatcher.cc:48-218.
// Rp / mps_r NULL: monocular / rectified frame (Nleft == -1); else the fisheye stereo form.
int orbr_search_by_projection_points_stereo(const FrameView* Fp, const FrameRight* Rp, const MapPointPod* mps,
                                            const MapPointRightPod* mps_r, const uint8_t* mp_desc, int m, float th,
                                            int bFarPoints, float thFarPoints, float mfNNratio, int32_t* kp_obs,
                                            int32_t* kp_match) {
    std::lock_guard<std::mutex> lk(g_mu);
    Frame F;
    fill_frame(F, *Fp, Rp);
    std::unique_ptr<MapPoint[]> mp(new MapPoint[m > 0 ? m : 1]), att(new MapPoint[F.N > 0 ? F.N : 1]);
    std::vector<MapPoint*> vp(m);
    for (int i = 0; i < m; ++i) {
        MapPoint& p = mp[i];
        p.mbTrackInView = mps[i].in_view != 0;
        p.mTrackProjX = mps[i].proj_x;
        p.mTrackProjY = mps[i].proj_y;
        p.mTrackProjXR = mps[i].proj_xr;
        p.mTrackViewCos = mps[i].view_cos;
        p.mTrackDepth = mps[i].depth;
        p.mnTrackScaleLevel = mps[i].level;
        p.nObs = mps[i].observations;
        p.mbBad = mps[i].bad != 0;
        p.mbTrackInViewR = false;
        p.mnTrackScaleLevelR = -1;
        if (mps_r) {
            p.mbTrackInViewR = mps_r[i].in_view != 0;
            p.mTrackProjXR = mps_r[i].proj_x;
            p.mTrackProjYR = mps_r[i].proj_y;
            p.mTrackViewCosR = mps_r[i].view_cos;
            p.mnTrackScaleLevelR = mps_r[i].level;
        }
        set_desc(p, mp_desc + (size_t)i * 32);
        vp[i] = &p;
    }
    attach_existing(F, att.get(), kp_obs);
    ORBmatcher matcher(mfNNratio, true);
    const int n = matcher.SearchByProjection(F, vp, th, bFarPoints != 0, thFarPoints);
    read_attachments(F, mp.get(), m, kp_obs, kp_match);
    return n;
}

int orbr_search_by_projection_points(const FrameView* Fp, const MapPointPod* mps, const uint8_t* mp_desc, int m, float th,
                                     int bFarPoints, float thFarPoints, float mfNNratio, int32_t* kp_obs, int32_t* kp_match) {
    return orbr_search_by_projection_points_stereo(Fp, nullptr, mps, nullptr, mp_desc, m, th, bFarPoints, thFarPoints, mfNNratio,
                                                   kp_obs, kp_match);
}

// SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono), src/ORBmatcher.cc:1965-2181.
// Both poses are the identity (LastFrame's translation carries the forward / backward decision of :1983-1986),
// map point i sits at (i, 0, 1 / inv_z): x3Dc == x3Dw, the camera returns (u, v) of observation i, and
// invzc = 1.0 / z reproduces inv_z -- the wrapper refuses (-100) inputs whose inv_z does not survive
// float(1 / float(1 / inv_z)); the tests canonicalise their inputs that way.  uv_right != NULL with Rp: the
// fisheye form (:2087-2155), reached through mTrl's translation (kRightBase, 0, 0).
static int sbp_frames(const FrameView* Cp, const FrameRight* Rp, const LastObs* last, const float* uv_right,
                      const uint8_t* last_desc, int n_last, float th, int bForward, int bBackward, float mbf,
                      int mbCheckOrientation, int32_t* kp_obs, int32_t* kp_match) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (n_last >= kRightBase) return -101;
    Frame Cur, Last;
    fill_frame(Cur, *Cp, Rp);
    Cur.mbf = mbf;
    Cur.mb = 1.f;
    Cur.SetPose(eye4());
    std::vector<float> uv(2 * (size_t)n_last);
    TableCamera cam;
    cam.uv = uv.data();
    cam.uv_r = uv_right;
    Cur.mpCamera = &cam;
    if (Rp) Cur.mTrl.at<float>(0, 3) = (float)kRightBase;
    // LastFrame: only N, mvpMapPoints, mvbOutlier, mvKeys[i].octave, mvKeysUn[i].angle, Nleft and mTcw are read
    Last.N = n_last;
    Last.Nleft = Last.Nright = -1;
    Last.mvKeys.resize(n_last);
    Last.mvpMapPoints.assign(n_last, static_cast<MapPoint*>(NULL));
    Last.mvbOutlier.assign(n_last, false);
    std::unique_ptr<MapPoint[]> mp(new MapPoint[n_last > 0 ? n_last : 1]), att(new MapPoint[Cur.N > 0 ? Cur.N : 1]);
    for (int i = 0; i < n_last; ++i) {
        const LastObs& o = last[i];
        uv[2 * i] = o.u;
        uv[2 * i + 1] = o.v;
        Last.mvKeys[i].octave = o.octave;
        Last.mvKeys[i].angle = o.angle;
        if (!o.valid) continue;
        const float z = (float)(1.0 / (double)o.inv_z);
        if ((float)(1.0 / z) != o.inv_z || !std::isfinite(z)) return -100;
        mp[i].mWorldPos = (cv::Mat_<float>(3, 1) << (float)i, 0.f, z);
        mp[i].nObs = o.observations;
        set_desc(mp[i], last_desc + (size_t)i * 32);
        Last.mvpMapPoints[i] = &mp[i];
    }
    Last.mvKeysUn = Last.mvKeys;
    cv::Mat Tlw = eye4();
    Tlw.at<float>(2, 3) = bForward ? 2.f : bBackward ? -2.f : 0.f;  // tlc.z vs CurrentFrame.mb = 1
    Last.SetPose(Tlw);
    attach_existing(Cur, att.get(), kp_obs);
    ORBmatcher matcher(0.9f, mbCheckOrientation != 0);
    const int n = matcher.SearchByProjection(Cur, Last, th, false);
    read_attachments(Cur, mp.get(), n_last, kp_obs, kp_match);
    return n;
}

int orbr_search_by_projection_frames(const FrameView* Cp, const LastObs* last, const uint8_t* last_desc, int n_last, float th,
                                     int bForward, int bBackward, float mbf, int mbCheckOrientation, int32_t* kp_obs,
                                     int32_t* kp_match) {
    return sbp_frames(Cp, nullptr, last, nullptr, last_desc, n_last, th, bForward, bBackward, mbf, mbCheckOrientation, kp_obs,
                      kp_match);
}
int orbr_search_by_projection_frames_stereo(const FrameView* Cp, const FrameRight* Rp, const LastObs* last, const float* uv_right,
                                            const uint8_t* last_desc, int n_last, float th, int bForward, int bBackward,
                                            int mbCheckOrientation, int32_t* kp_obs, int32_t* kp_match) {
    return sbp_frames(Cp, Rp, last, uv_right, last_desc, n_last, th, bForward, bBackward, 0.f, mbCheckOrientation, kp_obs, kp_match);
}

// SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), src/ORBmatcher.cc:710-825
int orbr_search_for_initialization(const FrameView* F1p, const FrameView* F2p, float* vbPrevMatched, int32_t* vnMatches12,
                                   int windowSize, float mfNNratio, int mbCheckOrientation) {
    std::lock_guard<std::mutex> lk(g_mu);
    Frame F1, F2;
    fill_frame(F1, *F1p);
    fill_frame(F2, *F2p);
    std::vector<cv::Point2f> prev(F1.N);
    for (int i = 0; i < F1.N; ++i) prev[i] = cv::Point2f(vbPrevMatched[2 * i], vbPrevMatched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const int n = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
    for (int i = 0; i < F1.N; ++i) {
        vnMatches12[i] = m12[i];
        vbPrevMatched[2 * i] = prev[i].x;
        vbPrevMatched[2 * i + 1] = prev[i].y;
    }
    return n;
}

// Frame::ComputeStereoMatches, src/Frame.cc:802-980, on two ORBextractor objects whose mvImagePyramid are
// views of the caller's padded pyramids (pyr_l[l] / pyr_r[l] point at the ROI origin; the 19-px border is
// addressable around it, as in the reference's buffers).
int orbr_compute_stereo_matches(const Kp* mvKeys, const uint8_t* mDescriptors, int N, const Kp* mvKeysRight,
                                const uint8_t* mDescriptorsRight, int Nr, const float* mvScaleFactors,
                                const float* mvInvScaleFactors, const uint8_t* const* pyr_l, const int* stride_l,
                                const uint8_t* const* pyr_r, const int* stride_r, const int* cols_r, int n_rows, float mb,
                                float mbf, float* mvuRight, float* mvDepth, int nlevels) {
    std::lock_guard<std::mutex> lk(g_mu);
    Frame F;
    F.N = N;
    F.mvKeys = to_keypoints(mvKeys, N);
    F.mvKeysRight = to_keypoints(mvKeysRight, Nr);
    F.mDescriptors = to_desc(mDescriptors, N);
    F.mDescriptorsRight = to_desc(mDescriptorsRight, Nr);
    set_scales(F, mvScaleFactors, nlevels);
    for (int i = 0; i < nlevels; ++i) F.mvInvScaleFactors[i] = mvInvScaleFactors[i];
    F.mb = mb;
    F.mbf = mbf;
    ORBextractor exl(1000, 1.2f, nlevels, 20, 7), exr(1000, 1.2f, nlevels, 20, 7);
    for (int l = 0; l < nlevels; ++l) {
        const int rows = l == 0 ? n_rows : 1 << 20;  // only level 0's rows and the right levels' cols are read as sizes
        exl.mvImagePyramid[l] = cv::Mat(rows, 1 << 20, CV_8UC1, (void*)pyr_l[l], (size_t)stride_l[l]);
        exr.mvImagePyramid[l] = cv::Mat(rows, cols_r[l], CV_8UC1, (void*)pyr_r[l], (size_t)stride_r[l]);
    }
    F.mpORBextractorLeft = &exl;
    F.mpORBextractorRight = &exr;
    F.ComputeStereoMatches();
    int n = 0;
    for (int i = 0; i < N; ++i) {
        mvuRight[i] = F.mvuRight[i];
        mvDepth[i] = F.mvDepth[i];
        n += F.mvDepth[i] > 0;
    }
    return n;
}

}  // extern "C"

// ORBR_DROPIN (oracle/Makefile target `mdrop` -> oracle/_ref/liborbmdrop.so): the same wrapper, the same reference
// Frame.cc / MapPoint.cc, but the class behind `ORBmatcher` is the PRODUCT's drop-in (orb-slam3-mac_b200/host/
// ORBmatcher.{h,cc} -> liborbx_b200.so) instead of src/ORBmatcher.cc.  It wraps the tracking searches above; what
// entry points below that need a method the drop-in does not wrap are left out of that build (#ifndef ORBR_DROPIN).
// ---- SearchByBoW x2, Sim3 / reloc projections, Fuse x2, SearchBySim3, SearchForTriangulation x2 ----
namespace {

const float kDefaultSF[8] = {1.f, 1.2f, 1.44f, 1.728f, 2.0736f, 2.48832f, 2.985985f, 3.583182f};

// a BowView has no image bounds: the BoW searches never query the grid, any bounds do
FrameView view_of(const BowView& B, int n, const float* sf = kDefaultSF, int nlevels = 8) {
    FrameView v;
    v.keys = B.keys;
    v.desc = B.desc;
    v.n = n;
    v.min_x = v.min_y = 0.f;
    v.max_x = v.max_y = 4096.f;
    v.grid_w_inv = 64.f / 4096.f;
    v.grid_h_inv = 48.f / 4096.f;
    v.scale_factors = sf;
    v.nlevels = nlevels;
    v.u_right = nullptr;
    return v;
}

// Frame with keys [0, nleft) in mvKeys and the rest in mvKeysRight (nleft < 0: everything in mvKeys / mvKeysUn)
void fill_bow_frame(Frame& F, const BowView& B, int nleft, const float* sf = kDefaultSF, int nlevels = 8) {
    if (nleft < 0 || nleft > B.n) {
        FrameView v = view_of(B, B.n, sf, nlevels);
        fill_frame(F, v);
    } else {
        FrameView v = view_of(B, nleft, sf, nlevels);
        FrameRight R;
        R.keys = B.keys + nleft;
        R.desc = B.desc + (size_t)nleft * 32;
        R.n = B.n - nleft;
        R.left_to_right = R.right_to_left = nullptr;
        fill_frame(F, v, &R);
    }
    F.mFeatVec.clear();
    for (int a = 0; a < B.nnodes; ++a)
        F.mFeatVec[B.node_ids[a]] = std::vector<unsigned int>(B.node_items + B.node_ptr[a], B.node_items + B.node_ptr[a + 1]);
}

// One map point per projected snapshot point: world position (i, 0, z) [or xyz], normal along it, no lower
// distance bound, and the upper bound chosen so that PredictScale(dist, .) (src/MapPoint.cc:514-546) lands in
// the middle of level `level`'s interval: ceil(log(max / dist) / log(sf)) == level.
void place_point(MapPoint& p, float x, float y, float z, int level, float sf, Map* map, long unsigned id) {
    p.mWorldPos = (cv::Mat_<float>(3, 1) << x, y, z);
    p.mNormalVector = (cv::Mat_<float>(3, 1) << x, y, z);
    const float dist = (float)std::sqrt((double)x * x + (double)y * y + (double)z * z);
    p.mfMinDistance = 0.f;
    p.mfMaxDistance = dist * (float)std::pow((double)sf, (double)level - 0.5);
    p.mpMap = map;
    p.mnId = id;
    p.mbBad = false;
    p.nObs = 0;
}

int index_in(MapPoint* p, MapPoint* base, int n) { return (p >= base && p < base + n) ? (int)(p - base) : -1; }

}  // namespace

extern "C" {

// SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches), src/ORBmatcher.cc:273-475.  Nleft: F.Nleft.
int orbr_search_by_bow(const BowView* KFp, const uint8_t* kf_mp_ok, const BowView* Fp, int Nleft, float mfNNratio,
                       int mbCheckOrientation, int32_t* matches) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    Frame KFf, F;
    fill_bow_frame(KFf, *KFp, -1);
    fill_bow_frame(F, *Fp, Nleft);
    TableCamera dummy;
    if (Nleft != -1) F.mpCamera2 = &dummy;  // a fisheye frame has a second camera (:421)
    KeyFrame kf(KFf, &map);
    std::unique_ptr<MapPoint[]> mp(new MapPoint[KFp->n > 0 ? KFp->n : 1]);
    for (int i = 0; i < KFp->n; ++i) kf.mvpMapPoints[i] = kf_mp_ok[i] ? &mp[i] : static_cast<MapPoint*>(NULL);
    std::vector<MapPoint*> vp;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const int n = matcher.SearchByBoW(&kf, F, vp);
    for (int j = 0; j < Fp->n; ++j) matches[j] = index_in(vp[j], mp.get(), KFp->n);
    return n;
}

// SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12), src/ORBmatcher.cc:827-967.
// n_un >= 0: that keyframe has NLeft != -1 and mvKeysUn.size() == n_un (:863-865, :887-889).
int orbr_search_by_bow_kf(const BowView* K1, const uint8_t* ok1, int n_un1, const BowView* K2, const uint8_t* ok2, int n_un2,
                          float mfNNratio, int mbCheckOrientation, int32_t* vpMatches12) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    Frame F1, F2;
    fill_bow_frame(F1, *K1, n_un1);
    fill_bow_frame(F2, *K2, n_un2);
    KeyFrame kf1(F1, &map), kf2(F2, &map);
    std::unique_ptr<MapPoint[]> mp1(new MapPoint[K1->n > 0 ? K1->n : 1]), mp2(new MapPoint[K2->n > 0 ? K2->n : 1]);
    for (int i = 0; i < K1->n; ++i) kf1.mvpMapPoints[i] = ok1[i] ? &mp1[i] : static_cast<MapPoint*>(NULL);
    for (int i = 0; i < K2->n; ++i) kf2.mvpMapPoints[i] = ok2[i] ? &mp2[i] : static_cast<MapPoint*>(NULL);
    std::vector<MapPoint*> vp;
    ORBmatcher matcher(mfNNratio, mbCheckOrientation != 0);
    const int n = matcher.SearchByBoW(&kf1, &kf2, vp);
    for (int i = 0; i < K1->n; ++i) vpMatches12[i] = index_in(vp[i], mp2.get(), K2->n);
    return n;
}


#ifndef ORBR_DROPIN
// SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming), src/ORBmatcher.cc:477-590, AND its overload
// with vpPointsKFs / vpMatchedKF, :592-708 (the wrapper runs both and refuses, -102, if they disagree).  Scw = I.
// The first overload projects through pKF->mpCamera (:523), the second with the keyframe's fx, fy, cx, cy (:635-640):
// the reference's own Pinhole with unit intrinsics and the point (u, v, 1) give exactly (u, v) in both.
int orbr_search_by_projection_sim3(const FrameView* KFp, const ProjPoint* pts, const uint8_t* mp_desc, int m, int th,
                                   float ratioHamming, const uint8_t* kp_taken, int32_t* kp_match) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    Frame F;
    fill_frame(F, *KFp);
    Pinhole cam(std::vector<float>{1.f, 1.f, 0.f, 0.f});
    F.mpCamera = &cam;
    KeyFrame kf(F, &map);
    std::unique_ptr<MapPoint[]> mp(new MapPoint[m > 0 ? m : 1]), taken(new MapPoint[F.N > 0 ? F.N : 1]);
    std::vector<MapPoint*> vpPoints(m);
    for (int i = 0; i < m; ++i) {
        place_point(mp[i], pts[i].u, pts[i].v, 1.f, pts[i].level, F.mfScaleFactor, &map, 1000 + i);
        mp[i].mbBad = !pts[i].valid;
        set_desc(mp[i], mp_desc + (size_t)i * 32);
        vpPoints[i] = &mp[i];
    }
    std::vector<MapPoint*> vpMatched(F.N), vpMatched2;
    for (int i = 0; i < F.N; ++i) vpMatched[i] = kp_taken[i] ? &taken[i] : static_cast<MapPoint*>(NULL);
    vpMatched2 = vpMatched;
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchByProjection(&kf, eye4(), vpPoints, vpMatched, th, ratioHamming);
    std::vector<KeyFrame*> vpPointsKFs(m, &kf), vpMatchedKF(F.N, static_cast<KeyFrame*>(NULL));
    const int n2 = matcher.SearchByProjection(&kf, eye4(), vpPoints, vpPointsKFs, vpMatched2, vpMatchedKF, th, ratioHamming);
    if (n2 != n || vpMatched2 != vpMatched) return -102;
    for (int i = 0; i < F.N; ++i) kp_match[i] = index_in(vpMatched[i], mp.get(), m);
    return n;
}
#endif  // ORBR_DROPIN

#ifndef ORBR_DROPIN
// SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:2183-2305.
// Point i = map point of keyframe feature i (its angle in pts[i].angle).
int orbr_search_by_projection_reloc(const FrameView* Fp, const ProjPoint* pts, const uint8_t* mp_desc, int m, float th,
                                    int ORBdist, int mbCheckOrientation, const uint8_t* kp_taken, int32_t* kp_match) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (m >= kRightBase) return -101;
    Map map;
    // the keyframe: m features, of which only mvKeysUn[i].angle and the map points are read
    std::vector<Kp> kkeys(m > 0 ? m : 1);
    std::vector<uint8_t> kdesc(32 * (size_t)(m > 0 ? m : 1), 0);
    memset(kkeys.data(), 0, kkeys.size() * sizeof(Kp));
    for (int i = 0; i < m; ++i) kkeys[i].angle = pts[i].angle;
    FrameView kv = *Fp;
    kv.keys = kkeys.data();
    kv.desc = kdesc.data();
    kv.n = m;
    kv.u_right = nullptr;
    Frame KFf, F;
    fill_frame(KFf, kv);
    fill_frame(F, *Fp);  // last: Frame's static bounds are the current frame's
    std::vector<float> uv(2 * (size_t)(m > 0 ? m : 1));
    TableCamera cam;
    cam.uv = uv.data();
    F.mpCamera = &cam;
    F.SetPose(eye4());
    KeyFrame kf(KFf, &map);
    std::unique_ptr<MapPoint[]> mp(new MapPoint[m > 0 ? m : 1]), taken(new MapPoint[F.N > 0 ? F.N : 1]);
    for (int i = 0; i < m; ++i) {
        uv[2 * i] = pts[i].u;
        uv[2 * i + 1] = pts[i].v;
        place_point(mp[i], (float)i, 0.f, 1.f, pts[i].level, F.mfScaleFactor, &map, 1000 + i);
        set_desc(mp[i], mp_desc + (size_t)i * 32);
        kf.mvpMapPoints[i] = pts[i].valid ? &mp[i] : static_cast<MapPoint*>(NULL);
    }
    for (int i = 0; i < F.N; ++i) F.mvpMapPoints[i] = kp_taken[i] ? &taken[i] : static_cast<MapPoint*>(NULL);
    std::set<MapPoint*> sAlreadyFound;
    ORBmatcher matcher(0.75f, mbCheckOrientation != 0);
    const int n = matcher.SearchByProjection(F, &kf, sAlreadyFound, th, ORBdist);
    for (int i = 0; i < F.N; ++i) kp_match[i] = index_in(F.mvpMapPoints[i], mp.get(), m);
    return n;
}
#endif  // ORBR_DROPIN

#ifndef ORBR_DROPIN
// ORBmatcher::Fuse, both overloads.  mvInvLevelSigma2 != NULL: Fuse(pKF, vpMapPoints, th, bRight = false),
// src/ORBmatcher.cc:1403-1613 -- point i sits at (i, 0, depth[i]) and the keyframe's mbf is `bf`, so the reference's
// ur = uv.x - bf * invz (:1490) must reproduce pts[i].ur (else -100: the tests derive ur from depth that way).
// mvInvLevelSigma2 == NULL: Fuse(pKF, Scw = I, vpPoints, th, vpReplacePoint), :1615-1737.
// The search of a point does not depend on what earlier points did to the map, so the per-point
// result is read by fusing ONE point at a time into a keyframe without map points: nFused is 0 or 1, the
// chosen keypoint is the observation the reference added, its distance is recomputed with DescriptorDistance.
// Then all points are fused in one call (Replace / AddObservation paths included) and the totals must agree (-103).
int orbr_fuse(const FrameView* KFp, const ProjPoint* pts, const uint8_t* mp_desc, int m, float th,
              const float* mvInvLevelSigma2, int idx_offset, int32_t* best_idx, int32_t* best_dist, const float* depth,
              float bf) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (m >= kRightBase) return -101;
    Map map;
    Frame F;
    fill_frame(F, *KFp);
    if (mvInvLevelSigma2) F.mvInvLevelSigma2.assign(mvInvLevelSigma2, mvInvLevelSigma2 + F.mnScaleLevels);
    F.mbf = bf;
    std::vector<float> uv(2 * (size_t)(m > 0 ? m : 1));
    TableCamera cam;
    cam.uv = uv.data();
    F.mpCamera = &cam;
    KeyFrame kf(F, &map);
    ORBmatcher matcher(0.6f, true);
    int total = 0;
    for (int pass = 0; pass < 2; ++pass) {  // 0: one point per call; 1: all points in one call
        std::unique_ptr<MapPoint[]> mp(new MapPoint[m > 0 ? m : 1]);
        std::vector<MapPoint*> all;
        for (int i = 0; i < m; ++i) {
            uv[2 * i] = pts[i].u;
            uv[2 * i + 1] = pts[i].v;
            const float z = (mvInvLevelSigma2 && depth) ? depth[i] : 1.f;
            if (mvInvLevelSigma2 && pts[i].valid) {
                const float invz = 1 / z;
                const float ur = pts[i].u - bf * invz;
                if (ur != pts[i].ur) return -100;
            }
            place_point(mp[i], (float)i, 0.f, z, pts[i].level, F.mfScaleFactor, &map, 1000 + i);
            mp[i].mbBad = !pts[i].valid;
            set_desc(mp[i], mp_desc + (size_t)i * 32);
            all.push_back(&mp[i]);
        }
        kf.mvpMapPoints.assign(F.N, static_cast<MapPoint*>(NULL));
        if (pass == 1) {
            std::vector<MapPoint*> repl(m, static_cast<MapPoint*>(NULL));
            const int nb = mvInvLevelSigma2 ? matcher.Fuse(&kf, all, th, false) : matcher.Fuse(&kf, eye4(), all, th, repl);
            if (nb != total) return -103;
            break;
        }
        for (int i = 0; i < m; ++i) {
            best_idx[i] = -1;
            best_dist[i] = 256;
            std::vector<MapPoint*> one(1, &mp[i]), repl(1, static_cast<MapPoint*>(NULL));
            const int nf = mvInvLevelSigma2 ? matcher.Fuse(&kf, one, th, false) : matcher.Fuse(&kf, eye4(), one, th, repl);
            if (nf != 1) continue;
            const int idx = get<0>(mp[i].GetIndexInKeyFrame(&kf));
            if (idx < 0 || kf.mvpMapPoints[idx] != &mp[i]) return -105;
            best_idx[i] = idx + idx_offset;
            best_dist[i] = ORBmatcher::DescriptorDistance(mp[i].mDescriptor, kf.mDescriptors.row(idx));
            kf.mvpMapPoints[idx] = static_cast<MapPoint*>(NULL);
            ++total;
        }
    }
    return total;
}
#endif  // ORBR_DROPIN

#ifndef ORBR_DROPIN
// SearchBySim3(pKF1, pKF2, vpMatches12, s12 = 1, R12 = I, t12 = 0, th), src/ORBmatcher.cc:1739-1963.  The projection there
// is pinhole arithmetic on the keyframe's fx, fy, cx, cy (:1802-1806, :1882-1886), not a camera call: with fx = fy = 1,
// cx = cy = 0 the point (u, v, 1) projects to exactly (u, v).
int orbr_search_by_sim3(const FrameView* KF1, const FrameView* KF2, const ProjPoint* pts1, const uint8_t* mp_desc1,
                        const ProjPoint* pts2, const uint8_t* mp_desc2, float th, int32_t* match12) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    Frame F1, F2;
    fill_frame(F1, *KF1);
    fill_frame(F2, *KF2);
    KeyFrame kf1(F1, &map), kf2(F2, &map);
    const int N1 = KF1->n, N2 = KF2->n;
    std::unique_ptr<MapPoint[]> mp1(new MapPoint[N1 > 0 ? N1 : 1]), mp2(new MapPoint[N2 > 0 ? N2 : 1]);
    for (int i = 0; i < N1; ++i) {
        place_point(mp1[i], pts1[i].u, pts1[i].v, 1.f, pts1[i].level, F2.mfScaleFactor, &map, 1000 + i);
        set_desc(mp1[i], mp_desc1 + (size_t)i * 32);
        kf1.mvpMapPoints[i] = pts1[i].valid ? &mp1[i] : static_cast<MapPoint*>(NULL);
    }
    for (int i = 0; i < N2; ++i) {
        place_point(mp2[i], pts2[i].u, pts2[i].v, 1.f, pts2[i].level, F1.mfScaleFactor, &map, 100000 + i);
        set_desc(mp2[i], mp_desc2 + (size_t)i * 32);
        kf2.mvpMapPoints[i] = pts2[i].valid ? &mp2[i] : static_cast<MapPoint*>(NULL);
    }
    std::vector<MapPoint*> vpMatches12(N1, static_cast<MapPoint*>(NULL));
    const float s12 = 1.f;
    cv::Mat R12 = cv::Mat::eye(3, 3, CV_32F), t12 = cv::Mat::zeros(3, 1, CV_32F);
    ORBmatcher matcher(0.75f, true);
    const int n = matcher.SearchBySim3(&kf1, &kf2, vpMatches12, s12, R12, t12, th);
    for (int i = 0; i < N1; ++i) match12[i] = index_in(vpMatches12[i], mp2.get(), N2);
    return n;
}
#endif  // ORBR_DROPIN

#ifndef ORBR_DROPIN
// SearchForTriangulation, src/ORBmatcher.cc:969-1210 (overload == 1: epipolarConstrain, bOnlyStereo, bCoarse) and
// :1212-1401 (overload == 2: matchAndtriangulate).  pair_ok != NULL: the camera model's pair test is that table;
// else F12 != NULL (a skew-symmetric [t]x, -104 otherwise): the reference's own Pinhole::epipolarConstrain with
// K = I, R12 = I, t12 = t; else bCoarse.  ep == NULL: the epipole is put far outside the image.
int orbr_search_for_triangulation(const BowView* K1, const uint8_t* has_mp1, const uint8_t* stereo1, const BowView* K2,
                                  const uint8_t* has_mp2, const uint8_t* stereo2, int bOnlyStereo, const float* ep,
                                  const float* mvScaleFactors2, const uint8_t* pair_ok, const float* F12,
                                  const float* mvLevelSigma2_2, int mbCheckOrientation, int32_t* vMatches12, int overload) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    Frame F1, F2;
    fill_bow_frame(F1, *K1, -1);
    fill_bow_frame(F2, *K2, -1, mvScaleFactors2 ? mvScaleFactors2 : kDefaultSF, 8);
    if (mvLevelSigma2_2) F2.mvLevelSigma2.assign(mvLevelSigma2_2, mvLevelSigma2_2 + 8);
    for (int i = 0; i < K1->n; ++i) F1.mvuRight[i] = (stereo1 && stereo1[i]) ? 1.f : -1.f;
    for (int i = 0; i < K2->n; ++i) F2.mvuRight[i] = (stereo2 && stereo2[i]) ? 1.f : -1.f;
    const float epv[2] = {ep ? ep[0] : 1e9f, ep ? ep[1] : 1e9f};
    TableCamera tcam;
    tcam.uv = epv;  // the epipole is "point 0"
    tcam.pair_ok = pair_ok;
    tcam.n2 = K2->n;
    std::unique_ptr<PinholeEp> pcam;
    cv::Mat T1 = eye4();
    const bool coarse = !pair_ok && !F12;
    if (!pair_ok && F12) {
        const float t[3] = {F12[7], F12[2], F12[3]};
        const float want[9] = {0.f, -t[2], t[1], t[2], 0.f, -t[0], -t[1], t[0], 0.f};
        for (int i = 0; i < 9; ++i)
            if (want[i] != F12[i]) return -104;
        pcam.reset(new PinholeEp(std::vector<float>{1.f, 1.f, 0.f, 0.f}, cv::Point2f(epv[0], epv[1])));
        for (int i = 0; i < 3; ++i) T1.at<float>(i, 3) = t[i];
        F1.mpCamera = F2.mpCamera = pcam.get();
    } else {
        F1.mpCamera = F2.mpCamera = &tcam;
    }
    F1.SetPose(T1);
    F2.SetPose(eye4());
    KeyFrame kf1(F1, &map), kf2(F2, &map);
    kf1.Ow = cv::Mat::zeros(3, 1, CV_32F);  // C2 = R2w * Cw + t2w = (0, 0, 1): "point 0" of the table camera
    kf1.Ow.at<float>(2) = 1.f;
    std::unique_ptr<MapPoint[]> mp1(new MapPoint[K1->n > 0 ? K1->n : 1]), mp2(new MapPoint[K2->n > 0 ? K2->n : 1]);
    for (int i = 0; i < K1->n; ++i) kf1.mvpMapPoints[i] = has_mp1[i] ? &mp1[i] : static_cast<MapPoint*>(NULL);
    for (int i = 0; i < K2->n; ++i) kf2.mvpMapPoints[i] = has_mp2[i] ? &mp2[i] : static_cast<MapPoint*>(NULL);
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher matcher(0.6f, mbCheckOrientation != 0);
    int n;
    cv::Mat F12m = cv::Mat::zeros(3, 3, CV_32F);  // the argument is not read by either overload
    if (overload == 2) {
        std::vector<cv::Mat> pts3d;
        n = matcher.SearchForTriangulation(&kf1, &kf2, F12m, pairs, bOnlyStereo != 0, pts3d);
    } else {
        n = matcher.SearchForTriangulation(&kf1, &kf2, F12m, pairs, bOnlyStereo != 0, coarse);
    }
    for (int i = 0; i < K1->n; ++i) vMatches12[i] = -1;
    for (size_t k = 0; k < pairs.size(); ++k) vMatches12[pairs[k].first] = (int32_t)pairs[k].second;
    return n;
}
#endif  // ORBR_DROPIN

}  // extern "C"

// ---- MapPoint::ComputeDistinctiveDescriptors, DBoW2 transform, Frame::UndistortKeyPoints ---------------
#include <cstdio>
#include <new>
#include <unistd.h>

extern "C" {

// MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:330-404, for a batch of map points.  Observation j of
// every group lives in keyframe j (keyframes are laid out contiguously, so the std::map<KeyFrame*, ..> of the
// reference iterates them in index order = the order of the snapshot's rows); its descriptor is row g of that
// keyframe's mDescriptors.  best[g] = first row of the group equal to the descriptor the reference chose
// (equal rows have equal medians, so this IS its BestIdx), -1 for an empty group.
int orbr_distinctive_descriptors(const uint8_t* desc, const int32_t* group_ptr, int ngroups, int32_t* best) {
    std::lock_guard<std::mutex> lk(g_mu);
    Map map;
    int K = 0;
    for (int g = 0; g < ngroups; ++g) K = std::max(K, group_ptr[g + 1] - group_ptr[g]);
    std::vector<Kp> keys(ngroups > 0 ? ngroups : 1);
    memset(keys.data(), 0, keys.size() * sizeof(Kp));
    KeyFrame* kfs = (KeyFrame*)operator new[]((size_t)(K > 0 ? K : 1) * sizeof(KeyFrame));
    std::vector<uint8_t> rows((size_t)(ngroups > 0 ? ngroups : 1) * 32);
    for (int j = 0; j < K; ++j) {
        std::fill(rows.begin(), rows.end(), 0);
        for (int g = 0; g < ngroups; ++g)
            if (j < group_ptr[g + 1] - group_ptr[g]) memcpy(&rows[(size_t)g * 32], desc + (size_t)(group_ptr[g] + j) * 32, 32);
        FrameView v;
        v.keys = keys.data();
        v.desc = rows.data();
        v.n = ngroups;
        v.min_x = v.min_y = 0.f;
        v.max_x = v.max_y = 4096.f;
        v.grid_w_inv = v.grid_h_inv = 0.01f;
        v.scale_factors = kDefaultSF;
        v.nlevels = 8;
        v.u_right = nullptr;
        Frame F;
        fill_frame(F, v);
        new (&kfs[j]) KeyFrame(F, &map);
    }
    for (int g = 0; g < ngroups; ++g) {
        const int n = group_ptr[g + 1] - group_ptr[g];
        MapPoint mp;
        mp.mpMap = &map;
        for (int j = 0; j < n; ++j) mp.mObservations[&kfs[j]] = std::tuple<int, int>(g, -1);
        mp.ComputeDistinctiveDescriptors();
        best[g] = -1;
        if (mp.mDescriptor.empty()) continue;
        for (int j = 0; j < n; ++j)
            if (!memcmp(mp.mDescriptor.data, desc + (size_t)(group_ptr[g] + j) * 32, 32)) {
                best[g] = j;
                break;
            }
    }
    for (int j = 0; j < K; ++j) kfs[j].~KeyFrame();
    operator delete[](kfs);
    return 0;
}

// DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>::transform(features, v, fv, levelsup),
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1136-1206 (per feature :1230-1266, FORB::distance FORB.cpp:85-105).
// The vocabulary is handed to the reference's own loader, loadFromTextFile (:1350-1436), through a temporary file
// in ORBvoc.txt's format (no newline after the last node: the loader would read one more, phantom, node).
// norm_kind 0 / 1 / 2 -> DOT_PRODUCT / L1_NORM / L2_NORM (ScoringObject.h:73-90: the scorings that do not
// normalise, normalise L1, normalise L2).
int orbr_bow_transform(int n_nodes, const int32_t* child_ptr, const uint32_t* child_ids, const uint8_t* node_desc,
                       const double* weight, const uint32_t* word_id_of, int m_L, int m_weighting, int norm_kind,
                       const uint8_t* features, int n, int levelsup, uint32_t* bow_ids, double* bow_vals, int32_t* n_words,
                       uint32_t* fv_node_ids, int32_t* fv_ptr, uint32_t* fv_items, int32_t* n_fv_nodes) {
    std::lock_guard<std::mutex> lk(g_mu);
    std::vector<int> parent(n_nodes, 0);
    int k = 0;
    for (int i = 0; i < n_nodes; ++i) {
        k = std::max(k, child_ptr[i + 1] - child_ptr[i]);
        for (int c = child_ptr[i]; c < child_ptr[i + 1]; ++c) parent[child_ids[c]] = i;
    }
    char path[] = "/tmp/orbmref_voc_XXXXXX";
    const int fd = mkstemp(path);
    if (fd < 0) return -110;
    FILE* f = fdopen(fd, "w");
    const int scoring = norm_kind == 1 ? 0 : norm_kind == 2 ? 1 : 5;
    fprintf(f, "%d %d %d %d", k, m_L, scoring, m_weighting);
    int next_word = 0;
    for (int i = 1; i < n_nodes; ++i) {  // node ids are assigned in file order, word ids in order of appearance
        const bool leaf = child_ptr[i + 1] == child_ptr[i];
        if (leaf && (int)word_id_of[i] != next_word++) {
            fclose(f);
            unlink(path);
            return -111;
        }
        fprintf(f, "\n%d %d", parent[i], leaf ? 1 : 0);
        for (int b = 0; b < 32; ++b) fprintf(f, " %d", node_desc[(size_t)i * 32 + b]);
        fprintf(f, " %.17g", weight[i]);
    }
    fclose(f);
    ORBVocabulary voc;
    const bool ok = voc.loadFromTextFile(path);
    unlink(path);
    if (!ok) return -112;
    std::vector<cv::Mat> feats = Converter::toDescriptorVector(to_desc(features, n));  // src/Frame.cc:733
    DBoW2::BowVector v;
    DBoW2::FeatureVector fv;
    voc.transform(feats, v, fv, levelsup);
    int a = 0;
    for (DBoW2::BowVector::const_iterator it = v.begin(); it != v.end(); ++it, ++a) {
        bow_ids[a] = it->first;
        bow_vals[a] = it->second;
    }
    *n_words = a;
    a = 0;
    int items = 0;
    fv_ptr[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
        fv_node_ids[a] = it->first;
        for (size_t j = 0; j < it->second.size(); ++j) fv_items[items++] = it->second[j];
        fv_ptr[++a] = items;
    }
    *n_fv_nodes = a;
    return 0;
}

// Frame::UndistortKeyPoints (src/Frame.cc:738-771) and Frame::ComputeImageBounds (:773-800) with the reference's own
// Pinhole::toK; cv::undistortPoints is the shim's restatement, pinned against cv2 by tests/test_oracle_cvshim.py.
// xy_out: n (x, y) pairs; bounds_out: mnMinX, mnMaxX, mnMinY, mnMaxY.
int orbr_undistort_keypoints(const Kp* keys, int n, float fx, float fy, float cx, float cy, const float* dist, int ndist,
                             int width, int height, float* xy_out, float* bounds_out) {
    std::lock_guard<std::mutex> lk(g_mu);
    Frame F;
    F.N = n;
    F.mvKeys = to_keypoints(keys, n);
    Pinhole cam(std::vector<float>{fx, fy, cx, cy});
    F.mpCamera = &cam;
    F.mK = cam.toK();
    F.mDistCoef = cv::Mat(ndist, 1, CV_32F);
    for (int i = 0; i < ndist; ++i) F.mDistCoef.at<float>(i) = dist[i];
    F.UndistortKeyPoints();
    for (int i = 0; i < n; ++i) {
        xy_out[2 * i] = F.mvKeysUn[i].pt.x;
        xy_out[2 * i + 1] = F.mvKeysUn[i].pt.y;
    }
    cv::Mat im(height, width, CV_8UC1, (void*)keys, (size_t)width);  // only cols / rows are read
    F.ComputeImageBounds(im);
    bounds_out[0] = Frame::mnMinX;
    bounds_out[1] = Frame::mnMaxX;
    bounds_out[2] = Frame::mnMinY;
    bounds_out[3] = Frame::mnMaxY;
    return 0;
}

// ---- hooks that expose the shim's OpenCV restatements to tests/test_oracle_cvshim.py (checked against cv2) ----
// D = alpha * op(A) * op(B) + beta * C through the shim's MatExpr machinery, float32
int orbr_shim_gemm(const float* A, int ar, int ac, const float* B, int br, int bc, double alpha, const float* C, double beta,
                   int flags, float* D) {
    cv::Mat a(ar, ac, CV_32F, (void*)A), b(br, bc, CV_32F, (void*)B);
    cv::Mat c;
    if (C) c = cv::Mat((flags & 1) ? ac : ar, (flags & 2) ? br : bc, CV_32F, (void*)C);
    cv::Mat d = cv::shim::gemm(a, b, alpha, flags, C ? &c : nullptr, beta);
    for (int y = 0; y < d.rows; ++y) memcpy(D + (size_t)y * d.cols, d.ptr(y), (size_t)d.cols * 4);
    return d.rows * d.cols;
}
// the expression forms the compiled reference files use on poses; which: 0 R*p+t, 1 -R.t()*t, 2 R1*R2.t(),
// 3 M/s, 4 K.t().inv()*S*R*K2.inv() (Pinhole.cpp:133), 5 -R1*R2.t()*t2+t1 (ORBmatcher.cc:996)
int orbr_shim_expr(int which, const float* m1, const float* m2, const float* v1, const float* v2, float s, float* out) {
    cv::Mat A(3, 3, CV_32F, (void*)m1), B(3, 3, CV_32F, (void*)m2), p(3, 1, CV_32F, (void*)v1), t(3, 1, CV_32F, (void*)v2);
    cv::Mat r;
    switch (which) {
        case 0: r = A * p + t; break;
        case 1: r = -A.t() * t; break;
        case 2: r = A * B.t(); break;
        case 3: r = A / s; break;
        case 4: { cv::Mat I = cv::Mat::eye(3, 3, CV_32F); r = A.t().inv() * B * I * A.inv(); break; }
        case 5: r = -A * B.t() * p + t; break;
        default: return -1;
    }
    for (int y = 0; y < r.rows; ++y) memcpy(out + (size_t)y * r.cols, r.ptr(y), (size_t)r.cols * 4);
    return r.rows * r.cols;
}
double orbr_shim_norm(const float* v, int n) { return cv::norm(cv::Mat(n, 1, CV_32F, (void*)v)); }
int orbr_shim_invert3(const float* m, float* out) {
    cv::Mat r = cv::Mat(3, 3, CV_32F, (void*)m).inv();
    memcpy(out, r.data, 36);
    return 0;
}
int orbr_shim_undistort(const float* xy, int n, const float* K, const float* dist, int ndist, const float* P, float* out) {
    cv::Mat src(n, 1, CV_32FC2, (void*)xy), k(3, 3, CV_32F, (void*)K), d(ndist, 1, CV_32F, (void*)dist), p(3, 3, CV_32F, (void*)P);
    cv::Mat dst;
    cv::undistortPoints(src, dst, k, d, cv::Mat(), p);
    memcpy(out, dst.data, (size_t)n * 8);
    return 0;
}
int orbr_shim_knn2(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, float* dist) {
    cv::BFMatcher bf(cv::NORM_HAMMING);
    std::vector<std::vector<cv::DMatch> > m;
    bf.knnMatch(to_desc(q, nq), to_desc(t, nt), m, 2);
    for (int i = 0; i < nq; ++i)
        for (int j = 0; j < 2; ++j) {
            const bool have = i < (int)m.size() && j < (int)m[i].size();
            idx[2 * i + j] = have ? m[i][j].trainIdx : -1;
            dist[2 * i + j] = have ? m[i][j].distance : -1.f;
        }
    return 0;
}

}  // extern "C"
