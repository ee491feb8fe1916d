// oracle/matchshim/facade.h -- TEST INFRASTRUCTURE ONLY.
//
// Pre-included (g++ -include) ahead of the reference files that oracle/Makefile compiles VERBATIM
// from /root/reference into oracle/_ref/liborbmref.so:
//     src/ORBmatcher.cc   src/Frame.cc   src/MapPoint.cc   src/CameraModels/Pinhole.cpp
//     Thirdparty/DBoW2/DBoW2/{FORB,BowVector,FeatureVector,ScoringObject}.cpp
// (src/ORBextractor.cc is compiled next to them with oracle/cvshim/list_arena.h, as for liborbref.so).
//
// The reference's own headers are used wherever they can be: include/Frame.h, MapPoint.h,
// ORBmatcher.h, ORBextractor.h, ORBVocabulary.h, CameraModels/GeometricCamera.h, CameraModels/Pinhole.h
// and all of Thirdparty/DBoW2.  What they pull in and this container does not have (Eigen, boost,
// OpenCV C++) or what drags in the rest of the SLAM system (KeyFrame.h -> KeyFrameDatabase, Atlas,
// ImuTypes -> Eigen; Map.h; G2oTypes.h -> g2o; Converter.h -> Eigen + g2o) is replaced HERE, by
// pre-defining the header's include guard and declaring a stand-in with exactly the members the
// compiled reference files touch:
//     KeyFrame   data members as include/KeyFrame.h declares them; the few methods the matcher calls
//                restate src/KeyFrame.cc (cited at each) -- the grid itself is copied from a real
//                Frame, i.e. it is built by the reference's own Frame::AssignFeaturesToGrid
//     Map        GetId / EraseMapPoint / mMutexPointCreation (what src/MapPoint.cc names)
//     IMU::*, ConstraintPoseImu, TwoViewReconstruction, Converter, KannalaBrandt8: inert stand-ins
//     Eigen::Matrix, boost::serialization: names only
// No reference source line is edited or copied; the stand-ins carry data, not matcher logic.
#ifndef ORACLE_MATCH_FACADE_H
#define ORACLE_MATCH_FACADE_H

#include <algorithm>
#include <cmath>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <mutex>
#include <numeric>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <tuple>
#include <vector>

#include "../cvshim/cvshim.h"

// ---- names only ---------------------------------------------------------------------------------
namespace boost {
namespace serialization {
class access;
template <class T>
inline T* make_array(T* p, size_t) { return p; }
template <class B, class D>
inline B& base_object(D& d) { return d; }
}  // namespace serialization
}  // namespace boost

namespace Eigen {
template <typename T, int R, int C>
struct Matrix {
    T v[R * C];
    Matrix() { for (int i = 0; i < R * C; ++i) v[i] = T(0); }
    T& operator[](int i) { return v[i]; }
    const T& operator[](int i) const { return v[i]; }
    T& operator()(int i) { return v[i]; }
    const T& operator()(int i) const { return v[i]; }
    T& operator()(int r, int c) { return v[r * C + c]; }
    const T& operator()(int r, int c) const { return v[r * C + c]; }
};
typedef Matrix<double, 2, 1> Vector2d;
typedef Matrix<double, 3, 1> Vector3d;
}  // namespace Eigen

// ---- headers replaced by the stand-ins below -------------------------------------------------------
#define KEYFRAME_H
#define MAP_H
#define IMUTYPES_H
#define G2OTYPES_H
#define CONVERTER_H
#define TwoViewReconstruction_H
#define CAMERAMODELS_KANNALABRANDT8_H
#define KEYFRAMEDATABASE_H
#define ATLAS_H

using namespace std;  // the reference headers rely on it leaking from their includes

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

namespace ORB_SLAM3 {

class MapPoint;
class Frame;
class Map;
class KeyFrame;
class GeometricCamera;
class KeyFrameDatabase;
class ConstraintPoseImu {};

namespace IMU {
class Bias {
public:
    float bax, bay, baz, bwx, bwy, bwz;
    Bias() : bax(0), bay(0), baz(0), bwx(0), bwy(0), bwz(0) {}
};
class Calib {
public:
    cv::Mat Tcb, Tbc, Cov, CovWalk;
    Calib() {}
};
class Preintegrated {
public:
    void SetNewBias(const Bias&) {}
};
}  // namespace IMU

// include/TwoViewReconstruction.h: only named by Pinhole::ReconstructWithTwoViews (monocular
// initialisation, not on the path)
class TwoViewReconstruction {
public:
    explicit TwoViewReconstruction(cv::Mat&, float = 1.0, int = 200) {}
    bool Reconstruct(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&, const std::vector<int>&,
                     cv::Mat&, cv::Mat&, std::vector<cv::Point3f>&, std::vector<bool>&) {
        fprintf(stderr, "matchshim: TwoViewReconstruction is not part of the oracle\n");
        abort();
    }
};

// include/Converter.h: Frame::ComputeBoW only needs toDescriptorVector (src/Converter.cc:25-34:
// one cv::Mat row view per descriptor)
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
        std::vector<cv::Mat> vDesc;
        vDesc.reserve(Descriptors.rows);
        for (int j = 0; j < Descriptors.rows; j++) vDesc.push_back(Descriptors.row(j));
        return vDesc;
    }
};

// include/Map.h
class Map {
public:
    long unsigned int mnId = 0;
    std::mutex mMutexPointCreation;
    std::set<MapPoint*> erased;
    long unsigned int GetId() { return mnId; }
    void EraseMapPoint(MapPoint* pMP) { erased.insert(pMP); }
};

}  // namespace ORB_SLAM3

// the reference's own camera interface and vocabulary (need the Eigen / boost names above)
#include "CameraModels/GeometricCamera.h"
#include "ORBVocabulary.h"

namespace ORB_SLAM3 {

// include/CameraModels/KannalaBrandt8.h: Frame::ComputeStereoFishEyeMatches casts the camera to it and
// calls TriangulateMatches (src/Frame.cc:1158).  The fisheye geometry (cv::fisheye, SVD) is out of
// scope; the stand-in keeps the interface and is implemented in oracle/ref_match_capi.cpp.
class KannalaBrandt8 : public GeometricCamera {
public:
    KannalaBrandt8() { mvParameters.resize(8); mnId = nNextId++; mnType = CAM_FISHEYE; }
    explicit KannalaBrandt8(const std::vector<float>& p) : GeometricCamera(p) { mnId = nNextId++; mnType = CAM_FISHEYE; }
    cv::Point2f project(const cv::Point3f& p3D);
    cv::Point2f project(const cv::Mat& m3D);
    Eigen::Vector2d project(const Eigen::Vector3d& v3D);
    cv::Mat projectMat(const cv::Point3f& p3D);
    float uncertainty2(const Eigen::Matrix<double, 2, 1>& p2D);
    cv::Point3f unproject(const cv::Point2f& p2D);
    cv::Mat unprojectMat(const cv::Point2f& p2D);
    cv::Mat projectJac(const cv::Point3f& p3D);
    Eigen::Matrix<double, 2, 3> projectJac(const Eigen::Vector3d& v3D);
    cv::Mat unprojectJac(const cv::Point2f& p2D);
    bool ReconstructWithTwoViews(const std::vector<cv::KeyPoint>&, const std::vector<cv::KeyPoint>&, const std::vector<int>&,
                                 cv::Mat&, cv::Mat&, std::vector<cv::Point3f>&, std::vector<bool>&);
    cv::Mat toK();
    bool epipolarConstrain(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12,
                           const cv::Mat& t12, const float sigmaLevel, const float unc);
    bool matchAndtriangulate(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, GeometricCamera* pOther, cv::Mat& Tcw1,
                             cv::Mat& Tcw2, const float sigmaLevel1, const float sigmaLevel2, cv::Mat& x3Dtriangulated);
    float TriangulateMatches(GeometricCamera* pCamera2, const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& R12,
                             const cv::Mat& t12, const float sigmaLevel, const float unc, cv::Mat& p3D);
    std::vector<int> mvLappingArea;
};

// include/KeyFrame.h: the data members the compiled reference files read, with the reference's names
// and types (const qualifiers included: mnMinX.. are ints there).  Constructed from a real Frame the
// way KeyFrame::KeyFrame(Frame&, Map*, KeyFrameDatabase*) does (src/KeyFrame.cc:49-100), so the grid,
// keypoints, descriptors and feature vector are whatever the reference's Frame produced.
class KeyFrame {
public:
    KeyFrame(Frame& F, Map* pMap);

    // src/KeyFrame.cc:137-189, 1176-1260
    cv::Mat GetPose() { return Tcw.clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }
    cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }
    cv::Mat GetRightPose();
    cv::Mat GetRightCameraCenter();
    cv::Mat GetRightRotation();
    cv::Mat GetRightTranslation();
    void SetPose(const cv::Mat& Tcw_);

    // src/KeyFrame.cc:676-760
    void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
    void EraseMapPointMatch(const int& idx) { mvpMapPoints[idx] = static_cast<MapPoint*>(NULL); }
    void EraseMapPointMatch(MapPoint* pMP);
    void ReplaceMapPointMatch(const int& idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
    std::set<MapPoint*> GetMapPoints();
    std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }

    // src/KeyFrame.cc:770-819
    std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const bool bRight = false) const;
    bool IsInImage(const float& x, const float& y) const;

    bool isBad() { return mbBad; }
    Map* GetMap() { return mpMap; }

    static long unsigned int nNextId;
    long unsigned int mnId;
    const long unsigned int mnFrameId;

    const int mnGridCols;
    const int mnGridRows;
    const float mfGridElementWidthInv;
    const float mfGridElementHeightInv;

    const float fx, fy, cx, cy, invfx, invfy, mbf, mb, mThDepth;
    const int N;
    const std::vector<cv::KeyPoint> mvKeys;
    const std::vector<cv::KeyPoint> mvKeysUn;
    const std::vector<float> mvuRight;
    const std::vector<float> mvDepth;
    const cv::Mat mDescriptors;
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;

    const int mnScaleLevels;
    const float mfScaleFactor;
    const float mfLogScaleFactor;
    const std::vector<float> mvScaleFactors;
    const std::vector<float> mvLevelSigma2;
    const std::vector<float> mvInvLevelSigma2;

    const int mnMinX;
    const int mnMinY;
    const int mnMaxX;
    const int mnMaxY;

    std::vector<MapPoint*> mvpMapPoints;
    std::vector<std::vector<std::vector<size_t> > > mGrid;
    bool mbBad;
    Map* mpMap;

    GeometricCamera *mpCamera, *mpCamera2;
    std::vector<int> mvLeftToRightMatch, mvRightToLeftMatch;
    cv::Mat mTlr;
    cv::Mat mTrl;
    const std::vector<cv::KeyPoint> mvKeysRight;
    const int NLeft, NRight;
    std::vector<std::vector<std::vector<size_t> > > mGridRight;

    cv::Mat Tcw, Twc, Ow;
};

}  // namespace ORB_SLAM3

#endif  // ORACLE_MATCH_FACADE_H
