// match_prelude.h — TEST INFRASTRUCTURE ONLY.  Force-included (-include) when the reference's own src/ORBmatcher.cc,
// src/Align.cc and src/SparseImageAlign.cc (with include/NLSSolver.h, NLSSolver_impl.hpp, RobustCost.h) are compiled, unmodified and in place, into oracle/_ref/libygz_ref_match.so (oracle/Makefile, refmatch).
//
// include/ORBmatcher.h pulls in Common.h (Eigen, sophus, g2o, pangolin, glog, OpenCV C++) and the full Frame / KeyFrame /
// MapPoint headers (DBoW2 vocabulary, IMU, Map, g2o types); none of that exists in this image and none of it is used by the
// matcher code.  Their include guards are defined here so the real headers become no-ops, and what the matcher code really
// touches comes from this file:
//   * OpenCV  -> oracle/minicv (bit-exact against cv2 4.13) + minicv_matops.hpp (the cv::Mat float expressions);
//   * Eigen / Sophus -> ref_shim/eigen_lite.h (fixed-size stand-in; it fixes only the order of float operations inside a
//     matrix / vector expression);
//   * Frame, KeyFrame, MapPoint, ORBextractor, Converter -> the shim classes below: exactly the members the matcher code
//     reads or writes.  The member FUNCTIONS on the path (GetFeaturesInArea, PosInGrid, AssignFeaturesToGrid,
//     ComputeStereoMatches, IsInImage, PredictScale, ComputeDistinctiveDescriptors, the pose getters, World2Pixel, ...) are
//     NOT restated: oracle/extract_ref_bodies.py copies their bodies out of the reference at build time and
//     oracle/ref_match_api.cpp compiles them against these classes.  Map bookkeeping the searches trigger (Replace,
//     AddObservation, AddMapPoint) is recorded, not performed: it is out of scope.
#ifndef YGZ_REF_MATCH_PRELUDE_H_
#define YGZ_REF_MATCH_PRELUDE_H_
#define YGZ_COMMON_H_      // include/Common.h:1-2
#define YGZ_MAPPOINT_H     // include/MapPoint.h:12-13
#define YGZ_KEYFRAME_H_    // include/KeyFrame.h:10-11
#define YGZ_FRAME_H_       // include/Frame.h:8-9
#define YGZ_CONVERTER_H_   // include/Converter.h:6-7

#include <algorithm>
#include <iostream>
#include <list>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>
#include <climits>
#include <cmath>
#include <cstdint>
#include <immintrin.h>   // _mm_popcnt_u64 of src/ORBmatcher.cc:1515 (the reference gets it through OpenCV / Eigen headers)
using namespace std;  // include/Common.h:19

#include "../minicv/minicv_matops.hpp"
using cv::Mat;        // include/Common.h:62
using cv::uchar;
using cv::CV_32F;     // macros in OpenCV
using cv::CV_8U;

#include "eigen_lite.h"
using Eigen::Vector2f;  // include/Common.h:24-36
using Eigen::Vector3f;
using Eigen::Matrix2f;
using Eigen::Matrix3f;
using Sophus::SE3f;     // include/Common.h:44

#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

struct YgzNullLog {   // glog stand-in (include/Common.h:70): messages are dropped
    template <class T> YgzNullLog &operator<<(const T &) { return *this; }
    YgzNullLog &operator<<(std::ostream &(*)(std::ostream &)) { return *this; }
};
#define LOG(severity) YgzNullLog()

// display-only statements of src/SparseImageAlign.cc (:139-140 resimg_, :201-203, :247-253 imshow): compiled, never executed
// (display_ is false in Tracking.cc:284 and here)
namespace cv {
struct DisplayExpr {};
inline DisplayExpr operator*(const Mat &, int) { return DisplayExpr(); }
inline void namedWindow(const char *, int) {}
inline void imshow(const char *, const DisplayExpr &) {}
inline int waitKey(int) { return 0; }
}  // namespace cv
#define CV_WINDOW_AUTOSIZE 1

#define FRAME_GRID_ROWS 48   // include/Frame.h:27-28
#define FRAME_GRID_COLS 64

namespace ygz {
using namespace Sophus;   // src/IMU/IMUPreintegrator.h:10 (reaches SparseImageAlign.cc through Frame.h): Vector6f

class Frame;
class KeyFrame;
class MapPoint;

// include/ORBextractor.h:112 — the public pyramid Frame::ComputeStereoMatches reads (src/Frame.cc:515, 604, 618, 623)
class ORBextractor {
public:
    std::vector<cv::Mat> mvImagePyramid;
};

// What a search did to the map, in call order (ORBmatcher::Fuse, src/ORBmatcher.cc:862-877, 989-998).
struct MapAction {
    int kind;            // 0 AddObservation(pKF, idx) on `self`, 1 self->Replace(other), 2 pKF->AddMapPoint(self, idx)
    MapPoint *self, *other;
    size_t idx;
};
extern std::vector<MapAction> g_mapActions;

// include/MapPoint.h: tracking scratch :95-102; position / descriptor / observations :150-185
class MapPoint {
public:
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;

    Vector3f GetWorldPos();
    Vector3f GetNormal();
    std::map<KeyFrame *, size_t> GetObservations();
    int Observations();
    bool isBad();
    void ComputeDistinctiveDescriptors();
    cv::Mat GetDescriptor();
    int GetIndexInKeyFrame(KeyFrame *pKF);
    bool IsInKeyFrame(KeyFrame *pKF);
    float GetMinDistanceInvariance();
    float GetMaxDistanceInvariance();
    int PredictScale(const float &currentDist, KeyFrame *pKF);
    int PredictScale(const float &currentDist, Frame *pF);
    // recorded, not performed (see the header comment)
    void AddObservation(KeyFrame *, size_t idx) { g_mapActions.push_back(MapAction{0, this, nullptr, idx}); }
    void Replace(MapPoint *pMP) { g_mapActions.push_back(MapAction{1, this, pMP, 0}); }

    Vector3f mWorldPos, mNormalVector;
    cv::Mat mDescriptor;
    std::map<KeyFrame *, size_t> mObservations;
    int nObs = 0;
    bool mbBad = false;
    float mfMinDistance = 0, mfMaxDistance = 0;
    std::mutex mMutexPos, mMutexFeatures;
};

// include/KeyFrame.h:208-211, 236-271, 305
class KeyFrame {
public:
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<std::vector<std::vector<size_t> > > mGrid;
    int mnScaleLevels = 0;
    float mfLogScaleFactor = 0;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    std::vector<cv::Mat> mvImagePyramid;
    std::vector<MapPoint *> mvpMapPoints;
    bool mbBad = false;
    SE3f Tcw;
    Vector3f Ow;
    std::mutex mMutexPose;

    std::vector<MapPoint *> GetMapPointMatches() { return mvpMapPoints; }                 // src/KeyFrame.cc:550-553
    MapPoint *GetMapPoint(const size_t &idx) { return mvpMapPoints[idx]; }                // :555-558
    std::set<MapPoint *> GetMapPoints() {                                                  // :507-521: non-null, not bad
        std::set<MapPoint *> s;
        for (MapPoint *p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
        return s;
    }
    void AddMapPoint(MapPoint *pMP, const size_t &idx) { g_mapActions.push_back(MapAction{2, pMP, nullptr, idx}); }
    bool isBad() { return mbBad; }
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r) const;
    bool IsInImage(const float &x, const float &y) const;
    SE3f GetPose();
    Vector3f GetCameraCenter();
    Matrix3f GetRotation();
    Vector3f GetTranslation();
#include "../_ref/gen/keyframe_members.inc"
};

// include/Frame.h:248-332
class Frame {
public:
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvuRight, mvDepth;
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    DBoW2::FeatureVector mFeatVec;
    SE3f mTcw;
    float mb = 0, mbf = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];
    bool mbGridSet = false;
    int mnScaleLevels = 0;
    float mfLogScaleFactor = 0;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    std::vector<cv::Mat> mvImagePyramid;
    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;

    void AssignFeaturesToGrid();
    std::vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r, const int minLevel = -1,
                                          const int maxLevel = -1) const;     // defaults: include/Frame.h:115-116
    bool PosInGrid(const cv::KeyPoint &kp, int &posX, int &posY);
    void ComputeStereoMatches();
#include "../_ref/gen/frame_members.inc"
};

// include/Converter.h:35-37 — the two conversions src/ORBmatcher.cc uses (src/Converter.cc: element copies)
class Converter {
public:
    static cv::Mat toCvMat(const Vector3f &m) {
        cv::Mat r(3, 1, cv::CV_32F);
        for (int i = 0; i < 3; i++) r.at<float>(i) = m(i);
        return r;
    }
    static cv::Mat toCvMat(const Matrix3f &m) {
        cv::Mat r(3, 3, cv::CV_32F);
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.at<float>(i, j) = m(i, j);
        return r;
    }
};

}  // namespace ygz
#endif
