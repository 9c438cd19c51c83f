// FrameLite.h — test double of the members of ygz::Frame (include/Frame.h:248-332) that the drop-in
// classes read or write.  In an integration the reference's own Frame.h is used instead (the member
// names below are the reference's).
#ifndef YGZ_HOST_FRAMELITE_H_
#define YGZ_HOST_FRAMELITE_H_
#include <cmath>
#include <map>
#include <vector>

#include "cvcompat.h"
#include "ORBVocabulary.h"
#include "../../include/ygzorb.h"

namespace ygz {
class ORBextractor;
struct KeyFrameLite;
struct MapPointLite {                       // include/MapPoint.h:95-102 + the calls SearchByProjection makes
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 1;
    int mnTrackScaleLevel = 0;
    bool mbTrackInView = true, mbBad = false;
    int nObs = 1;
    cv::Mat mDescriptor;
    float mWorldPos[3] = {0, 0, 0};
    const float *GetWorldPos() const { return mWorldPos; }       // Vector3f in the reference
    float mNormal[3] = {0, 0, 1}, mfMaxDistance = 1, mfMinDistance = 1;
    const float *GetNormal() const { return mNormal; }
    float GetMaxDistance() const { return mfMaxDistance; }      // the two getters INTEGRATION.md adds to MapPoint
    float GetMinDistance() const { return mfMinDistance; }
    std::map<KeyFrameLite *, size_t> mObservations;              // include/MapPoint.h:129
    MapPointLite *mpReplaced = nullptr;
    bool IsInKeyFrame(KeyFrameLite *pKF) const { return mObservations.count(pKF) != 0; }
    void AddObservation(KeyFrameLite *pKF, size_t idx) { if (!mObservations.count(pKF)) { mObservations[pKF] = idx; nObs++; } }
    inline void Replace(MapPointLite *pMP);                      // src/MapPoint.cc:159-197
    bool isBad() const { return mbBad; }
    int Observations() const { return nObs; }
    cv::Mat GetDescriptor() const { return mDescriptor.clone(); }
};
struct PoseLite {                           // the two accessors of Sophus::SE3f the matcher uses (mTcw)
    float R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, t[3] = {0, 0, 0};
    struct Rot { const float *p; float operator()(int r, int c) const { return p[3 * r + c]; } };
    Rot rotationMatrix() const { return Rot{R}; }
    const float *translation() const { return t; }
};
// Stand-in for Sophus::SE3f in the direct-tracking drop-ins (host/DirectTracking.h): unit quaternion (x, y, z, w) + translation
// with the accessor names of Sophus / Eigen, the group product (quaternion product + normalisation, translation through the
// rotated vector) and inverse.  An integration uses Sophus::SE3f itself.
struct Se3Lite {
    float q[4] = {0, 0, 0, 1}, t[3] = {0, 0, 0};
    struct Coeffs { const float *p; const float *data() const { return p; } };
    struct Quat { const float *p; Coeffs coeffs() const { return Coeffs{p}; } };
    struct Vec { const float *p; const float *data() const { return p; } };
    Quat unit_quaternion() const { return Quat{q}; }
    Vec translation() const { return Vec{t}; }
    static Se3Lite FromFloats(const float *T) { Se3Lite r; for (int i = 0; i < 4; i++) r.q[i] = T[i]; for (int i = 0; i < 3; i++) r.t[i] = T[4 + i]; return r; }
    void rotate(const float *v, float *o) const {          // Eigen's QuaternionBase::_transformVector
        const float ux = q[1] * v[2] - q[2] * v[1], uy = q[2] * v[0] - q[0] * v[2], uz = q[0] * v[1] - q[1] * v[0];
        const float ax = ux + ux, ay = uy + uy, az = uz + uz;
        o[0] = (v[0] + q[3] * ax) + (q[1] * az - q[2] * ay);
        o[1] = (v[1] + q[3] * ay) + (q[2] * ax - q[0] * az);
        o[2] = (v[2] + q[3] * az) + (q[0] * ay - q[1] * ax);
    }
    Se3Lite operator*(const Se3Lite &b) const {
        Se3Lite r;
        float rt[3];
        rotate(b.t, rt);
        for (int i = 0; i < 3; i++) r.t[i] = t[i] + rt[i];
        const float ax = q[0], ay = q[1], az = q[2], aw = q[3], bx = b.q[0], by = b.q[1], bz = b.q[2], bw = b.q[3];
        const float p[4] = {aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx,
                            aw * bw - ax * bx - ay * by - az * bz};
        const float len = std::sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2] + p[3] * p[3]);
        for (int i = 0; i < 4; i++) r.q[i] = p[i] / len;
        return r;
    }
    Se3Lite inverse() const {
        Se3Lite r;
        r.q[0] = -q[0]; r.q[1] = -q[1]; r.q[2] = -q[2]; r.q[3] = q[3];
        float rt[3];
        r.rotate(t, rt);
        for (int i = 0; i < 3; i++) r.t[i] = -rt[i];
        return r;
    }
};
// The members the direct tracker reads (include/Frame.h, include/KeyFrame.h) + the one an integration adds: mpDevPyramid.
struct DirectKeyFrameLite {
    std::vector<cv::KeyPoint> mvKeys;
    cv::Mat mDescriptors;
    std::vector<float> mvInvLevelSigma2;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    Se3Lite mTcw;
    bool mbBad = false;
    ygzorb_pyramid *mpDevPyramid = nullptr;
    Se3Lite GetPose() const { return mTcw; }
    bool isBad() const { return mbBad; }
};
struct DirectMapPointLite {
    float mWorldPos[3] = {0, 0, 0};
    bool mbBad = false;
    cv::Mat mDescriptor;
    std::map<DirectKeyFrameLite *, size_t> mObservations;
    const float *GetWorldPos() const { return mWorldPos; }
    bool isBad() const { return mbBad; }
};
struct DirectFrameLite {
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys;
    std::vector<DirectMapPointLite *> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    Se3Lite mTcw;
    ygzorb_pyramid *mpDevPyramid = nullptr;
};
class Frame {
public:
    PoseLite mTcw;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    std::vector<bool> mvbOutlier;
    std::vector<cv::Mat> mvImagePyramid;
    cv::Mat mImRight;
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysRight;
    cv::Mat mDescriptors, mDescriptorsRight;
    std::vector<float> mvuRight, mvDepth;
    std::vector<MapPointLite *> mvpMapPoints;
    DBoW2::BowVector mBowVec;
    DBoW2::FeatureVector mFeatVec;
    std::vector<float> mvScaleFactors, mvInvScaleFactors;
    float mnMinX = 0, mnMaxX = 0, mnMinY = 0, mnMaxY = 0;
    float mbf = 0, mb = 0, mfLogScaleFactor = 0;
    int mnScaleLevels = 0;
    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
};
struct KeyFrameLite {                        // include/KeyFrame.h: the members the matcher reads
    std::vector<cv::KeyPoint> mvKeys;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<MapPointLite *> mvpMapPoints;
    std::vector<MapPointLite *> GetMapPointMatches() const { return mvpMapPoints; }
    MapPointLite *GetMapPoint(size_t idx) const { return mvpMapPoints[idx]; }
    void AddMapPoint(MapPointLite *pMP, size_t idx) { mvpMapPoints[idx] = pMP; }
    std::vector<float> mvuRight, mvScaleFactors, mvInvLevelSigma2;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0, mfLogScaleFactor = 0;
    int mnScaleLevels = 0, mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    PoseLite mTcw;
    float mOw[3] = {0, 0, 0};
    PoseLite::Rot GetRotation() const { return mTcw.rotationMatrix(); }
    const float *GetTranslation() const { return mTcw.translation(); }
    const float *GetCameraCenter() const { return mOw; }
};
inline void MapPointLite::Replace(MapPointLite *pMP) {
    if (pMP == this) return;
    std::map<KeyFrameLite *, size_t> obs;
    obs.swap(mObservations);
    mbBad = true; mpReplaced = pMP;
    for (auto &o : obs) {
        if (!pMP->IsInKeyFrame(o.first)) { o.first->mvpMapPoints[o.second] = pMP; pMP->AddObservation(o.first, o.second); }
        else o.first->mvpMapPoints[o.second] = nullptr;
    }
}
}  // namespace ygz
#endif
