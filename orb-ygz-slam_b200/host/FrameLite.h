// FrameLite.h — test double of the members of ygz::Frame (include/Frame.h:248-332) that the drop-in
// classes read or write.  In an integration the reference's own Frame.h is used instead (the member
// names below are the reference's).
#ifndef YGZ_HOST_FRAMELITE_H_
#define YGZ_HOST_FRAMELITE_H_
#include <map>
#include <vector>

#include "cvcompat.h"
#include "ORBVocabulary.h"

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
