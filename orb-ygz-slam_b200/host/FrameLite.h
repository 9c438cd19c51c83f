// FrameLite.h — test double of the members of ygz::Frame (include/Frame.h:248-332) that the drop-in
// classes read or write.  In an integration the reference's own Frame.h is used instead (the member
// names below are the reference's).
#ifndef YGZ_HOST_FRAMELITE_H_
#define YGZ_HOST_FRAMELITE_H_
#include <vector>

#include "cvcompat.h"
#include "ORBVocabulary.h"

namespace ygz {
class ORBextractor;
struct MapPointLite {                       // include/MapPoint.h:95-102 + the calls SearchByProjection makes
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 1;
    int mnTrackScaleLevel = 0;
    bool mbTrackInView = true, mbBad = false;
    int nObs = 1;
    cv::Mat mDescriptor;
    float mWorldPos[3] = {0, 0, 0};
    const float *GetWorldPos() const { return mWorldPos; }       // Vector3f in the reference
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
    float mbf = 0, mb = 0;
    ORBextractor *mpORBextractorLeft = nullptr, *mpORBextractorRight = nullptr;
};
struct KeyFrameLite {                        // include/KeyFrame.h: the members SearchByBoW reads
    std::vector<cv::KeyPoint> mvKeys;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<MapPointLite *> mvpMapPoints;
    std::vector<MapPointLite *> GetMapPointMatches() const { return mvpMapPoints; }
};
}  // namespace ygz
#endif
