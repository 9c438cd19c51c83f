// ORBextractor.h — drop-in for the reference's include/ORBextractor.h: same namespace, class name,
// constructor, both operator() overloads, ComputePyramid, the public mvImagePyramid and the getters
// (include/ORBextractor.h:49-175).  The work is done by libygzorb_cuda.so through include/ygzorb.h;
// there is no CPU implementation behind this class.
#ifndef YGZ_HOST_ORBEXTRACTOR_H_
#define YGZ_HOST_ORBEXTRACTOR_H_

#include <vector>

#include "cvcompat.h"
#include "../../include/ygzorb.h"

namespace ygz {

class Frame;   // the reference's include/Frame.h (or host/FrameLite.h in the tests)

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };                                        // ORBextractor.h:43-45
    typedef enum { FAST_KEYPOINT, ORBSLAM_KEYPOINT, DSO_KEYPOINT } KeyPointMethod;    // :47-51

    // :53-57 — device = CUDA device ordinal (new, defaulted: existing call sites compile unchanged)
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0);
    ~ORBextractor();
    ORBextractor(const ORBextractor &) = delete;
    ORBextractor &operator=(const ORBextractor &) = delete;

    // :71-75 — ORB on an image; the mask is ignored, as in the reference.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints,
                    cv::OutputArray descriptors);
    // :79-85 — the Frame overload (src/ORBextractor.cc:1031-1127): ORBSLAM_KEYPOINT on a fresh Frame, DSO_KEYPOINT on a
    // Frame holding the keypoints of direct tracking.
    void operator()(Frame *frame, std::vector<cv::KeyPoint> &keypoints, cv::OutputArray descriptors,
                    KeyPointMethod method, bool leftEye = true);
    // New: the maps of Frame::ComputeImagePyramid (src/Frame.cc:776-784, CV_16SC2 + CV_16UC1, continuous).  Once set, every image
    // handed to this extractor is remapped on the device first (:788) and mvImagePyramid[0] is the undistorted image.
    void SetUndistortMaps(const cv::Mat &map1, const cv::Mat &map2);
    // :114 — fills mvImagePyramid (continuous host images, as Frame clones them, src/Frame.cc:810-813).
    void ComputePyramid(cv::Mat image);

    int inline GetLevels() { return nlevels; }                                          // :87-109
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;                                                // :112

    ygzorb_extractor *handle() const { return mHandle; }   // for Frame::ComputeStereoMatches (device-resident pyramids)

protected:
    void downloadPyramid(int w, int h);
    void preparePyramidSink(int w, int h);
    void ensureStage(size_t imageBytes, int cap);
    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
    ygzorb_extractor *mHandle;
    int32_t mnGridSize = -1;                              // :191, dynamic grid size of the DSO keypoint mode
    std::vector<ygzorb_keypoint> mKpBuf;
    std::vector<uint8_t> mDescBuf;
    uint8_t *mPyrSlab = nullptr;                          // page-locked slab behind mvImagePyramid
    size_t mPyrSlabBytes = 0;
    std::vector<int> mLw, mLh;                            // level sizes of the current image size
    uint8_t *mStage = nullptr; size_t mStageBytes = 0;    // page-locked staging of one call (image in, keypoints + descriptors out)
    ygzorb_keypoint *mStageKps = nullptr; uint8_t *mStageDesc = nullptr; int32_t *mStageCount = nullptr;
};

}  // namespace ygz
#endif
