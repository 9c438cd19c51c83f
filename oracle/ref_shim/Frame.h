// Frame.h — TEST INFRASTRUCTURE ONLY: the four members of ygz::Frame (reference include/Frame.h:248-332)
// that src/ORBextractor.cc touches (:1038-1040, 1047, 1056, 1089-1098).  Shadows the real header,
// which needs Eigen/sophus/DBoW2.
#if !defined(YGZ_REF_SHIM_FRAME_H_) && !defined(YGZ_FRAME_H_)   // YGZ_FRAME_H_: match_prelude.h brings its own, larger shim
#define YGZ_REF_SHIM_FRAME_H_
namespace ygz {
class Frame {
public:
    std::vector<cv::Mat> mvImagePyramid;  // Frame.h: pyramid cloned from the extractor (Frame.cc:810-813)
    cv::Mat mImRight;
    std::vector<cv::KeyPoint> mvKeys;
    int N = 0;  // uninitialised in the reference (Frame.h:285); canonicalised to 0 for a fresh Frame
};
}  // namespace ygz
#endif
