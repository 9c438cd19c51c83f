// ORBextractor.cc — host side of the drop-in ygz::ORBextractor: marshals cv::Mat / std::vector to the
// C ABI of libygzorb_cuda.so.  Mirrors src/ORBextractor.cc:412-470 (constructor state), :970-1028 and
// :1031-1127 (the two operator() overloads, ORBSLAM_KEYPOINT branch) and :1129-1150 (ComputePyramid).
#include "ORBextractor.h"

#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>

#ifndef YGZ_FRAME_H_
#include "FrameLite.h"
#endif

namespace ygz {

static void fail(const char *what) {
    throw std::runtime_error(std::string(what) + ": " + ygzorb_last_error());
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST, int device)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mHandle(nullptr) {
    ygzorb_params p{_nfeatures, _scaleFactor, _nlevels, _iniThFAST, _minThFAST};
    if (ygzorb_extractor_create(&p, 0, 0, 1, device, &mHandle) != 0) fail("ygzorb_extractor_create");
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
    mvImagePyramid.resize(nlevels);
    if (ygzorb_extractor_tables(mHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                                mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), nullptr) != 0)
        fail("ygzorb_extractor_tables");
}

ORBextractor::~ORBextractor() { ygzorb_extractor_destroy(mHandle); }

void ORBextractor::downloadPyramid(int w, int h) {
    std::vector<int> lw(nlevels), lh(nlevels);
    if (ygzorb_extractor_level_sizes(mHandle, w, h, lw.data(), lh.data()) != 0) fail("ygzorb_extractor_level_sizes");
    for (int l = 0; l < nlevels; l++) {
        mvImagePyramid[l].create(lh[l], lw[l], cv::CV_8UC1);
        if (ygzorb_get_pyramid_level(mHandle, 0, l, mvImagePyramid[l].data) != 0) fail("ygzorb_get_pyramid_level");
    }
}

void ORBextractor::ComputePyramid(cv::Mat image) {
    if (image.empty()) return;
    if (ygzorb_compute_pyramid_batch(mHandle, image.data, 1, image.cols, image.rows, (int)image.step, 0, 0) != 0)
        fail("ygzorb_compute_pyramid_batch");
    downloadPyramid(image.cols, image.rows);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint> &_keypoints,
                              cv::OutputArray _descriptors) {
    if (_image.empty()) return;                                       // :972-973
    cv::Mat image = _image.getMat();
    const int cap = ygzorb_extractor_max_keypoints(mHandle, image.cols, image.rows);
    if (cap < 0) fail("ygzorb_extractor_max_keypoints");
    mKpBuf.resize(cap); mDescBuf.resize((size_t)cap * 32);
    int32_t n = 0;
    if (ygzorb_extract_batch(mHandle, image.data, 1, image.cols, image.rows, (int)image.step, 0, 0, cap, &n, mKpBuf.data(),
                             mDescBuf.data()) != 0)
        fail("ygzorb_extract_batch");
    downloadPyramid(image.cols, image.rows);                          // ComputePyramid(image) side effect (:979)
    _keypoints.clear();
    if (n == 0) { _descriptors.release(); return; }                   // :990-991
    _descriptors.create(n, 32, cv::CV_8U);                            // :993
    cv::Mat descriptors = _descriptors.getMat();
    _keypoints.resize(n);
    static_assert(sizeof(cv::KeyPoint) == sizeof(ygzorb_keypoint), "layout");
    std::memcpy((void *)_keypoints.data(), mKpBuf.data(), (size_t)n * sizeof(ygzorb_keypoint));
    for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), &mDescBuf[(size_t)i * 32], 32);
}

void ORBextractor::operator()(Frame *frame, std::vector<cv::KeyPoint> &_keypoints, cv::OutputArray _descriptors,
                              KeyPointMethod method, bool leftEye) {
    if (method != ORBSLAM_KEYPOINT || frame->N > 0) {
        // FAST_KEYPOINT / DSO_KEYPOINT (:1040-1056) and descriptors for pre-existing keys (:1088-1099)
        // are the "next" row f3 of SURVEY.md §8(f): not on the named path, not built yet.
        std::fprintf(stderr, "ygz::ORBextractor: only ORBSLAM_KEYPOINT on a fresh Frame is implemented on the GPU path\n");
        _keypoints.clear();
        _descriptors.release();
        return;
    }
    // :1037-1040 — right eye: pyramid of mImRight; left eye: the Frame's own pyramid (level 0 = the image)
    const cv::Mat &src = leftEye ? frame->mvImagePyramid[0] : frame->mImRight;
    (*this)(src, cv::_InputArray(), _keypoints, _descriptors);
}

}  // namespace ygz
