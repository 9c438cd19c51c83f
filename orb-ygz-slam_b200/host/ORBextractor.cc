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

ORBextractor::~ORBextractor() {
    ygzorb_extractor_destroy(mHandle);
    ygzorb_host_free(mPyrSlab);
    ygzorb_host_free(mStage);
}

// Page-locked staging of one call: [image w*h][keypoints cap*28][descriptors cap*32].  The caller's cv::Mat / std::vector are
// pageable; one memcpy into / out of this block costs less than a pageable cudaMemcpy's own staging, and lets every transfer
// of the call run asynchronously beside the kernels.
void ORBextractor::ensureStage(size_t imageBytes, int cap) {
    const size_t need = ((imageBytes + 255) & ~(size_t)255) + (size_t)cap * (sizeof(ygzorb_keypoint) + 32) + 256;
    if (need > mStageBytes) {
        ygzorb_host_free(mStage);
        mStage = nullptr; mStageBytes = 0;
        void *p = nullptr;
        if (ygzorb_host_alloc(need, &p) != 0) fail("ygzorb_host_alloc");
        mStage = (uint8_t *)p; mStageBytes = need;
    }
    mStageKps = (ygzorb_keypoint *)(mStage + ((imageBytes + 255) & ~(size_t)255));
    mStageDesc = (uint8_t *)(mStageKps + cap);
    mStageCount = (int32_t *)(mStageDesc + (size_t)cap * 32 + ((4 - ((size_t)cap * 32) % 4) % 4));
}

// ComputePyramid's visible effect (:1129-1150): mvImagePyramid holds the levels.  They arrive in a page-locked slab through the
// extractor's pyramid sink (copied on the download stream while the rest of the extraction runs); the Mats are headers over it
// (continuous, step == cols), valid until the next call on this extractor like the reference's own mvImagePyramid.  Frame
// deep-copies them (src/Frame.cc:810-813).
void ORBextractor::preparePyramidSink(int w, int h) {
    mLw.resize(nlevels); mLh.resize(nlevels);
    if (ygzorb_extractor_level_sizes(mHandle, w, h, mLw.data(), mLh.data()) != 0) fail("ygzorb_extractor_level_sizes");
    size_t total = 0;
    for (int l = 0; l < nlevels; l++) total += (size_t)mLw[l] * mLh[l];
    if (total > mPyrSlabBytes) {
        ygzorb_extractor_set_pyramid_sink(mHandle, nullptr);
        ygzorb_host_free(mPyrSlab);
        mPyrSlab = nullptr; mPyrSlabBytes = 0;
        void *p = nullptr;
        if (ygzorb_host_alloc(total, &p) != 0) fail("ygzorb_host_alloc");
        mPyrSlab = (uint8_t *)p; mPyrSlabBytes = total;
    }
    if (ygzorb_extractor_set_pyramid_sink(mHandle, mPyrSlab) != 0) fail("ygzorb_extractor_set_pyramid_sink");
}

void ORBextractor::downloadPyramid(int /*w*/, int /*h*/) {   // the sink has been filled by the call that just returned
    size_t off = 0;
    for (int l = 0; l < nlevels; l++) {
        mvImagePyramid[l] = cv::Mat(mLh[l], mLw[l], cv::CV_8UC1, mPyrSlab + off, (size_t)mLw[l]);
        off += (size_t)mLw[l] * mLh[l];
    }
}

void ORBextractor::SetUndistortMaps(const cv::Mat &map1, const cv::Mat &map2) {
    const bool off = map1.empty() || map2.empty();
    if (ygzorb_extractor_set_undistort_maps(mHandle, off ? nullptr : (const int16_t *)map1.data, off ? nullptr : (const uint16_t *)map2.data,
                                            map2.cols, map2.rows) != 0)
        fail("ygzorb_extractor_set_undistort_maps");
}

void ORBextractor::ComputePyramid(cv::Mat image) {
    if (image.empty()) return;
    preparePyramidSink(image.cols, image.rows);
    ensureStage((size_t)image.cols * image.rows, 1);
    for (int r = 0; r < image.rows; r++) std::memcpy(mStage + (size_t)r * image.cols, image.ptr(r), (size_t)image.cols);
    if (ygzorb_compute_pyramid_batch(mHandle, mStage, 1, image.cols, image.rows, image.cols, 0, 0) != 0)
        fail("ygzorb_compute_pyramid_batch");
    downloadPyramid(image.cols, image.rows);
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint> &_keypoints,
                              cv::OutputArray _descriptors) {
    if (_image.empty()) return;                                       // :972-973
    cv::Mat image = _image.getMat();
    const int cap = ygzorb_extractor_max_keypoints(mHandle, image.cols, image.rows);
    if (cap < 0) fail("ygzorb_extractor_max_keypoints");
    preparePyramidSink(image.cols, image.rows);
    ensureStage((size_t)image.cols * image.rows, cap);
    for (int r = 0; r < image.rows; r++) std::memcpy(mStage + (size_t)r * image.cols, image.ptr(r), (size_t)image.cols);
    // the staging slab is page-locked (ygzorb_host_alloc): the last kernel writes the results into it directly
    if (ygzorb_extract_batch(mHandle, mStage, 1, image.cols, image.rows, image.cols, 0, YGZORB_OUT_PINNED, cap, mStageCount, mStageKps, mStageDesc) != 0)
        fail("ygzorb_extract_batch");
    const int32_t n = *mStageCount;
    downloadPyramid(image.cols, image.rows);                          // ComputePyramid(image) side effect (:979)
    _keypoints.clear();
    if (n == 0) { _descriptors.release(); return; }                   // :990-991
    _descriptors.create(n, 32, cv::CV_8U);                            // :993
    cv::Mat descriptors = _descriptors.getMat();
    _keypoints.resize(n);
    static_assert(sizeof(cv::KeyPoint) == sizeof(ygzorb_keypoint), "layout");
    std::memcpy((void *)_keypoints.data(), mStageKps, (size_t)n * sizeof(ygzorb_keypoint));
    for (int i = 0; i < n; i++) std::memcpy(descriptors.ptr(i), mStageDesc + (size_t)i * 32, 32);
}

void ORBextractor::operator()(Frame *frame, std::vector<cv::KeyPoint> &_keypoints, cv::OutputArray _descriptors,
                              KeyPointMethod method, bool leftEye) {
    if (method == DSO_KEYPOINT && leftEye) {
        // :1052-1058 + :1079-1125: grid FAST-10 / Shi-Tomasi keypoints on level 0 next to the keypoints direct tracking left in
        // frame->mvKeys, orientation and descriptors for both (Frame::ExtractORB, src/Frame.cc:335-337)
        const cv::Mat &img = frame->mvImagePyramid[0];
        mvImagePyramid = frame->mvImagePyramid;                       // :1039
        const int nExist = frame->N;
        const int cap = nExist + 3 * (img.rows / 7) * (img.cols / 7);
        mKpBuf.resize(cap); mDescBuf.resize((size_t)cap * 32);
        if (nExist) std::memcpy((void *)mKpBuf.data(), frame->mvKeys.data(), (size_t)nExist * sizeof(ygzorb_keypoint));
        int32_t n = 0, ne = nExist;
        if (ygzorb_extract_dso_batch(mHandle, img.data, 1, img.cols, img.rows, (int)img.step, 0, 0, &mnGridSize, &ne, cap, &n,
                                     mKpBuf.data(), mDescBuf.data()) != 0)
            fail("ygzorb_extract_dso_batch");
        const cv::KeyPoint *rec = (const cv::KeyPoint *)mKpBuf.data();
        for (int i = 0; i < nExist; i++) frame->mvKeys[i].angle = rec[i].angle;      // :1383-1385
        const bool same = &_keypoints == &frame->mvKeys;
        const size_t before = _keypoints.size();
        _keypoints.insert(_keypoints.end(), rec + nExist, rec + n);                   // :1124
        const int rows = (int)(same ? _keypoints.size() : before + (size_t)(n - nExist));
        if (rows == 0) { _descriptors.release(); return; }
        _descriptors.create(rows, 32, cv::CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        for (int i = 0; i < std::min(rows, (int)n); i++) std::memcpy(descriptors.ptr(i), &mDescBuf[(size_t)i * 32], 32);
        return;
    }
    if (method != ORBSLAM_KEYPOINT) {
        // FAST_KEYPOINT (:1045-1051, ComputeKeyPointsFast): the reference's own header says not to call it; not built.
        std::fprintf(stderr, "ygz::ORBextractor: FAST_KEYPOINT is not implemented on the GPU path\n");
        _keypoints.clear();
        _descriptors.release();
        return;
    }
    if (!leftEye) {
        // :1037-1038 + :1059-1060 — right eye: ComputePyramid(frame->mImRight) + the plain OctTree, whatever the Frame already
        // holds (frame->N belongs to the left image; the existing-key rows :1086-1099 are left-eye only).  This is what
        // Frame::ExtractORB(1, ...) runs on every stereo frame, also those that come out of direct tracking (src/Frame.cc:345).
        // Rows: the reference sizes the Mat for _keypoints.size() + new keys and fills the new keys' rows from 0 (offset = 0).
        const size_t before = _keypoints.size();
        std::vector<cv::KeyPoint> fresh;
        cv::Mat d;
        (*this)(frame->mImRight, cv::_InputArray(), fresh, d);
        const int rows = (int)(before + fresh.size());
        _keypoints.insert(_keypoints.end(), fresh.begin(), fresh.end());
        if (rows == 0) { _descriptors.release(); return; }
        _descriptors.create(rows, 32, cv::CV_8U);
        cv::Mat descriptors = _descriptors.getMat();
        for (size_t i = 0; i < fresh.size(); i++) std::memcpy(descriptors.ptr((int)i), d.ptr((int)i), 32);
        for (int i = (int)fresh.size(); i < rows; i++) std::memset(descriptors.ptr(i), 0, 32);   // uninitialised in the reference
        return;
    }
    // left eye: the Frame's own pyramid (:1039-1040; level 0 = the image, the levels above are what ComputePyramid gives)
    if (frame->N <= 0) {
        (*this)(frame->mvImagePyramid[0], cv::_InputArray(), _keypoints, _descriptors);
        return;
    }
    // :1086-1099 — a Frame that already holds keypoints: rows [0, N) = descriptors of frame->mvKeys at
    // pt * mvInvScaleFactor[octave] with the angle each key carries, then the new OctTree keys
    const int nExist = frame->N;
    std::vector<cv::KeyPoint> fresh;
    cv::Mat d;
    (*this)(frame->mvImagePyramid[0], cv::_InputArray(), fresh, d);
    std::vector<uint8_t> ed((size_t)nExist * 32);
    if (ygzorb_describe_existing(mHandle, 0, nExist, (const ygzorb_keypoint *)frame->mvKeys.data(), ed.data()) != 0)
        fail("ygzorb_describe_existing");
    const int rows = (int)(_keypoints.size() + fresh.size());                       // :1069 nkeypoints += _keypoints.size()
    _keypoints.insert(_keypoints.end(), fresh.begin(), fresh.end());
    _descriptors.create(rows, 32, cv::CV_8U);
    cv::Mat descriptors = _descriptors.getMat();
    for (int i = 0; i < rows; i++) std::memset(descriptors.ptr(i), 0, 32);
    for (int i = 0; i < std::min(nExist, rows); i++) std::memcpy(descriptors.ptr(i), &ed[(size_t)i * 32], 32);
    for (size_t i = 0; i < fresh.size() && nExist + (int)i < rows; i++) std::memcpy(descriptors.ptr(nExist + (int)i), d.ptr((int)i), 32);
}

}  // namespace ygz
