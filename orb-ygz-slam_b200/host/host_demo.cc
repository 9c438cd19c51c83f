// host_demo.cc — exercises the drop-in C++ classes the way the reference's callers do
// (Frame::ExtractORB src/Frame.cc:332-348, Tracking::SearchLocalPoints src/Tracking.cc:1662-1674,
// Frame::ComputeStereoMatches src/Frame.cc:747) and dumps the results for tests/test_gpu_host.py.
// usage: host_demo <in.bin> <out.bin>
//   in.bin : int32 w, h, nfeatures, nMP | left image w*h | right image w*h |
//            nMP x {float projX, projY, viewCos; int32 level; 32 B descriptor}
//   out.bin: int32 nL | nL x 28 B keypoints | nL x 32 B desc | int32 nR | nR keypoints | nR desc |
//            nL x float uRight | nL x float depth | int32 nmatches | nL x int32 (MapPoint index per keypoint, -1 none)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "FrameLite.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ygz;

template <class T> static void rd(FILE *f, T *p, size_t n) { if (fread(p, sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); } }
template <class T> static void wr(FILE *f, const T *p, size_t n) { fwrite(p, sizeof(T), n, f); }

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: host_demo in.bin out.bin\n"); return 2; }
    FILE *fi = fopen(argv[1], "rb");
    if (!fi) { perror("in"); return 2; }
    int32_t hdr[4];
    rd(fi, hdr, 4);
    const int w = hdr[0], h = hdr[1], nfeat = hdr[2], nMP = hdr[3];
    cv::Mat left(h, w, cv::CV_8UC1), right(h, w, cv::CV_8UC1);
    rd(fi, left.data, (size_t)w * h); rd(fi, right.data, (size_t)w * h);
    std::vector<MapPointLite> mps(nMP);
    for (int i = 0; i < nMP; i++) {
        float v[3]; int32_t lvl;
        rd(fi, v, 3); rd(fi, &lvl, 1);
        mps[i].mTrackProjX = v[0]; mps[i].mTrackProjY = v[1]; mps[i].mTrackViewCos = v[2]; mps[i].mnTrackScaleLevel = lvl;
        mps[i].mDescriptor.create(1, 32, cv::CV_8U);
        rd(fi, mps[i].mDescriptor.data, 32);
    }
    fclose(fi);

    ORBextractor exL(nfeat, 1.2f, 8, 20, 7), exR(nfeat, 1.2f, 8, 20, 7);
    Frame F;
    F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
    // Frame ctor: ComputeImagePyramid -> extractor->ComputePyramid + clone (src/Frame.cc:807-813)
    exL.ComputePyramid(left);
    F.mvImagePyramid.resize(exL.GetLevels());
    for (int l = 0; l < exL.GetLevels(); l++) F.mvImagePyramid[l] = exL.mvImagePyramid[l].clone();
    F.mImRight = right;
    // Frame::ExtractORB (src/Frame.cc:332-348)
    exL(&F, F.mvKeys, F.mDescriptors, ORBextractor::ORBSLAM_KEYPOINT, true);
    exR(&F, F.mvKeysRight, F.mDescriptorsRight, ORBextractor::ORBSLAM_KEYPOINT, false);
    F.N = (int)F.mvKeys.size();
    F.mvScaleFactors = exL.GetScaleFactors(); F.mvInvScaleFactors = exL.GetInverseScaleFactors();
    F.mnMinX = 0; F.mnMinY = 0; F.mnMaxX = (float)w; F.mnMaxY = (float)h;      // src/Frame.cc:502-507 (no distortion)
    F.mbf = 386.1448f; F.mb = F.mbf / 718.856f;
    F.mvpMapPoints.assign(F.N, nullptr);
    ORBmatcher stereoMatcher;
    stereoMatcher.ComputeStereoMatches(F);
    std::vector<float> uRight = F.mvuRight, depth = F.mvDepth;
    // Tracking::SearchLocalPoints: ORBmatcher matcher(0.8); matcher.SearchByProjection(F, mps, th, false)
    F.mvuRight.assign(F.N, -1.0f);     // monocular search, like config 2
    std::vector<MapPointLite *> vp(nMP);
    for (int i = 0; i < nMP; i++) vp[i] = &mps[i];
    ORBmatcher matcher(0.8f);
    const int nmatches = matcher.SearchByProjection(F, vp, 1.0f, false);

    FILE *fo = fopen(argv[2], "wb");
    int32_t nL = F.N, nR = (int32_t)F.mvKeysRight.size();
    wr(fo, &nL, 1); wr(fo, F.mvKeys.data(), nL);
    for (int i = 0; i < nL; i++) wr(fo, F.mDescriptors.ptr(i), 32);
    wr(fo, &nR, 1); wr(fo, F.mvKeysRight.data(), nR);
    for (int i = 0; i < nR; i++) wr(fo, F.mDescriptorsRight.ptr(i), 32);
    wr(fo, uRight.data(), nL); wr(fo, depth.data(), nL);
    int32_t nm = nmatches;
    wr(fo, &nm, 1);
    for (int i = 0; i < nL; i++) { int32_t idx = F.mvpMapPoints[i] ? (int32_t)(F.mvpMapPoints[i] - mps.data()) : -1; wr(fo, &idx, 1); }
    fclose(fo);
    printf("host_demo: %d left / %d right keypoints, %d projection matches, ORBmatcher::DescriptorDistance(self)=%d\n", nL, nR, nmatches,
           nL ? ORBmatcher::DescriptorDistance(F.mDescriptors.row(0), F.mDescriptors.row(0)) : 0);
    return 0;
}
