// host_demo.cc — exercises the drop-in C++ classes the way the reference's callers do
// (Frame::ExtractORB src/Frame.cc:332-348, Tracking::SearchLocalPoints src/Tracking.cc:1662-1674,
// Frame::ComputeStereoMatches src/Frame.cc:747) and dumps the results for tests/test_gpu_host.py.
// usage: host_demo <in.bin> <out.bin> [voc.bin bow_out.bin]
//   with a vocabulary (binary format): Frame::ComputeBoW (src/Frame.cc:495-500) on the left descriptors ->
//   bow_out.bin: int32 nb | nb x {int32 word, double value} | int32 nn | nn x {int32 node, int32 cnt, cnt x int32 feature}
//   in.bin : int32 w, h, nfeatures, nMP | left image w*h | right image w*h |
//            nMP x {float projX, projY, viewCos; int32 level; 32 B descriptor} |
//            int32 nLast | 30 float pose | nLast x {float wx, wy, wz, angle; int32 octave, valid, hasObs; 32 B descriptor}
//   out.bin: int32 nL | nL x 28 B keypoints | nL x 32 B desc | int32 nR | nR keypoints | nR desc |
//            nL x float uRight | nL x float depth | int32 nmatches | nL x int32 (MapPoint index per keypoint, -1 none) |
//            int32 nmatchesLast | nL x int32 (LastFrame index per keypoint, -1 none)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "FrameLite.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"
#include "ORBVocabulary.h"

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
    int32_t nLast = 0;
    rd(fi, &nLast, 1);
    float pose[30];
    rd(fi, pose, 30);
    std::vector<MapPointLite> lastMps(nLast);
    Frame Last;
    Last.N = nLast; Last.mvKeys.resize(nLast); Last.mvpMapPoints.assign(nLast, nullptr); Last.mvbOutlier.assign(nLast, false);
    for (int i = 0; i < nLast; i++) {
        float v[4]; int32_t q[3];
        rd(fi, v, 4); rd(fi, q, 3);
        for (int r = 0; r < 3; r++) lastMps[i].mWorldPos[r] = v[r];
        Last.mvKeys[i].angle = v[3]; Last.mvKeys[i].octave = q[0];
        lastMps[i].nObs = q[2];
        lastMps[i].mDescriptor.create(1, 32, cv::CV_8U);
        rd(fi, lastMps[i].mDescriptor.data, 32);
        if (q[1]) Last.mvpMapPoints[i] = &lastMps[i];
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

    std::vector<int32_t> assign(F.N);
    for (int i = 0; i < F.N; i++) assign[i] = F.mvpMapPoints[i] ? (int32_t)(F.mvpMapPoints[i] - mps.data()) : -1;
    // Tracking::TrackWithMotionModel (src/Tracking.cc:1017-1026): clear, then SearchByProjection(Current, Last, th, mono)
    for (int r = 0; r < 9; r++) { F.mTcw.R[r] = pose[r]; Last.mTcw.R[r] = pose[12 + r]; }
    for (int r = 0; r < 3; r++) { F.mTcw.t[r] = pose[9 + r]; Last.mTcw.t[r] = pose[21 + r]; }
    F.fx = pose[24]; F.fy = pose[25]; F.cx = pose[26]; F.cy = pose[27]; F.mbf = pose[28]; F.mb = pose[29];
    F.mvpMapPoints.assign(F.N, nullptr);
    ORBmatcher motionMatcher(0.9f, true);
    const int nmatchesLast = motionMatcher.SearchByProjection(F, Last, 15.0f, true);

    FILE *fo = fopen(argv[2], "wb");
    int32_t nL = F.N, nR = (int32_t)F.mvKeysRight.size();
    wr(fo, &nL, 1); wr(fo, F.mvKeys.data(), nL);
    for (int i = 0; i < nL; i++) wr(fo, F.mDescriptors.ptr(i), 32);
    wr(fo, &nR, 1); wr(fo, F.mvKeysRight.data(), nR);
    for (int i = 0; i < nR; i++) wr(fo, F.mDescriptorsRight.ptr(i), 32);
    wr(fo, uRight.data(), nL); wr(fo, depth.data(), nL);
    int32_t nm = nmatches;
    wr(fo, &nm, 1);
    wr(fo, assign.data(), nL);
    int32_t nml = nmatchesLast;
    wr(fo, &nml, 1);
    for (int i = 0; i < nL; i++) { int32_t idx = F.mvpMapPoints[i] ? (int32_t)(F.mvpMapPoints[i] - lastMps.data()) : -1; wr(fo, &idx, 1); }
    {   // Frame::ExtractORB on a frame that comes out of direct tracking (src/Frame.cc:335-337): every 7th keypoint of the
        // ORB extraction plays the tracked features, two consecutive frames share the extractor's grid size
        for (int rep = 0; rep < 2; rep++) {
            Frame D;
            D.mvImagePyramid = F.mvImagePyramid;
            for (int i = 0; i < nL; i += 7) D.mvKeys.push_back(F.mvKeys[i]);
            D.N = (int)D.mvKeys.size();
            cv::Mat dd;
            exL(&D, D.mvKeys, dd, ORBextractor::DSO_KEYPOINT, true);
            int32_t nd = (int32_t)D.mvKeys.size();
            wr(fo, &nd, 1); wr(fo, D.mvKeys.data(), nd);
            for (int i = 0; i < nd; i++) wr(fo, dd.ptr(i), 32);
            if (rep == 1) {
                // ... and the right image of that stereo frame: Frame::ExtractORB(1, ...) calls the Right extractor with
                // ORBSLAM_KEYPOINT, leftEye = false on a Frame whose N is already > 0 (src/Frame.cc:345, :728-731)
                D.N = nd;
                D.mImRight = right;
                exR(&D, D.mvKeysRight, D.mDescriptorsRight, ORBextractor::ORBSLAM_KEYPOINT, false);
                int32_t nr = (int32_t)D.mvKeysRight.size();
                wr(fo, &nr, 1); wr(fo, D.mvKeysRight.data(), nr);
                for (int i = 0; i < nr; i++) wr(fo, D.mDescriptorsRight.ptr(i), 32);
                for (int l = 0; l < exR.GetLevels(); l++) {        // the Right extractor's pyramid was refreshed (:1037-1038)
                    int32_t wh[2] = {exR.mvImagePyramid[l].cols, exR.mvImagePyramid[l].rows};
                    wr(fo, wh, 2);
                    for (int r = 0; r < wh[1]; r++) wr(fo, exR.mvImagePyramid[l].ptr(r), (size_t)wh[0]);
                }
            }
        }
        // ORBSLAM_KEYPOINT, leftEye = true on a Frame that already holds keypoints (src/ORBextractor.cc:1086-1099)
        Frame E;
        E.mvImagePyramid = F.mvImagePyramid;
        for (int i = 0; i < nL; i += 7) E.mvKeys.push_back(F.mvKeys[i]);
        E.N = (int)E.mvKeys.size();
        cv::Mat de;
        exL(&E, E.mvKeys, de, ORBextractor::ORBSLAM_KEYPOINT, true);
        int32_t ne = (int32_t)E.mvKeys.size();
        wr(fo, &ne, 1); wr(fo, E.mvKeys.data(), ne);
        for (int i = 0; i < ne; i++) wr(fo, de.ptr(i), 32);
    }
    fclose(fo);
    if (argc >= 5) {
        ORBVocabulary voc;
        if (!voc.loadFromBinaryFile(argv[3])) { fprintf(stderr, "cannot load vocabulary\n"); return 2; }
        std::vector<cv::Mat> vCurrentDesc;                    // Converter::toDescriptorVector (src/Frame.cc:497)
        for (int i = 0; i < F.mDescriptors.rows; i++) vCurrentDesc.push_back(F.mDescriptors.row(i));
        DBoW2::BowVector bowVec;
        DBoW2::FeatureVector featVec;
        voc.transform(vCurrentDesc, bowVec, featVec, 4);
        FILE *fb = fopen(argv[4], "wb");
        int32_t nb = (int32_t)bowVec.size();
        wr(fb, &nb, 1);
        for (const auto &e : bowVec) { int32_t w = (int32_t)e.first; double v = e.second; wr(fb, &w, 1); wr(fb, &v, 1); }
        int32_t nn = (int32_t)featVec.size();
        wr(fb, &nn, 1);
        for (const auto &e : featVec) {
            int32_t a[2] = {(int32_t)e.first, (int32_t)e.second.size()};
            wr(fb, a, 2);
            for (unsigned idx : e.second) { int32_t ii = (int32_t)idx; wr(fb, &ii, 1); }
        }
        // Tracking::TrackReferenceKeyFrame (src/Tracking.cc:955-964): a keyframe made of the right image's features, every
        // third one without a MapPoint and every 7th one bad; then LoopClosing's keyframe-to-keyframe overload.
        F.mFeatVec = featVec;
        KeyFrameLite kfR, kfL;
        std::vector<MapPointLite> mpsR(F.mvKeysRight.size()), mpsL(F.N);
        auto fill = [&](KeyFrameLite &kf, const std::vector<cv::KeyPoint> &keys, const cv::Mat &desc, std::vector<MapPointLite> &pts) {
            kf.mvKeys = keys; kf.mDescriptors = desc;
            std::vector<cv::Mat> v;
            for (int i = 0; i < desc.rows; i++) v.push_back(desc.row(i));
            DBoW2::BowVector bv;
            voc.transform(v, bv, kf.mFeatVec, 4);
            kf.mvpMapPoints.assign(keys.size(), nullptr);
            for (size_t i = 0; i < keys.size(); i++) {
                pts[i].mbBad = i % 7 == 3;
                if (i % 3 != 1) kf.mvpMapPoints[i] = &pts[i];
            }
        };
        fill(kfR, F.mvKeysRight, F.mDescriptorsRight, mpsR);
        fill(kfL, F.mvKeys, F.mDescriptors, mpsL);
        ORBmatcher bowMatcher(0.7f, true);
        std::vector<MapPointLite *> vpMatches, vpMatches12;
        int32_t nbow = bowMatcher.SearchByBoW(&kfR, F, vpMatches);
        wr(fb, &nbow, 1);
        for (int i = 0; i < F.N; i++) { int32_t idx = vpMatches[i] ? (int32_t)(vpMatches[i] - mpsR.data()) : -1; wr(fb, &idx, 1); }
        ORBmatcher loopMatcher(0.75f, true);
        int32_t nbow12 = loopMatcher.SearchByBoW(&kfL, &kfR, vpMatches12);
        wr(fb, &nbow12, 1);
        for (int i = 0; i < F.N; i++) { int32_t idx = vpMatches12[i] ? (int32_t)(vpMatches12[i] - mpsR.data()) : -1; wr(fb, &idx, 1); }
        fclose(fb);
        printf("host_demo: SearchByBoW keyframe->frame %d, keyframe->keyframe %d matches\n", nbow, nbow12);
        printf("host_demo: vocabulary with %u words, BowVector %d words, FeatureVector %d nodes\n", voc.size(), nb, nn);
    }
    printf("host_demo: %d left / %d right keypoints, %d projection matches, %d motion-model matches, DescriptorDistance(self)=%d\n", nL, nR, nmatches, nmatchesLast,
           nL ? ORBmatcher::DescriptorDistance(F.mDescriptors.row(0), F.mDescriptors.row(0)) : 0);
    return 0;
}
