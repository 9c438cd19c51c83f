// host_demo3.cc — drives the drop-ins of SURVEY 8(f) row f4 (host/DirectTracking.h) on Frame / KeyFrame / MapPoint objects the
// way Tracking::TrackWithSparseAlignment (src/Tracking.cc:2171), Tracking::SearchLocalPointsDirect (:2284-2305) and
// LocalMapping (MapPoint::ComputeDistinctiveDescriptors) use them, and dumps the outcome for tests/test_gpu_host.py.
// usage: host_demo3 <in.bin> <out.bin>
//   in.bin : int32 nlevels | ref pyramid: per level int32 w, h, w*h bytes | cur pyramid likewise |
//            float scale[nlevels] | float invScale[nlevels] | float invSigma2[nlevels] | float intr[4] |
//            float refPose[7] | float curPose[7] | int32 n | n x 28 B keypoints | n x uint8 valid | n x 3 float world |
//            [direct]   int32 m | m x { int32 kp index, float px_curr[2] } |
//            [distinct] int32 nPoints | int32 obsStart[nPoints + 1] | rows x 32 B
//   out.bin: float TCR[7] | int64 tracked | m x { float px[2], int32 level, int32 ok } | nPoints x 32 B mDescriptor
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "DirectTracking.h"
#include "FrameLite.h"

using namespace ygz;

template <class T> static void rd(FILE *f, T *p, size_t n) { if (n && fread(p, sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); } }
template <class T> static void wr(FILE *f, const T *p, size_t n) { fwrite(p, sizeof(T), n, f); }

static ygzorb_pyramid *readPyramid(FILE *fi, int nlevels) {
    std::vector<std::vector<uint8_t>> px(nlevels);
    std::vector<const uint8_t *> ptr(nlevels);
    std::vector<int32_t> w(nlevels), h(nlevels);
    for (int l = 0; l < nlevels; l++) {
        rd(fi, &w[l], 1); rd(fi, &h[l], 1);
        px[l].resize((size_t)w[l] * h[l]);
        rd(fi, px[l].data(), px[l].size());
        ptr[l] = px[l].data();
    }
    ygzorb_pyramid *p = nullptr;
    if (ygzorb_pyramid_create(0, ptr.data(), w.data(), h.data(), w.data(), nlevels, 0, &p) != 0) { fprintf(stderr, "%s\n", ygzorb_last_error()); exit(3); }
    return p;
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: host_demo3 in.bin out.bin\n"); return 2; }
    FILE *fi = fopen(argv[1], "rb"), *fo = fopen(argv[2], "wb");
    if (!fi || !fo) { perror("open"); return 2; }
    ygzorb_matcher *handle = nullptr;
    if (ygzorb_matcher_create(0, &handle) != 0) { fprintf(stderr, "%s\n", ygzorb_last_error()); return 3; }
    int32_t nlevels;
    rd(fi, &nlevels, 1);
    DirectFrameLite ref, cur;
    DirectKeyFrameLite kf;
    ref.mpDevPyramid = readPyramid(fi, nlevels);
    cur.mpDevPyramid = readPyramid(fi, nlevels);
    std::vector<float> scale(nlevels), invScale(nlevels), invSigma2(nlevels);
    rd(fi, scale.data(), nlevels); rd(fi, invScale.data(), nlevels); rd(fi, invSigma2.data(), nlevels);
    float intr[4], poseR[7], poseC[7];
    rd(fi, intr, 4); rd(fi, poseR, 7); rd(fi, poseC, 7);
    for (DirectFrameLite *F : {&ref, &cur}) {
        F->mvScaleFactors = scale; F->mvInvScaleFactors = invScale;
        F->fx = intr[0]; F->fy = intr[1]; F->cx = intr[2]; F->cy = intr[3];
    }
    ref.mTcw = Se3Lite::FromFloats(poseR); cur.mTcw = Se3Lite::FromFloats(poseC);
    int32_t n;
    rd(fi, &n, 1);
    ref.N = n; ref.mvKeys.resize(n);
    rd(fi, ref.mvKeys.data(), n);
    std::vector<uint8_t> valid(n);
    rd(fi, valid.data(), n);
    std::vector<DirectMapPointLite> pts(n);
    ref.mvpMapPoints.assign(n, nullptr); ref.mvbOutlier.assign(n, false);
    for (int i = 0; i < n; i++) {
        rd(fi, pts[i].mWorldPos, 3);
        if (valid[i] == 1) ref.mvpMapPoints[i] = &pts[i];
        else if (valid[i] == 2) { ref.mvpMapPoints[i] = &pts[i]; ref.mvbOutlier[i] = true; }     // outlier flag
        else if (valid[i] == 3) { ref.mvpMapPoints[i] = &pts[i]; pts[i].mbBad = true; }          // bad MapPoint
    }
    // ---- SparseImgAlign::run, as Tracking::TrackWithSparseAlignment calls it
    size_t tracked = 0;
    int alignIterations = 0;
    Se3Lite TCR;
    {
        SparseImgAlign align(handle, nlevels - 1, 1);     // lives shorter than the matcher handle it runs on
        tracked = align.run(&ref, &cur, TCR);
        alignIterations = align.iterations();
    }
    float T[7];
    PoseToFloats(TCR, T);
    wr(fo, T, 7);
    const int64_t tr = (int64_t)tracked;
    wr(fo, &tr, 1);
    // ---- FindDirectProjection: the reference frame promoted to a keyframe (shares the pyramid), requests from the file
    kf.mvKeys = ref.mvKeys; kf.mvInvLevelSigma2 = invSigma2; kf.fx = intr[0]; kf.fy = intr[1]; kf.cx = intr[2]; kf.cy = intr[3];
    kf.mTcw = ref.mTcw; kf.mpDevPyramid = ref.mpDevPyramid;
    int32_t m;
    rd(fi, &m, 1);
    std::vector<DirectRequest<DirectKeyFrameLite, DirectMapPointLite>> reqs(m);
    for (int i = 0; i < m; i++) {
        int32_t idx;
        rd(fi, &idx, 1); rd(fi, reqs[i].px_curr, 2);
        pts[idx].mObservations[&kf] = (size_t)idx;
        reqs[i].ref = &kf; reqs[i].mp = &pts[idx];
    }
    FindDirectProjectionBatch(handle, &cur, reqs);
    for (int i = 0; i < m; i++) {
        wr(fo, reqs[i].px_curr, 2);
        const int32_t lv = reqs[i].search_level, ok = reqs[i].success ? 1 : 0;
        wr(fo, &lv, 1); wr(fo, &ok, 1);
    }
    // ---- ComputeDistinctiveDescriptors: point p observes keyframes 0 .. N_p - 1 (vector order = map order) at row p
    int32_t nPoints;
    rd(fi, &nPoints, 1);
    std::vector<int32_t> start(nPoints + 1);
    rd(fi, start.data(), nPoints + 1);
    std::vector<uint8_t> rows((size_t)start[nPoints] * 32);
    rd(fi, rows.data(), rows.size());
    int maxObs = 0;
    for (int p = 0; p < nPoints; p++) maxObs = std::max(maxObs, start[p + 1] - start[p]);
    std::vector<DirectKeyFrameLite> kfs(maxObs);
    for (auto &k : kfs) k.mDescriptors.create(std::max(nPoints, 1), 32, cv::CV_8U);
    std::vector<DirectMapPointLite> mps(nPoints);
    std::vector<DirectMapPointLite *> mpp(nPoints);
    for (int p = 0; p < nPoints; p++) {
        mpp[p] = &mps[p];
        mps[p].mDescriptor.create(1, 32, cv::CV_8U);
        std::memset(mps[p].mDescriptor.ptr(0), 0xA5, 32);
        for (int k = 0; k < start[p + 1] - start[p]; k++) {
            std::memcpy(kfs[k].mDescriptors.ptr(p), &rows[(size_t)(start[p] + k) * 32], 32);
            mps[p].mObservations[&kfs[k]] = (size_t)p;
        }
    }
    ComputeDistinctiveDescriptors(handle, mpp);
    for (int p = 0; p < nPoints; p++) wr(fo, mps[p].mDescriptor.ptr(0), 32);
    printf("host_demo3: sparse align tracked %zu features in %d iterations, %d direct projections, %d distinctive descriptors\n",
           tracked, alignIterations, m, nPoints);
    ygzorb_pyramid_destroy(ref.mpDevPyramid); ygzorb_pyramid_destroy(cur.mpDevPyramid);
    ygzorb_matcher_destroy(handle);
    fclose(fi); fclose(fo);
    return 0;
}
