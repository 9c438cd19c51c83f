// host_demo2.cc — drives the ORBmatcher overloads of SURVEY 8(f) f2 that work on Frame / KeyFrame / MapPoint objects the
// way their callers do (Tracking::MonocularInitialization src/Tracking.cc:825-833, Tracking::Relocalization :1882-1886,
// LocalMapping::SearchInNeighbors src/LocalMapping.cc:1259-1270) and dumps the outcome for tests/test_gpu_host.py.
// usage: host_demo2 <in.bin> <out.bin>
//   in.bin : int32 n | n x 28 B keypoints | n x 32 B desc | float bounds[4] | int32 nlevels | float scale[nlevels] |
//            float invSigma2[nlevels] | float logScale |
//            [init]  int32 n2 | n2 keypoints | n2 desc | int32 window | float ratio |
//            [reloc] int32 m | float pose[30] | m x rec | n x uint8 claimed | float th | int32 ORBdist |
//            [fuse]  int32 m | float pose[30] | m x rec | n x float uRight | n x int32 existingObs (-1: no MapPoint) | float th
//            rec = float wpos[3], normal[3], angle, maxDistance, minDistance; int32 valid, nObs; 32 B descriptor
//   out.bin: [init]  int32 nmatches | n x int32 vnMatches12 | n x 2 float vbPrevMatched |
//            [reloc] int32 nmatches | n x int32 (index of the KeyFrame point now in mvpMapPoints[i]; -1 null; -2 held before) |
//            [fuse]  int32 nFused | n x int32 slot (id of the MapPoint in the keyframe slot: new point i -> i, original -> m + slot,
//                    -1 empty) | (m + n) x int32 replacedBy (id or -1) | (m + n) x int32 Observations()
#include <cstdio>
#include <cstdlib>
#include <set>
#include <vector>

#include "FrameLite.h"
#include "ORBmatcher.h"

using namespace ygz;

template <class T> static void rd(FILE *f, T *p, size_t n) { if (fread(p, sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); } }
template <class T> static void wr(FILE *f, const T *p, size_t n) { fwrite(p, sizeof(T), n, f); }

struct Rec { float wpos[3], normal[3], angle, maxD, minD; int32_t valid, nObs; uint8_t desc[32]; };

static void readPoints(FILE *fi, int m, std::vector<MapPointLite> &pts, std::vector<Rec> &recs) {
    pts.resize(m); recs.resize(m);
    rd(fi, recs.data(), m);
    for (int i = 0; i < m; i++) {
        MapPointLite &p = pts[i];
        for (int r = 0; r < 3; r++) { p.mWorldPos[r] = recs[i].wpos[r]; p.mNormal[r] = recs[i].normal[r]; }
        p.mfMaxDistance = recs[i].maxD; p.mfMinDistance = recs[i].minD; p.nObs = recs[i].nObs;
        p.mDescriptor.create(1, 32, cv::CV_8U);
        std::memcpy(p.mDescriptor.data, recs[i].desc, 32);
    }
}

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: host_demo2 in.bin out.bin\n"); return 2; }
    FILE *fi = fopen(argv[1], "rb"), *fo = fopen(argv[2], "wb");
    if (!fi || !fo) { perror("open"); return 2; }
    int32_t n;
    rd(fi, &n, 1);
    Frame F;
    F.N = n; F.mvKeys.resize(n);
    rd(fi, F.mvKeys.data(), n);
    F.mDescriptors.create(n, 32, cv::CV_8U);
    rd(fi, F.mDescriptors.data, (size_t)n * 32);
    float bounds[4];
    rd(fi, bounds, 4);
    F.mnMinX = bounds[0]; F.mnMinY = bounds[1]; F.mnMaxX = bounds[2]; F.mnMaxY = bounds[3];
    int32_t nlevels;
    rd(fi, &nlevels, 1);
    std::vector<float> invSigma2(nlevels);
    F.mvScaleFactors.resize(nlevels);
    rd(fi, F.mvScaleFactors.data(), nlevels); rd(fi, invSigma2.data(), nlevels);
    rd(fi, &F.mfLogScaleFactor, 1);
    F.mnScaleLevels = nlevels;
    F.mvuRight.assign(n, -1.0f);
    F.mvpMapPoints.assign(n, nullptr);

    // ---- Tracking::MonocularInitialization: mvbPrevMatched = the initial frame's keypoint positions (src/Tracking.cc:806-808)
    {
        int32_t n2, window; float ratio;
        rd(fi, &n2, 1);
        Frame F2;
        F2.N = n2; F2.mvKeys.resize(n2);
        rd(fi, F2.mvKeys.data(), n2);
        F2.mDescriptors.create(n2, 32, cv::CV_8U);
        rd(fi, F2.mDescriptors.data, (size_t)n2 * 32);
        F2.mnMinX = bounds[0]; F2.mnMinY = bounds[1]; F2.mnMaxX = bounds[2]; F2.mnMaxY = bounds[3];
        rd(fi, &window, 1); rd(fi, &ratio, 1);
        std::vector<cv::Point2f> vbPrevMatched(n);
        for (int i = 0; i < n; i++) vbPrevMatched[i] = F.mvKeys[i].pt;
        std::vector<int> vnMatches12;
        ORBmatcher matcher(ratio, true);
        int32_t nmatches = matcher.SearchForInitialization(F, F2, vbPrevMatched, vnMatches12, window);
        wr(fo, &nmatches, 1);
        for (int i = 0; i < n; i++) { int32_t v = vnMatches12[i]; wr(fo, &v, 1); }
        for (int i = 0; i < n; i++) { float xy[2] = {vbPrevMatched[i].x, vbPrevMatched[i].y}; wr(fo, xy, 2); }
        printf("host_demo2: SearchForInitialization %d matches\n", nmatches);
    }
    // ---- Tracking::Relocalization: SearchByProjection(CurrentFrame, pKF, sFound, th, ORBdist)
    {
        int32_t m, orbDist; float pose[30], th;
        rd(fi, &m, 1); rd(fi, pose, 30);
        std::vector<MapPointLite> pts; std::vector<Rec> recs;
        readPoints(fi, m, pts, recs);
        std::vector<uint8_t> claimed(n);
        rd(fi, claimed.data(), n);
        rd(fi, &th, 1); rd(fi, &orbDist, 1);
        for (int r = 0; r < 9; r++) F.mTcw.R[r] = pose[r];
        for (int r = 0; r < 3; r++) F.mTcw.t[r] = pose[9 + r];
        F.fx = pose[24]; F.fy = pose[25]; F.cx = pose[26]; F.cy = pose[27];
        KeyFrameLite kf;
        kf.mvKeys.resize(m); kf.mvpMapPoints.assign(m, nullptr);
        std::set<MapPointLite *> sFound;
        MapPointLite held;
        for (int i = 0; i < m; i++) {
            kf.mvKeys[i].angle = recs[i].angle;
            kf.mvpMapPoints[i] = &pts[i];
            if (!recs[i].valid) { if (i % 2) pts[i].mbBad = true; else sFound.insert(&pts[i]); }   // both ways of being skipped
        }
        for (int i = 0; i < n; i++) F.mvpMapPoints[i] = claimed[i] ? &held : nullptr;
        ORBmatcher matcher(0.9f, true);
        int32_t nmatches = matcher.SearchByProjection(F, &kf, sFound, th, orbDist);
        wr(fo, &nmatches, 1);
        for (int i = 0; i < n; i++) {
            MapPointLite *p = F.mvpMapPoints[i];
            int32_t v = !p ? -1 : (p == &held ? -2 : (int32_t)(p - pts.data()));
            wr(fo, &v, 1);
        }
        printf("host_demo2: relocalisation SearchByProjection %d matches\n", nmatches);
    }
    // ---- LocalMapping::SearchInNeighbors: Fuse(pKF, vpMapPoints)
    {
        int32_t m; float pose[30], th;
        rd(fi, &m, 1); rd(fi, pose, 30);
        std::vector<MapPointLite> pts; std::vector<Rec> recs;
        readPoints(fi, m, pts, recs);
        KeyFrameLite kf;
        kf.mvKeys = F.mvKeys; kf.mDescriptors = F.mDescriptors;
        kf.mvuRight.resize(n);
        rd(fi, kf.mvuRight.data(), n);
        std::vector<int32_t> existingObs(n);
        rd(fi, existingObs.data(), n);
        rd(fi, &th, 1);
        kf.mvScaleFactors = F.mvScaleFactors; kf.mvInvLevelSigma2 = invSigma2;
        kf.mfLogScaleFactor = F.mfLogScaleFactor; kf.mnScaleLevels = nlevels;
        kf.mnMinX = (int)bounds[0]; kf.mnMinY = (int)bounds[1]; kf.mnMaxX = (int)bounds[2]; kf.mnMaxY = (int)bounds[3];
        for (int r = 0; r < 9; r++) kf.mTcw.R[r] = pose[r];
        for (int r = 0; r < 3; r++) { kf.mTcw.t[r] = pose[9 + r]; kf.mOw[r] = pose[12 + r]; }
        kf.fx = pose[24]; kf.fy = pose[25]; kf.cx = pose[26]; kf.cy = pose[27]; kf.mbf = pose[28];
        std::vector<MapPointLite> original(n);
        kf.mvpMapPoints.assign(n, nullptr);
        for (int i = 0; i < n; i++)
            if (existingObs[i] >= 0) { original[i].nObs = existingObs[i]; original[i].mObservations[&kf] = i; kf.mvpMapPoints[i] = &original[i]; }
        std::vector<MapPointLite *> vp(m);
        for (int i = 0; i < m; i++) { vp[i] = &pts[i]; if (!recs[i].valid) pts[i].mbBad = true; }
        ORBmatcher matcher;
        int32_t nFused = matcher.Fuse(&kf, vp, th);
        auto idOf = [&](MapPointLite *p) -> int32_t {
            if (!p) return -1;
            if (p >= pts.data() && p < pts.data() + m) return (int32_t)(p - pts.data());
            return m + (int32_t)(p - original.data());
        };
        wr(fo, &nFused, 1);
        for (int i = 0; i < n; i++) { int32_t v = idOf(kf.mvpMapPoints[i]); wr(fo, &v, 1); }
        for (int i = 0; i < m; i++) { int32_t v = idOf(pts[i].mpReplaced); wr(fo, &v, 1); }
        for (int i = 0; i < n; i++) { int32_t v = idOf(original[i].mpReplaced); wr(fo, &v, 1); }
        for (int i = 0; i < m; i++) { int32_t v = pts[i].Observations(); wr(fo, &v, 1); }
        for (int i = 0; i < n; i++) { int32_t v = original[i].Observations(); wr(fo, &v, 1); }
        printf("host_demo2: Fuse %d\n", nFused);
    }
    fclose(fi); fclose(fo);
    return 0;
}
