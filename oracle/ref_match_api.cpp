// ref_match_api.cpp — TEST INFRASTRUCTURE ONLY.  C entry points (ygzr_*) over the REFERENCE'S OWN matcher code, compiled
// unmodified into oracle/_ref/libygz_ref_match.so together with this file (oracle/Makefile, target refmatch):
//   * /root/reference/src/ORBmatcher.cc (every search, DescriptorDistance, ComputeThreeMaxima, CheckDistEpipolarLine,
//     FindDirectProjection, GetWarpAffineMatrix, WarpAffine) and /root/reference/src/Align.cc (Align2D), whole files;
//   * the bodies of Frame::{AssignFeaturesToGrid, GetFeaturesInArea, PosInGrid, ComputeStereoMatches, World2Pixel, ...},
//     KeyFrame::{GetFeaturesInArea, IsInImage, pose getters, Pixel2Camera}, MapPoint::{PredictScale x2,
//     ComputeDistinctiveDescriptors, Get*DistanceInvariance, isBad, ...}, copied out of the reference at build time
//     (oracle/extract_ref_bodies.py -> oracle/_ref/gen/*.inc, git-ignored) and included below.
// Each entry point takes the same flat arrays as its restatement in oracle/match_oracle.cpp / direct_oracle.cpp (ygzo_*),
// builds the Frame / KeyFrame / MapPoint object graph the reference method expects (shim classes, ref_shim/match_prelude.h),
// calls the reference method, and flattens what the method left in its outputs.  tests/test_ref_match.py asserts ygzo_* ==
// ygzr_* on every case; tools/make_golden_match.py commits the ygzr_* outputs as tests/golden/match_*.npz.
#include "ORBmatcher.h"
#include "Align.h"
#include "SparseImageAlign.h"

#include <cstring>
#include <deque>
#include <memory>

namespace ygz {
std::vector<MapAction> g_mapActions;
#include "_ref/gen/frame_bodies.inc"
#include "_ref/gen/keyframe_bodies.inc"
#include "_ref/gen/mappoint_bodies.inc"
}  // namespace ygz

using namespace ygz;

#define YGZR_API extern "C" __attribute__((visibility("default")))

namespace {

struct KP { float x, y, size, angle, response; int octave, class_id; };
static_assert(sizeof(KP) == sizeof(cv::KeyPoint), "cv::KeyPoint layout");

std::vector<cv::KeyPoint> toKeys(const void *kps, int n) {
    std::vector<cv::KeyPoint> v((size_t)std::max(n, 0));
    if (n > 0) std::memcpy((void *)v.data(), kps, (size_t)n * sizeof(cv::KeyPoint));
    return v;
}
cv::Mat descMat(const uint8_t *d, int n) { return cv::Mat(n, 32, cv::CV_8U, (void *)d, 32); }
cv::Mat descRow(const uint8_t *d) { return cv::Mat(1, 32, cv::CV_8U, (void *)d, 32).clone(); }
int maxOctave(const std::vector<cv::KeyPoint> &k) { int m = 0; for (auto &p : k) m = std::max(m, p.octave); return m; }
std::vector<float> levels(const float *p, int n) { return p ? std::vector<float>(p, p + n) : std::vector<float>(); }

// a Frame as the constructor leaves it as far as the searches read it: keys, descriptors, grid (src/Frame.cc:296-297, 314-330)
void fillFrame(Frame &F, const void *kps, const uint8_t *desc, const float *uRight, int n, float minX, float minY, float maxX,
               float maxY, float invW, float invH, const float *scaleFactors, int nLevels) {
    F.N = n;
    F.mvKeys = toKeys(kps, n);
    F.mDescriptors = descMat(desc, n);
    F.mvuRight.assign((size_t)n, -1.0f);
    if (uRight) for (int i = 0; i < n; i++) F.mvuRight[i] = uRight[i];
    F.mvpMapPoints.assign((size_t)n, nullptr);
    F.mvbOutlier.assign((size_t)n, false);
    F.mnMinX = minX; F.mnMinY = minY; F.mnMaxX = maxX; F.mnMaxY = maxY;
    F.mfGridElementWidthInv = invW; F.mfGridElementHeightInv = invH;
    F.mvScaleFactors = levels(scaleFactors, nLevels);
    F.AssignFeaturesToGrid();                                   // reference body
}

// KeyFrame::KeyFrame copies the Frame's grid (src/KeyFrame.cc:230-236): same assignment rule, built through a Frame
void fillKeyFrame(KeyFrame &K, const void *kps, const uint8_t *desc, const float *uRight, int n, float minX, float minY, float maxX,
                  float maxY, float invW, float invH, const float *scaleFactors, int nLevels) {
    std::unique_ptr<Frame> F(new Frame());
    fillFrame(*F, kps, desc, uRight, n, minX, minY, maxX, maxY, invW, invH, scaleFactors, nLevels);
    K.N = n; K.mvKeys = F->mvKeys; K.mDescriptors = F->mDescriptors; K.mvuRight = F->mvuRight;
    K.mvpMapPoints.assign((size_t)n, nullptr);
    K.mnMinX = minX; K.mnMinY = minY; K.mnMaxX = maxX; K.mnMaxY = maxY;
    K.mfGridElementWidthInv = invW; K.mfGridElementHeightInv = invH;
    K.mvScaleFactors = F->mvScaleFactors;
    K.mGrid.resize(K.mnGridCols);
    for (int i = 0; i < K.mnGridCols; i++) {
        K.mGrid[i].resize(K.mnGridRows);
        for (int j = 0; j < K.mnGridRows; j++) K.mGrid[i][j] = F->mGrid[i][j];
    }
}

SE3f poseFromRt(const float *R, const float *t) {   // the rotation matrix is kept verbatim (eigen_lite.h)
    Matrix3f M;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) M(r, c) = R[3 * r + c];
    return SE3f(M, Vector3f(t[0], t[1], t[2]));
}
SE3f poseFromQt(const float *q) { return SE3f(Eigen::Quaternionf(q[3], q[0], q[1], q[2]), Vector3f(q[4], q[5], q[6])); }

void fillFeatVec(DBoW2::FeatureVector &fv, const uint32_t *nodeIds, const int32_t *start, const uint32_t *feats, int nNodes) {
    for (int k = 0; k < nNodes; k++)
        for (int i = start[k]; i < start[k + 1]; i++) fv.addFeature(nodeIds[k], feats[i]);   // insertion order = index order
}

}  // namespace

// ---- ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1507-1523
YGZR_API int ygzr_descriptor_distance(const uint8_t *a, const uint8_t *b) {
    return ORBmatcher::DescriptorDistance(cv::Mat(1, 32, cv::CV_8U, (void *)a, 32), cv::Mat(1, 32, cv::CV_8U, (void *)b, 32));
}

// ---- ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, checkLevel), src/ORBmatcher.cc:44-126.
// Arguments as ygzo_search_by_projection.  The reference leaves no per-MapPoint record: mpBestIdx is not produced.
YGZR_API int ygzr_search_by_projection(const void *kps, const uint8_t *desc, const float *uRight, const uint8_t *claimed, int n,
                                       float minX, float minY, float invW, float invH, const float *scaleFactors,
                                       const float *projX, const float *projY, const float *projXR, const float *viewCos,
                                       const int32_t *level, const uint8_t *inView, const uint8_t *hasObs, const uint8_t *mpDesc,
                                       int m, float th, int checkLevel, float nnratio, int32_t *kpToMp) {
    std::unique_ptr<Frame> F(new Frame());
    int nl = 1;
    for (int i = 0; i < m; i++) nl = std::max(nl, level[i] + 1);
    fillFrame(*F, kps, desc, uRight, n, minX, minY, 0, 0, invW, invH, scaleFactors, nl);
    std::vector<MapPoint> mps((size_t)std::max(m, 1));
    MapPoint occupied;                      // a MapPoint with observations already in the frame (:89-91)
    occupied.nObs = 1;
    if (claimed) for (int i = 0; i < n; i++) if (claimed[i]) F->mvpMapPoints[i] = &occupied;
    std::vector<MapPoint *> vp((size_t)m);
    for (int i = 0; i < m; i++) {
        MapPoint &p = mps[i];
        p.mbTrackInView = inView[i] != 0;
        p.mnTrackScaleLevel = level[i]; p.mTrackViewCos = viewCos[i];
        p.mTrackProjX = projX[i]; p.mTrackProjY = projY[i]; p.mTrackProjXR = projXR ? projXR[i] : 0.0f;
        p.mDescriptor = descRow(mpDesc + (size_t)i * 32);
        p.nObs = hasObs[i] ? 1 : 0;
        vp[i] = &p;
    }
    ORBmatcher matcher(nnratio);
    const int nm = matcher.SearchByProjection(*F, vp, th, checkLevel != 0);
    for (int i = 0; i < n; i++) {
        MapPoint *p = F->mvpMapPoints[i];
        kpToMp[i] = (p && p != &occupied) ? (int)(p - mps.data()) : -1;
    }
    return nm;
}

// ---- ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono, checkLevel), :1218-1350.
// Arguments as ygzo_search_by_projection_last (pose[30] = Rcw, tcw, Rlw, tlw, fx, fy, cx, cy, mbf, mb).
YGZR_API int ygzr_search_by_projection_last(const void *kps, const uint8_t *desc, const float *uRight, const uint8_t *claimed,
                                            int n, float minX, float minY, float maxX, float maxY, float invW, float invH,
                                            const float *scaleFactors, const float *pose, const uint8_t *valid,
                                            const uint8_t *hasObs, const float *worldPos, const uint8_t *ptDesc,
                                            const int32_t *octaveLast, const float *angleLast, int m, float th, int bMono,
                                            int checkLevel, int checkOrientation, int32_t *kpToPt) {
    std::unique_ptr<Frame> Cur(new Frame()), Last(new Frame());
    int nl = 1;
    for (int i = 0; i < m; i++) nl = std::max(nl, octaveLast[i] + 1);
    fillFrame(*Cur, kps, desc, uRight, n, minX, minY, maxX, maxY, invW, invH, scaleFactors, nl);
    Cur->mTcw = poseFromRt(pose, pose + 9);
    Last->mTcw = poseFromRt(pose + 12, pose + 21);
    Cur->fx = pose[24]; Cur->fy = pose[25]; Cur->cx = pose[26]; Cur->cy = pose[27]; Cur->mbf = pose[28]; Cur->mb = pose[29];
    MapPoint occupied;
    occupied.nObs = 1;
    if (claimed) for (int i = 0; i < n; i++) if (claimed[i]) Cur->mvpMapPoints[i] = &occupied;
    std::vector<MapPoint> mps((size_t)std::max(m, 1));
    Last->N = m;
    Last->mvKeys.resize((size_t)m);
    Last->mvpMapPoints.assign((size_t)m, nullptr);
    Last->mvbOutlier.assign((size_t)m, false);
    for (int i = 0; i < m; i++) {
        Last->mvKeys[i].octave = octaveLast[i]; Last->mvKeys[i].angle = angleLast[i];
        MapPoint &p = mps[i];
        p.mWorldPos = Vector3f(worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]);
        p.mDescriptor = descRow(ptDesc + (size_t)i * 32);
        p.nObs = hasObs[i] ? 1 : 0;
        if (valid[i]) Last->mvpMapPoints[i] = &p;
    }
    ORBmatcher matcher(0.9f, checkOrientation != 0);
    const int nm = matcher.SearchByProjection(*Cur, *Last, th, bMono != 0, checkLevel != 0);
    for (int i = 0; i < n; i++) {
        MapPoint *p = Cur->mvpMapPoints[i];
        kpToPt[i] = (p && p != &occupied) ? (int)(p - mps.data()) : -1;
    }
    return nm;
}

// Scw (4x4, row-major) -> what src/ORBmatcher.cc:275-278 (= :898-901) derives from it, with the same cv::Mat expressions:
// out[0..8] = Rcw, out[9..11] = tcw, out[12..14] = Ow.  The restatement and the kernels take these as inputs.
YGZR_API void ygzr_decompose_scw(const float *scw16, float *out15) {
    cv::Mat Scw(4, 4, cv::CV_32F, (void *)scw16, 16);
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) out15[3 * r + c] = Rcw.at<float>(r, c);
    for (int r = 0; r < 3; r++) { out15[9 + r] = tcw.at<float>(r); out15[12 + r] = Ow.at<float>(r); }
}

// ---- the searches over MapPoints that predict the level from the distance.  Arguments as ygzo_search_by_projection_keyframe
// plus scw16: the Sim3 matrix the reference method is given (variants 1, 3; the restatement receives ygzr_decompose_scw of it).
//   variant 0: SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist), :1352-1469
//   variant 1: SearchByProjection(KeyFrame *pKF, cv::Mat Scw, vpPoints, vpMatched, th), :265-373
//   variant 2: Fuse(KeyFrame *pKF, vpMapPoints, th), :748-886            (map actions recorded)
//   variant 3: Fuse(KeyFrame *pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint), :888-1004
// kpToPt[idx] = point left at the keypoint (variants 0, 1) / last point that asked for it (2, 3); ptBestIdx[i] = keypoint of
// point i where the reference exposes it (2, 3: AddObservation(pKF, bestIdx)), else untouched.
YGZR_API int ygzr_search_by_projection_keyframe(const void *kps, const uint8_t *desc, const uint8_t *claimed, const float *uRight,
                                                int n, float minX, float minY, float maxX, float maxY, float invW, float invH,
                                                const float *scaleFactors, const float *invSigma2, const float *pose,
                                                const uint8_t *valid, const float *worldPos, const float *normal,
                                                const uint8_t *ptDesc, const float *angleKF, const float *maxDistance,
                                                const float *minDistance, int m, float logScaleFactor, int nScaleLevels, float th,
                                                int bestDistMax, int checkOrientation, int variant, const float *scw16,
                                                int32_t *kpToPt, int32_t *ptBestIdx) {
    std::vector<MapPoint> mps((size_t)std::max(m, 1));
    for (int i = 0; i < m; i++) {
        MapPoint &p = mps[i];
        p.mWorldPos = Vector3f(worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]);
        if (normal) p.mNormalVector = Vector3f(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
        p.mDescriptor = descRow(ptDesc + (size_t)i * 32);
        p.mfMaxDistance = maxDistance[i]; p.mfMinDistance = minDistance[i];
        p.nObs = 1;
    }
    for (int i = 0; i < n; i++) kpToPt[i] = -1;
    MapPoint occupied;
    occupied.nObs = 1;
    g_mapActions.clear();
    int nm = 0;
    if (variant == 0) {
        std::unique_ptr<Frame> Cur(new Frame());
        fillFrame(*Cur, kps, desc, nullptr, n, minX, minY, maxX, maxY, invW, invH, scaleFactors, nScaleLevels);
        Cur->mTcw = poseFromRt(pose, pose + 9);
        Cur->fx = pose[24]; Cur->fy = pose[25]; Cur->cx = pose[26]; Cur->cy = pose[27];
        Cur->mfLogScaleFactor = logScaleFactor; Cur->mnScaleLevels = nScaleLevels;
        if (claimed) for (int i = 0; i < n; i++) if (claimed[i]) Cur->mvpMapPoints[i] = &occupied;
        KeyFrame KF;                                            // the candidate keyframe: one keypoint per point
        KF.mvKeys.resize((size_t)m);
        KF.mvpMapPoints.assign((size_t)m, nullptr);
        for (int i = 0; i < m; i++) { KF.mvKeys[i].angle = angleKF[i]; if (valid[i]) KF.mvpMapPoints[i] = &mps[i]; }
        std::set<MapPoint *> found;
        ORBmatcher matcher(0.9f, checkOrientation != 0);
        nm = matcher.SearchByProjection(*Cur, &KF, found, th, bestDistMax);
        for (int i = 0; i < n; i++) {
            MapPoint *p = Cur->mvpMapPoints[i];
            kpToPt[i] = (p && p != &occupied) ? (int)(p - mps.data()) : -1;
        }
        return nm;
    }
    KeyFrame KF;
    fillKeyFrame(KF, kps, desc, uRight, n, minX, minY, maxX, maxY, invW, invH, scaleFactors, nScaleLevels);
    KF.fx = pose[24]; KF.fy = pose[25]; KF.cx = pose[26]; KF.cy = pose[27]; KF.mbf = pose[28];
    KF.mfLogScaleFactor = logScaleFactor; KF.mnScaleLevels = nScaleLevels;
    if (invSigma2) KF.mvInvLevelSigma2 = levels(invSigma2, nScaleLevels);
    cv::Mat Scw;
    if (scw16) Scw = cv::Mat(4, 4, cv::CV_32F, (void *)scw16, 16).clone();
    ORBmatcher matcher(0.9f, checkOrientation != 0);
    if (variant == 1) {
        std::vector<MapPoint *> vpPoints((size_t)m), vpMatched((size_t)n, nullptr);
        for (int i = 0; i < m; i++) { vpPoints[i] = &mps[i]; mps[i].mbBad = !valid[i]; }   // valid = !isBad && !alreadyFound
        if (claimed) for (int i = 0; i < n; i++) if (claimed[i]) vpMatched[i] = &occupied;
        nm = matcher.SearchByProjection(&KF, Scw, vpPoints, vpMatched, (int)th);
        for (int i = 0; i < n; i++) kpToPt[i] = (vpMatched[i] && vpMatched[i] != &occupied) ? (int)(vpMatched[i] - mps.data()) : -1;
        return nm;
    }
    if (variant == 2) {
        KF.Tcw = poseFromRt(pose, pose + 9);
        KF.Ow = Vector3f(pose[12], pose[13], pose[14]);
        std::vector<MapPoint *> vp((size_t)m, nullptr);
        for (int i = 0; i < m; i++) if (valid[i]) vp[i] = &mps[i];                          // valid = non-null, !isBad, !IsInKeyFrame
        nm = matcher.Fuse(&KF, vp, th);
    } else {
        std::vector<MapPoint *> vp((size_t)m), vpReplace((size_t)m, nullptr);
        for (int i = 0; i < m; i++) { vp[i] = &mps[i]; mps[i].mbBad = !valid[i]; }
        nm = matcher.Fuse(&KF, Scw, vp, th, vpReplace);
    }
    for (const MapAction &a : g_mapActions)
        if (a.kind == 0) {
            const int i = (int)(a.self - mps.data());
            if (ptBestIdx) ptBestIdx[i] = (int)a.idx;
            kpToPt[a.idx] = i;
        }
    return nm;
}

// ---- ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th), :1006-1216, whole (both directions + agreement).
// Keyframe k (k = 1, 2): keys / descriptors / grid as above, one MapPoint per keypoint (valid, world position, descriptor,
// max / min distance), pose Rkw, tkw (12 floats).  already12[i1] = index in KF2 of a match vpMatches12 holds on entry, or -1.
// derived[21] (out) = sR12, sR21 (9 + 9... see below) as :1022-1025 computes them, for the restatement:
//   derived[0..8] = sR12, [9..17] = sR21, [18..20] = t21.
YGZR_API int ygzr_search_by_sim3(const void *kps1, const uint8_t *desc1, int n1, const void *kps2, const uint8_t *desc2, int n2,
                                 float minX, float minY, float maxX, float maxY, float invW, float invH, const float *scaleFactors,
                                 int nScaleLevels, float logScaleFactor, const float *intr, const float *pose1, const float *pose2,
                                 const uint8_t *valid1, const float *world1, const uint8_t *mpDesc1, const float *maxD1,
                                 const float *minD1, const uint8_t *valid2, const float *world2, const uint8_t *mpDesc2,
                                 const float *maxD2, const float *minD2, float s12, const float *R12f, const float *t12f, float th,
                                 const int32_t *already12, int32_t *matches12, float *derived) {
    KeyFrame K1, K2;
    fillKeyFrame(K1, kps1, desc1, nullptr, n1, minX, minY, maxX, maxY, invW, invH, scaleFactors, nScaleLevels);
    fillKeyFrame(K2, kps2, desc2, nullptr, n2, minX, minY, maxX, maxY, invW, invH, scaleFactors, nScaleLevels);
    for (KeyFrame *K : {&K1, &K2}) {
        K->fx = intr[0]; K->fy = intr[1]; K->cx = intr[2]; K->cy = intr[3];
        K->mfLogScaleFactor = logScaleFactor; K->mnScaleLevels = nScaleLevels;
    }
    K1.Tcw = poseFromRt(pose1, pose1 + 9);
    K2.Tcw = poseFromRt(pose2, pose2 + 9);
    std::vector<MapPoint> m1((size_t)std::max(n1, 1)), m2((size_t)std::max(n2, 1));
    auto fill = [](std::vector<MapPoint> &mp, KeyFrame &K, int n, const uint8_t *valid, const float *world, const uint8_t *d,
                   const float *maxD, const float *minD) {
        for (int i = 0; i < n; i++) {
            MapPoint &p = mp[i];
            p.mWorldPos = Vector3f(world[3 * i], world[3 * i + 1], world[3 * i + 2]);
            p.mDescriptor = descRow(d + (size_t)i * 32);
            p.mfMaxDistance = maxD[i]; p.mfMinDistance = minD[i];
            p.nObs = 1;
            p.mObservations[&K] = (size_t)i;
            if (valid[i]) K.mvpMapPoints[i] = &p;
        }
    };
    fill(m1, K1, n1, valid1, world1, mpDesc1, maxD1, minD1);
    fill(m2, K2, n2, valid2, world2, mpDesc2, maxD2, minD2);
    std::vector<MapPoint *> vpMatches12((size_t)n1, nullptr);
    if (already12) for (int i = 0; i < n1; i++) if (already12[i] >= 0) vpMatches12[i] = &m2[already12[i]];
    cv::Mat R12(3, 3, cv::CV_32F, (void *)R12f, 12), t12(3, 1, cv::CV_32F, (void *)t12f, 4);
    if (derived) {   // :1022-1025, the same three expressions
        cv::Mat sR12 = s12 * R12;
        cv::Mat sR21 = (1.0 / s12) * R12.t();
        cv::Mat t21 = -sR21 * t12;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { derived[3 * r + c] = sR12.at<float>(r, c); derived[9 + 3 * r + c] = sR21.at<float>(r, c); }
        for (int r = 0; r < 3; r++) derived[18 + r] = t21.at<float>(r);
    }
    ORBmatcher matcher(0.75f, true);
    const int nFound = matcher.SearchBySim3(&K1, &K2, vpMatches12, s12, R12.clone(), t12.clone(), th);
    for (int i = 0; i < n1; i++) matches12[i] = vpMatches12[i] ? (int)(vpMatches12[i] - m2.data()) : -1;
    return nFound;
}

// ---- ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), :375-478.  As ygzo_*.
YGZR_API int ygzr_search_for_initialization(const void *kps1, const uint8_t *desc1, int n1, const void *kps2, const uint8_t *desc2,
                                            int n2, float minX, float minY, float invW, float invH, float *prevMatched,
                                            int windowSize, float nnratio, int checkOrientation, int32_t *vnMatches12) {
    std::unique_ptr<Frame> F1(new Frame()), F2(new Frame());
    F1->N = n1; F1->mvKeys = toKeys(kps1, n1); F1->mDescriptors = descMat(desc1, n1);
    fillFrame(*F2, kps2, desc2, nullptr, n2, minX, minY, 0, 0, invW, invH, nullptr, 0);
    std::vector<cv::Point2f> prev((size_t)n1);
    for (int i = 0; i < n1; i++) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, checkOrientation != 0);
    const int nm = matcher.SearchForInitialization(*F1, *F2, prev, m12, windowSize);
    for (int i = 0; i < n1; i++) { vnMatches12[i] = m12[i]; prevMatched[2 * i] = prev[i].x; prevMatched[2 * i + 1] = prev[i].y; }
    return nm;
}

// ---- ORBmatcher::SearchByBoW, both overloads (:155-263, :480-596).  As ygzo_search_by_bow.
YGZR_API int ygzr_search_by_bow(const void *kps1, const uint8_t *desc1, const uint8_t *hasMP1, int n1, const uint32_t *nodeIds1,
                                const int32_t *start1, const uint32_t *feats1, int nNodes1, const void *kps2, const uint8_t *desc2,
                                const uint8_t *hasMP2, int n2, const uint32_t *nodeIds2, const int32_t *start2,
                                const uint32_t *feats2, int nNodes2, int secondIsFrame, float nnratio, int checkOrientation,
                                int32_t *matches) {
    KeyFrame K1;
    K1.N = n1; K1.mvKeys = toKeys(kps1, n1); K1.mDescriptors = descMat(desc1, n1);
    fillFeatVec(K1.mFeatVec, nodeIds1, start1, feats1, nNodes1);
    std::vector<MapPoint> m1((size_t)std::max(n1, 1)), m2((size_t)std::max(n2, 1));
    K1.mvpMapPoints.assign((size_t)n1, nullptr);
    for (int i = 0; i < n1; i++) if (hasMP1[i]) K1.mvpMapPoints[i] = &m1[i];
    ORBmatcher matcher(nnratio, checkOrientation != 0);
    std::vector<MapPoint *> out;
    int nm;
    if (secondIsFrame) {
        std::unique_ptr<Frame> F(new Frame());
        F->N = n2; F->mvKeys = toKeys(kps2, n2); F->mDescriptors = descMat(desc2, n2);
        fillFeatVec(F->mFeatVec, nodeIds2, start2, feats2, nNodes2);
        nm = matcher.SearchByBoW(&K1, *F, out);
        for (int i = 0; i < n2; i++) matches[i] = out[i] ? (int)(out[i] - m1.data()) : -1;
    } else {
        KeyFrame K2;
        K2.N = n2; K2.mvKeys = toKeys(kps2, n2); K2.mDescriptors = descMat(desc2, n2);
        fillFeatVec(K2.mFeatVec, nodeIds2, start2, feats2, nNodes2);
        K2.mvpMapPoints.assign((size_t)n2, nullptr);
        for (int i = 0; i < n2; i++) if (hasMP2[i]) K2.mvpMapPoints[i] = &m2[i];
        nm = matcher.SearchByBoW(&K1, &K2, out);
        for (int i = 0; i < n1; i++) matches[i] = out[i] ? (int)(out[i] - m2.data()) : -1;
    }
    return nm;
}

// ---- ORBmatcher::SearchForTriangulation (:597-746).  As ygzo_search_for_triangulation.  The reference computes the epipole
// itself (:604-610) from KF1's centre and KF2's pose and intrinsics; they are chosen so that it comes out as (ex, ey)
// exactly: Cw = (ex, ey, 1), R2w = I, t2w = 0, fx = fy = 1, cx = cy = 0.
YGZR_API int ygzr_search_for_triangulation(const void *kps1, const uint8_t *desc1, const float *uRight1, const uint8_t *hasMP1, int n1,
                                           const uint32_t *nodeIds1, const int32_t *start1, const uint32_t *feats1, int nNodes1,
                                           const void *kps2, const uint8_t *desc2, const float *uRight2, const uint8_t *hasMP2, int n2,
                                           const uint32_t *nodeIds2, const int32_t *start2, const uint32_t *feats2, int nNodes2,
                                           const float *scaleFactors2, const float *levelSigma2_2, const float *F12f, float ex,
                                           float ey, int bOnlyStereo, int checkOrientation, int32_t *matches12) {
    KeyFrame K1, K2;
    MapPoint some;
    auto fill = [&](KeyFrame &K, const void *kps, const uint8_t *desc, const float *uRight, const uint8_t *hasMP, int n,
                    const uint32_t *nodeIds, const int32_t *start, const uint32_t *feats, int nNodes) {
        K.N = n; K.mvKeys = toKeys(kps, n); K.mDescriptors = descMat(desc, n);
        K.mvuRight.assign(uRight, uRight + n);
        K.mvpMapPoints.assign((size_t)n, nullptr);
        for (int i = 0; i < n; i++) if (hasMP[i]) K.mvpMapPoints[i] = &some;
        fillFeatVec(K.mFeatVec, nodeIds, start, feats, nNodes);
    };
    fill(K1, kps1, desc1, uRight1, hasMP1, n1, nodeIds1, start1, feats1, nNodes1);
    fill(K2, kps2, desc2, uRight2, hasMP2, n2, nodeIds2, start2, feats2, nNodes2);
    const int nl = std::max(maxOctave(K2.mvKeys) + 1, 1);
    K2.mvScaleFactors = levels(scaleFactors2, nl); K2.mvLevelSigma2 = levels(levelSigma2_2, nl);
    K1.Ow = Vector3f(ex, ey, 1.0f);
    K2.fx = K2.fy = 1.0f; K2.cx = K2.cy = 0.0f;               // K2.Tcw = identity
    Matrix3f F12;
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) F12(r, c) = F12f[3 * r + c];
    std::vector<std::pair<size_t, size_t> > pairs;
    ORBmatcher matcher(0.6f, checkOrientation != 0);
    const int nm = matcher.SearchForTriangulation(&K1, &K2, F12, pairs, bOnlyStereo != 0);
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    for (auto &p : pairs) matches12[p.first] = (int)p.second;
    return nm;
}

// ---- Frame::ComputeStereoMatches (src/Frame.cc:509-682).  As ygzo_stereo_match (no bestDistOut: the reference keeps no
// per-keypoint SAD).  A frame without any match at all reads vDistIdx[0] of an empty vector in the reference (:671); the
// vector has reserved storage, so the read is harmless here and the invalidation loop does not run.
YGZR_API void ygzr_stereo_match(const void *kpsL, const uint8_t *descL, int nL, const void *kpsR, const uint8_t *descR, int nR,
                                const uint8_t *pyrL, const uint8_t *pyrR, const int64_t *off, const int32_t *lw, const int32_t *lh,
                                int nLevels, const float *scaleFactors, const float *invScaleFactors, float mbf, float mb,
                                float *uRightOut, float *depthOut) {
    std::unique_ptr<Frame> F(new Frame());
    ORBextractor exL, exR;
    for (int l = 0; l < nLevels; l++) {
        exL.mvImagePyramid.push_back(cv::Mat(lh[l], lw[l], cv::CV_8U, (void *)(pyrL + off[l]), (size_t)lw[l]));
        exR.mvImagePyramid.push_back(cv::Mat(lh[l], lw[l], cv::CV_8U, (void *)(pyrR + off[l]), (size_t)lw[l]));
    }
    F->mpORBextractorLeft = &exL; F->mpORBextractorRight = &exR;
    F->N = nL; F->mvKeys = toKeys(kpsL, nL); F->mvKeysRight = toKeys(kpsR, nR);
    F->mDescriptors = descMat(descL, nL); F->mDescriptorsRight = descMat(descR, nR);
    F->mvScaleFactors = levels(scaleFactors, nLevels); F->mvInvScaleFactors = levels(invScaleFactors, nLevels);
    F->mbf = mbf; F->mb = mb;
    F->ComputeStereoMatches();
    for (int i = 0; i < nL; i++) { uRightOut[i] = F->mvuRight[i]; depthOut[i] = F->mvDepth[i]; }
}

// ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:211-271).  As ygzo_distinctive_descriptors (no median out:
// the reference keeps only the descriptor).  The observations of a point live in a std::map keyed by KeyFrame*: the
// keyframes are allocated in one array, so pointer order = observation order of the flat input.
YGZR_API void ygzr_distinctive_descriptors(const uint8_t *obsDesc, const int32_t *obsStart, int nPoints, uint8_t *bestDesc,
                                           uint8_t *changed) {
    for (int p = 0; p < nPoints; p++) {
        const int N = obsStart[p + 1] - obsStart[p];
        std::vector<KeyFrame> kfs((size_t)std::max(N, 1));
        MapPoint mp;
        mp.mDescriptor = cv::Mat(1, 32, cv::CV_8U);
        std::memset(mp.mDescriptor.data, 0xAB, 32);
        for (int i = 0; i < N; i++) {
            kfs[i].mDescriptors = descMat(obsDesc + (size_t)(obsStart[p] + i) * 32, 1);
            mp.mObservations[&kfs[i]] = 0;
        }
        const uchar *before = mp.mDescriptor.data;
        mp.ComputeDistinctiveDescriptors();
        changed[p] = mp.mDescriptor.data != before;
        if (changed[p]) std::memcpy(bestDesc + (size_t)p * 32, mp.mDescriptor.data, 32);
    }
}

// ---- ORBmatcher::FindDirectProjection (:1573-1602) = GetWarpAffineMatrix + GetBestSearchLevel + WarpAffine + ygz::Align2D
// (src/Align.cc:8-105), point by point as Tracking::SearchLocalPointsDirect calls it.  Layout of YgzrDirectKeyFrame =
// YgzoDirectKeyFrame (oracle/direct_oracle.cpp) with tCurRef as an OUTPUT: the reference composes TCR = curr->mTcw *
// pose_ref.inverse() itself (:1582); what it used is handed back so that the restatement can be given the same TCR.
struct YgzrDirectKeyFrame {
    const uint8_t *levels[8];
    int32_t w[8], h[8], pitch[8];
    float poseRef[7];      // ref->GetPose(): unit quaternion x, y, z, w + translation
    float tCurRef[7];      // out
    float fx, fy, cx, cy;
};
YGZR_API void ygzr_find_direct_projection(YgzrDirectKeyFrame *kfs, int nKfs, const uint8_t *const *curLevels, const int32_t *curW,
                                          const int32_t *curH, const int32_t *curPitch, int nLevels, const float *curIntr,
                                          const float *curPose, const float *scale, const float *invScale, const float *invLevelSigma2,
                                          int n, const int32_t *kfIndex, const float *pxRef, const int32_t *octave,
                                          const float *worldPos, float *pxCurr, int32_t *searchLevel, uint8_t *success) {
    std::unique_ptr<Frame> Cur(new Frame());
    for (int l = 0; l < nLevels; l++)
        Cur->mvImagePyramid.push_back(cv::Mat(curH[l], curW[l], cv::CV_8U, (void *)curLevels[l], (size_t)curPitch[l]));
    Cur->fx = curIntr[0]; Cur->fy = curIntr[1]; Cur->cx = curIntr[2]; Cur->cy = curIntr[3];
    Cur->mTcw = poseFromQt(curPose);
    Cur->mvScaleFactors = levels(scale, nLevels); Cur->mvInvScaleFactors = levels(invScale, nLevels);
    std::vector<KeyFrame> K((size_t)std::max(nKfs, 1));
    for (int k = 0; k < nKfs; k++) {
        for (int l = 0; l < nLevels; l++)
            K[k].mvImagePyramid.push_back(cv::Mat(kfs[k].h[l], kfs[k].w[l], cv::CV_8U, (void *)kfs[k].levels[l], (size_t)kfs[k].pitch[l]));
        K[k].fx = kfs[k].fx; K[k].fy = kfs[k].fy; K[k].cx = kfs[k].cx; K[k].cy = kfs[k].cy;
        K[k].Tcw = poseFromQt(kfs[k].poseRef);
        K[k].mvScaleFactors = levels(scale, nLevels); K[k].mvInvLevelSigma2 = levels(invLevelSigma2, nLevels);
        const SE3f TCR = Cur->mTcw * K[k].GetPose().inverse();                              // :1581-1582
        kfs[k].tCurRef[0] = TCR.q.qx; kfs[k].tCurRef[1] = TCR.q.qy; kfs[k].tCurRef[2] = TCR.q.qz; kfs[k].tCurRef[3] = TCR.q.qw;
        for (int r = 0; r < 3; r++) kfs[k].tCurRef[4 + r] = TCR.t[r];
    }
    ORBmatcher matcher;
    for (int i = 0; i < n; i++) {
        KeyFrame &ref = K[kfIndex[i]];
        MapPoint mp;
        mp.mWorldPos = Vector3f(worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]);
        ref.mvKeys.assign(1, cv::KeyPoint());
        ref.mvKeys[0].pt = cv::Point2f(pxRef[2 * i], pxRef[2 * i + 1]); ref.mvKeys[0].octave = octave[i];
        mp.mObservations[&ref] = 0;
        Vector2f px(pxCurr[2 * i], pxCurr[2 * i + 1]);
        int level = 0;
        success[i] = matcher.FindDirectProjection(&ref, Cur.get(), &mp, px, level) ? 1 : 0;
        pxCurr[2 * i] = px[0]; pxCurr[2 * i + 1] = px[1];
        searchLevel[i] = level;
    }
}

// ---- SparseImgAlign (src/SparseImageAlign.cc, whole file) driven by NLLSSolver<6, SE3f>::optimizeGaussNewton
// (include/NLSSolver_impl.hpp:17-92), constructed as Tracking does (src/Tracking.cc:284: SparseImgAlign(nLevels - 1, 1, 30,
// GaussNewton, false, false)).  Layout of the pyramids = YgzoPyramid (oracle/direct_oracle.cpp); levels must be continuous
// (the reference takes stride = img.cols, src/SparseImageAlign.cc:60, 147).
struct YgzrPyramid { const uint8_t *levels[8]; int32_t w[8], h[8], pitch[8]; };
namespace {
struct SparseProbe : public SparseImgAlign {   // a derived class may read the protected state and call the protected steps
    SparseProbe(int maxLevel, int minLevel) : SparseImgAlign(maxLevel, minLevel, 30, GaussNewton, false, false) {}
    void prepare(Frame *ref, Frame *cur, int level) {   // what run() does before optimize() (:22-45), for ONE level
        reset();
        ref_frame_ = ref; cur_frame_ = cur;
        ref_patch_cache_ = cv::Mat(ref->N, 16, cv::CV_32F);
        std::memset(ref_patch_cache_.data, 0, (size_t)ref->N * 64);   // the probe reports the cache: start it defined
        jacobian_cache_.resize(Eigen::NoChange, ref_patch_cache_.rows * 16);
        visible_fts_ = vector<bool>(ref_patch_cache_.rows, false);
        level_ = level;
        jacobian_cache_.setZero();
        have_ref_patch_cache_ = false;
    }
    float residuals(const SE3f &T) { H_.setZero(); Jres_.setZero(); n_meas_ = 0; return computeResiduals(T, true, false); }
    const Matrix<float, 6, 6> &H() const { return H_; }
    const Matrix<float, 6, 1> &Jres() const { return Jres_; }
    size_t nMeas() const { return n_meas_; }
    bool visible(int i) const { return visible_fts_[i]; }
    const float *patch() const { return reinterpret_cast<const float *>(ref_patch_cache_.data); }
    const float *jac() { return &jacobian_cache_.d[0]; }
    float chi2() const { return chi2_; }
};
void fillDirectFrames(Frame &Ref, Frame &Cur, std::deque<MapPoint> &mps, const YgzrPyramid *ref, const YgzrPyramid *cur, int nLevels,
                      const float *invScale, float focal, const float *curIntr, const float *refPose, const float *curPose, int n,
                      const float *kpXY, const uint8_t *valid, const float *worldPos) {
    for (int l = 0; l < nLevels; l++) {
        Ref.mvImagePyramid.push_back(cv::Mat(ref->h[l], ref->w[l], cv::CV_8U, (void *)ref->levels[l], (size_t)ref->pitch[l]));
        Cur.mvImagePyramid.push_back(cv::Mat(cur->h[l], cur->w[l], cv::CV_8U, (void *)cur->levels[l], (size_t)cur->pitch[l]));
    }
    Ref.mvInvScaleFactors = levels(invScale, nLevels); Cur.mvInvScaleFactors = Ref.mvInvScaleFactors;
    Ref.fx = focal; Ref.fy = focal;
    Cur.fx = curIntr[0]; Cur.fy = curIntr[1]; Cur.cx = curIntr[2]; Cur.cy = curIntr[3];
    Ref.mTcw = poseFromQt(refPose);
    if (curPose) Cur.mTcw = poseFromQt(curPose);
    Ref.N = n;
    Ref.mvKeys.assign((size_t)n, cv::KeyPoint());
    Ref.mvpMapPoints.assign((size_t)n, nullptr);
    Ref.mvbOutlier.assign((size_t)n, false);
    mps.resize((size_t)n);
    for (int i = 0; i < n; i++) {
        Ref.mvKeys[i].pt = cv::Point2f(kpXY[2 * i], kpXY[2 * i + 1]);
        mps[i].mWorldPos = Vector3f(worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]);
        if (valid[i]) Ref.mvpMapPoints[i] = &mps[i];      // valid = MapPoint present, not bad, not an outlier (:72-74)
    }
}
void storeQt(const SE3f &T, float *o) { o[0] = T.q.qx; o[1] = T.q.qy; o[2] = T.q.qz; o[3] = T.q.qw; for (int r = 0; r < 3; r++) o[4 + r] = T.t[r]; }
}  // namespace

// One precomputeReferencePatches + computeResiduals(linearize) at `level` with the given T_cur_from_ref.
YGZR_API void ygzr_sparse_align_residuals(const YgzrPyramid *ref, const YgzrPyramid *cur, int nLevels, int level, const float *invScale,
                                          float focal, const float *curIntr, const float *refPose, const float *tCurRef, int n,
                                          const float *kpXY, const uint8_t *valid, const float *worldPos, float *H36, float *Jres6,
                                          float *chi2mean, int64_t *nMeas, uint8_t *visible, float *refPatch, float *jac) {
    std::unique_ptr<Frame> Ref(new Frame()), Cur(new Frame());
    std::deque<MapPoint> mps;   // MapPoint holds mutexes: not movable
    fillDirectFrames(*Ref, *Cur, mps, ref, cur, nLevels, invScale, focal, curIntr, refPose, nullptr, n, kpXY, valid, worldPos);
    SparseProbe P(nLevels - 1, 1);
    P.prepare(Ref.get(), Cur.get(), level);
    *chi2mean = P.residuals(poseFromQt(tCurRef));
    for (int r = 0; r < 6; r++) { Jres6[r] = P.Jres()(r); for (int c = 0; c < 6; c++) H36[6 * r + c] = P.H()(r, c); }
    *nMeas = (int64_t)P.nMeas();
    for (int i = 0; i < n; i++) visible[i] = P.visible(i) ? 1 : 0;
    std::memcpy(refPatch, P.patch(), (size_t)n * 64);
    std::memcpy(jac, P.jac(), (size_t)n * 384);
}

// SparseImgAlign::run(ref, cur, TCR).  tCurRef0 (out) = cur->mTcw * ref->mTcw.inverse() as the reference composed it (:35), the
// starting point the restatement is given; tCurRef (out) = TCR; stats (out): [0] final chi2_, [1] n_meas_.
YGZR_API int64_t ygzr_sparse_align_run(const YgzrPyramid *ref, const YgzrPyramid *cur, int nLevels, int maxLevel, int minLevel,
                                       const float *invScale, float focal, const float *curIntr, const float *refPose,
                                       const float *curPose, int n, const float *kpXY, const uint8_t *valid, const float *worldPos,
                                       float *tCurRef0, float *tCurRef, float *stats) {
    std::unique_ptr<Frame> Ref(new Frame()), Cur(new Frame());
    std::deque<MapPoint> mps;   // MapPoint holds mutexes: not movable
    fillDirectFrames(*Ref, *Cur, mps, ref, cur, nLevels, invScale, focal, curIntr, refPose, curPose, n, kpXY, valid, worldPos);
    storeQt(Cur->mTcw * Ref->mTcw.inverse(), tCurRef0);
    SparseProbe P(maxLevel, minLevel);
    SE3f TCR;
    const size_t ret = P.run(Ref.get(), Cur.get(), TCR);
    storeQt(TCR, tCurRef);
    if (stats) { stats[0] = P.chi2(); stats[1] = (float)P.nMeas(); }
    return (int64_t)ret;
}

// ---- bench.py's CPU leg: ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, checkLevel) of the reference on a batch
// of frames at fixed strides (the layout of ygzo_search_by_projection_batch / ygzorb_search_by_projection), `threads` host
// threads over frames, `repeat` passes.  The object graph (Frames, MapPoints) is built before the clock starts; the timed
// region is, per frame, what the reference's tracker pays for this step: Frame::AssignFeaturesToGrid (src/Frame.cc:314-330,
// part of the Frame constructor; the kernel builds the same grid) + the search.  Returns the seconds of the timed region.
#include <atomic>
#include <chrono>
#include <thread>
YGZR_API double ygzr_search_by_projection_batch(int nFrames, int kpStride, int mpStride, const int32_t *nKp, const int32_t *nMp,
                                                const void *kps, const uint8_t *desc, float minX, float minY, float invW,
                                                float invH, const float *scaleFactors, int nLevels, const float *projX,
                                                const float *projY, const float *projXR, const float *viewCos,
                                                const int32_t *level, const uint8_t *inView, const uint8_t *hasObs,
                                                const uint8_t *mpDesc, float th, int checkLevel, float nnratio, int threads,
                                                int repeat, int32_t *kpToMp, int32_t *nMatches) {
    struct Item { std::unique_ptr<Frame> F; std::vector<MapPoint> mps; std::vector<MapPoint *> vp; };
    std::vector<Item> items((size_t)nFrames);
    for (int f = 0; f < nFrames; f++) {
        Item &it = items[f];
        const int n = nKp[f], m = nMp[f];
        it.F.reset(new Frame());
        Frame &F = *it.F;
        F.N = n;
        F.mvKeys = toKeys((const uint8_t *)kps + (size_t)f * kpStride * sizeof(cv::KeyPoint), n);
        F.mDescriptors = descMat(desc + (size_t)f * kpStride * 32, n);
        F.mvuRight.assign((size_t)n, -1.0f);
        F.mnMinX = minX; F.mnMinY = minY; F.mfGridElementWidthInv = invW; F.mfGridElementHeightInv = invH;
        F.mvScaleFactors = levels(scaleFactors, nLevels);
        it.mps = std::vector<MapPoint>((size_t)std::max(m, 1));
        it.vp.resize((size_t)m);
        for (int i = 0; i < m; i++) {
            const size_t o = (size_t)f * mpStride + i;
            MapPoint &p = it.mps[i];
            p.mbTrackInView = inView[o] != 0; p.mnTrackScaleLevel = level[o]; p.mTrackViewCos = viewCos[o];
            p.mTrackProjX = projX[o]; p.mTrackProjY = projY[o]; p.mTrackProjXR = projXR ? projXR[o] : 0.0f;
            p.mDescriptor = descRow(mpDesc + o * 32);
            p.nObs = hasObs[o] ? 1 : 0;
            it.vp[i] = &p;
        }
    }
    std::atomic<int> next{0};
    // passes are kept separate so that a frame is never touched by two threads at once
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < std::max(repeat, 1); r++) {
        next = r * nFrames;
        const int stop = (r + 1) * nFrames;
        auto pass = [&]() {
            ORBmatcher matcher(nnratio);
            for (;;) {
                const int k = next.fetch_add(1);
                if (k >= stop) break;
                Item &it = items[k - r * nFrames];
                Frame &F = *it.F;
                for (int i = 0; i < FRAME_GRID_COLS; i++) for (int j = 0; j < FRAME_GRID_ROWS; j++) F.mGrid[i][j].clear();
                F.mvpMapPoints.assign((size_t)F.N, nullptr);
                F.AssignFeaturesToGrid();
                nMatches[k - r * nFrames] = matcher.SearchByProjection(F, it.vp, th, checkLevel != 0);
            }
        };
        if (threads <= 1) pass();
        else {
            std::vector<std::thread> th_;
            for (int t = 0; t < threads; t++) th_.emplace_back(pass);
            for (auto &t : th_) t.join();
        }
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    for (int f = 0; f < nFrames; f++)
        for (int i = 0; i < nKp[f]; i++) {
            MapPoint *p = items[f].F->mvpMapPoints[i];
            kpToMp[(size_t)f * kpStride + i] = p ? (int)(p - items[f].mps.data()) : -1;
        }
    return sec;
}

// ---- probe of the cv::Mat float expressions of oracle/minicv/minicv_matops.hpp, for tests/test_minicv_vs_cv2.py (pinned to
// cv2.gemm / cv2.norm of the installed OpenCV): out[0..2] = R * x + t, [3..5] = -R.t() * t, [6] = norm(x), [7] = x.dot(t),
// [8..16] = (1.0 / s) * R.t(), [17..25] = s * R, [26..28] = -R * t, [29..37] = R / s.
YGZR_API void ygzr_matops_probe(const float *R9, const float *x3, const float *t3, float s, float *out) {
    cv::Mat R(3, 3, cv::CV_32F, (void *)R9, 12), x(3, 1, cv::CV_32F, (void *)x3, 4), t(3, 1, cv::CV_32F, (void *)t3, 4);
    cv::Mat a = R * x + t;
    cv::Mat b = -R.t() * t;
    cv::Mat c = (1.0 / s) * R.t();
    cv::Mat e = s * R;
    cv::Mat f = -R * t;
    cv::Mat g = R / s;
    for (int i = 0; i < 3; i++) { out[i] = a.at<float>(i); out[3 + i] = b.at<float>(i); out[26 + i] = f.at<float>(i); }
    out[6] = (float)cv::norm(x);
    out[7] = (float)x.dot(t);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { out[8 + 3 * i + j] = c.at<float>(i, j); out[17 + 3 * i + j] = e.at<float>(i, j); out[29 + 3 * i + j] = g.at<float>(i, j); }
}
