// ygz_match.cuh — declarations shared by match_kernels.cu / matcher.cu.
#pragma once
#include "ygz_common.cuh"

namespace ygz {

constexpr int kGridCols = 64, kGridRows = 48;          // FRAME_GRID_COLS/ROWS, include/Frame.h:27-28
constexpr int kThHigh = 100, kThLow = 50, kHistoLength = 30;  // src/ORBmatcher.cc:36-38

// Device-side view of the arguments of SearchByProjection for a batch of frames (all device pointers).
struct ProjArgs {
    int nFrames, kpStride, mpStride;
    const int32_t *nKp, *nMp;
    const ygzorb_keypoint *kps;
    const uint8_t *desc;
    const float *uRight;       // may be null
    const uint8_t *claimed;    // may be null
    float minX, minY, invW, invH;
    float scale[kMaxLevels];
    const float *projX, *projY, *projXR, *viewCos;
    const int32_t *level;
    const uint8_t *inView, *hasObs, *mpDesc;
    float th, nnratio;
    int checkLevel;
    // LastFrame variant (src/ORBmatcher.cc:1218-1350): the "MapPoints" are LastFrame's keypoints; their projection
    // is computed in the kernel from the pose (30 floats per frame pair: Rcw, tcw, Rlw, tlw, fx, fy, cx, cy, mbf, mb)
    const float *pose, *worldPos, *angleLast;
    const int32_t *octaveLast;
    float maxX, maxY;
    int bMono, checkOrientation;
    // KeyFrame-point variants: relocalisation (:1352-1469, MODE 2; pose = Rcw, tcw, -, -, fx, fy, cx, cy) and Sim3
    // (:265-373, MODE 3; pose = Rcw, tcw, Ow, -, fx, fy, cx, cy).  maxDist / minDist = mfMaxDistance / mfMinDistance.
    // No-claim variants: Fuse (:748-886, MODE 4, Eigen; pose[12..14] = Ow, pose[28] = bf), Fuse with Sim3 (:888-1004, MODE 5,
    // cv::Mat), one direction of SearchBySim3 (:1006-1216, MODE 6; pose = Rw, tw, sR, t of the target camera).
    const float *maxDist, *minDist, *normal;
    float invSigma2[kMaxLevels];
    float logScale;
    int nScaleLevels, bestDistMax;
    int32_t *kpToMp, *mpBestIdx, *nMatches;   // outputs (mpBestIdx is also the iteration state)
    int32_t *claimTime;                       // scratch [nFrames * kpStride]
    int32_t *gridList;                        // scratch [nFrames * kpStride]
    uint32_t *candCache;                      // scratch [nFrames * 12 * mpStride]
    uint8_t *candCnt;                         // scratch [nFrames * mpStride]
    int stageSmem;                            // set by the launcher: frame-side arrays staged in dynamic shared memory
};

// ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:375-478), batched over frame pairs
struct InitArgs {
    int nFrames, stride1, stride2;
    const int32_t *n1, *n2;
    const ygzorb_keypoint *kps1, *kps2;
    const uint8_t *desc1, *desc2;
    float minX, minY, invW, invH;      // grid of F2
    float *prevMatched;                // vbPrevMatched, (x, y) per keypoint of F1, updated in place
    float window, nnratio;
    int checkOrientation;
    int32_t *matches12, *nMatches;
    int32_t *gridList;                 // scratch [nFrames * stride2]
    int32_t *state;                    // scratch [nFrames * (5 * stride1 + 2 * stride2)]
};

struct KfView {          // one KeyFrame's arrays for SearchForTriangulation (device pointers)
    int n, nNodes;
    const ygzorb_keypoint *kps;
    const uint8_t *desc;
    const float *uRight;
    const uint8_t *hasMP;
    const uint32_t *nodeIds;
    const int32_t *nodeStart;
    const uint32_t *nodeFeats;
};
struct TriArgs {
    KfView kf1, kf2;
    float scale2[kMaxLevels], sigma2_2[kMaxLevels];
    float F12[9];
    float ex, ey;
    int onlyStereo, checkOrientation;
    float nnratio;          // SearchByBoW only
    int32_t *matches12;     // [kf1.n]
    int32_t *nodeOf1;       // scratch [kf1.n]: index into kf2's node list, or -1
    int32_t *nMatches;      // [1]
    int32_t *hist;          // scratch [kHistoLength]
};

struct StereoArgs {
    const ygzorb_keypoint *kpsL, *kpsR;
    const uint8_t *descL, *descR;
    int nL, nR, nlevels;
    const uint8_t *lvL[kMaxLevels], *lvR[kMaxLevels];   // level base pointers (left / right pyramids)
    int pitchL[kMaxLevels], pitchR[kMaxLevels], lw[kMaxLevels], lh[kMaxLevels];
    float scale[kMaxLevels], invScale[kMaxLevels];
    float mbf, mb;
    float *uRight, *depth;      // outputs [nL]
    int32_t *bestDist;          // scratch/output [nL]: integer SAD of accepted keypoints, -1 otherwise
    int32_t *rowStart;          // scratch [rows + 1]
    int32_t *rowList;           // scratch [rowListCap]
    int rowListCap;
    int32_t *sortBuf;           // scratch [nL]
    int32_t *status;
};

int launchPopcPeak(unsigned *out, int blocks, int threads, int iters, cudaStream_t s);
int launchDescriptorDistance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist, cudaStream_t s);
int knnSplits(int nq, int nt, int smCount, size_t maxPartialBytes);
size_t knnPartialBytes(int nq, int splits);
int launchKnn2(const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio, int th, int splits, void *partial,
               int variant, int32_t *bestIdx, int32_t *d1, int32_t *d2, int32_t *match, cudaStream_t s);
int launchSearchByProjection(const ProjArgs &A, int maxKp, cudaStream_t s);
int launchSearchByProjectionLast(const ProjArgs &A, cudaStream_t s);
int launchSearchForInitialization(const InitArgs &A, cudaStream_t s);
int launchSearchByProjectionKeyFrame(const ProjArgs &A, int mode, cudaStream_t s);
int launchSim3Agreement(const int32_t *match1, const int32_t *match2, const int32_t *n1, const int32_t *n2, int stride1, int stride2,
                        int nFrames, int32_t *matches12, int32_t *nFound, cudaStream_t s);

#ifdef __CUDACC__
// ORBmatcher::DescriptorDistance / FORB::distance on two 256-bit rows held as 2 x uint4
__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
#endif

// Bag-of-words transform (bow_kernels.cu).  Vocabulary arrays are device pointers owned by a ygzorb_vocabulary.
struct BowArgs {
    int nFrames, stride, L, levelsup;
    const int32_t *nFeat;
    const uint8_t *desc;
    const int32_t *childStart, *childList;    // CSR of the tree; children of node i: childList[childStart[i] .. childStart[i+1])
    const uint8_t *childDesc;                 // 32 B per child-list entry (descriptor of childList[c])
    const int32_t *wordId;                    // per node, -1 for inner nodes
    const double *nodeWeight;                 // per node
    int32_t *word, *node;                     // per feature
    double *weight;
    int32_t *bowCount, *bowWord; double *bowValue;
    int32_t *fvCount; uint32_t *fvNode; int32_t *fvStart; uint32_t *fvFeat;
};
int launchBowTransform(const BowArgs &A, cudaStream_t s);
int launchTriangulation(const TriArgs &A, cudaStream_t s);
int launchSearchByBoW(const TriArgs &A, cudaStream_t s);
int launchStereo(const StereoArgs &A, cudaStream_t s);
// ORBmatcher::FindDirectProjection (direct_kernels.cu); all pointers are device pointers
struct DirectKf {
    const uint8_t *lv[kMaxLevels];
    int w[kMaxLevels], h[kMaxLevels], pitch[kMaxLevels];
    float poseRef[7], tCurRef[7];
    float fx, fy, cx, cy;
};
struct DirectArgs {
    const DirectKf *kfs;
    const uint8_t *curLv[kMaxLevels];
    int curW[kMaxLevels], curH[kMaxLevels], curPitch[kMaxLevels];
    int nLevels, n;
    float fx, fy, cx, cy;
    float scale[kMaxLevels], invScale[kMaxLevels];
    float invLevelSigma2_1;
    const int32_t *kfIndex, *octave;
    const float *pxRef, *worldPos;
    float *pxCurr;
    int32_t *searchLevel;
    uint8_t *success;
};
int launchFindDirectProjection(const DirectArgs &A, cudaStream_t s);
// SparseImgAlign (direct_kernels.cu); device pointers.  refPatch / jac / visible are the handle's caches
// (ref_patch_cache_, jacobian_cache_, visible_fts_ of include/SparseImageAlign.h:82-86).
struct SparseAlignArgs {
    const uint8_t *refLv[kMaxLevels], *curLv[kMaxLevels];
    int w[kMaxLevels], h[kMaxLevels], refPitch[kMaxLevels], curPitch[kMaxLevels];
    float invScale[kMaxLevels];
    float focal, fx, fy, cx, cy;
    int n, stride;                      // caches are feature-minor: element e of feature i at e * stride + i
    const float *kpXY, *worldPos;
    const uint8_t *valid;
    const float *refPose, *tCurRef;     // 7 floats each, device
    float *refPatch, *jac;
    uint8_t *visible;
};
int launchSparseAlignPrecompute(const SparseAlignArgs &A, int level, cudaStream_t s);
int launchSparseAlignResiduals(const SparseAlignArgs &A, int level, float *out, cudaStream_t s);
int launchSparseAlignRun(const SparseAlignArgs &A, int maxLevel, int minLevel, float *out, cudaStream_t s);
int launchDistinctive(const uint8_t *obsDesc, const int32_t *obsStart, int nPoints, int32_t *bestIdx, int32_t *bestMedian,
                      uint8_t *bestDesc, cudaStream_t s);
int configureMatchKernels();

}  // namespace ygz
