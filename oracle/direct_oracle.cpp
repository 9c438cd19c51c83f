// direct_oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of SURVEY.md 8(f) row f4: the pyramid consumers of the
// direct tracker and the descriptor post-processing of MapPoint (part 4 of the oracle).  Linked into
// oracle/libygz_oracle.so; used by tests/ only — never by the product.
//
// PARITY UNPINNED: src/MapPoint.cc, src/ORBmatcher.cc, src/Align.cc and src/SparseImageAlign.cc need Eigen, sophus, OpenCV C++
// and glog, none of which exist in this image, and the reference ships no test vectors for them.  These are line-by-line
// restatements over flat arrays, checked against independent numpy re-derivations in tests/test_direct_oracle.py.
// Citations are file:line in /root/reference.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#define YGZO_API extern "C" __attribute__((visibility("default")))

namespace {
inline int descriptorDistance(const uint8_t *a, const uint8_t *b) {   // src/ORBmatcher.cc:1507-1523
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4);
        dist += __builtin_popcount(x ^ y);
    }
    return dist;
}
}  // namespace

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:211-271) for nPoints MapPoints.  The descriptors of point p are
// rows obsStart[p] .. obsStart[p+1]-1 of obsDesc: mObservations in map order with bad keyframes already skipped (:226-232).
// N x N distances (:243-250), per row the sorted distances' element [0.5 * (N - 1)] (:256-258), first row with the least
// median wins (:260-263).  No observation -> the function returns early (:222, :234): bestIdx = -1, descriptor untouched.
YGZO_API void ygzo_distinctive_descriptors(const uint8_t *obsDesc, const int32_t *obsStart, int nPoints, int32_t *bestIdx,
                                           int32_t *bestMedian, uint8_t *bestDesc) {
    std::vector<int> row;
    for (int p = 0; p < nPoints; p++) {
        const uint8_t *D = obsDesc + (size_t)obsStart[p] * 32;
        const int N = obsStart[p + 1] - obsStart[p];
        bestIdx[p] = -1; bestMedian[p] = INT_MAX;
        if (N <= 0) continue;
        int BestMedian = INT_MAX, BestIdx = 0;
        for (int i = 0; i < N; i++) {
            row.assign(N, 0);
            for (int j = 0; j < N; j++) row[j] = i == j ? 0 : descriptorDistance(D + 32 * i, D + 32 * j);
            std::sort(row.begin(), row.end());
            const int median = row[(size_t)(0.5 * (N - 1))];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        bestIdx[p] = BestIdx; bestMedian[p] = BestMedian;
        std::memcpy(bestDesc + (size_t)p * 32, D + 32 * BestIdx, 32);
    }
}

// =====================================================================================================================
// ORBmatcher::FindDirectProjection (src/ORBmatcher.cc:1573-1602) = GetWarpAffineMatrix (:1525-1547) + GetBestSearchLevel
// (include/ORBmatcher.h:226-238) + WarpAffine (:1549-1571, GetBilateralInterpUchar include/ORBmatcher.h:241-252) +
// ygz::Align2D (src/Align.cc:8-105), over flat arrays.
//
// Eigen / Sophus expressions are evaluated the way Eigen 3.3 evaluates them for these fixed sizes (no vectorisation for
// 2- and 3-vectors of float): sums of three products as a0 + (a1 + a2) (redux_novec_unroller splits [0,1) | [1,3)), sums of two
// as a0 + a1, Matrix2f::inverse / determinant and Matrix3f::inverse by the cofactor formulas of Eigen/src/LU/InverseImpl.h,
// SE3f * Vector3f = unit_quaternion()._transformVector(p) + translation() (Thirdparty/sophus/sophus/so3.hpp:332-335,
// se3.hpp) with Eigen's "uv = q.vec x v; uv += uv; v + w * uv + q.vec x uv".  Poses are passed the way Sophus stores them:
// unit quaternion coefficients x, y, z, w then the translation (7 floats); composing TCR = curr->mTcw * pose_ref.inverse()
// (:1582) is left to the caller (it is per keyframe, not per point).  Floats uncontracted (-ffp-contract=off).
namespace {
struct V3 { float x, y, z; };
inline V3 cross3(const V3 &a, const V3 &b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline V3 se3Mul(const float *T, const V3 &v) {   // T = qx qy qz qw tx ty tz
    const V3 q{T[0], T[1], T[2]};
    V3 uv = cross3(q, v);
    uv.x += uv.x; uv.y += uv.y; uv.z += uv.z;
    const V3 c = cross3(q, uv);
    const float w = T[3];
    V3 r{(v.x + w * uv.x) + c.x, (v.y + w * uv.y) + c.y, (v.z + w * uv.z) + c.z};
    r.x += T[4]; r.y += T[5]; r.z += T[6];
    return r;
}
inline float sum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }

// ygz::Align2D, src/Align.cc:8-105 (the scalar code is the only implementation there).  Returns converged; u, v in/out.
bool align2D(const uint8_t *cur, int curCols, int curRows, int curStep, const uint8_t *refPatchWithBorder, const uint8_t *refPatch,
             int nIter, float *uv) {
    const int halfpatch = 4, patchSize = 8;
    bool converged = false;
    float dxs[64], dys[64];
    float H[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    const int refStep = patchSize + 2;
    int k = 0;
    for (int y = 0; y < patchSize; ++y) {
        const uint8_t *it = refPatchWithBorder + (y + 1) * refStep + 1;
        for (int x = 0; x < patchSize; ++x, ++it, ++k) {
            float J[3];
            J[0] = (float)(0.5 * (it[1] - it[-1]));
            J[1] = (float)(0.5 * (it[refStep] - it[-refStep]));
            J[2] = 1;
            dxs[k] = J[0]; dys[k] = J[1];
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) H[r][c] += J[r] * J[c];
        }
    }
    // Matrix3f::inverse(): cofactors of column 0, determinant = sum(cofactors_col0 .* col0), every cofactor * invdet
    float Hinv[3][3];
    {
#define COF(i, j) (H[((i) + 1) % 3][((j) + 1) % 3] * H[((i) + 2) % 3][((j) + 2) % 3] - H[((i) + 1) % 3][((j) + 2) % 3] * H[((i) + 2) % 3][((j) + 1) % 3])
        const float c0 = COF(0, 0), c1 = COF(1, 0), c2 = COF(2, 0);
        const float det = sum3(c0 * H[0][0], c1 * H[1][0], c2 * H[2][0]);
        const float invdet = 1.0f / det;
        Hinv[0][0] = c0 * invdet; Hinv[0][1] = c1 * invdet; Hinv[0][2] = c2 * invdet;
        Hinv[1][0] = COF(0, 1) * invdet; Hinv[1][1] = COF(1, 1) * invdet; Hinv[1][2] = COF(2, 1) * invdet;
        Hinv[2][0] = COF(0, 2) * invdet; Hinv[2][1] = COF(1, 2) * invdet; Hinv[2][2] = COF(2, 2) * invdet;
#undef COF
    }
    float meanDiff = 0;
    float u = uv[0], v = uv[1];
    const float minUpdateSquared = 0.03 * 0.03;
    for (int iter = 0; iter < nIter; ++iter) {
        const int u_r = (int)std::floor(u), v_r = (int)std::floor(v);
        if (u_r < halfpatch || v_r < halfpatch || u_r >= curCols - halfpatch || v_r >= curRows - halfpatch) break;
        if (std::isnan(u) || std::isnan(v)) return false;
        const float sx = u - u_r, sy = v - v_r;
        const float wTL = (1.0 - sx) * (1.0 - sy), wTR = sx * (1.0 - sy), wBL = (1.0 - sx) * sy, wBR = sx * sy;
        float Jres[3] = {0, 0, 0};
        int kk = 0;
        for (int y = 0; y < patchSize; ++y) {
            const uint8_t *it = cur + (long)(v_r + y - halfpatch) * curStep + u_r - halfpatch;
            for (int x = 0; x < patchSize; ++x, ++it, ++kk) {
                const float searchPixel = wTL * it[0] + wTR * it[1] + wBL * it[curStep] + wBR * it[curStep + 1];
                const float res = searchPixel - refPatch[kk] + meanDiff;
                Jres[0] -= res * dxs[kk];
                Jres[1] -= res * dys[kk];
                Jres[2] -= res;
            }
        }
        float upd[3];
        for (int r = 0; r < 3; r++) upd[r] = sum3(Hinv[r][0] * Jres[0], Hinv[r][1] * Jres[1], Hinv[r][2] * Jres[2]);
        u += upd[0]; v += upd[1]; meanDiff += upd[2];
        if (upd[0] * upd[0] + upd[1] * upd[1] < minUpdateSquared) { converged = true; break; }
    }
    uv[0] = u; uv[1] = v;
    return converged;
}
}  // namespace

// One reference keyframe as FindDirectProjection reads it.  levels[l] = ref->mvImagePyramid[l] (continuous or pitched).
struct YgzoDirectKeyFrame {
    const uint8_t *levels[8];
    int32_t w[8], h[8], pitch[8];
    float poseRef[7];      // ref->GetPose()
    float tCurRef[7];      // TCR = curr->mTcw * pose_ref.inverse()
    float fx, fy, cx, cy;  // ref intrinsics
};

// Batched FindDirectProjection: point i is observed in keyframe kfIndex[i] at keypoint (pxRef[2i], pxRef[2i+1]) of octave[i],
// world position worldPos[3i..]; pxCurr[2i..] in: the projection estimate (mTrackProjX/Y), out: the aligned position;
// searchLevel[i], success[i] as the reference returns them.  cur* = curr->mvImagePyramid and intrinsics; scale / invScale =
// mvScaleFactors / mvInvScaleFactors (shared by all frames of a System); invLevelSigma2_1 = ref->mvInvLevelSigma2[1].
YGZO_API void ygzo_find_direct_projection(const YgzoDirectKeyFrame *kfs, int nKfs, const uint8_t *const *curLevels, const int32_t *curW,
                                          const int32_t *curH, const int32_t *curPitch, int nLevels, const float *curIntr,
                                          const float *scale, const float *invScale, float invLevelSigma2_1, int n,
                                          const int32_t *kfIndex, const float *pxRef, const int32_t *octave, const float *worldPos,
                                          float *pxCurr, int32_t *searchLevel, uint8_t *success) {
    const int WarpHalfPatchSize = 4, WarpPatchSize = 8;   // include/ORBmatcher.h:32-33
    for (int i = 0; i < n; i++) {
        const YgzoDirectKeyFrame &K = kfs[kfIndex[i]];
        const float prx = pxRef[2 * i], pry = pxRef[2 * i + 1];
        const int level = octave[i];
        // GetWarpAffineMatrix (:1525-1547)
        const V3 ptRef = se3Mul(K.poseRef, V3{worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]});
        const float depth = ptRef.z;
        const float dux = prx + (float)WarpHalfPatchSize * scale[level], duy = pry + 0.0f * scale[level];
        const float dvx = prx + 0.0f * scale[level], dvy = pry + (float)WarpHalfPatchSize * scale[level];
        const V3 ptDu{(dux - K.cx) * depth / K.fx, (duy - K.cy) * depth / K.fy, depth};   // KeyFrame::Pixel2Camera, include/KeyFrame.h:146-152
        const V3 ptDv{(dvx - K.cx) * depth / K.fx, (dvy - K.cy) * depth / K.fy, depth};
        const float fx = curIntr[0], fy = curIntr[1], cx = curIntr[2], cy = curIntr[3];
        auto world2pixel = [&](const V3 &p, float *o) {   // Frame::World2Pixel, include/Frame.h:170-183,197-199
            const V3 c = se3Mul(K.tCurRef, p);
            o[0] = fx * c.x / c.z + cx; o[1] = fy * c.y / c.z + cy;
        };
        float pc[2], pdu[2], pdv[2];
        world2pixel(ptRef, pc); world2pixel(ptDu, pdu); world2pixel(ptDv, pdv);
        const float A00 = (pdu[0] - pc[0]) / WarpHalfPatchSize, A10 = (pdu[1] - pc[1]) / WarpHalfPatchSize;   // ACR.col(0)
        const float A01 = (pdv[0] - pc[0]) / WarpHalfPatchSize, A11 = (pdv[1] - pc[1]) / WarpHalfPatchSize;   // ACR.col(1)
        // GetBestSearchLevel (include/ORBmatcher.h:226-238)
        int sl = 0;
        float D = A00 * A11 - A10 * A01;
        while (D > 3.0 && sl < nLevels - 1) { sl += 1; D *= invLevelSigma2_1; }
        searchLevel[i] = sl;
        // WarpAffine with half_patch_size = 5 (:1549-1571)
        uint8_t patchWithBorder[(WarpPatchSize + 2) * (WarpPatchSize + 2)], patch[WarpPatchSize * WarpPatchSize];
        {
            const float det = A00 * A11 - A10 * A01, invdet = 1.0f / det;   // Matrix2f::inverse (compute_inverse<.., 2>)
            const float R00 = A11 * invdet, R10 = -A10 * invdet, R01 = -A01 * invdet, R11 = A00 * invdet;
            const int half = WarpHalfPatchSize + 1, size = 2 * half;
            const uint8_t *img = K.levels[level];
            const int cols = K.w[level], rows = K.h[level], step = K.pitch[level];
            const float rpx = prx / scale[level], rpy = pry / scale[level];
            uint8_t *pp = patchWithBorder;
            for (int y = 0; y < size; y++)
                for (int x = 0; x < size; x++, ++pp) {
                    float ppx = (float)(x - half), ppy = (float)(y - half);
                    ppx *= scale[sl]; ppy *= scale[sl];
                    const float px = (R00 * ppx + R01 * ppy) + rpx, py = (R10 * ppx + R11 * ppy) + rpy;
                    if (px < 0 || py < 0 || px >= cols - 1 || py >= rows - 1) {
                        *pp = 0;
                    } else {
                        const double X = px, Y = py;
                        const double xx = X - std::floor(X), yy = Y - std::floor(Y);
                        const uint8_t *d = img + (long)(int)Y * step + (int)X;
                        *pp = (uint8_t)((1 - xx) * (1 - yy) * d[0] + xx * (1 - yy) * d[1] + (1 - xx) * yy * d[step] + xx * yy * d[step + 1]);
                    }
                }
        }
        for (int y = 1; y < WarpPatchSize + 1; ++y)   // remove the border (:1590-1596)
            for (int x = 0; x < WarpPatchSize; ++x) patch[(y - 1) * WarpPatchSize + x] = patchWithBorder[y * (WarpPatchSize + 2) + 1 + x];
        float ps[2] = {pxCurr[2 * i] * invScale[sl], pxCurr[2 * i + 1] * invScale[sl]};
        const bool ok = align2D(curLevels[sl], curW[sl], curH[sl], curPitch[sl], patchWithBorder, patch, 10, ps);
        pxCurr[2 * i] = ps[0] * scale[sl]; pxCurr[2 * i + 1] = ps[1] * scale[sl];
        success[i] = ok ? 1 : 0;
    }
}
