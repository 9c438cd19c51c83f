// direct_oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of SURVEY.md 8(f) row f4: the pyramid consumers of the
// direct tracker and the descriptor post-processing of MapPoint (part 4 of the oracle).  Linked into
// oracle/libygz_oracle.so; used by tests/ only — never by the product.
//
// PINNED to the reference's own code: src/MapPoint.cc:211-271, src/ORBmatcher.cc:1525-1602, src/Align.cc and
// src/SparseImageAlign.cc (+ include/NLSSolver.h, NLSSolver_impl.hpp) compiled unmodified into oracle/_ref/libygz_ref_match.so
// (oracle/Makefile target refmatch, oracle/ref_match_api.cpp); tests/test_ref_match.py asserts this file == that library, live and
// through tests/golden/match_golden.npz, bit for bit (SparseImgAlign: caches, H_, Jres_, chi2, n_meas_ and whole run() calls over
// levels 5..1 — run() indexes a 6-entry iterations[] with the level, so levels 6 and 7 are undefined behaviour in the reference).
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

// =====================================================================================================================
// SparseImgAlign (src/SparseImageAlign.cc:20-256, include/SparseImageAlign.h:95-116) driven by NLLSSolver<6, SE3f>::
// optimizeGaussNewton (include/NLSSolver_impl.hpp:17-92), as Tracking::TrackWithSparseAlignment runs it
// (src/Tracking.cc:284, :2171: max_level = nLevels - 1, min_level = 1, 10 iterations per level, Gauss-Newton, no robust
// weights, no prior).  Sums over features are sequential in float like the reference's H_ / Jres_ / chi2 (accumulate = 0) or,
// for the comparison with the device's tree reduction, in double over per-feature float partials (accumulate = 1).
// Sophus pieces (Thirdparty/sophus/sophus/se3.hpp:239-271, 407-428, so3.hpp:268-276, 426-456) restated on (q, t) storage;
// H_.ldlt().solve(Jres_) as a diagonally pivoted LDL^T (Eigen/src/Cholesky/LDLT.h).
namespace {
struct Se3 { float q[4]; float t[3]; };   // q = x y z w
inline Se3 se3Load(const float *p) { Se3 T; std::memcpy(T.q, p, 16); std::memcpy(T.t, p + 4, 12); return T; }
inline void se3Store(const Se3 &T, float *p) { std::memcpy(p, T.q, 16); std::memcpy(p + 4, T.t, 12); }
inline V3 se3Apply(const Se3 &T, const V3 &v) { float b[7]; se3Store(T, b); return se3Mul(b, v); }
Se3 se3Compose(const Se3 &A, const Se3 &B) {   // A *= B: fastMultiply (translation first, then quaternion product) + normalize
    Se3 R;
    const V3 rt = [&] { float b[7]; Se3 Rot = A; Rot.t[0] = Rot.t[1] = Rot.t[2] = 0; se3Store(Rot, b); return se3Mul(b, V3{B.t[0], B.t[1], B.t[2]}); }();
    R.t[0] = A.t[0] + rt.x; R.t[1] = A.t[1] + rt.y; R.t[2] = A.t[2] + rt.z;
    const float ax = A.q[0], ay = A.q[1], az = A.q[2], aw = A.q[3], bx = B.q[0], by = B.q[1], bz = B.q[2], bw = B.q[3];
    float q[4] = {aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx,
                  aw * bw - ax * bx - ay * by - az * bz};
    const float len = std::sqrt((q[0] * q[0] + q[1] * q[1]) + (q[2] * q[2] + q[3] * q[3]));   // coeffs().norm(): one packet of four, pairwise
    for (int i = 0; i < 4; i++) R.q[i] = q[i] / len;
    return R;
}
Se3 se3Exp(const float *a) {   // a = (upsilon, omega); Thirdparty/sophus/sophus/se3.hpp:407-428 over so3.hpp:426-456
    const float wx = a[3], wy = a[4], wz = a[5];
    const float thetaSq = sum3(wx * wx, wy * wy, wz * wz), theta = std::sqrt(thetaSq), half = 0.5f * theta;   // omega.squaredNorm()
    const float eps = 1e-5f;   // SophusConstants<float>::epsilon()
    float imag, real;
    if (theta < eps) {
        const float po4 = thetaSq * thetaSq;
        imag = 0.5f - (float)(1.0 / 48.0) * thetaSq + (float)(1.0 / 3840.0) * po4;
        real = 1.0f - 0.5f * thetaSq + (float)(1.0 / 384.0) * po4;
    } else {
        imag = std::sin(half) / theta; real = std::cos(half);
    }
    Se3 T;
    T.q[0] = imag * wx; T.q[1] = imag * wy; T.q[2] = imag * wz; T.q[3] = real;
    {   // the SO3Group(Quaternion) constructor normalises (so3.hpp:739-743 -> :268-276)
        const float len = std::sqrt((T.q[0] * T.q[0] + T.q[1] * T.q[1]) + (T.q[2] * T.q[2] + T.q[3] * T.q[3]));
        for (int i = 0; i < 4; i++) T.q[i] = T.q[i] / len;
    }
    const float O[3][3] = {{0, -wz, wy}, {wz, 0, -wx}, {-wy, wx, 0}};
    float O2[3][3], V[3][3];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) O2[r][c] = sum3(O[r][0] * O[0][c], O[r][1] * O[1][c], O[r][2] * O[2][c]);
    if (theta < eps) {   // V = so3.matrix() = Quaternion::toRotationMatrix()
        const float x = T.q[0], y = T.q[1], z = T.q[2], w = T.q[3];
        const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
        const float twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
        const float M[3][3] = {{1.0f - (tyy + tzz), txy - twz, txz + twy}, {txy + twz, 1.0f - (txx + tzz), tyz - twx},
                               {txz - twy, tyz + twx, 1.0f - (txx + tyy)}};
        std::memcpy(V, M, sizeof V);
    } else {
        const float thetaSq2 = theta * theta;   // se3.hpp:422 squares the root again
        const float A = (1.0f - std::cos(theta)) / thetaSq2, B = (theta - std::sin(theta)) / (thetaSq2 * theta);
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) V[r][c] = ((r == c ? 1.0f : 0.0f) + A * O[r][c]) + B * O2[r][c];
    }
    for (int r = 0; r < 3; r++) T.t[r] = sum3(V[r][0] * a[0], V[r][1] * a[1], V[r][2] * a[2]);
    return T;
}
// x = H.ldlt().solve(b), 6 x 6: LDL^T with the largest remaining diagonal entry as pivot
void ldltSolve6(const float *Hin, const float *b, float *x) {
    const int n = 6;
    float A[6][6]; int perm[6];
    for (int r = 0; r < n; r++) { perm[r] = r; for (int c = 0; c < n; c++) A[r][c] = Hin[r * 6 + c]; }
    for (int k = 0; k < n; k++) {
        int p = k;
        for (int i = k + 1; i < n; i++) if (std::fabs(A[i][i]) > std::fabs(A[p][p])) p = i;
        if (p != k) {
            for (int c = 0; c < n; c++) std::swap(A[k][c], A[p][c]);
            for (int r = 0; r < n; r++) std::swap(A[r][k], A[r][p]);
            std::swap(perm[k], perm[p]);
        }
        float d = A[k][k];
        for (int j = 0; j < k; j++) d -= A[k][j] * A[k][j] * A[j][j];
        A[k][k] = d;
        for (int i = k + 1; i < n; i++) {
            float v = A[i][k];
            for (int j = 0; j < k; j++) v -= A[i][j] * A[k][j] * A[j][j];
            A[i][k] = d != 0 ? v / d : v;
        }
    }
    float y[6];
    for (int i = 0; i < n; i++) { float v = b[perm[i]]; for (int j = 0; j < i; j++) v -= A[i][j] * y[j]; y[i] = v; }
    for (int i = 0; i < n; i++) y[i] = std::fabs(A[i][i]) > 1.17549435e-38f ? y[i] / A[i][i] : 0.0f;
    for (int i = n - 1; i >= 0; i--) { float v = y[i]; for (int j = i + 1; j < n; j++) v -= A[j][i] * y[j]; y[i] = v; }
    for (int i = 0; i < n; i++) x[perm[i]] = y[i];
}

struct SparseAlignState {
    int n = 0;
    std::vector<float> refPatch;   // ref_patch_cache_, n x 16 (persists over the levels of a run)
    std::vector<float> jac;        // jacobian_cache_, n x 16 x 6 (zeroed per level)
    std::vector<uint8_t> visible;  // visible_fts_ (set per level, never cleared inside a run)
};

// SparseImgAlign::precomputeReferencePatches (src/SparseImageAlign.cc:57-128) at one level
void precomputeRef(SparseAlignState &S, const uint8_t *img, int cols, int rows, int stride, float scale, float focal, const float *refPose,
                   const float *kpXY, const uint8_t *valid, const float *worldPos) {
    const int border = 3, half = 2, size = 4;
    for (int i = 0; i < S.n; i++) {
        if (!valid[i]) continue;
        const float u = kpXY[2 * i] * scale, v = kpXY[2 * i + 1] * scale;
        const int ui = (int)floorf(u), vi = (int)floorf(v);
        if (ui - border < 0 || vi - border < 0 || ui + border >= cols || vi + border >= rows) continue;
        S.visible[i] = 1;
        const V3 p = se3Mul(refPose, V3{worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]});
        // JacobXYZ2Cam (include/SparseImageAlign.h:95-116)
        float J[2][6];
        const float x = p.x, y = p.y, zInv = (float)(1. / p.z), zInv2 = zInv * zInv;
        J[0][0] = -zInv; J[0][1] = 0.0f; J[0][2] = x * zInv2; J[0][3] = y * J[0][2]; J[0][4] = (float)(-(1.0 + x * J[0][2])); J[0][5] = y * zInv;
        J[1][0] = 0.0f; J[1][1] = -zInv; J[1][2] = y * zInv2; J[1][3] = (float)(1.0 + y * J[1][2]); J[1][4] = -J[0][3]; J[1][5] = -x * zInv;
        const float su = u - ui, sv = v - vi;
        const float wTL = (1.0 - su) * (1.0 - sv), wTR = su * (1.0 - sv), wBL = (1.0 - su) * sv, wBR = su * sv;
        const float fs = focal * scale;
        int k = 0;
        for (int yy = 0; yy < size; ++yy) {
            const uint8_t *r = img + (long)(vi + yy - half) * stride + (ui - half);
            for (int xx = 0; xx < size; ++xx, ++r, ++k) {
                S.refPatch[i * 16 + k] = wTL * r[0] + wTR * r[1] + wBL * r[stride] + wBR * r[stride + 1];
                const float dx = 0.5f * ((wTL * r[1] + wTR * r[2] + wBL * r[stride + 1] + wBR * r[stride + 2]) -
                                         (wTL * r[-1] + wTR * r[0] + wBL * r[stride - 1] + wBR * r[stride]));
                const float dy = 0.5f * ((wTL * r[stride] + wTR * r[1 + stride] + wBL * r[stride * 2] + wBR * r[stride * 2 + 1]) -
                                         (wTL * r[-stride] + wTR * r[1 - stride] + wBL * r[0] + wBR * r[1]));
                for (int c = 0; c < 6; c++) S.jac[(i * 16 + k) * 6 + c] = (dx * J[0][c] + dy * J[1][c]) * fs;
            }
        }
    }
}

// SparseImgAlign::computeResiduals (:130-225) with linearize_system, unit weights.  Returns the chi2 SUM; *nMeas counts.
template <class Acc> float computeRes(const SparseAlignState &S, const uint8_t *cur, int cols, int rows, int stride, float scale,
                                      const float *curIntr, const float *refPose, const float *tCurRef, const float *worldPos,
                                      float *H36, float *Jres6, long *nMeas) {
    const int border = 3, half = 2, size = 4;
    Acc H[6][6], Jr[6], chi2 = 0;
    for (int r = 0; r < 6; r++) { Jr[r] = 0; for (int c = 0; c < 6; c++) H[r][c] = 0; }
    long meas = 0;
    const bool perFeature = sizeof(Acc) == sizeof(double);
    for (int i = 0; i < S.n; i++) {
        if (!S.visible[i]) continue;
        const V3 pr = se3Mul(refPose, V3{worldPos[3 * i], worldPos[3 * i + 1], worldPos[3 * i + 2]});
        const V3 pc = se3Mul(tCurRef, pr);
        const float uc = (curIntr[0] * pc.x / pc.z + curIntr[2]) * scale, vc = (curIntr[1] * pc.y / pc.z + curIntr[3]) * scale;
        const int ui = (int)floorf(uc), vi = (int)floorf(vc);
        if (ui < 0 || vi < 0 || ui - border < 0 || vi - border < 0 || ui + border >= cols || vi + border >= rows) continue;
        const float su = uc - ui, sv = vc - vi;
        const float wTL = (1.0 - su) * (1.0 - sv), wTR = su * (1.0 - sv), wBL = (1.0 - su) * sv, wBR = su * sv;
        float fH[6][6], fJ[6], fChi = 0;   // per-feature float partials (accumulate = 1)
        for (int r = 0; r < 6; r++) { fJ[r] = 0; for (int c = 0; c < 6; c++) fH[r][c] = 0; }
        int k = 0;
        for (int yy = 0; yy < size; ++yy) {
            const uint8_t *p = cur + (long)(vi + yy - half) * stride + (ui - half);
            for (int xx = 0; xx < size; ++xx, ++k, ++p) {
                const float inten = wTL * p[0] + wTR * p[1] + wBL * p[stride] + wBR * p[stride + 1];
                const float res = inten - S.refPatch[i * 16 + k];
                const float *J = &S.jac[(i * 16 + k) * 6];
                meas++;
                if (perFeature) {
                    fChi += res * res * 1.0f;
                    for (int r = 0; r < 6; r++) { fJ[r] -= J[r] * res * 1.0f; for (int c = 0; c < 6; c++) fH[r][c] += J[r] * J[c] * 1.0f; }
                } else {
                    chi2 += res * res * 1.0f;
                    for (int r = 0; r < 6; r++) { Jr[r] -= J[r] * res * 1.0f; for (int c = 0; c < 6; c++) H[r][c] += J[r] * J[c] * 1.0f; }
                }
            }
        }
        if (perFeature) {
            chi2 += fChi;
            for (int r = 0; r < 6; r++) { Jr[r] += fJ[r]; for (int c = 0; c < 6; c++) H[r][c] += fH[r][c]; }
        }
    }
    for (int r = 0; r < 6; r++) { Jres6[r] = (float)Jr[r]; for (int c = 0; c < 6; c++) H36[r * 6 + c] = (float)H[r][c]; }
    *nMeas = meas;
    return (float)chi2;
}
}  // namespace

struct YgzoPyramid { const uint8_t *levels[8]; int32_t w[8], h[8], pitch[8]; };

// One computeResiduals call at `level` after precomputeReferencePatches at that level on a fresh state (unit test of the two
// hot loops).  Outputs H (36), Jres (6), chi2 sum, n_meas, visible[n], refPatch[n*16], jac[n*96].
YGZO_API void ygzo_sparse_align_residuals(const YgzoPyramid *ref, const YgzoPyramid *cur, int level, float scale, float focal,
                                          const float *curIntr, const float *refPose, const float *tCurRef, int n, const float *kpXY,
                                          const uint8_t *valid, const float *worldPos, int accumulate, float *H36, float *Jres6,
                                          float *chi2, int64_t *nMeas, uint8_t *visible, float *refPatch, float *jac) {
    SparseAlignState S;
    S.n = n; S.refPatch.assign((size_t)n * 16, 0.f); S.jac.assign((size_t)n * 96, 0.f); S.visible.assign(n, 0);
    precomputeRef(S, ref->levels[level], ref->w[level], ref->h[level], ref->pitch[level], scale, focal, refPose, kpXY, valid, worldPos);
    long m = 0;
    *chi2 = accumulate ? computeRes<double>(S, cur->levels[level], cur->w[level], cur->h[level], cur->pitch[level], scale, curIntr, refPose, tCurRef, worldPos, H36, Jres6, &m)
                       : computeRes<float>(S, cur->levels[level], cur->w[level], cur->h[level], cur->pitch[level], scale, curIntr, refPose, tCurRef, worldPos, H36, Jres6, &m);
    *nMeas = m;
    if (visible) std::memcpy(visible, S.visible.data(), n);
    if (refPatch) std::memcpy(refPatch, S.refPatch.data(), (size_t)n * 64);
    if (jac) std::memcpy(jac, S.jac.data(), (size_t)n * 384);
}

// SparseImgAlign::run (:20-55): tCurRef in = cur->mTcw * ref->mTcw.inverse(), out = TCR.  Returns n_meas_ / patch_area_.
// stats (may be NULL): [0] total iterations, [1] final chi2_ (mean).
YGZO_API int64_t ygzo_sparse_align_run(const YgzoPyramid *ref, const YgzoPyramid *cur, int maxLevel, int minLevel, const float *invScale,
                                       float focal, const float *curIntr, const float *refPose, int n, const float *kpXY,
                                       const uint8_t *valid, const float *worldPos, int accumulate, float *tCurRef, float *stats) {
    SparseAlignState S;
    S.n = n; S.refPatch.assign((size_t)n * 16, 0.f); S.jac.assign((size_t)n * 96, 0.f); S.visible.assign(n, 0);
    if (n == 0) return 0;   // "no features to track" (:24-27)
    Se3 model = se3Load(tCurRef);
    float chi2_ = 1e10f;      // reset() (include/NLSSolver_impl.hpp:289-298)
    bool stop = false;
    long nMeas = 0, iters = 0;
    const float eps = 0.000001f;
    for (int level = maxLevel; level >= minLevel; level -= 1) {
        std::fill(S.jac.begin(), S.jac.end(), 0.f);
        precomputeRef(S, ref->levels[level], ref->w[level], ref->h[level], ref->pitch[level], invScale[level], focal, refPose, kpXY, valid, worldPos);
        Se3 oldModel = model;
        for (int iter = 0; iter < 10; ++iter) {
            float H[36], Jres[6], T[7], x[6];
            se3Store(model, T);
            const float sum = accumulate ? computeRes<double>(S, cur->levels[level], cur->w[level], cur->h[level], cur->pitch[level], invScale[level], curIntr, refPose, T, worldPos, H, Jres, &nMeas)
                                         : computeRes<float>(S, cur->levels[level], cur->w[level], cur->h[level], cur->pitch[level], invScale[level], curIntr, refPose, T, worldPos, H, Jres, &nMeas);
            const float newChi2 = sum / nMeas;
            iters++;
            ldltSolve6(H, Jres, x);
            if (std::isnan(x[0])) stop = true;
            if ((iter > 0 && newChi2 > 1.2 * chi2_) || stop) { model = oldModel; break; }
            float mx[6];
            for (int i = 0; i < 6; i++) mx[i] = -x[i];
            const Se3 newModel = se3Compose(model, se3Exp(mx));
            oldModel = model; model = newModel;
            chi2_ = newChi2;
            float nm = -1;
            for (int i = 0; i < 6; i++) nm = std::max(nm, std::fabs(x[i]));
            if (nm <= eps) break;
        }
    }
    se3Store(model, tCurRef);
    if (stats) { stats[0] = (float)iters; stats[1] = chi2_; }
    return nMeas / 16;
}
