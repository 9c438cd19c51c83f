// direct_kernels.cu — SURVEY.md §8(f) row f4: the descriptor post-processing of MapPoint and the pyramid consumers of the
// direct tracker (MapPoint::ComputeDistinctiveDescriptors src/MapPoint.cc:211-271; ORBmatcher::FindDirectProjection
// src/ORBmatcher.cc:1525-1602 with ygz::Align2D src/Align.cc:8-105; SparseImgAlign src/SparseImageAlign.cc:60-256).
#include "ygz_match.cuh"

namespace ygz {

// ---- K13: MapPoint::ComputeDistinctiveDescriptors, one CTA per MapPoint, one thread per observation row.
// The reference sorts each row of the N x N distance matrix and reads element k = (int)(0.5 * (N - 1)); distances are
// integers in [0, 256], so that element is the smallest v with #{j : d_ij <= v} >= k + 1: a 9-step bisection on v, each
// step one sweep over the point's descriptors (in shared memory up to kDistinctRows rows, else from L2).  The first row with
// the least median wins (strict <, :260-263) = the minimum of the key (median << 32 | row).
constexpr int kDistinctThreads = 64, kDistinctRows = 512;

__global__ void __launch_bounds__(kDistinctThreads) k_distinctive(const uint8_t *obsDesc, const int32_t *obsStart, int nPoints,
                                                                  int32_t *bestIdx, int32_t *bestMedian, uint8_t *bestDesc) {
    __shared__ uint4 rows[2 * kDistinctRows];
    __shared__ unsigned long long best;
    const int p = blockIdx.x, tid = threadIdx.x;
    const int s = obsStart[p], N = obsStart[p + 1] - s;
    if (N <= 0) {
        if (tid == 0) { bestIdx[p] = -1; bestMedian[p] = INT_MAX; }
        return;
    }
    const uint4 *D = reinterpret_cast<const uint4 *>(obsDesc + (size_t)s * 32);
    const bool staged = N <= kDistinctRows;
    if (staged)
        for (int i = tid; i < 2 * N; i += kDistinctThreads) rows[i] = D[i];
    if (tid == 0) best = ~0ull;
    __syncthreads();
    const uint4 *R = staged ? rows : D;
    const int need = ((N - 1) >> 1) + 1;   // k + 1 with k = (int)(0.5 * (N - 1))
    for (int i = tid; i < N; i += kDistinctThreads) {
        const uint4 a0 = R[2 * i], a1 = R[2 * i + 1];
        int lo = 0, hi = 256;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < N; j++) cnt += hamming256(a0, a1, R[2 * j], R[2 * j + 1]) <= mid;   // d_ii = 0 like Distances[i][i]
            if (cnt >= need) hi = mid; else lo = mid + 1;
        }
        atomicMin(&best, ((unsigned long long)lo << 32) | (unsigned)i);
    }
    __syncthreads();
    const int bi = (int)(best & 0xffffffffu);
    if (tid == 0) { bestIdx[p] = bi; bestMedian[p] = (int)(best >> 32); }
    if (tid < 2) reinterpret_cast<uint4 *>(bestDesc + (size_t)p * 32)[tid] = R[2 * bi + tid];
}

int launchDistinctive(const uint8_t *obsDesc, const int32_t *obsStart, int nPoints, int32_t *bestIdx, int32_t *bestMedian,
                      uint8_t *bestDesc, cudaStream_t s) {
    if (nPoints <= 0) return 0;
    k_distinctive<<<nPoints, kDistinctThreads, 0, s>>>(obsDesc, obsStart, nPoints, bestIdx, bestMedian, bestDesc);
    return 1;
}

// ---- K14: ORBmatcher::FindDirectProjection, one thread per (MapPoint, reference keyframe) pair.  Everything a point does
// is sequential float arithmetic whose order fixes the result (the Hessian and Jres sums of Align2D run over the 64 patch
// pixels in raster order), so a point is not split over lanes; the per-thread patch and gradient arrays live in shared
// memory, lane-interleaved (element k of lane l at k * 32 + l: conflict-free).  Eigen / Sophus expressions are evaluated the
// way Eigen 3.3 evaluates them for these fixed sizes: three-term sums as a0 + (a1 + a2), Matrix2f / Matrix3f inverse by the
// cofactor formulas of Eigen/src/LU/InverseImpl.h, SE3f * point = q._transformVector(p) + t with Eigen's
// "uv = q.vec x v; uv += uv; v + w * uv + q.vec x uv"; --fmad=false keeps every product and sum separate.
constexpr int kDirectThreads = 32;

struct DV3 { float x, y, z; };
__device__ __forceinline__ DV3 dCross(const DV3 &a, const DV3 &b) { return DV3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
__device__ __forceinline__ DV3 dSe3Mul(const float *T, const DV3 &v) {   // Sophus SE3f (qx qy qz qw tx ty tz) * point
    const DV3 q{T[0], T[1], T[2]};
    DV3 uv = dCross(q, v);
    uv.x += uv.x; uv.y += uv.y; uv.z += uv.z;
    const DV3 c = dCross(q, uv);
    const float w = T[3];
    DV3 r{(v.x + w * uv.x) + c.x, (v.y + w * uv.y) + c.y, (v.z + w * uv.z) + c.z};
    r.x += T[4]; r.y += T[5]; r.z += T[6];
    return r;
}
__device__ __forceinline__ float dSum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }

__global__ void __launch_bounds__(kDirectThreads) k_find_direct_projection(const __grid_constant__ DirectArgs A) {
    __shared__ uint8_t sPatch[100 * kDirectThreads];
    __shared__ float sDx[64 * kDirectThreads], sDy[64 * kDirectThreads];
    const int lane = threadIdx.x, i = blockIdx.x * kDirectThreads + lane;
    if (i >= A.n) return;
    const DirectKf &K = A.kfs[A.kfIndex[i]];
    const float prx = A.pxRef[2 * i], pry = A.pxRef[2 * i + 1];
    const int level = A.octave[i];
    // GetWarpAffineMatrix (src/ORBmatcher.cc:1525-1547)
    const DV3 ptRef = dSe3Mul(K.poseRef, DV3{A.worldPos[3 * i], A.worldPos[3 * i + 1], A.worldPos[3 * i + 2]});
    const float depth = ptRef.z, sRef = A.scale[level];
    const float dux = prx + 4.0f * sRef, duy = pry + 0.0f * sRef, dvx = prx + 0.0f * sRef, dvy = pry + 4.0f * sRef;
    const DV3 ptDu{(dux - K.cx) * depth / K.fx, (duy - K.cy) * depth / K.fy, depth};
    const DV3 ptDv{(dvx - K.cx) * depth / K.fx, (dvy - K.cy) * depth / K.fy, depth};
    float pc[2], pdu[2], pdv[2];
    {
        DV3 c = dSe3Mul(K.tCurRef, ptRef);
        pc[0] = A.fx * c.x / c.z + A.cx; pc[1] = A.fy * c.y / c.z + A.cy;
        c = dSe3Mul(K.tCurRef, ptDu);
        pdu[0] = A.fx * c.x / c.z + A.cx; pdu[1] = A.fy * c.y / c.z + A.cy;
        c = dSe3Mul(K.tCurRef, ptDv);
        pdv[0] = A.fx * c.x / c.z + A.cx; pdv[1] = A.fy * c.y / c.z + A.cy;
    }
    const float A00 = (pdu[0] - pc[0]) / 4.0f, A10 = (pdu[1] - pc[1]) / 4.0f;
    const float A01 = (pdv[0] - pc[0]) / 4.0f, A11 = (pdv[1] - pc[1]) / 4.0f;
    // GetBestSearchLevel (include/ORBmatcher.h:226-238)
    int sl = 0;
    const float det2 = A00 * A11 - A10 * A01;
    {
        float D = det2;
        while ((double)D > 3.0 && sl < A.nLevels - 1) { sl += 1; D *= A.invLevelSigma2_1; }
    }
    A.searchLevel[i] = sl;
    // WarpAffine, half_patch_size 5 (:1549-1571) + GetBilateralInterpUchar (include/ORBmatcher.h:241-252)
    {
        const float invdet = 1.0f / det2;
        const float R00 = A11 * invdet, R10 = -A10 * invdet, R01 = -A01 * invdet, R11 = A00 * invdet;
        const uint8_t *img = K.lv[level];
        const int cols = K.w[level], rows = K.h[level], step = K.pitch[level];
        const float rpx = prx / sRef, rpy = pry / sRef, sS = A.scale[sl];
        const float colsM1 = (float)(cols - 1), rowsM1 = (float)(rows - 1);
        for (int y = 0; y < 10; y++)
            for (int x = 0; x < 10; x++) {
                const float ppx = (float)(x - 5) * sS, ppy = (float)(y - 5) * sS;
                const float px = (R00 * ppx + R01 * ppy) + rpx, py = (R10 * ppx + R11 * ppy) + rpy;
                uint8_t out = 0;
                if (!(px < 0 || py < 0 || px >= colsM1 || py >= rowsM1)) {   // NaN falls through to the interpolation like the reference
                    const double X = px, Y = py;
                    const double xx = X - floor(X), yy = Y - floor(Y);
                    const uint8_t *d = img + (long long)(int)Y * step + (int)X;
                    out = (uint8_t)(int)((1 - xx) * (1 - yy) * d[0] + xx * (1 - yy) * d[1] + (1 - xx) * yy * d[step] + xx * yy * d[step + 1]);
                }
                sPatch[(y * 10 + x) * kDirectThreads + lane] = out;
            }
    }
    // ygz::Align2D (src/Align.cc:8-105) on level sl of the current frame, 10 iterations
    float H[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
            const int c = (y + 1) * 10 + x + 1;
            float J[3];
            J[0] = (float)(0.5 * ((int)sPatch[(c + 1) * kDirectThreads + lane] - (int)sPatch[(c - 1) * kDirectThreads + lane]));
            J[1] = (float)(0.5 * ((int)sPatch[(c + 10) * kDirectThreads + lane] - (int)sPatch[(c - 10) * kDirectThreads + lane]));
            J[2] = 1.0f;
            sDx[(y * 8 + x) * kDirectThreads + lane] = J[0];
            sDy[(y * 8 + x) * kDirectThreads + lane] = J[1];
#pragma unroll
            for (int r = 0; r < 3; r++)
#pragma unroll
                for (int c2 = 0; c2 < 3; c2++) H[r][c2] += J[r] * J[c2];
        }
    float Hinv[3][3];
    {
#define YGZ_COF(i, j) (H[((i) + 1) % 3][((j) + 1) % 3] * H[((i) + 2) % 3][((j) + 2) % 3] - H[((i) + 1) % 3][((j) + 2) % 3] * H[((i) + 2) % 3][((j) + 1) % 3])
        const float c0 = YGZ_COF(0, 0), c1 = YGZ_COF(1, 0), c2 = YGZ_COF(2, 0);
        const float det = dSum3(c0 * H[0][0], c1 * H[1][0], c2 * H[2][0]);
        const float invdet = 1.0f / det;
        Hinv[0][0] = c0 * invdet; Hinv[0][1] = c1 * invdet; Hinv[0][2] = c2 * invdet;
        Hinv[1][0] = YGZ_COF(0, 1) * invdet; Hinv[1][1] = YGZ_COF(1, 1) * invdet; Hinv[1][2] = YGZ_COF(2, 1) * invdet;
        Hinv[2][0] = YGZ_COF(0, 2) * invdet; Hinv[2][1] = YGZ_COF(1, 2) * invdet; Hinv[2][2] = YGZ_COF(2, 2) * invdet;
#undef YGZ_COF
    }
    const uint8_t *cur = A.curLv[sl];
    const int curCols = A.curW[sl], curRows = A.curH[sl], curStep = A.curPitch[sl];
    float u = A.pxCurr[2 * i] * A.invScale[sl], v = A.pxCurr[2 * i + 1] * A.invScale[sl];
    const float u0 = u, v0 = v;
    float meanDiff = 0;
    const float minUpdateSquared = (float)(0.03 * 0.03);
    bool converged = false, nanExit = false;
    for (int iter = 0; iter < 10; ++iter) {
        const int u_r = (int)floor((double)u), v_r = (int)floor((double)v);
        if (u_r < 4 || v_r < 4 || u_r >= curCols - 4 || v_r >= curRows - 4) break;
        if (isnan(u) || isnan(v)) { nanExit = true; break; }
        const float sx = u - (float)u_r, sy = v - (float)v_r;
        const float wTL = (float)((1.0 - sx) * (1.0 - sy)), wTR = (float)(sx * (1.0 - sy));
        const float wBL = (float)((1.0 - sx) * sy), wBR = sx * sy;
        float J0 = 0, J1 = 0, J2 = 0;
        for (int y = 0; y < 8; ++y) {
            const uint8_t *it = cur + (long long)(v_r + y - 4) * curStep + u_r - 4;
            float top[9], bot[9];
#pragma unroll
            for (int x = 0; x < 9; x++) { top[x] = (float)it[x]; bot[x] = (float)it[curStep + x]; }
#pragma unroll
            for (int x = 0; x < 8; ++x) {
                const int k = y * 8 + x;
                const float searchPixel = ((wTL * top[x] + wTR * top[x + 1]) + wBL * bot[x]) + wBR * bot[x + 1];
                const float res = (searchPixel - (float)sPatch[((y + 1) * 10 + x + 1) * kDirectThreads + lane]) + meanDiff;
                J0 -= res * sDx[k * kDirectThreads + lane];
                J1 -= res * sDy[k * kDirectThreads + lane];
                J2 -= res;
            }
        }
        const float up0 = dSum3(Hinv[0][0] * J0, Hinv[0][1] * J1, Hinv[0][2] * J2);
        const float up1 = dSum3(Hinv[1][0] * J0, Hinv[1][1] * J1, Hinv[1][2] * J2);
        const float up2 = dSum3(Hinv[2][0] * J0, Hinv[2][1] * J1, Hinv[2][2] * J2);
        u += up0; v += up1; meanDiff += up2;
        if (up0 * up0 + up1 * up1 < minUpdateSquared) { converged = true; break; }
    }
    if (nanExit) { u = u0; v = v0; }   // `return false` before cur_px_estimate is written (src/Align.cc:64-66)
    A.pxCurr[2 * i] = u * A.scale[sl]; A.pxCurr[2 * i + 1] = v * A.scale[sl];
    A.success[i] = converged ? 1 : 0;
}

int launchFindDirectProjection(const DirectArgs &A, cudaStream_t s) {
    if (A.n <= 0) return 0;
    k_find_direct_projection<<<(A.n + kDirectThreads - 1) / kDirectThreads, kDirectThreads, 0, s>>>(A);
    return 1;
}

}  // namespace ygz
