// direct_kernels.cu — SURVEY.md §8(f) row f4: the descriptor post-processing of MapPoint and the pyramid consumers of the
// direct tracker (MapPoint::ComputeDistinctiveDescriptors src/MapPoint.cc:211-271; ORBmatcher::FindDirectProjection
// src/ORBmatcher.cc:1525-1602 with ygz::Align2D src/Align.cc:8-105; SparseImgAlign src/SparseImageAlign.cc:60-256).
#include <cooperative_groups.h>

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

// ---- K15: SparseImgAlign (src/SparseImageAlign.cc) — the two hot loops as device functions, used by the step kernels
// (precomputeReferencePatches / computeResiduals behind the reference's own NLLSSolver driver) and by the fused kernel that
// runs the whole coarse-to-fine Gauss-Newton of SparseImgAlign::run in one launch on a cluster of 8 CTAs (a frame has a few
// thousand features x 16 pixels: one thread per feature, and what bounds an iteration is the scattered 5 x 5 byte gathers, so
// they are spread over 8 SMs' load/store units; the 6 x 6 solve runs in one thread of CTA 0 between two cluster barriers).
// The caches are feature-minor (element e of feature i at e * stride + i), so a warp's loads and stores of the same element
// coalesce.  Per feature the arithmetic is the reference's, in its order, in float; ACROSS features the 21 + 6 + 1 sums are
// reduced by a fixed tree in double (warp shuffles, warps in order, CTAs in rank order: deterministic; the reference adds them up
// one after the other in float), so H, Jres and chi2 agree with the sequential float sums to rounding, not bit for bit.
constexpr int kAlignThreads = 256, kAlignCluster = 8, kAlignSums = 29;   // 21 upper-triangle H + 6 Jres + chi2 + n_meas

__device__ void alignPrecompute(const SparseAlignArgs &A, const uint8_t *__restrict__ img, int cols, int rows, int stride, float scale,
                                int first, int step) {
    const size_t S = (size_t)A.stride;
    for (int i = first; i < A.n; i += step) {
        float *__restrict__ jac = A.jac + i;
        float *__restrict__ patch = A.refPatch + i;
#pragma unroll 8
        for (int k = 0; k < 96; k++) jac[k * S] = 0.0f;                // jacobian_cache_.setZero() (:42)
        if (!A.valid[i]) continue;
        const float u = A.kpXY[2 * i] * scale, v = A.kpXY[2 * i + 1] * scale;
        const int ui = (int)floorf(u), vi = (int)floorf(v);
        if (ui - 3 < 0 || vi - 3 < 0 || ui + 3 >= cols || vi + 3 >= rows) continue;
        A.visible[i] = 1;
        float px[7][7];                                                // rows vi-3 .. vi+3, columns ui-3 .. ui+3: every load in flight at once
#pragma unroll
        for (int yy = 0; yy < 7; yy++)
#pragma unroll
            for (int xx = 0; xx < 7; xx++) px[yy][xx] = (float)__ldg(img + (long long)(vi + yy - 3) * stride + (ui + xx - 3));
        const DV3 p = dSe3Mul(A.refPose, DV3{A.worldPos[3 * i], A.worldPos[3 * i + 1], A.worldPos[3 * i + 2]});
        float J[2][6];
        const float x = p.x, y = p.y, zInv = (float)(1. / (double)p.z), zInv2 = zInv * zInv;
        J[0][0] = -zInv; J[0][1] = 0.0f; J[0][2] = x * zInv2; J[0][3] = y * J[0][2]; J[0][4] = (float)(-(1.0 + (double)(x * J[0][2]))); J[0][5] = y * zInv;
        J[1][0] = 0.0f; J[1][1] = -zInv; J[1][2] = y * zInv2; J[1][3] = (float)(1.0 + (double)(y * J[1][2])); J[1][4] = -J[0][3]; J[1][5] = -x * zInv;
        const float su = u - (float)ui, sv = v - (float)vi;
        const float wTL = (float)((1.0 - su) * (1.0 - sv)), wTR = (float)(su * (1.0 - sv)), wBL = (float)((1.0 - su) * sv), wBR = su * sv;
        const float fs = A.focal * scale;
#pragma unroll
        for (int yy = 0; yy < 4; ++yy)
#pragma unroll
            for (int xx = 0; xx < 4; ++xx) {
                const int k = yy * 4 + xx, r = yy + 1, c0 = xx + 1;   // r[0] of the reference = px[r][c0]
#define YGZ_BL(dy, dx) (((wTL * px[r + (dy)][c0 + (dx)] + wTR * px[r + (dy)][c0 + (dx) + 1]) + wBL * px[r + (dy) + 1][c0 + (dx)]) + wBR * px[r + (dy) + 1][c0 + (dx) + 1])
                patch[k * S] = YGZ_BL(0, 0);
                const float dx = 0.5f * (YGZ_BL(0, 1) - YGZ_BL(0, -1));
                const float dy = 0.5f * (YGZ_BL(1, 0) - YGZ_BL(-1, 0));
#undef YGZ_BL
#pragma unroll
                for (int c = 0; c < 6; c++) jac[(k * 6 + c) * S] = (dx * J[0][c] + dy * J[1][c]) * fs;
            }
    }
}

// computeResiduals (:130-225): this thread's features -> acc[29] (float partials per feature, added up in double)
__device__ void alignResiduals(const SparseAlignArgs &A, const float *T, const uint8_t *__restrict__ cur, int cols, int rows, int stride,
                               float scale, int first, int step, double *acc) {
    const size_t S = (size_t)A.stride;
#pragma unroll
    for (int k = 0; k < kAlignSums; k++) acc[k] = 0.0;
    for (int i = first; i < A.n; i += step) {
        if (!A.visible[i]) continue;
        const DV3 pr = dSe3Mul(A.refPose, DV3{A.worldPos[3 * i], A.worldPos[3 * i + 1], A.worldPos[3 * i + 2]});
        const DV3 pc = dSe3Mul(T, pr);
        const float uc = (A.fx * pc.x / pc.z + A.cx) * scale, vc = (A.fy * pc.y / pc.z + A.cy) * scale;
        const int ui = (int)floorf(uc), vi = (int)floorf(vc);
        if (ui < 0 || vi < 0 || ui - 3 < 0 || vi - 3 < 0 || ui + 3 >= cols || vi + 3 >= rows) continue;
        float px[5][5];
#pragma unroll
        for (int yy = 0; yy < 5; yy++)
#pragma unroll
            for (int xx = 0; xx < 5; xx++) px[yy][xx] = (float)__ldg(cur + (long long)(vi + yy - 2) * stride + (ui + xx - 2));
        const float *__restrict__ jac = A.jac + i;
        const float *__restrict__ patch = A.refPatch + i;
        const float su = uc - (float)ui, sv = vc - (float)vi;
        const float wTL = (float)((1.0 - su) * (1.0 - sv)), wTR = (float)(su * (1.0 - sv)), wBL = (float)((1.0 - su) * sv), wBR = su * sv;
        float f[kAlignSums - 1];
#pragma unroll
        for (int k = 0; k < kAlignSums - 1; k++) f[k] = 0.0f;
#pragma unroll
        for (int yy = 0; yy < 4; ++yy)
#pragma unroll
            for (int xx = 0; xx < 4; ++xx) {
                const int k = yy * 4 + xx;
                const float inten = ((wTL * px[yy][xx] + wTR * px[yy][xx + 1]) + wBL * px[yy + 1][xx]) + wBR * px[yy + 1][xx + 1];
                const float res = inten - patch[k * S];
                float J[6];
#pragma unroll
                for (int c = 0; c < 6; c++) J[c] = jac[(k * 6 + c) * S];
                f[27] += res * res;
                int q = 0;
#pragma unroll
                for (int r = 0; r < 6; r++) {
                    f[21 + r] -= J[r] * res;
#pragma unroll
                    for (int c = r; c < 6; c++) f[q++] += J[r] * J[c];
                }
            }
#pragma unroll
        for (int q = 0; q < kAlignSums - 1; q++) acc[q] += (double)f[q];
        acc[28] += 16.0;
    }
}

// block-wide fixed-order tree sum of acc[29] -> sums[29] in shared memory (valid after the trailing barrier)
__device__ void alignReduce(double *acc, double *warpBuf, double *sums) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int k = 0; k < kAlignSums; k++) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) warpBuf[wid * kAlignSums + k] = v;
    }
    __syncthreads();
    if (threadIdx.x < kAlignSums) {
        double v = 0.0;
        for (int w = 0; w < nw; w++) v += warpBuf[w * kAlignSums + threadIdx.x];
        sums[threadIdx.x] = v;
    }
    __syncthreads();
}

__device__ void alignUnpack(const double *sums, float *H36, float *Jres6) {
    int q = 0;
    for (int r = 0; r < 6; r++)
        for (int c = r; c < 6; c++) { H36[r * 6 + c] = H36[c * 6 + r] = (float)sums[q++]; }
    for (int r = 0; r < 6; r++) Jres6[r] = (float)sums[21 + r];
}

__global__ void __launch_bounds__(kAlignThreads) k_sparse_align_precompute(const __grid_constant__ SparseAlignArgs A, int level) {
    alignPrecompute(A, A.refLv[level], A.w[level], A.h[level], A.refPitch[level], A.invScale[level],
                    blockIdx.x * kAlignThreads + threadIdx.x, gridDim.x * kAlignThreads);
}

// out: H (36), Jres (6), chi2 sum, n_meas
__global__ void __launch_bounds__(kAlignThreads) k_sparse_align_residuals(const __grid_constant__ SparseAlignArgs A, int level, float *out) {
    __shared__ double warpBuf[(kAlignThreads / 32) * kAlignSums], sums[kAlignSums];
    __shared__ float T[7];
    if (threadIdx.x < 7) T[threadIdx.x] = A.tCurRef[threadIdx.x];
    __syncthreads();
    double acc[kAlignSums];
    alignResiduals(A, T, A.curLv[level], A.w[level], A.h[level], A.curPitch[level], A.invScale[level], threadIdx.x, kAlignThreads, acc);
    alignReduce(acc, warpBuf, sums);
    if (threadIdx.x == 0) { alignUnpack(sums, out, out + 36); out[42] = (float)sums[27]; out[43] = (float)sums[28]; }
}

// ---- host-free Gauss-Newton driver pieces (thread 0): Sophus SE3f on (q, t) storage and H.ldlt().solve(b)
__device__ void dSe3Compose(const float *A, const float *B, float *R) {   // SE3Group::operator*= : fastMultiply + normalize
    float rot[7] = {A[0], A[1], A[2], A[3], 0.f, 0.f, 0.f};
    const DV3 rt = dSe3Mul(rot, DV3{B[4], B[5], B[6]});
    const float ax = A[0], ay = A[1], az = A[2], aw = A[3], bx = B[0], by = B[1], bz = B[2], bw = B[3];
    const float q0 = aw * bx + ax * bw + ay * bz - az * by, q1 = aw * by + ay * bw + az * bx - ax * bz;
    const float q2 = aw * bz + az * bw + ax * by - ay * bx, q3 = aw * bw - ax * bx - ay * by - az * bz;
    const float len = sqrtf((q0 * q0 + q1 * q1) + (q2 * q2 + q3 * q3));   // coeffs().norm(): one packet of four, pairwise
    R[4] = A[4] + rt.x; R[5] = A[5] + rt.y; R[6] = A[6] + rt.z;
    R[0] = q0 / len; R[1] = q1 / len; R[2] = q2 / len; R[3] = q3 / len;
}
__device__ void dSe3Exp(const float *a, float *T) {   // SE3Group::exp (se3.hpp:407-428), SO3Group::expAndTheta (so3.hpp:426-456)
    const float wx = a[3], wy = a[4], wz = a[5];
    const float thetaSq = dSum3(wx * wx, wy * wy, wz * wz), theta = sqrtf(thetaSq), half = 0.5f * theta;   // omega.squaredNorm()
    const float eps = 1e-5f;
    float imag, real;
    if (theta < eps) {
        const float po4 = thetaSq * thetaSq;
        imag = 0.5f - (float)(1.0 / 48.0) * thetaSq + (float)(1.0 / 3840.0) * po4;
        real = 1.0f - 0.5f * thetaSq + (float)(1.0 / 384.0) * po4;
    } else {
        imag = (float)sin((double)half) / theta; real = (float)cos((double)half);
    }
    T[0] = imag * wx; T[1] = imag * wy; T[2] = imag * wz; T[3] = real;
    {   // the SO3Group(Quaternion) constructor normalises (so3.hpp:739-743 -> :268-276)
        const float len = sqrtf((T[0] * T[0] + T[1] * T[1]) + (T[2] * T[2] + T[3] * T[3]));
        T[0] = T[0] / len; T[1] = T[1] / len; T[2] = T[2] / len; T[3] = T[3] / len;
    }
    const float O[3][3] = {{0, -wz, wy}, {wz, 0, -wx}, {-wy, wx, 0}};
    float V[3][3];
    if (theta < eps) {   // V = so3.matrix() = Quaternion::toRotationMatrix()
        const float x = T[0], y = T[1], z = T[2], w = T[3];
        const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
        const float twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
        V[0][0] = 1.0f - (tyy + tzz); V[0][1] = txy - twz; V[0][2] = txz + twy;
        V[1][0] = txy + twz; V[1][1] = 1.0f - (txx + tzz); V[1][2] = tyz - twx;
        V[2][0] = txz - twy; V[2][1] = tyz + twx; V[2][2] = 1.0f - (txx + tyy);
    } else {
        const float thetaSq2 = theta * theta;   // se3.hpp:422 squares the root again
        const float Ac = (1.0f - (float)cos((double)theta)) / thetaSq2, Bc = (theta - (float)sin((double)theta)) / (thetaSq2 * theta);
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) {
                const float o2 = dSum3(O[r][0] * O[0][c], O[r][1] * O[1][c], O[r][2] * O[2][c]);
                V[r][c] = ((r == c ? 1.0f : 0.0f) + Ac * O[r][c]) + Bc * o2;
            }
    }
    for (int r = 0; r < 3; r++) T[4 + r] = dSum3(V[r][0] * a[0], V[r][1] * a[1], V[r][2] * a[2]);
}
__device__ void dLdltSolve6(const float *Hin, const float *b, float *x) {   // LDL^T, largest remaining diagonal entry as pivot
    float A[6][6]; int perm[6];
    for (int r = 0; r < 6; r++) { perm[r] = r; for (int c = 0; c < 6; c++) A[r][c] = Hin[r * 6 + c]; }
    for (int k = 0; k < 6; k++) {
        int p = k;
        for (int i = k + 1; i < 6; i++) if (fabsf(A[i][i]) > fabsf(A[p][p])) p = i;
        if (p != k) {
            for (int c = 0; c < 6; c++) { const float t = A[k][c]; A[k][c] = A[p][c]; A[p][c] = t; }
            for (int r = 0; r < 6; r++) { const float t = A[r][k]; A[r][k] = A[r][p]; A[r][p] = t; }
            const int t = perm[k]; perm[k] = perm[p]; perm[p] = t;
        }
        float d = A[k][k];
        for (int j = 0; j < k; j++) d -= A[k][j] * A[k][j] * A[j][j];
        A[k][k] = d;
        for (int i = k + 1; i < 6; i++) {
            float v = A[i][k];
            for (int j = 0; j < k; j++) v -= A[i][j] * A[k][j] * A[j][j];
            A[i][k] = d != 0 ? v / d : v;
        }
    }
    float y[6];
    for (int i = 0; i < 6; i++) { float v = b[perm[i]]; for (int j = 0; j < i; j++) v -= A[i][j] * y[j]; y[i] = v; }
    for (int i = 0; i < 6; i++) y[i] = fabsf(A[i][i]) > 1.17549435e-38f ? y[i] / A[i][i] : 0.0f;
    for (int i = 5; i >= 0; i--) { float v = y[i]; for (int j = i + 1; j < 6; j++) v -= A[j][i] * y[j]; y[i] = v; }
    for (int i = 0; i < 6; i++) x[perm[i]] = y[i];
}

// SparseImgAlign::run (:20-55) + NLLSSolver::optimizeGaussNewton (include/NLSSolver_impl.hpp:17-92) in one launch on one
// cluster.  out: TCR (7), n_meas (as float), iterations, final chi2_
__global__ void __cluster_dims__(kAlignCluster, 1, 1) __launch_bounds__(kAlignThreads)
k_sparse_align_run(const __grid_constant__ SparseAlignArgs A, int maxLevel, int minLevel, float *out) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ double warpBuf[(kAlignThreads / 32) * kAlignSums], sums[kAlignSums];
    __shared__ float ctl[16];          // CTA 0's copy is the master: [0..6] model, [7] break flag; every CTA mirrors it after the barrier
    __shared__ float oldModel[7];
    __shared__ float chi2S, nMeasS, itersS;
    __shared__ int stopS;
    const int tid = threadIdx.x, rank = (int)cluster.block_rank();
    const int first = rank * kAlignThreads + tid, step = kAlignCluster * kAlignThreads;
    for (int i = first; i < A.n; i += step) A.visible[i] = 0;               // visible_fts_ = vector<bool>(N, false) (:34)
    if (tid < 7) ctl[tid] = A.tCurRef[tid];
    if (tid == 0) { chi2S = 1e10f; nMeasS = 0.f; itersS = 0.f; stopS = 0; ctl[7] = 0.f; }
    __syncthreads();
    const double *remoteSums[kAlignCluster];
    for (int r = 0; r < kAlignCluster; r++) remoteSums[r] = cluster.map_shared_rank(sums, r);
    const float *masterCtl = cluster.map_shared_rank(ctl, 0);
    for (int level = maxLevel; level >= minLevel; --level) {
        alignPrecompute(A, A.refLv[level], A.w[level], A.h[level], A.refPitch[level], A.invScale[level], first, step);
        if (rank == 0 && tid < 7) oldModel[tid] = ctl[tid];
        __syncthreads();                                                    // a thread only reads back its own features' caches
        for (int iter = 0; iter < 10; ++iter) {
            double acc[kAlignSums];
            alignResiduals(A, ctl, A.curLv[level], A.w[level], A.h[level], A.curPitch[level], A.invScale[level], first, step, acc);
            alignReduce(acc, warpBuf, sums);
            cluster.sync();
            if (rank == 0 && tid == 0) {
                double tot[kAlignSums];
                for (int k = 0; k < kAlignSums; k++) {
                    double v = 0.0;
                    for (int r = 0; r < kAlignCluster; r++) v += remoteSums[r][k];
                    tot[k] = v;
                }
                float H[36], Jres[6], x[6];
                alignUnpack(tot, H, Jres);
                const float nMeas = (float)tot[28];
                const float newChi2 = (float)tot[27] / nMeas;
                nMeasS = nMeas; itersS += 1.0f;
                dLdltSolve6(H, Jres, x);
                if (isnan(x[0])) stopS = 1;
                float brk = 0.f;
                if ((iter > 0 && (double)newChi2 > 1.2 * (double)chi2S) || stopS) {
                    for (int k = 0; k < 7; k++) ctl[k] = oldModel[k];       // rollback
                    brk = 1.f;
                } else {
                    float mx[6], E[7], N[7];
                    for (int k = 0; k < 6; k++) mx[k] = -x[k];
                    dSe3Exp(mx, E);
                    dSe3Compose(ctl, E, N);
                    for (int k = 0; k < 7; k++) { oldModel[k] = ctl[k]; ctl[k] = N[k]; }
                    chi2S = newChi2;
                    float nm = -1.0f;
                    for (int k = 0; k < 6; k++) nm = fmaxf(nm, fabsf(x[k]));
                    if (nm <= 0.000001f) brk = 1.f;
                }
                ctl[7] = brk;
            }
            cluster.sync();
            float mine = 0.f;
            if (tid < 8) mine = masterCtl[tid];
            __syncthreads();
            if (rank != 0 && tid < 8) ctl[tid] = mine;
            __syncthreads();
            const bool brk = ctl[7] != 0.f;
            cluster.sync();                                                 // CTA 0 may overwrite ctl / sums only after everyone has read them
            if (brk) break;
        }
    }
    if (rank == 0) {
        if (tid < 7) out[tid] = ctl[tid];
        if (tid == 0) { out[7] = nMeasS; out[8] = itersS; out[9] = chi2S; }
    }
}

int launchSparseAlignPrecompute(const SparseAlignArgs &A, int level, cudaStream_t s) {
    k_sparse_align_precompute<<<(A.n + kAlignThreads - 1) / kAlignThreads, kAlignThreads, 0, s>>>(A, level);
    return 1;
}
int launchSparseAlignResiduals(const SparseAlignArgs &A, int level, float *out, cudaStream_t s) {
    k_sparse_align_residuals<<<1, kAlignThreads, 0, s>>>(A, level, out);
    return 1;
}
int launchSparseAlignRun(const SparseAlignArgs &A, int maxLevel, int minLevel, float *out, cudaStream_t s) {
    k_sparse_align_run<<<kAlignCluster, kAlignThreads, 0, s>>>(A, maxLevel, minLevel, out);
    return 1;
}

}  // namespace ygz
