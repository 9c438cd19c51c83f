// match_kernels.cu — sm_100a kernels of the Hamming matching paths:
// ORBmatcher::DescriptorDistance, brute-force kNN-2 + ratio test, windowed SearchByProjection,
// BoW-guided SearchForTriangulation and Frame::ComputeStereoMatches.  Integer work on the XOR/POPC
// pipes; no tensor cores (no part of this path is a dense float contraction).
#include <climits>

#include "ygz_match.cuh"

namespace ygz {

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1507-1523): sum over 8 words of popc(a ^ b).

// Same sum with 5 POPCs: carry-save adders (LOP3 0x96 = a^b^c, 0xE8 = majority) fold six of the
// eight XOR words into two weight-1 and two weight-2 words, then once more, so the slow POPC pipe
// and the LOP3/IADD pipe are about equally loaded.  Bit-exact: it is the same integer sum.
__device__ __forceinline__ uint32_t lop3_xor3(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ uint32_t lop3_maj(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r; asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
template <int VARIANT>
__device__ __forceinline__ int hamming256v(const uint32_t (&q)[8], const uint4 b0, const uint4 b1) {
    const uint32_t w0 = q[0] ^ b0.x, w1 = q[1] ^ b0.y, w2 = q[2] ^ b0.z, w3 = q[3] ^ b0.w;
    const uint32_t w4 = q[4] ^ b1.x, w5 = q[5] ^ b1.y, w6 = q[6] ^ b1.z, w7 = q[7] ^ b1.w;
    if (VARIANT == 0) {        // 8 POPC
        return __popc(w0) + __popc(w1) + __popc(w2) + __popc(w3) + __popc(w4) + __popc(w5) + __popc(w6) + __popc(w7);
    } else if (VARIANT == 1) { // 6 POPC
        const uint32_t s0 = lop3_xor3(w0, w1, w2), c0 = lop3_maj(w0, w1, w2);
        const uint32_t s1 = lop3_xor3(w3, w4, w5), c1 = lop3_maj(w3, w4, w5);
        return __popc(s0) + __popc(s1) + __popc(w6) + __popc(w7) + 2 * (__popc(c0) + __popc(c1));
    } else if (VARIANT == 2) { // 5 POPC
        const uint32_t s0 = lop3_xor3(w0, w1, w2), c0 = lop3_maj(w0, w1, w2);
        const uint32_t s1 = lop3_xor3(w3, w4, w5), c1 = lop3_maj(w3, w4, w5);
        const uint32_t s2 = lop3_xor3(s0, s1, w6), c2 = lop3_maj(s0, s1, w6);
        return __popc(s2) + __popc(w7) + 2 * (__popc(c0) + __popc(c1) + __popc(c2));
    } else {                   // 4 POPC
        const uint32_t s0 = lop3_xor3(w0, w1, w2), c0 = lop3_maj(w0, w1, w2);
        const uint32_t s1 = lop3_xor3(w3, w4, w5), c1 = lop3_maj(w3, w4, w5);
        const uint32_t s2 = lop3_xor3(s0, s1, w6), c2 = lop3_maj(s0, s1, w6);
        const uint32_t t = lop3_xor3(c0, c1, c2), d = lop3_maj(c0, c1, c2);
        return __popc(s2) + __popc(w7) + 2 * __popc(t) + 4 * __popc(d);
    }
}

// -------------------------------------------------------------------------------------------------
// Register-only POPC/XOR/IADD microkernel: the measured integer-pipe peak the Hamming roofline is
// quoted against.  Each thread runs `iters` x 8 dependent-free popc(x ^ k) + add.
__global__ void k_popc_peak(unsigned *out, int iters) {
    unsigned x0 = threadIdx.x * 2654435761u, x1 = x0 ^ 0x9e3779b9u, x2 = x0 + 0x7f4a7c15u, x3 = ~x0;
    unsigned a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 8; k++) {
            a0 += __popc(x0 ^ a1); a1 += __popc(x1 ^ a2); a2 += __popc(x2 ^ a3); a3 += __popc(x3 ^ a0);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}

int launchPopcPeak(unsigned *out, int blocks, int threads, int iters, cudaStream_t s) {
    k_popc_peak<<<blocks, threads, 0, s>>>(out, iters);
    return 1;
}

// Row-wise distances of n descriptor pairs.
__global__ void k_descriptor_distance(const uint4 *a, const uint4 *b, int n, int32_t *dist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    dist[i] = hamming256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}
int launchDescriptorDistance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist, cudaStream_t s) {
    if (n <= 0) return 0;
    k_descriptor_distance<<<(n + 255) / 256, 256, 0, s>>>((const uint4 *)a, (const uint4 *)b, n, dist);
    return 1;
}

// =================================================================================================
// K7 — brute-force kNN (k = 2) + ratio test with the best / second-best rule of SearchByBoW
// (src/ORBmatcher.cc:193-217) applied to all train rows in index order:
//     d < best1 -> (best2 = best1, best1 = d, idx = j);  else d < best2 -> best2 = d.
// Each thread owns kKnnQ queries in registers and scans the train set in index order; train rows are
// staged through shared memory (every lane reads the same row: one broadcast LDS.128 pair per row).
// Inside a stage the running top-2 is kept on keys (d << 16 | row-in-stage): keys are unique, so the
// smallest key is the first row with the minimum distance and the second smallest key carries the
// second-best distance with multiplicity — exactly the sequential rule.
// gridDim.y splits the train set for small query counts; partial results are merged in split order.
constexpr int kKnnThreads = 128, kKnnQ = 4, kKnnStage = 512;   // stage = 16 KB of train rows

struct KnnPartial { int d1, d2, idx; };

template <int VARIANT>
__global__ void __launch_bounds__(kKnnThreads) k_knn2(const uint4 *__restrict__ query, int nq, const uint4 *__restrict__ train,
                                                      int nt, int rowsPerSplit, KnnPartial *partial) {
    __shared__ uint4 stage[2][kKnnStage * 2];
    const int tid = threadIdx.x;
    const int q0 = (blockIdx.x * kKnnThreads + tid) * kKnnQ;
    const int tBegin = blockIdx.y * rowsPerSplit, tEnd = min(nt, tBegin + rowsPerSplit);
    uint32_t q[kKnnQ][8];
#pragma unroll
    for (int k = 0; k < kKnnQ; k++) {
        const int qi = min(q0 + k, nq - 1);
        const uint4 a = query[2 * qi], b = query[2 * qi + 1];
        q[k][0] = a.x; q[k][1] = a.y; q[k][2] = a.z; q[k][3] = a.w; q[k][4] = b.x; q[k][5] = b.y; q[k][6] = b.z; q[k][7] = b.w;
    }
    int d1[kKnnQ], d2[kKnnQ], idx[kKnnQ];
#pragma unroll
    for (int k = 0; k < kKnnQ; k++) { d1[k] = 256; d2[k] = 256; idx[k] = -1; }
    const int nStages = (tEnd - tBegin + kKnnStage - 1) / kKnnStage;
    auto loadStage = [&](int st, int buf) {   // cp.async (LDGSTS): 16 B per request, overlapped with the scan of the other buffer
        const int base = tBegin + st * kKnnStage;
        const int rows = min(kKnnStage, tEnd - base);
        for (int i = tid; i < rows * 2; i += kKnnThreads) {
            const unsigned dst = (unsigned)__cvta_generic_to_shared(&stage[buf][i]);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(train + 2 * (size_t)base + i) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (nStages > 0) loadStage(0, 0);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    for (int st = 0; st < nStages; st++) {
        const int buf = st & 1;
        if (st + 1 < nStages) loadStage(st + 1, buf ^ 1);
        const int base = tBegin + st * kKnnStage;
        const int rows = min(kKnnStage, tEnd - base);
        uint32_t k1[kKnnQ], k2[kKnnQ];
#pragma unroll
        for (int k = 0; k < kKnnQ; k++) { k1[k] = 0xffffffffu; k2[k] = 0xffffffffu; }
#pragma unroll 2
        for (int r = 0; r < rows; r++) {
            const uint4 b0 = stage[buf][2 * r], b1 = stage[buf][2 * r + 1];
#pragma unroll
            for (int k = 0; k < kKnnQ; k++) {
                const uint32_t key = ((uint32_t)hamming256v<VARIANT>(q[k], b0, b1) << 16) | (uint32_t)r;
                k2[k] = min(k2[k], max(k1[k], key));
                k1[k] = min(k1[k], key);
            }
        }
        // fold the stage into the running result (earlier rows win ties: strict <)
#pragma unroll
        for (int k = 0; k < kKnnQ; k++) {
            if (rows > 0) {
                const int sd1 = (int)(k1[k] >> 16), sidx = base + (int)(k1[k] & 0xffffu);
                const int sd2 = rows > 1 ? (int)(k2[k] >> 16) : 256;
                if (sd1 < d1[k]) { d2[k] = min(d1[k], sd2); d1[k] = sd1; idx[k] = sidx; }
                else d2[k] = min(d2[k], sd1);
            }
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < kKnnQ; k++)
        if (q0 + k < nq) partial[(size_t)blockIdx.y * nq + q0 + k] = KnnPartial{d1[k], d2[k], idx[k]};
}

// Merge the train splits in order and apply the acceptance rule (:207-209):
// bestDist1 <= th and (float)bestDist1 < ratio * (float)bestDist2.
__global__ void k_knn2_merge(const KnnPartial *partial, int nq, int splits, float ratio, int th, int32_t *bestIdx,
                             int32_t *outD1, int32_t *outD2, int32_t *match) {
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    int d1 = 256, d2 = 256, idx = -1;
    for (int s = 0; s < splits; s++) {
        const KnnPartial p = partial[(size_t)s * nq + qi];
        if (p.d1 < d1) { d2 = min(d1, p.d2); d1 = p.d1; idx = p.idx; }
        else d2 = min(d2, p.d1);
    }
    if (bestIdx) bestIdx[qi] = idx;
    if (outD1) outD1[qi] = d1;
    if (outD2) outD2[qi] = d2;
    if (match) match[qi] = (d1 <= th && __fmul_rn(ratio, (float)d2) > (float)d1) ? idx : -1;
}

// Train splits of a kNN call.  Few query blocks leave SMs idle, so the train set is split (gridDim.y) until the grid holds at
// least 4 CTAs per SM; beyond that the split count is chosen so that the CTA count is (close to) a multiple of the SM count:
// every CTA does the same work, so an SM that is handed one CTA more than the others sets the kernel's time (768 CTAs on 148
// SMs = 5 or 6 per SM ran at 0.86 of the balanced rate).  Bounded by one stage of train rows per split, 64 splits, and
// maxPartialBytes of scratch for the per-split partial results.
int knnSplits(int nq, int nt, int smCount, size_t maxPartialBytes) {
    const int qBlocks = (nq + kKnnThreads * kKnnQ - 1) / (kKnnThreads * kKnnQ);
    const int maxByRows = max(1, (nt + kKnnStage - 1) / kKnnStage);
    long long maxByBytes = (long long)(maxPartialBytes / (sizeof(KnnPartial) * (size_t)max(nq, 1)));
    if (maxByBytes < 1) maxByBytes = 1;
    long long hiLL = min(maxByRows, 64);
    if (maxByBytes < hiLL) hiLL = maxByBytes;
    const int hi = (int)hiLL;
    const int want = 4 * smCount;
    int lo = 1;
    if (qBlocks < want) lo = min((want + qBlocks - 1) / qBlocks, hi);
    if (smCount <= 0 || qBlocks >= 32 * smCount) return lo;   // many waves: the tail is negligible
    int best = lo;
    double bestEff = 0.0;
    for (int sp = lo; sp <= hi; sp++) {
        const long long c = (long long)qBlocks * sp;
        const double eff = (double)c / (double)(((c + smCount - 1) / smCount) * smCount);
        if (eff > bestEff + 1e-9) { bestEff = eff; best = sp; }
        if (eff >= 0.995) break;
    }
    return best;
}

int launchKnn2(const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio, int th, int splits, void *partial,
               int variant, int32_t *bestIdx, int32_t *d1, int32_t *d2, int32_t *match, cudaStream_t s) {
    if (nq <= 0) return 0;
    int rowsPerSplit = (nt + splits - 1) / splits;
    rowsPerSplit = max(kKnnStage, (rowsPerSplit + kKnnStage - 1) / kKnnStage * kKnnStage);
    const dim3 grid((nq + kKnnThreads * kKnnQ - 1) / (kKnnThreads * kKnnQ), splits);
    const uint4 *q = (const uint4 *)query, *t = (const uint4 *)train;
    KnnPartial *p = (KnnPartial *)partial;
    switch (variant) {
        case 0: k_knn2<0><<<grid, kKnnThreads, 0, s>>>(q, nq, t, nt, rowsPerSplit, p); break;
        case 1: k_knn2<1><<<grid, kKnnThreads, 0, s>>>(q, nq, t, nt, rowsPerSplit, p); break;
        case 3: k_knn2<3><<<grid, kKnnThreads, 0, s>>>(q, nq, t, nt, rowsPerSplit, p); break;
        default: k_knn2<2><<<grid, kKnnThreads, 0, s>>>(q, nq, t, nt, rowsPerSplit, p); break;
    }
    k_knn2_merge<<<(nq + 255) / 256, 256, 0, s>>>(p, nq, splits, ratio, th, bestIdx, d1, d2, match);
    return 2;
}
size_t knnPartialBytes(int nq, int splits) { return (size_t)nq * splits * sizeof(KnnPartial); }


// =================================================================================================
// K8 — ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, checkLevel)
// (src/ORBmatcher.cc:44-126) with Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea
// (src/Frame.cc:314-330, 424-493).  One CTA per frame.
//
// The reference walks the MapPoints serially and a keypoint claimed by an earlier MapPoint (one with
// observations) is skipped by later ones (:89-91, 120), so results depend on the order.  The kernel
// evaluates all MapPoints in parallel against a table claimTime[kp] = smallest MapPoint index that
// claimed the keypoint, and repeats until nothing changes: MapPoint j only looks at claims of
// indices < j, so after t rounds the first t MapPoints are final and the fixed point IS the serial
// result.  Candidates are visited in the reference's order (cell column, cell row, insertion order)
// by one thread per MapPoint, which keeps the first-wins / strict-< top-2 rule literal.
// CTA size: 1024 threads (one CTA per SM at 64 registers) measured best on 512 frames x 3000 MapPoints: 0.434 ms; 768: 0.473,
// 512 (two CTAs per SM): 0.445, 384: 0.526
#ifndef YGZ_PROJ_THREADS
#define YGZ_PROJ_THREADS 1024
#endif
constexpr int kProjThreads = YGZ_PROJ_THREADS;

__device__ int projScan(int *a, int n, int *ws) {   // in-place exclusive scan by the CTA, returns the total
    const int T = kProjThreads, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int chunk = (n + T - 1) / T;
    const int b = min(tid * chunk, n), e = min(b + chunk, n);
    int sum = 0;
    for (int i = b; i < e; i++) sum += a[i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) ws[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int w = lane < T / 32 ? ws[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += v; }
        if (lane < T / 32) ws[lane] = winc - w;
        if (lane == 31) ws[32] = winc;
    }
    __syncthreads();
    int run = ws[wid] + inc - sum;
    for (int i = b; i < e; i++) { const int v = a[i]; a[i] = run; run += v; }
    const int total = ws[32];
    __syncthreads();
    return total;
}

// rotation-histogram bin of a match (src/ORBmatcher.cc:1318-1323, :697-702): round() = half away from zero
__device__ __forceinline__ int rotationBinDev(float a1, float a2) {
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / kHistoLength));
    if (bin == kHistoLength) bin = 0;
    return bin;
}

__device__ __forceinline__ bool posInGrid(const ProjArgs &A, float x, float y, int *cell) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(x, A.minX), A.invW));   // round(): half away from zero
    const int py = (int)roundf(__fmul_rn(__fsub_rn(y, A.minY), A.invH));
    if (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) return false;
    *cell = px * kGridRows + py;   // mGrid[x][y]
    return true;
}

// One MapPoint against the frame, given the claim table.  Returns the accepted keypoint or -1.
// Candidate cache: the candidates of a MapPoint (keypoints inside its window that pass the level and
// uRight filters) and their distances do not depend on the claims, so round 0 records them —
// idx | dist << 16 | octave << 25, in visit order, entry k of MapPoint j at cache[(f*kProjCache + k)*mpStride + j]
// (coalesced over j) — and later rounds only re-apply the claim table to the cached list.  MapPoints
// with more than kProjCache candidates (large th) are re-walked in every round instead.
constexpr int kProjCache = 12;
constexpr int kProjOverflow = 255;

__device__ __forceinline__ int projAccept(const ProjArgs &A, int bestDist, int bestDist2, int bestLevel, int bestLevel2, int bestIdx) {
    if (bestDist <= kThHigh) {
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(A.nnratio, (float)bestDist2)) return -1;
        return bestIdx;
    }
    return -1;
}

template <int MODE> __device__ int projMatchCached(const ProjArgs &A, int f, int j, int cnt, const int32_t *claimTime) {
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    const uint32_t *c = A.candCache + (size_t)f * kProjCache * A.mpStride + j;
    for (int k = 0; k < cnt; k++) {
        const uint32_t e = c[(size_t)k * A.mpStride];
        const int idx = e & 0xffffu;
        if (claimTime[idx] < j) continue;
        const int dist = (e >> 16) & 0x1ff, oct = (int)(e >> 25);
        if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx; }
        else if (dist < bestDist2) { bestLevel2 = oct; bestDist2 = dist; }
    }
    if (MODE == 1) return bestDist <= kThHigh ? bestIdx : -1;     // :1311 (no second-best rule in the LastFrame variant)
    if (MODE >= 2) return bestDist <= A.bestDistMax ? bestIdx : -1;   // :1430 (ORBdist), :363 (TH_LOW)
    return projAccept(A, bestDist, bestDist2, bestLevel, bestLevel2, bestIdx);
}

// The per-point query of the window search.  MODE 0: a MapPoint after isInFrustum (:61-77).  MODE 1: keypoint j of
// LastFrame projected with the current pose (:1242-1282); dir = 1 forward, 2 backward, 0 neither (:1238-1239).
struct ProjQuery { float x, y, rad, xr; int minLevel, maxLevel; };

__device__ __forceinline__ float projDot3(float a0, float a1, float a2) { return __fadd_rn(a0, __fadd_rn(a1, a2)); }

// MapPoint::PredictScale (src/MapPoint.cc:343-373): ceil(log(mfMaxDistance / dist) / mfLogScaleFactor), clamped.  log() of
// a float argument resolves to logf there; here it is the double logarithm rounded to float, i.e. the correctly rounded
// logf (the two can differ by one ulp on rare arguments, which moves the level only when the quotient sits on an integer).
__device__ __forceinline__ int predictScale(const ProjArgs &A, float maxDistance, float dist) {
    const float ratio = __fdiv_rn(maxDistance, dist);
    const float q = ceilf(__fdiv_rn(__double2float_rn(log((double)ratio)), A.logScale));
    int nScale = (q >= -2147483648.0f && q < 2147483648.0f) ? (int)q : INT_MIN;   // x86 cvttss2si on NaN / overflow
    if (nScale < 0) nScale = 0;
    else if (nScale >= A.nScaleLevels) nScale = A.nScaleLevels - 1;
    return nScale;
}

// The projection variants over MapPoints with a predicted scale level.
//   MODE 2: SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist) (:1352-1469), Eigen
//   MODE 3: SearchByProjection(KeyFrame *pKF, Scw, vpPoints, vpMatched, th) (:265-373), cv::Mat
//   MODE 4: Fuse(KeyFrame *pKF, vpMapPoints, th) (:748-886), Eigen     MODE 5: Fuse(pKF, Scw, vpPoints, th, ...) (:888-1004), cv::Mat
//   MODE 6: one direction of SearchBySim3 (:1043-1110 / :1113-1180), cv::Mat, two chained transforms
// Eigen arithmetic: 3-term sums as a0 + (a1 + a2).  cv::Mat arithmetic (OpenCV 4.13): gemm = ((a0 + a1) + a2) + c in
// float, cv::norm and Mat::dot accumulate in double.
template <int MODE> __device__ __forceinline__ bool projQueryKeyFramePoint(const ProjArgs &A, const float *P, size_t mo, float X0,
                                                                            float X1, float X2, ProjQuery &q) {
    constexpr bool kEigen = MODE == 2 || MODE == 4;
    float c[3], Ow[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float a0 = __fmul_rn(P[3 * r], X0), a1 = __fmul_rn(P[3 * r + 1], X1), a2 = __fmul_rn(P[3 * r + 2], X2);
        c[r] = kEigen ? __fadd_rn(projDot3(a0, a1, a2), P[9 + r]) : __fadd_rn(__fadd_rn(__fadd_rn(a0, a1), a2), P[9 + r]);
        Ow[r] = MODE == 2 ? -projDot3(__fmul_rn(P[r], P[9]), __fmul_rn(P[3 + r], P[10]), __fmul_rn(P[6 + r], P[11])) : P[12 + r];
    }
    if (MODE == 6) {   // p3Dc2 = sR21 * p3Dc1 + t21
        float d[3];
#pragma unroll
        for (int r = 0; r < 3; r++)
            d[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(P[12 + 3 * r], c[0]), __fmul_rn(P[13 + 3 * r], c[1])), __fmul_rn(P[14 + 3 * r], c[2])), P[21 + r]);
        c[0] = d[0]; c[1] = d[1]; c[2] = d[2];
    }
    float u, v, invz;
    if (MODE == 2) {
        invz = __double2float_rn(__ddiv_rn(1.0, (double)c[2]));
        u = __fadd_rn(__fmul_rn(__fmul_rn(P[24], c[0]), invz), P[26]);
        v = __fadd_rn(__fmul_rn(__fmul_rn(P[25], c[1]), invz), P[27]);
        if (!isfinite(u) || !isfinite(v)) return false;
        if (u < A.minX || u > A.maxX) return false;
        if (v < A.minY || v > A.maxY) return false;
    } else {
        if (c[2] < 0.0f) return false;
        invz = __fdiv_rn(1.0f, c[2]);   // 1 / z and (float)(1.0 / z) round alike
        u = __fadd_rn(__fmul_rn(P[24], __fmul_rn(c[0], invz)), P[26]);
        v = __fadd_rn(__fmul_rn(P[25], __fmul_rn(c[1], invz)), P[27]);
        if (!(u >= A.minX && u < A.maxX && v >= A.minY && v < A.maxY)) return false;   // KeyFrame::IsInImage
    }
    const float PO0 = MODE == 6 ? c[0] : __fsub_rn(X0, Ow[0]), PO1 = MODE == 6 ? c[1] : __fsub_rn(X1, Ow[1]),
                PO2 = MODE == 6 ? c[2] : __fsub_rn(X2, Ow[2]);
    float dist;
    if (kEigen) dist = __fsqrt_rn(projDot3(__fmul_rn(PO0, PO0), __fmul_rn(PO1, PO1), __fmul_rn(PO2, PO2)));
    else {
        const double s = __dadd_rn(__dadd_rn(__dmul_rn(PO0, PO0), __dmul_rn(PO1, PO1)), __dmul_rn(PO2, PO2));
        dist = __double2float_rn(__dsqrt_rn(s));
    }
    const float maxD = A.maxDist[mo];
    if (dist < __fmul_rn(0.8f, A.minDist[mo]) || dist > __fmul_rn(1.2f, maxD)) return false;
    if (MODE >= 3 && MODE <= 5) {   // viewing angle below 60 degrees (:322-325, :801-804, :951-954)
        const float *Pn = A.normal + 3 * mo;
        const double d = MODE == 4 ? (double)projDot3(__fmul_rn(PO0, Pn[0]), __fmul_rn(PO1, Pn[1]), __fmul_rn(PO2, Pn[2]))
                                   : __dadd_rn(__dadd_rn(__dmul_rn(PO0, Pn[0]), __dmul_rn(PO1, Pn[1])), __dmul_rn(PO2, Pn[2]));
        if (d < __dmul_rn(0.5, (double)dist)) return false;
    }
    const int lvl = predictScale(A, maxD, dist);
    q.rad = __fmul_rn(A.th, A.scale[lvl]);
    q.x = u; q.y = v;
    q.xr = MODE == 4 ? __fsub_rn(u, __fmul_rn(P[28], invz)) : 0.f;
    q.minLevel = lvl - 1; q.maxLevel = MODE == 2 ? lvl + 1 : lvl;
    return true;
}

template <int MODE> __device__ __forceinline__ bool projQuery(const ProjArgs &A, int f, int j, int dir, ProjQuery &q) {
    const size_t mo = (size_t)f * A.mpStride + j;
    if (!A.inView[mo]) return false;
    if (MODE == 0) {
        const int lvl = A.level[mo];
        float r = ((double)A.viewCos[mo] > 0.998) ? 2.5f : 4.0f;      // RadiusByViewingCos (:128-133)
        if (A.th != 1.0f) r = __fmul_rn(r, A.th);
        q.rad = __fmul_rn(r, A.scale[lvl]);
        q.x = A.projX[mo]; q.y = A.projY[mo];
        q.xr = A.projXR ? A.projXR[mo] : 0.f;
        q.minLevel = A.checkLevel ? lvl - 1 : -1; q.maxLevel = A.checkLevel ? lvl : -1;
        return true;
    }
    const float *P = A.pose + (size_t)f * 30;
    const float *X = A.worldPos + 3 * mo;
    const float X0 = X[0], X1 = X[1], X2 = X[2];
    if (MODE >= 2) return projQueryKeyFramePoint<MODE>(A, P, mo, X0, X1, X2, q);
    const float xc = __fadd_rn(projDot3(__fmul_rn(P[0], X0), __fmul_rn(P[1], X1), __fmul_rn(P[2], X2)), P[9]);
    const float yc = __fadd_rn(projDot3(__fmul_rn(P[3], X0), __fmul_rn(P[4], X1), __fmul_rn(P[5], X2)), P[10]);
    const float zc = __fadd_rn(projDot3(__fmul_rn(P[6], X0), __fmul_rn(P[7], X1), __fmul_rn(P[8], X2)), P[11]);
    const float invzc = __double2float_rn(__ddiv_rn(1.0, (double)zc));   // const float invzc = 1.0 / x3Dc[2] (:1254)
    if (invzc < 0) return false;
    const float u = __fadd_rn(__fmul_rn(__fmul_rn(P[24], xc), invzc), P[26]);
    const float v = __fadd_rn(__fmul_rn(__fmul_rn(P[25], yc), invzc), P[27]);
    if (!isfinite(u) || !isfinite(v)) return false;
    if (u < A.minX || u > A.maxX) return false;
    if (v < A.minY || v > A.maxY) return false;
    const int oct = A.octaveLast[mo];
    q.rad = __fmul_rn(A.th, A.scale[oct]);
    q.x = u; q.y = v;
    q.xr = __fsub_rn(u, __fmul_rn(P[28], invzc));
    if (!A.checkLevel) { q.minLevel = -1; q.maxLevel = -1; }
    else if (dir == 1) { q.minLevel = oct; q.maxLevel = -1; }
    else if (dir == 2) { q.minLevel = 0; q.maxLevel = oct; }
    else { q.minLevel = oct - 1; q.maxLevel = oct + 1; }
    return true;
}

// Frame-side data of the window walk staged in shared memory when it fits (ProjStage::xy != nullptr): keypoint position and
// (index | octave << 16) in GRID-LIST order, so a thread walking a cell range reads consecutive entries instead of chasing
// gridList[k] -> kps[idx] through L2, plus the frame's descriptors and the claim table.
struct ProjStage { const float2 *xy; const uint32_t *meta; const uint4 *desc; };

template <bool RECORD, int MODE>
__device__ int projMatchOne(const ProjArgs &A, int f, int j, int n, const int *cellStart, const int32_t *gridList,
                            const int32_t *claimTime, int *recorded, int dir, const ProjStage &S) {
    const size_t mo = (size_t)f * A.mpStride + j;
    if (RECORD) *recorded = 0;
    ProjQuery q;
    if (!projQuery<MODE>(A, f, j, dir, q)) return -1;
    const float rad = q.rad, x = q.x, y = q.y;
    const int minLevel = q.minLevel, maxLevel = q.maxLevel;
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), rad), A.invW)));
    if (cx0 >= kGridCols) return -1;
    const int cx1 = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), rad), A.invW)));
    if (cx1 < 0) return -1;
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), rad), A.invH)));
    if (cy0 >= kGridRows) return -1;
    const int cy1 = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), rad), A.invH)));
    if (cy1 < 0) return -1;
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
    const uint4 *md = reinterpret_cast<const uint4 *>(A.mpDesc + mo * 32);
    const uint4 m0 = md[0], m1 = md[1];
    const ygzorb_keypoint *kps = A.kps + (size_t)f * A.kpStride;
    const bool staged = S.xy != nullptr;
    const uint4 *desc = staged ? S.desc : reinterpret_cast<const uint4 *>(A.desc + (size_t)f * A.kpStride * 32);
    const float *uRight = A.uRight ? A.uRight + (size_t)f * A.kpStride : nullptr;
    const float projXR = q.xr;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    (void)n;
    int nrec = 0;
    uint32_t *cache = RECORD ? A.candCache + (size_t)f * kProjCache * A.mpStride + j : nullptr;
    for (int ix = cx0; ix <= cx1; ix++) {
        const int cb = ix * kGridRows;
        for (int k = cellStart[cb + cy0], ke = cellStart[cb + cy1 + 1]; k < ke; k++) {   // cells (ix, cy0..cy1) are contiguous
            int idx, oct;
            float kx, ky;
            if (staged) {
                const float2 p = S.xy[k];
                const uint32_t me = S.meta[k];
                idx = (int)(me & 0xffffu); oct = (int)(me >> 16); kx = p.x; ky = p.y;
            } else {
                idx = gridList[k];
                const ygzorb_keypoint &kp0 = kps[idx];
                oct = kp0.octave; kx = kp0.x; ky = kp0.y;
            }
            if (checkLevels) {
                if (oct < minLevel) continue;
                if (maxLevel >= 0 && oct > maxLevel) continue;
            }
            if (!(fabsf(__fsub_rn(kx, x)) < rad && fabsf(__fsub_rn(ky, y)) < rad)) continue;
            const bool skip = claimTime[idx] < j;                      // mvpMapPoints[idx]->Observations() > 0 (:89-91)
            if (!RECORD && skip) continue;
            if (MODE <= 1 && uRight) {
                const float ur = uRight[idx];
                if (ur > 0 && fabsf(__fsub_rn(projXR, ur)) > rad) continue;   // :93-97
            }
            if (MODE == 4) {   // reprojection error gate of Fuse (:835-858)
                const float ex = __fsub_rn(x, kx), ey = __fsub_rn(y, ky);
                float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                const float kpr = uRight ? uRight[idx] : -1.0f;
                double lim = 5.99;
                if (kpr >= 0) { const float er = __fsub_rn(projXR, kpr); e2 = __fadd_rn(e2, __fmul_rn(er, er)); lim = 7.8; }
                if ((double)__fmul_rn(e2, A.invSigma2[oct]) > lim) continue;
            }
            const int dist = hamming256(m0, m1, desc[2 * idx], desc[2 * idx + 1]);
            if (RECORD) {
                if (nrec < kProjCache) cache[(size_t)nrec * A.mpStride] = (uint32_t)idx | ((uint32_t)dist << 16) | ((uint32_t)oct << 25);
                nrec++;
                if (skip) continue;
            }
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = oct; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = oct; bestDist2 = dist; }
        }
    }
    if (RECORD) *recorded = nrec;
    if (MODE == 1) return bestDist <= kThHigh ? bestIdx : -1;
    if (MODE >= 2) return bestDist <= A.bestDistMax ? bestIdx : -1;
    return projAccept(A, bestDist, bestDist2, bestLevel, bestLevel2, bestIdx);
}

template <int MODE> __global__ void __launch_bounds__(kProjThreads) k_search_by_projection(const __grid_constant__ ProjArgs A) {
    __shared__ int cellStart[kGridCols * kGridRows + 1];
    __shared__ int dirS, hist[kHistoLength], keep[3];
    __shared__ int cellFill[kGridCols * kGridRows];
    __shared__ int ws[40];
    extern __shared__ __align__(16) uint8_t projSmem[];   // A.stageSmem: [desc 32 B][xy 8 B][meta 4 B][claim 4 B] x kpStride
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n = min(A.nKp[f], A.kpStride), m = min(A.nMp[f], A.mpStride);
    const ygzorb_keypoint *kps = A.kps + (size_t)f * A.kpStride;
    int32_t *gridList = A.gridList + (size_t)f * A.kpStride;
    uint4 *sDesc = reinterpret_cast<uint4 *>(projSmem);
    float2 *sXY = reinterpret_cast<float2 *>(projSmem + (size_t)A.kpStride * 32);
    uint32_t *sMeta = reinterpret_cast<uint32_t *>(projSmem + (size_t)A.kpStride * 40);
    int32_t *claimTime = A.stageSmem ? reinterpret_cast<int32_t *>(projSmem + (size_t)A.kpStride * 44) : A.claimTime + (size_t)f * A.kpStride;
    ProjStage S{nullptr, nullptr, nullptr};
    int32_t *best = A.mpBestIdx + (size_t)f * A.mpStride;
    int32_t *kpToMp = A.kpToMp + (size_t)f * A.kpStride;
    const uint8_t *claimed = A.claimed ? A.claimed + (size_t)f * A.kpStride : nullptr;
    constexpr int kCells = kGridCols * kGridRows;
    // ---- AssignFeaturesToGrid: CSR of the 64x48 grid, cell lists in ascending keypoint index
    for (int c = tid; c < kCells; c += kProjThreads) { cellStart[c] = 0; cellFill[c] = 0; }
    __syncthreads();
    for (int i = tid; i < n; i += kProjThreads) {
        int cell;
        if (posInGrid(A, kps[i].x, kps[i].y, &cell)) atomicAdd(&cellStart[cell], 1);
    }
    __syncthreads();
    const int total = projScan(cellStart, kCells, ws);
    if (tid == 0) cellStart[kCells] = total;
    __syncthreads();
    for (int i = tid; i < n; i += kProjThreads) {
        int cell;
        if (posInGrid(A, kps[i].x, kps[i].y, &cell)) gridList[cellStart[cell] + atomicAdd(&cellFill[cell], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < kCells; c += kProjThreads) {   // insertion order == index order
        const int b = cellStart[c], e = cellStart[c + 1];
        for (int i = b + 1; i < e; i++) {
            const int v = gridList[i];
            int k = i - 1;
            while (k >= b && gridList[k] > v) { gridList[k + 1] = gridList[k]; k--; }
            gridList[k + 1] = v;
        }
    }
    if (A.stageSmem) {
        __syncthreads();
        const uint4 *gdesc = reinterpret_cast<const uint4 *>(A.desc + (size_t)f * A.kpStride * 32);
        for (int i = tid; i < 2 * n; i += kProjThreads) sDesc[i] = gdesc[i];
        for (int k = tid; k < total; k += kProjThreads) {
            const int idx = gridList[k];
            const ygzorb_keypoint &kp = kps[idx];
            sXY[k] = make_float2(kp.x, kp.y);
            sMeta[k] = (uint32_t)idx | ((uint32_t)kp.octave << 16);
        }
        S.xy = sXY; S.meta = sMeta; S.desc = sDesc;
    }
    uint8_t *candCnt = A.candCnt + (size_t)f * A.mpStride;
    if (MODE == 1 && tid == 0) {   // tlc = Rlw * (-Rcw^T tcw) + tlw; forward / backward motion along the optical axis (:1227-1239)
        const float *P = A.pose + (size_t)f * 30;
        float twc[3];
#pragma unroll
        for (int r = 0; r < 3; r++) twc[r] = -projDot3(__fmul_rn(P[r], P[9]), __fmul_rn(P[3 + r], P[10]), __fmul_rn(P[6 + r], P[11]));
        const float tlcz = __fadd_rn(projDot3(__fmul_rn(P[18], twc[0]), __fmul_rn(P[19], twc[1]), __fmul_rn(P[20], twc[2])), P[23]);
        const float mb = P[29];
        dirS = A.bMono ? 0 : (tlcz > mb ? 1 : (-tlcz > mb ? 2 : 0));
    }
    if (MODE >= 1 && tid < kHistoLength) hist[tid] = 0;
    __syncthreads();
    const int dir = MODE == 1 ? dirS : 0;
    // ---- round 0: full window walk against the entry claims, recording the candidate cache
    for (int i = tid; i < n; i += kProjThreads) claimTime[i] = (claimed && claimed[i]) ? -1 : INT_MAX;
    __syncthreads();
    for (int j = tid; j < m; j += kProjThreads) {
        int nrec;
        best[j] = projMatchOne<true, MODE>(A, f, j, n, cellStart, gridList, claimTime, &nrec, dir, S);
        candCnt[j] = (uint8_t)(nrec > kProjCache ? kProjOverflow : nrec);
    }
    __syncthreads();
    // ---- fixed-point iteration over the claim table
    for (int round = 1; MODE < 4 && round <= m + 1; round++) {
        for (int i = tid; i < n; i += kProjThreads) claimTime[i] = (claimed && claimed[i]) ? -1 : INT_MAX;
        __syncthreads();
        for (int j = tid; j < m; j += kProjThreads) {
            const int b = best[j];
            if (b >= 0 && (MODE >= 2 || A.hasObs[(size_t)f * A.mpStride + j])) atomicMin(&claimTime[b], j);
        }
        __syncthreads();
        int changed = 0;
        for (int j = tid; j < m; j += kProjThreads) {
            const int cnt = candCnt[j];
            const int b = cnt == kProjOverflow ? projMatchOne<false, MODE>(A, f, j, n, cellStart, gridList, claimTime, nullptr, dir, S)
                                               : projMatchCached<MODE>(A, f, j, cnt, claimTime);
            if (b != best[j]) { best[j] = b; changed = 1; }
        }
        if (!__syncthreads_or(changed)) break;
    }
    // ---- F.mvpMapPoints[bestIdx] = pMP: the last writer stays; nmatches counts every accepted MapPoint
    for (int i = tid; i < n; i += kProjThreads) kpToMp[i] = -1;
    if (tid == 0) ws[0] = 0;
    __syncthreads();
    int cnt = 0;
    for (int j = tid; j < m; j += kProjThreads) {
        const int b = best[j];
        if (b >= 0) { atomicMax(&kpToMp[b], j); cnt++; }
    }
    if (cnt) atomicAdd(&ws[0], cnt);
    if (MODE >= 1 && A.checkOrientation) {
        // ---- rotation consistency (:1316-1346): every accepted point votes; the keypoints of all entries outside the
        // three dominant bins are cleared afterwards (also when a later point of a kept bin had overwritten them)
        const float *angleLast = A.angleLast + (size_t)f * A.mpStride;
        for (int j = tid; j < m; j += kProjThreads) {
            const int b = best[j];
            if (b >= 0) atomicAdd(&hist[rotationBinDev(angleLast[j], kps[b].angle)], 1);
        }
        __syncthreads();
        if (tid == 0) {   // ComputeThreeMaxima (:1471-1502)
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < kHistoLength; i++) {
                const int sz = hist[i];
                if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
                else if (sz > max3) { max3 = sz; ind3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncthreads();   // (also orders the kpToMp atomicMax writes above before the clears below)
        int dropped = 0;
        for (int j = tid; j < m; j += kProjThreads) {
            const int b = best[j];
            if (b < 0) continue;
            const int bin = rotationBinDev(angleLast[j], kps[b].angle);
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { kpToMp[b] = -1; dropped++; }
        }
        if (dropped) atomicSub(&ws[0], dropped);
    }
    __syncthreads();
    if (tid == 0) A.nMatches[f] = ws[0];
}

// Shared-memory staging of the frame side (48 B per keypoint slot) when it leaves room for two CTAs per SM; octaves must fit
// 16 bits next to a 16-bit keypoint index (kpStride <= 65536 is assumed by the candidate cache already).
constexpr int kProjStageMax = 96 * 1024;
static int projStageBytes(ProjArgs &A) {
    const size_t b = (size_t)A.kpStride * 48;
    A.stageSmem = (A.kpStride <= 65536 && b <= (size_t)kProjStageMax) ? 1 : 0;
    return A.stageSmem ? (int)b : 0;
}
template <int MODE> static void launchProj(ProjArgs A, cudaStream_t s) {
    static bool configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaFuncSetAttribute(k_search_by_projection<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, kProjStageMax);
        configured[dev] = true;
    }
    const int smem = projStageBytes(A);
    k_search_by_projection<MODE><<<A.nFrames, kProjThreads, smem, s>>>(A);
}

int launchSearchByProjection(const ProjArgs &A, int maxKp, cudaStream_t s) {
    (void)maxKp;
    if (A.nFrames <= 0) return 0;
    launchProj<0>(A, s);
    return 1;
}

int launchSearchByProjectionKeyFrame(const ProjArgs &A, int mode, cudaStream_t s) {
    if (A.nFrames <= 0) return 0;
    switch (mode) {
        case 2: launchProj<2>(A, s); break;
        case 3: launchProj<3>(A, s); break;
        case 4: launchProj<4>(A, s); break;
        case 5: launchProj<5>(A, s); break;
        default: launchProj<6>(A, s); break;
    }
    return 1;
}

// SearchBySim3 "check agreement" (:1182-1196): vnMatch1[i1] = idx2 stays when vnMatch2[idx2] == i1.
__global__ void k_sim3_agreement(const int32_t *match1, const int32_t *match2, const int32_t *n1, const int32_t *n2, int stride1,
                                 int stride2, int32_t *matches12, int32_t *nFound) {
    const int f = blockIdx.x;
    const int32_t *m1 = match1 + (size_t)f * stride1, *m2 = match2 + (size_t)f * stride2;
    int cnt = 0;
    for (int i1 = threadIdx.x; i1 < min(n1[f], stride1); i1 += blockDim.x) {
        const int idx2 = m1[i1];
        const bool ok = idx2 >= 0 && idx2 < min(n2[f], stride2) && m2[idx2] == i1;
        matches12[(size_t)f * stride1 + i1] = ok ? idx2 : -1;
        cnt += ok;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    __shared__ int total;
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&total, cnt);
    __syncthreads();
    if (threadIdx.x == 0) nFound[f] = total;
}

int launchSim3Agreement(const int32_t *match1, const int32_t *match2, const int32_t *n1, const int32_t *n2, int stride1, int stride2,
                        int nFrames, int32_t *matches12, int32_t *nFound, cudaStream_t s) {
    if (nFrames <= 0) return 0;
    k_sim3_agreement<<<nFrames, 256, 0, s>>>(match1, match2, n1, n2, stride1, stride2, matches12, nFound);
    return 1;
}

int launchSearchByProjectionLast(const ProjArgs &A, cudaStream_t s) {
    if (A.nFrames <= 0) return 0;
    launchProj<1>(A, s);
    return 1;
}


// =================================================================================================
// K8b — ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:375-478).  One CTA per frame pair.
//
// The serial loop lets keypoint i1 of F1 skip a candidate i2 whose current match is at least as close
// (vMatchedDistance[i2] <= dist, :412) and lets a later i1 take i2 away from an earlier one (:426-429).  As in K8 the
// kernel evaluates every i1 in parallel against the accepted matches of the previous round restricted to i1' < i1 —
// vMatchedDistance[i2] at the time of i1 is the smallest accepted distance among them, because every new match of i2 is
// strictly closer than the one before — and iterates to the fixed point, which is the serial result.  Accepted matches of
// a round are chained per i2 (head / next), so the query walks a list of typically one or two entries.
__device__ __forceinline__ bool posInGridXY(float minX, float minY, float invW, float invH, float x, float y, int *cell) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(x, minX), invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(y, minY), invH));
    if (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) return false;
    *cell = px * kGridRows + py;
    return true;
}

__global__ void __launch_bounds__(kProjThreads) k_search_for_initialization(const __grid_constant__ InitArgs A) {
    __shared__ int cellStart[kGridCols * kGridRows + 1];
    __shared__ int cellFill[kGridCols * kGridRows];
    __shared__ int ws[40], hist[kHistoLength], keep[3];
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n1 = min(A.n1[f], A.stride1), n2 = min(A.n2[f], A.stride2);
    const ygzorb_keypoint *kps1 = A.kps1 + (size_t)f * A.stride1, *kps2 = A.kps2 + (size_t)f * A.stride2;
    const uint4 *desc1 = reinterpret_cast<const uint4 *>(A.desc1 + (size_t)f * A.stride1 * 32);
    const uint4 *desc2 = reinterpret_cast<const uint4 *>(A.desc2 + (size_t)f * A.stride2 * 32);
    float *prev = A.prevMatched + (size_t)f * A.stride1 * 2;
    int32_t *gridList = A.gridList + (size_t)f * A.stride2;
    int32_t *st = A.state + (size_t)f * (5 * (size_t)A.stride1 + 2 * (size_t)A.stride2);
    int32_t *best[2] = {st, st + A.stride1}, *bdist[2] = {st + 2 * A.stride1, st + 3 * A.stride1};
    int32_t *next = st + 4 * A.stride1, *head = st + 5 * A.stride1, *owner = head + A.stride2;
    int32_t *m12 = A.matches12 + (size_t)f * A.stride1;
    constexpr int kCells = kGridCols * kGridRows;
    // ---- F2.AssignFeaturesToGrid (as in K8)
    for (int c = tid; c < kCells; c += kProjThreads) { cellStart[c] = 0; cellFill[c] = 0; }
    if (tid < kHistoLength) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n2; i += kProjThreads) {
        int cell;
        if (posInGridXY(A.minX, A.minY, A.invW, A.invH, kps2[i].x, kps2[i].y, &cell)) atomicAdd(&cellStart[cell], 1);
    }
    __syncthreads();
    const int total = projScan(cellStart, kCells, ws);
    if (tid == 0) cellStart[kCells] = total;
    __syncthreads();
    for (int i = tid; i < n2; i += kProjThreads) {
        int cell;
        if (posInGridXY(A.minX, A.minY, A.invW, A.invH, kps2[i].x, kps2[i].y, &cell)) gridList[cellStart[cell] + atomicAdd(&cellFill[cell], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < kCells; c += kProjThreads) {
        const int b = cellStart[c], e = cellStart[c + 1];
        for (int i = b + 1; i < e; i++) {
            const int v = gridList[i];
            int k = i - 1;
            while (k >= b && gridList[k] > v) { gridList[k + 1] = gridList[k]; k--; }
            gridList[k + 1] = v;
        }
    }
    for (int i = tid; i < n1; i += kProjThreads) { best[0][i] = -1; bdist[0][i] = INT_MAX; }
    __syncthreads();
    // ---- fixed-point iteration; buffer `cur` holds the previous round's matches
    int cur = 0;
    for (int round = 0; round <= n1 + 1; round++) {
        for (int i = tid; i < n2; i += kProjThreads) head[i] = -1;
        __syncthreads();
        for (int i = tid; i < n1; i += kProjThreads) {
            const int b = best[cur][i];
            if (b >= 0) next[i] = atomicExch(&head[b], i);
        }
        __syncthreads();
        int changed = 0;
        for (int i1 = tid; i1 < n1; i1 += kProjThreads) {
            int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
            if (kps1[i1].octave <= 0) {
                const float x = prev[2 * i1], y = prev[2 * i1 + 1], r = A.window;
                const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), r), A.invW)));
                const int cx1 = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), r), A.invW)));
                const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), r), A.invH)));
                const int cy1 = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), r), A.invH)));
                if (cx0 < kGridCols && cx1 >= 0 && cy0 < kGridRows && cy1 >= 0) {
                    const uint4 q0 = desc1[2 * i1], q1 = desc1[2 * i1 + 1];
                    for (int ix = cx0; ix <= cx1; ix++) {
                        const int cb = ix * kGridRows;
                        for (int k = cellStart[cb + cy0], ke = cellStart[cb + cy1 + 1]; k < ke; k++) {
                            const int i2 = gridList[k];
                            const ygzorb_keypoint &kp = kps2[i2];
                            if (kp.octave != 0) continue;                     // GetFeaturesInArea(.., level1, level1) with level1 = 0
                            if (!(fabsf(__fsub_rn(kp.x, x)) < r && fabsf(__fsub_rn(kp.y, y)) < r)) continue;
                            const int dist = hamming256(q0, q1, desc2[2 * i2], desc2[2 * i2 + 1]);
                            int matched = INT_MAX;                            // vMatchedDistance[i2] when the serial loop reaches i1
                            for (int e = head[i2]; e >= 0; e = next[e])
                                if (e < i1) matched = min(matched, bdist[cur][e]);
                            if (matched <= dist) continue;
                            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
                            else if (dist < bestDist2) bestDist2 = dist;
                        }
                    }
                }
            }
            const bool ok = bestDist <= kThLow && (float)bestDist < __fmul_rn((float)bestDist2, A.nnratio);
            const int nb = ok ? bestIdx2 : -1, nd = ok ? bestDist : INT_MAX;
            if (nb != best[cur][i1] || nd != bdist[cur][i1]) changed = 1;
            best[cur ^ 1][i1] = nb; bdist[cur ^ 1][i1] = nd;
        }
        cur ^= 1;
        if (!__syncthreads_or(changed)) break;
    }
    // ---- vnMatches12: the last taker of i2 keeps it; every accepted match votes in the rotation histogram (:436-445)
    for (int i = tid; i < n2; i += kProjThreads) owner[i] = -1;
    if (tid == 0) ws[0] = 0;
    __syncthreads();
    for (int i = tid; i < n1; i += kProjThreads) {
        const int b = best[cur][i];
        if (b >= 0) {
            atomicMax(&owner[b], i);
            if (A.checkOrientation) atomicAdd(&hist[rotationBinDev(kps1[i].angle, kps2[b].angle)], 1);
        }
    }
    __syncthreads();
    if (tid == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < kHistoLength; i++) {
            const int sz = hist[i];
            if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
            else if (sz > max3) { max3 = sz; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
        keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
    }
    __syncthreads();
    int cnt = 0;
    for (int i = tid; i < n1; i += kProjThreads) {
        const int b = best[cur][i];
        int mt = (b >= 0 && owner[b] == i) ? b : -1;
        if (mt >= 0 && A.checkOrientation) {
            const int bin = rotationBinDev(kps1[i].angle, kps2[b].angle);
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) mt = -1;
        }
        m12[i] = mt;
        if (mt >= 0) { cnt++; prev[2 * i] = kps2[mt].x; prev[2 * i + 1] = kps2[mt].y; }   // :472-475
    }
    if (cnt) atomicAdd(&ws[0], cnt);
    __syncthreads();
    if (tid == 0) A.nMatches[f] = ws[0];
}

int launchSearchForInitialization(const InitArgs &A, cudaStream_t s) {
    if (A.nFrames <= 0) return 0;
    k_search_for_initialization<<<A.nFrames, kProjThreads, 0, s>>>(A);
    return 1;
}

// K9 — ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:597-746) + CheckDistEpipolarLine
// (:136-153) + the rotation-histogram vote (:695-731, ComputeThreeMaxima :1471-1502).  One CTA:
// the call is small (~1-2k keypoints) and latency-bound.  The merge-walk over the two FeatureVectors
// pairs equal NodeIds; every keypoint of KF1 sits in exactly one node, and vbMatched2 is never set
// by the reference, so keypoints of KF1 are independent: one thread per keypoint scans its node's
// list in KF2 in order with the reference's "dist > bestDist -> continue" rule (last one wins ties).
constexpr int kTriThreads = 1024;

__global__ void __launch_bounds__(kTriThreads) k_triangulation(const __grid_constant__ TriArgs A) {
    __shared__ int hist[kHistoLength];
    __shared__ int keep[3];
    __shared__ int total;
    const int tid = threadIdx.x;
    const KfView &K1 = A.kf1, &K2 = A.kf2;
    for (int i = tid; i < K1.n; i += kTriThreads) { A.nodeOf1[i] = -1; A.matches12[i] = -1; }
    if (tid < kHistoLength) hist[tid] = 0;
    if (tid == 0) total = 0;
    __syncthreads();
    // node of KF1 -> equal node of KF2 (binary search = the lower_bound jumps of :743-745)
    for (int a = tid; a < K1.nNodes; a += kTriThreads) {
        const uint32_t id = K1.nodeIds[a];
        int lo = 0, hi = K2.nNodes;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (K2.nodeIds[mid] < id) lo = mid + 1; else hi = mid; }
        if (lo < K2.nNodes && K2.nodeIds[lo] == id)
            for (int k = K1.nodeStart[a]; k < K1.nodeStart[a + 1]; k++) A.nodeOf1[K1.nodeFeats[k]] = lo;
    }
    __syncthreads();
    const uint4 *desc1 = reinterpret_cast<const uint4 *>(K1.desc), *desc2 = reinterpret_cast<const uint4 *>(K2.desc);
    for (int idx1 = tid; idx1 < K1.n; idx1 += kTriThreads) {
        const int b = A.nodeOf1[idx1];
        if (b < 0 || K1.hasMP[idx1]) continue;
        const bool bStereo1 = K1.uRight[idx1] >= 0;
        if (A.onlyStereo && !bStereo1) continue;
        const float x1 = K1.kps[idx1].x, y1 = K1.kps[idx1].y;
        // epipolar line l = x1' F12 (:139-141)
        const float la = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F12[0]), __fmul_rn(y1, A.F12[3])), A.F12[6]);
        const float lb = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F12[1]), __fmul_rn(y1, A.F12[4])), A.F12[7]);
        const float lc = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F12[2]), __fmul_rn(y1, A.F12[5])), A.F12[8]);
        const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
        const uint4 d0 = desc1[2 * idx1], d1 = desc1[2 * idx1 + 1];
        int bestDist = kThLow, bestIdx2 = -1;
        for (int k = K2.nodeStart[b], ke = K2.nodeStart[b + 1]; k < ke; k++) {
            const int idx2 = K2.nodeFeats[k];
            if (K2.hasMP[idx2]) continue;
            const bool bStereo2 = K2.uRight[idx2] >= 0;
            if (A.onlyStereo && !bStereo2) continue;
            const int dist = hamming256(d0, d1, desc2[2 * idx2], desc2[2 * idx2 + 1]);
            if (dist > kThLow || dist > bestDist) continue;
            const float x2 = K2.kps[idx2].x, y2 = K2.kps[idx2].y;
            const int oct2 = K2.kps[idx2].octave;
            if (!bStereo1 && !bStereo2) {
                const float dx = __fsub_rn(A.ex, x2), dy = __fsub_rn(A.ey, y2);
                if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, A.scale2[oct2])) continue;
            }
            const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, x2), __fmul_rn(lb, y2)), lc);
            if (den == 0) continue;
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            if ((double)dsqr < __dmul_rn(3.84, (double)A.sigma2_2[oct2])) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
            A.matches12[idx1] = bestIdx2;
            atomicAdd(&total, 1);
            if (A.checkOrientation) atomicAdd(&hist[rotationBinDev(K1.kps[idx1].angle, K2.kps[bestIdx2].angle)], 1);
        }
    }
    __syncthreads();
    if (A.checkOrientation) {
        if (tid == 0) {   // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < kHistoLength; i++) {
                const int sz = hist[i];
                if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
                else if (sz > max3) { max3 = sz; ind3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncthreads();
        for (int idx1 = tid; idx1 < K1.n; idx1 += kTriThreads) {
            const int m = A.matches12[idx1];
            if (m < 0) continue;
            const int bin = rotationBinDev(K1.kps[idx1].angle, K2.kps[m].angle);
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { A.matches12[idx1] = -1; atomicSub(&total, 1); }
        }
        __syncthreads();
    }
    if (tid == 0) *A.nMatches = total;
}

// K9b — ORBmatcher::SearchByBoW, both overloads (src/ORBmatcher.cc:155-263 keyframe -> frame, :480-596 keyframe ->
// keyframe).  A feature sits in exactly one vocabulary node, so the "already matched" flags of side 2 only couple the
// side-1 features of the SAME node: one thread per common node runs the reference's two nested loops in order, nodes
// run in parallel.  Rotation histogram and ComputeThreeMaxima as in K9.  One CTA (the call is ~1-2k keypoints).
__global__ void __launch_bounds__(kTriThreads) k_search_by_bow(const __grid_constant__ TriArgs A) {
    __shared__ int hist[kHistoLength];
    __shared__ int keep[3];
    __shared__ int total;
    const int tid = threadIdx.x;
    const KfView &K1 = A.kf1, &K2 = A.kf2;
    const bool frame2 = A.onlyStereo != 0;          // secondIsFrame rides in the onlyStereo slot
    const int nOut = frame2 ? K2.n : K1.n;
    for (int i = tid; i < nOut; i += kTriThreads) A.matches12[i] = -1;
    for (int i = tid; i < K2.n; i += kTriThreads) A.nodeOf1[i] = 0;     // taken2 flags
    if (tid < kHistoLength) hist[tid] = 0;
    if (tid == 0) total = 0;
    __syncthreads();
    const uint4 *desc1 = reinterpret_cast<const uint4 *>(K1.desc), *desc2 = reinterpret_cast<const uint4 *>(K2.desc);
    int mine = 0;
    for (int a = tid; a < K1.nNodes; a += kTriThreads) {
        const uint32_t id = K1.nodeIds[a];
        int lo = 0, hi = K2.nNodes;
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (K2.nodeIds[mid] < id) lo = mid + 1; else hi = mid; }
        if (lo >= K2.nNodes || K2.nodeIds[lo] != id) continue;
        for (int i1 = K1.nodeStart[a]; i1 < K1.nodeStart[a + 1]; i1++) {
            const int idx1 = K1.nodeFeats[i1];
            if (!K1.hasMP[idx1]) continue;
            const uint4 d0 = desc1[2 * idx1], d1 = desc1[2 * idx1 + 1];
            int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
            for (int i2 = K2.nodeStart[lo]; i2 < K2.nodeStart[lo + 1]; i2++) {
                const int idx2 = K2.nodeFeats[i2];
                if (A.nodeOf1[idx2]) continue;
                if (!frame2 && !K2.hasMP[idx2]) continue;
                const int dist = hamming256(d0, d1, desc2[2 * idx2], desc2[2 * idx2 + 1]);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                else if (dist < bestDist2) bestDist2 = dist;
            }
            const bool near = frame2 ? bestDist1 <= kThLow : bestDist1 < kThLow;
            if (near && (float)bestDist1 < __fmul_rn(A.nnratio, (float)bestDist2)) {
                A.nodeOf1[bestIdx2] = 1;
                const int slot = frame2 ? bestIdx2 : idx1;
                A.matches12[slot] = frame2 ? idx1 : bestIdx2;
                if (A.checkOrientation) atomicAdd(&hist[rotationBinDev(K1.kps[idx1].angle, K2.kps[bestIdx2].angle)], 1);
                mine++;
            }
        }
    }
    if (mine) atomicAdd(&total, mine);
    __syncthreads();
    if (A.checkOrientation) {
        if (tid == 0) {   // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < kHistoLength; i++) {
                const int sz = hist[i];
                if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
                else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
                else if (sz > max3) { max3 = sz; ind3 = i; }
            }
            if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
            else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
            keep[0] = ind1; keep[1] = ind2; keep[2] = ind3;
        }
        __syncthreads();
        int dropped = 0;
        for (int slot = tid; slot < nOut; slot += kTriThreads) {
            const int m = A.matches12[slot];
            if (m < 0) continue;
            const int idx1 = frame2 ? m : slot, idx2 = frame2 ? slot : m;
            const int bin = rotationBinDev(K1.kps[idx1].angle, K2.kps[idx2].angle);
            if (bin != keep[0] && bin != keep[1] && bin != keep[2]) { A.matches12[slot] = -1; dropped++; }
        }
        if (dropped) atomicSub(&total, dropped);
        __syncthreads();
    }
    if (tid == 0) *A.nMatches = total;
}

int launchSearchByBoW(const TriArgs &A, cudaStream_t s) {
    k_search_by_bow<<<1, kTriThreads, 0, s>>>(A);
    return 1;
}

int launchTriangulation(const TriArgs &A, cudaStream_t s) {
    k_triangulation<<<1, kTriThreads, 0, s>>>(A);
    return 1;
}


// =================================================================================================
// K10 — Frame::ComputeStereoMatches (src/Frame.cc:509-682), three launches:
//  k_stereo_rows : the row table vRowIndices (:517-536) as CSR, lists in ascending right index;
//  k_stereo_match: per left keypoint, best Hamming candidate on its row (octave +-1, uR in
//                  [uL - maxD, uL], strict < from TH_HIGH), then the 11x11 L1 patch search over
//                  11 shifts on the keypoint's pyramid level, parabola sub-pixel fit, disparity / depth;
//  k_stereo_cut  : median of the accepted SADs, outliers at >= 1.5 * 1.4 * median invalidated (:670-681).
// The SAD is integer (centre-subtracted 8-bit patches), exactly what cv::norm(NORM_L1) returns on
// the reference's integer-valued float patches.
constexpr int kStereoThreads = 1024;

__global__ void __launch_bounds__(kStereoThreads) k_stereo_rows(const __grid_constant__ StereoArgs A) {
    __shared__ int ws[40];
    const int tid = threadIdx.x, nRows = A.lh[0];
    int *rowStart = A.rowStart;
    for (int i = tid; i <= nRows; i += kStereoThreads) rowStart[i] = 0;
    __syncthreads();
    for (int iR = tid; iR < A.nR; iR += kStereoThreads) {
        const float kpY = A.kpsR[iR].y, r = __fmul_rn(2.0f, A.scale[A.kpsR[iR].octave]);
        const int maxr = (int)ceilf(__fadd_rn(kpY, r)), minr = (int)floorf(__fsub_rn(kpY, r));
        for (int yi = max(minr, 0); yi <= min(maxr, nRows - 1); yi++) atomicAdd(&rowStart[yi], 1);
    }
    __syncthreads();
    // exclusive scan over rows (nRows <= 4095): chunked like projScan, 1024 threads
    {
        const int T = kStereoThreads, lane = tid & 31, wid = tid >> 5, n = nRows;
        const int chunk = (n + T - 1) / T;
        const int b = min(tid * chunk, n), e = min(b + chunk, n);
        int sum = 0;
        for (int i = b; i < e; i++) sum += rowStart[i];
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        if (lane == 31) ws[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const int w = ws[lane];
            int winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += v; }
            ws[lane] = winc - w;
            if (lane == 31) ws[32] = winc;
        }
        __syncthreads();
        int run = ws[wid] + inc - sum;
        for (int i = b; i < e; i++) { const int v = rowStart[i]; rowStart[i] = run; run += v; }
        __syncthreads();
        if (tid == 0) { rowStart[n] = ws[32]; if (ws[32] > A.rowListCap) atomicOr(A.status, 4); }
        __syncthreads();
    }
    if (rowStart[nRows] > A.rowListCap) return;
    // fill: rows of one right keypoint are distinct, so a per-row cursor kept in sortBuf-free scratch
    // is avoided by letting each row collect its members itself (ascending iR = push_back order)
    for (int y = tid; y < nRows; y += kStereoThreads) {
        int k = rowStart[y];
        const int ke = rowStart[y + 1];
        if (k == ke) continue;
        for (int iR = 0; iR < A.nR && k < ke; iR++) {
            const float kpY = A.kpsR[iR].y, r = __fmul_rn(2.0f, A.scale[A.kpsR[iR].octave]);
            const int maxr = (int)ceilf(__fadd_rn(kpY, r)), minr = (int)floorf(__fsub_rn(kpY, r));
            if (y >= minr && y <= maxr) A.rowList[k++] = iR;
        }
    }
}

__global__ void __launch_bounds__(128) k_stereo_match(const __grid_constant__ StereoArgs A) {
    const int iL = blockIdx.x * blockDim.x + threadIdx.x;
    if (iL >= A.nL) return;
    A.uRight[iL] = -1.0f; A.depth[iL] = -1.0f; A.bestDist[iL] = -1;
    const int nRows = A.lh[0];
    if (A.rowStart[nRows] > A.rowListCap) return;
    const ygzorb_keypoint kpL = A.kpsL[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const int row = (int)vL;
    if (row < 0 || row >= nRows) return;
    const int cb = A.rowStart[row], ce = A.rowStart[row + 1];
    if (cb == ce) return;
    const float minZ = A.mb, minD = 0.f, maxD = __fdiv_rn(A.mbf, minZ);
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;
    const uint4 *dL = reinterpret_cast<const uint4 *>(A.descL) + 2 * iL;
    const uint4 l0 = dL[0], l1 = dL[1];
    const uint4 *descR = reinterpret_cast<const uint4 *>(A.descR);
    int bestDist = kThHigh, bestIdxR = 0;
    for (int k = cb; k < ce; k++) {
        const int iR = A.rowList[k];
        const int octR = A.kpsR[iR].octave;
        if (octR < levelL - 1 || octR > levelL + 1) continue;
        const float uR = A.kpsR[iR].x;
        if (uR >= minU && uR <= maxU) {
            const int dist = hamming256(l0, l1, descR[2 * iR], descR[2 * iR + 1]);
            if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
        }
    }
    if (!(bestDist < (kThHigh + kThLow) / 2)) return;
    // sub-pixel match by correlation (:594-668)
    const float uR0 = A.kpsR[bestIdxR].x;
    const float scaleFactor = A.invScale[levelL];
    const float scaleduL = roundf(__fmul_rn(kpL.x, scaleFactor));
    const float scaledvL = roundf(__fmul_rn(kpL.y, scaleFactor));
    const float scaleduR0 = roundf(__fmul_rn(uR0, scaleFactor));
    constexpr int w = 5, L = 5;
    const float iniu = __fsub_rn(__fadd_rn(scaleduR0, (float)L), (float)w);
    const float endu = __fadd_rn(__fadd_rn(__fadd_rn(scaleduR0, (float)L), (float)w), 1.f);
    if (iniu < 0 || endu >= (float)A.lw[levelL]) return;
    const int y0 = (int)__fsub_rn(scaledvL, (float)w), xL0 = (int)__fsub_rn(scaleduL, (float)w);
    const uint8_t *imL = A.lvL[levelL] + (long long)y0 * A.pitchL[levelL] + xL0;
    const int pL = A.pitchL[levelL], pR = A.pitchR[levelL];
    const int cL = imL[w * pL + w];
    int bestSad = INT_MAX, bestincR = 0;
    float vDists[2 * L + 1];
#pragma unroll 1
    for (int incR = -L; incR <= L; incR++) {
        const int xR0 = (int)__fsub_rn(__fadd_rn(scaleduR0, (float)incR), (float)w);
        const uint8_t *imR = A.lvR[levelL] + (long long)y0 * pR + xR0;
        const int cR = imR[w * pR + w];
        const int dc = cR - cL;     // (l - cL) - (r - cR) = l - r + dc
        int acc = 0;
        for (int yy = 0; yy < 2 * w + 1; yy++) {
#pragma unroll
            for (int xx = 0; xx < 2 * w + 1; xx++) acc += abs((int)imL[yy * pL + xx] - (int)imR[yy * pR + xx] + dc);
        }
        const float dist = (float)acc;
        if (dist < (float)bestSad) { bestSad = (int)dist; bestincR = incR; }
        vDists[L + incR] = dist;
    }
    if (bestincR == -L || bestincR == L) return;
    float dist1 = 0.f, dist2 = 0.f, dist3 = 0.f;
#pragma unroll
    for (int k = 0; k < 2 * L + 1; k++) {   // register-friendly indexing of vDists[L + bestincR - 1 .. + 1]
        if (k == L + bestincR - 1) dist1 = vDists[k];
        if (k == L + bestincR) dist2 = vDists[k];
        if (k == L + bestincR + 1) dist3 = vDists[k];
    }
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3),
                                   __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2))));
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(A.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
            disparity = 0.01f;                               // `disparity = 0.01` (double literal to float)
            bestuR = (float)__dsub_rn((double)uL, 0.01);     // `bestuR = uL - 0.01` evaluated in double
        }
        A.depth[iL] = __fdiv_rn(A.mbf, disparity);
        A.uRight[iL] = bestuR;
        A.bestDist[iL] = bestSad;
    }
}

__global__ void __launch_bounds__(kStereoThreads) k_stereo_cut(const __grid_constant__ StereoArgs A) {
    __shared__ int count, median;
    const int tid = threadIdx.x;
    if (tid == 0) { count = 0; median = -1; }
    __syncthreads();
    for (int i = tid; i < A.nL; i += kStereoThreads)
        if (A.bestDist[i] >= 0) A.sortBuf[atomicAdd(&count, 1)] = i;
    __syncthreads();
    const int n = count;
    if (n == 0) return;   // the reference indexes an empty vector here (UB): nothing to cut
    // element of rank n/2 in the order of sort(vector<pair<int,int>>): (dist, iL) ascending
    for (int a = tid; a < n; a += kStereoThreads) {
        const int ia = A.sortBuf[a], da = A.bestDist[ia];
        int rank = 0;
        for (int b = 0; b < n; b++) {
            const int ib = A.sortBuf[b], db = A.bestDist[ib];
            rank += (db < da) || (db == da && ib < ia);
        }
        if (rank == n / 2) median = da;
    }
    __syncthreads();
    const float thDist = __fmul_rn(1.5f * 1.4f, (float)median);
    for (int a = tid; a < n; a += kStereoThreads) {
        const int ia = A.sortBuf[a];
        if (!((float)A.bestDist[ia] < thDist)) { A.uRight[ia] = -1.f; A.depth[ia] = -1.f; }
    }
}

int launchStereo(const StereoArgs &A, cudaStream_t s) {
    if (A.nL <= 0) return 0;
    k_stereo_rows<<<1, kStereoThreads, 0, s>>>(A);
    k_stereo_match<<<(A.nL + 127) / 128, 128, 0, s>>>(A);
    k_stereo_cut<<<1, kStereoThreads, 0, s>>>(A);
    return 3;
}

int configureMatchKernels() { return 0; }

}  // namespace ygz
