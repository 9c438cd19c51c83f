// match_oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's Hamming matching
// paths on flat arrays (part 2 of the oracle; part 1 is orb_oracle.cpp).  Linked into
// oracle/libygz_oracle.so; used by tests/, smoke() and bench.py's CPU legs — never by the product.
//
// PINNED to the reference's own code: src/ORBmatcher.cc (whole, unmodified) and the hot-path member functions of Frame /
// KeyFrame / MapPoint are compiled into oracle/_ref/libygz_ref_match.so (oracle/Makefile target refmatch,
// oracle/ref_match_api.cpp); tests/test_ref_match.py asserts every function here equal to that library live and to the golden
// vectors it produced (tests/golden/match_golden.npz).  What stays a restatement is the order of float operations inside an
// Eigen / cv::Mat expression (a fixed-size stand-in replaces Eigen there, oracle/ref_shim/eigen_lite.h), written out the same
// way on both sides.  Independent brute-force re-derivations: tests/test_match_oracle.py.  Citations are file:line in
// /root/reference.
#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <utility>
#include <vector>

#define YGZO_API extern "C" __attribute__((visibility("default")))

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;   // src/ORBmatcher.cc:36-38
const int GRID_COLS = 64, GRID_ROWS = 48;                  // include/Frame.h:27-28

struct KeyPoint { float x, y, size, angle, response; int octave, class_id; };  // cv::KeyPoint, 28 B
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

template <class F> void parallelFor(long n, int threads, long chunk, F body) {
    if (threads <= 1 || n <= chunk) { body(0, n); return; }
    std::atomic<long> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++)
        th.emplace_back([&]() { for (;;) { const long b = next.fetch_add(chunk); if (b >= n) break; body(b, std::min(n, b + chunk)); } });
    for (auto &t : th) t.join();
}

// ORBmatcher::DescriptorDistance — src/ORBmatcher.cc:1507-1523 (the __SSE2__ branch: popcount per 32-bit word).
inline int descriptorDistance(const uint8_t *a, const uint8_t *b) {
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4);
        dist += __builtin_popcount(x ^ y);
    }
    return dist;
}

// Frame::PosInGrid + AssignFeaturesToGrid — src/Frame.cc:314-330, 483-493.  round() = half away from zero.
struct Grid {
    std::vector<int> cell[GRID_COLS][GRID_ROWS];
    float minX, minY, invW, invH;
};
void assignGrid(Grid &g, const KeyPoint *kps, int n, float minX, float minY, float invW, float invH) {
    g.minX = minX; g.minY = minY; g.invW = invW; g.invH = invH;
    for (int i = 0; i < n; i++) {
        const int px = (int)std::round((kps[i].x - minX) * invW);
        const int py = (int)std::round((kps[i].y - minY) * invH);
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
        g.cell[px][py].push_back(i);
    }
}

// Frame::GetFeaturesInArea — src/Frame.cc:424-481.
void featuresInArea(const Grid &g, const KeyPoint *kps, float x, float y, float r, int minLevel, int maxLevel,
                    std::vector<int> &out) {
    out.clear();
    const int nMinCellX = std::max(0, (int)std::floor((x - g.minX - r) * g.invW));
    if (nMinCellX >= GRID_COLS) return;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - g.minX + r) * g.invW));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - g.minY - r) * g.invH));
    if (nMinCellY >= GRID_ROWS) return;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - g.minY + r) * g.invH));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (int idx : g.cell[ix][iy]) {
                const KeyPoint &kp = kps[idx];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float dx = kp.x - x, dy = kp.y - y;
                if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(idx);
            }
}

// ORBmatcher::ComputeThreeMaxima — src/ORBmatcher.cc:1471-1502, on bin sizes.
void threeMaxima(const int *sizes, int L, int &ind1, int &ind2, int &ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

inline int rotationBin(float a1, float a2) {   // src/ORBmatcher.cc:697-702 (same at :221-226, 1318-1323)
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * (1.0f / HISTO_LENGTH));
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

}  // namespace

// ================================================================================================
YGZO_API int ygzo_descriptor_distance(const uint8_t *a, const uint8_t *b) { return descriptorDistance(a, b); }

// Brute-force kNN (k = 2) + ratio test, BASELINE.json config 5: the best / second-best rule of
// ORBmatcher::SearchByBoW (src/ORBmatcher.cc:193-217) over ALL train rows in index order.
// match[q] = best index if bestDist1 <= th and (float)bestDist1 < ratio * (float)bestDist2, else -1.
YGZO_API void ygzo_hamming_knn2(const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio, int th,
                                int threads, int32_t *bestIdx, int32_t *d1, int32_t *d2, int32_t *match) {
    parallelFor(nq, threads, 64, [&](long b, long e) {
        for (long q = b; q < e; q++) {
            int bestDist1 = 256, bestDist2 = 256, best = -1;
            const uint8_t *dq = query + q * 32;
            for (int j = 0; j < nt; j++) {
                const int dist = descriptorDistance(dq, train + (size_t)j * 32);
                if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; best = j; }
                else if (dist < bestDist2) bestDist2 = dist;
            }
            bestIdx[q] = best; d1[q] = bestDist1; d2[q] = bestDist2;
            match[q] = (bestDist1 <= th && (float)bestDist1 < ratio * (float)bestDist2) ? best : -1;
        }
    });
}

// ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, checkLevel) — src/ORBmatcher.cc:44-126
// with RadiusByViewingCos (:128-133).  MapPoint state is flattened: inView = mbTrackInView && !isBad(),
// hasObs = Observations() > 0 (decides whether a claimed keypoint blocks later MapPoints, :89-91).
// claimed[idx] = the frame already holds a MapPoint with observations at idx on entry.
// kpToMp[idx] = index of the MapPoint written to F.mvpMapPoints[idx] by this call (last writer), else -1.
YGZO_API int ygzo_search_by_projection(const void *kpsV, const uint8_t *desc, const float *uRight, const uint8_t *claimed,
                                       int n, float minX, float minY, float invW, float invH, const float *scaleFactors,
                                       const float *projX, const float *projY, const float *projXR, const float *viewCos,
                                       const int32_t *level, const uint8_t *inView, const uint8_t *hasObs,
                                       const uint8_t *mpDesc, int m, float th, int checkLevel, float nnratio,
                                       int32_t *kpToMp, int32_t *mpBestIdx) {
    const KeyPoint *kps = (const KeyPoint *)kpsV;
    Grid g;
    assignGrid(g, kps, n, minX, minY, invW, invH);
    std::vector<int> owner(n, -1);               // MapPoint index written this call
    std::vector<uint8_t> blocked(n, 0);
    for (int i = 0; i < n; i++) { kpToMp[i] = -1; blocked[i] = claimed ? claimed[i] : 0; }
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < m; iMP++) {
        if (mpBestIdx) mpBestIdx[iMP] = -1;
        if (!inView[iMP]) continue;
        const int nPredictedLevel = level[iMP];
        float r = viewCos[iMP] > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        const float rad = r * scaleFactors[nPredictedLevel];
        if (checkLevel) featuresInArea(g, kps, projX[iMP], projY[iMP], rad, nPredictedLevel - 1, nPredictedLevel, vIndices);
        else featuresInArea(g, kps, projX[iMP], projY[iMP], rad, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (blocked[idx]) continue;
            if (uRight && uRight[idx] > 0) {
                const float er = std::fabs(projXR[iMP] - uRight[idx]);
                if (er > rad) continue;
            }
            const int dist = descriptorDistance(mpDesc + (size_t)iMP * 32, desc + (size_t)idx * 32);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kps[idx].octave; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = kps[idx].octave; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
            kpToMp[bestIdx] = iMP;
            blocked[bestIdx] = hasObs[iMP] ? 1 : 0;
            if (mpBestIdx) mpBestIdx[iMP] = bestIdx;
            nmatches++;
        }
    }
    return nmatches;
}

// ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono, checkLevel) —
// src/ORBmatcher.cc:1218-1350 (the motion-model tracker's matcher) with ComputeThreeMaxima (:1471-1502).
// LastFrame side flattened per keypoint i: valid = mvpMapPoints[i] != NULL && !mvbOutlier[i]; hasObs, worldPos
// (GetWorldPos), ptDesc (GetDescriptor) of that MapPoint; octaveLast / angleLast = LastFrame.mvKeys[i].
// pose[30] = Rcw (row-major 9), tcw (3), Rlw (9), tlw (3), fx, fy, cx, cy, mbf, mb.
// Float evaluation order (the reference leaves it to Eigen + the compiler; canonical here and in the kernel):
// a 3-term dot product is a0 + (a1 + a2) (Eigen >= 3.3 redux), products and sums uncontracted;
// invzc = (float)(1.0 / (double)z) (:1254); a projection that is not finite counts as out of the image.
// kpToPt[i2] = LastFrame index left in CurrentFrame.mvpMapPoints[i2] by this call, else -1.
YGZO_API int ygzo_search_by_projection_last(const void *kpsV, const uint8_t *desc, const float *uRight,
                                            const uint8_t *claimed, int n, float minX, float minY, float maxX, float maxY,
                                            float invW, float invH, const float *scaleFactors, const float *pose,
                                            const uint8_t *valid, const uint8_t *hasObs, const float *worldPos,
                                            const uint8_t *ptDesc, const int32_t *octaveLast, const float *angleLast, int m,
                                            float th, int bMono, int checkLevel, int checkOrientation, int32_t *kpToPt,
                                            int32_t *ptBestIdx) {
    const KeyPoint *kps = (const KeyPoint *)kpsV;
    Grid g;
    assignGrid(g, kps, n, minX, minY, invW, invH);
    std::vector<uint8_t> blocked(n, 0);
    for (int i = 0; i < n; i++) { kpToPt[i] = -1; blocked[i] = claimed ? claimed[i] : 0; }
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float *Rcw = pose, *tcw = pose + 9, *Rlw = pose + 12, *tlw = pose + 21;
    const float fx = pose[24], fy = pose[25], cx = pose[26], cy = pose[27], mbf = pose[28], mb = pose[29];
    auto dot3 = [](float a0, float a1, float a2) { const float s = a1 + a2; return a0 + s; };
    float twc[3], tlc[3];
    for (int r = 0; r < 3; r++) twc[r] = -dot3(Rcw[r] * tcw[0], Rcw[3 + r] * tcw[1], Rcw[6 + r] * tcw[2]);   // :1231
    for (int r = 0; r < 3; r++) tlc[r] = dot3(Rlw[3 * r] * twc[0], Rlw[3 * r + 1] * twc[1], Rlw[3 * r + 2] * twc[2]) + tlw[r];
    const bool bForward = tlc[2] > mb && !bMono, bBackward = -tlc[2] > mb && !bMono;   // :1238-1239
    std::vector<int> vIndices2;
    for (int i = 0; i < m; i++) {
        if (ptBestIdx) ptBestIdx[i] = -1;
        if (!valid[i]) continue;
        const float *X = worldPos + 3 * (size_t)i;
        float c[3];
        for (int r = 0; r < 3; r++) c[r] = dot3(Rcw[3 * r] * X[0], Rcw[3 * r + 1] * X[1], Rcw[3 * r + 2] * X[2]) + tcw[r];
        const float xc = c[0], yc = c[1];
        const float invzc = (float)(1.0 / (double)c[2]);
        if (invzc < 0) continue;
        const float u = fx * xc * invzc + cx, v = fy * yc * invzc + cy;
        if (!std::isfinite(u) || !std::isfinite(v)) continue;
        if (u < minX || u > maxX) continue;
        if (v < minY || v > maxY) continue;
        const int nLastOctave = octaveLast[i];
        const float radius = th * scaleFactors[nLastOctave];
        if (!checkLevel) featuresInArea(g, kps, u, v, radius, -1, -1, vIndices2);
        else if (bForward) featuresInArea(g, kps, u, v, radius, nLastOctave, -1, vIndices2);
        else if (bBackward) featuresInArea(g, kps, u, v, radius, 0, nLastOctave, vIndices2);
        else featuresInArea(g, kps, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (blocked[i2]) continue;
            if (uRight && uRight[i2] > 0) {
                const float ur = u - mbf * invzc;
                const float er = std::fabs(ur - uRight[i2]);
                if (er > radius) continue;
            }
            const int dist = descriptorDistance(ptDesc + (size_t)i * 32, desc + (size_t)i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            kpToPt[bestIdx2] = i;
            blocked[bestIdx2] = hasObs[i] ? 1 : 0;
            if (ptBestIdx) ptBestIdx[i] = bestIdx2;
            nmatches++;
            if (checkOrientation) rotHist[rotationBin(angleLast[i], kps[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
        threeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { kpToPt[idx] = -1; nmatches--; }
    }
    return nmatches;
}

// MapPoint::PredictScale (src/MapPoint.cc:343-373); log(float) resolves to the float overload under <cmath>.
static int predictScale(float maxDistance, float currentDist, float logScaleFactor, int nScaleLevels) {
    const float ratio = maxDistance / currentDist;
    const float q = std::ceil(std::log(ratio) / logScaleFactor);
    int nScale = (q >= -2147483648.0f && q < 2147483648.0f) ? (int)q : INT_MIN;   // what cvttss2si returns otherwise
    if (nScale < 0) nScale = 0;
    else if (nScale >= nScaleLevels) nScale = nScaleLevels - 1;
    return nScale;
}

// The projection searches over MapPoints that predict the level from the distance (MapPoint::PredictScale):
//   variant 0: SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1352-1469.
//     Eigen arithmetic, canonical order as in ygzo_search_by_projection_last.  pose = Rcw(9), tcw(3), -, fx, fy, cx, cy at 24..27.
//   variant 1: SearchByProjection(KeyFrame *pKF, cv::Mat Scw, vpPoints, vpMatched, th), :265-373.  pose = Rcw, tcw, Ow (12..14)
//     as decomposed by the caller (:275-278).
//   variant 2: the search of Fuse(KeyFrame *pKF, vpMapPoints, th), :748-886.  Eigen; pose[12..14] = Ow, pose[28] = bf; uRight =
//     mvuRight, invSigma2 = mvInvLevelSigma2.  No claims: ptBestIdx[i] is the keypoint Fuse would act on, the return value nFused.
//   variant 3: the search of Fuse(KeyFrame *pKF, cv::Mat Scw, vpPoints, th, vpReplacePoint), :888-1004.  pose as variant 1.
//   variant 4: one direction of SearchBySim3, :1043-1110 (the other, :1113-1180, is the same code with the roles swapped).
//     pose = R1w (0..8), t1w (9..11), sR21 (12..20), t21 (21..23), fx, fy, cx, cy.  Accepts bestDist <= TH_HIGH.
// cv::Mat arithmetic as OpenCV 4.13 evaluates it (tests/test_match_oracle.py checks this against cv2.gemm / cv2.norm):
// gemm 3x3 * 3x1 + c = ((a0 + a1) + a2) + c in float; cv::norm and Mat::dot accumulate in double in index order.
// blocked = CurrentFrame.mvpMapPoints[i2] / vpMatched[idx] non-null on entry (variants 0, 1).
YGZO_API int ygzo_search_by_projection_keyframe(const void *kpsV, const uint8_t *desc, const uint8_t *claimed, const float *uRight,
                                                int n, float minX, float minY, float maxX, float maxY, float invW, float invH,
                                                const float *scaleFactors, const float *invSigma2, const float *pose,
                                                const uint8_t *valid, const float *worldPos, const float *normal,
                                                const uint8_t *ptDesc, const float *angleKF, const float *maxDistance,
                                                const float *minDistance, int m, float logScaleFactor, int nScaleLevels, float th,
                                                int bestDistMax, int checkOrientation, int variant, int32_t *kpToPt,
                                                int32_t *ptBestIdx) {
    const KeyPoint *kps = (const KeyPoint *)kpsV;
    const bool eigen = variant == 0 || variant == 2, claims = variant <= 1;
    Grid g;
    assignGrid(g, kps, n, minX, minY, invW, invH);
    std::vector<uint8_t> blocked(n, 0);
    for (int i = 0; i < n; i++) { kpToPt[i] = -1; blocked[i] = (claims && claimed) ? claimed[i] : 0; }
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float *Rcw = pose, *tcw = pose + 9;
    const float fx = pose[24], fy = pose[25], cx = pose[26], cy = pose[27], bf = pose[28];
    auto dot3 = [](float a0, float a1, float a2) { const float s = a1 + a2; return a0 + s; };
    auto gemm3 = [](const float *R, const float *x, const float *t, float *out) {
        for (int r = 0; r < 3; r++) {
            float t0 = R[3 * r] * x[0]; t0 = t0 + R[3 * r + 1] * x[1]; t0 = t0 + R[3 * r + 2] * x[2];
            out[r] = t0 + t[r];
        }
    };
    float Ow[3];
    for (int r = 0; r < 3; r++) Ow[r] = variant == 0 ? -dot3(Rcw[r] * tcw[0], Rcw[3 + r] * tcw[1], Rcw[6 + r] * tcw[2]) : pose[12 + r];   // :1358
    std::vector<int> vIndices;
    for (int i = 0; i < m; i++) {
        if (ptBestIdx) ptBestIdx[i] = -1;
        if (!valid[i]) continue;
        const float *X = worldPos + 3 * (size_t)i;
        float c[3], u, v, invz;
        if (eigen) for (int r = 0; r < 3; r++) c[r] = dot3(Rcw[3 * r] * X[0], Rcw[3 * r + 1] * X[1], Rcw[3 * r + 2] * X[2]) + tcw[r];
        else gemm3(Rcw, X, tcw, c);
        if (variant == 4) { float c2[3]; gemm3(pose + 12, c, pose + 21, c2); c[0] = c2[0]; c[1] = c2[1]; c[2] = c2[2]; }
        if (variant == 0) {
            invz = (float)(1.0 / (double)c[2]);
            u = fx * c[0] * invz + cx; v = fy * c[1] * invz + cy;
            if (!std::isfinite(u) || !std::isfinite(v)) continue;
            if (u < minX || u > maxX) continue;
            if (v < minY || v > maxY) continue;
        } else {
            if (c[2] < 0.0f) continue;
            invz = variant == 1 || variant == 2 ? 1 / c[2] : (float)(1.0 / c[2]);
            const float x = c[0] * invz, y = c[1] * invz;
            u = fx * x + cx; v = fy * y + cy;
            if (!(u >= minX && u < maxX && v >= minY && v < maxY)) continue;   // KeyFrame::IsInImage
        }
        const float ur = u - bf * invz;
        float PO[3] = {X[0] - Ow[0], X[1] - Ow[1], X[2] - Ow[2]};
        if (variant == 4) { PO[0] = c[0]; PO[1] = c[1]; PO[2] = c[2]; }   // dist3D = cv::norm(p3Dc2)
        float dist3D;
        if (eigen) dist3D = std::sqrt(dot3(PO[0] * PO[0], PO[1] * PO[1], PO[2] * PO[2]));
        else {
            double s = 0;
            for (int r = 0; r < 3; r++) s += (double)PO[r] * (double)PO[r];
            dist3D = (float)std::sqrt(s);
        }
        const float maxD = 1.2f * maxDistance[i], minD = 0.8f * minDistance[i];
        if (dist3D < minD || dist3D > maxD) continue;
        if (variant >= 1 && variant <= 3) {
            const float *Pn = normal + 3 * (size_t)i;
            double d = 0;
            if (variant == 2) d = dot3(PO[0] * Pn[0], PO[1] * Pn[1], PO[2] * Pn[2]);
            else for (int r = 0; r < 3; r++) d += (double)PO[r] * (double)Pn[r];
            if (d < 0.5 * dist3D) continue;
        }
        const int nPredictedLevel = predictScale(maxDistance[i], dist3D, logScaleFactor, nScaleLevels);
        const float radius = th * scaleFactors[nPredictedLevel];
        featuresInArea(g, kps, u, v, radius, nPredictedLevel - 1, variant == 0 ? nPredictedLevel + 1 : nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices) {
            if (blocked[i2]) continue;
            if (variant == 2) {   // :835-858
                const int kpLevel = kps[i2].octave;
                const float ex = u - kps[i2].x, ey = v - kps[i2].y;
                if (uRight && uRight[i2] >= 0) {
                    const float er = ur - uRight[i2];
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * invSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * invSigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = descriptorDistance(ptDesc + (size_t)i * 32, desc + (size_t)i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= bestDistMax) {
            kpToPt[bestIdx2] = i;
            if (claims) blocked[bestIdx2] = 1;
            if (ptBestIdx) ptBestIdx[i] = bestIdx2;
            nmatches++;
            if (checkOrientation) rotHist[rotationBin(angleKF[i], kps[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
        threeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { kpToPt[idx] = -1; nmatches--; }
    }
    return nmatches;
}

// SearchBySim3 "check agreement" (src/ORBmatcher.cc:1182-1196) over the two directions' results.
YGZO_API int ygzo_sim3_agreement(const int32_t *vnMatch1, int n1, const int32_t *vnMatch2, int n2, int32_t *matches12) {
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        matches12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && idx2 < n2 && vnMatch2[idx2] == i1) { matches12[i1] = idx2; nFound++; }
    }
    return nFound;
}

// ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) — src/ORBmatcher.cc:375-478.
// prevMatched = vbPrevMatched as (x, y) pairs, updated in place (:472-475).  Returns nmatches.
YGZO_API int ygzo_search_for_initialization(const void *kps1V, const uint8_t *desc1, int n1, const void *kps2V, const uint8_t *desc2,
                                            int n2, float minX, float minY, float invW, float invH, float *prevMatched,
                                            int windowSize, float nnratio, int checkOrientation, int32_t *vnMatches12) {
    const KeyPoint *kps1 = (const KeyPoint *)kps1V, *kps2 = (const KeyPoint *)kps2V;
    Grid g;
    assignGrid(g, kps2, n2, minX, minY, invW, invH);
    int nmatches = 0;
    for (int i = 0; i < n1; i++) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(std::max(n2, 1), INT_MAX), vnMatches21(std::max(n2, 1), -1);
    std::vector<int> vIndices2;
    for (int i1 = 0; i1 < n1; i1++) {
        const int level1 = kps1[i1].octave;
        if (level1 > 0) continue;
        featuresInArea(g, kps2, prevMatched[2 * i1], prevMatched[2 * i1 + 1], (float)windowSize, level1, level1, vIndices2);
        if (vIndices2.empty()) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = descriptorDistance(desc1 + (size_t)i1 * 32, desc2 + (size_t)i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOrientation) rotHist[rotationBin(kps1[i1].angle, kps2[bestIdx2].angle)].push_back(i1);
            }
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
        threeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (vnMatches12[i1] >= 0) { prevMatched[2 * i1] = kps2[vnMatches12[i1]].x; prevMatched[2 * i1 + 1] = kps2[vnMatches12[i1]].y; }
    return nmatches;
}

// ORBmatcher::SearchByBoW — both overloads over flattened FeatureVectors (CSR, ascending NodeId):
//   secondIsFrame = 1: SearchByBoW(KeyFrame *pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches), src/ORBmatcher.cc:155-263.
//     hasMP1 = the keyframe's MapPoint exists and !isBad(); frame features already matched in this call are skipped;
//     accept bestDist1 <= TH_LOW and ratio; matches[iF] (length n2) = keyframe index whose MapPoint the frame feature got.
//   secondIsFrame = 0: SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12), :480-596.
//     hasMP = MapPoint exists and !isBad() on both sides; vbMatched2 blocks reuse; accept bestDist1 < TH_LOW (strict) and
//     ratio; matches[idx1] (length n1) = matched keypoint of KF2.
// Returns nmatches (after the rotation-histogram rejection).
YGZO_API int ygzo_search_by_bow(const void *kps1V, const uint8_t *desc1, const uint8_t *hasMP1, int n1, const uint32_t *nodeIds1,
                                const int32_t *start1, const uint32_t *feats1, int nNodes1, const void *kps2V,
                                const uint8_t *desc2, const uint8_t *hasMP2, int n2, const uint32_t *nodeIds2,
                                const int32_t *start2, const uint32_t *feats2, int nNodes2, int secondIsFrame, float nnratio,
                                int checkOrientation, int32_t *matches) {
    const KeyPoint *kps1 = (const KeyPoint *)kps1V, *kps2 = (const KeyPoint *)kps2V;
    const int nOut = secondIsFrame ? n2 : n1;
    for (int i = 0; i < nOut; i++) matches[i] = -1;
    std::vector<uint8_t> taken2(std::max(n2, 1), 0);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0, a = 0, b = 0;
    while (a < nNodes1 && b < nNodes2) {
        if (nodeIds1[a] == nodeIds2[b]) {
            for (int i1 = start1[a]; i1 < start1[a + 1]; i1++) {
                const int idx1 = (int)feats1[i1];
                if (!hasMP1[idx1]) continue;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int i2 = start2[b]; i2 < start2[b + 1]; i2++) {
                    const int idx2 = (int)feats2[i2];
                    if (taken2[idx2]) continue;
                    if (!secondIsFrame && !hasMP2[idx2]) continue;
                    const int dist = descriptorDistance(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                const bool near = secondIsFrame ? bestDist1 <= TH_LOW : bestDist1 < TH_LOW;
                if (near && (float)bestDist1 < nnratio * (float)bestDist2) {
                    taken2[bestIdx2] = 1;
                    const int slot = secondIsFrame ? bestIdx2 : idx1;
                    matches[slot] = secondIsFrame ? idx1 : bestIdx2;
                    if (checkOrientation) rotHist[rotationBin(kps1[idx1].angle, kps2[bestIdx2].angle)].push_back(slot);
                    nmatches++;
                }
            }
            a++; b++;
        } else if (nodeIds1[a] < nodeIds2[b]) {
            a = (int)(std::lower_bound(nodeIds1, nodeIds1 + nNodes1, nodeIds2[b]) - nodeIds1);
        } else {
            b = (int)(std::lower_bound(nodeIds2, nodeIds2 + nNodes2, nodeIds1[a]) - nodeIds2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
        threeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int slot : rotHist[i]) { matches[slot] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ORBmatcher::SearchForTriangulation — src/ORBmatcher.cc:597-746 with CheckDistEpipolarLine (:136-153).
// The two DBoW2::FeatureVectors (std::map<NodeId, vector<unsigned>>, ascending NodeId) are flattened
// to CSR: nodeIds[k] ascending, features of node k = feats[start[k] .. start[k+1]).  hasMP = the
// keyframe already holds a MapPoint at that keypoint (GetMapPoint != NULL).  (ex, ey) = epipole of
// KF1's centre in KF2 (:604-610, computed by the caller).  F12 row-major: F12[3*r + c] = F12(r, c).
// matches12[idx1] = idx2 or -1 (vMatches12); returns nmatches.  vbMatched2 is never set by the
// reference (:616, 660), so keypoints of KF2 can be matched more than once — reproduced.
YGZO_API int ygzo_search_for_triangulation(
    const void *kps1V, const uint8_t *desc1, const float *uRight1, const uint8_t *hasMP1, int n1,
    const uint32_t *nodeIds1, const int32_t *start1, const uint32_t *feats1, int nNodes1,
    const void *kps2V, const uint8_t *desc2, const float *uRight2, const uint8_t *hasMP2, int n2,
    const uint32_t *nodeIds2, const int32_t *start2, const uint32_t *feats2, int nNodes2,
    const float *scaleFactors2, const float *levelSigma2_2, const float *F12, float ex, float ey,
    int bOnlyStereo, int checkOrientation, int32_t *matches12) {
    const KeyPoint *kps1 = (const KeyPoint *)kps1V, *kps2 = (const KeyPoint *)kps2V;
    (void)n2;
    int nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int a = 0, b = 0;
    while (a < nNodes1 && b < nNodes2) {
        if (nodeIds1[a] == nodeIds2[b]) {
            for (int i1 = start1[a]; i1 < start1[a + 1]; i1++) {
                const int idx1 = (int)feats1[i1];
                if (hasMP1[idx1]) continue;
                const bool bStereo1 = uRight1[idx1] >= 0;
                if (bOnlyStereo && !bStereo1) continue;
                const KeyPoint &kp1 = kps1[idx1];
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int i2 = start2[b]; i2 < start2[b + 1]; i2++) {
                    const int idx2 = (int)feats2[i2];
                    if (hasMP2[idx2]) continue;
                    const bool bStereo2 = uRight2[idx2] >= 0;
                    if (bOnlyStereo && !bStereo2) continue;
                    const int dist = descriptorDistance(desc1 + (size_t)idx1 * 32, desc2 + (size_t)idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const KeyPoint &kp2 = kps2[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x, distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * scaleFactors2[kp2.octave]) continue;
                    }
                    // CheckDistEpipolarLine
                    const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
                    const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
                    const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
                    const float num = la * kp2.x + lb * kp2.y + lc;
                    const float den = la * la + lb * lb;
                    if (den == 0) continue;
                    const float dsqr = num * num / den;
                    if (dsqr < 3.84 * levelSigma2_2[kp2.octave]) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    matches12[idx1] = bestIdx2;
                    nmatches++;
                    if (checkOrientation) rotHist[rotationBin(kp1.angle, kps2[bestIdx2].angle)].push_back(idx1);
                }
            }
            a++; b++;
        } else if (nodeIds1[a] < nodeIds2[b]) {
            a = (int)(std::lower_bound(nodeIds1, nodeIds1 + nNodes1, nodeIds2[b]) - nodeIds1);
        } else {
            b = (int)(std::lower_bound(nodeIds2, nodeIds2 + nNodes2, nodeIds1[a]) - nodeIds2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1, sizes[HISTO_LENGTH];
        for (int i = 0; i < HISTO_LENGTH; i++) sizes[i] = (int)rotHist[i].size();
        threeMaxima(sizes, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i]) { matches12[idx1] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// Frame::ComputeStereoMatches — src/Frame.cc:509-682.  Pyramids: level l of the left / right
// extractor at pyrL + offL[l], continuous (step == cols == lw[l]).  Outputs mvuRight / mvDepth
// (-1 = no match) and, for tests, the integer SAD kept per accepted keypoint (bestDistOut, -1 if none).
// UB in the reference when no keypoint matches at all (:671 indexes an empty vector) -> no cut here.
YGZO_API void ygzo_stereo_match(const void *kpsLV, const uint8_t *descL, int nL, const void *kpsRV, const uint8_t *descR,
                                int nR, const uint8_t *pyrL, const uint8_t *pyrR, const int64_t *off, const int32_t *lw,
                                const int32_t *lh, const float *scaleFactors, const float *invScaleFactors,
                                float mbf, float mb, float *uRightOut, float *depthOut, int32_t *bestDistOut) {
    const KeyPoint *kpsL = (const KeyPoint *)kpsLV, *kpsR = (const KeyPoint *)kpsRV;
    for (int i = 0; i < nL; i++) { uRightOut[i] = -1.0f; depthOut[i] = -1.0f; if (bestDistOut) bestDistOut[i] = -1; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = lh[0];
    std::vector<std::vector<int>> vRowIndices(nRows);
    for (int iR = 0; iR < nR; iR++) {
        const float kpY = kpsR[iR].y;
        const float r = 2.0f * scaleFactors[kpsR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);   // out of range = UB in the reference
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < nL; iL++) {
        const KeyPoint &kpL = kpsL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        const std::vector<int> &vCandidates = vRowIndices[row];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH, bestIdxR = 0;
        for (int iR : vCandidates) {
            const KeyPoint &kpR = kpsR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptorDistance(descL + (size_t)iL * 32, descR + (size_t)iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kpsR[bestIdxR].x;
            const float scaleFactor = invScaleFactors[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5, L = 5;
            const int lvW = lw[kpL.octave];
            const uint8_t *imL = pyrL + off[kpL.octave], *imR = pyrR + off[kpL.octave];
            const int y0 = (int)(scaledvL - w), xL0 = (int)(scaleduL - w);
            const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= lvW) continue;
            const int cL = imL[(size_t)(y0 + w) * lvW + xL0 + w];
            int bestSad = INT_MAX, bestincR = 0;
            float vDists[2 * 5 + 1];
            for (int incR = -L; incR <= L; incR++) {
                const int xR0 = (int)(scaleduR0 + incR - w);
                const int cR = imR[(size_t)(y0 + w) * lvW + xR0 + w];
                double acc = 0;   // cv::norm(NORM_L1) of two CV_32F patches accumulates |a - b| in double
                for (int yy = 0; yy < 2 * w + 1; yy++)
                    for (int xx = 0; xx < 2 * w + 1; xx++) {
                        const float a = (float)imL[(size_t)(y0 + yy) * lvW + xL0 + xx] - (float)cL;
                        const float b = (float)imR[(size_t)(y0 + yy) * lvW + xR0 + xx] - (float)cR;
                        acc += std::fabs(a - b);
                    }
                const float dist = (float)acc;
                if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scaleFactors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depthOut[iL] = mbf / disparity;
                uRightOut[iL] = bestuR;
                if (bestDistOut) bestDistOut[iL] = bestSad;
                vDistIdx.push_back(std::make_pair(bestSad, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        uRightOut[vDistIdx[i].second] = -1;
        depthOut[vDistIdx[i].second] = -1;
    }
}

// Batch of frames at fixed strides (same layout as ygzorb_search_by_projection), threads over frames:
// the timed CPU port of bench.py's SearchByProjection leg.
YGZO_API void ygzo_search_by_projection_batch(int nFrames, int kpStride, int mpStride, const int32_t *nKp, const int32_t *nMp,
                                              const void *kps, const uint8_t *desc, float minX, float minY, float invW,
                                              float invH, const float *scaleFactors, const float *projX, const float *projY,
                                              const float *projXR, const float *viewCos, const int32_t *level,
                                              const uint8_t *inView, const uint8_t *hasObs, const uint8_t *mpDesc, float th,
                                              int checkLevel, float nnratio, int threads, int32_t *kpToMp, int32_t *nMatches) {
    parallelFor(nFrames, threads, 1, [&](long b, long e) {
        for (long f = b; f < e; f++) {
            const size_t ko = (size_t)f * kpStride, mo = (size_t)f * mpStride;
            nMatches[f] = ygzo_search_by_projection((const char *)kps + ko * 28, desc + ko * 32, nullptr, nullptr, nKp[f], minX, minY,
                                                    invW, invH, scaleFactors, projX + mo, projY + mo, projXR + mo, viewCos + mo,
                                                    level + mo, inView + mo, hasObs + mo, mpDesc + mo * 32, nMp[f], th, checkLevel,
                                                    nnratio, kpToMp + ko, nullptr);
        }
    });
}
