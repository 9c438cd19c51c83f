// extractor.cu — host side of the extractor: the constructor's tables (src/ORBextractor.cc:412-470),
// per-image geometry, workspace, stage orchestration on the handle's stream, and the C ABI of
// include/ygzorb.h.  No CPU compute path: everything below either sizes buffers / fills lookup
// tables or enqueues kernels.
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#include "ygz_common.cuh"
#include "ygz_dso.cuh"

namespace ygz {

static thread_local std::string gLastError;
void setError(const std::string &msg) { gLastError = msg; }
int cudaFail(cudaError_t e, const char *what, const char *file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    gLastError = buf;
    return YGZORB_E_CUDA;
}
static int argFail(const char *msg) { gLastError = msg; return YGZORB_E_ARG; }

// OpenCV's cvRound / cvFloor / cvCeil on the host (round-half-to-even under the default FP mode).
static inline int cvRoundF(float v) { return (int)lrintf(v); }
static inline int cvRoundD(double v) { return (int)lrint(v); }
static inline int cvFloorF(float v) { return (int)floorf(v); }
static inline int alignUp(int v, int a) { return (v + a - 1) / a * a; }
static inline long long alignUpLL(long long v, long long a) { return (v + a - 1) / a * a; }

struct Tables {
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> nPerLevel;
    int umax[16];
};

// src/ORBextractor.cc:412-470.  `scaleFactor` is a double member initialised from the float
// argument (include/ORBextractor.h:157), so the products are taken in double and rounded to float.
static Tables makeTables(const ygzorb_params &p) {
    Tables t;
    const int L = p.nlevels;
    const double sf = (double)p.scaleFactor;
    t.scale.assign(L, 1.f); t.sigma2.assign(L, 1.f); t.invScale.resize(L); t.invSigma2.resize(L);
    for (int i = 1; i < L; i++) {
        t.scale[i] = (float)((double)t.scale[i - 1] * sf);
        t.sigma2[i] = t.scale[i] * t.scale[i];
    }
    for (int i = 0; i < L; i++) { t.invScale[i] = 1.0f / t.scale[i]; t.invSigma2[i] = 1.0f / t.sigma2[i]; }
    t.nPerLevel.assign(L, 0);
    const float factor = (float)(1.0f / sf);
    float nDesired = (float)(p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L)));
    int sum = 0;
    for (int l = 0; l < L - 1; l++) {
        t.nPerLevel[l] = cvRoundF(nDesired);
        sum += t.nPerLevel[l];
        nDesired *= factor;
    }
    t.nPerLevel[L - 1] = std::max(p.nfeatures - sum, 0);
    const int vmax = cvFloorF(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) t.umax[v] = cvRoundD(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (t.umax[v0] == t.umax[v0 + 1]) ++v0;
        t.umax[v] = v0;
        ++v0;
    }
    return t;
}

// Coefficient tables of cv::resize INTER_LINEAR 8U (OpenCV imgproc/resize.cpp, resizeGeneric_ +
// HResizeLinear<uchar,int,short,2048>): fx = (float)((dx+0.5)*scale-0.5), sx = floor(fx), fx -= sx,
// columns clamped with the weight forced to 0/1, rows clamped with the weight kept; weights are
// saturate_cast<short>(w*2048).  Entries packed for k_resize.
static inline short satShort(float v) {
    const int i = cvRoundF(v);
    return (short)std::min(std::max(i, (int)SHRT_MIN), (int)SHRT_MAX);
}
static inline int packS(short lo, short hi) { return (int)((uint32_t)(uint16_t)lo | ((uint32_t)(uint16_t)hi << 16)); }

static bool makeResizeTables(int sw, int sh, int dw, int dh, int padW, std::vector<int2> &X, std::vector<int2> &Y) {
    const double isx = (double)dw / sw, isy = (double)dh / sh;
    const double scx = 1. / isx, scy = 1. / isy;
    const int ix = (int)std::nearbyint(scx), iy = (int)std::nearbyint(scy);
    const bool area2x = std::abs(scx - ix) < DBL_EPSILON && std::abs(scy - iy) < DBL_EPSILON && ix == 2 && iy == 2;
    X.resize(padW); Y.resize(dh);
    for (int dx = 0; dx < padW; dx++) {
        const int d = std::min(dx, dw - 1);
        float fx = (float)((d + 0.5) * scx - 0.5);
        int sx = cvFloorF(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        X[dx].x = (int)((uint32_t)sx | ((uint32_t)std::min(sx + 1, sw - 1) << 16));
        X[dx].y = packS(satShort((1.f - fx) * 2048.f), satShort(fx * 2048.f));
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scy - 0.5);
        const int sy = cvFloorF(fy);
        fy -= sy;
        const int y0 = std::min(std::max(sy, 0), sh - 1), y1 = std::min(std::max(sy + 1, 0), sh - 1);
        Y[dy].x = (int)((uint32_t)y0 | ((uint32_t)y1 << 16));
        Y[dy].y = packS(satShort((1.f - fy) * 2048.f), satShort(fy * 2048.f));
    }
    return area2x;
}

// Everything that depends on the image size.  Returns 0 or an error code.
static int makeGeometry(const ygzorb_params &P, const Tables &T, int w, int h, int stride, FrameGeom &G,
                        std::vector<int2> &tabs, long long blurOff[kMaxLevels], int blurPitch[kMaxLevels],
                        long long *blurSlab) {
    memset(&G, 0, sizeof G);
    G.nlevels = P.nlevels; G.w = w; G.h = h; G.iniTh = P.iniThFAST; G.minTh = P.minThFAST;
    for (int v = 0; v < 16; v++) G.umax[v] = T.umax[v];
    for (int a = 0; a < 16; a++) {
        G.vmaxOfU[a] = 0;
        for (int v = 1; v <= 15; v++) G.vmaxOfU[a] += a <= T.umax[v];
    }
    tabs.clear();
    long long off = 0, candOff = 0, bOff = 0;
    int cells = 0, tiles = 0, keptOff = 0;
    std::vector<int2> cellTab, tileTab;
    int maxDh = 1, maxDw = 1, maxNpx = 1;
    for (int l = 0; l < P.nlevels; l++) {
        LevelGeom &L = G.lv[l];
        L.w = cvRoundF((float)w * T.invScale[l]);   // :1131-1133, from LEVEL-0 dims
        L.h = cvRoundF((float)h * T.invScale[l]);
        if (L.w < 1 || L.h < 1) return argFail("pyramid level collapses to zero size (cv::resize would throw)");
        if (L.w > 4095 || L.h > 4095) return argFail("image larger than 4095 px is not supported");
        L.pitch = l == 0 ? stride : alignUp(L.w, 64);
        L.scale = T.scale[l];
        L.kpSize = (float)(int)(31 * T.scale[l]);  // :787 PATCH_SIZE*mvScaleFactor[level], truncated to int
        L.nTarget = T.nPerLevel[l];
        if (l > 0) {
            L.off = off;
            off += alignUpLL((long long)L.pitch * L.h, 256);
            std::vector<int2> X, Y;
            L.area2x = makeResizeTables(G.lv[l - 1].w, G.lv[l - 1].h, L.w, L.h, L.pitch, X, Y) ? 1 : 0;
            // word-based kernel: non-negative weights, second tap = next column (or weight 0), a thread's 4 pixels
            // within 7 source columns; the X table must be 16-byte aligned for its int4 loads
            bool fast = (L.pitch % 4) == 0;
            for (size_t dx = 0; dx < X.size() && fast; dx++) {
                const int x0 = X[dx].x & 0xffff, x1 = (uint32_t)X[dx].x >> 16;
                const int a0 = (short)(X[dx].y & 0xffff), a1 = (short)((uint32_t)X[dx].y >> 16);
                fast = a0 >= 0 && a1 >= 0 && a0 + a1 == 2048 && (x1 == x0 + 1 || a1 == 0);
                if (fast && dx % 4 == 3) fast = x0 - (X[dx - 3].x & 0xffff) <= 6 && x0 >= (X[dx - 3].x & 0xffff);
            }
            for (size_t dy = 0; dy < Y.size() && fast; dy++) {
                const int b0 = (short)(Y[dy].y & 0xffff), b1 = (short)((uint32_t)Y[dy].y >> 16);
                fast = b0 >= 0 && b1 >= 0 && b0 + b1 == 2048;
            }
            L.fastResize = fast ? 1 : 0;
            // k_resize_band: source rows touched by a 64-row destination tile (the Y table must be monotone: it is for every
            // down-scale; checked so that an exotic table falls back to k_resize)
            L.bandRows = 0;
            if (fast && G.lv[l - 1].pitch >= 8) {
                bool mono = true;
                int band = 0;
                for (int dy0 = 0; dy0 < L.h && mono; dy0 += 64) {
                    const int dy1 = std::min(dy0 + 63, L.h - 1);
                    const int yFirst = Y[dy0].x & 0xffff, yLast = (int)((uint32_t)Y[dy1].x >> 16);
                    for (int dy = dy0; dy <= dy1 && mono; dy++) {
                        const int y0 = Y[dy].x & 0xffff, y1 = (int)((uint32_t)Y[dy].x >> 16);
                        mono = y0 >= yFirst && y1 <= yLast && y0 <= y1;
                    }
                    band = std::max(band, yLast - yFirst + 1);
                }
                if (mono) L.bandRows = band;
            }
            std::vector<int2> XG;
            L.rawPitch = 0;
            if (L.bandRows > 0) {
                // per column group of 4 destination pixels: {coef[4]}, {sel[4]}, {byte offset of the window's first word,
                // funnel shift, offset of the third word, 0}
                const int sp = G.lv[l - 1].pitch;
                int stripXs = 0, rawNeed = 0;
                for (int g = 0; g < L.pitch / 4; g++) {
                    const int2 *e = &X[4 * g];
                    const int base = e[0].x & 0xffff;
                    const int bs = std::min(base, sp - 8), delta = base - bs;
                    int sel[4];
                    for (int i = 0; i < 4; i++) {
                        const int d = (e[i].x & 0xffff) - base;
                        sel[i] = (d | ((d + 1) << 4)) + 0x11 * delta;
                    }
                    XG.push_back(make_int2(e[0].y, e[1].y)); XG.push_back(make_int2(e[2].y, e[3].y));
                    XG.push_back(make_int2(sel[0], sel[1])); XG.push_back(make_int2(sel[2], sel[3]));
                    if (g % 16 == 0) stripXs = bs & ~15;       // k_resize_band<STAGED>: the strip's staged columns start here
                    rawNeed = std::max(rawNeed, (bs & ~3) + 12 - stripXs);   // the three words of this group's window
                    XG.push_back(make_int2(bs & ~3, 8 * (bs & 3))); XG.push_back(make_int2((bs & ~3) + 12 <= sp ? 8 : 4, stripXs));
                }
                L.rawPitch = alignUp(rawNeed, 16) <= 256 ? alignUp(rawNeed, 16) : 0;
            }
            if (tabs.size() % 2) tabs.push_back(make_int2(0, 0));
            L.tabX = (int)tabs.size(); tabs.insert(tabs.end(), X.begin(), X.end());
            L.tabY = (int)tabs.size(); tabs.insert(tabs.end(), Y.begin(), Y.end());
            if (tabs.size() % 2) tabs.push_back(make_int2(0, 0));
            L.tabXG = (int)tabs.size(); tabs.insert(tabs.end(), XG.begin(), XG.end());
        }
        blurPitch[l] = alignUp(L.w, 64);
        blurOff[l] = bOff;
        bOff += alignUpLL((long long)blurPitch[l] * L.h, 256);
        // cells (:728-745)
        const int maxBX = L.w - kMinBorder, maxBY = L.h - kMinBorder;
        const float width = (float)(maxBX - kMinBorder), height = (float)(maxBY - kMinBorder);
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
        L.cellStart = cells;
        L.candOff = candOff;
        L.keptOff = keptOff;
        if (nCols > 0 && nRows > 0) {
            L.nCols = nCols; L.nRows = nRows;
            L.wCell = (int)std::ceil(width / nCols);
            L.hCell = (int)std::ceil(height / nRows);
            if (L.wCell > 64 || L.hCell > 64) return argFail("internal: FAST cell larger than 64 px");
            // NMS leaves at most one survivor per 2x2 block of a cell's detection area
            long long cap = 0;
            for (int i = 0; i < nRows; i++) {
                const int iniY = kMinBorder + i * L.hCell;
                if (iniY >= maxBY - 3) continue;
                const int dh = std::min(iniY + L.hCell + 6, maxBY) - iniY - 6;
                if (dh <= 0) continue;
                for (int j = 0; j < nCols; j++) {
                    const int iniX = kMinBorder + j * L.wCell;
                    if (iniX >= maxBX - 6) continue;
                    const int dw = std::min(iniX + L.wCell + 6, maxBX) - iniX - 6;
                    if (dw <= 0) continue;
                    cap += (long long)((dw + 1) / 2) * ((dh + 1) / 2);
                    int2 ce;
                    ce.x = iniX | (iniY << 12) | (l << 24);
                    ce.y = dw | (dh << 7) | ((i * nCols + j) << 14);
                    cellTab.push_back(ce);
                    maxDh = std::max(maxDh, dh); maxDw = std::max(maxDw, dw); maxNpx = std::max(maxNpx, dw * dh);
                    cells++;
                }
            }
            L.candCap = (int)cap;
            // quadtree roots (:537-543)
            L.nIni = (int)std::round((float)(maxBX - kMinBorder) / (maxBY - kMinBorder));
            if (L.nIni < 1) return argFail("level narrower than half its height: the reference's quadtree has no root (nIni == 0)");
            L.hX = (float)(maxBX - kMinBorder) / L.nIni;
            L.listCap = std::max(L.nTarget + 3, 4 * L.nIni) + 1;
            if (L.listCap > 16000) return argFail("nfeatures per level too large (quadtree list bound > 16000)");
        } else {
            L.listCap = 0;  // canonicalised: nRows == 0 -> ceil(x/0) in the reference, no cells run
        }
        candOff += alignUpLL(L.candCap, 32);
        keptOff += L.listCap;
        G.maxListCap = std::max(G.maxListCap, L.listCap);
        L.tilesX = (L.w + kBlurTW - 1) / kBlurTW; L.tilesY = (L.h + kBlurTH - 1) / kBlurTH;
        L.tileStart = tiles;
        tiles += L.tilesX * L.tilesY;
        for (int ty = 0; ty < L.tilesY; ty++)
            for (int tx = 0; tx < L.tilesX; tx++) {
                int2 te;
                te.x = (tx * kBlurTW) | ((ty * kBlurTH) << 12) | (l << 24); te.y = 0;
                tileTab.push_back(te);
            }
    }
    G.totalCells = cells; G.totalTiles = tiles;
    // k_fast shared memory: [image tile] [score tile] [per-warp lists]; a list holds every pixel its warp visits
    G.fastWarpCap = (maxNpx + kFastCtaThreads - 1) / kFastCtaThreads * 32;
    // tile pitch: the narrow instantiation when every cell (dw + 6 window + up to 3 alignment bytes) fits 48 bytes — the usual
    // 30..39-px cells — which brings the CTA's shared memory under the 32-CTAs-per-SM line; else the general 76
    G.fastPitch = (maxDw + 9 <= kFastTilePitchNarrow) ? kFastTilePitchNarrow : kFastTilePitch;
    G.fastRegionA = alignUp((maxDh + 6) * G.fastPitch, 16);
    G.fastScoreBytes = alignUp((maxDh + 2) * G.fastPitch, 16);
    G.cellTab = (int)tabs.size(); tabs.insert(tabs.end(), cellTab.begin(), cellTab.end());
    G.tileTab = (int)tabs.size(); tabs.insert(tabs.end(), tileTab.begin(), tileTab.end());
    G.slabBytes = std::max(off, 256LL);
    G.candPerFrame = std::max(candOff, 32LL);
    G.keptCap = std::max(keptOff, 1);
    G.maxListCap = std::max(G.maxListCap, 8);
    *blurSlab = bOff;
    if (quadtreeSmemBytes(G.maxListCap) > 200 * 1024) return argFail("nfeatures per level too large for the quadtree kernel's shared memory");
    return 0;
}


// -------------------------------------------------------------------------------------------------
#ifndef YGZ_IN_SLOTS
#define YGZ_IN_SLOTS 3
#endif
constexpr int kInSlots = YGZ_IN_SLOTS;   // host-input staging slots

struct Extractor {
    ygzorb_params params;
    cudaStream_t side = nullptr; cudaEvent_t sideGo = nullptr, sideDone = nullptr;   // blur beside the quadtree (batches)
    Tables tables;
    int device = 0, maxBatch = 1;
    cudaStream_t stream = nullptr;
    // geometry of the current image size
    int gw = -1, gh = -1, gstride = -1;
    FrameGeom G;
    DevBuffers B;
    long long blurOff[kMaxLevels]; int blurPitch[kMaxLevels]; long long blurSlab = 0;
    // device workspace (grown on demand)
    // host-input staging: kInSlots slots filled by the upload stream while the compute stream works on an earlier one (three, so
    // that the upload of chunk k+2 need not wait for the kernels of chunk k: the PCIe link stays busy)
    uint8_t *dIn[kInSlots] = {}; size_t inBytes[kInSlots] = {};
    cudaStream_t h2d = nullptr, d2h = nullptr;
    cudaEvent_t inReady[kInSlots] = {}, inFree[kInSlots] = {}, outReady[2] = {}, outFree[2] = {};
    unsigned inSeq = 0, outSeq = 0;
    int pipeFrames = 256;      // frames per pipelined chunk of a host-input call (0: one chunk per max_batch)
    uint8_t *dPyr = nullptr; size_t pyrBytes = 0;
    uint8_t *dBlur = nullptr; size_t blurBytes = 0;
    int2 *dTabs = nullptr; size_t tabBytes = 0;
    uint2 *dCand = nullptr; size_t candBytes = 0;
    uint16_t *dNodeOf = nullptr; size_t nodeBytes = 0;
    uint32_t *dKept = nullptr; size_t keptBytes = 0;
    int *dCounts = nullptr;   // [maxBatch][kMaxLevels] x2 (cand, kept) + status
    // host-output staging, two slots drained by the download stream
    int32_t *dOutCounts[2] = {}; ygzorb_keypoint *dOutKps[2] = {}; uint8_t *dOutDesc[2] = {};
    size_t outCap = 0;  // records per frame the dOut* buffers hold
    int lastFrames = 0; bool lastInDevice = false; bool stagesValid = false;
    const uint8_t *lastIn = nullptr; long long lastInFrameStride = 0;
    // undistortion in front of the pyramid (ygzorb_extractor_set_undistort_maps)
    short2 *dMap1 = nullptr; uint16_t *dMap2 = nullptr; size_t map1Bytes = 0, map2Bytes = 0;
    int mapW = 0, mapH = 0;
    uint8_t *dUndist = nullptr; size_t undistBytes = 0;
    // DSO keypoint mode (ygzorb_extract_dso_batch): workspace grown on demand
    uint32_t *dsoOcc = nullptr; size_t dsoOccBytes = 0;
    uint32_t *dsoCellKp = nullptr; size_t dsoCellKpBytes = 0;
    uint8_t *dsoCellCnt = nullptr; size_t dsoCellCntBytes = 0;
    int32_t *dsoSmall = nullptr; size_t dsoSmallBytes = 0;     // grid, nExist, total, counts [maxBatch] + active bytes
    int4 *dsoCentre = nullptr; size_t dsoCentreBytes = 0;
    ygzorb_keypoint *dsoKps = nullptr; size_t dsoKpsBytes = 0;
    uint8_t *dsoDesc = nullptr; size_t dsoDescBytes = 0;
    // ygzorb_extractor_set_pyramid_sink: host slab that receives the pyramid of single-frame calls on the download stream
    uint8_t *pyrSink = nullptr;
    cudaEvent_t pyrReady = nullptr, pyrCopied = nullptr;
    long long launches = 0;
    bool profiling = false;
    cudaEvent_t ev[8] = {};
    float stageMs[8] = {};
};

template <class T> static int growBuf(T *&p, size_t &have, size_t need) {
    if (need <= have && p) return 0;
    if (p) YGZ_CUDA(cudaFree(p));
    p = nullptr; have = 0;
    YGZ_CUDA(cudaMalloc((void **)&p, std::max<size_t>(need, 256)));
    have = need;
    return 0;
}

static int ensureGeometry(Extractor *ex, int w, int h, int stride, int nFrames, bool inDevice) {
    if (w != ex->gw || h != ex->gh || stride != ex->gstride) {
        std::vector<int2> tabs;
        FrameGeom G;
        long long bo[kMaxLevels]; int bp[kMaxLevels]; long long bs = 0;
        // host frames are packed into the staging buffer with stride == w
        const int rc = makeGeometry(ex->params, ex->tables, w, h, stride, G, tabs, bo, bp, &bs);
        if (rc) return rc;
        ex->G = G; memcpy(ex->blurOff, bo, sizeof bo); memcpy(ex->blurPitch, bp, sizeof bp); ex->blurSlab = bs;
        ex->gw = w; ex->gh = h; ex->gstride = stride;
        ex->stagesValid = false;
        YGZ_CUDA(cudaStreamSynchronize(ex->stream));
        if (int rc2 = growBuf(ex->dTabs, ex->tabBytes, std::max<size_t>(tabs.size(), 1) * sizeof(int2))) return rc2;
        if (!tabs.empty()) YGZ_CUDA(cudaMemcpy(ex->dTabs, tabs.data(), tabs.size() * sizeof(int2), cudaMemcpyHostToDevice));
    }
    const size_t nb = (size_t)ex->maxBatch;
    (void)nFrames;
    int rc;
    if (!inDevice)
        for (int sl = 0; sl < kInSlots; sl++)
            if ((rc = growBuf(ex->dIn[sl], ex->inBytes[sl], nb * (size_t)w * h + 64))) return rc;
    if ((rc = growBuf(ex->dPyr, ex->pyrBytes, nb * (size_t)ex->G.slabBytes))) return rc;
    if ((rc = growBuf(ex->dBlur, ex->blurBytes, nb * (size_t)ex->blurSlab))) return rc;
    if ((rc = growBuf(ex->dCand, ex->candBytes, nb * (size_t)ex->G.candPerFrame * sizeof(uint2)))) return rc;
    if ((rc = growBuf(ex->dNodeOf, ex->nodeBytes, nb * (size_t)ex->G.candPerFrame * sizeof(uint16_t)))) return rc;
    if ((rc = growBuf(ex->dKept, ex->keptBytes, nb * (size_t)ex->G.keptCap * sizeof(uint32_t)))) return rc;
    DevBuffers &B = ex->B;
    B.pyr = ex->dPyr; B.blur = ex->dBlur; B.blurSlab = ex->blurSlab;
    for (int l = 0; l < kMaxLevels; l++) { B.blurOff[l] = ex->blurOff[l]; B.blurPitch[l] = ex->blurPitch[l]; }
    B.tabs = ex->dTabs; B.cand = ex->dCand; B.nodeOf = ex->dNodeOf; B.kept = ex->dKept;
    B.candCount = ex->dCounts; B.keptCount = ex->dCounts + nb * kMaxLevels; B.status = ex->dCounts + 2 * nb * kMaxLevels;
    return 0;
}

static int ensureOut(Extractor *ex, int cap) {
    if ((size_t)cap <= ex->outCap && ex->dOutKps[0]) return 0;
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    YGZ_CUDA(cudaStreamSynchronize(ex->d2h));
    for (int sl = 0; sl < 2; sl++) {
        if (ex->dOutKps[sl]) YGZ_CUDA(cudaFree(ex->dOutKps[sl]));
        if (ex->dOutDesc[sl]) YGZ_CUDA(cudaFree(ex->dOutDesc[sl]));
        ex->dOutKps[sl] = nullptr; ex->dOutDesc[sl] = nullptr;
    }
    ex->outCap = 0;
    for (int sl = 0; sl < 2; sl++) {
        YGZ_CUDA(cudaMalloc((void **)&ex->dOutKps[sl], (size_t)ex->maxBatch * cap * sizeof(ygzorb_keypoint)));
        YGZ_CUDA(cudaMalloc((void **)&ex->dOutDesc[sl], (size_t)ex->maxBatch * cap * 32));
    }
    ex->outCap = cap;
    return 0;
}

// Host input: enqueue the upload of a chunk of frames into the next staging slot on the upload stream
// (which first waits until the kernels that last read that slot are done).  Returns the slot.
static int uploadInput(Extractor *ex, const uint8_t *frames, int n, int w, int h, int stride, size_t frameStride, int *slotOut) {
    const int sl = (int)(ex->inSeq++ % (unsigned)kInSlots);
    uint8_t *dst = ex->dIn[sl];
    YGZ_CUDA(cudaStreamWaitEvent(ex->h2d, ex->inFree[sl], 0));
    // staging buffer is continuous: pitch == w (the geometry was built with stride == w for host input)
    if (stride == w && frameStride == (size_t)w * h) {
        YGZ_CUDA(cudaMemcpyAsync(dst, frames, (size_t)n * w * h, cudaMemcpyHostToDevice, ex->h2d));
    } else if (frameStride == (size_t)stride * h) {
        YGZ_CUDA(cudaMemcpy2DAsync(dst, w, frames, stride, w, (size_t)h * n, cudaMemcpyHostToDevice, ex->h2d));
    } else {
        for (int f = 0; f < n; f++)
            YGZ_CUDA(cudaMemcpy2DAsync(dst + (size_t)f * w * h, w, frames + f * frameStride, stride, w, h,
                                       cudaMemcpyHostToDevice, ex->h2d));
    }
    YGZ_CUDA(cudaEventRecord(ex->inReady[sl], ex->h2d));
    *slotOut = sl;
    return 0;
}

// Make level 0 of the chunk visible to the kernels: alias device input, or wait for the slot's upload.
static int bindInput(Extractor *ex, const uint8_t *frames, int w, int h, size_t frameStride, bool inDevice, int slot) {
    if (inDevice) {
        ex->B.in = frames; ex->B.inFrameStride = (long long)frameStride;
        return 0;
    }
    YGZ_CUDA(cudaStreamWaitEvent(ex->stream, ex->inReady[slot], 0));
    ex->B.in = ex->dIn[slot]; ex->B.inFrameStride = (long long)w * h;
    return 0;
}

// cv::remap of the bound level 0 into the extractor's own buffer (Frame::ComputeImagePyramid, src/Frame.cc:775-790), when maps are set
static int applyUndistort(Extractor *ex, int n, int w, int h, int srcPitch) {
    if (!ex->dMap1) return 0;
    if (w != ex->mapW || h != ex->mapH) { setError("image size differs from the undistortion maps"); return YGZORB_E_ARG; }
    if (int rc = growBuf(ex->dUndist, ex->undistBytes, (size_t)ex->maxBatch * w * h + 64)) return rc;
    ex->launches += launchRemap(ex->B.in, ex->B.inFrameStride, srcPitch, w, h, n, ex->dMap1, ex->dMap2, ex->dUndist, ex->stream);
    ex->B.in = ex->dUndist; ex->B.inFrameStride = (long long)w * h;
    return 0;
}

constexpr int kPdlMaxFrames = 8;

static int runPyramid(Extractor *ex, int n) {
    for (int l = 1; l < ex->G.nlevels; l++) ex->launches += launchResizeLevel(ex->G, ex->B, l, n, ex->stream);
    return 0;
}

// Single-frame calls with a pyramid sink: the levels go to the host on the download stream while FAST / quadtree / blur /
// descriptors run (the reference's callers read mvImagePyramid right after operator(), src/Frame.cc:810-813).
static int sinkPyramid(Extractor *ex, int n) {
    if (!ex->pyrSink || n != 1) return 0;
    YGZ_CUDA(cudaEventRecord(ex->pyrReady, ex->stream));
    YGZ_CUDA(cudaStreamWaitEvent(ex->d2h, ex->pyrReady, 0));
    size_t off = 0;
    for (int l = 0; l < ex->G.nlevels; l++) {
        const LevelGeom &L = ex->G.lv[l];
        const uint8_t *src = l == 0 ? ex->B.in : ex->B.pyr + L.off;
        YGZ_CUDA(cudaMemcpy2DAsync(ex->pyrSink + off, L.w, src, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost, ex->d2h));
        off += (size_t)L.w * L.h;
    }
    YGZ_CUDA(cudaEventRecord(ex->pyrCopied, ex->d2h));
    YGZ_CUDA(cudaStreamWaitEvent(ex->stream, ex->pyrCopied, 0));   // level 0's staging slot and the slab are reused by later work of this stream
    return 0;
}

#define YGZ_MARK(i) do { if (ex->profiling) YGZ_CUDA(cudaEventRecord(ex->ev[i], ex->stream)); } while (0)

// Kernels of one chunk (its input already uploaded into `slot` when it comes from the host).
static int extractChunk(Extractor *ex, const uint8_t *frames, int n, int w, int h, size_t frameStride, int srcPitch,
                        unsigned flags, int cap, int slot, int32_t *counts, ygzorb_keypoint *kps, uint8_t *desc) {
    const bool inDev = flags & YGZORB_IN_DEVICE;
    // page-locked host outputs of a small call are written by k_describe itself (UVA: the host pointer is a device pointer)
    const bool outDev = (flags & YGZORB_OUT_DEVICE) || ((flags & YGZORB_OUT_PINNED) && n <= kPdlMaxFrames);
    int rc;
    YGZ_MARK(0);
    if ((rc = bindInput(ex, frames, w, h, frameStride, inDev, slot))) return rc;
    if ((rc = applyUndistort(ex, n, w, h, srcPitch))) return rc;
    // the counters only: the status word behind them is sticky across the chunks of a call and cleared when it is read
    YGZ_CUDA(cudaMemsetAsync(ex->dCounts, 0, 2 * (size_t)ex->maxBatch * kMaxLevels * sizeof(int), ex->stream));
    YGZ_MARK(1);
    const int pdl = (n <= kPdlMaxFrames && !ex->profiling) ? 1 : 0;   // live tracking: the chain is launch-latency bound
    ex->B.pdl = ex->profiling ? 0 : 1;   // the 7 dependent pyramid launches are short at any batch size
    runPyramid(ex, n);
    ex->B.pdl = pdl;
    YGZ_MARK(2);
    const bool sink = ex->pyrSink && n == 1;
    if (sink) {   // the copies start now; the stream takes their completion in only at the end of the chunk
        YGZ_CUDA(cudaEventRecord(ex->pyrReady, ex->stream));
        YGZ_CUDA(cudaStreamWaitEvent(ex->d2h, ex->pyrReady, 0));
        size_t off = 0;
        for (int l = 0; l < ex->G.nlevels; l++) {
            const LevelGeom &L = ex->G.lv[l];
            const uint8_t *src = l == 0 ? ex->B.in : ex->B.pyr + L.off;
            YGZ_CUDA(cudaMemcpy2DAsync(ex->pyrSink + off, L.w, src, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost, ex->d2h));
            off += (size_t)L.w * L.h;
        }
        YGZ_CUDA(cudaEventRecord(ex->pyrCopied, ex->d2h));
    }
#ifndef YGZ_BLUR_SIDE
#define YGZ_BLUR_SIDE 1
#endif
    if (YGZ_BLUR_SIDE && !pdl && !ex->profiling) {
        // batches: the blur (which only needs the pyramid) runs on a side stream beside the quadtree, whose CTAs leave half of
        // the issue slots idle (barrier-separated passes)
        if (!ex->side) {
            YGZ_CUDA(cudaStreamCreateWithFlags(&ex->side, cudaStreamNonBlocking));
            YGZ_CUDA(cudaEventCreateWithFlags(&ex->sideGo, cudaEventDisableTiming));
            YGZ_CUDA(cudaEventCreateWithFlags(&ex->sideDone, cudaEventDisableTiming));
        }
        if (YGZ_BLUR_SIDE == 2) YGZ_CUDA(cudaEventRecord(ex->sideGo, ex->stream));
        ex->launches += launchFast(ex->G, ex->B, n, ex->stream);
        YGZ_MARK(3);
        if (YGZ_BLUR_SIDE == 1) YGZ_CUDA(cudaEventRecord(ex->sideGo, ex->stream));
        YGZ_CUDA(cudaStreamWaitEvent(ex->side, ex->sideGo, 0));
        ex->launches += launchBlur(ex->G, ex->B, n, ex->side);
        YGZ_CUDA(cudaEventRecord(ex->sideDone, ex->side));
        ex->launches += launchQuadtree(ex->G, ex->B, n, ex->stream);
        YGZ_MARK(4);
        YGZ_CUDA(cudaStreamWaitEvent(ex->stream, ex->sideDone, 0));
        YGZ_MARK(5);
    } else {
        ex->launches += launchFast(ex->G, ex->B, n, ex->stream);
        YGZ_MARK(3);
        ex->launches += launchQuadtree(ex->G, ex->B, n, ex->stream);
        YGZ_MARK(4);
        ex->launches += launchBlur(ex->G, ex->B, n, ex->stream);
        YGZ_MARK(5);
    }
    const int os = outDev ? 0 : (int)(ex->outSeq++ & 1u);
    int32_t *oc = outDev ? counts : ex->dOutCounts[os];
    ygzorb_keypoint *ok = outDev ? kps : ex->dOutKps[os];
    uint8_t *od = outDev ? desc : ex->dOutDesc[os];
    if (!outDev) YGZ_CUDA(cudaStreamWaitEvent(ex->stream, ex->outFree[os], 0));
    ex->launches += launchDescribe(ex->G, ex->B, n, cap, oc, ok, od, ex->stream);
    ex->B.pdl = 0;
    YGZ_MARK(6);
    if (sink) YGZ_CUDA(cudaStreamWaitEvent(ex->stream, ex->pyrCopied, 0));   // slot and slab are reused only after the copies
    if (!inDev) YGZ_CUDA(cudaEventRecord(ex->inFree[slot], ex->stream));   // IC_Angle was the last reader of level 0
    if (!outDev) {   // drain on the download stream so that the next chunk's kernels overlap it
        YGZ_CUDA(cudaEventRecord(ex->outReady[os], ex->stream));
        YGZ_CUDA(cudaStreamWaitEvent(ex->d2h, ex->outReady[os], 0));
        YGZ_CUDA(cudaMemcpyAsync(counts, oc, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, ex->d2h));
        YGZ_CUDA(cudaMemcpyAsync(kps, ok, (size_t)n * cap * sizeof(ygzorb_keypoint), cudaMemcpyDeviceToHost, ex->d2h));
        YGZ_CUDA(cudaMemcpyAsync(desc, od, (size_t)n * cap * 32, cudaMemcpyDeviceToHost, ex->d2h));
        YGZ_CUDA(cudaEventRecord(ex->outFree[os], ex->d2h));
    }
    YGZ_MARK(7);
    YGZ_CUDA(cudaGetLastError());
    ex->lastFrames = n; ex->stagesValid = true;
    return 0;
}

static int finishProfile(Extractor *ex) {
    if (!ex->profiling) return 0;
    for (int i = 0; i < 7; i++) YGZ_CUDA(cudaEventElapsedTime(&ex->stageMs[i], ex->ev[i], ex->ev[i + 1]));
    return 0;
}

static int checkStatus(Extractor *ex) {
    int st = 0;
    YGZ_CUDA(cudaMemcpy(&st, ex->B.status, sizeof st, cudaMemcpyDeviceToHost));
    if (st) {
        YGZ_CUDA(cudaMemset(ex->B.status, 0, sizeof st));
        setError("internal capacity overflow in the FAST/quadtree stage (status " + std::to_string(st) + ")");
        return YGZORB_E_CAPACITY;
    }
    return 0;
}

}  // namespace ygz

namespace ygz {

int extractorLevelView(ygzorb_extractor *h, int frame, int level, LevelView *out) {
    Extractor *ex = (Extractor *)h;
    if (!ex || !out) { setError("null argument"); return YGZORB_E_ARG; }
    if (ex->lastFrames <= 0) { setError("extractor holds no pyramid yet"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames || level < 0 || level >= ex->G.nlevels) { setError("frame/level out of range"); return YGZORB_E_ARG; }
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    const LevelGeom &L = ex->G.lv[level];
    out->ptr = level == 0 ? ex->B.in + (long long)frame * ex->B.inFrameStride : ex->B.pyr + (long long)frame * ex->G.slabBytes + L.off;
    out->pitch = L.pitch; out->w = L.w; out->h = L.h;
    return 0;
}
int extractorTables(ygzorb_extractor *h, int *nlevels, const float **scale, const float **invScale) {
    Extractor *ex = (Extractor *)h;
    if (!ex) { setError("null extractor"); return YGZORB_E_ARG; }
    *nlevels = ex->params.nlevels; *scale = ex->tables.scale.data(); *invScale = ex->tables.invScale.data();
    return 0;
}

}  // namespace ygz

// =================================================================================================
// C ABI
using namespace ygz;

extern "C" {

const char *ygzorb_last_error(void) { return gLastError.c_str(); }
const char *ygzorb_version(void) { return "ygzorb-b200 0.1 (sm_100a)"; }

int ygzorb_device_count(void) {
    int n = 0;
    YGZ_CUDA(cudaGetDeviceCount(&n));
    return n;
}

int ygzorb_extractor_create(const ygzorb_params *params, int max_w, int max_h, int max_batch, int device,
                            ygzorb_extractor **out) {
    if (!params || !out) return argFail("null argument");
    if (params->nlevels < 1 || params->nlevels > kMaxLevels) return argFail("nlevels must be in [1, 16]");
    if (params->nfeatures < 1 || !(params->scaleFactor > 1.0f)) return argFail("nfeatures must be >= 1 and scaleFactor > 1");
    if (params->iniThFAST < 1 || params->iniThFAST > 255 || params->minThFAST < 1 || params->minThFAST > params->iniThFAST)
        return argFail("FAST thresholds must satisfy 1 <= minThFAST <= iniThFAST <= 255");
    if (max_batch < 1 || max_batch > 65535) return argFail("max_batch must be in [1, 65535]");
    (void)max_w; (void)max_h;  // workspace is sized on first use and grown on demand
    YGZ_CUDA(cudaSetDevice(device));
    Extractor *ex = new Extractor();
    ex->params = *params; ex->tables = makeTables(*params); ex->device = device; ex->maxBatch = max_batch;
    memset(&ex->B, 0, sizeof ex->B);
    int rc = configureKernels();
    if (rc) { delete ex; return rc; }
    cudaError_t e = cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void **)&ex->dCounts, (2 * (size_t)max_batch * kMaxLevels + 1) * sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(ex->dCounts, 0, (2 * (size_t)max_batch * kMaxLevels + 1) * sizeof(int));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ex->h2d, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ex->d2h, cudaStreamNonBlocking);
    for (int sl = 0; sl < 2 && e == cudaSuccess; sl++) {
        e = cudaMalloc((void **)&ex->dOutCounts[sl], (size_t)max_batch * sizeof(int32_t));
        cudaEvent_t *evs[] = {&ex->outReady[sl], &ex->outFree[sl]};
        for (cudaEvent_t *pe : evs) if (e == cudaSuccess) e = cudaEventCreateWithFlags(pe, cudaEventDisableTiming);
    }
    for (int sl = 0; sl < kInSlots && e == cudaSuccess; sl++) {
        cudaEvent_t *evs[] = {&ex->inReady[sl], &ex->inFree[sl]};
        for (cudaEvent_t *pe : evs) if (e == cudaSuccess) e = cudaEventCreateWithFlags(pe, cudaEventDisableTiming);
    }
    for (int i = 0; i < 8 && e == cudaSuccess; i++) e = cudaEventCreate(&ex->ev[i]);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->pyrReady, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ex->pyrCopied, cudaEventDisableTiming);
    if (e != cudaSuccess) { ygzorb_extractor_destroy((ygzorb_extractor *)ex); return cudaFail(e, "extractor_create", __FILE__, __LINE__); }
    *out = (ygzorb_extractor *)ex;
    return 0;
}

void ygzorb_extractor_destroy(ygzorb_extractor *h) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return;
    cudaSetDevice(ex->device);
    cudaStream_t streams[] = {ex->h2d, ex->stream, ex->d2h, ex->side};
    for (cudaStream_t st : streams) if (st) cudaStreamSynchronize(st);
    if (ex->sideGo) cudaEventDestroy(ex->sideGo);
    if (ex->sideDone) cudaEventDestroy(ex->sideDone);
    for (int sl = 0; sl < kInSlots; sl++) if (ex->dIn[sl]) cudaFree(ex->dIn[sl]);
    void *bufs[] = {ex->dPyr, ex->dBlur, ex->dTabs, ex->dCand, ex->dNodeOf, ex->dKept, ex->dCounts,
                    ex->dOutCounts[0], ex->dOutCounts[1], ex->dOutKps[0], ex->dOutKps[1], ex->dOutDesc[0], ex->dOutDesc[1],
                    ex->dsoOcc, ex->dsoCellKp, ex->dsoCellCnt, ex->dsoSmall, ex->dsoCentre, ex->dsoKps, ex->dsoDesc,
                    ex->dMap1, ex->dMap2, ex->dUndist};
    for (void *p : bufs) if (p) cudaFree(p);
    for (int i = 0; i < 8; i++) if (ex->ev[i]) cudaEventDestroy(ex->ev[i]);
    if (ex->pyrReady) cudaEventDestroy(ex->pyrReady);
    if (ex->pyrCopied) cudaEventDestroy(ex->pyrCopied);
    for (int sl = 0; sl < 2; sl++) {
        cudaEvent_t evs[] = {ex->outReady[sl], ex->outFree[sl]};
        for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    }
    for (int sl = 0; sl < kInSlots; sl++) {
        cudaEvent_t evs[] = {ex->inReady[sl], ex->inFree[sl]};
        for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
    }
    for (cudaStream_t st : streams) if (st) cudaStreamDestroy(st);
    delete ex;
}

int ygzorb_extractor_tables(const ygzorb_extractor *h, float *scale, float *inv_scale, float *sigma2,
                            float *inv_sigma2, int *n_per_level, int *umax16) {
    const Extractor *ex = (const Extractor *)h;
    if (!ex) return argFail("null extractor");
    for (int l = 0; l < ex->params.nlevels; l++) {
        if (scale) scale[l] = ex->tables.scale[l];
        if (inv_scale) inv_scale[l] = ex->tables.invScale[l];
        if (sigma2) sigma2[l] = ex->tables.sigma2[l];
        if (inv_sigma2) inv_sigma2[l] = ex->tables.invSigma2[l];
        if (n_per_level) n_per_level[l] = ex->tables.nPerLevel[l];
    }
    if (umax16) for (int v = 0; v < 16; v++) umax16[v] = ex->tables.umax[v];
    return 0;
}

static int geometryOnly(const Extractor *ex, int w, int h, FrameGeom &G) {
    std::vector<int2> tabs;
    long long bo[kMaxLevels]; int bp[kMaxLevels]; long long bs;
    return makeGeometry(ex->params, ex->tables, w, h, w, G, tabs, bo, bp, &bs);
}

int ygzorb_extractor_level_sizes(const ygzorb_extractor *h, int w, int hgt, int *lw, int *lh) {
    const Extractor *ex = (const Extractor *)h;
    if (!ex || !lw || !lh) return argFail("null argument");
    for (int l = 0; l < ex->params.nlevels; l++) {
        lw[l] = cvRoundF((float)w * ex->tables.invScale[l]);
        lh[l] = cvRoundF((float)hgt * ex->tables.invScale[l]);
    }
    return 0;
}

int ygzorb_extractor_max_keypoints(const ygzorb_extractor *h, int w, int hgt) {
    const Extractor *ex = (const Extractor *)h;
    if (!ex) return argFail("null extractor");
    if (w <= 0 || hgt <= 0) return 1;
    FrameGeom G;
    const int rc = geometryOnly(ex, w, hgt, G);
    return rc ? rc : G.keptCap;
}

static int checkFrames(Extractor *ex, const uint8_t *frames, int n, int w, int h, int stride) {
    if (!ex) return argFail("null extractor");
    if (n < 0) return argFail("n_frames < 0");
    if (n > 0 && w > 0 && h > 0 && !frames) return argFail("null frames pointer");
    if (w > 0 && stride < w) return argFail("stride < w");
    return 0;
}

int ygzorb_extract_batch(ygzorb_extractor *h, const uint8_t *frames, int n_frames, int w, int hgt, int stride,
                         size_t frame_stride, unsigned flags, int cap, int32_t *counts, ygzorb_keypoint *keypoints,
                         uint8_t *desc) {
    Extractor *ex = (Extractor *)h;
    int rc = checkFrames(ex, frames, n_frames, w, hgt, stride);
    if (rc) return rc;
    if (!counts || (n_frames > 0 && cap > 0 && (!keypoints || !desc))) return argFail("null output pointer");
    YGZ_CUDA(cudaSetDevice(ex->device));
    const bool outDev = flags & YGZORB_OUT_DEVICE;
    if (w <= 0 || hgt <= 0 || n_frames == 0) {  // empty image: silent return, nothing extracted (:972-973)
        if (outDev) { if (n_frames) YGZ_CUDA(cudaMemsetAsync(counts, 0, (size_t)n_frames * 4, ex->stream)); }
        else for (int f = 0; f < n_frames; f++) counts[f] = 0;
        return 0;
    }
    if (frame_stride == 0) frame_stride = (size_t)stride * hgt;
    const bool inDev = flags & YGZORB_IN_DEVICE;
    // Chunking.  Device input: chunks of max_batch (the workspace size).  Host input: chunks of the pipeline
    // size, the upload of chunk k+1 (upload stream, other staging slot) overlapping the kernels of chunk k and
    // the download of chunk k-1 (download stream); YGZORB_KEEP_STAGES asks for one chunk so that the stage
    // taps cover every frame of the call.
    int chunk = ex->maxBatch;
    if (!inDev && !(flags & YGZORB_KEEP_STAGES) && !ex->profiling && ex->pipeFrames > 0) chunk = std::min(chunk, ex->pipeFrames);
    if ((rc = ensureGeometry(ex, w, hgt, (inDev && !ex->dMap1) ? stride : w, std::min(chunk, n_frames), inDev))) return rc;
    if (cap < ex->G.keptCap) { setError("cap is smaller than ygzorb_extractor_max_keypoints()"); return YGZORB_E_CAPACITY; }
    if (!outDev && (rc = ensureOut(ex, cap))) return rc;
    int slot = 0, nextSlot = 0;
    if (!inDev && (rc = uploadInput(ex, frames, std::min(chunk, n_frames), w, hgt, stride, frame_stride, &nextSlot))) return rc;
    for (int f0 = 0; f0 < n_frames; f0 += chunk) {
        const int n = std::min(chunk, n_frames - f0);
        slot = nextSlot;
        if (!inDev && f0 + n < n_frames) {   // enqueue the next upload before this chunk's kernels
            const int n2 = std::min(chunk, n_frames - f0 - n);
            if ((rc = uploadInput(ex, frames + (size_t)(f0 + n) * frame_stride, n2, w, hgt, stride, frame_stride, &nextSlot))) return rc;
        }
        rc = extractChunk(ex, frames + (size_t)f0 * frame_stride, n, w, hgt, frame_stride, inDev ? stride : w, flags, cap, slot,
                          counts + f0, keypoints + (size_t)f0 * cap, desc + (size_t)f0 * cap * 32);
        if (rc) return rc;
    }
    if (!(flags & YGZORB_NO_SYNC)) return ygzorb_extractor_sync(h);
    return 0;
}

int ygzorb_extract_dso_batch(ygzorb_extractor *h, const uint8_t *frames, int n_frames, int w, int hgt, int stride,
                             size_t frame_stride, unsigned flags, int32_t *grid_size, const int32_t *n_exist, int cap,
                             int32_t *counts, ygzorb_keypoint *keypoints, uint8_t *desc) {
    Extractor *ex = (Extractor *)h;
    int rc = checkFrames(ex, frames, n_frames, w, hgt, stride);
    if (rc) return rc;
    if (!grid_size || !n_exist || !counts || !keypoints || !desc || cap < 1) return argFail("null argument");
    if (n_frames == 0) return 0;
    if (w < 64 || hgt < 64) return argFail("image too small for the DSO keypoint grid");
    if (n_frames > ex->maxBatch) return argFail("n_frames exceeds max_batch");
    YGZ_CUDA(cudaSetDevice(ex->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    if (frame_stride == 0) frame_stride = (size_t)stride * hgt;
    if ((rc = ensureGeometry(ex, w, hgt, (inDev && !ex->dMap1) ? stride : w, n_frames, inDev))) return rc;
    const int nfeat = ex->params.nfeatures;
    // ---- host-side state of the while (cnt < n) loop (src/ORBextractor.cc:1295-1376), one entry per frame
    std::vector<int32_t> state(n_frames), used(n_frames), cnt(n_frames, 0), nE(n_frames);
    std::vector<uint8_t> active(n_frames, 1);
    int maxExist = 0;
    for (int f = 0; f < n_frames; f++) {
        state[f] = grid_size[f] < 0 ? (int)std::sqrt(1.0 * hgt * w / nfeat) : grid_size[f];   // :1295-1296
        if (state[f] < 7 || state[f] > kDsoMaxGrid) return argFail("grid size outside 7..96 (nfeatures too small for this image?)");
        if (n_exist[f] < 0 || n_exist[f] > cap) return argFail("n_exist outside 0..cap");
        nE[f] = n_exist[f]; maxExist = std::max(maxExist, nE[f]);
    }
    const int maxCells = (hgt / 7) * (w / 7);
    const int occPitch = (w + 31) / 32;
    const size_t nb = (size_t)n_frames;
    if ((rc = growBuf(ex->dsoOcc, ex->dsoOccBytes, nb * occPitch * hgt * 4))) return rc;
    if ((rc = growBuf(ex->dsoCellKp, ex->dsoCellKpBytes, nb * maxCells * 3 * 4))) return rc;
    if ((rc = growBuf(ex->dsoCellCnt, ex->dsoCellCntBytes, nb * maxCells))) return rc;
    if ((rc = growBuf(ex->dsoSmall, ex->dsoSmallBytes, (size_t)ex->maxBatch * 20))) return rc;
    if ((rc = growBuf(ex->dsoCentre, ex->dsoCentreBytes, nb * cap * sizeof(int4)))) return rc;
    if (!outDev) {
        if ((rc = growBuf(ex->dsoKps, ex->dsoKpsBytes, nb * cap * sizeof(ygzorb_keypoint)))) return rc;
        if ((rc = growBuf(ex->dsoDesc, ex->dsoDescBytes, nb * cap * 32))) return rc;
    }
    int32_t *dGrid = ex->dsoSmall, *dNE = dGrid + ex->maxBatch, *dTotal = dNE + ex->maxBatch, *dCounts = dTotal + ex->maxBatch;
    uint8_t *dActive = (uint8_t *)(dCounts + ex->maxBatch);
    ygzorb_keypoint *dKps = outDev ? keypoints : ex->dsoKps;
    uint8_t *dDesc = outDev ? desc : ex->dsoDesc;
    int32_t *dCnt = outDev ? counts : dCounts;
    cudaStream_t st = ex->stream;
    // ---- level 0, pyramid and blurred pyramid (Frame::ComputeImagePyramid + :1080-1084)
    int slot = 0;
    if (!inDev && (rc = uploadInput(ex, frames, n_frames, w, hgt, stride, frame_stride, &slot))) return rc;
    if ((rc = bindInput(ex, frames, w, hgt, frame_stride, inDev, slot))) return rc;
    if ((rc = applyUndistort(ex, n_frames, w, hgt, inDev ? stride : w))) return rc;
    runPyramid(ex, n_frames);
    ex->launches += launchBlur(ex->G, ex->B, n_frames, st);
    if (!outDev)
        for (int f = 0; f < n_frames; f++)
            if (nE[f]) YGZ_CUDA(cudaMemcpyAsync(dKps + (size_t)f * cap, keypoints + (size_t)f * cap, (size_t)nE[f] * sizeof(ygzorb_keypoint),
                                                cudaMemcpyHostToDevice, st));
    YGZ_CUDA(cudaMemcpyAsync(dNE, nE.data(), nb * 4, cudaMemcpyHostToDevice, st));
    DsoArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = n_frames; A.w = w; A.h = hgt; A.pitch = ex->G.lv[0].pitch; A.nlevels = ex->G.nlevels; A.cap = cap;
    A.frameStride = ex->B.inFrameStride; A.img = ex->B.in;
    A.grid = dGrid; A.active = dActive; A.nExist = dNE; A.kps = dKps;
    A.occ = ex->dsoOcc; A.occPitch = occPitch; A.occWords = (long long)occPitch * hgt;
    A.maxCells = maxCells; A.cellKp = ex->dsoCellKp; A.cellCnt = ex->dsoCellCnt; A.total = dTotal; A.counts = dCnt;
    A.centre = ex->dsoCentre;
    for (int l = 0; l < ex->G.nlevels; l++) A.invScale[l] = ex->tables.invScale[l];
    YGZ_CUDA(cudaMemsetAsync(ex->dsoOcc, 0, nb * occPitch * hgt * 4, st));
    ex->launches += launchDsoOccupancy(A, maxExist, st);
    // ---- shrink the grid until a frame has its nfeatures corners (or the grid bottoms out at 7)
    std::vector<int32_t> total(n_frames);
    for (;;) {
        int maxGrid = 0, maxCellsNow = 0, nActive = 0;
        for (int f = 0; f < n_frames; f++)
            if (active[f]) {
                used[f] = state[f];
                maxGrid = std::max(maxGrid, state[f]);
                maxCellsNow = std::max(maxCellsNow, (hgt / state[f]) * (w / state[f]));
                nActive++;
            }
        if (!nActive) break;
        YGZ_CUDA(cudaMemcpyAsync(dGrid, used.data(), nb * 4, cudaMemcpyHostToDevice, st));
        YGZ_CUDA(cudaMemcpyAsync(dActive, active.data(), nb, cudaMemcpyHostToDevice, st));
        YGZ_CUDA(cudaMemsetAsync(dTotal, 0, nb * 4, st));
        ex->launches += launchDsoCells(A, maxGrid, maxCellsNow, st);
        YGZ_CUDA(cudaMemcpyAsync(total.data(), dTotal, nb * 4, cudaMemcpyDeviceToHost, st));
        YGZ_CUDA(cudaStreamSynchronize(st));
        for (int f = 0; f < n_frames; f++) {
            if (!active[f]) continue;
            cnt[f] = total[f];
            // cnt == 0 would repeat the same grid forever in the reference (:1301 shrinks only when cnt > 0): stop instead
            if (cnt[f] >= nfeat || cnt[f] == 0) { active[f] = 0; continue; }
            state[f] -= 5;
            if (state[f] < 7) { state[f] = 7; active[f] = 0; }      // :1303-1307: keeps the keypoints of the previous grid
        }
    }
    for (int f = 0; f < n_frames; f++) {
        if (cnt[f] > nfeat) state[f] += 5;                          // :1378-1380
        grid_size[f] = state[f];
    }
    YGZ_CUDA(cudaMemcpyAsync(dGrid, used.data(), nb * 4, cudaMemcpyHostToDevice, st));
    ex->launches += launchDsoAssemble(A, st);
    ex->launches += launchDescribeList(ex->G, ex->B, n_frames, cap, ex->dsoCentre, dCnt, dKps, dDesc, st);
    if (!inDev) YGZ_CUDA(cudaEventRecord(ex->inFree[slot], st));
    if (!outDev) {
        YGZ_CUDA(cudaMemcpyAsync(counts, dCnt, nb * 4, cudaMemcpyDeviceToHost, st));
        YGZ_CUDA(cudaStreamSynchronize(st));
        for (int f = 0; f < n_frames; f++) {
            const size_t c = (size_t)std::min(counts[f], cap);
            if (!c) continue;
            YGZ_CUDA(cudaMemcpyAsync(keypoints + (size_t)f * cap, dKps + (size_t)f * cap, c * sizeof(ygzorb_keypoint), cudaMemcpyDeviceToHost, st));
            YGZ_CUDA(cudaMemcpyAsync(desc + (size_t)f * cap * 32, dDesc + (size_t)f * cap * 32, c * 32, cudaMemcpyDeviceToHost, st));
        }
    }
    YGZ_CUDA(cudaGetLastError());
    ex->lastFrames = n_frames; ex->stagesValid = false;
    YGZ_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int ygzorb_describe_existing(ygzorb_extractor *h, int frame, int n, const ygzorb_keypoint *keypoints, uint8_t *desc) {
    Extractor *ex = (Extractor *)h;
    if (!ex || (n > 0 && (!keypoints || !desc))) return argFail("null argument");
    if (n < 0) return argFail("n < 0");
    if (n == 0) return 0;
    if (!ex->stagesValid || ex->lastFrames <= 0) { setError("no blurred pyramid: call ygzorb_extract_batch first"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames) return argFail("frame out of range");
    YGZ_CUDA(cudaSetDevice(ex->device));
    // :1091-1096 — tmp.pt *= mvInvScaleFactor[tmp.octave] (float), then computeOrbDescriptor reads at cvRound(pt) (:111)
    std::vector<int4> centre((size_t)n);
    for (int i = 0; i < n; i++) {
        const int oct = keypoints[i].octave;
        if (oct < 0 || oct >= ex->G.nlevels) return argFail("keypoint octave outside the pyramid");
        const float inv = ex->tables.invScale[oct];
        centre[i] = make_int4(oct, cvRoundF(keypoints[i].x * inv), cvRoundF(keypoints[i].y * inv), 0);
    }
    int rc;
    const size_t nb = (size_t)n;
    if ((rc = growBuf(ex->dsoCentre, ex->dsoCentreBytes, nb * sizeof(int4)))) return rc;
    if ((rc = growBuf(ex->dsoKps, ex->dsoKpsBytes, nb * sizeof(ygzorb_keypoint)))) return rc;
    if ((rc = growBuf(ex->dsoDesc, ex->dsoDescBytes, nb * 32))) return rc;
    if ((rc = growBuf(ex->dsoSmall, ex->dsoSmallBytes, (size_t)ex->maxBatch * 20))) return rc;
    cudaStream_t st = ex->stream;
    YGZ_CUDA(cudaMemcpyAsync(ex->dsoCentre, centre.data(), nb * sizeof(int4), cudaMemcpyHostToDevice, st));
    YGZ_CUDA(cudaMemcpyAsync(ex->dsoKps, keypoints, nb * sizeof(ygzorb_keypoint), cudaMemcpyHostToDevice, st));
    YGZ_CUDA(cudaMemcpyAsync(ex->dsoSmall, &n, sizeof(int32_t), cudaMemcpyHostToDevice, st));
    // one "frame" of n records whose level buffers are those of `frame`: shift the frame bases instead of the kernel's index
    DevBuffers B = ex->B;
    B.in = ex->B.in + (long long)frame * ex->B.inFrameStride;
    B.pyr = ex->B.pyr + (long long)frame * ex->G.slabBytes;
    B.blur = ex->B.blur + (long long)frame * ex->B.blurSlab;
    ex->launches += launchDescribeList(ex->G, B, 1, n, ex->dsoCentre, ex->dsoSmall, ex->dsoKps, ex->dsoDesc, st, true);
    YGZ_CUDA(cudaMemcpyAsync(desc, ex->dsoDesc, nb * 32, cudaMemcpyDeviceToHost, st));
    YGZ_CUDA(cudaGetLastError());
    YGZ_CUDA(cudaStreamSynchronize(st));
    return 0;
}

int ygzorb_compute_pyramid_batch(ygzorb_extractor *h, const uint8_t *frames, int n_frames, int w, int hgt,
                                 int stride, size_t frame_stride, unsigned flags) {
    Extractor *ex = (Extractor *)h;
    int rc = checkFrames(ex, frames, n_frames, w, hgt, stride);
    if (rc) return rc;
    if (w <= 0 || hgt <= 0 || n_frames == 0) return 0;
    if (n_frames > ex->maxBatch) return argFail("n_frames exceeds max_batch (the pyramid of every frame must stay resident)");
    YGZ_CUDA(cudaSetDevice(ex->device));
    if (frame_stride == 0) frame_stride = (size_t)stride * hgt;
    const bool inDev = flags & YGZORB_IN_DEVICE;
    if ((rc = ensureGeometry(ex, w, hgt, (inDev && !ex->dMap1) ? stride : w, n_frames, inDev))) return rc;
    int slot = 0;
    if (!inDev && (rc = uploadInput(ex, frames, n_frames, w, hgt, stride, frame_stride, &slot))) return rc;
    if ((rc = bindInput(ex, frames, w, hgt, frame_stride, inDev, slot))) return rc;
    if ((rc = applyUndistort(ex, n_frames, w, hgt, inDev ? stride : w))) return rc;
    runPyramid(ex, n_frames);
    if ((rc = sinkPyramid(ex, n_frames))) return rc;
    if (!inDev) YGZ_CUDA(cudaEventRecord(ex->inFree[slot], ex->stream));
    YGZ_CUDA(cudaGetLastError());
    ex->lastFrames = n_frames; ex->stagesValid = false;
    if (!(flags & YGZORB_NO_SYNC)) YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    return 0;
}

static int copyLevelOut(Extractor *ex, const uint8_t *src, int pitch, int w, int hh, uint8_t *dst) {
    YGZ_CUDA(cudaMemcpy2DAsync(dst, w, src, pitch, w, hh, cudaMemcpyDeviceToHost, ex->stream));
    return 0;
}

int ygzorb_get_pyramid_level(ygzorb_extractor *h, int frame, int level, uint8_t *dst) {
    Extractor *ex = (Extractor *)h;
    if (!ex || !dst) return argFail("null argument");
    if (ex->lastFrames <= 0) { setError("no pyramid computed yet"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames || level < 0 || level >= ex->G.nlevels) return argFail("frame/level out of range");
    YGZ_CUDA(cudaSetDevice(ex->device));
    const LevelGeom &L = ex->G.lv[level];
    const uint8_t *src = level == 0 ? ex->B.in + (long long)frame * ex->B.inFrameStride
                                    : ex->B.pyr + (long long)frame * ex->G.slabBytes + L.off;
    int rc = copyLevelOut(ex, src, L.pitch, L.w, L.h, dst);
    if (rc) return rc;
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    return 0;
}

int ygzorb_get_pyramid(ygzorb_extractor *h, int frame, uint8_t *dst) {
    Extractor *ex = (Extractor *)h;
    if (!ex || !dst) return argFail("null argument");
    if (ex->lastFrames <= 0) { setError("no pyramid computed yet"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames) return argFail("frame out of range");
    YGZ_CUDA(cudaSetDevice(ex->device));
    size_t off = 0;
    for (int l = 0; l < ex->G.nlevels; l++) {
        const LevelGeom &L = ex->G.lv[l];
        const uint8_t *src = l == 0 ? ex->B.in + (long long)frame * ex->B.inFrameStride
                                    : ex->B.pyr + (long long)frame * ex->G.slabBytes + L.off;
        int rc = copyLevelOut(ex, src, L.pitch, L.w, L.h, dst + off);
        if (rc) return rc;
        off += (size_t)L.w * L.h;
    }
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    return 0;
}

int ygzorb_extractor_set_pyramid_sink(ygzorb_extractor *h, uint8_t *host_dst) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return argFail("null extractor");
    YGZ_CUDA(cudaSetDevice(ex->device));
    YGZ_CUDA(cudaStreamSynchronize(ex->d2h));
    ex->pyrSink = host_dst;
    return 0;
}

int ygzorb_host_alloc(size_t bytes, void **out) {
    if (!out || bytes == 0) return argFail("bad argument");
    YGZ_CUDA(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
    return 0;
}

void ygzorb_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

int ygzorb_get_blurred_level(ygzorb_extractor *h, int frame, int level, uint8_t *dst) {
    Extractor *ex = (Extractor *)h;
    if (!ex || !dst) return argFail("null argument");
    if (!ex->stagesValid) { setError("no extraction ran yet"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames || level < 0 || level >= ex->G.nlevels) return argFail("frame/level out of range");
    YGZ_CUDA(cudaSetDevice(ex->device));
    const LevelGeom &L = ex->G.lv[level];
    int rc = copyLevelOut(ex, ex->B.blur + (long long)frame * ex->blurSlab + ex->blurOff[level], ex->blurPitch[level], L.w, L.h, dst);
    if (rc) return rc;
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    return 0;
}

int ygzorb_get_candidates(ygzorb_extractor *h, int frame, int level, int cap, int16_t *x, int16_t *y,
                          uint8_t *response, uint32_t *order_key) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return argFail("null extractor");
    if (!ex->stagesValid) { setError("no extraction ran yet"); return YGZORB_E_STATE; }
    if (frame < 0 || frame >= ex->lastFrames || level < 0 || level >= ex->G.nlevels) return argFail("frame/level out of range");
    YGZ_CUDA(cudaSetDevice(ex->device));
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    int n = 0;
    YGZ_CUDA(cudaMemcpy(&n, ex->B.candCount + frame * kMaxLevels + level, sizeof n, cudaMemcpyDeviceToHost));
    n = std::min(n, ex->G.lv[level].candCap);
    std::vector<uint2> tmp((size_t)std::max(n, 1));
    if (n > 0)
        YGZ_CUDA(cudaMemcpy(tmp.data(), ex->B.cand + (long long)frame * ex->G.candPerFrame + ex->G.lv[level].candOff,
                            (size_t)n * sizeof(uint2), cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && i < cap; i++) {
        if (x) x[i] = (int16_t)((int)(tmp[i].x & 0xfff) - kMinBorder);
        if (y) y[i] = (int16_t)((int)((tmp[i].x >> 12) & 0xfff) - kMinBorder);
        if (response) response[i] = (uint8_t)(tmp[i].x >> 24);
        if (order_key) order_key[i] = tmp[i].y;
    }
    return n;
}

int ygzorb_extractor_set_undistort_maps(ygzorb_extractor *h, const int16_t *map1, const uint16_t *map2, int w, int hgt) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return argFail("null extractor");
    YGZ_CUDA(cudaSetDevice(ex->device));
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    const bool had = ex->dMap1 != nullptr;
    if (!map1 || !map2) {
        if (ex->dMap1) cudaFree(ex->dMap1);
        if (ex->dMap2) cudaFree(ex->dMap2);
        ex->dMap1 = nullptr; ex->dMap2 = nullptr; ex->map1Bytes = ex->map2Bytes = 0; ex->mapW = ex->mapH = 0;
    } else {
        if (w < 2 || hgt < 2) return argFail("bad map size");
        const size_t n = (size_t)w * hgt;
        int rc;
        if ((rc = growBuf(ex->dMap1, ex->map1Bytes, n * sizeof(short2)))) return rc;
        if ((rc = growBuf(ex->dMap2, ex->map2Bytes, n * sizeof(uint16_t)))) return rc;
        YGZ_CUDA(cudaMemcpy(ex->dMap1, map1, n * sizeof(short2), cudaMemcpyHostToDevice));
        YGZ_CUDA(cudaMemcpy(ex->dMap2, map2, n * sizeof(uint16_t), cudaMemcpyHostToDevice));
        ex->mapW = w; ex->mapH = hgt;
    }
    if (had != (ex->dMap1 != nullptr)) { ex->gw = -1; ex->stagesValid = false; }   // level-0 pitch changes: rebuild the geometry
    return 0;
}

int ygzorb_extractor_sync(ygzorb_extractor *h) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return argFail("null extractor");
    YGZ_CUDA(cudaSetDevice(ex->device));
    YGZ_CUDA(cudaStreamSynchronize(ex->stream));
    YGZ_CUDA(cudaStreamSynchronize(ex->d2h));      // host outputs of the last call(s) are complete after this
    int rc = finishProfile(ex);
    if (rc) return rc;
    return ex->B.status ? checkStatus(ex) : 0;
}

int ygzorb_extractor_set_pipeline(ygzorb_extractor *h, int frames_per_chunk) {
    Extractor *ex = (Extractor *)h;
    if (!ex) return argFail("null extractor");
    if (frames_per_chunk < 0) return argFail("frames_per_chunk < 0");
    ex->pipeFrames = frames_per_chunk;
    return 0;
}

void *ygzorb_extractor_stream(ygzorb_extractor *h) { return h ? (void *)((Extractor *)h)->stream : nullptr; }
long long ygzorb_extractor_launch_count(const ygzorb_extractor *h) { return h ? ((const Extractor *)h)->launches : 0; }

int ygzorb_extractor_set_profiling(ygzorb_extractor *h, int on) {
    if (!h) return argFail("null extractor");
    ((Extractor *)h)->profiling = on != 0;
    return 0;
}
int ygzorb_extractor_stage_ms(ygzorb_extractor *h, float *ms, int n) {
    Extractor *ex = (Extractor *)h;
    if (!ex || !ms) return argFail("null argument");
    for (int i = 0; i < n && i < 8; i++) ms[i] = ex->stageMs[i];
    return 0;
}

}  // extern "C"

// ---- stand-alone stages (single host image; allocate, run one kernel, free) --------------------
namespace {
struct TmpDev {
    std::vector<void *> ptrs;
    ~TmpDev() { for (void *p : ptrs) cudaFree(p); }
    template <class T> cudaError_t alloc(T **p, size_t bytes) {
        cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(bytes, 256));
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};
}  // namespace

extern "C" int ygzorb_resize_linear_u8(const uint8_t *src, int sw, int sh, uint8_t *dst, int dw, int dh, int device) {
    if (!src || !dst) return argFail("null argument");
    if (sw < 1 || sh < 1 || dw < 1 || dh < 1 || sw > 65535 || sh > 65535) return argFail("bad size");
    YGZ_CUDA(cudaSetDevice(device));
    FrameGeom G; memset(&G, 0, sizeof G);
    DevBuffers B; memset(&B, 0, sizeof B);
    G.nlevels = 2; G.w = sw; G.h = sh;
    G.lv[0].w = sw; G.lv[0].h = sh; G.lv[0].pitch = sw;
    LevelGeom &D = G.lv[1];
    D.w = dw; D.h = dh; D.pitch = alignUp(dw, 64); D.off = 0;
    std::vector<int2> X, Y;
    D.area2x = makeResizeTables(sw, sh, dw, dh, D.pitch, X, Y) ? 1 : 0;
    D.tabX = 0; D.tabY = (int)X.size();
    X.insert(X.end(), Y.begin(), Y.end());
    G.slabBytes = (long long)D.pitch * dh;
    TmpDev tmp;
    uint8_t *dS = nullptr, *dD = nullptr; int2 *dT = nullptr;
    YGZ_CUDA(tmp.alloc(&dS, (size_t)sw * sh));
    YGZ_CUDA(tmp.alloc(&dD, (size_t)G.slabBytes));
    YGZ_CUDA(tmp.alloc(&dT, X.size() * sizeof(int2)));
    YGZ_CUDA(cudaMemcpy(dS, src, (size_t)sw * sh, cudaMemcpyHostToDevice));
    YGZ_CUDA(cudaMemcpy(dT, X.data(), X.size() * sizeof(int2), cudaMemcpyHostToDevice));
    B.in = dS; B.inFrameStride = (long long)sw * sh; B.pyr = dD; B.tabs = dT;
    launchResizeLevel(G, B, 1, 1, 0);
    YGZ_CUDA(cudaGetLastError());
    YGZ_CUDA(cudaMemcpy2D(dst, dw, dD, D.pitch, dw, dh, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ygzorb_gaussian_blur7_u8(const uint8_t *src, int w, int h, uint8_t *dst, int device) {
    if (!src || !dst) return argFail("null argument");
    if (w < 1 || h < 1) return argFail("bad size");
    YGZ_CUDA(cudaSetDevice(device));
    FrameGeom G; memset(&G, 0, sizeof G);
    DevBuffers B; memset(&B, 0, sizeof B);
    G.nlevels = 1; G.w = w; G.h = h;
    LevelGeom &L = G.lv[0];
    L.w = w; L.h = h; L.pitch = w;
    L.tilesX = (w + kBlurTW - 1) / kBlurTW; L.tilesY = (h + kBlurTH - 1) / kBlurTH; L.tileStart = 0;
    G.totalTiles = L.tilesX * L.tilesY;
    TmpDev tmp;
    uint8_t *dS = nullptr, *dD = nullptr;
    const int dp = alignUp(w, 64);
    YGZ_CUDA(tmp.alloc(&dS, (size_t)w * h));
    YGZ_CUDA(tmp.alloc(&dD, (size_t)dp * h));
    YGZ_CUDA(cudaMemcpy(dS, src, (size_t)w * h, cudaMemcpyHostToDevice));
    std::vector<int2> tileTab;
    for (int ty = 0; ty < L.tilesY; ty++)
        for (int tx = 0; tx < L.tilesX; tx++) tileTab.push_back(make_int2((tx * kBlurTW) | ((ty * kBlurTH) << 12), 0));
    int2 *dT = nullptr;
    YGZ_CUDA(tmp.alloc(&dT, tileTab.size() * sizeof(int2)));
    YGZ_CUDA(cudaMemcpy(dT, tileTab.data(), tileTab.size() * sizeof(int2), cudaMemcpyHostToDevice));
    G.tileTab = 0; B.tabs = dT;
    B.in = dS; B.inFrameStride = (long long)w * h; B.blur = dD; B.blurSlab = (long long)dp * h; B.blurPitch[0] = dp;
    launchBlur(G, B, 1, 0);
    YGZ_CUDA(cudaGetLastError());
    YGZ_CUDA(cudaMemcpy2D(dst, w, dD, dp, w, h, cudaMemcpyDeviceToHost));
    return 0;
}
