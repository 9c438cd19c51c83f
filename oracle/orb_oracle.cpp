// orb_oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement ("oracle") of the reference's ORB
// front-end hot path on flat arrays, exported with a C ABI for ctypes (tests/, smoke(), bench.py's
// cpu_baseline leg).  The product library never links, loads or calls this file.
//
// Pinned: (1) the OpenCV primitives in minicv.hpp against cv2 4.13.0 (tests/test_minicv_vs_cv2.py);
// (2) the extractor control flow below against the reference's own src/ORBextractor.cc compiled
// unmodified into oracle/_ref/ygz_ref_orb (tests/test_oracle_vs_ref.py, golden vectors under
// tests/golden/).  The matcher restatements (part 2) have no compilable reference here (ORBmatcher.cc
// / Frame.cc need Eigen, sophus, DBoW2::Vocabulary, OpenCV C++) -> "parity unpinned" for those.
//
// All citations are file:line in /root/reference.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <atomic>
#include <functional>
#include <numeric>
#include <thread>
#include <utility>
#include <vector>

#include "minicv.hpp"
#include "orb_pattern.inc"  // static const signed char kOrbPattern[256*4] — see tools/gen_pattern.py

#define YGZO_API extern "C" __attribute__((visibility("default")))

namespace {

// Dynamic work-sharing loop over [0, n) on `threads` std::threads (no OpenMP runtime needed).
template <class F> void parallelFor(int n, int threads, F body) {
    if (threads <= 1 || n <= 1) { for (int i = 0; i < n; i++) body(i); return; }
    std::atomic<int> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < std::min(threads, n); t++)
        th.emplace_back([&]() { for (;;) { const int i = next.fetch_add(1); if (i >= n) break; body(i); } });
    for (auto &t : th) t.join();
}

// ------------------------------------------------------------------------------------------------
// glibc 2.39 cosf/sinf, FMA variant (sysdeps/ieee754/flt-32/s_cosf.c, s_sinf.c, sincosf.h — the
// ARM optimized-routines algorithm).  The reference calls cos(float)/sin(float) under
// `using namespace std` (src/ORBextractor.cc:67,108-109), which resolves to std::cos(float) = cosf,
// NOT to a double cosine.  Restated here op for op (double arithmetic with the same fused
// multiply-adds the x86-64 FMA build executes) so that the CUDA kernel, which has IEEE double fma,
// can reproduce it bit for bit.  Valid for |x| < 120 (angles are radians in [0, 2pi]).
const double kHpiInv = 0x1.45f306dc9c883p+23;  // 2/pi * 2^24
const double kHpi = 0x1.921fb54442d18p+0;
const double kC0 = 1.0, kC1 = -0x1.ffffffd0c621cp-2, kC2 = 0x1.55553e1068f19p-5,
             kC3 = -0x1.6c087e89a359dp-10, kC4 = 0x1.99343027bf8c3p-16;
const double kS1 = -0x1.555545995a603p-3, kS2 = 0x1.1107605230bc4p-7, kS3 = -0x1.994eb3774cf24p-13;

inline uint32_t abstop12(float x) {
    uint32_t u;
    std::memcpy(&u, &x, 4);
    return (u >> 20) & 0x7ff;
}
// sinf_poly(x, x2, p, n): n even -> sine polynomial, n odd -> cosine polynomial; `neg` selects the
// second table (cosine coefficients negated, sincosf.h __sincosf_table[1]).
inline float sincosPoly(double x, double x2, int n, bool neg) {
    if ((n & 1) == 0) {
        const double x3 = x * x2;
        const double s1 = std::fma(x2, kS3, kS2);
        const double x7 = x3 * x2;
        const double s = std::fma(x3, kS1, x);
        return (float)std::fma(x7, s1, s);
    }
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = x2 * x2;
    const double c2 = std::fma(x2, sg * kC4, sg * kC3);
    const double c1 = std::fma(x2, sg * kC1, sg * kC0);
    const double x6 = x4 * x2;
    const double c = std::fma(x4, sg * kC2, c1);
    return (float)std::fma(x6, c2, c);
}
inline double reduceFast(double x, int *np) {
    const double r = x * kHpiInv;
    const int n = ((int32_t)r + 0x800000) >> 24;
    *np = n;
    return std::fma(-(double)n, kHpi, x);
}
const double kSign[4] = {1.0, -1.0, -1.0, 1.0};

float refCosf(float y) {
    double x = y;
    if (abstop12(y) < 0x3f4) {  // |y| < pi/4
        const double x2 = x * x;
        if (abstop12(y) < 0x398) return 1.0f;  // |y| < 2^-12
        return sincosPoly(x, x2, 1, false);
    }
    int n;
    x = reduceFast(x, &n);
    const double s = kSign[n & 3];
    return sincosPoly(x * s, x * x, n ^ 1, (n & 2) != 0);
}
float refSinf(float y) {
    double x = y;
    if (abstop12(y) < 0x3f4) {
        const double x2 = x * x;
        if (abstop12(y) < 0x398) return y;
        return sincosPoly(x, x2, 0, false);
    }
    int n;
    x = reduceFast(x, &n);
    const double s = kSign[n & 3];
    return sincosPoly(x * s, x * x, n, (n & 2) != 0);
}

// ------------------------------------------------------------------------------------------------
struct Params { int nfeatures; float scaleFactor; int nlevels; int iniThFAST; int minThFAST; };

const int kEdge = 19, kHalfPatch = 15, kPatch = 31;

// src/ORBextractor.cc:412-470 — per-level tables.  NB the member `scaleFactor` is a double
// initialised from the float argument (include/ORBextractor.h:157).
struct Tables {
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> nPerLevel;
    int umax[kHalfPatch + 1];
};

Tables makeTables(const Params &p) {
    Tables t;
    const int L = p.nlevels;
    const double sf = (double)p.scaleFactor;
    t.scale.assign(L, 1.f); t.sigma2.assign(L, 1.f);
    for (int i = 1; i < L; i++) {
        t.scale[i] = (float)((double)t.scale[i - 1] * sf);
        t.sigma2[i] = t.scale[i] * t.scale[i];
    }
    t.invScale.resize(L); t.invSigma2.resize(L);
    for (int i = 0; i < L; i++) { t.invScale[i] = 1.0f / t.scale[i]; t.invSigma2[i] = 1.0f / t.sigma2[i]; }
    t.nPerLevel.assign(L, 0);
    const float factor = (float)(1.0f / sf);
    float nDesired = (float)(p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L)));
    int sum = 0;
    for (int l = 0; l < L - 1; l++) {
        t.nPerLevel[l] = cv::cvRound(nDesired);
        sum += t.nPerLevel[l];
        nDesired *= factor;
    }
    t.nPerLevel[L - 1] = std::max(p.nfeatures - sum, 0);
    // :451-466 umax
    const int vmax = cv::cvFloor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = cv::cvCeil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) t.umax[v] = cv::cvRound(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (t.umax[v0] == t.umax[v0 + 1]) ++v0;
        t.umax[v] = v0;
        ++v0;
    }
    return t;
}

// src/ORBextractor.cc:1129-1133 — level sizes from LEVEL-0 dims.
void levelSize(const Tables &t, int w, int h, int l, int *lw, int *lh) {
    *lw = cv::cvRound((float)w * t.invScale[l]);
    *lh = cv::cvRound((float)h * t.invScale[l]);
}

// src/ORBextractor.cc:1129-1150 — pyramid; the 19-px REFLECT_101 frame is not produced (never read
// by the OctTree path, dropped by Frame.cc:810-813).
std::vector<cv::Mat> buildPyramid(const Tables &t, const Params &p, const cv::Mat &img) {
    std::vector<cv::Mat> pyr(p.nlevels);
    for (int l = 0; l < p.nlevels; l++) {
        int lw, lh;
        levelSize(t, img.cols, img.rows, l, &lw, &lh);
        if (l == 0) pyr[0] = img.clone();
        else cv::resize(pyr[l - 1], pyr[l], cv::Size(lw, lh), 0, 0, cv::INTER_LINEAR);
    }
    return pyr;
}

struct Cand { short x, y; unsigned char resp; };  // coordinates relative to (16,16), like vToDistributeKeys

// src/ORBextractor.cc:728-781 — per-cell FAST with threshold fallback, cell-major / raster order.
void fastCandidates(const cv::Mat &lvl, int iniTh, int minTh, std::vector<Cand> &out) {
    out.clear();
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = lvl.cols - kEdge + 3, maxBY = lvl.rows - kEdge + 3;
    const float W = 30;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / W), nRows = (int)(height / W);
    if (nCols <= 0 || nRows <= 0) return;  // canonicalised: ceil(x/0) in the reference, no cells run
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    std::vector<cv::KeyPoint> cellKeys;
    for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = (float)maxBY;
        for (int j = 0; j < nCols; j++) {
            const float iniX = (float)(minBX + j * wCell);
            float maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = (float)maxBX;
            cv::Mat cell = lvl.rowRange((int)iniY, (int)maxY).colRange((int)iniX, (int)maxX);
            cv::FAST(cell, cellKeys, iniTh, true);
            if (cellKeys.empty()) cv::FAST(cell, cellKeys, minTh, true);
            for (const cv::KeyPoint &k : cellKeys) {
                Cand c;
                c.x = (short)(k.pt.x + j * wCell);
                c.y = (short)(k.pt.y + i * hCell);
                c.resp = (unsigned char)k.response;
                out.push_back(c);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Quadtree distribution — src/ORBextractor.cc:479-723 (DivideNode, DistributeOctTree).
// Nodes live in a pool; pool index == creation order, which is the canonical stand-in for the
// reference's pointer tie-break (:653-657; see oracle/ref_cli.cpp).  Each node owns a contiguous
// slice of `perm` (candidate indices); a split is a stable 4-way partition of that slice, so the
// children keep the parent's key order exactly like the push_back loop at :512-523.
struct QNode {
    int x0, x1, y0, y1;   // UL.x, UR.x, UL.y, BL.y
    int begin, end;       // slice of perm
    int prev, next;       // list links (-1 = none)
};

struct QTree {
    std::vector<QNode> pool;
    std::vector<int> perm, tmp;
    int head = -1, tail = -1, count = 0;
    const short *xs, *ys;

    int pushFront(const QNode &n) {
        pool.push_back(n);
        const int id = (int)pool.size() - 1;
        pool[id].prev = -1; pool[id].next = head;
        if (head >= 0) pool[head].prev = id; else tail = id;
        head = id; count++;
        return id;
    }
    int pushBack(const QNode &n) {
        pool.push_back(n);
        const int id = (int)pool.size() - 1;
        pool[id].next = -1; pool[id].prev = tail;
        if (tail >= 0) pool[tail].next = id; else head = id;
        tail = id; count++;
        return id;
    }
    void erase(int id) {
        const int p = pool[id].prev, n = pool[id].next;
        if (p >= 0) pool[p].next = n; else head = n;
        if (n >= 0) pool[n].prev = p; else tail = p;
        count--;
    }
    // DivideNode (:479-531).  Children are appended to `kids` in n1..n4 order (empty ones included).
    void divide(int id, QNode kids[4]) {
        const QNode p = pool[id];
        const int halfX = (int)std::ceil((float)(p.x1 - p.x0) / 2);
        const int halfY = (int)std::ceil((float)(p.y1 - p.y0) / 2);
        const int mx = p.x0 + halfX, my = p.y0 + halfY;
        int cnt[4] = {0, 0, 0, 0};
        for (int i = p.begin; i < p.end; i++) {
            const int k = perm[i];
            const int q = ((float)xs[k] < (float)mx ? 0 : 1) + ((float)ys[k] < (float)my ? 0 : 2);
            cnt[q]++;
        }
        int off[4] = {p.begin, p.begin + cnt[0], p.begin + cnt[0] + cnt[1], p.begin + cnt[0] + cnt[1] + cnt[2]};
        int cur[4] = {off[0], off[1], off[2], off[3]};
        for (int i = p.begin; i < p.end; i++) tmp[i] = perm[i];
        for (int i = p.begin; i < p.end; i++) {
            const int k = tmp[i];
            const int q = ((float)xs[k] < (float)mx ? 0 : 1) + ((float)ys[k] < (float)my ? 0 : 2);
            perm[cur[q]++] = k;
        }
        kids[0] = QNode{p.x0, mx, p.y0, my, off[0], off[0] + cnt[0], -1, -1};    // n1 top-left
        kids[1] = QNode{mx, p.x1, p.y0, my, off[1], off[1] + cnt[1], -1, -1};    // n2 top-right
        kids[2] = QNode{p.x0, mx, my, p.y1, off[2], off[2] + cnt[2], -1, -1};    // n3 bottom-left
        kids[3] = QNode{mx, p.x1, my, p.y1, off[3], off[3] + cnt[3], -1, -1};    // n4 bottom-right
    }
};

// Returns indices (into the candidate arrays) of the kept keypoints, in the reference's list order.
std::vector<int> distributeQuadtree(const short *xs, const short *ys, const unsigned char *resp, int n,
                                    int minX, int maxX, int minY, int maxY, int N) {
    std::vector<int> result;
    if (n == 0) return result;
    QTree t;
    t.xs = xs; t.ys = ys;
    t.pool.reserve((size_t)4 * n + 16);
    t.perm.resize(n); t.tmp.resize(n);
    const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));  // :537
    const float hX = (float)(maxX - minX) / nIni;
    // roots (:545-556) and initial assignment (:559-562), stable per root
    std::vector<int> rootOf(n), rootCnt(std::max(nIni, 1), 0);
    for (int i = 0; i < n; i++) { rootOf[i] = (int)((float)xs[i] / hX); rootCnt[rootOf[i]]++; }
    std::vector<int> rootOff(nIni + 1, 0);
    for (int r = 0; r < nIni; r++) rootOff[r + 1] = rootOff[r] + rootCnt[r];
    { std::vector<int> cur(rootOff.begin(), rootOff.end() - 1);
      for (int i = 0; i < n; i++) t.perm[cur[rootOf[i]]++] = i; }
    for (int r = 0; r < nIni; r++) {
        QNode q{(int)(hX * (float)r), (int)(hX * (float)(r + 1)), 0, maxY - minY, rootOff[r], rootOff[r + 1], -1, -1};
        t.pushBack(q);
    }
    for (int id = t.head; id >= 0;) {  // :566-574 drop empty roots
        const int nx = t.pool[id].next;
        if (t.pool[id].begin == t.pool[id].end) t.erase(id);
        id = nx;
    }
    auto sz = [&](int id) { return t.pool[id].end - t.pool[id].begin; };
    std::vector<std::pair<int, int>> sizeAndNode, prevList;  // (size, creation index)
    bool finish = false;
    QNode kids[4];
    auto splitInto = [&](int id, bool countExpand, int *nToExpand) {
        t.divide(id, kids);
        for (int c = 0; c < 4; c++) {
            const int s = kids[c].end - kids[c].begin;
            if (s > 0) {
                const int cid = t.pushFront(kids[c]);
                if (s > 1) {
                    if (countExpand) (*nToExpand)++;
                    sizeAndNode.push_back(std::make_pair(s, cid));
                }
            }
        }
        t.erase(id);
    };
    while (!finish) {
        const int prevSize = t.count;
        int nToExpand = 0;
        sizeAndNode.clear();
        for (int id = t.head; id >= 0;) {       // :591-643; children go to the front, never revisited
            const int nx = t.pool[id].next;
            if (sz(id) > 1) splitInto(id, true, &nToExpand);
            id = nx;
        }
        if (t.count >= N || t.count == prevSize) {
            finish = true;
        } else if (t.count + nToExpand * 3 > N) {  // :649-699
            while (!finish) {
                const int prevSize2 = t.count;
                prevList = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prevList.begin(), prevList.end());
                for (int j = (int)prevList.size() - 1; j >= 0; j--) {
                    splitInto(prevList[j].second, false, nullptr);
                    if (t.count >= N) break;
                }
                if (t.count >= N || t.count == prevSize2) finish = true;
            }
        }
    }
    // :703-720 best response per node, first wins on ties
    result.reserve(t.count);
    for (int id = t.head; id >= 0; id = t.pool[id].next) {
        const QNode &q = t.pool[id];
        int best = t.perm[q.begin];
        for (int i = q.begin + 1; i < q.end; i++)
            if (resp[t.perm[i]] > resp[best]) best = t.perm[i];
        result.push_back(best);
    }
    return result;
}

// ------------------------------------------------------------------------------------------------
// IC_Angle — src/ORBextractor.cc:77-101, on the UNBLURRED level at integer (x, y).
float icAngle(const cv::Mat &img, int x, int y, const int *umax) {
    int m01 = 0, m10 = 0;
    const unsigned char *c = img.ptr(y) + x;
    const int step = (int)img.step;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int p = c[u + v * step], m = c[u - v * step];
            vsum += p - m;
            m10 += u * (p + m);
        }
        m01 += v * vsum;
    }
    return cv::fastAtan2((float)m01, (float)m10);
}

// computeOrbDescriptor — src/ORBextractor.cc:103-149, on the BLURRED level.
void orbDescriptor(const cv::Mat &blur, int x, int y, float angleDeg, unsigned char *desc) {
    const float factorPI = (float)(CV_PI / 180.f);
    const float angle = angleDeg * factorPI;
    const float a = refCosf(angle), b = refSinf(angle);
    const unsigned char *c = blur.ptr(y) + x;
    const int step = (int)blur.step;
    const signed char *pat = kOrbPattern;
    for (int i = 0; i < 32; i++, pat += 32) {
        int val = 0;
        for (int j = 0; j < 8; j++) {
            const float x0 = pat[4 * j], y0 = pat[4 * j + 1], x1 = pat[4 * j + 2], y1 = pat[4 * j + 3];
            const int t0 = c[cv::cvRound(x0 * b + y0 * a) * step + cv::cvRound(x0 * a - y0 * b)];
            const int t1 = c[cv::cvRound(x1 * b + y1 * a) * step + cv::cvRound(x1 * a - y1 * b)];
            val |= (t0 < t1) << j;
        }
        desc[i] = (unsigned char)val;
    }
}

struct KeyOut { float x, y, size, angle, response; int octave, class_id; };
static_assert(sizeof(KeyOut) == 28, "cv::KeyPoint layout");

// operator() — src/ORBextractor.cc:970-1028 (ComputePyramid -> ComputeKeyPointsOctTree ->
// per-level blur + descriptors -> scale + append).
int extractFrame(const Params &p, const cv::Mat &img, std::vector<KeyOut> &keys, std::vector<unsigned char> &desc) {
    keys.clear(); desc.clear();
    if (img.empty()) return 0;
    const Tables t = makeTables(p);
    std::vector<cv::Mat> pyr = buildPyramid(t, p, img);
    std::vector<Cand> cand;
    std::vector<short> xs, ys;
    std::vector<unsigned char> rs;
    for (int l = 0; l < p.nlevels; l++) {
        const cv::Mat &lvl = pyr[l];
        fastCandidates(lvl, p.iniThFAST, p.minThFAST, cand);
        const int n = (int)cand.size();
        xs.resize(n); ys.resize(n); rs.resize(n);
        for (int i = 0; i < n; i++) { xs[i] = cand[i].x; ys[i] = cand[i].y; rs[i] = cand[i].resp; }
        const int minB = kEdge - 3;
        std::vector<int> keep = distributeQuadtree(xs.data(), ys.data(), rs.data(), n, minB, lvl.cols - kEdge + 3,
                                                   minB, lvl.rows - kEdge + 3, t.nPerLevel[l]);
        if (keep.empty()) continue;
        const int scaledPatch = (int)(kPatch * t.scale[l]);  // :787
        cv::Mat blur;
        cv::GaussianBlur(lvl, blur, cv::Size(7, 7), 2, 2, cv::BORDER_REFLECT_101);  // :1009-1010
        for (int k : keep) {
            KeyOut o;
            const int x = xs[k] + minB, y = ys[k] + minB;
            o.angle = icAngle(lvl, x, y, t.umax);
            o.size = (float)scaledPatch; o.response = (float)rs[k]; o.octave = l; o.class_id = -1;
            desc.resize(desc.size() + 32);
            orbDescriptor(blur, x, y, o.angle, &desc[desc.size() - 32]);
            o.x = (float)x; o.y = (float)y;
            if (l != 0) { o.x *= t.scale[l]; o.y *= t.scale[l]; }  // :1019-1024
            keys.push_back(o);
        }
    }
    return (int)keys.size();
}

cv::Mat wrap(const unsigned char *img, int w, int h, int stride) {
    return cv::Mat(h, w, cv::CV_8UC1, (void *)img, (size_t)stride);
}
void packTo(const cv::Mat &m, unsigned char *out) {
    for (int r = 0; r < m.rows; r++) std::memcpy(out + (size_t)r * m.cols, m.ptr(r), (size_t)m.cols);
}

}  // namespace

// ================================================================================================
// C ABI (ctypes).  All images 8-bit single channel; outputs are caller-allocated.
struct ygzo_params { int nfeatures; float scaleFactor; int nlevels; int iniThFAST; int minThFAST; };
static Params toP(const ygzo_params *p) { return Params{p->nfeatures, p->scaleFactor, p->nlevels, p->iniThFAST, p->minThFAST}; }

YGZO_API void ygzo_tables(const ygzo_params *p, float *scale, float *invScale, float *sigma2, float *invSigma2,
                          int *nPerLevel, int *umax16) {
    Tables t = makeTables(toP(p));
    for (int l = 0; l < p->nlevels; l++) {
        scale[l] = t.scale[l]; invScale[l] = t.invScale[l]; sigma2[l] = t.sigma2[l]; invSigma2[l] = t.invSigma2[l];
        nPerLevel[l] = t.nPerLevel[l];
    }
    for (int v = 0; v < 16; v++) umax16[v] = t.umax[v];
}
YGZO_API void ygzo_level_sizes(const ygzo_params *p, int w, int h, int *lw, int *lh) {
    Tables t = makeTables(toP(p));
    for (int l = 0; l < p->nlevels; l++) levelSize(t, w, h, l, &lw[l], &lh[l]);
}
YGZO_API void ygzo_resize(const unsigned char *src, int sw, int sh, int sstride, unsigned char *dst, int dw, int dh) {
    cv::Mat d;
    cv::resize(wrap(src, sw, sh, sstride), d, cv::Size(dw, dh), 0, 0, cv::INTER_LINEAR);
    packTo(d, dst);
}
// cv::remap(src, dst, map1 CV_16SC2, map2 CV_16UC1, INTER_LINEAR), BORDER_CONSTANT 0, 8UC1 — the undistortion of
// Frame::ComputeImagePyramid (src/Frame.cc:775-790) — restated from OpenCV's fixed-point remapBilinear: 5 fractional bits per
// axis, bilinear weights of 15 bits summing to 1 << 15 ((32-fx)(32-fy)*32, ...), FixedPtCast (sum + (1 << 14)) >> 15; a tap
// outside the source is the border value.  tests/test_minicv_vs_cv2.py checks it against cv2.remap.
YGZO_API void ygzo_remap_u8(const unsigned char *src, int w, int h, int stride, const short *map1, const unsigned short *map2,
                            int dw, int dh, unsigned char *dst) {
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            const size_t o = (size_t)y * dw + x;
            const int sx = map1[2 * o], sy = map1[2 * o + 1], idx = map2[o] & 1023, fx = idx & 31, fy = idx >> 5;
            const int wt[4] = {(32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32, (32 - fx) * fy * 32, fx * fy * 32};
            int acc = 0;
            for (int k = 0; k < 4; k++) {
                const int xx = sx + (k & 1), yy = sy + (k >> 1);
                const int p = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? src[(size_t)yy * stride + xx] : 0;
                acc += wt[k] * p;
            }
            dst[o] = (unsigned char)((acc + (1 << 14)) >> 15);
        }
}

YGZO_API void ygzo_blur7(const unsigned char *src, int w, int h, int stride, unsigned char *dst) {
    cv::Mat d;
    cv::GaussianBlur(wrap(src, w, h, stride), d, cv::Size(7, 7), 2, 2, cv::BORDER_REFLECT_101);
    packTo(d, dst);
}
YGZO_API void ygzo_border101(const unsigned char *src, int w, int h, int stride, int b, unsigned char *dst) {
    cv::Mat d;
    cv::copyMakeBorder(wrap(src, w, h, stride), d, b, b, b, b, cv::BORDER_REFLECT_101);
    packTo(d, dst);
}
YGZO_API int ygzo_fast(const unsigned char *img, int w, int h, int stride, int threshold, int nms, int cap,
                       int *xs, int *ys, int *resp) {
    std::vector<cv::KeyPoint> k;
    cv::FAST(wrap(img, w, h, stride), k, threshold, nms != 0);
    for (int i = 0; i < (int)k.size() && i < cap; i++) { xs[i] = (int)k[i].pt.x; ys[i] = (int)k[i].pt.y; resp[i] = (int)k[i].response; }
    return (int)k.size();
}
YGZO_API void ygzo_fast_atan2(const float *y, const float *x, int n, float *out) {
    for (int i = 0; i < n; i++) out[i] = cv::fastAtan2(y[i], x[i]);
}
YGZO_API void ygzo_cossin(const float *a, int n, float *c, float *s) {
    for (int i = 0; i < n; i++) { c[i] = refCosf(a[i]); s[i] = refSinf(a[i]); }
}
// Compares the restated cosf/sinf against libm's over every float in [lo, hi]; returns mismatches.
YGZO_API long ygzo_cossin_check_libm(float lo, float hi, long stride) {
    uint32_t a, b;
    std::memcpy(&a, &lo, 4); std::memcpy(&b, &hi, 4);
    long bad = 0;
    for (uint64_t u = a; u <= b; u += (uint64_t)stride) {
        const uint32_t uu = (uint32_t)u;
        float f;
        std::memcpy(&f, &uu, 4);
        if (refCosf(f) != cosf(f) || refSinf(f) != sinf(f)) bad++;
    }
    return bad;
}
YGZO_API int ygzo_round_half_even(float v) { return cv::cvRound(v); }

// ---- stage-level entry points (one per CUDA kernel of the product, for stage-wise parity) ----
// Pyramid: levels packed back to back, each continuous (step == cols).
YGZO_API void ygzo_pyramid(const ygzo_params *p, const unsigned char *img, int w, int h, int stride, unsigned char *out) {
    const Params pp = toP(p);
    const Tables t = makeTables(pp);
    std::vector<cv::Mat> pyr = buildPyramid(t, pp, wrap(img, w, h, stride));
    size_t off = 0;
    for (int l = 0; l < pp.nlevels; l++) { packTo(pyr[l], out + off); off += (size_t)pyr[l].rows * pyr[l].cols; }
}
// Per-cell FAST candidates of one level image (coordinates relative to (16,16)); returns the count.
YGZO_API int ygzo_fast_candidates(const unsigned char *lvl, int w, int h, int stride, int iniTh, int minTh, int cap,
                                  short *xs, short *ys, unsigned char *resp) {
    std::vector<Cand> c;
    fastCandidates(wrap(lvl, w, h, stride), iniTh, minTh, c);
    for (int i = 0; i < (int)c.size() && i < cap; i++) { xs[i] = c[i].x; ys[i] = c[i].y; resp[i] = c[i].resp; }
    return (int)c.size();
}
YGZO_API int ygzo_distribute(const short *xs, const short *ys, const unsigned char *resp, int n, int minX, int maxX,
                             int minY, int maxY, int N, int cap, int *outIdx) {
    std::vector<int> k = distributeQuadtree(xs, ys, resp, n, minX, maxX, minY, maxY, N);
    for (int i = 0; i < (int)k.size() && i < cap; i++) outIdx[i] = k[i];
    return (int)k.size();
}
YGZO_API void ygzo_ic_angles(const unsigned char *lvl, int w, int h, int stride, const int *xs, const int *ys, int n, float *out) {
    ygzo_params dp{1000, 1.2f, 8, 20, 7};
    const Tables t = makeTables(toP(&dp));
    cv::Mat m = wrap(lvl, w, h, stride);
    for (int i = 0; i < n; i++) out[i] = icAngle(m, xs[i], ys[i], t.umax);
}
YGZO_API void ygzo_descriptors(const unsigned char *blurred, int w, int h, int stride, const int *xs, const int *ys,
                               const float *angles, int n, unsigned char *out) {
    cv::Mat m = wrap(blurred, w, h, stride);
    for (int i = 0; i < n; i++) orbDescriptor(m, xs[i], ys[i], angles[i], out + (size_t)i * 32);
}
// Whole extractor on one frame; keypoints as 28-byte cv::KeyPoint records.  Returns n (may exceed cap:
// only the first cap are written).
YGZO_API int ygzo_extract(const ygzo_params *p, const unsigned char *img, int w, int h, int stride, int cap,
                          void *keypoints, unsigned char *desc) {
    std::vector<KeyOut> k;
    std::vector<unsigned char> d;
    const int n = extractFrame(toP(p), wrap(img, w, h, stride), k, d);
    const int m = std::min(n, cap);
    if (m > 0) { std::memcpy(keypoints, k.data(), (size_t)m * 28); std::memcpy(desc, d.data(), (size_t)m * 32); }
    return n;
}
// Batch of same-size frames, OpenMP over frames (the timed "port" CPU baseline).  counts[f] keypoints per
// frame, records at frame stride cap.
YGZO_API void ygzo_extract_batch(const ygzo_params *p, const unsigned char *frames, int nframes, int w, int h, int cap,
                                 int threads, int *counts, void *keypoints, unsigned char *desc) {
    const Params pp = toP(p);
    parallelFor(nframes, threads, [&](int f) {
        std::vector<KeyOut> k;
        std::vector<unsigned char> d;
        const int n = extractFrame(pp, wrap(frames + (size_t)f * w * h, w, h, w), k, d);
        const int m = std::min(n, cap);
        counts[f] = n;
        if (keypoints && m > 0) {
            std::memcpy((char *)keypoints + (size_t)f * cap * 28, k.data(), (size_t)m * 28);
            std::memcpy(desc + (size_t)f * cap * 32, d.data(), (size_t)m * 32);
        }
    });
}
