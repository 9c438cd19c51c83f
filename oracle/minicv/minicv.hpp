// minicv.hpp — TEST INFRASTRUCTURE ONLY (CPU oracle). Never linked into the product library.
//
// A CPU restatement of exactly the OpenCV surface that the reference's ORB front-end calls
// (reference: src/ORBextractor.cc:58-61 includes; call sites :80,100,111,115-116,441,455-459,
// 765,768,1009-1010,1082-1083,1132-1146).  OpenCV itself is an un-vendored, un-pinned dependency of
// the reference (CMakeLists.txt:40-46) and its C++ headers/libs are absent from this image, so the
// pixel arithmetic is restated here from OpenCV's published algorithms and pinned, bit for bit,
// against cv2 4.13.0 (the OpenCV that *is* in the image, Python only) by tests/test_minicv_vs_cv2.py:
//   resize 8UC1 INTER_LINEAR (11-bit fixed point; exact 2x -> INTER_AREA), GaussianBlur 8U 7x7
//   (8.8 fixed point), FAST-9/16 + 3x3 NMS, copyMakeBorder REFLECT_101, fastAtan2, cvRound.
// The type surface (Mat, KeyPoint, Point_, InputArray ...) is just wide enough for the reference's
// src/ORBextractor.cc to compile UNMODIFIED against it (oracle/Makefile -> oracle/_ref/).
#ifndef YGZ_ORACLE_MINICV_HPP_
#define YGZ_ORACLE_MINICV_HPP_

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

namespace cv {

typedef unsigned char uchar;
typedef unsigned short ushort;

enum { CV_8U = 0, CV_8UC1 = 0, CV_32F = 5, CV_32FC1 = 5 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_AREA = 3 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_REFLECT_101 = 4,
       BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
#ifndef CV_PI
#define CV_PI 3.1415926535897932384626433832795
#endif

// cvRound: round-half-to-even (SSE cvtss2si / cvtsd2si under the default rounding mode).
static inline int cvRound(double v) { return (int)std::nearbyint(v); }
static inline int cvRound(float v) { return (int)std::nearbyintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(int v) { return v; }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(int v) { return v; }

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U> &p) : x((T)p.x), y((T)p.y) {}
    Point_ &operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_ &operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_ &operator*=(int s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
template <typename T> static inline Point_<T> operator*(const Point_<T> &p, float s) { return Point_<T>((T)(p.x * s), (T)(p.y * s)); }
template <typename T> static inline Point_<T> operator*(const Point_<T> &p, double s) { return Point_<T>((T)(p.x * s), (T)(p.y * s)); }
template <typename T> static inline Point_<T> operator*(const Point_<T> &p, int s) { return Point_<T>((T)(p.x * s), (T)(p.y * s)); }
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Scalar {   // only the single-value form (src/SparseImageAlign.cc:140)
    double v0;
    Scalar(double v = 0) : v0(v) {}
};
struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

class KeyPoint {
public:
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f _pt, float _size, float _angle = -1, float _response = 0, int _octave = 0,
             int _class_id = -1)
        : pt(_pt), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
             int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout is 28 bytes");

struct MatStep {
    size_t p[2];
    MatStep() { p[0] = p[1] = 0; }
    operator size_t() const { return p[0]; }
};

struct MatZerosExpr { int rows, cols, type; };
struct MatTransposed;

class Mat {
public:
    int rows, cols;
    uchar *data;
    MatStep step;

    Mat() : rows(0), cols(0), data(nullptr), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), type_(type) { create(r, c, type); }
    Mat(Size sz, int type) : rows(0), cols(0), data(nullptr), type_(type) { create(sz.height, sz.width, type); }
    Mat(Size sz, int type, const Scalar &s) : rows(0), cols(0), data(nullptr), type_(type) {
        create(sz.height, sz.width, type);
        for (int r = 0; r < rows; r++) for (int c = 0; c < cols; c++) { if (type == CV_32F) at<float>(r, c) = (float)s.v0; else at<uchar>(r, c) = (uchar)s.v0; }
    }
    // External (non-owned) data, like cv::Mat(rows, cols, type, void* data, step).
    Mat(int r, int c, int type, void *ext, size_t stepBytes = 0)
        : rows(r), cols(c), data((uchar *)ext), type_(type) {
        step.p[0] = stepBytes ? stepBytes : (size_t)c * elemSize();
        step.p[1] = elemSize();
    }
    Mat(const Mat &m, const Rect &r) { *this = m; adjustROI(r); }
    Mat(const MatZerosExpr &e) : rows(0), cols(0), data(nullptr), type_(e.type) { *this = e; }

    // OpenCV semantics of `m = Mat::zeros(r,c,t)`: create() is a no-op when size/type already match,
    // then the EXISTING storage is zero-filled.  src/ORBextractor.cc:962-964 relies on this to write
    // descriptors straight into a row-range view of the output matrix.
    Mat &operator=(const MatZerosExpr &e) {
        if (!(data && rows == e.rows && cols == e.cols && type_ == e.type)) create(e.rows, e.cols, e.type);
        for (int r = 0; r < rows; r++) std::memset(data + r * step.p[0], 0, (size_t)cols * elemSize());
        return *this;
    }

    static MatZerosExpr zeros(int r, int c, int type) { return MatZerosExpr{r, c, type}; }

    void create(int r, int c, int type) {
        if (data && rows == r && cols == c && type_ == type) return;
        type_ = type;
        rows = r; cols = c;
        step.p[1] = elemSize();
        step.p[0] = (size_t)c * elemSize();
        buf_ = std::make_shared<std::vector<uchar>>((size_t)r * step.p[0] + 64);
        data = buf_->data();
    }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; }

    size_t elemSize() const { return type_ == CV_32F ? 4 : 1; }
    size_t step1() const { return step.p[0] / elemSize(); }
    int type() const { return type_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step.p[0] == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }

    template <typename T> T &at(int r, int c) { return *(T *)(data + (size_t)r * step.p[0] + (size_t)c * sizeof(T)); }
    template <typename T> const T &at(int r, int c) const {
        return *(const T *)(data + (size_t)r * step.p[0] + (size_t)c * sizeof(T));
    }
    uchar *ptr(int r = 0) { return data + (size_t)r * step.p[0]; }
    const uchar *ptr(int r = 0) const { return data + (size_t)r * step.p[0]; }
    template <typename T> T *ptr(int r = 0) { return (T *)(data + (size_t)r * step.p[0]); }
    template <typename T> const T *ptr(int r = 0) const { return (const T *)(data + (size_t)r * step.p[0]); }

    Mat rowRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * step.p[0]; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.data = data + (size_t)a * elemSize(); m.cols = b - a; return m; }
    Mat row(int r) const { return rowRange(r, r + 1); }
    Mat col(int c) const { return colRange(c, c + 1); }
    Mat operator()(const Rect &r) const { return Mat(*this, r); }
    // single-index access of a vector (n x 1 or 1 x n), as cv::Mat::at<T>(int i0)
    template <typename T> T &at(int i) { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    template <typename T> const T &at(int i) const { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    // float-matrix surface of src/ORBmatcher.cc / src/Frame.cc, defined in minicv_matops.hpp
    struct MatTransposed t() const;   // lazy, like cv::MatExpr: -A.t() * B becomes a gemm with the GEMM_1_T flag
    double dot(const Mat &o) const;
    void convertTo(Mat &dst, int type) const;
    static Mat ones(int r, int c, int type);

    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * elemSize());
        return m;
    }
    void copyTo(Mat &dst) const {
        dst.create(rows, cols, type_);
        for (int r = 0; r < rows; r++) std::memmove(dst.ptr(r), ptr(r), (size_t)cols * elemSize());
    }

private:
    void adjustROI(const Rect &r) {
        data = data + (size_t)r.y * step.p[0] + (size_t)r.x * elemSize();
        rows = r.height; cols = r.width;
    }
    int type_;
    std::shared_ptr<std::vector<uchar>> buf_;
};

// InputArray / OutputArray proxies: thin wrappers over a Mat, enough for the reference's operator().
class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat &m) : m_(const_cast<Mat *>(&m)) {}
    bool empty() const { return !m_ || m_->empty(); }
    Mat getMat() const { return m_ ? *m_ : Mat(); }
protected:
    Mat *m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat &m) { m_ = &m; }
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { if (m_) m_->release(); }
};
typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;
static inline _InputArray noArray() { return _InputArray(); }

// ---------------------------------------------------------------------------------------------
// borderInterpolate for BORDER_REFLECT_101 (gfedcb|abcdefgh|gfedcba), OpenCV core/copy.cpp.
static inline int borderReflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p;
        else p = 2 * len - 2 - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// copyMakeBorder, 8UC1, BORDER_REFLECT_101 (optionally | BORDER_ISOLATED).  dst is expected to be
// pre-sized (src/ORBextractor.cc:1141-1146 passes the whole padded buffer whose centre *is* src),
// so rows are moved with memmove and the centre is left untouched when it aliases.
static inline void copyMakeBorder(const Mat &src, Mat &dst, int top, int bottom, int left, int right, int borderType) {
    (void)borderType;  // only REFLECT_101 is used on the path; ISOLATED is what we always do
    dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
    const int w = src.cols, h = src.rows;
    for (int r = 0; r < h; r++) {
        uchar *d = dst.ptr(r + top);
        const uchar *s = src.ptr(r);
        if (d + left != s) std::memmove(d + left, s, (size_t)w);
        for (int x = 0; x < left; x++) d[x] = d[left + borderReflect101(x - left, w)];
        for (int x = 0; x < right; x++) d[left + w + x] = d[left + borderReflect101(w + x, w)];
    }
    const int W = w + left + right;
    for (int r = 0; r < top; r++)
        std::memcpy(dst.ptr(r), dst.ptr(top + borderReflect101(r - top, h)), (size_t)W);
    for (int r = 0; r < bottom; r++)
        std::memcpy(dst.ptr(top + h + r), dst.ptr(top + borderReflect101(h + r, h)), (size_t)W);
}

// ---------------------------------------------------------------------------------------------
// resize, 8UC1, INTER_LINEAR — OpenCV imgproc/resize.cpp (resizeGeneric_ with
// HResizeLinear<uchar,int,short,2048> and VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>).
// Exact 2x decimation is silently switched to INTER_AREA (ResizeAreaFast 2x2 mean) by cv::resize.
static inline short saturateShort(float v) {
    int i = cvRound(v);
    return (short)(i < SHRT_MIN ? SHRT_MIN : i > SHRT_MAX ? SHRT_MAX : i);
}

struct ResizeLinearTab {
    std::vector<int> xofs, xofs1, yofs0, yofs1;
    std::vector<short> alpha, beta;  // 2 per index
    bool area2x;
};

static inline ResizeLinearTab makeResizeLinearTab(int sw, int sh, int dw, int dh) {
    ResizeLinearTab t;
    const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
    const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    const int iscale_x = (int)std::nearbyint(scale_x), iscale_y = (int)std::nearbyint(scale_y);
    t.area2x = std::abs(scale_x - iscale_x) < DBL_EPSILON && std::abs(scale_y - iscale_y) < DBL_EPSILON &&
               iscale_x == 2 && iscale_y == 2;
    t.xofs.resize(dw); t.xofs1.resize(dw); t.alpha.resize(2 * dw);
    t.yofs0.resize(dh); t.yofs1.resize(dh); t.beta.resize(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cvFloor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        t.xofs[dx] = sx;
        t.xofs1[dx] = std::min(sx + 1, sw - 1);
        t.alpha[2 * dx] = saturateShort((1.f - fx) * 2048.f);
        t.alpha[2 * dx + 1] = saturateShort(fx * 2048.f);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cvFloor(fy);
        fy -= sy;
        // rows are clipped, the weights are not (resizeGeneric_Invoker: clip(sy0 + k, 0, ssize.height))
        t.yofs0[dy] = std::min(std::max(sy, 0), sh - 1);
        t.yofs1[dy] = std::min(std::max(sy + 1, 0), sh - 1);
        t.beta[2 * dy] = saturateShort((1.f - fy) * 2048.f);
        t.beta[2 * dy + 1] = saturateShort(fy * 2048.f);
    }
    return t;
}

static inline void resize(const Mat &src, Mat &dst, Size dsize, double fx = 0, double fy = 0,
                          int interpolation = INTER_LINEAR) {
    (void)fx; (void)fy; (void)interpolation;
    assert(src.type() == CV_8UC1);
    dst.create(dsize.height, dsize.width, CV_8UC1);
    const int sw = src.cols, sh = src.rows, dw = dsize.width, dh = dsize.height;
    ResizeLinearTab t = makeResizeLinearTab(sw, sh, dw, dh);
    if (t.area2x) {
        for (int dy = 0; dy < dh; dy++) {
            const uchar *s0 = src.ptr(2 * dy), *s1 = src.ptr(2 * dy + 1);
            uchar *d = dst.ptr(dy);
            for (int dx = 0; dx < dw; dx++)
                d[dx] = (uchar)((s0[2 * dx] + s0[2 * dx + 1] + s1[2 * dx] + s1[2 * dx + 1] + 2) >> 2);
        }
        return;
    }
    std::vector<int> r0(dw), r1(dw);
    int prev0 = -1, prev1 = -1;
    for (int dy = 0; dy < dh; dy++) {
        const int y0 = t.yofs0[dy], y1 = t.yofs1[dy];
        if (y0 == prev1) { r0.swap(r1); prev0 = prev1; prev1 = -1; }
        if (y0 != prev0) {
            const uchar *s = src.ptr(y0);
            for (int dx = 0; dx < dw; dx++)
                r0[dx] = s[t.xofs[dx]] * t.alpha[2 * dx] + s[t.xofs1[dx]] * t.alpha[2 * dx + 1];
            prev0 = y0;
        }
        if (y1 != prev1) {
            if (y1 == y0) r1 = r0;
            else {
                const uchar *s = src.ptr(y1);
                for (int dx = 0; dx < dw; dx++)
                    r1[dx] = s[t.xofs[dx]] * t.alpha[2 * dx] + s[t.xofs1[dx]] * t.alpha[2 * dx + 1];
            }
            prev1 = y1;
        }
        const int b0 = t.beta[2 * dy], b1 = t.beta[2 * dy + 1];
        uchar *d = dst.ptr(dy);
        for (int dx = 0; dx < dw; dx++) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            d[dx] = (uchar)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur, 8UC1, 7x7, sigma 2 — OpenCV >= 3.4/4.x fixed-point path (imgproc/smooth.simd.hpp,
// ufixedpoint16 = 8.8).  Kernel bits: getGaussianKernel(7,2) rounded to 8 fractional bits with error
// diffusion -> {18,34,48,56,48,34,18} (sum 256).  Horizontal sum is 8.8 (<= 65280, fits 16 bits),
// vertical sum 16.16, one rounding: (v + 32768) >> 16.  Border: REFLECT_101 about the true edge.
static const int kGauss7[7] = {18, 34, 48, 56, 48, 34, 18};

static inline void GaussianBlur(const Mat &srcIn, Mat &dst, Size ksize, double sx, double sy = 0,
                                int borderType = BORDER_DEFAULT) {
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    assert(ksize.width == 7 && ksize.height == 7 && srcIn.type() == CV_8UC1);
    Mat src = (srcIn.data == dst.data) ? srcIn.clone() : srcIn;
    dst.create(src.rows, src.cols, CV_8UC1);
    const int w = src.cols, h = src.rows;
    std::vector<uint16_t> hbuf((size_t)w * h);
    std::vector<int> xi((size_t)w + 6);
    for (int x = -3; x < w + 3; x++) xi[x + 3] = borderReflect101(x, w);
    for (int y = 0; y < h; y++) {
        const uchar *s = src.ptr(y);
        uint16_t *hb = &hbuf[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            int acc = 0;
            for (int k = 0; k < 7; k++) acc += kGauss7[k] * s[xi[x + k]];
            hb[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; y++) {
        const uint16_t *r[7];
        for (int k = 0; k < 7; k++) r[k] = &hbuf[(size_t)borderReflect101(y + k - 3, h) * w];
        uchar *d = dst.ptr(y);
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int k = 0; k < 7; k++) acc += (uint32_t)kGauss7[k] * r[k][x];
            d[x] = (uchar)((acc + 32768u) >> 16);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// fastAtan2 (degrees, [0,360)) — OpenCV core/mathfuncs_core.simd.hpp atan_f32: 7th-order odd
// polynomial, all-float, no contraction.
static inline float fastAtan2(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / CV_PI);
    static const float p3 = -0.3258083974640975f * (float)(180 / CV_PI);
    static const float p5 = 0.1555786518463281f * (float)(180 / CV_PI);
    static const float p7 = -0.04432655554792128f * (float)(180 / CV_PI);
    float ax = std::abs(x), ay = std::abs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---------------------------------------------------------------------------------------------
// FAST-9/16 with optional 3x3 non-maximum suppression — OpenCV features2d/fast.cpp FAST_t<16> and
// fast_score.cpp cornerScore<16>.  Detection area: rows/cols [3, dim-3).  Score of a corner is the
// largest threshold at which it is still a corner (= m-1, m = max over the 16 arcs of 9 contiguous
// circle pixels of min(v-p) resp. min(p-v)); a corner survives NMS iff its score is strictly greater
// than the scores of all 8 neighbours, where non-corners count as 0.  Output in raster order,
// KeyPoint(x, y, size 7, angle -1, response score).
static inline int fastCornerScore16(const uchar *ptr, const int pixel[25], int threshold) {
    const int K = 8, N = 25;
    int d[N];
    const int v = ptr[0];
    for (int k = 0; k < N; k++) d[k] = v - ptr[pixel[k]];
    int a0 = threshold;
    for (int k = 0; k < 16; k += 2) {
        int a = std::min(d[k + 1], d[k + 2]);
        a = std::min(a, d[k + 3]);
        if (a <= a0) continue;
        a = std::min(a, d[k + 4]); a = std::min(a, d[k + 5]); a = std::min(a, d[k + 6]);
        a = std::min(a, d[k + 7]); a = std::min(a, d[k + 8]);
        a0 = std::max(a0, std::min(a, d[k]));
        a0 = std::max(a0, std::min(a, d[k + 9]));
    }
    int b0 = -a0;
    for (int k = 0; k < 16; k += 2) {
        int b = std::max(d[k + 1], d[k + 2]);
        b = std::max(b, d[k + 3]); b = std::max(b, d[k + 4]); b = std::max(b, d[k + 5]);
        if (b >= b0) continue;
        b = std::max(b, d[k + 6]); b = std::max(b, d[k + 7]); b = std::max(b, d[k + 8]);
        b0 = std::min(b0, std::max(b, d[k]));
        b0 = std::min(b0, std::max(b, d[k + 9]));
    }
    (void)K;
    return -b0 - 1;
}

static inline void FAST(const Mat &img, std::vector<KeyPoint> &keypoints, int threshold, bool nonmaxSuppression = true) {
    keypoints.clear();
    const int K = 8, N = 25;
    static const int offs[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                    {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};
    const int step = (int)img.step;
    int pixel[25];
    for (int k = 0; k < 16; k++) pixel[k] = offs[k][0] + offs[k][1] * step;
    for (int k = 16; k < 25; k++) pixel[k] = pixel[k - 16];
    threshold = std::min(std::max(threshold, 0), 255);
    uchar tab[512];
    for (int i = -255; i <= 255; i++) tab[i + 255] = (uchar)(i < -threshold ? 1 : i > threshold ? 2 : 0);

    const int cols = img.cols, rows = img.rows;
    if (cols < 7 || rows < 7) return;
    std::vector<uchar> bufv((size_t)3 * cols, 0);
    std::vector<int> cpv((size_t)3 * (cols + 1), 0);
    uchar *buf[3] = {&bufv[0], &bufv[cols], &bufv[2 * cols]};
    int *cpbuf[3] = {&cpv[1], &cpv[cols + 2], &cpv[2 * cols + 3]};

    for (int i = 3; i < rows - 2; i++) {
        const uchar *ptr = img.ptr(i) + 3;
        uchar *curr = buf[(i - 3) % 3];
        int *cornerpos = cpbuf[(i - 3) % 3];
        std::memset(curr, 0, (size_t)cols);
        int ncorners = 0;
        if (i < rows - 3) {
            for (int j = 3; j < cols - 3; j++, ptr++) {
                const int v = ptr[0];
                const uchar *tb = &tab[0] - v + 255;
                int d = tb[ptr[pixel[0]]] | tb[ptr[pixel[8]]];
                if (d == 0) continue;
                d &= tb[ptr[pixel[2]]] | tb[ptr[pixel[10]]];
                d &= tb[ptr[pixel[4]]] | tb[ptr[pixel[12]]];
                d &= tb[ptr[pixel[6]]] | tb[ptr[pixel[14]]];
                if (d == 0) continue;
                d &= tb[ptr[pixel[1]]] | tb[ptr[pixel[9]]];
                d &= tb[ptr[pixel[3]]] | tb[ptr[pixel[11]]];
                d &= tb[ptr[pixel[5]]] | tb[ptr[pixel[13]]];
                d &= tb[ptr[pixel[7]]] | tb[ptr[pixel[15]]];
                bool found = false;
                if (d & 1) {
                    const int vt = v - threshold;
                    int count = 0;
                    for (int k = 0; k < N; k++) {
                        if (ptr[pixel[k]] < vt) { if (++count > K) { found = true; break; } }
                        else count = 0;
                    }
                }
                if (!found && (d & 2)) {
                    const int vt = v + threshold;
                    int count = 0;
                    for (int k = 0; k < N; k++) {
                        if (ptr[pixel[k]] > vt) { if (++count > K) { found = true; break; } }
                        else count = 0;
                    }
                }
                if (found) {
                    cornerpos[ncorners++] = j;
                    if (nonmaxSuppression) curr[j] = (uchar)fastCornerScore16(ptr, pixel, threshold);
                }
            }
        }
        cornerpos[-1] = ncorners;
        if (i == 3) continue;
        const uchar *prev = buf[(i - 4 + 3) % 3];
        const uchar *pprev = buf[(i - 5 + 3) % 3];
        cornerpos = cpbuf[(i - 4 + 3) % 3];
        ncorners = cornerpos[-1];
        for (int k = 0; k < ncorners; k++) {
            const int j = cornerpos[k];
            const int score = prev[j];
            if (!nonmaxSuppression ||
                (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] && score > pprev[j] &&
                 score > pprev[j + 1] && score > curr[j - 1] && score > curr[j] && score > curr[j + 1])) {
                keypoints.push_back(KeyPoint((float)j, (float)(i - 1), 7.f, -1, (float)score));
            }
        }
    }
}

// KeyPointsFilter::retainBest — only reached from ComputeKeyPointsOld (src/ORBextractor.cc:806-960),
// which no caller uses; provided so the reference file links.  Keeps the n strongest responses
// (plus ties with the n-th), like OpenCV's nth_element + partition.
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint> &kps, int n) {
        if (n >= 0 && (int)kps.size() > n) {
            if (n == 0) { kps.clear(); return; }
            std::nth_element(kps.begin(), kps.begin() + n - 1, kps.end(),
                             [](const KeyPoint &a, const KeyPoint &b) { return a.response > b.response; });
            const float amb = kps[n - 1].response;
            auto newEnd = std::partition(kps.begin() + n, kps.end(),
                                         [amb](const KeyPoint &k) { return k.response >= amb; });
            kps.resize(newEnd - kps.begin());
        }
    }
};

}  // namespace cv

using cv::cvRound;
using cv::cvFloor;
using cv::cvCeil;

#endif  // YGZ_ORACLE_MINICV_HPP_
