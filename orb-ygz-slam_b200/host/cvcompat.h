// cvcompat.h — the handful of OpenCV types the drop-in classes use.  With OpenCV installed (the
// reference's build: find_package(OpenCV), CMakeLists.txt:40-46) the real headers are used; this image
// has no OpenCV C++, so the host tests compile against the minimal stand-ins below (same member names
// and memory layout: cv::KeyPoint is the 28-byte record of include/ygzorb.h).
#ifndef YGZ_HOST_CVCOMPAT_H_
#define YGZ_HOST_CVCOMPAT_H_

#if defined(YGZ_USE_OPENCV) || (defined(__has_include) && __has_include(<opencv2/core/core.hpp>))
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#else
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

namespace cv {
typedef unsigned char uchar;
enum { CV_8U = 0, CV_8UC1 = 0 };
struct Point2f { float x, y; Point2f() : x(0), y(0) {} Point2f(float a, float b) : x(a), y(b) {} };
struct KeyPoint {
    Point2f pt; float size, angle, response; int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};
class Mat {
public:
    int rows, cols; size_t step; uchar *data;
    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int /*type*/) { create(r, c, CV_8UC1); }
    Mat(int r, int c, int /*type*/, void *p, size_t s = 0) : rows(r), cols(c), step(s ? s : (size_t)c), data((uchar *)p) {}
    void create(int r, int c, int /*type*/) {
        rows = r; cols = c; step = (size_t)c;
        buf_.reset(new std::vector<uchar>((size_t)r * c)); data = buf_->data();
    }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    bool isContinuous() const { return step == (size_t)cols; }
    int type() const { return CV_8UC1; }
    uchar *ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar *ptr(int r = 0) const { return data + (size_t)r * step; }
    template <class T> T *ptr(int r = 0) { return (T *)(data + (size_t)r * step); }
    template <class T> const T *ptr(int r = 0) const { return (const T *)(data + (size_t)r * step); }
    Mat row(int r) const { Mat m(*this); m.data = data + (size_t)r * step; m.rows = 1; return m; }
    Mat clone() const { Mat m(rows, cols, CV_8UC1); for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols); return m; }
private:
    std::shared_ptr<std::vector<uchar>> buf_;
};
class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat &m) : m_(const_cast<Mat *>(&m)) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
protected:
    Mat *m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat &m) { m_ = &m; }
    void create(int r, int c, int t) const { m_->create(r, c, t); }
    void release() const { if (m_) m_->release(); }
    Mat &getMatRef() const { return *m_; }
};
typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;
}  // namespace cv
#endif
static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint must be the 28-byte record of ygzorb_keypoint");
#endif  // YGZ_HOST_CVCOMPAT_H_
