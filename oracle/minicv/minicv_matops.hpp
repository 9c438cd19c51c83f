// minicv_matops.hpp — TEST INFRASTRUCTURE ONLY.  The float-matrix surface of cv::Mat that the reference's src/ORBmatcher.cc
// (Sim3 overloads :265-373, :888-1004, :1006-1216) and src/Frame.cc (ComputeStereoMatches :604-627) use, restated from
// OpenCV's core/matmul, matop and norm code paths for the shapes that occur there (3x3, 3x1, 11x11, CV_32F) and pinned to
// cv2 4.13 by tests/test_minicv_vs_cv2.py:
//   * A * B (+ C), no transposition flag, inner dimension 3: OpenCV's unrolled small-matrix path of gemm — the three
//     products are summed left to right in float, then d = (float)(t * alpha + c * beta) with alpha, beta double;
//     with a transposed operand (-Rcw.t() * tcw): the general GEMMSingleMul<float, double> path, accumulation in double;
//   * s * A, A / s: MatExpr scale, evaluated by convertTo: each element times (float)alpha (alpha = s resp. 1.0 / s, double);
//   * A + B, A - B element-wise in float; -A exact;
//   * Mat::dot and cv::norm (NORM_L2, NORM_L1 of a difference) accumulate in double in index order.
#ifndef YGZ_ORACLE_MINICV_MATOPS_HPP_
#define YGZ_ORACLE_MINICV_MATOPS_HPP_
#include "minicv.hpp"

namespace cv {

enum { NORM_L1 = 2, NORM_L2 = 4 };

struct MatTransposed {
    Mat a;
    operator Mat() const {
        Mat r(a.cols, a.rows, CV_32F);
        for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) r.at<float>(j, i) = a.at<float>(i, j);
        return r;
    }
};
inline MatTransposed Mat::t() const { return MatTransposed{*this}; }
inline double Mat::dot(const Mat &o) const {
    double s = 0;
    for (int i = 0; i < rows; i++) for (int j = 0; j < cols; j++) s += (double)at<float>(i, j) * (double)o.at<float>(i, j);
    return s;
}
inline void Mat::convertTo(Mat &dst, int type) const {   // 8U -> 32F (exact) or same type
    Mat r(rows, cols, type);
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) {
            const float v = type_ == CV_32F ? at<float>(i, j) : (float)at<uchar>(i, j);
            if (type == CV_32F) r.at<float>(i, j) = v; else r.at<uchar>(i, j) = (uchar)v;
        }
    dst = r;
}
inline Mat Mat::ones(int r, int c, int type) {
    Mat m(r, c, type);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) { if (type == CV_32F) m.at<float>(i, j) = 1.0f; else m.at<uchar>(i, j) = 1; }
    return m;
}

// gemm: D = alpha * op(A) * B + beta * C.  aT: A is used transposed (flag GEMM_1_T).
inline Mat gemmF(const Mat &A, bool aT, const Mat &B, double alpha, const Mat *C, double beta) {
    const int M = aT ? A.cols : A.rows, K = aT ? A.rows : A.cols, N = B.cols;
    assert(B.rows == K);
    Mat D(M, N, CV_32F);
    const bool small = !aT && K >= 2 && K <= 4 && (K == N || K == M);   // matmul.simd.hpp: flags == 0 && 2 <= len <= 4
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) {
            const float c = C ? C->at<float>(i, C->cols == 1 ? 0 : j) : 0.0f;
            if (small) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < K; k++) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                D.at<float>(i, j) = (float)((double)t * alpha + (double)c * beta);
            } else {
                double s = 0;
                for (int k = 0; k < K; k++) s += (double)(aT ? A.at<float>(k, i) : A.at<float>(i, k)) * (double)B.at<float>(k, j);
                D.at<float>(i, j) = (float)(s * alpha + (double)c * beta);
            }
        }
    return D;
}

// Lazy expression just deep enough for the forms used: [-]A[.t()] * B [+ C],  s * A,  s * A.t(),  A / s.
struct MatScaled {     // alpha * op(A), not yet evaluated
    Mat a; bool t; double alpha;
    operator Mat() const {
        Mat src = t ? (Mat)a.t() : a;
        Mat r(src.rows, src.cols, CV_32F);
        const float f = (float)alpha;
        for (int i = 0; i < src.rows; i++) for (int j = 0; j < src.cols; j++) r.at<float>(i, j) = src.at<float>(i, j) * f;
        return r;
    }
};
struct MatProduct {    // alpha * op(A) * B, not yet added to C
    Mat a, b; bool t; double alpha;
    operator Mat() const { return gemmF(a, t, b, alpha, nullptr, 0.0); }
};

inline MatScaled operator-(const Mat &a) { return MatScaled{a, false, -1.0}; }
inline MatScaled operator-(const MatTransposed &a) { return MatScaled{a.a, true, -1.0}; }
inline MatScaled operator*(double s, const MatTransposed &a) { return MatScaled{a.a, true, s}; }
inline MatScaled operator*(double s, const Mat &a) { return MatScaled{a, false, s}; }
inline MatScaled operator*(float s, const Mat &a) { return MatScaled{a, false, (double)s}; }
inline MatScaled operator/(const Mat &a, double s) { return MatScaled{a, false, 1.0 / s}; }
inline MatScaled operator/(const Mat &a, float s) { return MatScaled{a, false, 1.0 / (double)s}; }
inline MatProduct operator*(const Mat &a, const Mat &b) { return MatProduct{a, b, false, 1.0}; }
inline MatProduct operator*(const MatScaled &a, const Mat &b) { return MatProduct{a.a, b, a.t, a.alpha}; }
inline Mat operator+(const MatProduct &p, const Mat &c) { return gemmF(p.a, p.t, p.b, p.alpha, &c, 1.0); }
inline Mat operator+(const Mat &a, const Mat &b) {
    Mat r(a.rows, a.cols, CV_32F);
    for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) r.at<float>(i, j) = a.at<float>(i, j) + b.at<float>(i, j);
    return r;
}
inline Mat operator-(const Mat &a, const Mat &b) {
    Mat r(a.rows, a.cols, CV_32F);
    for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) r.at<float>(i, j) = a.at<float>(i, j) - b.at<float>(i, j);
    return r;
}
inline Mat operator-(const Mat &a, const MatScaled &b) { return a - (Mat)b; }

inline double norm(const Mat &a) {   // NORM_L2, CV_32F
    double s = 0;
    for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) { const double v = a.at<float>(i, j); s += v * v; }
    return std::sqrt(s);
}
inline double norm(const Mat &a, const Mat &b, int type) {   // NORM_L1 of the difference, CV_32F
    (void)type;
    double s = 0;
    for (int i = 0; i < a.rows; i++) for (int j = 0; j < a.cols; j++) s += std::fabs((double)a.at<float>(i, j) - (double)b.at<float>(i, j));
    return s;
}

}  // namespace cv
#endif
