// eigen_lite.h — TEST INFRASTRUCTURE ONLY.  A fixed-size stand-in for the slice of Eigen 3.3 and Sophus that the reference's
// src/ORBmatcher.cc, src/Align.cc, src/SparseImageAlign.cc and the Frame / KeyFrame / MapPoint members on the hot path use
// (Eigen and sophus' dependency on it are absent from this image; include/Common.h:20-45 pulls them in).  It exists so that
// those files compile UNMODIFIED into oracle/_ref/ (oracle/Makefile, target refmatch): the control flow, the choice of float
// vs double in every scalar expression and the order of the scalar statements are then the reference's own text.  What this
// header decides is the order of float operations INSIDE a matrix / vector expression.  It follows Eigen 3.3 for these sizes
// (no vectorisation for 2- and 3-vectors of float, DefaultTraversal + CompleteUnrolling):
//   * a sum of three terms (dot, norm, row of a 3x3 product) = a0 + (a1 + a2)   (redux_novec_unroller splits [0,1) | [1,3));
//     a sum of two = a0 + a1; a sum of six (Matrix<float,6,1>) = (a0 + (a1 + a2)) + (a3 + (a4 + a5));
//   * Matrix2f / Matrix3f inverse and determinant by the cofactor formulas of Eigen/src/LU/InverseImpl.h / Determinant.h;
//   * Sophus::SE3f stores (unit quaternion, translation); SE3 * point = q._transformVector(p) + t with Eigen's
//     "uv = q.vec x v; uv += uv; v + w * uv + q.vec x uv"; rotationMatrix() = Quaternion::toRotationMatrix();
//     inverse() = (q.conjugate(), q.conjugate() * (t * -1)); product = translation += q * other.t, quaternion product, then
//     Sophus' normalize() (Thirdparty/sophus/sophus/so3.hpp:268-276: coeffs() /= norm()).  Pose PRODUCTS are not on the
//     device path (they stay in the caller's Sophus code); the wrappers hand the product they computed to
//     the restatement as an input, so nothing downstream depends on how the four squares of norm() are summed.
//   * round 2 (SparseImageAlign / NLSSolver): Matrix<float,6,1>, <6,6>, <2,6>, <6,Dynamic>; (J * J^T) * w and (J * r) * w element
//     by element in that order; LDLT as written in MatF::LDLT::solve; SE3f::exp from the vendored sophus text.
// All of it uncontracted (-ffp-contract=off).  The same conventions are written out in oracle/match_oracle.cpp and
// oracle/direct_oracle.cpp (restatement) and in the kernels; DESIGN.md §2 names them as the canonical float order.
#ifndef YGZ_REF_SHIM_EIGEN_LITE_H_
#define YGZ_REF_SHIM_EIGEN_LITE_H_
#include <cmath>
#include <cstring>
#include <ostream>
#include <vector>

namespace Eigen {

template <int N> struct VecF;
template <int N> struct MatF;

// comma initialiser: v << a, b;
template <int N> struct CommaInit {
    float *p; int i;
    CommaInit &operator,(float v) { if (i < N) p[i++] = v; return *this; }
};

inline float sum2(float a0, float a1) { return a0 + a1; }
inline float sum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }
template <int N> inline float sumN(const float *a);
template <> inline float sumN<2>(const float *a) { return sum2(a[0], a[1]); }
template <> inline float sumN<3>(const float *a) { return sum3(a[0], a[1], a[2]); }
template <> inline float sumN<6>(const float *a) { return sum3(a[0], a[1], a[2]) + sum3(a[3], a[4], a[5]); }

template <int N> struct RowVecF {   // v.transpose(): only as the right operand of an outer product
    float d[N];
};

template <int N> struct VecF {
    float d[N];
    VecF() {}
    VecF(float x, float y) { static_assert(N == 2, "2-vector"); d[0] = x; d[1] = y; }
    VecF(float x, float y, float z) { static_assert(N == 3, "3-vector"); d[0] = x; d[1] = y; d[2] = z; }
    float &operator[](int i) { return d[i]; }
    const float &operator[](int i) const { return d[i]; }
    float &operator()(int i) { return d[i]; }
    const float &operator()(int i) const { return d[i]; }
    float &operator()(int i, int) { return d[i]; }               // column vector: (i, 0)
    const float &operator()(int i, int) const { return d[i]; }
    float &x() { return d[0]; } float &y() { return d[1]; } float &z() { return d[2]; }
    const float &x() const { return d[0]; } const float &y() const { return d[1]; } const float &z() const { return d[2]; }
    void setZero() { for (int i = 0; i < N; i++) d[i] = 0.0f; }
    static VecF Zero() { VecF r; r.setZero(); return r; }
    float dot(const VecF &o) const { float p[N]; for (int i = 0; i < N; i++) p[i] = d[i] * o.d[i]; return sumN<N>(p); }
    float squaredNorm() const { return dot(*this); }
    float norm() const { return std::sqrt(squaredNorm()); }
    RowVecF<N> transpose() const { RowVecF<N> r; for (int i = 0; i < N; i++) r.d[i] = d[i]; return r; }
    CommaInit<N> operator<<(float v) { d[0] = v; return CommaInit<N>{d, 1}; }
    VecF &operator+=(const VecF &o) { for (int i = 0; i < N; i++) d[i] = d[i] + o.d[i]; return *this; }
    VecF &operator-=(const VecF &o) { for (int i = 0; i < N; i++) d[i] = d[i] - o.d[i]; return *this; }
    VecF &operator*=(float s) { for (int i = 0; i < N; i++) d[i] = d[i] * s; return *this; }
    VecF operator-() const { VecF r; for (int i = 0; i < N; i++) r.d[i] = -d[i]; return r; }
    template <int K> VecF<K> head() const { VecF<K> r; for (int i = 0; i < K; i++) r.d[i] = d[i]; return r; }
    VecF &noalias() { return *this; }                            // v.noalias() -= expr: the expression is evaluated first either way
    struct Diag { const float *d; };                             // v.asDiagonal(), only as the right operand of M += ...
    Diag asDiagonal() const { return Diag{d}; }
};
template <int N> inline VecF<N> operator+(const VecF<N> &a, const VecF<N> &b) { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = a.d[i] + b.d[i]; return r; }
template <int N> inline VecF<N> operator-(const VecF<N> &a, const VecF<N> &b) { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = a.d[i] - b.d[i]; return r; }
template <int N> inline VecF<N> operator*(const VecF<N> &a, float s) { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = a.d[i] * s; return r; }
template <int N> inline VecF<N> operator*(float s, const VecF<N> &a) { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = s * a.d[i]; return r; }
template <int N> inline VecF<N> operator/(const VecF<N> &a, float s) { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = a.d[i] / s; return r; }

template <int N> struct ColRef {   // M.col(j) = v
    float *p;   // first element of the column, stride N (row-major storage below)
    ColRef &operator=(const VecF<N> &v) { for (int i = 0; i < N; i++) p[i * N] = v.d[i]; return *this; }
    operator VecF<N>() const { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = p[i * N]; return r; }
};

template <int N> struct MatF {
    float m[N][N];   // m[row][col]; the layout is private to this header
    MatF() {}
    float &operator()(int r, int c) { return m[r][c]; }
    const float &operator()(int r, int c) const { return m[r][c]; }
    void setZero() { for (int r = 0; r < N; r++) for (int c = 0; c < N; c++) m[r][c] = 0.0f; }
    static MatF Zero() { MatF r; r.setZero(); return r; }
    static MatF Identity() { MatF r; r.setZero(); for (int i = 0; i < N; i++) r.m[i][i] = 1.0f; return r; }
    MatF transpose() const { MatF r; for (int a = 0; a < N; a++) for (int b = 0; b < N; b++) r.m[a][b] = m[b][a]; return r; }
    ColRef<N> col(int j) { return ColRef<N>{&m[0][j]}; }
    VecF<N> col(int j) const { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = m[i][j]; return r; }
    MatF &operator+=(const MatF &o) { for (int r = 0; r < N; r++) for (int c = 0; c < N; c++) m[r][c] = m[r][c] + o.m[r][c]; return *this; }
    float determinant() const;
    MatF inverse() const;
    // --- what include/NLSSolver_impl.hpp and src/SparseImageAlign.cc add (N = 6)
    MatF &noalias() { return *this; }
    VecF<N> diagonal() const { VecF<N> r; for (int i = 0; i < N; i++) r.d[i] = m[i][i]; return r; }
    MatF &operator+=(const typename VecF<N>::Diag &dg) { for (int i = 0; i < N; i++) m[i][i] = m[i][i] + dg.d[i]; return *this; }
    struct LDLT { const MatF *A; VecF<N> solve(const VecF<N> &b) const; };
    LDLT ldlt() const { return LDLT{this}; }
};
template <int N> inline MatF<N> operator/(const MatF<N> &A, float s) { MatF<N> r; for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) r.m[i][j] = A.m[i][j] / s; return r; }
template <int N> inline std::ostream &operator<<(std::ostream &o, const MatF<N> &A) {
    for (int i = 0; i < N; i++) { for (int j = 0; j < N; j++) o << A.m[i][j] << (j + 1 < N ? " " : ""); o << "\n"; }
    return o;
}
template <int N> inline std::ostream &operator<<(std::ostream &o, const VecF<N> &v) { for (int i = 0; i < N; i++) o << v.d[i] << "\n"; return o; }
// x = A.ldlt().solve(b): LDL^T with the largest remaining |diagonal| as pivot (Eigen/src/Cholesky/LDLT.h: unblocked in-place
// factorisation with transpositions, then P, L, D (entries below the smallest normal float count as zero), L^T, P^T).  The
// order of the float operations inside the column updates is a choice of this stand-in (sequential, as written below); it is
// the one oracle/direct_oracle.cpp (ldltSolve6) and the kernel use.
template <int N> inline VecF<N> MatF<N>::LDLT::solve(const VecF<N> &b) const {
    float Aw[N][N]; int perm[N];
    for (int r = 0; r < N; r++) { perm[r] = r; for (int c = 0; c < N; c++) Aw[r][c] = A->m[r][c]; }
    for (int k = 0; k < N; k++) {
        int p = k;
        for (int i = k + 1; i < N; i++) if (std::fabs(Aw[i][i]) > std::fabs(Aw[p][p])) p = i;
        if (p != k) {
            for (int c = 0; c < N; c++) { const float t = Aw[k][c]; Aw[k][c] = Aw[p][c]; Aw[p][c] = t; }
            for (int r = 0; r < N; r++) { const float t = Aw[r][k]; Aw[r][k] = Aw[r][p]; Aw[r][p] = t; }
            const int t = perm[k]; perm[k] = perm[p]; perm[p] = t;
        }
        float d = Aw[k][k];
        for (int j = 0; j < k; j++) d -= Aw[k][j] * Aw[k][j] * Aw[j][j];
        Aw[k][k] = d;
        for (int i = k + 1; i < N; i++) {
            float v = Aw[i][k];
            for (int j = 0; j < k; j++) v -= Aw[i][j] * Aw[k][j] * Aw[j][j];
            Aw[i][k] = d != 0 ? v / d : v;
        }
    }
    float y[N];
    for (int i = 0; i < N; i++) { float v = b.d[perm[i]]; for (int j = 0; j < i; j++) v -= Aw[i][j] * y[j]; y[i] = v; }
    for (int i = 0; i < N; i++) y[i] = std::fabs(Aw[i][i]) > 1.17549435e-38f ? y[i] / Aw[i][i] : 0.0f;
    for (int i = N - 1; i >= 0; i--) { float v = y[i]; for (int j = i + 1; j < N; j++) v -= Aw[j][i] * y[j]; y[i] = v; }
    VecF<N> x;
    for (int i = 0; i < N; i++) x.d[perm[i]] = y[i];
    return x;
}
// outer product J * J.transpose()
template <int N> inline MatF<N> operator*(const VecF<N> &a, const RowVecF<N> &b) {
    MatF<N> r; for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) r.m[i][j] = a.d[i] * b.d[j]; return r;
}
template <int N> inline VecF<N> operator*(const MatF<N> &A, const VecF<N> &v) {
    VecF<N> r;
    for (int i = 0; i < N; i++) { float p[N]; for (int k = 0; k < N; k++) p[k] = A.m[i][k] * v.d[k]; r.d[i] = sumN<N>(p); }
    return r;
}
template <int N> inline MatF<N> operator*(const MatF<N> &A, const MatF<N> &B) {
    MatF<N> r;
    for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) { float p[N]; for (int k = 0; k < N; k++) p[k] = A.m[i][k] * B.m[k][j]; r.m[i][j] = sumN<N>(p); }
    return r;
}
template <int N> inline MatF<N> operator*(float s, const MatF<N> &A) { MatF<N> r; for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) r.m[i][j] = s * A.m[i][j]; return r; }
template <int N> inline MatF<N> operator*(int s, const MatF<N> &A) { return (float)s * A; }
template <int N> inline MatF<N> operator*(const MatF<N> &A, float s) { MatF<N> r; for (int i = 0; i < N; i++) for (int j = 0; j < N; j++) r.m[i][j] = A.m[i][j] * s; return r; }

// Eigen/src/LU/Determinant.h, InverseImpl.h
template <> inline float MatF<2>::determinant() const { return m[0][0] * m[1][1] - m[1][0] * m[0][1]; }
template <> inline MatF<2> MatF<2>::inverse() const {
    const float invdet = 1.0f / determinant();
    MatF<2> r;
    r.m[0][0] = m[1][1] * invdet; r.m[1][0] = -m[1][0] * invdet; r.m[0][1] = -m[0][1] * invdet; r.m[1][1] = m[0][0] * invdet;
    return r;
}
inline float cofactor3(const MatF<3> &M, int i, int j) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return M.m[i1][j1] * M.m[i2][j2] - M.m[i1][j2] * M.m[i2][j1];
}
template <> inline float MatF<3>::determinant() const {   // bruteforce_det3_helper: a0 + (a1 + a2) of m(0,k) * minor
    auto h = [&](int a, int b, int c) { return m[0][a] * (m[1][b] * m[2][c] - m[1][c] * m[2][b]); };
    return h(0, 1, 2) - h(1, 0, 2) + h(2, 0, 1);
}
template <> inline MatF<3> MatF<3>::inverse() const {
    // compute_inverse<Matrix3f>: cofactors of column 0, det = (cofactors_col0 .* col(0)).sum(), result = cofactor^T * invdet
    const float c0 = cofactor3(*this, 0, 0), c1 = cofactor3(*this, 1, 0), c2 = cofactor3(*this, 2, 0);
    const float det = sum3(c0 * m[0][0], c1 * m[1][0], c2 * m[2][0]);
    const float invdet = 1.0f / det;
    MatF<3> r;
    r.m[0][0] = c0 * invdet; r.m[0][1] = c1 * invdet; r.m[0][2] = c2 * invdet;
    r.m[1][0] = cofactor3(*this, 0, 1) * invdet; r.m[1][1] = cofactor3(*this, 1, 1) * invdet; r.m[1][2] = cofactor3(*this, 2, 1) * invdet;
    r.m[2][0] = cofactor3(*this, 0, 2) * invdet; r.m[2][1] = cofactor3(*this, 1, 2) * invdet; r.m[2][2] = cofactor3(*this, 2, 2) * invdet;
    return r;
}

typedef VecF<2> Vector2f;
typedef VecF<3> Vector3f;
typedef MatF<2> Matrix2f;
typedef MatF<3> Matrix3f;

// ---- Matrix<float, R, C> spellings used by include/NLSSolver.h and include/SparseImageAlign.h
enum { Dynamic = -1, ColMajor = 0, RowMajor = 1 };
enum NoChange_t { NoChange };
struct Mat26 {                      // Matrix<float, 2, 6>: JacobXYZ2Cam
    float m[2][6];
    float &operator()(int r, int c) { return m[r][c]; }
    const float &operator()(int r, int c) const { return m[r][c]; }
    VecF<6> row(int r) const { VecF<6> v; for (int i = 0; i < 6; i++) v.d[i] = m[r][i]; return v; }   // only inside (a * row + b * row) * s
};
struct Col6Ref {                    // jacobian_cache_.col(j): six contiguous floats (column-major 6 x n)
    float *p;
    Col6Ref &operator=(const VecF<6> &v) { for (int i = 0; i < 6; i++) p[i] = v.d[i]; return *this; }
    operator VecF<6>() const { VecF<6> r; for (int i = 0; i < 6; i++) r.d[i] = p[i]; return r; }
};
struct Mat6X {                      // Matrix<float, 6, Dynamic, ColMajor>
    std::vector<float> d;
    void resize(NoChange_t, long cols) { d.resize((size_t)6 * cols); }
    void setZero() { for (float &v : d) v = 0.0f; }
    long cols() const { return (long)(d.size() / 6); }
    Col6Ref col(long j) { return Col6Ref{&d[(size_t)6 * j]}; }
};
struct VectorXf {                   // only as the parameter type of norm_max (include/NLSSolver.h)
    std::vector<float> d;
    template <int N> VectorXf(const VecF<N> &v) : d(v.d, v.d + N) {}
    int size() const { return (int)d.size(); }
    float operator[](int i) const { return d[i]; }
};
template <typename T, int R, int C> struct MatrixSel;
template <int N> struct MatrixSel<float, N, 1> { typedef VecF<N> type; };
template <int N> struct MatrixSel<float, N, N> { typedef MatF<N> type; };
template <> struct MatrixSel<float, 2, 6> { typedef Mat26 type; };
template <> struct MatrixSel<float, 6, Dynamic> { typedef Mat6X type; };
template <typename T, int R, int C, int Options = 0> using Matrix = typename MatrixSel<T, R, C>::type;
#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

inline Vector3f cross(const Vector3f &a, const Vector3f &b) {
    return Vector3f(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}

struct Quaternionf {
    float qx, qy, qz, qw;   // coeffs() order x, y, z, w
    Quaternionf() {}
    Quaternionf(float w, float x, float y, float z) : qx(x), qy(y), qz(z), qw(w) {}
    Vector3f vec() const { return Vector3f(qx, qy, qz); }
    float w() const { return qw; }
    Quaternionf conjugate() const { return Quaternionf(qw, -qx, -qy, -qz); }
    float squaredNorm() const { return (qx * qx + qy * qy) + (qz * qz + qw * qw); }   // 4 floats: one SSE packet, hadd pairs
    // QuaternionBase::_transformVector
    Vector3f _transformVector(const Vector3f &v) const {
        Vector3f uv = cross(vec(), v);
        uv += uv;
        const Vector3f c = cross(vec(), uv);
        return Vector3f((v[0] + qw * uv[0]) + c[0], (v[1] + qw * uv[1]) + c[1], (v[2] + qw * uv[2]) + c[2]);
    }
    // QuaternionBase::toRotationMatrix
    Matrix3f toRotationMatrix() const {
        Matrix3f R;
        const float tx = 2.0f * qx, ty = 2.0f * qy, tz = 2.0f * qz;
        const float twx = tx * qw, twy = ty * qw, twz = tz * qw;
        const float txx = tx * qx, txy = ty * qx, txz = tz * qx;
        const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
        R(0, 0) = 1.0f - (tyy + tzz); R(0, 1) = txy - twz; R(0, 2) = txz + twy;
        R(1, 0) = txy + twz; R(1, 1) = 1.0f - (txx + tzz); R(1, 2) = tyz - twx;
        R(2, 0) = txz - twy; R(2, 1) = tyz + twx; R(2, 2) = 1.0f - (txx + tyy);
        return R;
    }
};
// quat_product<Architecture::Generic> (the SSE float specialisation computes the same products with the same signs)
inline Quaternionf operator*(const Quaternionf &a, const Quaternionf &b) {
    return Quaternionf(a.qw * b.qw - a.qx * b.qx - a.qy * b.qy - a.qz * b.qz,
                       a.qw * b.qx + a.qx * b.qw + a.qy * b.qz - a.qz * b.qy,
                       a.qw * b.qy + a.qy * b.qw + a.qz * b.qx - a.qx * b.qz,
                       a.qw * b.qz + a.qz * b.qw + a.qx * b.qy - a.qy * b.qx);
}

}  // namespace Eigen

namespace Sophus {
typedef Eigen::VecF<6> Vector6f;   // Thirdparty/sophus/sophus/se3.hpp:40
// Sophus::SE3Group<float> on (unit quaternion, translation) storage — Thirdparty/sophus/sophus/se3.hpp, so3.hpp.
class SE3f {
public:
    Eigen::Quaternionf q;
    Eigen::Vector3f t;
    // A pose given as a rotation MATRIX keeps that matrix verbatim (real Sophus converts to a quaternion and back): the
    // test vectors name Rcw / tcw as float matrices and the restatement must see the very same numbers.
    bool hasR = false;
    Eigen::Matrix3f R;
    SE3f() : q(1, 0, 0, 0), t(0, 0, 0) {}
    SE3f(const Eigen::Quaternionf &q_, const Eigen::Vector3f &t_) : q(q_), t(t_) {}
    SE3f(const Eigen::Matrix3f &R_, const Eigen::Vector3f &t_) : q(1, 0, 0, 0), t(t_), hasR(true), R(R_) {}
    Eigen::Matrix3f rotationMatrix() const { return hasR ? R : q.toRotationMatrix(); }
    const Eigen::Vector3f &translation() const { return t; }
    Eigen::Vector3f &translation() { return t; }
    const Eigen::Quaternionf &unit_quaternion() const { return q; }
    Eigen::Vector3f operator*(const Eigen::Vector3f &p) const { return q._transformVector(p) + t; }   // se3.hpp: so3() * p + translation()
    SE3f inverse() const {   // se3.hpp:  invR = so3().inverse();  SE3(invR, invR * (translation() * -1))
        const Eigen::Quaternionf c = q.conjugate();
        return SE3f(c, c._transformVector(t * -1.0f));
    }
    SE3f operator*(const SE3f &o) const {   // se3.hpp operator*=: translation() += so3() * other.translation(); so3() *= other.so3()
        SE3f r(*this);
        r.t += q._transformVector(o.t);
        Eigen::Quaternionf p = q * o.q;
        // so3.hpp:268-276 normalize(): coeffs() /= norm()
        const float length = std::sqrt(p.squaredNorm());
        p.qx /= length; p.qy /= length; p.qz /= length; p.qw /= length;
        r.q = p;
        return r;
    }
    // SE3Group::exp (Thirdparty/sophus/sophus/se3.hpp:407-428) over SO3Group::expAndTheta (so3.hpp:426-456); a = (upsilon, omega).
    // The SO3Group(Quaternion) constructor normalises (so3.hpp:739-743 -> :268-276); V uses theta * theta, not the squared
    // norm it took the root of (se3.hpp:422).
    static SE3f exp(const Eigen::VecF<6> &a) {
        const Eigen::Vector3f omega(a[3], a[4], a[5]), upsilon(a[0], a[1], a[2]);
        const float theta_sq = omega.squaredNorm();
        const float theta = std::sqrt(theta_sq);
        const float half_theta = 0.5f * theta;
        const float eps = 1e-5f;   // SophusConstants<float>::epsilon() (sophus.hpp)
        float imag_factor, real_factor;
        if (theta < eps) {
            const float theta_po4 = theta_sq * theta_sq;
            imag_factor = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * theta_po4;
            real_factor = 1.0f - 0.5f * theta_sq + (float)(1.0 / 384.0) * theta_po4;
        } else {
            const float sin_half_theta = std::sin(half_theta);
            imag_factor = sin_half_theta / theta;
            real_factor = std::cos(half_theta);
        }
        Eigen::Quaternionf q(real_factor, imag_factor * omega.x(), imag_factor * omega.y(), imag_factor * omega.z());
        const float length = std::sqrt(q.squaredNorm());
        q.qx /= length; q.qy /= length; q.qz /= length; q.qw /= length;
        Eigen::Matrix3f Omega;   // SO3Group::hat
        Omega(0, 0) = 0.0f; Omega(0, 1) = -omega[2]; Omega(0, 2) = omega[1];
        Omega(1, 0) = omega[2]; Omega(1, 1) = 0.0f; Omega(1, 2) = -omega[0];
        Omega(2, 0) = -omega[1]; Omega(2, 1) = omega[0]; Omega(2, 2) = 0.0f;
        const Eigen::Matrix3f Omega_sq = Omega * Omega;
        Eigen::Matrix3f V;
        if (theta < eps) {
            V = q.toRotationMatrix();
        } else {
            const float theta_sq2 = theta * theta;
            const float c1 = (1.0f - std::cos(theta)) / theta_sq2, c2 = (theta - std::sin(theta)) / (theta_sq2 * theta);
            for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) V(r, c) = ((r == c ? 1.0f : 0.0f) + c1 * Omega(r, c)) + c2 * Omega_sq(r, c);
        }
        return SE3f(q, V * upsilon);
    }
};
}  // namespace Sophus
#endif
