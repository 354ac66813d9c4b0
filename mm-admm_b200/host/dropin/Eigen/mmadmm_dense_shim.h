// Minimal dense-matrix types under the Eigen names the MM-ADMM driver layer uses.
//
// The reference's main.cpp, src/MeshUtils.h callers and Experiments/TestMonitors/MEx*.h are written against
// Eigen 3.4 (lib/eigen @ f0f1d79, an un-fetched submodule: SURVEY.md F4).  None of that code is on the hot
// path -- it builds the initial mesh and evaluates the user's monitor tensor once per vertex -- so this
// header only has to make those translation units compile and produce the same VALUES.  It is NOT Eigen
// and has none of Eigen's expression templates: every operator returns by value.  Storage is column-major
// like Eigen's default (the driver never depends on the storage order, only on (i,j) values).
//
// Surface provided: Matrix<T,R,C>, Vector<T,N>, MatrixXd, MatrixXi, VectorXd; (rows,cols) /
// (size) constructors; rows/cols/size/resize/conservativeResize/setZero; operator()(i,j), (i), [i];
// operator()(i, placeholders::all) row proxies (read, assign, +=); row(i); Identity / Constant / Random;
// norm / squaredNorm / normalize / transpose / determinant (<= 3x3); + - * / with scalars and matrices;
// *= /= += -=; the comma initialiser `v << a, b;`.
#ifndef MMADMM_EIGEN_DENSE_SHIM_H
#define MMADMM_EIGEN_DENSE_SHIM_H
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <type_traits>
#include <vector>

namespace Eigen {

constexpr int Dynamic = -1;
namespace placeholders { struct all_t {}; static constexpr all_t all{}; }

template <typename T, int R, int C> class Matrix;

namespace shim_detail {
// fixed-size storage lives inline, dynamic storage on the heap
template <typename T, int N> struct Store {
    T v[N > 0 ? N : 1];
    void alloc(size_t) {}
    T *ptr() { return v; }
    const T *ptr() const { return v; }
};
template <typename T> struct Store<T, Dynamic> {
    std::vector<T> v;
    void alloc(size_t n) { v.resize(n); }
    T *ptr() { return v.data(); }
    const T *ptr() const { return v.data(); }
};
template <int R, int C> struct Count { static constexpr int value = (R == Dynamic || C == Dynamic) ? Dynamic : R * C; };

template <typename M> class CommaInit {
    M &m; int k;
public:
    CommaInit(M &m_, typename M::Scalar first) : m(m_), k(0) { put(first); }
    void put(typename M::Scalar s) {
        assert(k < m.size());
        // Eigen fills row by row
        int i = k / m.cols(), j = k % m.cols();
        m(i, j) = s; k++;
    }
    CommaInit &operator,(typename M::Scalar s) { put(s); return *this; }
};
}  // namespace shim_detail

// one row of a (dynamic) matrix: what `A(i, Eigen::placeholders::all)` and `A.row(i)` return
template <typename M> class RowRef {
    M *m; int i;
public:
    using Scalar = typename M::Scalar;
    RowRef(M *m_, int i_) : m(m_), i(i_) {}
    int size() const { return m->cols(); }
    Scalar operator()(int j) const { return (*m)(i, j); }
    Scalar operator[](int j) const { return (*m)(i, j); }
    template <int N, int K> RowRef &operator=(const Matrix<Scalar, N, K> &v) {
        assert(v.size() == size());
        for (int j = 0; j < size(); j++) (*m)(i, j) = v[j];
        return *this;
    }
    template <typename M2> RowRef &assign(const RowRef<M2> &o) {
        assert(o.size() == size());
        for (int j = 0; j < size(); j++) (*m)(i, j) = o(j);
        return *this;
    }
    RowRef &operator=(const RowRef &o) { return assign(o); }
    template <typename M2> RowRef &operator=(const RowRef<M2> &o) { return assign(o); }
    template <int N, int K> RowRef &operator+=(const Matrix<Scalar, N, K> &v) {
        for (int j = 0; j < size(); j++) (*m)(i, j) += v[j];
        return *this;
    }
    template <int N, int K> RowRef &operator-=(const Matrix<Scalar, N, K> &v) {
        for (int j = 0; j < size(); j++) (*m)(i, j) -= v[j];
        return *this;
    }
};

template <typename T, int R, int C> class Matrix {
    shim_detail::Store<T, shim_detail::Count<R, C>::value> s;
    int r_, c_;
public:
    using Scalar = T;
    Matrix() : r_(R == Dynamic ? 0 : R), c_(C == Dynamic ? 0 : C) { s.alloc((size_t)r_ * c_); }
    Matrix(int rows, int cols) : r_(rows), c_(cols) {
        assert((R == Dynamic || R == rows) && (C == Dynamic || C == cols));
        s.alloc((size_t)r_ * c_);
    }
    // vector of a given length (Eigen: VectorXd v(n)); for fixed sizes the argument must match
    explicit Matrix(int n) : r_(R == Dynamic ? n : R), c_(C == Dynamic ? (R == Dynamic ? 1 : n) : C) {
        static_assert(R == 1 || C == 1 || R == Dynamic || C == Dynamic, "size constructor is for vectors");
        s.alloc((size_t)r_ * c_);
    }
    Matrix(const Matrix &) = default;
    Matrix &operator=(const Matrix &) = default;
    template <int R2, int C2> Matrix(const Matrix<T, R2, C2> &o) : r_(o.rows()), c_(o.cols()) {
        assert((R == Dynamic || R == r_) && (C == Dynamic || C == c_));
        s.alloc((size_t)r_ * c_);
        for (int k = 0; k < size(); k++) s.ptr()[k] = o.data()[k];
    }
    // a matrix row converts into a (column) vector, as Eigen allows for vectors
    template <typename M> Matrix(const RowRef<M> &row) : r_(R == Dynamic ? row.size() : R), c_(C == Dynamic ? 1 : C) {
        s.alloc((size_t)r_ * c_);
        assert(size() == row.size());
        for (int k = 0; k < size(); k++) s.ptr()[k] = row(k);
    }
    template <typename M> Matrix &operator=(const RowRef<M> &row) {
        if (R == Dynamic || C == Dynamic) resize(R == 1 ? 1 : row.size(), R == 1 ? row.size() : 1);
        assert(size() == row.size());
        for (int k = 0; k < size(); k++) s.ptr()[k] = row(k);
        return *this;
    }

    int rows() const { return r_; }
    int cols() const { return c_; }
    int size() const { return r_ * c_; }
    T *data() { return s.ptr(); }
    const T *data() const { return s.ptr(); }
    void resize(int rows, int cols) {
        assert((R == Dynamic || R == rows) && (C == Dynamic || C == cols));
        r_ = rows; c_ = cols; s.alloc((size_t)r_ * c_);
    }
    void resize(int n) { if (C == 1 || R == Dynamic) resize(n, 1); else resize(1, n); }
    void conservativeResize(int rows, int cols) {
        Matrix old(*this);
        resize(rows, cols);
        for (int j = 0; j < cols && j < old.cols(); j++)
            for (int i = 0; i < rows && i < old.rows(); i++) (*this)(i, j) = old(i, j);
    }
    void setZero() { for (int k = 0; k < size(); k++) s.ptr()[k] = T(0); }
    void setConstant(T v) { for (int k = 0; k < size(); k++) s.ptr()[k] = v; }

    T &operator()(int i, int j) { assert(i >= 0 && i < r_ && j >= 0 && j < c_); return s.ptr()[(size_t)j * r_ + i]; }
    const T &operator()(int i, int j) const { assert(i >= 0 && i < r_ && j >= 0 && j < c_); return s.ptr()[(size_t)j * r_ + i]; }
    T &operator()(int i) { assert(i >= 0 && i < size()); return s.ptr()[i]; }
    const T &operator()(int i) const { assert(i >= 0 && i < size()); return s.ptr()[i]; }
    T &operator[](int i) { return (*this)(i); }
    const T &operator[](int i) const { return (*this)(i); }
    RowRef<Matrix> operator()(int i, placeholders::all_t) { return RowRef<Matrix>(this, i); }
    RowRef<const Matrix> operator()(int i, placeholders::all_t) const { return RowRef<const Matrix>(this, i); }
    RowRef<Matrix> row(int i) { return RowRef<Matrix>(this, i); }
    RowRef<const Matrix> row(int i) const { return RowRef<const Matrix>(this, i); }

    static Matrix Zero(int rows, int cols) { Matrix m(rows, cols); m.setZero(); return m; }
    static Matrix Identity(int rows, int cols) {
        Matrix m(rows, cols); m.setZero();
        for (int i = 0; i < rows && i < cols; i++) m(i, i) = T(1);
        return m;
    }
    static Matrix Identity() { static_assert(R != Dynamic && C != Dynamic, "fixed size only"); return Identity(R, C); }
    static Matrix Constant(T v) { Matrix m; m.setConstant(v); return m; }
    static Matrix Constant(int n, T v) { Matrix m(n); m.setConstant(v); return m; }
    static Matrix Constant(int rows, int cols, T v) { Matrix m(rows, cols); m.setConstant(v); return m; }
    // uniform in [-1, 1] from rand(), like Eigen's default random generator
    static Matrix Random(int rows, int cols) {
        Matrix m(rows, cols);
        for (int k = 0; k < m.size(); k++) m.data()[k] = T(2.0 * std::rand() / (double)RAND_MAX - 1.0);
        return m;
    }
    static Matrix Random(int n) { Matrix m(n); return Random(m.rows(), m.cols()); }
    static Matrix Random() { Matrix m; return Random(m.rows(), m.cols()); }

    T squaredNorm() const { T a = T(0); for (int k = 0; k < size(); k++) a += data()[k] * data()[k]; return a; }
    T norm() const { return std::sqrt(squaredNorm()); }
    T sum() const { T a = T(0); for (int k = 0; k < size(); k++) a += data()[k]; return a; }
    void normalize() { T n = norm(); for (int k = 0; k < size(); k++) data()[k] /= n; }
    Matrix<T, C, R> transpose() const {
        Matrix<T, C, R> t(c_, r_);
        for (int j = 0; j < c_; j++) for (int i = 0; i < r_; i++) t(j, i) = (*this)(i, j);
        return t;
    }
    T determinant() const {
        assert(r_ == c_ && r_ >= 1 && r_ <= 3);
        const Matrix &a = *this;
        if (r_ == 1) return a(0, 0);
        if (r_ == 2) return a(0, 0) * a(1, 1) - a(0, 1) * a(1, 0);
        return a(0, 0) * (a(1, 1) * a(2, 2) - a(1, 2) * a(2, 1)) - a(0, 1) * (a(1, 0) * a(2, 2) - a(1, 2) * a(2, 0)) +
               a(0, 2) * (a(1, 0) * a(2, 1) - a(1, 1) * a(2, 0));
    }

    Matrix &operator*=(T v) { for (int k = 0; k < size(); k++) data()[k] *= v; return *this; }
    Matrix &operator/=(T v) { for (int k = 0; k < size(); k++) data()[k] /= v; return *this; }
    template <int R2, int C2> Matrix &operator+=(const Matrix<T, R2, C2> &o) {
        assert(o.rows() == r_ && o.cols() == c_);
        for (int k = 0; k < size(); k++) data()[k] += o.data()[k];
        return *this;
    }
    template <int R2, int C2> Matrix &operator-=(const Matrix<T, R2, C2> &o) {
        assert(o.rows() == r_ && o.cols() == c_);
        for (int k = 0; k < size(); k++) data()[k] -= o.data()[k];
        return *this;
    }
    Matrix operator-() const { Matrix m(*this); for (int k = 0; k < size(); k++) m.data()[k] = -m.data()[k]; return m; }
    shim_detail::CommaInit<Matrix> operator<<(T first) { return shim_detail::CommaInit<Matrix>(*this, first); }
};

template <typename T, int N> using Vector = Matrix<T, N, 1>;
template <typename T, int N> using RowVector = Matrix<T, 1, N>;
using MatrixXd = Matrix<double, Dynamic, Dynamic>;
using MatrixXi = Matrix<int, Dynamic, Dynamic>;
using VectorXd = Matrix<double, Dynamic, 1>;
using VectorXi = Matrix<int, Dynamic, 1>;
using Matrix2d = Matrix<double, 2, 2>;
using Matrix3d = Matrix<double, 3, 3>;
using Vector2d = Matrix<double, 2, 1>;
using Vector3d = Matrix<double, 3, 1>;

template <typename T, int R, int C, int R2, int C2>
Matrix<T, R, C> operator+(const Matrix<T, R, C> &a, const Matrix<T, R2, C2> &b) { Matrix<T, R, C> m(a); m += b; return m; }
template <typename T, int R, int C, int R2, int C2>
Matrix<T, R, C> operator-(const Matrix<T, R, C> &a, const Matrix<T, R2, C2> &b) { Matrix<T, R, C> m(a); m -= b; return m; }
template <typename T, int R, int C, typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
Matrix<T, R, C> operator*(const Matrix<T, R, C> &a, S s) { Matrix<T, R, C> m(a); m *= T(s); return m; }
template <typename T, int R, int C, typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
Matrix<T, R, C> operator*(S s, const Matrix<T, R, C> &a) { Matrix<T, R, C> m(a); m *= T(s); return m; }
template <typename T, int R, int C, typename S, typename = std::enable_if_t<std::is_arithmetic<S>::value>>
Matrix<T, R, C> operator/(const Matrix<T, R, C> &a, S s) { Matrix<T, R, C> m(a); m /= T(s); return m; }
template <typename T, int R, int K, int K2, int C>
Matrix<T, R, C> operator*(const Matrix<T, R, K> &a, const Matrix<T, K2, C> &b) {
    assert(a.cols() == b.rows());
    Matrix<T, R, C> m(a.rows(), b.cols());
    for (int i = 0; i < a.rows(); i++)
        for (int j = 0; j < b.cols(); j++) {
            T acc = T(0);
            for (int k = 0; k < a.cols(); k++) acc += a(i, k) * b(k, j);
            m(i, j) = acc;
        }
    return m;
}
// rows on either side of + and - (e.g. Xp(a, all) - Xp(b, all)) give a column vector
template <typename M1, typename M2>
Matrix<typename M1::Scalar, Dynamic, 1> operator-(const RowRef<M1> &a, const RowRef<M2> &b) {
    Matrix<typename M1::Scalar, Dynamic, 1> m(a.size());
    for (int k = 0; k < a.size(); k++) m[k] = a(k) - b(k);
    return m;
}
template <typename M1, typename M2>
Matrix<typename M1::Scalar, Dynamic, 1> operator+(const RowRef<M1> &a, const RowRef<M2> &b) {
    Matrix<typename M1::Scalar, Dynamic, 1> m(a.size());
    for (int k = 0; k < a.size(); k++) m[k] = a(k) + b(k);
    return m;
}

}  // namespace Eigen
#endif
