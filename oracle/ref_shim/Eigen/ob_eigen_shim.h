// ob_eigen_shim.h — a minimal stand-in for the subset of Eigen that libopenbeam's static-solve path uses, so that the
// reference's UNMODIFIED sources (/root/reference/libopenbeam/src/*.cpp) can be compiled here, where Eigen itself is
// not available (TEST INFRASTRUCTURE: oracle/_ref; see oracle/Makefile.ref and DESIGN.md).
//
// Everything is evaluated eagerly with textbook, strictly sequential summation orders:
//   dense products        c(i,j) = ((a(i,0) b(0,j) + a(i,1) b(1,j)) + ...)            (k ascending)
//   setFromTriplets       duplicates summed in triplet order, explicit zeros kept      (Eigen <= 3.4 semantics)
//   sparse * dense vector y(i) accumulates column by column, j ascending               (column-major storage)
//   LLT                   unblocked lower Cholesky, fails on a pivot <= 0              (Eigen's llt_inplace::unblocked)
// Real Eigen vectorises some reductions (SSE2 packets), so values may differ from a true Eigen build in the last
// bits; integer artefacts (patterns, index lists) cannot.
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <iostream>
#include <map>
#include <stdexcept>
#include <type_traits>
#include <utility>
#include <vector>

namespace Eigen {

constexpr int Dynamic = -1;
enum ComputationInfo { Success = 0, NumericalIssue = 1, NoConvergence = 2, InvalidInput = 3 };
using Index = std::ptrdiff_t;

template <class T, int R, int C> class Matrix;
template <class T, int BR, int BC, bool Const> class Block;
template <class T> class SparseMatrix;
template <class D> struct traits;

namespace detail {
template <class T, int R, int C, bool Fixed = (R != Dynamic && C != Dynamic)> struct Storage;
template <class T, int R, int C> struct Storage<T, R, C, true> {
  T d[R * C];
  Storage() { for (int i = 0; i < R * C; i++) d[i] = T(0); }
  void resize(Index r, Index c) { assert(r == R && c == C); (void)r; (void)c; }
  Index rows() const { return R; }
  Index cols() const { return C; }
  T* data() { return d; }
  const T* data() const { return d; }
};
template <class T, int R, int C> struct Storage<T, R, C, false> {
  std::vector<T> d;
  Index r_ = (R == Dynamic ? 0 : R), c_ = (C == Dynamic ? 0 : C);
  void resize(Index r, Index c) { r_ = r; c_ = c; d.assign(size_t(r * c), T(0)); }
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  T* data() { return d.data(); }
  const T* data() const { return d.data(); }
};
}  // namespace detail

template <class T> struct ArrayWrap;

// CRTP base: anything with rows(), cols(), coeff(i,j)
template <class D> struct MatrixBase {
  using Scalar = typename traits<D>::Scalar;
  using Plain = Matrix<Scalar, traits<D>::Rows, traits<D>::Cols>;
  using PlainT = Matrix<Scalar, traits<D>::Cols, traits<D>::Rows>;
  const D& derived() const { return static_cast<const D&>(*this); }
  D& derived() { return static_cast<D&>(*this); }
  Index rows() const { return derived().rows(); }
  Index cols() const { return derived().cols(); }
  Index size() const { return rows() * cols(); }
  Scalar coeff(Index i, Index j) const { return derived().coeff(i, j); }
  Scalar coeff(Index i) const { return cols() == 1 ? coeff(i, 0) : coeff(0, i); }
  Scalar operator()(Index i, Index j) const { return coeff(i, j); }
  Scalar operator()(Index i) const { return coeff(i); }
  Scalar operator[](Index i) const { return coeff(i); }
  Scalar x() const { return coeff(0); }
  Scalar y() const { return coeff(1); }
  Scalar z() const { return coeff(2); }
  Plain eval() const {
    Plain m(rows(), cols());
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) m(i, j) = coeff(i, j);
    return m;
  }
  PlainT transpose() const {
    PlainT m(cols(), rows());
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) m(j, i) = coeff(i, j);
    return m;
  }
  PlainT adjoint() const { return transpose(); }
  ArrayWrap<Scalar> array() const;
  Scalar squaredNorm() const {
    Scalar s = 0;
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) s += coeff(i, j) * coeff(i, j);
    return s;
  }
  Scalar norm() const { return std::sqrt(squaredNorm()); }
  Scalar sum() const {
    Scalar s = 0;
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) s += coeff(i, j);
    return s;
  }
  Block<Scalar, Dynamic, Dynamic, true> block(Index i0, Index j0, Index r, Index c) const;
};

// coefficient-wise view (a copy): just what the reference asks of .array()
template <class T> struct BoolArray {
  std::vector<bool> v;
  bool all() const { for (bool b : v) if (!b) return false; return true; }
  bool any() const { for (bool b : v) if (b) return true; return false; }
};
template <class T> struct ArrayWrap {
  std::vector<T> v;
  BoolArray<T> operator==(T x) const { BoolArray<T> b; for (T e : v) b.v.push_back(e == x); return b; }
  BoolArray<T> operator!=(T x) const { BoolArray<T> b; for (T e : v) b.v.push_back(e != x); return b; }
  ArrayWrap square() const { ArrayWrap a; for (T e : v) a.v.push_back(e * e); return a; }
  ArrayWrap abs() const { ArrayWrap a; for (T e : v) a.v.push_back(std::abs(e)); return a; }
  T sum() const { T s = 0; for (T e : v) s += e; return s; }
  T maxCoeff() const { T m = v.empty() ? T(0) : v[0]; for (T e : v) m = std::max(m, e); return m; }
};
template <class D> ArrayWrap<typename MatrixBase<D>::Scalar> MatrixBase<D>::array() const {
  ArrayWrap<Scalar> a;
  for (Index j = 0; j < cols(); j++)
    for (Index i = 0; i < rows(); i++) a.v.push_back(coeff(i, j));
  return a;
}

template <class T, int R, int C> struct traits<Matrix<T, R, C>> {
  using Scalar = T;
  static constexpr int Rows = R, Cols = C;
};
template <class T, int BR, int BC, bool Const> struct traits<Block<T, BR, BC, Const>> {
  using Scalar = T;
  static constexpr int Rows = BR, Cols = BC;
};

template <class MatrixType> class LLT;

template <class T, int R, int C> class Matrix : public MatrixBase<Matrix<T, R, C>> {
  detail::Storage<T, R, C> s_;

 public:
  using Scalar = T;
  using Base = MatrixBase<Matrix<T, R, C>>;
  static constexpr int RowsAtCompileTime = R, ColsAtCompileTime = C;
  Matrix() {}
  explicit Matrix(Index n) {  // vector of size n (or fixed-size no-op)
    if (R == Dynamic && C == 1) s_.resize(n, 1);
    else if (C == Dynamic && R == 1) s_.resize(1, n);
    else s_.resize(R == Dynamic ? n : R, C == Dynamic ? n : C);
  }
  Matrix(Index r, Index c) { s_.resize(r, c); }
  // 3-vector / 2-vector from coefficients (Eigen's fixed-size vector constructors)
  template <int RR = R, int CC = C, class = typename std::enable_if<RR == 3 && CC == 1>::type>
  Matrix(T x, T y, T z) { s_.d[0] = x; s_.d[1] = y; s_.d[2] = z; }
  template <class O> Matrix(const MatrixBase<O>& o) { *this = o; }
  explicit Matrix(const SparseMatrix<T>& sp);
  template <class O> Matrix& operator=(const MatrixBase<O>& o) {
    s_.resize(o.rows(), o.cols());
    // evaluate first when the source may alias this matrix (it is eager: o is already a plain value or a block)
    std::vector<T> tmp(size_t(o.rows() * o.cols()));
    for (Index j = 0; j < o.cols(); j++)
      for (Index i = 0; i < o.rows(); i++) tmp[size_t(j * o.rows() + i)] = o.coeff(i, j);
    std::copy(tmp.begin(), tmp.end(), s_.data());
    return *this;
  }
  Index rows() const { return s_.rows(); }
  Index cols() const { return s_.cols(); }
  Index size() const { return rows() * cols(); }
  T coeff(Index i, Index j) const { return s_.data()[j * rows() + i]; }
  T coeff(Index i) const { return s_.data()[i]; }
  T& coeffRef(Index i, Index j) { return s_.data()[j * rows() + i]; }
  T& operator()(Index i, Index j) { return s_.data()[j * rows() + i]; }
  T operator()(Index i, Index j) const { return s_.data()[j * rows() + i]; }
  T& operator()(Index i) { return s_.data()[i]; }
  T operator()(Index i) const { return s_.data()[i]; }
  T& operator[](Index i) { return s_.data()[i]; }
  T operator[](Index i) const { return s_.data()[i]; }
  T* data() { return s_.data(); }
  const T* data() const { return s_.data(); }
  using Base::x;
  using Base::y;
  using Base::z;
  T& x() { return s_.data()[0]; }
  T& y() { return s_.data()[1]; }
  T& z() { return s_.data()[2]; }
  void resize(Index n) {
    if (C == 1 || (C == Dynamic && R != 1)) s_.resize(n, C == Dynamic ? 1 : C);
    else s_.resize(1, n);
  }
  void resize(Index r, Index c) { s_.resize(r, c); }
  Matrix& setZero() { for (Index i = 0; i < size(); i++) s_.data()[i] = T(0); return *this; }
  Matrix& setZero(Index n) { resize(n); return setZero(); }
  Matrix& setZero(Index r, Index c) { resize(r, c); return setZero(); }
  Matrix& setConstant(T v) { for (Index i = 0; i < size(); i++) s_.data()[i] = v; return *this; }
  Matrix& fill(T v) { return setConstant(v); }
  Matrix& setIdentity() {
    setZero();
    for (Index i = 0; i < std::min(rows(), cols()); i++) (*this)(i, i) = T(1);
    return *this;
  }
  static Matrix Zero() { Matrix m; m.setZero(); return m; }
  static Matrix Zero(Index n) { Matrix m(n); m.setZero(); return m; }
  static Matrix Zero(Index r, Index c) { Matrix m(r, c); m.setZero(); return m; }
  static Matrix Identity() { Matrix m; m.setIdentity(); return m; }
  static Matrix Identity(Index r, Index c) { Matrix m(r, c); m.setIdentity(); return m; }
  static Matrix Constant(T v) { Matrix m; m.setConstant(v); return m; }
  Matrix& noalias() { return *this; }
  using Base::block;
  Block<T, Dynamic, Dynamic, false> block(Index i0, Index j0, Index r, Index c);
  template <int BR, int BC> Block<T, BR, BC, false> block(Index i0, Index j0);
  template <int BR, int BC> Block<T, BR, BC, true> block(Index i0, Index j0) const;
  template <class O> Matrix& operator+=(const MatrixBase<O>& o) {
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) (*this)(i, j) += o.coeff(i, j);
    return *this;
  }
  template <class O> Matrix& operator-=(const MatrixBase<O>& o) {
    for (Index j = 0; j < cols(); j++)
      for (Index i = 0; i < rows(); i++) (*this)(i, j) -= o.coeff(i, j);
    return *this;
  }
  Matrix& operator*=(T v) { for (Index i = 0; i < size(); i++) s_.data()[i] *= v; return *this; }
  Matrix& operator/=(T v) { for (Index i = 0; i < size(); i++) s_.data()[i] /= v; return *this; }
  LLT<Matrix> llt() const;
  T determinant() const {
    assert(rows() == cols());
    if (rows() == 2) return coeff(0, 0) * coeff(1, 1) - coeff(0, 1) * coeff(1, 0);
    if (rows() == 3)
      return coeff(0, 0) * (coeff(1, 1) * coeff(2, 2) - coeff(1, 2) * coeff(2, 1)) -
             coeff(0, 1) * (coeff(1, 0) * coeff(2, 2) - coeff(1, 2) * coeff(2, 0)) +
             coeff(0, 2) * (coeff(1, 0) * coeff(2, 1) - coeff(1, 1) * coeff(2, 0));
    throw std::logic_error("ob_eigen_shim: determinant only for 2x2 / 3x3");
  }
};

// a rectangular window into a Matrix (reference semantics on the left-hand side, value semantics when read)
template <class T, int BR, int BC, bool Const> class Block : public MatrixBase<Block<T, BR, BC, Const>> {
  using Ptr = typename std::conditional<Const, const T*, T*>::type;
  Ptr base_;
  Index ld_, r_, c_;

 public:
  Block(Ptr base, Index ld, Index r, Index c) : base_(base), ld_(ld), r_(r), c_(c) {}
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  T coeff(Index i, Index j) const { return base_[j * ld_ + i]; }
  T coeff(Index i) const { return c_ == 1 ? coeff(i, 0) : coeff(0, i); }
  template <bool CC = Const, class = typename std::enable_if<!CC>::type> T& operator()(Index i, Index j) { return base_[j * ld_ + i]; }
  T operator()(Index i, Index j) const { return coeff(i, j); }
  Block& noalias() { return *this; }
  template <class O> Block& operator=(const MatrixBase<O>& o) {
    static_assert(!Const, "assignment to a const block");
    assert(o.rows() == r_ && o.cols() == c_);
    std::vector<T> tmp(size_t(r_ * c_));
    for (Index j = 0; j < c_; j++)
      for (Index i = 0; i < r_; i++) tmp[size_t(j * r_ + i)] = o.coeff(i, j);
    for (Index j = 0; j < c_; j++)
      for (Index i = 0; i < r_; i++) base_[j * ld_ + i] = tmp[size_t(j * r_ + i)];
    return *this;
  }
  Block& operator=(const Block& o) { return this->template operator=<Block>(o); }
  template <class O> Block& operator+=(const MatrixBase<O>& o) {
    for (Index j = 0; j < c_; j++)
      for (Index i = 0; i < r_; i++) base_[j * ld_ + i] += o.coeff(i, j);
    return *this;
  }
  template <class O> Block& operator-=(const MatrixBase<O>& o) {
    for (Index j = 0; j < c_; j++)
      for (Index i = 0; i < r_; i++) base_[j * ld_ + i] -= o.coeff(i, j);
    return *this;
  }
  Block& setZero() {
    for (Index j = 0; j < c_; j++)
      for (Index i = 0; i < r_; i++) base_[j * ld_ + i] = T(0);
    return *this;
  }
};
template <class D>
Block<typename MatrixBase<D>::Scalar, Dynamic, Dynamic, true> MatrixBase<D>::block(Index i0, Index j0, Index r, Index c) const {
  // only plain matrices hand out windows (a block of a block is not needed by the reference)
  const auto& m = derived();
  return Block<Scalar, Dynamic, Dynamic, true>(m.data() + j0 * m.rows() + i0, m.rows(), r, c);
}
template <class T, int R, int C> Block<T, Dynamic, Dynamic, false> Matrix<T, R, C>::block(Index i0, Index j0, Index r, Index c) {
  return Block<T, Dynamic, Dynamic, false>(data() + j0 * rows() + i0, rows(), r, c);
}
template <class T, int R, int C> template <int BR, int BC> Block<T, BR, BC, false> Matrix<T, R, C>::block(Index i0, Index j0) {
  return Block<T, BR, BC, false>(data() + j0 * rows() + i0, rows(), BR, BC);
}
template <class T, int R, int C> template <int BR, int BC> Block<T, BR, BC, true> Matrix<T, R, C>::block(Index i0, Index j0) const {
  return Block<T, BR, BC, true>(data() + j0 * rows() + i0, rows(), BR, BC);
}

// ---- eager arithmetic ----
template <class A, class B>
Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<B>::Cols> operator*(const MatrixBase<A>& a, const MatrixBase<B>& b) {
  assert(a.cols() == b.rows());
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<B>::Cols> c(a.rows(), b.cols());
  for (Index j = 0; j < b.cols(); j++)
    for (Index i = 0; i < a.rows(); i++) {
      typename traits<A>::Scalar s = 0;
      for (Index k = 0; k < a.cols(); k++) {
        const auto t = a.coeff(i, k) * b.coeff(k, j);
        s = k == 0 ? t : s + t;
      }
      c(i, j) = s;
    }
  return c;
}
template <class A> typename MatrixBase<A>::Plain operator*(const MatrixBase<A>& a, typename traits<A>::Scalar v) {
  auto m = a.eval();
  m *= v;
  return m;
}
template <class A> typename MatrixBase<A>::Plain operator*(typename traits<A>::Scalar v, const MatrixBase<A>& a) {
  auto m = a.eval();
  for (Index i = 0; i < m.size(); i++) m.data()[i] = v * m.data()[i];
  return m;
}
template <class A> typename MatrixBase<A>::Plain operator/(const MatrixBase<A>& a, typename traits<A>::Scalar v) {
  auto m = a.eval();
  m /= v;
  return m;
}
template <class A, class B> typename MatrixBase<A>::Plain operator+(const MatrixBase<A>& a, const MatrixBase<B>& b) {
  auto m = a.eval();
  m += b;
  return m;
}
template <class A, class B> typename MatrixBase<A>::Plain operator-(const MatrixBase<A>& a, const MatrixBase<B>& b) {
  auto m = a.eval();
  m -= b;
  return m;
}
template <class A> typename MatrixBase<A>::Plain operator-(const MatrixBase<A>& a) {
  auto m = a.eval();
  for (Index i = 0; i < m.size(); i++) m.data()[i] = -m.data()[i];
  return m;
}
template <class A> std::ostream& operator<<(std::ostream& o, const MatrixBase<A>& a) {
  for (Index i = 0; i < a.rows(); i++) {
    for (Index j = 0; j < a.cols(); j++) o << (j ? " " : "") << a.coeff(i, j);
    if (i + 1 < a.rows()) o << "\n";
  }
  return o;
}

using MatrixXd = Matrix<double, Dynamic, Dynamic>;
using VectorXd = Matrix<double, Dynamic, 1>;
using Matrix3d = Matrix<double, 3, 3>;
using Vector3d = Matrix<double, 3, 1>;

// ---- dense Cholesky (lower), Eigen::LLT ----
template <class MatrixType> class LLT {
  MatrixType L_;
  ComputationInfo info_ = Success;

 public:
  using T = typename MatrixType::Scalar;
  LLT() {}
  explicit LLT(const MatrixType& a) { compute(a); }
  LLT& compute(const MatrixType& a) {
    // llt_inplace<Scalar, Lower>::unblocked: reads the LOWER triangle only.  Worked on a row-major copy so that the two
    // rows of every inner product are contiguous (same operations in the same order, only the memory layout differs;
    // Eigen's own kernel is blocked and vectorised, i.e. faster still)
    const Index n = a.rows();
    std::vector<T> R(size_t(n) * size_t(n), T(0));
    for (Index i = 0; i < n; i++)
      for (Index j = 0; j <= i; j++) R[size_t(i) * n + j] = a(i, j);
    info_ = Success;
    for (Index k = 0; k < n && info_ == Success; k++) {
      const T* rk = &R[size_t(k) * n];
      T x = rk[k];
      for (Index j = 0; j < k; j++) x -= rk[j] * rk[j];
      if (!(x > T(0))) { info_ = NumericalIssue; break; }
      x = std::sqrt(x);
      R[size_t(k) * n + k] = x;
      for (Index i = k + 1; i < n; i++) {
        T* ri = &R[size_t(i) * n];
        T s = ri[k];
        for (Index j = 0; j < k; j++) s -= ri[j] * rk[j];
        ri[k] = s / x;
      }
    }
    L_.resize(n, n);
    for (Index j = 0; j < n; j++)
      for (Index i = 0; i < n; i++) L_(i, j) = i >= j ? R[size_t(i) * n + j] : T(0);
    return *this;
  }
  ComputationInfo info() const { return info_; }
  const MatrixType& matrixL() const { return L_; }
  template <class B> Matrix<T, Dynamic, traits<B>::Cols> solve(const MatrixBase<B>& b) const {
    const Index n = L_.rows();
    Matrix<T, Dynamic, traits<B>::Cols> x(b.rows(), b.cols());
    for (Index c = 0; c < b.cols(); c++) {
      for (Index i = 0; i < n; i++) {  // L y = b
        T s = b.coeff(i, c);
        for (Index j = 0; j < i; j++) s -= L_(i, j) * x(j, c);
        x(i, c) = s / L_(i, i);
      }
      for (Index i = n - 1; i >= 0; i--) {  // L^T x = y
        T s = x(i, c);
        for (Index j = i + 1; j < n; j++) s -= L_(j, i) * x(j, c);
        x(i, c) = s / L_(i, i);
      }
    }
    return x;
  }
};
template <class T, int R, int C> LLT<Matrix<T, R, C>> Matrix<T, R, C>::llt() const { return LLT<Matrix<T, R, C>>(*this); }

// ---- sparse ----
template <class T, class I = int> class Triplet {
  I r_, c_;
  T v_;

 public:
  Triplet() : r_(0), c_(0), v_(0) {}
  Triplet(I r, I c, T v = T(0)) : r_(r), c_(c), v_(v) {}
  I row() const { return r_; }
  I col() const { return c_; }
  T value() const { return v_; }
};

// Column-major sparse matrix.  Columns are kept as sorted (row, value) lists at all times; compressed arrays are
// produced on demand (outerIndexPtr / innerIndexPtr / valuePtr after makeCompressed(), as in Eigen).
template <class T> class SparseMatrix {
 public:
  using Scalar = T;
  using Index = int;
  using StorageIndex = int;

 private:
  Index rows_ = 0, cols_ = 0;
  std::vector<std::vector<std::pair<Index, T>>> col_;
  mutable std::vector<Index> outer_, inner_;
  mutable std::vector<T> val_;
  void flatten() const {
    outer_.assign(size_t(cols_) + 1, 0);
    inner_.clear();
    val_.clear();
    for (Index j = 0; j < cols_; j++) {
      for (const auto& e : col_[size_t(j)]) { inner_.push_back(e.first); val_.push_back(e.second); }
      outer_[size_t(j) + 1] = Index(inner_.size());
    }
  }

 public:
  SparseMatrix() {}
  SparseMatrix(Index r, Index c) { resize(r, c); }
  void resize(Index r, Index c) { rows_ = r; cols_ = c; col_.assign(size_t(c), {}); }
  Index rows() const { return rows_; }
  Index cols() const { return cols_; }
  Index outerSize() const { return cols_; }
  Index nonZeros() const { size_t n = 0; for (const auto& c : col_) n += c.size(); return Index(n); }
  void setZero() { for (auto& c : col_) c.clear(); }
  void reserve(Index) {}
  template <class It> void setFromTriplets(It b, It e) {
    setZero();
    for (It it = b; it != e; ++it) {
      auto& c = col_[size_t(it->col())];
      auto p = std::lower_bound(c.begin(), c.end(), it->row(), [](const std::pair<Index, T>& a, Index r) { return a.first < r; });
      if (p != c.end() && p->first == it->row()) p->second += it->value();  // duplicates summed in triplet order
      else c.insert(p, std::make_pair(Index(it->row()), T(it->value())));     // zeros are stored too
    }
  }
  T coeff(Index r, Index c) const {
    const auto& cc = col_[size_t(c)];
    auto p = std::lower_bound(cc.begin(), cc.end(), r, [](const std::pair<Index, T>& a, Index rr) { return a.first < rr; });
    return (p != cc.end() && p->first == r) ? p->second : T(0);
  }
  T& insert(Index r, Index c) {
    auto& cc = col_[size_t(c)];
    auto p = std::lower_bound(cc.begin(), cc.end(), r, [](const std::pair<Index, T>& a, Index rr) { return a.first < rr; });
    assert(!(p != cc.end() && p->first == r) && "SparseMatrix::insert: the entry already exists");
    return cc.insert(p, std::make_pair(r, T(0)))->second;
  }
  T& coeffRef(Index r, Index c) {
    auto& cc = col_[size_t(c)];
    auto p = std::lower_bound(cc.begin(), cc.end(), r, [](const std::pair<Index, T>& a, Index rr) { return a.first < rr; });
    if (p != cc.end() && p->first == r) return p->second;
    return cc.insert(p, std::make_pair(r, T(0)))->second;
  }
  void makeCompressed() { flatten(); }
  bool isCompressed() const { return true; }
  const Index* outerIndexPtr() const { flatten(); return outer_.data(); }
  const Index* innerIndexPtr() const { flatten(); return inner_.data(); }
  const T* valuePtr() const { flatten(); return val_.data(); }
  const std::vector<std::pair<Index, T>>& column(Index j) const { return col_[size_t(j)]; }
  // y = A x, column by column
  template <class B> Matrix<T, Dynamic, 1> operator*(const MatrixBase<B>& x) const {
    assert(x.rows() == cols_ && x.cols() == 1);
    Matrix<T, Dynamic, 1> y(rows_);
    y.setZero();
    for (Index j = 0; j < cols_; j++) {
      const T xj = x.coeff(j, 0);
      for (const auto& e : col_[size_t(j)]) y[e.first] += e.second * xj;
    }
    return y;
  }
  struct TransposeView {
    const SparseMatrix& m;
    template <class B> Matrix<T, Dynamic, 1> operator*(const MatrixBase<B>& x) const {
      Matrix<T, Dynamic, 1> y(m.cols_);
      for (Index j = 0; j < m.cols_; j++) {
        T s = 0;
        for (const auto& e : m.col_[size_t(j)]) s += e.second * x.coeff(e.first, 0);
        y[j] = s;
      }
      return y;
    }
  };
  TransposeView transpose() const { return TransposeView{*this}; }
};
template <class T, int R, int C> Matrix<T, R, C>::Matrix(const SparseMatrix<T>& sp) {
  s_.resize(sp.rows(), sp.cols());
  setZero();
  for (int j = 0; j < sp.cols(); j++)
    for (const auto& e : sp.column(j)) (*this)(e.first, j) = e.second;
}

// Eigen::SimplicialLLT stand-in: the reference hands it the lower-stored, structurally dense K_ff
// (assemble.cpp:264-270), so a dense factorisation of that lower triangle is the same mathematics; no fill-reducing
// ordering is applied (AMD only permutes the elimination order).
enum { Lower = 1, Upper = 2 };
template <class SpMat, int UpLo = Lower> class SimplicialLLT {
  using T = typename SpMat::Scalar;
  LLT<Matrix<T, Dynamic, Dynamic>> llt_;

 public:
  void analyzePattern(const SpMat&) {}
  void factorize(const SpMat& a) { llt_.compute(Matrix<T, Dynamic, Dynamic>(a)); }
  void compute(const SpMat& a) { factorize(a); }
  ComputationInfo info() const { return llt_.info(); }
  template <class B> Matrix<T, Dynamic, traits<B>::Cols> solve(const MatrixBase<B>& b) const { return llt_.solve(b); }
};

template <class T> struct aligned_allocator : std::allocator<T> {
  template <class U> struct rebind { using other = aligned_allocator<U>; };
};

}  // namespace Eigen
