// Stand-in for <mrpt/math/TPoint3D.h>: the arithmetic libopenbeam's hot path uses (norm, sqrNorm, operators)
#pragma once
#include <cmath>
namespace mrpt::math {
template <class T> struct TPoint3D_ {
  T x = 0, y = 0, z = 0;
  constexpr TPoint3D_() = default;
  constexpr TPoint3D_(T xx, T yy, T zz) : x(xx), y(yy), z(zz) {}
  T& operator[](size_t i) { return i == 0 ? x : i == 1 ? y : z; }
  const T& operator[](size_t i) const { return i == 0 ? x : i == 1 ? y : z; }
  T sqrNorm() const { return x * x + y * y + z * z; }
  T norm() const { return std::sqrt(sqrNorm()); }
  T distanceTo(const TPoint3D_& o) const { return (*this - o).norm(); }
  TPoint3D_ unitarize() const { const T n = norm(); return TPoint3D_(x / n, y / n, z / n); }
  TPoint3D_ operator-(const TPoint3D_& o) const { return TPoint3D_(x - o.x, y - o.y, z - o.z); }
  TPoint3D_ operator+(const TPoint3D_& o) const { return TPoint3D_(x + o.x, y + o.y, z + o.z); }
  TPoint3D_ operator*(T f) const { return TPoint3D_(x * f, y * f, z * f); }
  TPoint3D_ operator/(T f) const { return TPoint3D_(x / f, y / f, z / f); }
  TPoint3D_ operator-() const { return TPoint3D_(-x, -y, -z); }
  TPoint3D_& operator+=(const TPoint3D_& o) { x += o.x; y += o.y; z += o.z; return *this; }
  TPoint3D_& operator-=(const TPoint3D_& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  TPoint3D_& operator*=(T f) { x *= f; y *= f; z *= f; return *this; }
};
template <class T> inline TPoint3D_<T> operator*(T f, const TPoint3D_<T>& p) { return p * f; }
using TPoint3D = TPoint3D_<double>;
using TVector3D = TPoint3D_<double>;
using TPoint3Df = TPoint3D_<float>;
template <class T> inline T dotProduct3(const TPoint3D_<T>& a, const TPoint3D_<T>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
template <class T> inline TPoint3D_<T> crossProduct3D(const TPoint3D_<T>& a, const TPoint3D_<T>& b) {
  return TPoint3D_<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
}  // namespace mrpt::math
