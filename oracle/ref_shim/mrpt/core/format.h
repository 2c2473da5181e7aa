// Stand-in for <mrpt/core/format.h>
#pragma once
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>
namespace mrpt {
inline std::string format(const char* fmt, ...) {
  va_list ap, ap2;
  va_start(ap, fmt);
  va_copy(ap2, ap);
  const int n = vsnprintf(nullptr, 0, fmt, ap);
  va_end(ap);
  std::string s(size_t(n > 0 ? n : 0), '\0');
  if (n > 0) vsnprintf(&s[0], size_t(n) + 1, fmt, ap2);
  va_end(ap2);
  return s;
}
template <class T> inline void keep_max(T& v, const T& x) { if (x > v) v = x; }
template <class T> inline void keep_min(T& v, const T& x) { if (x < v) v = x; }
template <class T> inline int round(const T x) { return int(x > 0 ? x + T(0.5) : x - T(0.5)); }
}  // namespace mrpt
