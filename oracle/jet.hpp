// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// Forward-mode dual numbers restating ceres::Jet<double, N>. Ceres is a
// third-party dependency of the reference that is NOT vendored under
// /root/reference (pinned: Ceres 1.13.0, scripts/install_ceres.sh:20;
// bazel/repositories.bzl:137-145). The arithmetic rules below are Ceres'
// published Jet rules (include/ceres/jet.h of 1.13): product rule, quotient
// via multiplication by the reciprocal of the denominator's scalar part,
// sqrt / sin / cos / acos chain rules, comparisons on the scalar part.
// Call sites in the reference: every CreateAutoDiffCostFunction of
// mapping/internal/3d/scan_matching/*_cost_function_3d.h.
#pragma once

#include <cmath>

namespace oracle {

template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) {
    for (int i = 0; i < N; ++i) v[i] = 0.0;
  }
  explicit Jet(const double value) : a(value) {
    for (int i = 0; i < N; ++i) v[i] = 0.0;
  }
  Jet(const double value, const int k) : a(value) {
    for (int i = 0; i < N; ++i) v[i] = 0.0;
    v[k] = 1.0;
  }
};

template <int N>
inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a + g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i];
  return h;
}
template <int N>
inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a - g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i];
  return h;
}
template <int N>
inline Jet<N> operator-(const Jet<N>& f) {
  Jet<N> h;
  h.a = -f.a;
  for (int i = 0; i < N; ++i) h.v[i] = -f.v[i];
  return h;
}
template <int N>
inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h;
  h.a = f.a * g.a;
  for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a;
  return h;
}
template <int N>
inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  const double g_a_inverse = 1.0 / g.a;
  const double f_a_by_g_a = f.a * g_a_inverse;
  Jet<N> h;
  h.a = f_a_by_g_a;
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - f_a_by_g_a * g.v[i]) * g_a_inverse;
  return h;
}
// Jet (op) scalar
template <int N>
inline Jet<N> operator+(const Jet<N>& f, const double s) {
  Jet<N> h = f;
  h.a = f.a + s;
  return h;
}
template <int N>
inline Jet<N> operator+(const double s, const Jet<N>& f) {
  Jet<N> h = f;
  h.a = f.a + s;
  return h;
}
template <int N>
inline Jet<N> operator-(const Jet<N>& f, const double s) {
  Jet<N> h = f;
  h.a = f.a - s;
  return h;
}
template <int N>
inline Jet<N> operator-(const double s, const Jet<N>& f) {
  Jet<N> h;
  h.a = s - f.a;
  for (int i = 0; i < N; ++i) h.v[i] = -f.v[i];
  return h;
}
template <int N>
inline Jet<N> operator*(const Jet<N>& f, const double s) {
  Jet<N> h;
  h.a = f.a * s;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s;
  return h;
}
template <int N>
inline Jet<N> operator*(const double s, const Jet<N>& f) {
  Jet<N> h;
  h.a = f.a * s;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s;
  return h;
}
template <int N>
inline Jet<N> operator/(const Jet<N>& f, const double s) {
  const double s_inverse = 1.0 / s;
  Jet<N> h;
  h.a = f.a * s_inverse;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * s_inverse;
  return h;
}
template <int N>
inline Jet<N> operator/(const double s, const Jet<N>& g) {
  const double minus_s_g_a_inverse2 = -s / (g.a * g.a);
  Jet<N> h;
  h.a = s / g.a;
  for (int i = 0; i < N; ++i) h.v[i] = g.v[i] * minus_s_g_a_inverse2;
  return h;
}
template <int N>
inline Jet<N>& operator+=(Jet<N>& f, const Jet<N>& g) {
  f = f + g;
  return f;
}

// Comparisons act on the scalar part.
template <int N> inline bool operator<(const Jet<N>& f, const Jet<N>& g) { return f.a < g.a; }
template <int N> inline bool operator>(const Jet<N>& f, const Jet<N>& g) { return f.a > g.a; }
template <int N> inline bool operator>=(const Jet<N>& f, const Jet<N>& g) { return f.a >= g.a; }
template <int N> inline bool operator<(const Jet<N>& f, const double s) { return f.a < s; }

template <int N>
inline Jet<N> abs(const Jet<N>& f) {
  return f.a < 0.0 ? -f : f;
}
template <int N>
inline Jet<N> sqrt(const Jet<N>& f) {
  const double tmp = std::sqrt(f.a);
  const double two_a_inverse = 1.0 / (2.0 * tmp);
  Jet<N> h;
  h.a = tmp;
  for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * two_a_inverse;
  return h;
}
template <int N>
inline Jet<N> sin(const Jet<N>& f) {
  const double c = std::cos(f.a);
  Jet<N> h;
  h.a = std::sin(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i];
  return h;
}
template <int N>
inline Jet<N> cos(const Jet<N>& f) {
  const double ms = -std::sin(f.a);
  Jet<N> h;
  h.a = std::cos(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = ms * f.v[i];
  return h;
}
template <int N>
inline Jet<N> acos(const Jet<N>& f) {
  const double tmp = -1.0 / std::sqrt(1.0 - f.a * f.a);
  Jet<N> h;
  h.a = std::acos(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = tmp * f.v[i];
  return h;
}

// Plain-double overloads so templated code can call oracle::sin(T) etc.
inline double abs(const double x) { return std::fabs(x); }
inline double sqrt(const double x) { return std::sqrt(x); }
inline double sin(const double x) { return std::sin(x); }
inline double cos(const double x) { return std::cos(x); }
inline double acos(const double x) { return std::acos(x); }

// Scalar part of a value (used for interval selection).
inline double ScalarPart(const double x) { return x; }
template <int N>
inline double ScalarPart(const Jet<N>& x) { return x.a; }

}  // namespace oracle
