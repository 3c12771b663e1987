// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// Rigid-transform arithmetic as the reference uses it through Eigen. Eigen is
// NOT vendored under /root/reference (pinned by bazel/repositories.bzl:149-157,
// commit f3a22f35b044, or the distro eigen3, package.xml:45); the formulas
// below restate Eigen 3.3's published Geometry module
// (Eigen/src/Geometry/Quaternion.h, arch/Geometry_SSE.h):
//   * QuaternionBase::_transformVector:  uv = 2 (u x v);  v + w uv + u x uv
//   * quaternion product (generic scalar formula, and the SSE float
//     specialisation's summation order — see QuatProductFloat)
//   * QuaternionBase::slerp (lerp branch when |d| >= 1 - epsilon)
//   * inverse() = conjugate() / squaredNorm();  normalized() = coeffs / norm()
// Reference call sites:
//   transform/rigid_transform.h:161-196  (inverse, operator*, point transform)
//   transform/timestamped_transform.h:39-48  (InterpolateTransform)
//   transform/transform.h:78-83, 121-135  (GetYaw, AngleAxisVectorToRotationQuaternion)
//
// PARITY NOTE: the float summation orders of the SSE code paths cannot be
// checked against a real Eigen in this environment (see DESIGN.md "unpinned").
#pragma once

#include <cmath>
#include <limits>

#include "jet.hpp"

namespace oracle {

template <typename T>
struct Vec3 {
  T x, y, z;
  Vec3() : x(T(0.0)), y(T(0.0)), z(T(0.0)) {}
  Vec3(const T& x_, const T& y_, const T& z_) : x(x_), y(y_), z(z_) {}
};
template <typename T>
inline Vec3<T> operator+(const Vec3<T>& a, const Vec3<T>& b) {
  return Vec3<T>(a.x + b.x, a.y + b.y, a.z + b.z);
}
template <typename T>
inline Vec3<T> operator-(const Vec3<T>& a, const Vec3<T>& b) {
  return Vec3<T>(a.x - b.x, a.y - b.y, a.z - b.z);
}
template <typename T>
inline Vec3<T> Cross(const Vec3<T>& a, const Vec3<T>& b) {
  return Vec3<T>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

template <typename T>
struct Quat {
  T w, x, y, z;
  Quat() : w(T(1.0)), x(T(0.0)), y(T(0.0)), z(T(0.0)) {}
  Quat(const T& w_, const T& x_, const T& y_, const T& z_) : w(w_), x(x_), y(y_), z(z_) {}
};

// Eigen _transformVector (no normalisation of q).
template <typename T>
inline Vec3<T> Rotate(const Quat<T>& q, const Vec3<T>& v) {
  const Vec3<T> u(q.x, q.y, q.z);
  Vec3<T> uv = Cross(u, v);
  uv = uv + uv;
  const Vec3<T> c = Cross(u, uv);
  return Vec3<T>(v.x + q.w * uv.x + c.x, v.y + q.w * uv.y + c.y, v.z + q.w * uv.z + c.z);
}

template <typename T>
struct Rigid3 {
  Vec3<T> translation;
  Quat<T> rotation;
  Rigid3() {}
  Rigid3(const Vec3<T>& t, const Quat<T>& q) : translation(t), rotation(q) {}
};

// rigid_transform.h:191-196
template <typename T>
inline Vec3<T> Apply(const Rigid3<T>& rigid, const Vec3<T>& p) {
  return Rotate(rigid.rotation, p) + rigid.translation;
}

// Eigen slerp. NumTraits<Jet>::epsilon() is numeric_limits<double>::epsilon().
template <typename T>
inline Quat<T> Slerp(const Quat<T>& q0, const T& t, const Quat<T>& q1) {
  const T one = T(1.0) - T(std::numeric_limits<double>::epsilon());
  // coeffs are stored (x, y, z, w); 4-term dot is summed pairwise.
  const T d = (q0.x * q1.x + q0.y * q1.y) + (q0.z * q1.z + q0.w * q1.w);
  const T abs_d = abs(d);
  T scale0, scale1;
  if (abs_d >= one) {
    scale0 = T(1.0) - t;
    scale1 = t;
  } else {
    const T theta = acos(abs_d);
    const T sin_theta = sin(theta);
    scale0 = sin((T(1.0) - t) * theta) / sin_theta;
    scale1 = sin(t * theta) / sin_theta;
  }
  if (d < T(0.0)) scale1 = -scale1;
  return Quat<T>(scale0 * q0.w + scale1 * q1.w, scale0 * q0.x + scale1 * q1.x,
                 scale0 * q0.y + scale1 * q1.y, scale0 * q0.z + scale1 * q1.z);
}

// timestamped_transform.h:39-48
template <typename T>
inline Rigid3<T> InterpolateTransform(const Rigid3<T>& start, const Rigid3<T>& end,
                                      const double factor) {
  const Vec3<T> origin(start.translation.x + (end.translation.x - start.translation.x) * factor,
                       start.translation.y + (end.translation.y - start.translation.y) * factor,
                       start.translation.z + (end.translation.z - start.translation.z) * factor);
  return Rigid3<T>(origin, Slerp(start.rotation, T(factor), end.rotation));
}

// ---------------------------------------------------------------------------
// Single-precision path (FastCorrelativeScanMatcher3D / LowResolutionMatcher).

using Vec3f = Vec3<float>;
using Quatf = Quat<float>;
using Rigid3f = Rigid3<float>;

// Flavor 1 (default) = Eigen's SSE2 float specialisations, which an x86-64
// build of the reference uses; flavor 0 = Eigen's generic scalar code
// (EIGEN_DONT_VECTORIZE / non-x86). They differ only in summation order.
struct FloatFlavor {
  static int& Get() {
    static int flavor = 1;
    return flavor;
  }
};

inline Quatf QuatProductFloat(const Quatf& a, const Quatf& b) {
  if (FloatFlavor::Get() == 1) {
    // arch/Geometry_SSE.h quat_product<SSE, float>:
    //   (a * b.wwww - a.zxyx * b.yzxx) + (+,+,+,-)(a.yzxz * b.zxyz + a.wwwy * b.xyzy)
    const float x = (a.x * b.w - a.z * b.y) + (a.y * b.z + a.w * b.x);
    const float y = (a.y * b.w - a.x * b.z) + (a.z * b.x + a.w * b.y);
    const float z = (a.z * b.w - a.y * b.x) + (a.x * b.y + a.w * b.z);
    const float w = (a.w * b.w - a.x * b.x) + -(a.z * b.z + a.y * b.y);
    return Quatf(w, x, y, z);
  }
  return Quatf(a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z,
               a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
               a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
               a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x);
}

inline float QuatSquaredNormFloat(const Quatf& q) {
  // coeffs (x, y, z, w). SSE2 predux: (c0 + c2) + (c1 + c3); scalar: (c0 + c1) + (c2 + c3).
  if (FloatFlavor::Get() == 1) {
    return (q.x * q.x + q.z * q.z) + (q.y * q.y + q.w * q.w);
  }
  return (q.x * q.x + q.y * q.y) + (q.z * q.z + q.w * q.w);
}

inline Quatf QuatNormalizedFloat(const Quatf& q) {
  const float n = std::sqrt(QuatSquaredNormFloat(q));
  return Quatf(q.w / n, q.x / n, q.y / n, q.z / n);
}

inline Quatf QuatConjugateFloat(const Quatf& q) { return Quatf(q.w, -q.x, -q.y, -q.z); }

// Eigen Quaternion::inverse(): conjugate / squaredNorm (a true division).
inline Quatf QuatInverseFloat(const Quatf& q) {
  const float n2 = QuatSquaredNormFloat(q);
  if (n2 > 0.f) {
    return Quatf(q.w / n2, -q.x / n2, -q.y / n2, -q.z / n2);
  }
  return Quatf(0.f, 0.f, 0.f, 0.f);
}

// 3-vector norm: Eigen's non-vectorised unrolled reduction sums c0 + (c1 + c2).
inline float Norm3f(const Vec3f& v) { return std::sqrt(v.x * v.x + (v.y * v.y + v.z * v.z)); }
inline float SquaredNorm3f(const Vec3f& v) { return v.x * v.x + (v.y * v.y + v.z * v.z); }

// rigid_transform.h:161-165  Rigid3::inverse(): rotation.conjugate(), -(rot * t)
inline Rigid3f RigidInverseFloat(const Rigid3f& r) {
  const Quatf rot = QuatConjugateFloat(r.rotation);
  const Vec3f t = Rotate(rot, r.translation);
  return Rigid3f(Vec3f(-t.x, -t.y, -t.z), rot);
}

// rigid_transform.h:183-189  (product renormalises the quaternion)
inline Rigid3f RigidProductFloat(const Rigid3f& lhs, const Rigid3f& rhs) {
  return Rigid3f(Rotate(lhs.rotation, rhs.translation) + lhs.translation,
                 QuatNormalizedFloat(QuatProductFloat(lhs.rotation, rhs.rotation)));
}

// transform/transform.h:121-135 for T = float. Mixed float/double steps as in
// the template: squaredNorm (float) > 1e-8 (double); norm / 2. is a double
// division, sin/cos evaluated in double, narrowed on assignment.
inline Quatf AngleAxisVectorToRotationQuaternionFloat(const Vec3f& angle_axis) {
  float scale = 0.5f;
  float w = 1.f;
  constexpr double kCutoffAngle = 1e-8;
  if (static_cast<double>(SquaredNorm3f(angle_axis)) > kCutoffAngle) {
    const float norm = Norm3f(angle_axis);
    scale = static_cast<float>(std::sin(norm / 2.) / norm);
    w = static_cast<float>(std::cos(norm / 2.));
  }
  return Quatf(w, scale * angle_axis.x, scale * angle_axis.y, scale * angle_axis.z);
}

// transform/transform.h:78-83 for T = float.
inline float GetYawFloat(const Quatf& rotation) {
  const Vec3f direction = Rotate(rotation, Vec3f(1.f, 0.f, 0.f));
  return std::atan2(direction.y, direction.x);
}

}  // namespace oracle
