// Single-precision pose arithmetic of the correlative matcher, host side.
//
// The reference does this through Eigen (not vendored in the reference tree):
//   transform/rigid_transform.h:161-165, 183-196   inverse, product (renormalised), point transform
//   transform/transform.h:78-83, 121-135           GetYaw, AngleAxisVectorToRotationQuaternion
//   mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.cc:276-326, 400-406
// Formulas are Eigen 3.3's (Geometry/Quaternion.h, arch/Geometry_SSE.h). Where
// Eigen's SSE2 float code sums in a different order than its generic code, both
// orders are provided; 'flavor' 1 selects the SSE2 order (what an x86-64 build
// of the reference runs), 0 the generic order. Compiled with -ffp-contract=off.
#pragma once

#include <cmath>

namespace hg {
namespace fgeom {

struct V3 {
  float x, y, z;
};
struct Q4 {
  float w, x, y, z;
};
struct Rigid {
  V3 t;
  Q4 q;
};

inline V3 Cross(const V3& a, const V3& b) {
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// Eigen _transformVector.
inline V3 Rotate(const Q4& q, const V3& v) {
  const V3 u{q.x, q.y, q.z};
  V3 uv = Cross(u, v);
  uv = V3{uv.x + uv.x, uv.y + uv.y, uv.z + uv.z};
  const V3 c = Cross(u, uv);
  return V3{v.x + q.w * uv.x + c.x, v.y + q.w * uv.y + c.y, v.z + q.w * uv.z + c.z};
}
inline V3 Apply(const Rigid& r, const V3& p) {
  const V3 v = Rotate(r.q, p);
  return V3{v.x + r.t.x, v.y + r.t.y, v.z + r.t.z};
}
inline Q4 Product(const Q4& a, const Q4& b, const int flavor) {
  if (flavor == 1) {
    const float x = (a.x * b.w - a.z * b.y) + (a.y * b.z + a.w * b.x);
    const float y = (a.y * b.w - a.x * b.z) + (a.z * b.x + a.w * b.y);
    const float z = (a.z * b.w - a.y * b.x) + (a.x * b.y + a.w * b.z);
    const float w = (a.w * b.w - a.x * b.x) + -(a.z * b.z + a.y * b.y);
    return Q4{w, x, y, z};
  }
  return Q4{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
            a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
}
inline float SquaredNorm(const Q4& q, const int flavor) {
  if (flavor == 1) return (q.x * q.x + q.z * q.z) + (q.y * q.y + q.w * q.w);
  return (q.x * q.x + q.y * q.y) + (q.z * q.z + q.w * q.w);
}
inline Q4 Normalized(const Q4& q, const int flavor) {
  const float n = std::sqrt(SquaredNorm(q, flavor));
  return Q4{q.w / n, q.x / n, q.y / n, q.z / n};
}
inline Q4 Conjugate(const Q4& q) { return Q4{q.w, -q.x, -q.y, -q.z}; }
inline Q4 Inverse(const Q4& q, const int flavor) {
  const float n2 = SquaredNorm(q, flavor);
  if (n2 > 0.f) return Q4{q.w / n2, -q.x / n2, -q.y / n2, -q.z / n2};
  return Q4{0.f, 0.f, 0.f, 0.f};
}
inline float SquaredNorm3(const V3& v) { return v.x * v.x + (v.y * v.y + v.z * v.z); }
inline float Norm3(const V3& v) { return std::sqrt(SquaredNorm3(v)); }

inline Rigid RigidInverse(const Rigid& r) {
  const Q4 rot = Conjugate(r.q);
  const V3 t = Rotate(rot, r.t);
  return Rigid{V3{-t.x, -t.y, -t.z}, rot};
}
inline Rigid RigidProduct(const Rigid& a, const Rigid& b, const int flavor) {
  const V3 v = Rotate(a.q, b.t);
  return Rigid{V3{v.x + a.t.x, v.y + a.t.y, v.z + a.t.z}, Normalized(Product(a.q, b.q, flavor), flavor)};
}
inline Q4 AngleAxisVectorToRotationQuaternion(const V3& angle_axis) {
  float scale = 0.5f;
  float w = 1.f;
  constexpr double kCutoffAngle = 1e-8;
  if (static_cast<double>(SquaredNorm3(angle_axis)) > kCutoffAngle) {
    const float norm = Norm3(angle_axis);
    scale = static_cast<float>(std::sin(norm / 2.) / norm);
    w = static_cast<float>(std::cos(norm / 2.));
  }
  return Q4{w, scale * angle_axis.x, scale * angle_axis.y, scale * angle_axis.z};
}
inline float GetYaw(const Q4& q) {
  const V3 d = Rotate(q, V3{1.f, 0.f, 0.f});
  return std::atan2(d.y, d.x);
}

}  // namespace fgeom
}  // namespace hg
