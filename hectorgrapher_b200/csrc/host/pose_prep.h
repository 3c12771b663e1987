// Host-side preparation of the effective pose of a residual block and of the
// chain-rule factors the device kernels need.
//
// One-pose functors (tsdf_space_cost_function_3d.h:47-55 and siblings) use the
// parameters directly. Two-pose functors evaluate
// InterpolateTransform(T0, T1, f) (transform/timestamped_transform.h:39-48):
//   t = t0 + (t1 - t0) f,   q = q0.slerp(f, q1)   (Eigen, not renormalised)
// on Ceres Jets. The pose depends only on the 14 parameters, not on the point,
// so its value and its 4x4 Jacobians dq/dq0, dq/dq1 are computed here once per
// block, in double, with forward-mode duals over the 8 quaternion parameters —
// the same derivative the reference's Jet<double,14> carries through slerp
// (acos / sin branch, or the lerp branch when |q0.q1| >= 1 - epsilon).
#pragma once

#include <cmath>
#include <limits>

namespace hg {

struct Dual8 {
  double a;
  double v[8];
};
inline Dual8 D8(double a) {
  Dual8 r;
  r.a = a;
  for (int i = 0; i < 8; ++i) r.v[i] = 0.0;
  return r;
}
inline Dual8 D8(double a, int k) {
  Dual8 r = D8(a);
  r.v[k] = 1.0;
  return r;
}
inline Dual8 operator+(const Dual8& f, const Dual8& g) {
  Dual8 r;
  r.a = f.a + g.a;
  for (int i = 0; i < 8; ++i) r.v[i] = f.v[i] + g.v[i];
  return r;
}
inline Dual8 operator-(const Dual8& f, const Dual8& g) {
  Dual8 r;
  r.a = f.a - g.a;
  for (int i = 0; i < 8; ++i) r.v[i] = f.v[i] - g.v[i];
  return r;
}
inline Dual8 operator-(const Dual8& f) {
  Dual8 r;
  r.a = -f.a;
  for (int i = 0; i < 8; ++i) r.v[i] = -f.v[i];
  return r;
}
inline Dual8 operator*(const Dual8& f, const Dual8& g) {
  Dual8 r;
  r.a = f.a * g.a;
  for (int i = 0; i < 8; ++i) r.v[i] = f.a * g.v[i] + f.v[i] * g.a;
  return r;
}
inline Dual8 operator/(const Dual8& f, const Dual8& g) {
  const double gi = 1.0 / g.a;
  const double fg = f.a * gi;
  Dual8 r;
  r.a = fg;
  for (int i = 0; i < 8; ++i) r.v[i] = (f.v[i] - fg * g.v[i]) * gi;
  return r;
}
inline Dual8 Sin(const Dual8& f) {
  Dual8 r;
  r.a = std::sin(f.a);
  const double c = std::cos(f.a);
  for (int i = 0; i < 8; ++i) r.v[i] = c * f.v[i];
  return r;
}
inline Dual8 Acos(const Dual8& f) {
  Dual8 r;
  r.a = std::acos(f.a);
  const double t = -1.0 / std::sqrt(1.0 - f.a * f.a);
  for (int i = 0; i < 8; ++i) r.v[i] = t * f.v[i];
  return r;
}

// Everything a kernel needs to know about one residual block's pose.
struct BlockPose {
  double t[3];
  double q[4];       // (w, x, y, z)
  double ct0, ct1;   // d t / d t0 = ct0 I,  d t / d t1 = ct1 I
  double m0[16];     // d q_i / d q0_j, row-major [i][j]
  double m1[16];     // d q_i / d q1_j
  // The same two matrices in the form slerp gives them: q = s0 q0 + s1 q1 with s0, s1 functions
  // of d = q0 . q1 only, so m0 = s0 I + c q1', m1 = s1 I + c q0' with c = s0'(d) q0 + s1'(d) q1.
  // A row vector a times them is s0 a + (a . c) q1' and s1 a + (a . c) q0': 20 multiply-adds per
  // point in the Jacobian kernel instead of 32.
  double s0, s1;
  double c[4];
  double q0[4], q1[4];
};

inline BlockPose OnePose(const double* t, const double* q) {
  BlockPose p;
  for (int i = 0; i < 3; ++i) p.t[i] = t[i];
  for (int i = 0; i < 4; ++i) p.q[i] = q[i];
  p.ct0 = 1.0;
  p.ct1 = 0.0;
  for (int i = 0; i < 16; ++i) {
    p.m0[i] = (i % 5 == 0) ? 1.0 : 0.0;
    p.m1[i] = 0.0;
  }
  p.s0 = 1.0;
  p.s1 = 0.0;
  for (int i = 0; i < 4; ++i) {
    p.c[i] = 0.0;
    p.q0[i] = q[i];
    p.q1[i] = 0.0;
  }
  return p;
}

inline BlockPose TwoPose(const double* t0, const double* q0, const double* t1, const double* q1,
                         const double f) {
  BlockPose p;
  for (int i = 0; i < 3; ++i) p.t[i] = t0[i] + (t1[i] - t0[i]) * f;
  p.ct0 = 1.0 + (-1.0) * f;
  p.ct1 = f;
  // Duals: q0 -> directions 0..3, q1 -> directions 4..7, order (w, x, y, z).
  Dual8 a[4], b[4];
  for (int i = 0; i < 4; ++i) {
    a[i] = D8(q0[i], i);
    b[i] = D8(q1[i], 4 + i);
  }
  const Dual8 one = D8(1.0 - std::numeric_limits<double>::epsilon());
  const Dual8 t = D8(f);
  // Eigen dot over coeffs (x, y, z, w), summed pairwise.
  const Dual8 d = (a[1] * b[1] + a[2] * b[2]) + (a[3] * b[3] + a[0] * b[0]);
  const Dual8 abs_d = d.a < 0.0 ? -d : d;
  Dual8 s0, s1;
  if (abs_d.a >= one.a) {
    s0 = D8(1.0) - t;
    s1 = t;
  } else {
    const Dual8 theta = Acos(abs_d);
    const Dual8 sin_theta = Sin(theta);
    s0 = Sin((D8(1.0) - t) * theta) / sin_theta;
    s1 = Sin(t * theta) / sin_theta;
  }
  if (d.a < 0.0) s1 = -s1;
  for (int i = 0; i < 4; ++i) {
    const Dual8 qi = s0 * a[i] + s1 * b[i];
    p.q[i] = qi.a;
    for (int j = 0; j < 4; ++j) {
      p.m0[i * 4 + j] = qi.v[j];
      p.m1[i * 4 + j] = qi.v[4 + j];
    }
  }
  // d s / d q0_j = s'(d) q1_j: read s'(d) off the largest component of q1 (a quaternion is never
  // all zero; in the lerp branch the derivatives are zero).
  int js = 0;
  for (int j = 1; j < 4; ++j) {
    if (std::abs(q1[j]) > std::abs(q1[js])) js = j;
  }
  const double ds0 = q1[js] != 0.0 ? s0.v[js] / q1[js] : 0.0;
  const double ds1 = q1[js] != 0.0 ? s1.v[js] / q1[js] : 0.0;
  p.s0 = s0.a;
  p.s1 = s1.a;
  for (int i = 0; i < 4; ++i) {
    p.c[i] = ds0 * q0[i] + ds1 * q1[i];
    p.q0[i] = q0[i];
    p.q1[i] = q1[i];
  }
  return p;
}

}  // namespace hg
