// CPU-only check of the host-side pose preparation the kernels rely on
// (hectorgrapher_b200/csrc/host/pose_prep.h):
//   * TwoPose reproduces transform/timestamped_transform_test.cc:37-80 (InterpolateTransform:
//     identity -> translation (10,10,10), 2 rad about z; the interpolant at factor f is
//     Translation(f t) * Rotation(2 f about z), tolerance 1e-6);
//   * its Jacobians d q / d q0, d q / d q1 (what the reference's Jet<double,14> carries through
//     Eigen's slerp) agree with central differences, in the acos/sin branch, the lerp branch
//     (|q0.q1| >= 1 - epsilon) and for q0.q1 < 0;
//   * the same Jacobians in the rank-one form the Jacobian kernel multiplies with
//     (m0 = s0 I + c q1', m1 = s1 I + c q0').
// Prints "POSE PREP OK" and exits 0 on success.
#include <cmath>
#include <cstdio>
#include <cstring>

#include "../../hectorgrapher_b200/csrc/host/pose_prep.h"

static int Fail(const char* what) {
  std::fprintf(stderr, "FAILED: %s\n", what);
  return 1;
}

static bool JacobiansMatch(const double* q0, const double* q1, double f) {
  const double t0[3] = {0, 0, 0}, t1[3] = {1, 2, 3};
  const hg::BlockPose p = hg::TwoPose(t0, q0, t1, q1, f);
  const double eps = 1e-6;
  for (int which = 0; which < 2; ++which) {
    for (int j = 0; j < 4; ++j) {
      double a0[4], a1[4], b0[4], b1[4];
      std::memcpy(a0, q0, sizeof(a0)); std::memcpy(a1, q1, sizeof(a1));
      std::memcpy(b0, q0, sizeof(b0)); std::memcpy(b1, q1, sizeof(b1));
      (which == 0 ? a0 : a1)[j] += eps;
      (which == 0 ? b0 : b1)[j] -= eps;
      const hg::BlockPose hi = hg::TwoPose(t0, a0, t1, a1, f), lo = hg::TwoPose(t0, b0, t1, b1, f);
      for (int i = 0; i < 4; ++i) {
        const double numeric = (hi.q[i] - lo.q[i]) / (2 * eps);
        const double analytic = (which == 0 ? p.m0 : p.m1)[i * 4 + j];
        if (std::fabs(numeric - analytic) > 1e-6 * (1.0 + std::fabs(analytic))) return false;
      }
    }
  }
  return std::fabs(p.ct0 - (1.0 - f)) < 1e-15 && std::fabs(p.ct1 - f) < 1e-15;
}

// The form the Jacobian kernel uses: m0 = s0 I + c q1', m1 = s1 I + c q0' (BlockPose::s0, s1, c, q0, q1).
static bool RankOneFormMatches(const double* q0, const double* q1, double f) {
  const double t0[3] = {0, 0, 0}, t1[3] = {1, 2, 3};
  const hg::BlockPose p = hg::TwoPose(t0, q0, t1, q1, f);
  for (int i = 0; i < 4; ++i) {
    if (p.q0[i] != q0[i] || p.q1[i] != q1[i]) return false;
    for (int j = 0; j < 4; ++j) {
      const double m0 = (i == j ? p.s0 : 0.0) + p.c[i] * q1[j], m1 = (i == j ? p.s1 : 0.0) + p.c[i] * q0[j];
      if (std::fabs(m0 - p.m0[i * 4 + j]) > 1e-14 || std::fabs(m1 - p.m1[i * 4 + j]) > 1e-14) return false;
    }
  }
  // and a row vector through both forms, as the kernel computes it
  const double a[4] = {0.3, -1.7, 0.6, 2.2};
  const double ac = a[0] * p.c[0] + a[1] * p.c[1] + a[2] * p.c[2] + a[3] * p.c[3];
  for (int j = 0; j < 4; ++j) {
    const double full0 = a[0] * p.m0[j] + a[1] * p.m0[4 + j] + a[2] * p.m0[8 + j] + a[3] * p.m0[12 + j];
    const double full1 = a[0] * p.m1[j] + a[1] * p.m1[4 + j] + a[2] * p.m1[8 + j] + a[3] * p.m1[12 + j];
    if (std::fabs(p.s0 * a[j] + ac * q1[j] - full0) > 1e-13) return false;
    if (std::fabs(p.s1 * a[j] + ac * q0[j] - full1) > 1e-13) return false;
  }
  return true;
}

int main() {
  const double t0[3] = {0, 0, 0}, q0[4] = {1, 0, 0, 0};
  const double t1[3] = {10, 10, 10}, q1[4] = {std::cos(1.0), 0, 0, std::sin(1.0)};  // 2 rad about z
  for (int k = 0; k <= 20; ++k) {
    const double f = k / 20.0;
    const hg::BlockPose p = hg::TwoPose(t0, q0, t1, q1, f);
    const double want_q[4] = {std::cos(f), 0, 0, std::sin(f)};
    for (int i = 0; i < 3; ++i)
      if (std::fabs(p.t[i] - f * 10.0) > 1e-6) return Fail("interpolated translation");
    for (int i = 0; i < 4; ++i)
      if (std::fabs(p.q[i] - want_q[i]) > 1e-6) return Fail("interpolated rotation");
  }
  const double a[4] = {0.9, 0.1, -0.3, 0.2}, b[4] = {0.7, -0.2, 0.4, 0.5};
  if (!JacobiansMatch(a, b, 0.37)) return Fail("slerp Jacobians (acos/sin branch)");
  const double nb[4] = {-0.7, 0.2, -0.4, -0.5};
  if (!JacobiansMatch(a, nb, 0.81)) return Fail("slerp Jacobians (negative dot)");
  if (!RankOneFormMatches(a, b, 0.37)) return Fail("rank-one form (acos/sin branch)");
  if (!RankOneFormMatches(a, nb, 0.81)) return Fail("rank-one form (negative dot)");
  const double zy[4] = {0.0, 0.6, 0.0, 0.8}, zw[4] = {0.0, 0.0, 1.0, 0.0};  // zero components: s'(d) is read off the largest one
  if (!RankOneFormMatches(zy, zw, 0.5)) return Fail("rank-one form (zero components)");
  // Identical unit quaternions: |d| = 1 >= 1 - epsilon -> lerp branch, value and derivative.
  const double u[4] = {0.5, 0.5, 0.5, 0.5};
  const hg::BlockPose same = hg::TwoPose(t0, u, t1, u, 0.25);
  for (int i = 0; i < 4; ++i) {
    if (std::fabs(same.q[i] - 0.5) > 1e-15) return Fail("lerp branch value");
    for (int j = 0; j < 4; ++j) {
      if (std::fabs(same.m0[i * 4 + j] - (i == j ? 0.75 : 0.0)) > 1e-15) return Fail("lerp branch d/dq0");
      if (std::fabs(same.m1[i * 4 + j] - (i == j ? 0.25 : 0.0)) > 1e-15) return Fail("lerp branch d/dq1");
    }
  }
  if (!RankOneFormMatches(u, u, 0.25)) return Fail("rank-one form (lerp branch)");
  if (same.s0 != 0.75 || same.s1 != 0.25 || same.c[0] != 0.0 || same.c[3] != 0.0) return Fail("lerp branch factors");
  const double t[3] = {1, 2, 3}, q[4] = {0.1, 0.2, 0.3, 0.4};
  const hg::BlockPose one = hg::OnePose(t, q);
  if (one.q[2] != 0.3 || one.t[1] != 2 || one.ct0 != 1.0 || one.ct1 != 0.0 || one.m0[5] != 1.0 || one.m0[1] != 0.0)
    return Fail("OnePose");
  if (one.s0 != 1.0 || one.s1 != 0.0 || one.c[1] != 0.0 || one.q0[3] != 0.4) return Fail("OnePose factors");
  std::printf("POSE PREP OK\n");
  return 0;
}
