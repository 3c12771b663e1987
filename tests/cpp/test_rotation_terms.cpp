// CPU-only check of the closed form behind RotationTerms (hectorgrapher_b200/csrc/interpolation.cuh):
// with world = p + w uv + u x uv + t, uv = 2 u x p (Eigen's quaternion * vector on the raw, not
// normalised, quaternion: transform/rigid_transform.h:191-196),
//   g . d world / d w = g . uv
//   g . d world / d u = 2 w (p x g) + 2 [(u . p) g + (g . u) p - 2 (g . p) u]
// against central differences of g . world(q). Prints "ROTATION TERMS OK" and exits 0 on success.
#include <cmath>
#include <cstdio>
#include <random>

static void World(const double* q, const double* p, double* w) {
  const double qw = q[0], qx = q[1], qy = q[2], qz = q[3];
  double uvx = qy * p[2] - qz * p[1], uvy = qz * p[0] - qx * p[2], uvz = qx * p[1] - qy * p[0];
  uvx += uvx; uvy += uvy; uvz += uvz;
  w[0] = p[0] + qw * uvx + (qy * uvz - qz * uvy);
  w[1] = p[1] + qw * uvy + (qz * uvx - qx * uvz);
  w[2] = p[2] + qw * uvz + (qx * uvy - qy * uvx);
}

static void ClosedForm(const double* q, const double* p, const double* g, double* a) {
  const double qw = q[0], qx = q[1], qy = q[2], qz = q[3];
  double uvx = qy * p[2] - qz * p[1], uvy = qz * p[0] - qx * p[2], uvz = qx * p[1] - qy * p[0];
  uvx += uvx; uvy += uvy; uvz += uvz;
  a[0] = g[0] * uvx + g[1] * uvy + g[2] * uvz;
  const double cx = p[1] * g[2] - p[2] * g[1], cy = p[2] * g[0] - p[0] * g[2], cz = p[0] * g[1] - p[1] * g[0];
  const double up = qx * p[0] + qy * p[1] + qz * p[2];
  const double gu = g[0] * qx + g[1] * qy + g[2] * qz;
  const double gp2 = -2.0 * (g[0] * p[0] + g[1] * p[1] + g[2] * p[2]);
  const double tx = qw * cx + (up * g[0] + gu * p[0] + gp2 * qx);
  const double ty = qw * cy + (up * g[1] + gu * p[1] + gp2 * qy);
  const double tz = qw * cz + (up * g[2] + gu * p[2] + gp2 * qz);
  a[1] = tx + tx;
  a[2] = ty + ty;
  a[3] = tz + tz;
}

int main() {
  std::mt19937_64 rng(7);
  std::uniform_real_distribution<double> unit(-1.0, 1.0), far(-40.0, 40.0);
  for (int trial = 0; trial < 20000; ++trial) {
    double q[4], p[3], g[3];
    for (double& v : q) v = unit(rng);  // raw quaternion: the functors do not normalise
    for (double& v : p) v = far(rng);
    for (double& v : g) v = unit(rng);
    double a[4];
    ClosedForm(q, p, g, a);
    for (int k = 0; k < 4; ++k) {
      const double h = 1e-6;
      double hi[4] = {q[0], q[1], q[2], q[3]}, lo[4] = {q[0], q[1], q[2], q[3]}, wh[3], wl[3];
      hi[k] += h;
      lo[k] -= h;
      World(hi, p, wh);
      World(lo, p, wl);
      const double numeric = (g[0] * (wh[0] - wl[0]) + g[1] * (wh[1] - wl[1]) + g[2] * (wh[2] - wl[2])) / (2 * h);
      // world is quadratic in q: the central difference is exact up to rounding
      if (std::fabs(numeric - a[k]) > 1e-6 * (1.0 + std::fabs(a[k]))) {
        std::fprintf(stderr, "FAILED: trial %d component %d: closed form %.12g, central difference %.12g\n", trial, k,
                     a[k], numeric);
        return 1;
      }
    }
  }
  std::printf("ROTATION TERMS OK\n");
  return 0;
}
