// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the three interpolated grid lookups, templated on
// T = double or Jet<N> exactly as the reference templates on Ceres Jets.
//
// Follows (under /root/reference/cartographer/mapping/internal/3d/scan_matching):
//   interpolated_grid.h:50-102, 106-138              tricubic occupancy
//   interpolated_tsdf.h:29-46, 71-116, 129-192       weight-aware trilinear TSDF
//   interpolated_multi_resolution_tsdf.h:29-46, 82-137, 139-171, 175-208
#pragma once

#include <vector>

#include "grid.hpp"
#include "jet.hpp"

namespace oracle {

// Lower-voxel selection shared by all three interpolators
// (interpolated_tsdf.h:159-192 and its two siblings): narrow the scalar part
// to float, round to the containing cell, take its float centre, step one
// float resolution down where the centre lies above the double coordinate;
// the upper data point is the float sum centre + resolution.
template <typename Grid>
inline void InterpolationDataPoints(const Grid& grid, const double x, const double y,
                                    const double z, double* x1, double* y1, double* z1,
                                    double* x2, double* y2, double* z2) {
  Vector3f center = grid.GetCenterOfCell(grid.GetCellIndex(
      Vector3f(static_cast<float>(x), static_cast<float>(y), static_cast<float>(z))));
  if (center.x > x) center.x -= grid.resolution();
  if (center.y > y) center.y -= grid.resolution();
  if (center.z > z) center.z -= grid.resolution();
  *x1 = center.x;
  *y1 = center.y;
  *z1 = center.z;
  *x2 = center.x + grid.resolution();  // float addition, widened afterwards
  *y2 = center.y + grid.resolution();
  *z2 = center.z + grid.resolution();
}

class InterpolatedGrid {
 public:
  explicit InterpolatedGrid(const HybridGrid& grid) : grid_(grid) {}

  template <typename T>
  T GetProbability(const T& x, const T& y, const T& z) const {
    double x1, y1, z1, x2, y2, z2;
    InterpolationDataPoints(grid_, ScalarPart(x), ScalarPart(y), ScalarPart(z), &x1, &y1, &z1,
                            &x2, &y2, &z2);
    const Array3i i1 = grid_.GetCellIndex(
        Vector3f(static_cast<float>(x1), static_cast<float>(y1), static_cast<float>(z1)));
    const double q111 = grid_.GetProbability(i1);
    const double q112 = grid_.GetProbability(i1 + Array3i(0, 0, 1));
    const double q121 = grid_.GetProbability(i1 + Array3i(0, 1, 0));
    const double q122 = grid_.GetProbability(i1 + Array3i(0, 1, 1));
    const double q211 = grid_.GetProbability(i1 + Array3i(1, 0, 0));
    const double q212 = grid_.GetProbability(i1 + Array3i(1, 0, 1));
    const double q221 = grid_.GetProbability(i1 + Array3i(1, 1, 0));
    const double q222 = grid_.GetProbability(i1 + Array3i(1, 1, 1));

    const T nx = (x - x1) / (x2 - x1);
    const T ny = (y - y1) / (y2 - y1);
    const T nz = (z - z1) / (z2 - z1);
    const T nxx = nx * nx;
    const T nxxx = nx * nxx;
    const T nyy = ny * ny;
    const T nyyy = ny * nyy;
    const T nzz = nz * nz;
    const T nzzz = nz * nzz;

    // Hermite blend A (2t^3 - 3t^2 + 1) + B (-2t^3 + 3t^2), z then y then x.
    const T q11 = (q111 - q112) * nzzz * 2. + (q112 - q111) * nzz * 3. + q111;
    const T q12 = (q121 - q122) * nzzz * 2. + (q122 - q121) * nzz * 3. + q121;
    const T q21 = (q211 - q212) * nzzz * 2. + (q212 - q211) * nzz * 3. + q211;
    const T q22 = (q221 - q222) * nzzz * 2. + (q222 - q221) * nzz * 3. + q221;
    const T q1 = (q11 - q12) * nyyy * 2. + (q12 - q11) * nyy * 3. + q11;
    const T q2 = (q21 - q22) * nyyy * 2. + (q22 - q21) * nyy * 3. + q21;
    return (q1 - q2) * nxxx * 2. + (q2 - q1) * nxx * 3. + q1;
  }

 private:
  const HybridGrid& grid_;
};

// interpolated_tsdf.h:29-46 (unknown constant -0.3) and
// interpolated_multi_resolution_tsdf.h:29-46 (unknown constant = that level's
// minimum TSD) share this shape; 'unknown' carries the constant.
template <typename T, typename T2>
inline void InterpolateLinear(const double unknown, const T2& q1, const T2& q2, const double w1,
                              const double w2, const T& ratio, T& q, double& w) {
  if (w1 == 0.0 && w2 == 0.0) {
    q = T(unknown);
    w = 0.0;
  } else if (w1 == 0.0) {
    q = T(q2);
    w = w2;
  } else if (w2 == 0.0) {
    q = T(q1);
    w = w1;
  } else {
    q = (q2 - q1) * ratio + q1;
    w = w1 + w2;
  }
}

struct TsdfCorners {
  double w[8];  // order 111,112,121,122,211,212,221,222 (x,y,z; 2 = +1)
  double q[8];
};

inline void FetchWeights(const HybridGridTSDF& g, const Array3i& i1, double* w) {
  w[0] = g.GetWeight(i1);
  w[1] = g.GetWeight(i1 + Array3i(0, 0, 1));
  w[2] = g.GetWeight(i1 + Array3i(0, 1, 0));
  w[3] = g.GetWeight(i1 + Array3i(0, 1, 1));
  w[4] = g.GetWeight(i1 + Array3i(1, 0, 0));
  w[5] = g.GetWeight(i1 + Array3i(1, 0, 1));
  w[6] = g.GetWeight(i1 + Array3i(1, 1, 0));
  w[7] = g.GetWeight(i1 + Array3i(1, 1, 1));
}
inline void FetchTsds(const HybridGridTSDF& g, const Array3i& i1, double* q) {
  q[0] = g.GetTSD(i1);
  q[1] = g.GetTSD(i1 + Array3i(0, 0, 1));
  q[2] = g.GetTSD(i1 + Array3i(0, 1, 0));
  q[3] = g.GetTSD(i1 + Array3i(0, 1, 1));
  q[4] = g.GetTSD(i1 + Array3i(1, 0, 0));
  q[5] = g.GetTSD(i1 + Array3i(1, 0, 1));
  q[6] = g.GetTSD(i1 + Array3i(1, 1, 0));
  q[7] = g.GetTSD(i1 + Array3i(1, 1, 1));
}

template <typename T>
inline T BlendTsd(const double unknown, const double* w8, const double* q8, const T& nx,
                  const T& ny, const T& nz) {
  T q11, q12, q21, q22, q1, q2, q;
  double w11, w12, w21, w22, w1, w2, w;
  InterpolateLinear(unknown, q8[0], q8[1], w8[0], w8[1], nz, q11, w11);
  InterpolateLinear(unknown, q8[2], q8[3], w8[2], w8[3], nz, q12, w12);
  InterpolateLinear(unknown, q8[4], q8[5], w8[4], w8[5], nz, q21, w21);
  InterpolateLinear(unknown, q8[6], q8[7], w8[6], w8[7], nz, q22, w22);
  InterpolateLinear(unknown, q11, q12, w11, w12, ny, q1, w1);
  InterpolateLinear(unknown, q21, q22, w21, w22, ny, q2, w2);
  InterpolateLinear(unknown, q1, q2, w1, w2, nx, q, w);
  return q;
}

class InterpolatedTSDF {
 public:
  explicit InterpolatedTSDF(const HybridGridTSDF& tsdf) : tsdf_(tsdf) {}

  template <typename T>
  T GetTSD(const T& x, const T& y, const T& z) const {
    double x1, y1, z1, x2, y2, z2;
    InterpolationDataPoints(tsdf_, ScalarPart(x), ScalarPart(y), ScalarPart(z), &x1, &y1, &z1,
                            &x2, &y2, &z2);
    const Array3i i1 = tsdf_.GetCellIndex(
        Vector3f(static_cast<float>(x1), static_cast<float>(y1), static_cast<float>(z1)));
    double w[8];
    FetchWeights(tsdf_, i1, w);
    if (w[0] == 0.0 && w[1] == 0.0 && w[2] == 0.0 && w[3] == 0.0 && w[4] == 0.0 && w[5] == 0.0 &&
        w[6] == 0.0 && w[7] == 0.0) {
      return T(tsdf_.ValueConverter().getMinTSD());
    }
    double q[8];
    FetchTsds(tsdf_, i1, q);
    const T nx = (x - x1) / (x2 - x1);
    const T ny = (y - y1) / (y2 - y1);
    const T nz = (z - z1) / (z2 - z1);
    // interpolated_tsdf.h:34 hard-codes -0.3 for a pair of unobserved voxels.
    return BlendTsd(-0.3, w, q, nx, ny, nz);
  }

  // interpolated_tsdf.h:129-156: trilinear blend of the corner weights, z then
  // y then x. interpolated_multi_resolution_tsdf.h:139-171 is the same lookup
  // on the finest grid (its loop returns in the first iteration).
  template <typename T>
  T GetWeight(const T& x, const T& y, const T& z) const {
    double x1, y1, z1, x2, y2, z2;
    InterpolationDataPoints(tsdf_, ScalarPart(x), ScalarPart(y), ScalarPart(z), &x1, &y1, &z1,
                            &x2, &y2, &z2);
    const Array3i i1 = tsdf_.GetCellIndex(
        Vector3f(static_cast<float>(x1), static_cast<float>(y1), static_cast<float>(z1)));
    double w[8];
    FetchWeights(tsdf_, i1, w);
    const T nx = (x - x1) / (x2 - x1);
    const T ny = (y - y1) / (y2 - y1);
    const T nz = (z - z1) / (z2 - z1);
    const T q11 = (w[1] - w[0]) * nz + w[0];
    const T q12 = (w[3] - w[2]) * nz + w[2];
    const T q21 = (w[5] - w[4]) * nz + w[4];
    const T q22 = (w[7] - w[6]) * nz + w[6];
    const T q1 = (q12 - q11) * ny + q11;
    const T q2 = (q22 - q21) * ny + q21;
    return (q2 - q1) * nx + q1;
  }

 private:
  const HybridGridTSDF& tsdf_;
};

class InterpolatedMultiResolutionTSDF {
 public:
  // The reference keeps a reference to the caller's vector
  // (interpolated_multi_resolution_tsdf.h:218); a copy is equivalent while the
  // caller's vector is alive and safe when it is not.
  explicit InterpolatedMultiResolutionTSDF(const std::vector<const HybridGridTSDF*>& pyramid)
      : pyramid_(pyramid) {
    double last = 0.;
    for (const HybridGridTSDF* g : pyramid_) {
      if (!(last < g->resolution())) std::abort();  // reference CHECK_LT, :62-67
      last = g->resolution();
    }
  }

  template <typename T>
  T GetTSD(const T& x, const T& y, const T& z) const {
    for (const HybridGridTSDF* tsdf : pyramid_) {
      double x1, y1, z1, x2, y2, z2;
      InterpolationDataPoints(*tsdf, ScalarPart(x), ScalarPart(y), ScalarPart(z), &x1, &y1, &z1,
                              &x2, &y2, &z2);
      const Array3i i1 = tsdf->GetCellIndex(
          Vector3f(static_cast<float>(x1), static_cast<float>(y1), static_cast<float>(z1)));
      double w[8];
      FetchWeights(*tsdf, i1, w);
      int invalid = 0;
      for (int k = 0; k < 8; ++k) invalid += int(w[k] == 0.0);
      if (invalid > 0) continue;  // any unobserved corner -> next coarser level
      double q[8];
      FetchTsds(*tsdf, i1, q);
      const T nx = (x - x1) / (x2 - x1);
      const T ny = (y - y1) / (y2 - y1);
      const T nz = (z - z1) / (z2 - z1);
      return BlendTsd(static_cast<double>(tsdf->ValueConverter().getMinTSD()), w, q, nx, ny, nz);
    }
    return T(pyramid_.front()->ValueConverter().getMinTSD());
  }

 private:
  const std::vector<const HybridGridTSDF*> pyramid_;
};

}  // namespace oracle
