// Interpolated lookups of the submap grids with analytic gradients — the device form of
// InterpolatedGrid::GetProbability, InterpolatedTSDF::GetTSD and one level of
// InterpolatedMultiResolutionTSDF (paths under
// /root/reference/cartographer/mapping/internal/3d/scan_matching):
//   interpolated_grid.h:50-102            tricubic-Hermite occupancy lookup
//   interpolated_tsdf.h:29-46, 71-116     weight-aware trilinear TSDF lookup
//   interpolated_multi_resolution_tsdf.h:82-137  one level of the fine-to-coarse fall-through
//   interpolated_tsdf.h:159-192           lower-voxel selection (float/double mix)
// plus the pose chain rule shared by the cost kernels (cost.cu) and the on-device scan
// matcher (scan_match.cu).
#pragma once

#include "device_math.cuh"

namespace hg {

// Lower-voxel selection for one axis (interpolated_tsdf.h:159-192): narrow to
// float, containing cell, its float centre, one float step down if the centre
// lies above the double coordinate; x2 is the float sum centre + resolution.
// The reference re-derives the cell index of the lower centre with a second
// GetCellIndex; that is i or i - 1 exactly (|fl(i*res)/res - i| << 0.5 for the
// +-8192 index range), so no second division is needed.
__device__ __forceinline__ void LowerVoxel(const GridDev& g, const double x, double* x1, double* x2,
                                           int* index) {
  const float xf = __double2float_rn(x);
  int i = CellIndexFast(xf, g.resolution, g.inv_resolution);
  float c = __fmul_rn(__int2float_rn(i), g.resolution);
  if (static_cast<double>(c) > x) {
    c = __fsub_rn(c, g.resolution);
    --i;
  }
  *x1 = static_cast<double>(c);
  *x2 = static_cast<double>(__fadd_rn(c, g.resolution));
  *index = i;
}

// The same selection without a branch per axis: q = fl(xf * fl(1/res)) rounded to the nearest
// integer decides the cell whenever q's fraction is outside the band around one half in which
// the product and the IEEE quotient can round differently (CellIndexFast, device_math.cuh);
// there |q - i| < 0.5 strictly, so nearest-even and half-away-from-zero agree. Returns true
// when the axis needs the exact path (LowerVoxel) instead — the caller takes one branch for
// all three axes.
__device__ __forceinline__ bool LowerVoxelNoBranch(const GridDev& g, const double x, double* x1, double* x2,
                                                   int* index) {
  const float xf = __double2float_rn(x);
  const float q = __fmul_rn(xf, g.inv_resolution);
  const int i = __float2int_rn(q);
  const float fi = __int2float_rn(i);
  const float d = __fsub_rn(q, fi);
  const bool exact_path = !(__fsub_rn(0.5f, fabsf(d)) > __fmul_rn(fabsf(q), 4.8e-7f));
  const float c = __fmul_rn(fi, g.resolution);
  const bool above = static_cast<double>(c) > x;
  const float lower = above ? __fsub_rn(c, g.resolution) : c;
  *x1 = static_cast<double>(lower);
  *x2 = static_cast<double>(__fadd_rn(lower, g.resolution));
  *index = above ? i - 1 : i;
  return exact_path;
}

__device__ __forceinline__ void LowerVoxel3(const GridDev& g, const double x, const double y, const double z,
                                            double* x1, double* y1, double* z1, double* x2, double* y2,
                                            double* z2, int* ix, int* iy, int* iz) {
  const bool ex = LowerVoxelNoBranch(g, x, x1, x2, ix);
  const bool ey = LowerVoxelNoBranch(g, y, y1, y2, iy);
  const bool ez = LowerVoxelNoBranch(g, z, z1, z2, iz);
  if (ex | ey | ez) {  // rare: a coordinate within ~5e-7 relative of a cell boundary, or NaN / huge
    LowerVoxel(g, x, x1, x2, ix);
    LowerVoxel(g, y, y1, y2, iy);
    LowerVoxel(g, z, z1, z2, iz);
  }
}

// 1 / d for d = x2 - x1, which equals the resolution up to float rounding of the
// cell centres: two Newton steps from 1 / double(resolution) converge to full
// double precision (the seed is already within ~1e-7 relative).
__device__ __forceinline__ double CellWidthReciprocal(const double seed, const double d) {
  double r = seed * (2.0 - d * seed);
  r = r * (2.0 - d * r);
  return r;
}

// Fetches the 2x2x2 stencil at (ix..ix+1, iy..iy+1, iz..iz+1) in the reference's
// corner order 111,112,121,122,211,212,221,222 (first digit x, last digit z).
// Submap bricks carry a one-voxel apron (hg_internal.cuh), so the whole stencil
// lives in the brick of its lower corner: one index load, eight voxel loads at
// fixed offsets, no branches on the brick faces.
// The voxels are the pre-decoded floats (GridDev::decoded); 'absent' is what a
// voxel of a missing brick reads.
// The packed stencil of one voxel in one 256-bit load (LDG.E.256, sm_100): half the load
// instructions of two float4 loads, and the L2 evict-last hint keeps the stencil tables resident
// under the kernels' streaming output (C2: 0.1851 -> 0.1831 ms for the wide load, 0.1811 with the hint).
__device__ __forceinline__ void LoadStencil256(const float4* p, float* v) {
  unsigned r0, r1, r2, r3, r4, r5, r6, r7;
  asm volatile("ld.global.nc.L2::evict_last.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
               : "l"(p));
  v[0] = __uint_as_float(r0); v[1] = __uint_as_float(r1); v[2] = __uint_as_float(r2); v[3] = __uint_as_float(r3);
  v[4] = __uint_as_float(r4); v[5] = __uint_as_float(r5); v[6] = __uint_as_float(r6); v[7] = __uint_as_float(r7);
}

__device__ __forceinline__ void FetchStencilOfBrick(const GridDev& g, const int id, const int ix, const int iy,
                                                    const int iz, float* v) {
  if (g.stencils != nullptr) {  // packed copy: the whole stencil is one 32-byte sector, one load
    const float4* s = g.stencils + 2 * (id * 512 + InBrick(ix, iy, iz));
    LoadStencil256(s, v);
    return;
  }
  const float* base = g.decoded + (id * kApronBrickVoxels + InApronBrick(ix, iy, iz));
  v[0] = __ldg(base);
  v[1] = __ldg(base + 81);
  v[2] = __ldg(base + 9);
  v[3] = __ldg(base + 90);
  v[4] = __ldg(base + 1);
  v[5] = __ldg(base + 82);
  v[6] = __ldg(base + 10);
  v[7] = __ldg(base + 91);
}
__device__ __forceinline__ void FetchStencil(const GridDev& g, const int ix, const int iy, const int iz,
                                             const float absent, float* v) {
  const int id = BrickId(g.table, ix, iy, iz);
  if (id < 0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = absent;
    return;
  }
  FetchStencilOfBrick(g, id, ix, iy, iz, v);
}

struct ValueGrad {
  double v, gx, gy, gz;  // value and gradient w.r.t. the world point
};

// interpolated_grid.h:50-102
__device__ __forceinline__ ValueGrad LookupOccupancy(const GridDev& g, const double x, const double y,
                                                     const double z) {
  double x1, y1, z1, x2, y2, z2;
  int ix, iy, iz;
  LowerVoxel3(g, x, y, z, &x1, &y1, &z1, &x2, &y2, &z2, &ix, &iy, &iz);
  float raw[8];
  FetchStencil(g, ix, iy, iz, g.a.unknown, raw);
  double q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) q[k] = static_cast<double>(raw[k]);
  // Ceres' Jet / double is a multiplication by the reciprocal (jet.h), which is
  // what the reference's Jacobian-evaluating instantiation computes.
  const double sx = CellWidthReciprocal(g.inv_resolution_d, x2 - x1);
  const double sy = CellWidthReciprocal(g.inv_resolution_d, y2 - y1);
  const double sz = CellWidthReciprocal(g.inv_resolution_d, z2 - z1);
  const double nx = (x - x1) * sx, ny = (y - y1) * sy, nz = (z - z1) * sz;
  const double nxx = nx * nx, nxxx = nx * nxx, nyy = ny * ny, nyyy = ny * nyy, nzz = nz * nz,
               nzzz = nz * nzz;
  // Derivatives of n^2, n^3 with respect to n.
  const double dzz = 2.0 * nz, dzzz = 3.0 * nzz, dyy = 2.0 * ny, dyyy = 3.0 * nyy, dxx = 2.0 * nx,
               dxxx = 3.0 * nxx;
  // z stage: value and d/dnz.
  double a[4], az[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const double qa = q[2 * k], qb = q[2 * k + 1];
    a[k] = (qa - qb) * nzzz * 2. + (qb - qa) * nzz * 3. + qa;
    az[k] = (qa - qb) * dzzz * 2. + (qb - qa) * dzz * 3.;
  }
  // y stage: pairs (a0, a1) -> x = 1 side, (a2, a3) -> x = 2 side.
  double b[2], by[2], bz[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const double qa = a[2 * k], qb = a[2 * k + 1];
    b[k] = (qa - qb) * nyyy * 2. + (qb - qa) * nyy * 3. + qa;
    by[k] = (qa - qb) * dyyy * 2. + (qb - qa) * dyy * 3.;
    const double za = az[2 * k], zb = az[2 * k + 1];
    bz[k] = (za - zb) * nyyy * 2. + (zb - za) * nyy * 3. + za;
  }
  ValueGrad r;
  r.v = (b[0] - b[1]) * nxxx * 2. + (b[1] - b[0]) * nxx * 3. + b[0];
  r.gx = ((b[0] - b[1]) * dxxx * 2. + (b[1] - b[0]) * dxx * 3.) * sx;
  r.gy = ((by[0] - by[1]) * nxxx * 2. + (by[1] - by[0]) * nxx * 3. + by[0]) * sy;
  r.gz = ((bz[0] - bz[1]) * nxxx * 2. + (bz[1] - bz[0]) * nxx * 3. + bz[0]) * sz;
  return r;
}

// interpolated_tsdf.h:29-46 on (value, partials) tuples, branch-free. The
// reference's weights only ever decide "observed or not" (w == 0.0 tests; sums
// of non-negative weights are zero iff both are), so a flag per voxel carries
// them exactly: a stored weight value of 0 or 1 decodes to 0.0f, anything larger
// to a positive weight (value_conversion_tables.cc:29-51 with lower bound 0).
// Copy branches pass the upstream partials through; both-unobserved is a constant.
struct Q1 { double v, dz; bool o; };
struct Q2 { double v, dy, dz; bool o; };

__device__ __forceinline__ Q1 BlendZ(const double unknown, const double q1, const double q2,
                                     const bool o1, const bool o2, const double n) {
  Q1 r;
  const double d = q2 - q1;
  const bool both = o1 && o2;
  r.v = both ? d * n + q1 : (o1 ? q1 : (o2 ? q2 : unknown));
  r.dz = both ? d : 0.0;
  r.o = o1 || o2;
  return r;
}
__device__ __forceinline__ Q2 BlendY(const double unknown, const Q1& a, const Q1& b, const double n) {
  Q2 r;
  const bool both = a.o && b.o;
  const double d = b.v - a.v;
  r.v = both ? d * n + a.v : (a.o ? a.v : (b.o ? b.v : unknown));
  r.dy = both ? d : 0.0;
  r.dz = both ? (b.dz - a.dz) * n + a.dz : (a.o ? a.dz : (b.o ? b.dz : 0.0));
  r.o = a.o || b.o;
  return r;
}

// Shared tail of the TSDF lookups: 7 weight-aware blends, z then y then x.
__device__ __forceinline__ ValueGrad BlendTsdf(const double unknown, const bool* o, const double* q,
                                               const double nx, const double ny, const double nz,
                                               const double sx, const double sy, const double sz) {
  const Q1 q11 = BlendZ(unknown, q[0], q[1], o[0], o[1], nz);
  const Q1 q12 = BlendZ(unknown, q[2], q[3], o[2], o[3], nz);
  const Q1 q21 = BlendZ(unknown, q[4], q[5], o[4], o[5], nz);
  const Q1 q22 = BlendZ(unknown, q[6], q[7], o[6], o[7], nz);
  const Q2 q1 = BlendY(unknown, q11, q12, ny);
  const Q2 q2 = BlendY(unknown, q21, q22, ny);
  ValueGrad r;
  const bool both = q1.o && q2.o;
  const double d = q2.v - q1.v;
  r.v = both ? d * nx + q1.v : (q1.o ? q1.v : (q2.o ? q2.v : unknown));
  r.gx = both ? d * sx : 0.0;
  r.gy = (both ? (q2.dy - q1.dy) * nx + q1.dy : (q1.o ? q1.dy : (q2.o ? q2.dy : 0.0))) * sy;
  r.gz = (both ? (q2.dz - q1.dz) * nx + q1.dz : (q1.o ? q1.dz : (q2.o ? q2.dz : 0.0))) * sz;
  return r;
}

// One level of the multi-resolution lookup, the form the Jacobian kernels run: a level is used
// only when all eight corners are observed (interpolated_multi_resolution_tsdf.h:99-106), i.e.
// when none of the eight pre-decoded floats is the NaN marker — which is what the sum of the
// eight says (decoded TSDs are finite, |tsd| <= truncation distance) — and a missing brick
// is eight unobserved corners. Then all seven blends take the lerp branch: plain trilinear.
__device__ __forceinline__ bool LookupTsdfLevelAllObserved(const GridDev& g, const double x, const double y,
                                                           const double z, ValueGrad* out) {
  double x1, y1, z1, x2, y2, z2;
  int ix, iy, iz;
  LowerVoxel3(g, x, y, z, &x1, &y1, &z1, &x2, &y2, &z2, &ix, &iy, &iz);
  const int id = BrickId(g.table, ix, iy, iz);
  if (id < 0) return false;
  float r[8];
  FetchStencilOfBrick(g, id, ix, iy, iz, r);
  const float r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3], r4 = r[4], r5 = r[5], r6 = r[6], r7 = r[7];
  const float all = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
  if (all != all) return false;
  const double q0 = r0, q1 = r1, q2 = r2, q3 = r3, q4 = r4, q5 = r5, q6 = r6, q7 = r7;
  const double sx = CellWidthReciprocal(g.inv_resolution_d, x2 - x1);
  const double sy = CellWidthReciprocal(g.inv_resolution_d, y2 - y1);
  const double sz = CellWidthReciprocal(g.inv_resolution_d, z2 - z1);
  const double nx = (x - x1) * sx, ny = (y - y1) * sy, nz = (z - z1) * sz;
  const double d0 = q1 - q0, d1 = q3 - q2, d2 = q5 - q4, d3 = q7 - q6;
  const double a0 = d0 * nz + q0, a1 = d1 * nz + q2, a2 = d2 * nz + q4, a3 = d3 * nz + q6;
  const double b0y = a1 - a0, b1y = a3 - a2;
  const double b0 = b0y * ny + a0, b0z = (d1 - d0) * ny + d0;
  const double b1 = b1y * ny + a2, b1z = (d3 - d2) * ny + d2;
  const double cx = b1 - b0;
  out->v = cx * nx + b0;
  out->gx = cx * sx;
  out->gy = ((b1y - b0y) * nx + b0y) * sy;
  out->gz = ((b1z - b0z) * nx + b0z) * sz;
  return true;
}

// InterpolatedTSDF::GetTSD (interpolated_tsdf.h:71-116): the single-grid lookup with its
// weight-aware blends. (The multi-resolution lookup runs LookupTsdfLevelAllObserved per level.)
__device__ __forceinline__ void LookupTsdf(const GridDev& g, const double x, const double y, const double z,
                                           ValueGrad* out) {
  double x1, y1, z1, x2, y2, z2;
  int ix, iy, iz;
  LowerVoxel3(g, x, y, z, &x1, &y1, &z1, &x2, &y2, &z2, &ix, &iy, &iz);
  float raw[8];
  FetchStencil(g, ix, iy, iz, __int_as_float(0x7fc00000), raw);
  bool o[8];
  int observed = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    o[k] = raw[k] == raw[k];  // NaN marks an unobserved voxel
    observed += o[k] ? 1 : 0;
  }
  if (observed == 0) {  // interpolated_tsdf.h:87-90
    out->v = static_cast<double>(g.min_tsd);
    out->gx = out->gy = out->gz = 0.0;
    return;
  }
  double q[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) q[k] = static_cast<double>(raw[k]);
  const double sx = CellWidthReciprocal(g.inv_resolution_d, x2 - x1);
  const double sy = CellWidthReciprocal(g.inv_resolution_d, y2 - y1);
  const double sz = CellWidthReciprocal(g.inv_resolution_d, z2 - z1);
  const double nx = (x - x1) * sx, ny = (y - y1) * sy, nz = (z - z1) * sz;
  // interpolated_tsdf.h:34 hard-codes -0.3 for a pair of unobserved voxels.
  *out = BlendTsdf(-0.3, o, q, nx, ny, nz, sx, sy, sz);
}

// LOOKUP: 0 occupancy, 1 TSDF, 2 multi-resolution TSDF. Returns the residual
// and its gradient w.r.t. the world point, both already scaled.

// world = R(q) p + t with Eigen's _transformVector formula on the raw (not normalised)
// quaternion (transform/rigid_transform.h:191-196): uv = 2 (u x p); world = p + w uv + u x uv + t.
struct WorldPoint {
  double x, y, z;
  double uvx, uvy, uvz;
};
__device__ __forceinline__ WorldPoint TransformPoint(const double qw, const double qx, const double qy,
                                                     const double qz, const double tx, const double ty,
                                                     const double tz, const double px, const double py,
                                                     const double pz) {
  WorldPoint w;
  double uvx = qy * pz - qz * py, uvy = qz * px - qx * pz, uvz = qx * py - qy * px;
  uvx += uvx; uvy += uvy; uvz += uvz;
  w.uvx = uvx; w.uvy = uvy; w.uvz = uvz;
  w.x = px + qw * uvx + (qy * uvz - qz * uvy) + tx;
  w.y = py + qw * uvy + (qz * uvx - qx * uvz) + ty;
  w.z = pz + qw * uvz + (qx * uvy - qy * uvx) + tz;
  return w;
}

// a_k = grad . d world / d q_k for k = w, x, y, z, g = grad. With u = (qx, qy, qz), w = qw and
// world = p + w uv + u x uv + t, uv = 2 u x p:
//   a_w = g . uv
//   a_u = 2 w (p x g) + 2 [(u . p) g + (g . u) p - 2 (g . p) u]
// (from g . (e x p) = e . (p x g), (u x p) x g = p (u . g) - u (p . g), p x (g x u) =
// g (p . u) - u (p . g)): 37 FP64 operations, where differentiating the formula term by term
// takes ~50.
__device__ __forceinline__ void RotationTerms(const double qw, const double qx, const double qy, const double qz,
                                              const double px, const double py, const double pz,
                                              const WorldPoint& w, const ValueGrad& v, double* aw, double* ax,
                                              double* ay, double* az) {
  *aw = v.gx * w.uvx + v.gy * w.uvy + v.gz * w.uvz;
  const double cx = py * v.gz - pz * v.gy, cy = pz * v.gx - px * v.gz, cz = px * v.gy - py * v.gx;
  const double up = qx * px + qy * py + qz * pz;
  const double gu = v.gx * qx + v.gy * qy + v.gz * qz;
  const double gp2 = -2.0 * (v.gx * px + v.gy * py + v.gz * pz);
  const double tx = qw * cx + (up * v.gx + gu * px + gp2 * qx);
  const double ty = qw * cy + (up * v.gy + gu * py + gp2 * qy);
  const double tz = qw * cz + (up * v.gz + gu * pz + gp2 * qz);
  *ax = tx + tx;
  *ay = ty + ty;
  *az = tz + tz;
}

}  // namespace hg
