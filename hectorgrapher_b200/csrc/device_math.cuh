// Device helpers: bit-exact single-precision steps (no FMA contraction), brick
// table lookups and value decoding.
#pragma once

#include "hg_internal.cuh"

namespace hg {

// common/port.h:40-42 RoundToInt == std::lround: half away from zero. Evaluated
// exactly without the 64-bit lroundf path: x - trunc(x) is exact in float, so
// comparing it with +-0.5 decides the rounding without any rounding error.
__device__ __forceinline__ int RoundToIntF(const float x) {
  const int i = __float2int_rz(x);
  const float f = __fsub_rn(x, __int2float_rn(i));
  return i + (f >= 0.5f ? 1 : 0) - (f <= -0.5f ? 1 : 0);
}

// HybridGridBase::GetCellIndex for one component (hybrid_grid_base.h:428-433):
// a true IEEE float division, then lround.
__device__ __forceinline__ int CellIndex1(const float p, const float resolution) {
  return RoundToIntF(__fdiv_rn(p, resolution));
}

// Same result as CellIndex1, bit for bit, with the IEEE division taken only
// when it can matter. q = p * fl(1/res) is within 1.5 * 2^-23 |q| of the
// correctly rounded quotient; lround of the two can only differ when q's
// fraction is that close to 0.5, so outside a 4 * 2^-23 |q| band around the
// half-way point the product decides, inside it the exact division does.
__device__ __forceinline__ int CellIndexFast(const float p, const float resolution,
                                             const float inv_resolution) {
  const float q = __fmul_rn(p, inv_resolution);
  const int i = __float2int_rz(q);
  const float f = fabsf(__fsub_rn(q, __int2float_rn(i)));
  if (fabsf(f - 0.5f) <= fabsf(q) * 4.8e-7f) return CellIndex1(p, resolution);
  return i + (f > 0.5f ? (q > 0.f ? 1 : -1) : 0);
}

// Brick id of the brick containing voxel (x, y, z), or -1.
__device__ __forceinline__ int BrickId(const BrickTable& t, const int x, const int y, const int z) {
  if (t.hash_mask != 0) {
    const int kx = x >> 3, ky = y >> 3, kz = z >> 3;
    const int4* entries = reinterpret_cast<const int4*>(t.index);
    unsigned h = BrickHash(kx, ky, kz) & t.hash_mask;
    for (;;) {
      const int4 e = __ldg(entries + h);
      if (e.w < 0) return -1;
      if (e.x == kx && e.y == ky && e.z == kz) return e.w;
      h = (h + 1) & t.hash_mask;
    }
  }
  const unsigned bx = static_cast<unsigned>((x >> 3) - t.bmin[0]);
  const unsigned by = static_cast<unsigned>((y >> 3) - t.bmin[1]);
  const unsigned bz = static_cast<unsigned>((z >> 3) - t.bmin[2]);
  if (bx >= static_cast<unsigned>(t.bdim[0]) || by >= static_cast<unsigned>(t.bdim[1]) ||
      bz >= static_cast<unsigned>(t.bdim[2])) {
    return -1;
  }
  return __ldg(t.index + ((static_cast<int>(bz) * t.bdim[1] + static_cast<int>(by)) * t.bdim[0] + static_cast<int>(bx)));
}

__device__ __forceinline__ int InBrick(const int x, const int y, const int z) {
  return ((z & 7) << 6) | ((y & 7) << 3) | (x & 7);
}

// Core position inside a 9x9x9 apron brick (submap grids).
__device__ __forceinline__ int InApronBrick(const int x, const int y, const int z) {
  return ((z & 7) * 9 + (y & 7)) * 9 + (x & 7);
}

__device__ __forceinline__ unsigned short LoadU16(const BrickTable& t, int x, int y, int z) {
  const int id = BrickId(t, x, y, z);
  if (id < 0) return 0;
  return __ldg(static_cast<const unsigned short*>(t.bricks) + static_cast<size_t>(id) * kApronBrickVoxels +
               InApronBrick(x, y, z));
}
__device__ __forceinline__ ushort2 LoadTsdf(const BrickTable& t, int x, int y, int z) {
  const int id = BrickId(t, x, y, z);
  if (id < 0) return make_ushort2(0, 0);
  return __ldg(static_cast<const ushort2*>(t.bricks) + static_cast<size_t>(id) * kApronBrickVoxels +
               InApronBrick(x, y, z));
}
__device__ __forceinline__ unsigned char LoadU8(const BrickTable& t, int x, int y, int z) {
  const int id = BrickId(t, x, y, z);
  if (id < 0) return 0;
  return __ldg(static_cast<const unsigned char*>(t.bricks) + static_cast<size_t>(id) * 512 +
               InBrick(x, y, z));
}

// value_conversion_tables.cc:29-51 evaluated directly instead of through the
// 65536-entry table: same float multiply, same float add.
__device__ __forceinline__ float Decode(const ValueDecode& d, const unsigned short raw) {
  const unsigned v = raw & 0x7FFFu;
  if (v == 0) return d.unknown;
  return __fadd_rn(__fmul_rn(static_cast<float>(v), d.scale), d.offset);
}

// Eigen float quaternion * vector (_transformVector), op by op, no contraction.
__device__ __forceinline__ void RotateF(const float qw, const float qx, const float qy,
                                        const float qz, const float px, const float py,
                                        const float pz, float* ox, float* oy, float* oz) {
  float uvx = __fsub_rn(__fmul_rn(qy, pz), __fmul_rn(qz, py));
  float uvy = __fsub_rn(__fmul_rn(qz, px), __fmul_rn(qx, pz));
  float uvz = __fsub_rn(__fmul_rn(qx, py), __fmul_rn(qy, px));
  uvx = __fadd_rn(uvx, uvx);
  uvy = __fadd_rn(uvy, uvy);
  uvz = __fadd_rn(uvz, uvz);
  const float cx = __fsub_rn(__fmul_rn(qy, uvz), __fmul_rn(qz, uvy));
  const float cy = __fsub_rn(__fmul_rn(qz, uvx), __fmul_rn(qx, uvz));
  const float cz = __fsub_rn(__fmul_rn(qx, uvy), __fmul_rn(qy, uvx));
  *ox = __fadd_rn(__fadd_rn(px, __fmul_rn(qw, uvx)), cx);
  *oy = __fadd_rn(__fadd_rn(py, __fmul_rn(qw, uvy)), cy);
  *oz = __fadd_rn(__fadd_rn(pz, __fmul_rn(qw, uvz)), cz);
}

}  // namespace hg
