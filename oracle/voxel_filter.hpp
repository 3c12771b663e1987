// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the voxel filters that feed the matcher.
// Follows (under /root/reference/cartographer/sensor/internal):
//   voxel_filter.cc:25-73            first point per voxel, input order kept
//   voxel_filter.cc:58-72            96-bit key from the three int32 cell indices, GetCellIndex = lround(p / size)
//   adaptive_voxel_filter.h:34-84    FilterByMaxRange, AdaptivelyVoxelFiltered (halving + bisection on the edge length)
#pragma once

#include <cstdint>
#include <unordered_set>
#include <vector>

#include "grid.hpp"

namespace oracle {

struct VoxelKey {
  int32_t x, y, z;
  bool operator==(const VoxelKey& o) const { return x == o.x && y == o.y && z == o.z; }
};
struct VoxelKeyHash {
  size_t operator()(const VoxelKey& k) const {
    uint64_t h = static_cast<uint32_t>(k.x) * 0x9E3779B97F4A7C15ull;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.y)) + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.z)) + 0x165667B1ull) * 0xD6E8FEB86659FD93ull;
    return static_cast<size_t>(h ^ (h >> 29));
  }
};

// Indices (into 'indices_in' order) of the points VoxelFilter(size).Filter keeps.
inline std::vector<int> VoxelFilterIndices(const float size, const float* pts, const int stride,
                                           const std::vector<int>& indices_in) {
  std::unordered_set<VoxelKey, VoxelKeyHash> voxels;
  std::vector<int> kept;
  for (const int i : indices_in) {
    const float* p = pts + static_cast<size_t>(i) * stride;
    const VoxelKey k{RoundToInt(p[0] / size), RoundToInt(p[1] / size), RoundToInt(p[2] / size)};
    if (voxels.insert(k).second) kept.push_back(i);
  }
  return kept;
}

// AdaptiveVoxelFilter::Filter: indices of the kept points in the original cloud.
inline std::vector<int> AdaptiveVoxelFilterIndices(const float max_length, const int min_num_points,
                                                   const float max_range, const float* pts, const int stride,
                                                   const int n) {
  std::vector<int> in_range;
  for (int i = 0; i < n; ++i) {
    const float* p = pts + static_cast<size_t>(i) * stride;
    const float norm = std::sqrt(p[0] * p[0] + (p[1] * p[1] + p[2] * p[2]));
    if (norm <= max_range) in_range.push_back(i);
  }
  if (static_cast<int>(in_range.size()) <= min_num_points) return in_range;
  std::vector<int> result = VoxelFilterIndices(max_length, pts, stride, in_range);
  if (static_cast<int>(result.size()) >= min_num_points) return result;
  for (float high_length = max_length; high_length > 1e-2f * max_length; high_length /= 2.f) {
    float low_length = high_length / 2.f;
    result = VoxelFilterIndices(low_length, pts, stride, in_range);
    if (static_cast<int>(result.size()) >= min_num_points) {
      while ((high_length - low_length) / low_length > 1e-1f) {
        const float mid_length = (low_length + high_length) / 2.f;
        const std::vector<int> candidate = VoxelFilterIndices(mid_length, pts, stride, in_range);
        if (static_cast<int>(candidate.size()) >= min_num_points) {
          low_length = mid_length;
          result = candidate;
        } else {
          high_length = mid_length;
        }
      }
      return result;
    }
  }
  return result;
}

}  // namespace oracle
