// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the correlative (branch-and-bound) 3D scan matcher and
// everything it needs: precomputation grids, rotational prefilter, low
// resolution leaf gate, and the occupancy / TSDF writers used to build the
// reference's own test fixtures.
//
// Follows (under /root/reference/cartographer):
//   mapping/internal/3d/scan_matching/precomputation_grid_3d.cc:49-97
//   mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.cc:60-94   stack
//   ... :96-127 (DiscreteScan3D, Candidate3D), :157-228 (Match*), :230-326 (discretise),
//   ... :328-398 (lowest-resolution candidates, ScoreCandidates), :400-471 (pose, B&B)
//   mapping/internal/3d/scan_matching/low_resolution_matcher.cc:24-69
//   mapping/internal/3d/scan_matching/rotational_scan_matcher.cc:121-185
//   sensor/point_cloud.cc:34-43 (TransformPointCloud)
//   mapping/3d/range_data_inserter_3d.cc:27-101 (occupancy writer, fixtures only)
//   mapping/3d/tsdf_range_data_inserter_3d.cc:294-343, 725-737 (TSDF writer, fixtures only)
#pragma once

#include <algorithm>
#include <cmath>
#include <functional>
#include <limits>
#include <memory>
#include <vector>

#include "geometry.hpp"
#include "grid.hpp"

namespace oracle {

// ---------------------------------------------------------------------------
// Writers (fixture builders).

struct OccupancyInserterOptions {
  float hit_probability = 0.55f;
  float miss_probability = 0.49f;
  int num_free_space_voxels = 2;
};

class OccupancyGridRangeDataInserter3D {
 public:
  explicit OccupancyGridRangeDataInserter3D(const OccupancyInserterOptions& o)
      : options_(o),
        hit_table_(ComputeLookupTableToApplyOdds(Odds(o.hit_probability))),
        miss_table_(ComputeLookupTableToApplyOdds(Odds(o.miss_probability))) {}

  void Insert(const Vector3f& origin, const std::vector<Vector3f>& returns,
              HybridGrid* grid) const {
    for (const Vector3f& hit : returns) {
      grid->ApplyLookupTable(grid->GetCellIndex(hit), hit_table_);
    }
    const Array3i origin_cell = grid->GetCellIndex(origin);
    for (const Vector3f& hit : returns) {
      const Array3i hit_cell = grid->GetCellIndex(hit);
      const Array3i delta = hit_cell - origin_cell;
      const int num_samples =
          std::max(std::abs(delta.x), std::max(std::abs(delta.y), std::abs(delta.z)));
      for (int position = std::max(0, num_samples - options_.num_free_space_voxels);
           position < num_samples; ++position) {
        const Array3i miss_cell(origin_cell.x + delta.x * position / num_samples,
                                origin_cell.y + delta.y * position / num_samples,
                                origin_cell.z + delta.z * position / num_samples);
        grid->ApplyLookupTable(miss_cell, miss_table_);
      }
    }
    grid->FinishUpdate();
  }

 private:
  const OccupancyInserterOptions options_;
  const std::vector<uint16> hit_table_;
  const std::vector<uint16> miss_table_;
};

// proto::TSDFRangeDataInserterOptions3D: the numeric options are doubles in the proto
// (mapping/proto/3d/tsdf_range_data_inserter_options_3d.proto) and are narrowed where the
// reference narrows them.
struct TsdfInserterOptions {
  double relative_truncation_distance = 2.5;
  double maximum_weight = 1000.;
  int num_free_space_voxels = 0;
  double weight_function_epsilon = 0.1;
  double weight_function_sigma = 4.;
  double min_range = 0.;
  double max_range = 1e30;
  double insertion_ratio = 1.;
};

// Default ray-cast branch of TSDFRangeDataInserter3D::InsertHit (:294-343) + UpdateCell (:725-737).
inline void TsdfInsertHit(const TsdfInserterOptions& options, const Vector3f& hit,
                          const Vector3f& origin, HybridGridTSDF* tsdf) {
  const Vector3f ray(hit.x - origin.x, hit.y - origin.y, hit.z - origin.z);
  const float range = std::sqrt(ray.x * ray.x + (ray.y * ray.y + ray.z * ray.z));
  const float truncation_distance =
      static_cast<float>(options.relative_truncation_distance * tsdf->resolution());
  if (range < truncation_distance) return;
  const float truncation_ratio = truncation_distance / range;
  const bool update_free_space = options.num_free_space_voxels > 0;
  const float fb = 1.0f - truncation_ratio;
  const float fe = 1.0f + truncation_ratio;
  const Vector3f ray_begin =
      update_free_space ? origin
                        : Vector3f(origin.x + fb * ray.x, origin.y + fb * ray.y, origin.z + fb * ray.z);
  const Vector3f ray_end(origin.x + fe * ray.x, origin.y + fe * ray.y, origin.z + fe * ray.z);
  const Array3i begin_cell = tsdf->GetCellIndex(ray_begin);
  const Array3i end_cell = tsdf->GetCellIndex(ray_end);
  const Array3i delta = end_cell - begin_cell;
  const int num_samples =
      std::max(std::abs(delta.x), std::max(std::abs(delta.y), std::abs(delta.z)));
  const float epsilon = static_cast<float>(options.weight_function_epsilon);
  const float sigma = static_cast<float>(options.weight_function_sigma);
  const float maximum_weight = static_cast<float>(options.maximum_weight);
  for (int position = 0; position <= num_samples; ++position) {
    const Array3i cell(
        begin_cell.x + static_cast<int>(std::round(float(delta.x) * float(position) / float(num_samples))),
        begin_cell.y + static_cast<int>(std::round(float(delta.y) * float(position) / float(num_samples))),
        begin_cell.z + static_cast<int>(std::round(float(delta.z) * float(position) / float(num_samples))));
    const Vector3f center = tsdf->GetCenterOfCell(cell);
    const Vector3f d(center.x - origin.x, center.y - origin.y, center.z - origin.z);
    const float distance_cell_to_origin = std::sqrt(d.x * d.x + (d.y * d.y + d.z * d.z));
    float update_tsd = range - distance_cell_to_origin;
    update_tsd = Clamp(update_tsd, -truncation_distance, truncation_distance);
    float update_weight = 1.0f;
    const float normalized = update_tsd / truncation_distance;
    if (normalized < -epsilon) {
      // std::pow(float, int) and the product with the float sigma are evaluated in double.
      update_weight = float(std::exp(-sigma * std::pow(-normalized - epsilon, 2)));
    }
    if (update_weight == 0.f) continue;
    const float old_weight = tsdf->GetWeight(cell);
    const float old_sdf = tsdf->GetTSD(cell);
    float updated_weight = old_weight + update_weight;
    const float updated_sdf = (old_sdf * old_weight + update_tsd * update_weight) / updated_weight;
    updated_weight = std::min(updated_weight, maximum_weight);
    tsdf->SetCell(cell, updated_sdf, updated_weight);
  }
}

// TSDFRangeDataInserter3D::Insert, the branch without normal projection (:700-722): the
// insertion-ratio counter, NaN and range gates, then InsertHit per return.
inline void TsdfInsert(const TsdfInserterOptions& options, const Vector3f& origin,
                       const std::vector<Vector3f>& returns, HybridGridTSDF* tsdf) {
  size_t num_inserted_points = 0;
  size_t num_omitted_points = 0;
  for (const Vector3f& hit : returns) {
    if (double(num_inserted_points) <=
        options.insertion_ratio * double(num_inserted_points + num_omitted_points)) {
      num_inserted_points++;
    } else {
      num_omitted_points++;
      continue;
    }
    if (std::isnan(hit.x) || std::isnan(hit.y) || std::isnan(hit.z)) continue;
    const Vector3f r(hit.x - origin.x, hit.y - origin.y, hit.z - origin.z);
    const float r0 = std::sqrt(r.x * r.x + (r.y * r.y + r.z * r.z));
    if (r0 < options.min_range) continue;
    if (r0 > options.max_range) continue;
    TsdfInsertHit(options, hit, origin, tsdf);
  }
}

// ---------------------------------------------------------------------------
// Precomputation grids.

// precomputation_grid_3d.cc:49-61
inline PrecomputationGrid3D ConvertToPrecomputationGrid(const HybridGrid& grid) {
  PrecomputationGrid3D result(grid.resolution());
  grid.ForEach([&result](const Array3i& index, const uint16 value) {
    const int cell_value = RoundToInt((ValueToProbability(value) - kMinProbability) *
                                      (255.f / (kMaxProbability - kMinProbability)));
    *result.mutable_value(index) = static_cast<uint8>(cell_value);
  });
  return result;
}

// precomputation_grid_3d.cc:62-77
inline PrecomputationGrid3D ConvertToPrecomputationGrid(const HybridGridTSDF& grid) {
  PrecomputationGrid3D result(grid.resolution());
  grid.ForEach([&result, &grid](const Array3i& index, const TSDFVoxel& voxel) {
    const float tsd = grid.ValueConverter().ValueToTSD(voxel.discrete_tsd);
    const float normalized_tsd = 1.f - std::abs(tsd) / grid.ValueConverter().getMaxTSD();
    const int cell_value = RoundToInt(normalized_tsd * 255.f);
    *result.mutable_value(index) = static_cast<uint8>(cell_value);
  });
  return result;
}

// precomputation_grid_3d.cc:79-97 (scatter-max into 8 cells; optional >>1 fold)
inline PrecomputationGrid3D PrecomputeGrid(const PrecomputationGrid3D& grid,
                                           const bool half_resolution, const Array3i& shift) {
  PrecomputationGrid3D result(grid.resolution());
  grid.ForEach([&](const Array3i& index, const uint8 value) {
    for (int i = 0; i != 8; ++i) {
      Array3i cell(index.x - shift.x * ((i & 1) ? 1 : 0), index.y - shift.y * ((i & 2) ? 1 : 0),
                   index.z - shift.z * ((i & 4) ? 1 : 0));
      if (half_resolution) cell = Array3i(cell.x >> 1, cell.y >> 1, cell.z >> 1);
      uint8* const dst = result.mutable_value(cell);
      *dst = std::max(value, *dst);
    }
  });
  return result;
}

struct FastCorrelativeScanMatcherOptions3D {
  int branch_and_bound_depth = 8;
  int full_resolution_depth = 3;
  bool use_rotational_scan_matcher = true;
  double min_rotational_score = 0.77;
  double min_low_resolution_score = 0.55;
  double linear_xy_search_window = 5.;
  double linear_z_search_window = 1.;
  double angular_search_window = 0.2617993877991494;  // 15 degrees
};

// fast_correlative_scan_matcher_3d.cc:60-94
class PrecomputationGridStack3D {
 public:
  PrecomputationGridStack3D(const GridInterface& grid,
                            const FastCorrelativeScanMatcherOptions3D& options) {
    grids_.reserve(options.branch_and_bound_depth);
    if (grid.GetGridType() == GridType::PROBABILITY_GRID) {
      grids_.push_back(ConvertToPrecomputationGrid(static_cast<const HybridGrid&>(grid)));
    } else {
      grids_.push_back(ConvertToPrecomputationGrid(static_cast<const HybridGridTSDF&>(grid)));
    }
    int last_width = 1;
    for (int depth = 1; depth != options.branch_and_bound_depth; ++depth) {
      const bool half_resolution = depth >= options.full_resolution_depth;
      const int next_width = 1 << depth;
      const int full_voxels_per_high_resolution_voxel =
          1 << std::max(0, depth - options.full_resolution_depth);
      const int shift = (next_width - last_width + (full_voxels_per_high_resolution_voxel - 1)) /
                        full_voxels_per_high_resolution_voxel;
      grids_.push_back(PrecomputeGrid(grids_.back(), half_resolution, Array3i(shift, shift, shift)));
      last_width = next_width;
    }
  }
  const PrecomputationGrid3D& Get(int depth) const { return grids_.at(depth); }
  int max_depth() const { return static_cast<int>(grids_.size()) - 1; }

 private:
  std::vector<PrecomputationGrid3D> grids_;
};

// ---------------------------------------------------------------------------
// Rotational scan matcher (Match / RotateHistogram / MatchHistograms only; the
// histogram itself is an input of the hot path).

inline std::vector<float> RotateHistogram(const std::vector<float>& histogram, const float angle) {
  const long size = static_cast<long>(histogram.size());
  // float * long -> float, / M_PI -> double, narrowed to float.
  const float rotate_by_buckets = static_cast<float>((-angle * size) / M_PI);
  int full_buckets = RoundToInt(rotate_by_buckets - 0.5f);
  const float fraction = rotate_by_buckets - full_buckets;
  while (full_buckets < 0) full_buckets += static_cast<int>(size);
  std::vector<float> out(histogram.size());
  for (long i = 0; i != size; ++i) {
    const float h0 = histogram[(i + full_buckets) % size];
    const float h1 = histogram[(i + 1 + full_buckets) % size];
    out[i] = fraction * h1 + (1.f - fraction) * h0;
  }
  return out;
}

// Eigen's sum of n floats term(0) .. term(n-1) over a VectorXf expression
// (Eigen 3.3 Core/Redux.h, restated from memory — Eigen is not vendored in the reference):
//   flavor 1, the SSE2 build an x86-64 reference runs: redux_impl<LinearVectorizedTraversal,
//     NoUnrolling> — packets of 4 floats from the (16-byte aligned) start, two packet accumulators
//     taking alternate packets, accumulator 0 + accumulator 1, one more packet if an odd number of
//     packets, then predux (a0 + a2) + (a1 + a3) (arch/SSE/PacketMath.h, SSE2 version), then the
//     scalar tail;
//   flavor 0, the generic scalar code: left to right.
template <typename Term>
inline float EigenSum(const long n, const Term& term) {
  if (FloatFlavor::Get() != 1 || n < 4) {
    float res = term(0);
    for (long i = 1; i < n; ++i) res = res + term(i);
    return res;
  }
  const long aligned_end2 = (n / 8) * 8, aligned_end = (n / 4) * 4;
  float p0[4], p1[4];
  for (int k = 0; k < 4; ++k) p0[k] = term(k);
  if (aligned_end > 4) {
    for (int k = 0; k < 4; ++k) p1[k] = term(4 + k);
    for (long index = 8; index < aligned_end2; index += 8) {
      for (int k = 0; k < 4; ++k) p0[k] = p0[k] + term(index + k);
      for (int k = 0; k < 4; ++k) p1[k] = p1[k] + term(index + 4 + k);
    }
    for (int k = 0; k < 4; ++k) p0[k] = p0[k] + p1[k];
    if (aligned_end > aligned_end2) {
      for (int k = 0; k < 4; ++k) p0[k] = p0[k] + term(aligned_end2 + k);
    }
  }
  float res = (p0[0] + p0[2]) + (p0[1] + p0[3]);
  for (long i = aligned_end; i < n; ++i) res = res + term(i);
  return res;
}

// rotational_scan_matcher.cc:121-132: VectorXf::norm() = sqrt(sum of squares), VectorXf::dot().
inline float MatchHistograms(const std::vector<float>& submap, const std::vector<float>& scan) {
  const long n = static_cast<long>(scan.size());
  if (n == 0) return 1.f;
  const float scan_norm = std::sqrt(EigenSum(n, [&](long i) { return scan[i] * scan[i]; }));
  const float submap_norm = std::sqrt(EigenSum(n, [&](long i) { return submap[i] * submap[i]; }));
  const float normalization = scan_norm * submap_norm;
  if (normalization < 1e-3f) return 1.f;
  return EigenSum(n, [&](long i) { return submap[i] * scan[i]; }) / normalization;
}

inline std::vector<float> RotationalMatch(const std::vector<float>& submap_histogram,
                                          const std::vector<float>& scan_histogram,
                                          const float initial_angle,
                                          const std::vector<float>& angles) {
  std::vector<float> result;
  result.reserve(angles.size());
  for (const float angle : angles) {
    result.push_back(
        MatchHistograms(submap_histogram, RotateHistogram(scan_histogram, initial_angle + angle)));
  }
  return result;
}

// ---------------------------------------------------------------------------
// Low resolution matcher (leaf gate).

inline float LowResolutionScore(const GridInterface* low_resolution_grid,
                                const std::vector<Vector3f>& points, const Rigid3f& pose) {
  float score = 0.f;
  if (low_resolution_grid->GetGridType() == GridType::PROBABILITY_GRID) {
    const HybridGrid* g = static_cast<const HybridGrid*>(low_resolution_grid);
    for (const Vector3f& p : points) {
      const Vec3f w = Apply(pose, Vec3f(p.x, p.y, p.z));
      score += g->GetProbability(g->GetCellIndex(Vector3f(w.x, w.y, w.z)));
    }
  } else {
    const HybridGridTSDF* g = static_cast<const HybridGridTSDF*>(low_resolution_grid);
    for (const Vector3f& p : points) {
      const Vec3f w = Apply(pose, Vec3f(p.x, p.y, p.z));
      score += 1.f - std::abs(g->GetTSD(g->GetCellIndex(Vector3f(w.x, w.y, w.z))) /
                              g->ValueConverter().getMaxTSD());
    }
  }
  return score / points.size();
}

// ---------------------------------------------------------------------------
// Fast correlative scan matcher.

struct DiscreteScan3D {
  Rigid3f pose;
  std::vector<std::vector<Array3i>> cell_indices_per_depth;
  float rotational_score;
};

struct Candidate3D {
  Candidate3D(const int scan_index_, const Array3i& offset_)
      : scan_index(scan_index_), offset(offset_) {}
  static Candidate3D Unsuccessful() { return Candidate3D(0, Array3i(0, 0, 0)); }
  int scan_index;
  Array3i offset;
  float score = -std::numeric_limits<float>::infinity();
  float low_resolution_score = 0.f;
  int sum = 0;  // oracle-only: the integer gather sum behind 'score'
  bool operator<(const Candidate3D& other) const { return score < other.score; }
  bool operator>(const Candidate3D& other) const { return score > other.score; }
};

struct NodeData {
  std::vector<Vector3f> high_resolution_point_cloud;
  std::vector<Vector3f> low_resolution_point_cloud;
  std::vector<float> rotational_scan_matcher_histogram;
  double gravity_alignment[4] = {1., 0., 0., 0.};  // (w, x, y, z)
};

struct MatchResult {
  bool found = false;
  float score = 0.f;
  double pose_t[3] = {0, 0, 0};
  double pose_q[4] = {1, 0, 0, 0};  // (w, x, y, z)
  float rotational_score = 0.f;
  float low_resolution_score = 0.f;
  // oracle-only diagnostics
  int best_scan_index = 0;
  Array3i best_offset;
  long long num_candidates_scored = 0;
  int num_discrete_scans = 0;
};

class FastCorrelativeScanMatcher3D {
 public:
  FastCorrelativeScanMatcher3D(const GridInterface& grid, const GridInterface* low_resolution_grid,
                               const std::vector<float>* rotational_scan_matcher_histogram,
                               const FastCorrelativeScanMatcherOptions3D& options)
      : options_(options),
        stack_(new PrecomputationGridStack3D(grid, options)),
        low_resolution_grid_(low_resolution_grid),
        submap_histogram_(rotational_scan_matcher_histogram) {
    if (grid.GetGridType() == GridType::PROBABILITY_GRID) {
      resolution_ = static_cast<const HybridGrid&>(grid).resolution();
      width_in_voxels_ = static_cast<const HybridGrid&>(grid).grid_size();
    } else {
      resolution_ = static_cast<const HybridGridTSDF&>(grid).resolution();
      width_in_voxels_ = static_cast<const HybridGridTSDF&>(grid).grid_size();
    }
  }

  struct SearchParameters {
    int linear_xy_window_size;
    int linear_z_window_size;
    double angular_search_window;
  };

  const PrecomputationGridStack3D& stack() const { return *stack_; }
  int width_in_voxels() const { return width_in_voxels_; }

  // :157-174. Poses are (t[3], q[4] = w,x,y,z) doubles, cast to float.
  MatchResult Match(const double* node_t, const double* node_q, const double* submap_t,
                    const double* submap_q, const NodeData& data, const float min_score) const {
    const SearchParameters sp{
        RoundToInt(options_.linear_xy_search_window / resolution_),
        RoundToInt(options_.linear_z_search_window / resolution_),
        options_.angular_search_window};
    return MatchWithSearchParameters(sp, ToRigidf(node_t, node_q), ToRigidf(submap_t, submap_q),
                                     data, min_score);
  }

  // :176-200
  MatchResult MatchFullSubmap(const double* node_q, const double* submap_q, const NodeData& data,
                              const float min_score) const {
    float max_point_distance = 0.f;
    for (const Vector3f& p : data.high_resolution_point_cloud) {
      max_point_distance = std::max(max_point_distance, Norm3f(Vec3f(p.x, p.y, p.z)));
    }
    const int linear_window_size =
        (width_in_voxels_ + 1) / 2 + RoundToInt(max_point_distance / resolution_ + 0.5f);
    const SearchParameters sp{linear_window_size, linear_window_size, M_PI};
    const double zero[3] = {0., 0., 0.};
    return MatchWithSearchParameters(sp, ToRigidf(zero, node_q), ToRigidf(zero, submap_q), data,
                                     min_score);
  }

  // Exposed for parity tests of the discretisation / scoring stages.
  std::vector<DiscreteScan3D> GenerateDiscreteScans(const SearchParameters& sp,
                                                    const NodeData& data,
                                                    const Rigid3f& global_node_pose,
                                                    const Rigid3f& global_submap_pose) const {
    std::vector<DiscreteScan3D> result;
    const std::vector<Vector3f>& cloud = data.high_resolution_point_cloud;
    float max_scan_range = 3.f * resolution_;
    for (const Vector3f& p : cloud) {
      const float range = Norm3f(Vec3f(p.x, p.y, p.z));
      max_scan_range = std::max(range, max_scan_range);
    }
    const float kSafetyMargin = 1.f - 1e-2f;
    const float angular_step_size =
        kSafetyMargin * std::acos(1.f - (resolution_ * resolution_) /
                                            (2.f * (max_scan_range * max_scan_range)));
    // double window / float step -> double -> lround
    const int angular_window_size = RoundToInt(sp.angular_search_window / angular_step_size);
    std::vector<float> angles;
    for (int rz = -angular_window_size; rz <= angular_window_size; ++rz) {
      angles.push_back(rz * angular_step_size);
    }
    const Rigid3f node_to_submap =
        RigidProductFloat(RigidInverseFloat(global_submap_pose), global_node_pose);
    std::vector<float> scores;
    if (options_.use_rotational_scan_matcher) {
      const Quatf gravity(static_cast<float>(data.gravity_alignment[0]),
                          static_cast<float>(data.gravity_alignment[1]),
                          static_cast<float>(data.gravity_alignment[2]),
                          static_cast<float>(data.gravity_alignment[3]));
      // gravity_alignment.inverse().cast<float>(): inverse in double, then cast.
      const double* ga = data.gravity_alignment;
      const double n2 = (ga[1] * ga[1] + ga[2] * ga[2]) + (ga[3] * ga[3] + ga[0] * ga[0]);
      const Quatf gravity_inverse(static_cast<float>(ga[0] / n2), static_cast<float>(-ga[1] / n2),
                                  static_cast<float>(-ga[2] / n2), static_cast<float>(-ga[3] / n2));
      (void)gravity;
      scores = RotationalMatch(
          *submap_histogram_, data.rotational_scan_matcher_histogram,
          GetYawFloat(QuatProductFloat(node_to_submap.rotation, gravity_inverse)), angles);
    }
    for (size_t i = 0; i != angles.size(); ++i) {
      if (options_.use_rotational_scan_matcher &&
          static_cast<double>(scores[i]) < options_.min_rotational_score) {
        continue;
      }
      const Vec3f angle_axis(0.f, 0.f, angles[i]);
      const Rigid3f pose(
          node_to_submap.translation,
          QuatProductFloat(QuatProductFloat(QuatInverseFloat(global_submap_pose.rotation),
                                            AngleAxisVectorToRotationQuaternionFloat(angle_axis)),
                           global_node_pose.rotation));
      const float score = options_.use_rotational_scan_matcher ? scores[i] : 1.0f;
      result.push_back(DiscretizeScan(sp, cloud, pose, score));
    }
    return result;
  }

  // :363-386
  void ScoreCandidates(const int depth, const std::vector<DiscreteScan3D>& discrete_scans,
                       std::vector<Candidate3D>* const candidates, const bool sort = true) const {
    const int reduction_exponent = std::max(0, depth - options_.full_resolution_depth + 1);
    const PrecomputationGrid3D& grid = stack_->Get(depth);
    for (Candidate3D& candidate : *candidates) {
      int sum = 0;
      const DiscreteScan3D& scan = discrete_scans[candidate.scan_index];
      const Array3i offset(candidate.offset.x >> reduction_exponent,
                           candidate.offset.y >> reduction_exponent,
                           candidate.offset.z >> reduction_exponent);
      for (const Array3i& cell : scan.cell_indices_per_depth[depth]) {
        sum += grid.value(cell + offset);
      }
      candidate.sum = sum;
      candidate.score = PrecomputationGrid3D::ToProbability(
          sum / static_cast<float>(scan.cell_indices_per_depth[depth].size()));
    }
    num_scored_ += static_cast<long long>(candidates->size());
    if (sort) std::sort(candidates->begin(), candidates->end(), std::greater<Candidate3D>());
  }

  // :328-361
  std::vector<Candidate3D> GenerateLowestResolutionCandidates(const SearchParameters& sp,
                                                              const int num_discrete_scans) const {
    const int linear_step_size = 1 << stack_->max_depth();
    std::vector<Candidate3D> candidates;
    for (int scan_index = 0; scan_index != num_discrete_scans; ++scan_index) {
      for (int z = -sp.linear_z_window_size; z <= sp.linear_z_window_size; z += linear_step_size) {
        for (int y = -sp.linear_xy_window_size; y <= sp.linear_xy_window_size;
             y += linear_step_size) {
          for (int x = -sp.linear_xy_window_size; x <= sp.linear_xy_window_size;
               x += linear_step_size) {
            candidates.emplace_back(scan_index, Array3i(x, y, z));
          }
        }
      }
    }
    return candidates;
  }

  // :202-228
  MatchResult MatchWithSearchParameters(const SearchParameters& sp, const Rigid3f& node_pose,
                                        const Rigid3f& submap_pose, const NodeData& data,
                                        const float min_score) const {
    num_scored_ = 0;
    const std::vector<DiscreteScan3D> discrete_scans =
        GenerateDiscreteScans(sp, data, node_pose, submap_pose);
    std::vector<Candidate3D> lowest =
        GenerateLowestResolutionCandidates(sp, static_cast<int>(discrete_scans.size()));
    ScoreCandidates(stack_->max_depth(), discrete_scans, &lowest);
    const Candidate3D best =
        BranchAndBound(sp, data, discrete_scans, lowest, stack_->max_depth(), min_score);
    MatchResult r;
    r.num_candidates_scored = num_scored_;
    r.num_discrete_scans = static_cast<int>(discrete_scans.size());
    if (best.score > min_score) {
      r.found = true;
      r.score = best.score;
      const Rigid3f pose = GetPoseFromCandidate(discrete_scans, best);
      r.pose_t[0] = pose.translation.x;
      r.pose_t[1] = pose.translation.y;
      r.pose_t[2] = pose.translation.z;
      r.pose_q[0] = pose.rotation.w;
      r.pose_q[1] = pose.rotation.x;
      r.pose_q[2] = pose.rotation.y;
      r.pose_q[3] = pose.rotation.z;
      r.rotational_score = discrete_scans[best.scan_index].rotational_score;
      r.low_resolution_score = best.low_resolution_score;
      r.best_scan_index = best.scan_index;
      r.best_offset = best.offset;
    }
    return r;
  }

  SearchParameters LocalSearchParameters() const {
    return SearchParameters{
        RoundToInt(options_.linear_xy_search_window / resolution_),
        RoundToInt(options_.linear_z_search_window / resolution_),
        options_.angular_search_window};
  }

  static Rigid3f ToRigidf(const double* t, const double* q) {
    return Rigid3f(Vec3f(static_cast<float>(t[0]), static_cast<float>(t[1]), static_cast<float>(t[2])),
                   Quatf(static_cast<float>(q[0]), static_cast<float>(q[1]),
                         static_cast<float>(q[2]), static_cast<float>(q[3])));
  }

 private:
  // :230-274
  DiscreteScan3D DiscretizeScan(const SearchParameters& sp, const std::vector<Vector3f>& cloud,
                                const Rigid3f& pose, const float rotational_score) const {
    std::vector<std::vector<Array3i>> per_depth;
    const PrecomputationGrid3D& original = stack_->Get(0);
    std::vector<Array3i> full;
    full.reserve(cloud.size());
    for (const Vector3f& p : cloud) {
      const Vec3f w = Apply(pose, Vec3f(p.x, p.y, p.z));
      full.push_back(original.GetCellIndex(Vector3f(w.x, w.y, w.z)));
    }
    const int full_resolution_depth =
        std::min(options_.full_resolution_depth, options_.branch_and_bound_depth);
    for (int i = 0; i != full_resolution_depth; ++i) per_depth.push_back(full);
    const int low_resolution_depth = options_.branch_and_bound_depth - full_resolution_depth;
    const Array3i start(-sp.linear_xy_window_size, -sp.linear_xy_window_size,
                        -sp.linear_z_window_size);
    for (int i = 0; i != low_resolution_depth; ++i) {
      const int e = i + 1;
      const Array3i low_start(start.x >> e, start.y >> e, start.z >> e);
      per_depth.emplace_back();
      per_depth.back().reserve(full.size());
      for (const Array3i& c : full) {
        const Array3i at_start = c + start;
        per_depth.back().push_back(Array3i((at_start.x >> e) - low_start.x,
                                           (at_start.y >> e) - low_start.y,
                                           (at_start.z >> e) - low_start.z));
      }
    }
    return DiscreteScan3D{pose, per_depth, rotational_score};
  }

  // :400-406
  Rigid3f GetPoseFromCandidate(const std::vector<DiscreteScan3D>& scans,
                               const Candidate3D& c) const {
    const Rigid3f translation(Vec3f(resolution_ * static_cast<float>(c.offset.x),
                                    resolution_ * static_cast<float>(c.offset.y),
                                    resolution_ * static_cast<float>(c.offset.z)),
                              Quatf(1.f, 0.f, 0.f, 0.f));
    return RigidProductFloat(translation, scans[c.scan_index].pose);
  }

  // :408-471
  Candidate3D BranchAndBound(const SearchParameters& sp, const NodeData& data,
                             const std::vector<DiscreteScan3D>& scans,
                             const std::vector<Candidate3D>& candidates, const int candidate_depth,
                             float min_score) const {
    if (candidate_depth == 0) {
      for (const Candidate3D& candidate : candidates) {
        if (candidate.score <= min_score) return Candidate3D::Unsuccessful();
        const float low_resolution_score =
            LowResolutionScore(low_resolution_grid_, data.low_resolution_point_cloud,
                               GetPoseFromCandidate(scans, candidate));
        if (static_cast<double>(low_resolution_score) >= options_.min_low_resolution_score) {
          Candidate3D best = candidate;
          best.low_resolution_score = low_resolution_score;
          return best;
        }
      }
      return Candidate3D::Unsuccessful();
    }
    Candidate3D best = Candidate3D::Unsuccessful();
    best.score = min_score;
    for (const Candidate3D& candidate : candidates) {
      if (candidate.score <= min_score) break;
      std::vector<Candidate3D> children;
      const int half_width = 1 << (candidate_depth - 1);
      for (int z : {0, half_width}) {
        if (candidate.offset.z + z > sp.linear_z_window_size) break;
        for (int y : {0, half_width}) {
          if (candidate.offset.y + y > sp.linear_xy_window_size) break;
          for (int x : {0, half_width}) {
            if (candidate.offset.x + x > sp.linear_xy_window_size) break;
            children.emplace_back(candidate.scan_index,
                                  Array3i(candidate.offset.x + x, candidate.offset.y + y,
                                          candidate.offset.z + z));
          }
        }
      }
      ScoreCandidates(candidate_depth - 1, scans, &children);
      best = std::max(best, BranchAndBound(sp, data, scans, children, candidate_depth - 1,
                                           best.score));
    }
    return best;
  }

  const FastCorrelativeScanMatcherOptions3D options_;
  float resolution_;
  int width_in_voxels_;
  std::unique_ptr<PrecomputationGridStack3D> stack_;
  const GridInterface* low_resolution_grid_;
  const std::vector<float>* submap_histogram_;
  mutable long long num_scored_ = 0;
};

}  // namespace oracle
