// hg_adapters.h — C++ host adapters that present the reference's own class surfaces over
// the C-ABI of include/hg_b200.h. Header-only, C++11, no third-party dependencies.
//
// A maintainer of the reference swaps these in where the CPU classes are constructed
// (see INTEGRATION.md); everything above them (CeresScanMatcher3D's problem set-up,
// OptimizingLocalTrajectoryBuilder, PoseGraph3D) stays as it is.
//
// Mirrors (paths under /root/reference/cartographer):
//   ceres::CostFunction                       third-party Ceres 1.13, cost_function.h
//   XxxSpaceCostFunction3D::CreateAutoDiffCostFunction
//       mapping/internal/3d/scan_matching/*_cost_function_3d.h  (8 factories)
//   CeresScanMatcher3D::Match                 mapping/internal/3d/scan_matching/ceres_scan_matcher_3d.h:52-71
//   FastCorrelativeScanMatcher3D              mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.h:68-100
//   ConstraintBuilder3D                       mapping/internal/constraints/constraint_builder_3d.h:51-114
//
// Error behaviour: the reference CHECK-fails (aborts) on invalid arguments; these adapters
// print the library's error text and abort() likewise. Evaluate() returns false on failure,
// which is how a ceres::CostFunction reports it.
#pragma once

#include <array>
#include <cmath>
#include <cstdint>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../../include/hg_b200.h"

#if defined(HG_USE_REAL_CERES)
#include "ceres/cost_function.h"
namespace hg {
using CostFunctionBase = ::ceres::CostFunction;
}
#else
namespace hg {
namespace ceres_compat {
// Same virtual interface and protected setters as ceres::CostFunction (Ceres 1.13
// include/ceres/cost_function.h), for builds without Ceres.
class CostFunction {
 public:
  CostFunction() : num_residuals_(0) {}
  virtual ~CostFunction() {}
  virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
  const std::vector<int>& parameter_block_sizes() const { return parameter_block_sizes_; }
  int num_residuals() const { return num_residuals_; }

 protected:
  std::vector<int>* mutable_parameter_block_sizes() { return &parameter_block_sizes_; }
  void set_num_residuals(int n) { num_residuals_ = n; }

 private:
  std::vector<int> parameter_block_sizes_;
  int num_residuals_;
};
}  // namespace ceres_compat
using CostFunctionBase = ceres_compat::CostFunction;
}  // namespace hg
#endif

namespace hg {

inline void CheckOk(int rc, const char* what) {
  if (rc != HG_OK) {
    std::fprintf(stderr, "[hg_b200] %s failed (%d): %s\n", what, rc, hg_last_error());
    std::abort();  // the reference CHECK-fails
  }
}

// transform::Rigid3d without Eigen: translation + rotation (w, x, y, z).
struct Rigid3d {
  std::array<double, 3> translation{{0., 0., 0.}};
  std::array<double, 4> rotation{{1., 0., 0., 0.}};
  void ToArray(double* out7) const {
    for (int i = 0; i < 3; ++i) out7[i] = translation[i];
    for (int i = 0; i < 4; ++i) out7[3 + i] = rotation[i];
  }
};

// sensor::PointCloud / sensor::TimedPointCloud as a view: 'stride' floats per point.
struct PointCloudView {
  const float* data = nullptr;
  int size = 0;
  int stride = 3;
};

// Owns a device copy of a HybridGrid / HybridGridTSDF.
class DeviceGrid {
 public:
  // From the content of proto::HybridGrid (mapping/proto/3d/hybrid_grid.proto:19-29).
  static std::shared_ptr<DeviceGrid> FromHybridGrid(float resolution, int grid_size, int64_t n,
                                                    const int32_t* xyz, const uint16_t* values) {
    hg_grid* g = nullptr;
    CheckOk(hg_grid_upload_prob(resolution, grid_size, n, xyz, values, &g), "hg_grid_upload_prob");
    return std::shared_ptr<DeviceGrid>(new DeviceGrid(g));
  }
  // From the content of proto::HybridGridTSDF (hybrid_grid_tsdf.proto:19-32).
  static std::shared_ptr<DeviceGrid> FromHybridGridTSDF(float resolution, float relative_truncation_distance,
                                                        float max_weight, int grid_size, int64_t n,
                                                        const int32_t* xyz, const uint16_t* tsd,
                                                        const uint16_t* weight) {
    hg_grid* g = nullptr;
    CheckOk(hg_grid_upload_tsdf(resolution, relative_truncation_distance, max_weight, grid_size, n, xyz, tsd,
                                weight, &g),
            "hg_grid_upload_tsdf");
    return std::shared_ptr<DeviceGrid>(new DeviceGrid(g));
  }
  // HybridGrid(const proto::HybridGrid&) / HybridGridTSDF(const proto::HybridGridTSDF&) from the
  // serialized message (proto.SerializeAsString()); hybrid_grid.h:61-73, hybrid_grid_tsdf.h:69-83.
  static std::shared_ptr<DeviceGrid> FromProto(const std::string& serialized, bool tsdf) {
    hg_grid* g = nullptr;
    CheckOk(hg_grid_from_proto(serialized.data(), static_cast<int64_t>(serialized.size()), tsdf ? 1 : 0, &g),
            "hg_grid_from_proto");
    return std::shared_ptr<DeviceGrid>(new DeviceGrid(g));
  }
  // ToProto().SerializeAsString() (hybrid_grid.h:88-99, hybrid_grid_tsdf.h:119-134).
  std::string ToProto() const {
    int64_t size = 0;
    CheckOk(hg_grid_to_proto(grid_, nullptr, 0, &size), "hg_grid_to_proto");
    std::string out(static_cast<size_t>(size), '\0');
    if (size > 0) CheckOk(hg_grid_to_proto(grid_, &out[0], size, &size), "hg_grid_to_proto");
    return out;
  }
  ~DeviceGrid() { hg_grid_destroy(grid_); }
  const hg_grid* get() const { return grid_; }
  hg_grid* mutable_get() { return grid_; }  // for the range-data inserters

 private:
  explicit DeviceGrid(hg_grid* g) : grid_(g) {}
  DeviceGrid(const DeviceGrid&) = delete;
  DeviceGrid& operator=(const DeviceGrid&) = delete;
  hg_grid* grid_;
};

// A ceres::CostFunction whose Evaluate runs on the GPU.
class GpuScanCostFunction : public CostFunctionBase {
 public:
  ~GpuScanCostFunction() override { hg_cost_destroy(cost_); }
  bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const override {
    return hg_cost_evaluate(cost_, parameters, residuals, jacobians) == HG_OK;
  }
  const hg_cost* handle() const { return cost_; }

 private:
  friend CostFunctionBase* CreateGpuCostFunction(int, double, const PointCloudView&,
                                                 const std::vector<const hg_grid*>&, double);
  explicit GpuScanCostFunction(hg_cost* c) : cost_(c) {
    set_num_residuals(hg_cost_num_residuals(c));
    const int blocks = hg_cost_num_parameter_blocks(c);
    for (int b = 0; b < blocks; ++b) mutable_parameter_block_sizes()->push_back(b % 2 == 0 ? 3 : 4);
  }
  hg_cost* cost_;
};

inline CostFunctionBase* CreateGpuCostFunction(int kind, double scaling_factor, const PointCloudView& cloud,
                                               const std::vector<const hg_grid*>& grids,
                                               double interpolation_ratio) {
  hg_cost* c = nullptr;
  CheckOk(hg_cost_create(kind, scaling_factor, cloud.data, cloud.stride, cloud.size, grids.data(),
                         static_cast<int>(grids.size()), interpolation_ratio, &c),
          "hg_cost_create");
  return new GpuScanCostFunction(c);
}

// The eight factories, named and argued like the reference's (the caller — Ceres — owns the result).
struct OccupiedSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const DeviceGrid& hybrid_grid) {
    return CreateGpuCostFunction(HG_COST_OCCUPIED_SPACE, scaling_factor, point_cloud, {hybrid_grid.get()}, 0.);
  }
};
struct InterpolatedOccupiedSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const DeviceGrid& hybrid_grid, double interpolation_ratio) {
    return CreateGpuCostFunction(HG_COST_INTERPOLATED_OCCUPIED_SPACE, scaling_factor, point_cloud,
                                 {hybrid_grid.get()}, interpolation_ratio);
  }
};
struct TSDFSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const DeviceGrid& tsdf) {
    return CreateGpuCostFunction(HG_COST_TSDF_SPACE, scaling_factor, point_cloud, {tsdf.get()}, 0.);
  }
};
struct InterpolatedTSDFSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const DeviceGrid& hybrid_grid, double interpolation_ratio) {
    return CreateGpuCostFunction(HG_COST_INTERPOLATED_TSDF_SPACE, scaling_factor, point_cloud,
                                 {hybrid_grid.get()}, interpolation_ratio);
  }
};
struct InterpolatedTSDFPerPointSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const float* point_xyz,
                                                      const DeviceGrid& hybrid_grid, double interpolation_ratio) {
    PointCloudView v;
    v.data = point_xyz;
    v.size = 1;
    return CreateGpuCostFunction(HG_COST_INTERPOLATED_TSDF_PER_POINT_SPACE, scaling_factor, v,
                                 {hybrid_grid.get()}, interpolation_ratio);
  }
};
inline std::vector<const hg_grid*> PyramidHandles(const std::vector<const DeviceGrid*>& pyramid) {
  std::vector<const hg_grid*> g;
  for (const DeviceGrid* d : pyramid) g.push_back(d->get());
  return g;
}
struct MultiResolutionTSDFSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const std::vector<const DeviceGrid*>& tsdf) {
    return CreateGpuCostFunction(HG_COST_MULTI_RESOLUTION_TSDF_SPACE, scaling_factor, point_cloud,
                                 PyramidHandles(tsdf), 0.);
  }
};
struct InterpolatedMultiResolutionTSDFSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const PointCloudView& point_cloud,
                                                      const std::vector<const DeviceGrid*>& hybrid_grid_pyramid,
                                                      double interpolation_ratio) {
    return CreateGpuCostFunction(HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_SPACE, scaling_factor, point_cloud,
                                 PyramidHandles(hybrid_grid_pyramid), interpolation_ratio);
  }
};
struct InterpolatedMultiResolutionTSDFPerPointSpaceCostFunction3D {
  static CostFunctionBase* CreateAutoDiffCostFunction(double scaling_factor, const float* point_xyz,
                                                      const std::vector<const DeviceGrid*>& tsdf_pyramid,
                                                      double interpolation_ratio) {
    PointCloudView v;
    v.data = point_xyz;
    v.size = 1;
    return CreateGpuCostFunction(HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_PER_POINT_SPACE, scaling_factor, v,
                                 PyramidHandles(tsdf_pyramid), interpolation_ratio);
  }
};

// sensor::VoxelFilter (sensor/internal/voxel_filter.h:30-60): Filter returns the first point of
// every voxel, in input order. Works on any point type that starts with float xyz.
class VoxelFilter {
 public:
  explicit VoxelFilter(float size) : size_(size) {}
  // Indices of the kept points.
  std::vector<int32_t> FilterIndices(const PointCloudView& cloud) const {
    std::vector<int32_t> kept(static_cast<size_t>(cloud.size));
    int64_t n = 0;
    CheckOk(hg_voxel_filter(size_, cloud.data, cloud.stride, cloud.size, &n, kept.data()), "hg_voxel_filter");
    kept.resize(static_cast<size_t>(n));
    return kept;
  }
  // The filtered cloud itself (same stride as the input).
  std::vector<float> Filter(const PointCloudView& cloud) const { return Gather(cloud, FilterIndices(cloud)); }
  static std::vector<float> Gather(const PointCloudView& cloud, const std::vector<int32_t>& kept) {
    std::vector<float> out;
    out.reserve(kept.size() * cloud.stride);
    for (const int32_t i : kept)
      out.insert(out.end(), cloud.data + static_cast<size_t>(i) * cloud.stride,
                 cloud.data + static_cast<size_t>(i + 1) * cloud.stride);
    return out;
  }

 private:
  float size_;
};

// sensor::AdaptiveVoxelFilter (sensor/internal/adaptive_voxel_filter.h:88-111);
// options = proto::AdaptiveVoxelFilterOptions {max_length, min_num_points, max_range}.
class AdaptiveVoxelFilter {
 public:
  AdaptiveVoxelFilter(float max_length, int min_num_points, float max_range)
      : max_length_(max_length), min_num_points_(min_num_points), max_range_(max_range) {}
  std::vector<float> Filter(const PointCloudView& cloud) const {
    std::vector<int32_t> kept(static_cast<size_t>(cloud.size));
    int64_t n = 0;
    CheckOk(hg_adaptive_voxel_filter(max_length_, min_num_points_, max_range_, cloud.data, cloud.stride, cloud.size,
                                     &n, kept.data()),
            "hg_adaptive_voxel_filter");
    kept.resize(static_cast<size_t>(n));
    return VoxelFilter::Gather(cloud, kept);
  }

 private:
  float max_length_;
  int min_num_points_;
  float max_range_;
};

// sensor::RangeData (sensor/range_data.h:45-57) as a view: origin + returns ('misses' are not
// read by the 3D inserters).
struct RangeDataView {
  std::array<float, 3> origin{{0.f, 0.f, 0.f}};
  PointCloudView returns;
};

// mapping::OccupancyGridRangeDataInserter3D (mapping/3d/range_data_inserter_3d.h:33-52); options =
// proto::ProbabilityGridRangeDataInserterOptions3D. Insert writes the device-resident grid in place.
class OccupancyGridRangeDataInserter3D {
 public:
  explicit OccupancyGridRangeDataInserter3D(const hg_occupancy_inserter_options& options) : options_(options) {}
  void Insert(const RangeDataView& range_data, DeviceGrid* grid) const {
    CheckOk(hg_grid_insert_occupancy(grid->mutable_get(), &options_, range_data.origin.data(),
                                     range_data.returns.data, range_data.returns.stride, range_data.returns.size),
            "hg_grid_insert_occupancy");
  }

 private:
  hg_occupancy_inserter_options options_;
};

// mapping::TSDFRangeDataInserter3D (mapping/3d/tsdf_range_data_inserter_3d.h:32-46) for
// project_sdf_distance_to_scan_normal = false; options = the fields of
// proto::TSDFRangeDataInserterOptions3D that branch reads.
class TSDFRangeDataInserter3D {
 public:
  explicit TSDFRangeDataInserter3D(const hg_tsdf_inserter_options& options) : options_(options) {}
  void Insert(const RangeDataView& range_data, DeviceGrid* grid) const {
    CheckOk(hg_grid_insert_tsdf(grid->mutable_get(), &options_, range_data.origin.data(), range_data.returns.data,
                                range_data.returns.stride, range_data.returns.size, nullptr),
            "hg_grid_insert_tsdf");
  }

 private:
  hg_tsdf_inserter_options options_;
};

// proto::CeresScanMatcherOptions3D (ceres_scan_matcher_3d.cc:40-62).
struct CeresScanMatcherOptions3D {
  std::vector<double> occupied_space_weight;
  double translation_weight = 0.;
  double rotation_weight = 0.;
  bool only_optimize_yaw = false;
  int max_num_iterations = 10;
  bool use_nonmonotonic_steps = false;
};

struct SolverSummary {
  double initial_cost = 0., final_cost = 0.;
  int num_iterations = 0, num_successful_steps = 0;
};

class CeresScanMatcher3D {
 public:
  using PointCloudAndHybridGridPointers = std::pair<PointCloudView, const DeviceGrid*>;
  explicit CeresScanMatcher3D(const CeresScanMatcherOptions3D& options) : options_(options) {}

  // ceres_scan_matcher_3d.cc:72-95
  void Match(const std::array<double, 3>& target_translation, const Rigid3d& initial_pose_estimate,
             const std::vector<PointCloudAndHybridGridPointers>& point_clouds_and_hybrid_grids,
             Rigid3d* pose_estimate, SolverSummary* summary) const {
    const int n = static_cast<int>(point_clouds_and_hybrid_grids.size());
    if (static_cast<int>(options_.occupied_space_weight.size()) != n) {
      std::fprintf(stderr, "[hg_b200] occupied_space_weight count != number of point clouds\n");
      std::abort();  // CHECK_EQ at ceres_scan_matcher_3d.cc:126-127
    }
    std::vector<const float*> clouds;
    std::vector<int32_t> sizes;
    std::vector<const hg_grid*> grids;
    for (const auto& p : point_clouds_and_hybrid_grids) {
      if (p.first.stride != 3) {
        std::fprintf(stderr, "[hg_b200] CeresScanMatcher3D::Match takes sensor::PointCloud (stride 3)\n");
        std::abort();
      }
      clouds.push_back(p.first.data);
      sizes.push_back(p.first.size);
      grids.push_back(p.second->get());
    }
    double init[7], out[7], sum[6];
    initial_pose_estimate.ToArray(init);
    CheckOk(hg_scan_match(n, clouds.data(), sizes.data(), grids.data(), options_.occupied_space_weight.data(),
                          options_.translation_weight, options_.rotation_weight, options_.only_optimize_yaw ? 1 : 0,
                          options_.max_num_iterations, options_.use_nonmonotonic_steps ? 1 : 0,
                          target_translation.data(), init, out, sum),
            "hg_scan_match");
    for (int i = 0; i < 3; ++i) pose_estimate->translation[i] = out[i];
    for (int i = 0; i < 4; ++i) pose_estimate->rotation[i] = out[3 + i];
    if (summary != nullptr) {
      summary->initial_cost = sum[0];
      summary->final_cost = sum[1];
      summary->num_iterations = static_cast<int>(sum[2]);
      summary->num_successful_steps = static_cast<int>(sum[3]);
    }
  }

  // ceres_scan_matcher_3d.cc:97-118: the assembled problem evaluated at the initial pose. The reference
  // hands its 'jacobians' vector to ceres::Problem::Evaluate as the GRADIENT argument: it receives the
  // gradient of the cost in the tangent space of the pose (3 translation + 3 rotation entries, or 3 + 1
  // with only_optimize_yaw), not a Jacobian. 'residuals': every cloud's residuals, then the translation
  // and rotation deltas (3 + 3).
  void Evaluate(const std::array<double, 3>& target_translation, const Rigid3d& initial_pose_estimate,
                const std::vector<PointCloudAndHybridGridPointers>& point_clouds_and_hybrid_grids, double* cost,
                std::vector<double>* residuals, std::vector<double>* jacobians) const {
    const int n = static_cast<int>(point_clouds_and_hybrid_grids.size());
    if (static_cast<int>(options_.occupied_space_weight.size()) != n) {
      std::fprintf(stderr, "[hg_b200] occupied_space_weight count != number of point clouds\n");
      std::abort();
    }
    std::vector<const float*> clouds;
    std::vector<int32_t> sizes;
    std::vector<const hg_grid*> grids;
    size_t num_residuals = 6;
    for (const auto& p : point_clouds_and_hybrid_grids) {
      clouds.push_back(p.first.data);
      sizes.push_back(p.first.size);
      grids.push_back(p.second->get());
      num_residuals += static_cast<size_t>(p.first.size);
    }
    double init[7], c = 0.;
    initial_pose_estimate.ToArray(init);
    std::vector<double> r(num_residuals), g(options_.only_optimize_yaw ? 4 : 6);
    CheckOk(hg_scan_match_evaluate(n, clouds.data(), sizes.data(), grids.data(), options_.occupied_space_weight.data(),
                                   options_.translation_weight, options_.rotation_weight,
                                   options_.only_optimize_yaw ? 1 : 0, target_translation.data(), init, &c, r.data(),
                                   g.data()),
            "hg_scan_match_evaluate");
    if (cost != nullptr) *cost = c;
    if (residuals != nullptr) *residuals = r;
    if (jacobians != nullptr) *jacobians = g;
  }

 private:
  const CeresScanMatcherOptions3D options_;
};

// mapping/trajectory_node.h:45-62 (the fields the matchers read).
struct TrajectoryNodeData {
  std::array<double, 4> gravity_alignment{{1., 0., 0., 0.}};
  std::vector<float> high_resolution_point_cloud;  // xyz
  std::vector<float> low_resolution_point_cloud;   // xyz
  std::vector<float> rotational_scan_matcher_histogram;
  hg_node_data View() const {
    hg_node_data d;
    d.high_resolution_point_cloud = high_resolution_point_cloud.data();
    d.num_high_resolution_points = static_cast<int32_t>(high_resolution_point_cloud.size() / 3);
    d.low_resolution_point_cloud = low_resolution_point_cloud.data();
    d.num_low_resolution_points = static_cast<int32_t>(low_resolution_point_cloud.size() / 3);
    d.rotational_scan_matcher_histogram = rotational_scan_matcher_histogram.data();
    d.histogram_size = static_cast<int32_t>(rotational_scan_matcher_histogram.size());
    for (int i = 0; i < 4; ++i) d.gravity_alignment[i] = gravity_alignment[i];
    return d;
  }
};

class FastCorrelativeScanMatcher3D {
 public:
  // fast_correlative_scan_matcher_3d.h:68-73
  struct Result {
    float score;
    Rigid3d pose_estimate;
    float rotational_score;
    float low_resolution_score;
  };

  // fast_correlative_scan_matcher_3d.cc:129-153. The grids must outlive the matcher.
  FastCorrelativeScanMatcher3D(const DeviceGrid& grid, const DeviceGrid* low_resolution_grid,
                               const std::vector<float>* rotational_scan_matcher_histogram,
                               const hg_fcsm_options& options) {
    CheckOk(hg_fcsm_create(grid.get(), low_resolution_grid->get(),
                           rotational_scan_matcher_histogram ? rotational_scan_matcher_histogram->data() : nullptr,
                           rotational_scan_matcher_histogram
                               ? static_cast<int>(rotational_scan_matcher_histogram->size())
                               : 0,
                           &options, &matcher_),
            "hg_fcsm_create");
  }
  ~FastCorrelativeScanMatcher3D() { hg_fcsm_destroy(matcher_); }
  FastCorrelativeScanMatcher3D(const FastCorrelativeScanMatcher3D&) = delete;
  FastCorrelativeScanMatcher3D& operator=(const FastCorrelativeScanMatcher3D&) = delete;

  // :157-174; nullptr when no candidate beats min_score and passes the low-resolution gate.
  std::unique_ptr<Result> Match(const Rigid3d& global_node_pose, const Rigid3d& global_submap_pose,
                                const TrajectoryNodeData& constant_data, float min_score) const {
    double node[7], submap[7];
    global_node_pose.ToArray(node);
    global_submap_pose.ToArray(submap);
    const hg_node_data d = constant_data.View();
    hg_match_result r;
    CheckOk(hg_fcsm_match(matcher_, node, submap, &d, min_score, &r), "hg_fcsm_match");
    return ToResult(r);
  }
  // :176-200
  std::unique_ptr<Result> MatchFullSubmap(const std::array<double, 4>& global_node_rotation,
                                          const std::array<double, 4>& global_submap_rotation,
                                          const TrajectoryNodeData& constant_data, float min_score) const {
    const hg_node_data d = constant_data.View();
    hg_match_result r;
    CheckOk(hg_fcsm_match_full_submap(matcher_, global_node_rotation.data(), global_submap_rotation.data(), &d,
                                      min_score, 0, 1, &r),
            "hg_fcsm_match_full_submap");
    return ToResult(r);
  }
  // MatchFullSubmap spread over num_shards processes, one GPU each (the matcher's discrete scans are
  // dealt round-robin; every process holds the same submap and node). The two collectives are the
  // caller's:
  //   all_reduce_max(float* device_value, void* cuda_stream) — all-reduce (max) of ONE float in device
  //     memory, in place, queued on cuda_stream, e.g. ncclAllReduce(p, p, 1, ncclFloat, ncclMax, comm,
  //     (cudaStream_t)cuda_stream); it is not called when the matcher searches on the host
  //     (strict_visit_order), where every shard searches under its own bound;
  //   all_gather(const double* mine, double* all, int doubles_per_rank) — all-gather of one host record
  //     per shard into all[num_shards * doubles_per_rank], in shard order.
  // Every process returns the same result: the reference's answer (fast_correlative_scan_matcher_3d.cc:
  // 176-200, 384, 437-468), equal leaves ordered the way its search meets them. *order_unresolved (may
  // be null) is set in the one case only an unsharded search in strict order can settle.
  template <typename AllReduceMax, typename AllGather>
  std::unique_ptr<Result> MatchFullSubmapSharded(const std::array<double, 4>& global_node_rotation,
                                                 const std::array<double, 4>& global_submap_rotation,
                                                 const TrajectoryNodeData& constant_data, float min_score,
                                                 int shard_index, int num_shards, AllReduceMax&& all_reduce_max,
                                                 AllGather&& all_gather, bool* order_unresolved = nullptr) const {
    const hg_node_data d = constant_data.View();
    hg_fcsm_search* search = nullptr;
    float* d_bound = nullptr;
    void* stream = nullptr;
    CheckOk(hg_fcsm_full_submap_begin_async(matcher_, global_node_rotation.data(), global_submap_rotation.data(), &d,
                                            min_score, shard_index, num_shards, &search, &d_bound, &stream),
            "hg_fcsm_full_submap_begin_async");
    if (d_bound != nullptr) all_reduce_max(d_bound, stream);
    hg_match_result mine;
    CheckOk(hg_fcsm_full_submap_finish(search, 0.f, &mine), "hg_fcsm_full_submap_finish");
    std::vector<double> record(HG_BEST_RECORD_DOUBLES), all(static_cast<size_t>(num_shards) * HG_BEST_RECORD_DOUBLES);
    CheckOk(hg_pack_best_record(&mine, record.data()), "hg_pack_best_record");
    all_gather(record.data(), all.data(), HG_BEST_RECORD_DOUBLES);
    int winner = -1, unresolved = 0;
    CheckOk(hg_select_best_record(all.data(), num_shards, &winner, &unresolved), "hg_select_best_record");
    if (order_unresolved != nullptr) *order_unresolved = unresolved != 0;
    if (winner < 0) return nullptr;
    hg_match_result best;
    CheckOk(hg_unpack_best_record(all.data() + static_cast<size_t>(winner) * HG_BEST_RECORD_DOUBLES, &best),
            "hg_unpack_best_record");
    return ToResult(best);
  }
  const hg_fcsm* handle() const { return matcher_; }

 private:
  static std::unique_ptr<Result> ToResult(const hg_match_result& r) {
    if (!r.found) return nullptr;
    std::unique_ptr<Result> out(new Result);
    out->score = r.score;
    for (int i = 0; i < 3; ++i) out->pose_estimate.translation[i] = r.pose_translation[i];
    for (int i = 0; i < 4; ++i) out->pose_estimate.rotation[i] = r.pose_rotation[i];
    out->rotational_score = r.rotational_score;
    out->low_resolution_score = r.low_resolution_score;
    return out;
  }
  hg_fcsm* matcher_ = nullptr;
};

// ConstraintBuilder3D (constraint_builder_3d.h:51-114), compute side. Pairs are queued by
// MaybeAddConstraint / MaybeAddGlobalConstraint and computed by a background worker in batches
// through hg_constraints_run_batch; WhenDone's callback receives the non-null constraints, like
// RunWhenDoneCallback (constraint_builder_3d.cc:297-323). MaybeAddConstraint applies the distance
// gate and the sampler of the reference (:82-86) before anything is queued.

// common::FixedRatioSampler (common/fixed_ratio_sampler.cc:23-43).
class FixedRatioSampler {
 public:
  explicit FixedRatioSampler(double ratio) : ratio_(ratio) {
    if (!(ratio >= 0.) || !(ratio <= 1.)) {
      std::fprintf(stderr, "[hg_b200] FixedRatioSampler: ratio must be in [0, 1]\n");  // CHECK_GE / CHECK_LE
      std::abort();
    }
    if (ratio == 0.) std::fprintf(stderr, "[hg_b200] FixedRatioSampler is dropping all data.\n");
  }
  bool Pulse() {
    ++num_pulses_;
    if (static_cast<double>(num_samples_) / num_pulses_ < ratio_) {
      ++num_samples_;
      return true;
    }
    return false;
  }

 private:
  const double ratio_;
  int64_t num_pulses_ = 0;
  int64_t num_samples_ = 0;
};

class ConstraintBuilder3D {
 public:
  struct SubmapId { int trajectory_id; int submap_index; };
  struct NodeId { int trajectory_id; int node_index; };
  struct Constraint {
    SubmapId submap_id;
    NodeId node_id;
    Rigid3d zbar_ij;
    double translation_weight;
    double rotation_weight;
  };
  using Result = std::vector<Constraint>;

  struct Options {
    hg_constraint_options compute;
    hg_fcsm_options fast_correlative_scan_matcher_options_3d;
    double loop_closure_translation_weight = 1.1e4;
    double loop_closure_rotation_weight = 1e5;
    // constraint_builder options (configuration_files/pose_graph.lua:18-19)
    double max_constraint_distance = 15.;
    double sampling_ratio = 0.3;
    // One process per GPU, every rank receiving the same sequence of MaybeAdd(Global)Constraint calls: this
    // builder computes the accepted pairs number shard_index, shard_index + num_shards, ... of that sequence
    // (0 / 1 = all); the ranks gather their constraints at WhenDone (fixed-size records).
    int shard_index = 0;
    int num_shards = 1;
  };

  explicit ConstraintBuilder3D(const Options& options)
      : options_(options), sampler_(options.sampling_ratio), worker_([this] { Work(); }) {}
  ~ConstraintBuilder3D() {
    {
      std::unique_lock<std::mutex> lock(mu_);
      // constraint_builder_3d.cc:68-75 CHECKs quiescence.
      if (!pending_.empty() || busy_ || when_done_) {
        std::fprintf(stderr, "[hg_b200] ConstraintBuilder3D destroyed while computations are pending\n");
        std::abort();
      }
      stop_ = true;
    }
    cv_.notify_all();
    worker_.join();
  }

  // Registers the grids of a submap; the matcher is built lazily on first use
  // (DispatchScanMatcherConstruction, constraint_builder_3d.cc:163-190).
  void AddSubmap(const SubmapId& id, std::shared_ptr<DeviceGrid> high, std::shared_ptr<DeviceGrid> low,
                 const std::vector<float>& rotational_scan_matcher_histogram) {
    std::unique_lock<std::mutex> lock(mu_);
    SubmapEntry& e = submaps_[Key(id)];
    e.high = std::move(high);
    e.low = std::move(low);
    e.histogram = rotational_scan_matcher_histogram;
  }
  void DeleteScanMatcher(const SubmapId& id) {  // :325-337
    std::unique_lock<std::mutex> lock(mu_);
    if (!pending_.empty() || busy_) {
      std::fprintf(stderr, "[hg_b200] DeleteScanMatcher while computations are pending\n");
      std::abort();
    }
    submaps_.erase(Key(id));
  }

  void MaybeAddConstraint(const SubmapId& submap_id, const NodeId& node_id,
                          const TrajectoryNodeData* constant_data, const Rigid3d& global_node_pose,
                          const Rigid3d& global_submap_pose) {
    // constraint_builder_3d.cc:82-86
    double d2 = 0.;
    for (int k = 0; k < 3; ++k) {
      const double d = global_node_pose.translation[k] - global_submap_pose.translation[k];
      d2 += d * d;
    }
    if (std::sqrt(d2) > options_.max_constraint_distance) return;
    {
      std::unique_lock<std::mutex> lock(sampler_mu_);
      if (!sampler_.Pulse()) return;
    }
    Enqueue(submap_id, node_id, constant_data, global_node_pose, global_submap_pose, false);
  }
  void MaybeAddGlobalConstraint(const SubmapId& submap_id, const NodeId& node_id,
                                const TrajectoryNodeData* constant_data, const Rigid3d& global_node_pose,
                                const Rigid3d& global_submap_pose) {
    Enqueue(submap_id, node_id, constant_data, global_node_pose, global_submap_pose, true);
  }
  void NotifyEndOfNode() {
    std::unique_lock<std::mutex> lock(mu_);
    ++num_finished_nodes_;
  }
  // Runs 'callback' (on the worker thread) once all queued computations are done.
  void WhenDone(const std::function<void(const Result&)>& callback) {
    {
      std::unique_lock<std::mutex> lock(mu_);
      if (when_done_) {
        std::fprintf(stderr, "[hg_b200] WhenDone called twice\n");  // CHECK(when_done_ == nullptr), :154
        std::abort();
      }
      when_done_.reset(new std::function<void(const Result&)>(callback));
    }
    cv_.notify_all();
  }
  int GetNumFinishedNodes() {
    std::unique_lock<std::mutex> lock(mu_);
    return num_finished_nodes_;
  }

 private:
  struct SubmapEntry {
    std::shared_ptr<DeviceGrid> high, low;
    std::vector<float> histogram;
    std::unique_ptr<FastCorrelativeScanMatcher3D> matcher;
  };
  struct Pending {
    SubmapId submap_id;
    NodeId node_id;
    const TrajectoryNodeData* data;
    Rigid3d node_pose, submap_pose;
    bool full;
  };
  static std::pair<int, int> Key(const SubmapId& id) { return {id.trajectory_id, id.submap_index}; }

  void Enqueue(const SubmapId& s, const NodeId& n, const TrajectoryNodeData* d, const Rigid3d& np,
               const Rigid3d& sp, bool full) {
    {
      std::unique_lock<std::mutex> lock(mu_);
      if (when_done_) {  // :88-91, :119-122: a warning, the constraint is still computed
        std::fprintf(stderr, "[hg_b200] MaybeAddConstraint was called while WhenDone was scheduled.\n");
      }
      const int64_t sequence = next_sequence_++;
      if (sequence % options_.num_shards != options_.shard_index) return;  // another rank computes this pair
      pending_.push_back(Pending{s, n, d, np, sp, full});
    }
    cv_.notify_all();
  }

  void Work() {
    for (;;) {
      std::vector<Pending> batch;
      {
        std::unique_lock<std::mutex> lock(mu_);
        cv_.wait(lock, [this] { return stop_ || !pending_.empty() || when_done_; });
        if (stop_) return;
        batch.assign(pending_.begin(), pending_.end());
        pending_.clear();
        busy_ = true;
      }
      if (!batch.empty()) Compute(batch);
      std::unique_ptr<std::function<void(const Result&)>> cb;
      Result out;
      {
        std::unique_lock<std::mutex> lock(mu_);
        busy_ = false;
        if (pending_.empty() && when_done_) {
          cb = std::move(when_done_);
          out.swap(constraints_);
        }
      }
      if (cb) (*cb)(out);
    }
  }

  void Compute(const std::vector<Pending>& batch) {
    // One matcher / grid triple per distinct submap in this batch.
    std::map<std::pair<int, int>, int> index;
    std::vector<const hg_fcsm*> matchers;
    std::vector<const hg_grid*> highs, lows;
    std::vector<hg_constraint_job> jobs(batch.size());
    for (size_t i = 0; i < batch.size(); ++i) {
      const Pending& p = batch[i];
      SubmapEntry* e;
      {
        std::unique_lock<std::mutex> lock(mu_);
        auto it = submaps_.find(Key(p.submap_id));
        if (it == submaps_.end()) {
          std::fprintf(stderr, "[hg_b200] constraint for an unknown submap\n");
          std::abort();
        }
        e = &it->second;
      }
      if (!e->matcher) {
        e->matcher.reset(new FastCorrelativeScanMatcher3D(*e->high, e->low.get(), &e->histogram,
                                                          options_.fast_correlative_scan_matcher_options_3d));
      }
      auto ins = index.emplace(Key(p.submap_id), static_cast<int>(matchers.size()));
      if (ins.second) {
        matchers.push_back(e->matcher->handle());
        highs.push_back(e->high->get());
        lows.push_back(e->low->get());
      }
      hg_constraint_job& j = jobs[i];
      j.submap_index = ins.first->second;
      j.match_full_submap = p.full ? 1 : 0;
      p.node_pose.ToArray(j.global_node_pose);
      p.submap_pose.ToArray(j.global_submap_pose);
      j.data = p.data->View();
    }
    std::vector<hg_constraint_result> results(batch.size());
    CheckOk(hg_constraints_run_batch(matchers.data(), highs.data(), lows.data(), static_cast<int>(matchers.size()),
                                     jobs.data(), static_cast<int64_t>(jobs.size()), &options_.compute, 0, 1,
                                     results.data()),
            "hg_constraints_run_batch");
    std::unique_lock<std::mutex> lock(mu_);
    for (size_t i = 0; i < batch.size(); ++i) {
      if (!results[i].found) continue;
      Constraint c;
      c.submap_id = batch[i].submap_id;
      c.node_id = batch[i].node_id;
      for (int k = 0; k < 3; ++k) c.zbar_ij.translation[k] = results[i].pose[k];
      for (int k = 0; k < 4; ++k) c.zbar_ij.rotation[k] = results[i].pose[3 + k];
      c.translation_weight = options_.loop_closure_translation_weight;
      c.rotation_weight = options_.loop_closure_rotation_weight;
      constraints_.push_back(c);
    }
  }

  const Options options_;
  std::mutex sampler_mu_;
  FixedRatioSampler sampler_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Pending> pending_;
  std::map<std::pair<int, int>, SubmapEntry> submaps_;
  Result constraints_;
  std::unique_ptr<std::function<void(const Result&)>> when_done_;
  int num_finished_nodes_ = 0;
  int64_t next_sequence_ = 0;
  bool busy_ = false;
  bool stop_ = false;
  std::thread worker_;  // last member: starts after everything above is constructed
};

}  // namespace hg
