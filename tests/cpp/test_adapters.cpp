// Exercises the C++ host adapters (hectorgrapher_b200/csrc/host/hg_adapters.h) the way the
// reference's own tests exercise its classes:
//   ceres_scan_matcher_3d_test.cc:36-117      (7-point cloud, occupancy grid, pose recovered)
//   fast_correlative_scan_matcher_3d_test.cc  (Match returns a result / a null result)
//   constraint_builder_3d_test.cc:62-72       (CallsBack)
//   range_data_inserter_3d_test.cc:43-98      (hit / miss probabilities after one insertion)
// Prints "ADAPTERS OK" and exits 0 on success. With argument "compile-only" it exits before
// touching the GPU (used by the CPU-side test to check that the header compiles and links).
#include <array>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <future>

#include "../../hectorgrapher_b200/csrc/host/hg_adapters.h"

static int Fail(const char* what) {
  std::fprintf(stderr, "FAILED: %s\n", what);
  return 1;
}

int main(int argc, char** argv) {
  if (argc > 1 && std::strcmp(argv[1], "compile-only") == 0) {
    std::printf("%s\n", hg_version());
    return 0;
  }
  // --- CeresScanMatcher3D fixture -------------------------------------------------------
  const float pts[7][3] = {{-3, 2, 0}, {-4, 2, 0}, {-5, 2, 0}, {-6, 2, 0}, {-6, 3, 1}, {-6, 4, 2}, {-7, 3, 1}};
  std::vector<int32_t> xyz;
  std::vector<uint16_t> values;
  for (auto& p : pts) {
    xyz.push_back(static_cast<int32_t>(std::lround(p[0] - 1.f)));
    xyz.push_back(static_cast<int32_t>(std::lround(p[1])));
    xyz.push_back(static_cast<int32_t>(std::lround(p[2])));
    values.push_back(32767);  // ProbabilityToValue(1.) clamps to kMaxProbability
  }
  auto grid = hg::DeviceGrid::FromHybridGrid(1.f, 0, 7, xyz.data(), values.data());
  hg::PointCloudView cloud;
  cloud.data = &pts[0][0];
  cloud.size = 7;
  // A ceres::CostFunction created through the reference-named factory.
  std::unique_ptr<hg::CostFunctionBase> cost(
      hg::OccupiedSpaceCostFunction3D::CreateAutoDiffCostFunction(1.0 / std::sqrt(7.0), cloud, *grid));
  if (cost->num_residuals() != 7 || cost->parameter_block_sizes().size() != 2) return Fail("cost function shape");
  double t[3] = {-1, 0, 0}, q[4] = {1, 0, 0, 0};
  const double* params[2] = {t, q};
  double r[7], jt[21], jq[28];
  double* jac[2] = {jt, jq};
  if (!cost->Evaluate(params, r, jac)) return Fail("Evaluate");
  for (int i = 0; i < 7; ++i)
    if (std::fabs(r[i] - (1.0 - 0.9) / std::sqrt(7.0)) > 1e-6) return Fail("residual at the optimum");
  if (!cost->Evaluate(params, r, nullptr)) return Fail("Evaluate without jacobians");

  hg::CeresScanMatcherOptions3D options;
  options.occupied_space_weight = {1.0};
  options.translation_weight = 0.01;
  options.rotation_weight = 0.1;
  options.max_num_iterations = 40;
  hg::CeresScanMatcher3D matcher(options);
  hg::Rigid3d initial, estimate;
  initial.translation = {{-0.8, 0., 0.}};
  hg::SolverSummary summary;
  matcher.Match(initial.translation, initial, {{cloud, grid.get()}}, &estimate, &summary);
  if (summary.final_cost > 1e-2) return Fail("final cost");
  if (std::fabs(estimate.translation[0] + 1.0) > 6e-2 || std::fabs(estimate.translation[1]) > 6e-2 ||
      std::fabs(estimate.translation[2]) > 6e-2)
    return Fail("CeresScanMatcher3D pose");

  // --- FastCorrelativeScanMatcher3D -----------------------------------------------------
  // A small L-shaped wall of occupied voxels at 5 cm, scan = the wall's voxel centres shifted.
  std::vector<int32_t> wall_xyz;
  std::vector<uint16_t> wall_values;
  hg::TrajectoryNodeData data;
  for (int i = 0; i < 40; ++i)
    for (int k = 0; k < 10; ++k) {
      const int cells[2][3] = {{i + 20, 20, k}, {20, i + 20, k}};
      for (auto& c : cells) {
        wall_xyz.insert(wall_xyz.end(), {c[0], c[1], c[2]});
        wall_values.push_back(30000);
        if ((i + k) % 3 == 0) {
          data.high_resolution_point_cloud.insert(data.high_resolution_point_cloud.end(),
                                                  {c[0] * 0.05f - 0.2f, c[1] * 0.05f + 0.15f, c[2] * 0.05f});
        }
      }
    }
  data.low_resolution_point_cloud = data.high_resolution_point_cloud;
  data.rotational_scan_matcher_histogram.assign(10, 0.f);
  auto wall = hg::DeviceGrid::FromHybridGrid(0.05f, 0, static_cast<int64_t>(wall_values.size()), wall_xyz.data(),
                                             wall_values.data());
  hg_fcsm_options fo;
  std::memset(&fo, 0, sizeof(fo));
  fo.branch_and_bound_depth = 5;
  fo.full_resolution_depth = 3;
  fo.use_rotational_scan_matcher = 1;
  fo.min_rotational_score = 0.1;
  fo.min_low_resolution_score = 0.15;
  fo.linear_xy_search_window = 0.5;
  fo.linear_z_search_window = 0.2;
  fo.angular_search_window = 0.05;
  fo.float_flavor = 1;
  const std::vector<float> submap_histogram(10, 0.f);
  {
    hg::FastCorrelativeScanMatcher3D fcsm(*wall, wall.get(), &submap_histogram, fo);
    hg::Rigid3d identity;
    auto result = fcsm.Match(identity, identity, data, 0.3f);
    if (!result) return Fail("FastCorrelativeScanMatcher3D::Match returned null");
    if (std::fabs(result->pose_estimate.translation[0] - 0.2) > 0.051 ||
        std::fabs(result->pose_estimate.translation[1] + 0.15) > 0.051)
      return Fail("FastCorrelativeScanMatcher3D pose");
    hg::TrajectoryNodeData far = data;
    far.low_resolution_point_cloud = {42.f, 42.f, 42.f};
    if (fcsm.Match(identity, identity, far, 0.3f)) return Fail("low resolution gate");
    // The sharded form with one shard and pass-through collectives is the unsharded call.
    const std::array<double, 4> no_rotation = {{1.0, 0.0, 0.0, 0.0}};
    auto whole = fcsm.MatchFullSubmap(no_rotation, no_rotation, data, 0.3f);
    int reduces = 0, gathers = 0;
    bool unresolved = true;
    auto sharded = fcsm.MatchFullSubmapSharded(
        no_rotation, no_rotation, data, 0.3f, 0, 1,
        [&reduces](float* device_value, void* stream) { reduces += device_value != nullptr && stream != nullptr; },
        [&gathers](const double* mine, double* all, int n) {
          ++gathers;
          std::memcpy(all, mine, sizeof(double) * n);
        },
        &unresolved);
    if (reduces != 1 || gathers != 1) return Fail("MatchFullSubmapSharded: collectives not called once each");
    if (static_cast<bool>(whole) != static_cast<bool>(sharded)) return Fail("MatchFullSubmapSharded: found differs");
    if (whole) {
      if (whole->score != sharded->score) return Fail("MatchFullSubmapSharded: score");
      for (int i = 0; i < 3; ++i)
        if (whole->pose_estimate.translation[i] != sharded->pose_estimate.translation[i])
          return Fail("MatchFullSubmapSharded: translation");
      for (int i = 0; i < 4; ++i)
        if (whole->pose_estimate.rotation[i] != sharded->pose_estimate.rotation[i])
          return Fail("MatchFullSubmapSharded: rotation");
    }
  }

  // --- ConstraintBuilder3D: CallsBack + finds the constraint ------------------------------
  {
    hg::ConstraintBuilder3D::Options co;
    std::memset(&co.compute, 0, sizeof(co.compute));
    co.compute.min_score = 0.3f;
    co.compute.global_localization_min_score = 0.4f;
    co.compute.occupied_space_weights[0] = 5.;
    co.compute.occupied_space_weights[1] = 30.;
    co.compute.translation_weight = 10.;
    co.compute.rotation_weight = 1.;
    co.compute.max_num_iterations = 10;
    co.fast_correlative_scan_matcher_options_3d = fo;
    co.sampling_ratio = 1.0;
    hg::ConstraintBuilder3D builder(co);
    if (builder.GetNumFinishedNodes() != 0) return Fail("GetNumFinishedNodes");
    builder.AddSubmap({0, 0}, wall, wall, submap_histogram);
    hg::Rigid3d identity;
    for (int n = 0; n < 3; ++n) {
      builder.MaybeAddConstraint({0, 0}, {0, n}, &data, identity, identity);
      builder.NotifyEndOfNode();
    }
    std::promise<hg::ConstraintBuilder3D::Result> done;
    builder.WhenDone([&done](const hg::ConstraintBuilder3D::Result& r) { done.set_value(r); });
    const hg::ConstraintBuilder3D::Result constraints = done.get_future().get();
    if (builder.GetNumFinishedNodes() != 3) return Fail("finished nodes");
    if (constraints.size() != 3) return Fail("expected 3 constraints");
    if (std::fabs(constraints[0].zbar_ij.translation[0] - 0.2) > 0.06) return Fail("constraint pose");
    // a second round with nothing queued still calls back
    std::promise<size_t> again;
    builder.WhenDone([&again](const hg::ConstraintBuilder3D::Result& r) { again.set_value(r.size()); });
    if (again.get_future().get() != 0) return Fail("empty WhenDone");
  }
  // --- MaybeAddConstraint's distance gate and sampler (constraint_builder_3d.cc:82-86) ------------
  {
    hg::FixedRatioSampler sampler(0.3);  // common/fixed_ratio_sampler.cc:33-40: 1, 0, 0, 1, 0, 0, 1, 0, 0, 0
    int taken = 0;
    const bool expect[10] = {true, false, false, true, false, false, true, false, false, false};
    for (int k = 0; k < 10; ++k) {
      const bool got = sampler.Pulse();
      if (got != expect[k]) return Fail("FixedRatioSampler sequence");
      taken += got ? 1 : 0;
    }
    if (taken != 3) return Fail("FixedRatioSampler count");
    hg::ConstraintBuilder3D::Options co;
    std::memset(&co.compute, 0, sizeof(co.compute));
    co.compute.min_score = 0.3f;
    co.compute.global_localization_min_score = 0.4f;
    co.compute.occupied_space_weights[0] = 5.;
    co.compute.occupied_space_weights[1] = 30.;
    co.compute.translation_weight = 10.;
    co.compute.rotation_weight = 1.;
    co.compute.max_num_iterations = 10;
    co.fast_correlative_scan_matcher_options_3d = fo;
    co.sampling_ratio = 0.5;
    co.max_constraint_distance = 15.;
    hg::ConstraintBuilder3D builder(co);
    builder.AddSubmap({0, 0}, wall, wall, submap_histogram);
    hg::Rigid3d identity, far_node;
    far_node.translation = {{16., 0., 0.}};
    // six near nodes at ratio 0.5 -> three are sampled; the far ones never reach the sampler
    for (int n = 0; n < 6; ++n) {
      builder.MaybeAddConstraint({0, 0}, {0, 100 + n}, &data, far_node, identity);
      builder.MaybeAddConstraint({0, 0}, {0, n}, &data, identity, identity);
      builder.NotifyEndOfNode();
    }
    std::promise<hg::ConstraintBuilder3D::Result> done;
    builder.WhenDone([&done](const hg::ConstraintBuilder3D::Result& r) { done.set_value(r); });
    const hg::ConstraintBuilder3D::Result constraints = done.get_future().get();
    if (constraints.size() != 3) return Fail("distance gate / sampler: expected 3 constraints");
    for (const auto& c : constraints)
      if (c.node_id.node_index >= 100) return Fail("a constraint beyond max_constraint_distance was computed");
  }
  // --- CeresScanMatcher3D::Evaluate (ceres_scan_matcher_3d.cc:97-118) ------------------------------
  {
    hg::CeresScanMatcherOptions3D so;
    so.occupied_space_weight = {1.0};
    so.translation_weight = 0.5;
    so.rotation_weight = 0.1;
    hg::CeresScanMatcher3D matcher(so);
    hg::Rigid3d pose;
    pose.translation = {{0.1, -0.05, 0.02}};
    hg::PointCloudView view;
    view.data = data.high_resolution_point_cloud.data();
    view.size = static_cast<int32_t>(data.high_resolution_point_cloud.size() / 3);
    double cost = -1.;
    std::vector<double> residuals, gradient;
    matcher.Evaluate({{0., 0., 0.}}, pose, {{view, wall.get()}}, &cost, &residuals, &gradient);
    if (residuals.size() != static_cast<size_t>(view.size) + 6 || gradient.size() != 6) return Fail("Evaluate sizes");
    double half = 0.;
    for (double r : residuals) half += 0.5 * r * r;
    if (std::fabs(half - cost) > 1e-12 * (1. + cost)) return Fail("Evaluate cost != 1/2 |r|^2");
    if (std::fabs(residuals[view.size] - 0.5 * 0.1) > 1e-12) return Fail("Evaluate translation delta");
  }
  // --- best-record exchange of a sharded search (hg_pack / select) ------------------------------------
  {
    hg_match_result a, b;
    std::memset(&a, 0, sizeof(a));
    std::memset(&b, 0, sizeof(b));
    a.found = b.found = 1;
    a.score = b.score = 0.8f;
    a.best_ancestor_score = 0.9f;
    b.best_ancestor_score = 0.95f;
    a.best_scan_index = 3;
    b.best_scan_index = 40;
    double records[2 * HG_BEST_RECORD_DOUBLES];
    if (hg_pack_best_record(&a, records) != HG_OK || hg_pack_best_record(&b, records + HG_BEST_RECORD_DOUBLES) != HG_OK)
      return Fail("hg_pack_best_record");
    int winner = -1, unresolved = -1;
    if (hg_select_best_record(records, 2, &winner, &unresolved) != HG_OK || winner != 1 || unresolved != 0)
      return Fail("hg_select_best_record: the leaf below the higher-scoring ancestor is met first");
    hg_match_result back;
    if (hg_unpack_best_record(records + HG_BEST_RECORD_DOUBLES, &back) != HG_OK || back.best_scan_index != 40)
      return Fail("hg_unpack_best_record");
  }
  // --- range-data inserters (range_data_inserter_3d_test.cc:43-98) ---------------------------
  {
    auto occupancy = hg::DeviceGrid::FromHybridGrid(1.f, 0, 0, nullptr, nullptr);
    hg_occupancy_inserter_options oo = {0.7, 0.4, 1000, 0};
    hg::OccupancyGridRangeDataInserter3D inserter(oo);
    const float returns[4][3] = {{-3.f, -1.f, 4.f}, {-2.f, 0.f, 4.f}, {-1.f, 1.f, 4.f}, {0.f, 2.f, 4.f}};
    hg::RangeDataView range_data;
    range_data.origin = {{0.f, 0.f, -4.f}};
    range_data.returns.data = &returns[0][0];
    range_data.returns.size = 4;
    inserter.Insert(range_data, occupancy.get());
    const int32_t cells[3][3] = {{0, 0, -4}, {-2, 0, 4}, {-2, 0, 3}};
    float probability[3];
    if (hg_grid_lookup(occupancy->get(), 0, 3, &cells[0][0], probability) != HG_OK) return Fail("lookup");
    if (std::fabs(probability[0] - 0.4f) > 1e-4f || std::fabs(probability[1] - 0.7f) > 1e-4f ||
        std::fabs(probability[2] - 0.4f) > 1e-4f)
      return Fail("OccupancyGridRangeDataInserter3D probabilities");

    auto tsdf = hg::DeviceGrid::FromHybridGridTSDF(0.1f, 2.5f, 1000.f, 0, 0, nullptr, nullptr, nullptr);
    hg_tsdf_inserter_options to = {2.5, 1000., 1.0, 4., 0.4, 15.0, 1.0, 0, 0};
    hg::TSDFRangeDataInserter3D tsdf_inserter(to);
    const float hit[1][3] = {{2.f, 0.f, 0.f}};
    range_data.origin = {{0.f, 0.f, 0.f}};
    range_data.returns.data = &hit[0][0];
    range_data.returns.size = 1;
    tsdf_inserter.Insert(range_data, tsdf.get());
    const int32_t surface[2][3] = {{20, 0, 0}, {19, 0, 0}};
    float tsd[2], weight[2];
    if (hg_grid_lookup(tsdf->get(), 1, 2, &surface[0][0], tsd) != HG_OK) return Fail("tsd lookup");
    if (hg_grid_lookup(tsdf->get(), 2, 2, &surface[0][0], weight) != HG_OK) return Fail("weight lookup");
    if (std::fabs(tsd[0]) > 1e-3f || std::fabs(tsd[1] - 0.1f) > 1e-3f || std::fabs(weight[0] - 1.f) > 0.05f)
      return Fail("TSDFRangeDataInserter3D voxels");
  }
  std::printf("ADAPTERS OK (%llu kernel launches)\n", static_cast<unsigned long long>(hg_kernel_launch_count()));
  return 0;
}
