// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// Plain-C surface over the CPU restatement so that tests (ctypes) and the
// cpu_baseline / --impl reference legs of bench.py can drive it. See the
// headers next to this file for the reference file:line each piece follows.
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../include/hg_dense_lm.h"
#include "cost_functions.hpp"
#include "fcsm.hpp"
#include "grid.hpp"
#include "interpolation.hpp"
#include "voxel_filter.hpp"

using namespace oracle;

namespace {
struct GridHandle {
  std::unique_ptr<HybridGrid> prob;
  std::unique_ptr<HybridGridTSDF> tsdf;
  const GridInterface* iface() const {
    return prob ? static_cast<const GridInterface*>(prob.get())
                : static_cast<const GridInterface*>(tsdf.get());
  }
};
struct CostHandle {
  std::unique_ptr<ScanCostFunction> f;
};
struct StackHandle {
  std::unique_ptr<PrecomputationGridStack3D> stack;
};
struct FcsmHandle {
  std::vector<float> histogram;
  std::unique_ptr<FastCorrelativeScanMatcher3D> m;
};
std::vector<Vector3f> ToCloud(const float* pts, int n, int stride) {
  std::vector<Vector3f> c;
  c.reserve(n);
  for (int i = 0; i < n; ++i) c.emplace_back(pts[i * stride], pts[i * stride + 1], pts[i * stride + 2]);
  return c;
}
}  // namespace

extern "C" {

void orc_set_float_flavor(int flavor) { FloatFlavor::Get() = flavor; }

// ---- grids ----------------------------------------------------------------
void* orc_prob_grid_new(float resolution) {
  auto* h = new GridHandle;
  h->prob.reset(new HybridGrid(resolution));
  return h;
}
void* orc_tsdf_grid_new(float resolution, float relative_truncation_distance, float max_weight) {
  auto* h = new GridHandle;
  h->tsdf.reset(new HybridGridTSDF(resolution, relative_truncation_distance, max_weight));
  return h;
}
void orc_grid_free(void* g) { delete static_cast<GridHandle*>(g); }
int orc_grid_is_tsdf(void* g) { return static_cast<GridHandle*>(g)->tsdf ? 1 : 0; }
int orc_grid_size(void* g) {
  auto* h = static_cast<GridHandle*>(g);
  return h->prob ? h->prob->grid_size() : h->tsdf->grid_size();
}
float orc_grid_resolution(void* g) {
  auto* h = static_cast<GridHandle*>(g);
  return h->prob ? h->prob->resolution() : h->tsdf->resolution();
}
float orc_tsdf_grid_max_tsd(void* g) {
  return static_cast<GridHandle*>(g)->tsdf->ValueConverter().getMaxTSD();
}
void orc_prob_grid_set_raw(void* g, int n, const int32_t* xyz, const uint16_t* values) {
  auto* h = static_cast<GridHandle*>(g);
  for (int i = 0; i < n; ++i)
    *h->prob->mutable_value(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = values[i];
}
void orc_prob_grid_set_probability(void* g, int n, const int32_t* xyz, const float* p) {
  auto* h = static_cast<GridHandle*>(g);
  for (int i = 0; i < n; ++i)
    h->prob->SetProbability(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), p[i]);
}
void orc_tsdf_grid_set_raw(void* g, int n, const int32_t* xyz, const uint16_t* tsd,
                           const uint16_t* weight) {
  auto* h = static_cast<GridHandle*>(g);
  for (int i = 0; i < n; ++i) {
    TSDFVoxel v;
    v.discrete_tsd = tsd[i];
    v.discrete_weight = weight[i];
    *h->tsdf->mutable_value(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2])) = v;
  }
}
void orc_tsdf_grid_set_cell(void* g, int n, const int32_t* xyz, const float* tsd,
                            const float* weight) {
  auto* h = static_cast<GridHandle*>(g);
  for (int i = 0; i < n; ++i)
    h->tsdf->SetCell(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), tsd[i], weight[i]);
}
long long orc_grid_num_voxels(void* g) {
  auto* h = static_cast<GridHandle*>(g);
  long long n = 0;
  if (h->prob) {
    h->prob->ForEach([&n](const Array3i&, const uint16) { ++n; });
  } else {
    h->tsdf->ForEach([&n](const Array3i&, const TSDFVoxel&) { ++n; });
  }
  return n;
}
// a = probability value or tsd value, b = weight value (TSDF only; may be null for prob).
long long orc_grid_export(void* g, int32_t* xyz, uint16_t* a, uint16_t* b) {
  auto* h = static_cast<GridHandle*>(g);
  long long n = 0;
  if (h->prob) {
    h->prob->ForEach([&](const Array3i& i, const uint16 v) {
      xyz[3 * n] = i.x; xyz[3 * n + 1] = i.y; xyz[3 * n + 2] = i.z;
      a[n] = v;
      ++n;
    });
  } else {
    h->tsdf->ForEach([&](const Array3i& i, const TSDFVoxel& v) {
      xyz[3 * n] = i.x; xyz[3 * n + 1] = i.y; xyz[3 * n + 2] = i.z;
      a[n] = v.discrete_tsd;
      b[n] = v.discrete_weight;
      ++n;
    });
  }
  return n;
}
void orc_get_cell_index(float resolution, int n, const float* pts, int32_t* out) {
  HybridGridBase<uint8> g(resolution);
  for (int i = 0; i < n; ++i) {
    const Array3i c = g.GetCellIndex(Vector3f(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]));
    out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
  }
}
void orc_get_center_of_cell(float resolution, int n, const int32_t* xyz, float* out) {
  HybridGridBase<uint8> g(resolution);
  for (int i = 0; i < n; ++i) {
    const Vector3f c = g.GetCenterOfCell(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
    out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
  }
}
// what: 0 probability, 1 tsd, 2 weight, 3 raw value a, 4 raw value b
void orc_grid_lookup(void* g, int what, int n, const int32_t* xyz, float* out) {
  auto* h = static_cast<GridHandle*>(g);
  for (int i = 0; i < n; ++i) {
    const Array3i c(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    switch (what) {
      case 0: out[i] = h->prob->GetProbability(c); break;
      case 1: out[i] = h->tsdf->GetTSD(c); break;
      case 2: out[i] = h->tsdf->GetWeight(c); break;
      case 3: out[i] = h->prob ? h->prob->value(c) : h->tsdf->value(c).discrete_tsd; break;
      case 4: out[i] = h->tsdf->value(c).discrete_weight; break;
    }
  }
}
void orc_value_tables(float unknown, float lo, float hi, float* out65536) {
  const std::vector<float> t = MakeValueToFloatTable(unknown, lo, hi);
  std::memcpy(out65536, t.data(), 65536 * sizeof(float));
}
uint16_t orc_probability_to_value(float p) { return ProbabilityToValue(p); }
float orc_value_to_probability(uint16_t v) { return ValueToProbability(v); }
void orc_apply_odds_table(float odds, uint16_t* out32768) {
  const std::vector<uint16> t = ComputeLookupTableToApplyOdds(odds);
  std::memcpy(out32768, t.data(), 32768 * sizeof(uint16_t));
}
float orc_odds(float p) { return Odds(p); }
int orc_prob_grid_apply_odds(void* g, const int32_t* xyz, float probability) {
  auto* h = static_cast<GridHandle*>(g);
  return h->prob->ApplyLookupTable(Array3i(xyz[0], xyz[1], xyz[2]),
                                   ComputeLookupTableToApplyOdds(Odds(probability)))
             ? 1 : 0;
}
void orc_prob_grid_finish_update(void* g) { static_cast<GridHandle*>(g)->prob->FinishUpdate(); }

void orc_insert_occupancy(void* g, const float* origin, int n, const float* pts, float hit_p,
                          float miss_p, int num_free_space_voxels) {
  auto* h = static_cast<GridHandle*>(g);
  OccupancyInserterOptions o;
  o.hit_probability = hit_p;
  o.miss_probability = miss_p;
  o.num_free_space_voxels = num_free_space_voxels;
  OccupancyGridRangeDataInserter3D inserter(o);
  inserter.Insert(Vector3f(origin[0], origin[1], origin[2]), ToCloud(pts, n, 3), h->prob.get());
}
void orc_insert_tsdf(void* g, const float* origin, int n, const float* pts,
                     double relative_truncation_distance, double maximum_weight,
                     int num_free_space_voxels, double epsilon, double sigma, double min_range,
                     double max_range, double insertion_ratio) {
  auto* h = static_cast<GridHandle*>(g);
  TsdfInserterOptions o;
  o.relative_truncation_distance = relative_truncation_distance;
  o.maximum_weight = maximum_weight;
  o.num_free_space_voxels = num_free_space_voxels;
  o.weight_function_epsilon = epsilon;
  o.weight_function_sigma = sigma;
  o.min_range = min_range;
  o.max_range = max_range;
  o.insertion_ratio = insertion_ratio;
  TsdfInsert(o, Vector3f(origin[0], origin[1], origin[2]), ToCloud(pts, n, 3), h->tsdf.get());
}

// ---- pose helpers pinned by the reference's own tests -----------------------
// transform/timestamped_transform.h:39-48; poses are t3 q4 (w, x, y, z).
void orc_interpolate_transform(const double* start, const double* end, double factor, double* out) {
  const Rigid3<double> a(Vec3<double>(start[0], start[1], start[2]), Quat<double>(start[3], start[4], start[5], start[6]));
  const Rigid3<double> b(Vec3<double>(end[0], end[1], end[2]), Quat<double>(end[3], end[4], end[5], end[6]));
  const Rigid3<double> r = InterpolateTransform(a, b, factor);
  out[0] = r.translation.x; out[1] = r.translation.y; out[2] = r.translation.z;
  out[3] = r.rotation.w; out[4] = r.rotation.x; out[5] = r.rotation.y; out[6] = r.rotation.z;
}
// rotation_delta_cost_functor_3d.h:35-54: residuals[3] (and the 3 x 4 Jacobian if asked).
void orc_rotation_delta(double scaling_factor, const double* target_wxyz, const double* rotation_wxyz,
                        double* residuals, double* jacobian) {
  RotationDeltaCostFunction f(scaling_factor, target_wxyz);
  const double* params[1] = {rotation_wxyz};
  double* jac[1] = {jacobian};
  f.Evaluate(params, residuals, jacobian ? jac : nullptr);
}
// translation_delta_cost_functor_3d.h:31-45
void orc_translation_delta(double scaling_factor, const double* target, const double* translation,
                           double* residuals, double* jacobian) {
  TranslationDeltaCostFunction f(scaling_factor, target);
  const double* params[1] = {translation};
  double* jac[1] = {jacobian};
  f.Evaluate(params, residuals, jacobian ? jac : nullptr);
}

// ---- interpolators (double path) -------------------------------------------
void orc_interpolate(void* g, int n, const double* xyz, double* out) {
  auto* h = static_cast<GridHandle*>(g);
  if (h->prob) {
    InterpolatedGrid ig(*h->prob);
    for (int i = 0; i < n; ++i) out[i] = ig.GetProbability(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  } else {
    InterpolatedTSDF it(*h->tsdf);
    for (int i = 0; i < n; ++i) out[i] = it.GetTSD(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
  }
}

void orc_interpolate_weight(void* g, int n, const double* xyz, double* out) {
  auto* h = static_cast<GridHandle*>(g);
  InterpolatedTSDF it(*h->tsdf);
  for (int i = 0; i < n; ++i) out[i] = it.GetWeight(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}

// ---- cost functions ---------------------------------------------------------
void* orc_cost_new(int kind, double scaling_factor, const float* pts, int n, int stride,
                   void* const* grids, int num_grids, double interpolation_ratio) {
  CloudView cloud;
  cloud.data = pts;
  cloud.n = n;
  cloud.stride = stride;
  const HybridGrid* prob = nullptr;
  std::vector<const HybridGridTSDF*> tsdfs;
  for (int i = 0; i < num_grids; ++i) {
    auto* h = static_cast<GridHandle*>(grids[i]);
    if (h->prob) prob = h->prob.get();
    if (h->tsdf) tsdfs.push_back(h->tsdf.get());
  }
  auto* c = new CostHandle;
  c->f.reset(new ScanCostFunction(kind, scaling_factor, cloud, prob, tsdfs, interpolation_ratio));
  return c;
}
void orc_cost_free(void* c) { delete static_cast<CostHandle*>(c); }
int orc_cost_num_residuals(void* c) { return static_cast<CostHandle*>(c)->f->num_residuals(); }
int orc_cost_num_blocks(void* c) {
  return static_cast<int>(static_cast<CostHandle*>(c)->f->parameter_block_sizes().size());
}
// params: array of block pointers; jacobians: array of block pointers (entries may be null) or null.
int orc_cost_evaluate(void* c, const double* const* params, double* residuals, double** jacobians) {
  return static_cast<CostHandle*>(c)->f->Evaluate(params, residuals, jacobians) ? 1 : 0;
}

// Times 'repeats' Evaluate calls (with Jacobians) on 'threads' std::threads,
// each thread running the same block; returns seconds of wall time.
double orc_cost_time(void* c, const double* const* params, int repeats, int threads) {
  auto* h = static_cast<CostHandle*>(c);
  const int n = h->f->num_residuals();
  const std::vector<int> sizes = h->f->parameter_block_sizes();
  auto work = [&]() {
    std::vector<double> r(n);
    std::vector<std::vector<double>> J(sizes.size());
    std::vector<double*> Jp(sizes.size());
    for (size_t b = 0; b < sizes.size(); ++b) {
      J[b].resize(static_cast<size_t>(n) * sizes[b]);
      Jp[b] = J[b].data();
    }
    for (int k = 0; k < repeats; ++k) h->f->Evaluate(params, r.data(), Jp.data());
  };
  const auto t0 = std::chrono::steady_clock::now();
  if (threads <= 1) {
    work();
  } else {
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) pool.emplace_back(work);
    for (auto& t : pool) t.join();
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int orc_scan_match_trace(int, const float* const*, const int*, void* const*, const double*, double, double, int, int,
                         int, const double*, const double*, double*, double*, double*, int);

// ---- CeresScanMatcher3D::Match restated on the dense LM of include/hg_dense_lm.h
// (ceres_scan_matcher_3d.cc:72-95, :120-172). clouds[i]/grids[i] pairs.
int orc_scan_match(int num_clouds, const float* const* clouds, const int* cloud_sizes,
                   void* const* grids, const double* occupied_space_weights,
                   double translation_weight, double rotation_weight, int only_optimize_yaw,
                   int max_num_iterations, int use_nonmonotonic_steps,
                   const double* target_translation, const double* initial_pose /* t3 q4 */,
                   double* pose_out /* t3 q4 */, double* summary_out /* initial, final, iters */) {
  return orc_scan_match_trace(num_clouds, clouds, cloud_sizes, grids, occupied_space_weights, translation_weight,
                              rotation_weight, only_optimize_yaw, max_num_iterations, use_nonmonotonic_steps,
                              target_translation, initial_pose, pose_out, summary_out, nullptr, 0);
}

// Same, and the per-iteration log (Solver::Summary::iterations): 7 doubles per entry —
// iteration, cost, candidate cost, model cost change, step quality, radius, successful.
int orc_scan_match_trace(int num_clouds, const float* const* clouds, const int* cloud_sizes,
                         void* const* grids, const double* occupied_space_weights,
                         double translation_weight, double rotation_weight, int only_optimize_yaw,
                         int max_num_iterations, int use_nonmonotonic_steps,
                         const double* target_translation, const double* initial_pose,
                         double* pose_out, double* summary_out, double* trace_out, int trace_capacity) {
  double t[3] = {initial_pose[0], initial_pose[1], initial_pose[2]};
  double q[4] = {initial_pose[3], initial_pose[4], initial_pose[5], initial_pose[6]};
  hg::DenseLmProblem<CostFunction> problem;
  const int tb = problem.AddParameterBlock(t, 3, hg::LocalParameterization::kIdentity);
  const int qb = problem.AddParameterBlock(
      q, 4, only_optimize_yaw ? hg::LocalParameterization::kYawOnlyQuaternion
                              : hg::LocalParameterization::kQuaternion);
  std::vector<std::unique_ptr<CostFunction>> owned;
  for (int i = 0; i < num_clouds; ++i) {
    auto* h = static_cast<GridHandle*>(grids[i]);
    CloudView cloud;
    cloud.data = clouds[i];
    cloud.n = cloud_sizes[i];
    cloud.stride = 3;
    const double scaling = occupied_space_weights[i] / std::sqrt(static_cast<double>(cloud.n));
    if (h->prob) {
      owned.emplace_back(new ScanCostFunction(kOccupiedSpace, scaling, cloud, h->prob.get(), {}, 0.));
    } else {
      owned.emplace_back(new ScanCostFunction(kTSDFSpace, scaling, cloud, nullptr, {h->tsdf.get()}, 0.));
    }
    problem.AddResidualBlock(owned.back().get(), {tb, qb});
  }
  owned.emplace_back(new TranslationDeltaCostFunction(translation_weight, target_translation));
  problem.AddResidualBlock(owned.back().get(), {tb});
  owned.emplace_back(new RotationDeltaCostFunction(rotation_weight, initial_pose + 3));
  problem.AddResidualBlock(owned.back().get(), {qb});
  hg::LmOptions options;
  options.max_num_iterations = max_num_iterations;
  options.use_nonmonotonic_steps = use_nonmonotonic_steps != 0;
  std::vector<hg::LmIteration> trace;
  if (trace_out != nullptr) options.trace = &trace;
  const hg::LmSummary s = problem.Solve(options);
  for (size_t i = 0; i < trace.size() && static_cast<int>(i) < trace_capacity; ++i) {
    double* o = trace_out + 7 * i;
    o[0] = trace[i].iteration; o[1] = trace[i].cost; o[2] = trace[i].candidate_cost;
    o[3] = trace[i].model_cost_change; o[4] = trace[i].relative_decrease; o[5] = trace[i].radius;
    o[6] = trace[i].successful ? 1.0 : 0.0;
  }
  for (int k = 0; k < 3; ++k) pose_out[k] = t[k];
  for (int k = 0; k < 4; ++k) pose_out[3 + k] = q[k];
  if (summary_out != nullptr) {
    summary_out[0] = s.initial_cost;
    summary_out[1] = s.final_cost;
    summary_out[2] = s.num_iterations;
    summary_out[3] = s.num_successful_steps;
    summary_out[4] = s.num_cost_evaluations;
    summary_out[5] = s.num_jacobian_evaluations;
  }
  return s.termination;
}

// ---- CeresScanMatcher3D::Evaluate (ceres_scan_matcher_3d.cc:97-118): cost, residuals and the gradient
// in the tangent space (the reference passes its 'jacobians' vector as Problem::Evaluate's gradient).
int orc_scan_match_evaluate(int num_clouds, const float* const* clouds, const int* cloud_sizes, void* const* grids,
                            const double* occupied_space_weights, double translation_weight, double rotation_weight,
                            int only_optimize_yaw, const double* target_translation, const double* initial_pose,
                            double* cost, double* residuals, double* gradient) {
  double t[3] = {initial_pose[0], initial_pose[1], initial_pose[2]};
  double q[4] = {initial_pose[3], initial_pose[4], initial_pose[5], initial_pose[6]};
  hg::DenseLmProblem<CostFunction> problem;
  const int tb = problem.AddParameterBlock(t, 3, hg::LocalParameterization::kIdentity);
  const int qb = problem.AddParameterBlock(
      q, 4, only_optimize_yaw ? hg::LocalParameterization::kYawOnlyQuaternion : hg::LocalParameterization::kQuaternion);
  std::vector<std::unique_ptr<CostFunction>> owned;
  for (int i = 0; i < num_clouds; ++i) {
    auto* h = static_cast<GridHandle*>(grids[i]);
    CloudView cloud;
    cloud.data = clouds[i];
    cloud.n = cloud_sizes[i];
    cloud.stride = 3;
    const double scaling = occupied_space_weights[i] / std::sqrt(static_cast<double>(cloud.n));
    if (h->prob) {
      owned.emplace_back(new ScanCostFunction(kOccupiedSpace, scaling, cloud, h->prob.get(), {}, 0.));
    } else {
      owned.emplace_back(new ScanCostFunction(kTSDFSpace, scaling, cloud, nullptr, {h->tsdf.get()}, 0.));
    }
    problem.AddResidualBlock(owned.back().get(), {tb, qb});
  }
  owned.emplace_back(new TranslationDeltaCostFunction(translation_weight, target_translation));
  problem.AddResidualBlock(owned.back().get(), {tb});
  owned.emplace_back(new RotationDeltaCostFunction(rotation_weight, initial_pose + 3));
  problem.AddResidualBlock(owned.back().get(), {qb});
  double c = 0.0;
  std::vector<double> r, J;
  if (!problem.Evaluate(&c, &r, &J)) return 1;
  *cost = c;
  std::copy(r.begin(), r.end(), residuals);
  const int n = problem.NumLocalParameters();
  for (int k = 0; k < n; ++k) gradient[k] = 0.0;
  for (size_t i = 0; i < r.size(); ++i)
    for (int k = 0; k < n; ++k) gradient[k] += J[i * n + k] * r[i];
  return 0;
}

// ---- precomputation stack ---------------------------------------------------
void* orc_stack_new(void* grid, int branch_and_bound_depth, int full_resolution_depth) {
  auto* h = static_cast<GridHandle*>(grid);
  FastCorrelativeScanMatcherOptions3D o;
  o.branch_and_bound_depth = branch_and_bound_depth;
  o.full_resolution_depth = full_resolution_depth;
  auto* s = new StackHandle;
  s->stack.reset(new PrecomputationGridStack3D(*h->iface(), o));
  return s;
}
void orc_stack_free(void* s) { delete static_cast<StackHandle*>(s); }
long long orc_stack_num_voxels(void* s, int depth) {
  long long n = 0;
  static_cast<StackHandle*>(s)->stack->Get(depth).ForEach([&n](const Array3i&, const uint8) { ++n; });
  return n;
}
long long orc_stack_export(void* s, int depth, int32_t* xyz, uint8_t* values) {
  long long n = 0;
  static_cast<StackHandle*>(s)->stack->Get(depth).ForEach([&](const Array3i& i, const uint8 v) {
    xyz[3 * n] = i.x; xyz[3 * n + 1] = i.y; xyz[3 * n + 2] = i.z;
    values[n] = v;
    ++n;
  });
  return n;
}
void orc_stack_lookup(void* s, int depth, int n, const int32_t* xyz, uint8_t* out) {
  const PrecomputationGrid3D& g = static_cast<StackHandle*>(s)->stack->Get(depth);
  for (int i = 0; i < n; ++i) out[i] = g.value(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
}
// Single PrecomputeGrid step exposed for the reference's known-answer test
// (precomputation_grid_3d_test.cc:30-77): convert, then 'levels' full-resolution
// steps with shift 2^(d-1); looks up 'n' cells at the last level.
void orc_precompute_fullres_lookup(void* grid, int levels, int n, const int32_t* xyz, uint8_t* out) {
  auto* h = static_cast<GridHandle*>(grid);
  std::vector<PrecomputationGrid3D> gs;
  gs.push_back(h->prob ? ConvertToPrecomputationGrid(*h->prob)
                       : ConvertToPrecomputationGrid(*h->tsdf));
  for (int d = 1; d <= levels; ++d) {
    const int s = 1 << (d - 1);
    gs.push_back(PrecomputeGrid(gs.back(), false, Array3i(s, s, s)));
  }
  const PrecomputationGrid3D& g = gs.back();
  for (int i = 0; i < n; ++i) out[i] = g.value(Array3i(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
}
float orc_precomp_to_probability(float v) { return PrecomputationGrid3D::ToProbability(v); }

// ---- rotational matcher -------------------------------------------------------
void orc_rotational_match(const float* submap_hist, const float* scan_hist, int size,
                          float initial_angle, const float* angles, int num_angles, float* out) {
  const std::vector<float> a(submap_hist, submap_hist + size), b(scan_hist, scan_hist + size);
  const std::vector<float> r =
      RotationalMatch(a, b, initial_angle, std::vector<float>(angles, angles + num_angles));
  std::memcpy(out, r.data(), r.size() * sizeof(float));
}
void orc_rotate_histogram(const float* hist, int size, float angle, float* out) {
  const std::vector<float> r = RotateHistogram(std::vector<float>(hist, hist + size), angle);
  std::memcpy(out, r.data(), r.size() * sizeof(float));
}

// ---- fast correlative scan matcher ---------------------------------------------
void* orc_fcsm_new(void* grid, void* low_resolution_grid, const float* submap_histogram,
                   int histogram_size, int branch_and_bound_depth, int full_resolution_depth,
                   int use_rotational_scan_matcher, double min_rotational_score,
                   double min_low_resolution_score, double linear_xy_search_window,
                   double linear_z_search_window, double angular_search_window) {
  FastCorrelativeScanMatcherOptions3D o;
  o.branch_and_bound_depth = branch_and_bound_depth;
  o.full_resolution_depth = full_resolution_depth;
  o.use_rotational_scan_matcher = use_rotational_scan_matcher != 0;
  o.min_rotational_score = min_rotational_score;
  o.min_low_resolution_score = min_low_resolution_score;
  o.linear_xy_search_window = linear_xy_search_window;
  o.linear_z_search_window = linear_z_search_window;
  o.angular_search_window = angular_search_window;
  auto* f = new FcsmHandle;
  f->histogram.assign(submap_histogram, submap_histogram + histogram_size);
  f->m.reset(new FastCorrelativeScanMatcher3D(
      *static_cast<GridHandle*>(grid)->iface(), static_cast<GridHandle*>(low_resolution_grid)->iface(),
      &f->histogram, o));
  return f;
}
void orc_fcsm_free(void* f) { delete static_cast<FcsmHandle*>(f); }

static NodeData MakeNodeData(const float* high, int n_high, const float* low, int n_low,
                             const float* hist, int n_hist, const double* gravity) {
  NodeData d;
  d.high_resolution_point_cloud = ToCloud(high, n_high, 3);
  d.low_resolution_point_cloud = ToCloud(low, n_low, 3);
  d.rotational_scan_matcher_histogram.assign(hist, hist + n_hist);
  for (int k = 0; k < 4; ++k) d.gravity_alignment[k] = gravity[k];
  return d;
}
static void PackResult(const MatchResult& r, double* out) {
  // found, score, t3, q4, rot_score, lowres_score, scan_index, ox, oy, oz, num_scored, num_scans
  out[0] = r.found ? 1.0 : 0.0;
  out[1] = r.score;
  for (int k = 0; k < 3; ++k) out[2 + k] = r.pose_t[k];
  for (int k = 0; k < 4; ++k) out[5 + k] = r.pose_q[k];
  out[9] = r.rotational_score;
  out[10] = r.low_resolution_score;
  out[11] = r.best_scan_index;
  out[12] = r.best_offset.x;
  out[13] = r.best_offset.y;
  out[14] = r.best_offset.z;
  out[15] = static_cast<double>(r.num_candidates_scored);
  out[16] = r.num_discrete_scans;
}
// full = 0: Match(node pose, submap pose); full = 1: MatchFullSubmap(rotations only).
void orc_fcsm_match(void* f, int full, const double* node_pose /* t3 q4 */,
                    const double* submap_pose /* t3 q4 */, const float* high, int n_high,
                    const float* low, int n_low, const float* hist, int n_hist,
                    const double* gravity, float min_score, double* out17) {
  auto* h = static_cast<FcsmHandle*>(f);
  const NodeData d = MakeNodeData(high, n_high, low, n_low, hist, n_hist, gravity);
  MatchResult r;
  if (full) {
    r = h->m->MatchFullSubmap(node_pose + 3, submap_pose + 3, d, min_score);
  } else {
    r = h->m->Match(node_pose, node_pose + 3, submap_pose, submap_pose + 3, d, min_score);
  }
  PackResult(r, out17);
}

// Discretisation stage for parity: window given directly. Returns the number of
// discrete scans S; fills poses (S x 7 floats: t3, q wxyz), rotational scores
// (S) and cells (S x D x N x 3) when the buffers are non-null and large enough.
int orc_fcsm_discrete_scans(void* f, int wxy, int wz, double angular_window,
                            const double* node_pose, const double* submap_pose, const float* high,
                            int n_high, const float* hist, int n_hist, const double* gravity,
                            int max_scans, float* poses, float* rot_scores, int32_t* cells) {
  auto* h = static_cast<FcsmHandle*>(f);
  const NodeData d = MakeNodeData(high, n_high, high, 0, hist, n_hist, gravity);
  const FastCorrelativeScanMatcher3D::SearchParameters sp{wxy, wz, angular_window};
  const auto scans = h->m->GenerateDiscreteScans(
      sp, d, FastCorrelativeScanMatcher3D::ToRigidf(node_pose, node_pose + 3),
      FastCorrelativeScanMatcher3D::ToRigidf(submap_pose, submap_pose + 3));
  const int S = static_cast<int>(scans.size());
  if (S > max_scans) return S;
  for (int s = 0; s < S; ++s) {
    if (poses != nullptr) {
      poses[7 * s] = scans[s].pose.translation.x;
      poses[7 * s + 1] = scans[s].pose.translation.y;
      poses[7 * s + 2] = scans[s].pose.translation.z;
      poses[7 * s + 3] = scans[s].pose.rotation.w;
      poses[7 * s + 4] = scans[s].pose.rotation.x;
      poses[7 * s + 5] = scans[s].pose.rotation.y;
      poses[7 * s + 6] = scans[s].pose.rotation.z;
    }
    if (rot_scores != nullptr) rot_scores[s] = scans[s].rotational_score;
    if (cells != nullptr) {
      const int D = static_cast<int>(scans[s].cell_indices_per_depth.size());
      for (int dd = 0; dd < D; ++dd)
        for (int i = 0; i < n_high; ++i) {
          const Array3i& c = scans[s].cell_indices_per_depth[dd][i];
          int32_t* o = cells + ((static_cast<size_t>(s) * D + dd) * n_high + i) * 3;
          o[0] = c.x; o[1] = c.y; o[2] = c.z;
        }
    }
  }
  return S;
}

// Scores explicit candidates (scan, ox, oy, oz) at 'depth' without sorting;
// writes integer sums and float scores. Returns seconds spent scoring
// (used as the CPU baseline for candidates/s); 'threads' > 1 splits the
// candidate list across std::threads.
double orc_fcsm_score_candidates(void* f, int wxy, int wz, double angular_window,
                                 const double* node_pose, const double* submap_pose,
                                 const float* high, int n_high, const float* hist, int n_hist,
                                 const double* gravity, int depth, long long num_candidates,
                                 const int32_t* cand /* n x 4 */, int32_t* sums, float* scores,
                                 int threads) {
  auto* h = static_cast<FcsmHandle*>(f);
  const NodeData d = MakeNodeData(high, n_high, high, 0, hist, n_hist, gravity);
  const FastCorrelativeScanMatcher3D::SearchParameters sp{wxy, wz, angular_window};
  const auto scans = h->m->GenerateDiscreteScans(
      sp, d, FastCorrelativeScanMatcher3D::ToRigidf(node_pose, node_pose + 3),
      FastCorrelativeScanMatcher3D::ToRigidf(submap_pose, submap_pose + 3));
  if (threads < 1) threads = 1;
  auto work = [&](long long begin, long long end) {
    std::vector<Candidate3D> c;
    c.reserve(end - begin);
    for (long long i = begin; i < end; ++i)
      c.emplace_back(cand[4 * i], Array3i(cand[4 * i + 1], cand[4 * i + 2], cand[4 * i + 3]));
    h->m->ScoreCandidates(depth, scans, &c, /*sort=*/false);
    for (long long i = begin; i < end; ++i) {
      if (sums != nullptr) sums[i] = c[i - begin].sum;
      if (scores != nullptr) scores[i] = c[i - begin].score;
    }
  };
  const auto t0 = std::chrono::steady_clock::now();
  if (threads == 1) {
    work(0, num_candidates);
  } else {
    std::vector<std::thread> pool;
    const long long chunk = (num_candidates + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
      const long long b = t * chunk, e = std::min(num_candidates, b + chunk);
      if (b < e) pool.emplace_back(work, b, e);
    }
    for (auto& t : pool) t.join();
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// ---- voxel filters (row N2) -----------------------------------------------------
// Returns the number of kept points; writes their indices (input order) to 'kept' (capacity n).
int orc_voxel_filter(float size, const float* pts, int stride, int n, int32_t* kept) {
  std::vector<int> all(n);
  for (int i = 0; i < n; ++i) all[i] = i;
  const std::vector<int> r = VoxelFilterIndices(size, pts, stride, all);
  for (size_t i = 0; i < r.size(); ++i) kept[i] = r[i];
  return static_cast<int>(r.size());
}
int orc_adaptive_voxel_filter(float max_length, int min_num_points, float max_range, const float* pts,
                              int stride, int n, int32_t* kept) {
  const std::vector<int> r = AdaptiveVoxelFilterIndices(max_length, min_num_points, max_range, pts, stride, n);
  for (size_t i = 0; i < r.size(); ++i) kept[i] = r[i];
  return static_cast<int>(r.size());
}

float orc_low_resolution_score(void* low_grid, const float* pts, int n, const float* pose /* t3 q wxyz */) {
  auto* h = static_cast<GridHandle*>(low_grid);
  const Rigid3f p(Vec3f(pose[0], pose[1], pose[2]), Quatf(pose[3], pose[4], pose[5], pose[6]));
  return LowResolutionScore(h->iface(), ToCloud(pts, n, 3), p);
}

}  // extern "C"
