// hg_dense_lm.h — small dense Levenberg-Marquardt solver for pose problems.
//
// Host-side stand-in for what the reference asks of Ceres in
// CeresScanMatcher3D::Match (cartographer/mapping/internal/3d/scan_matching/
// ceres_scan_matcher_3d.cc:65-95): a trust-region LM solve with a dense linear
// solver over a handful of 3- and 4-sized parameter blocks, the quaternion
// blocks moving on their 3-dimensional (or yaw-only, 1-dimensional) tangent
// space. Ceres itself (pinned 1.13.0, scripts/install_ceres.sh:20) is a
// third-party dependency that is not vendored in the reference; where a real
// Ceres is linked, the cost functions of this library plug into it directly
// (they implement ceres::CostFunction::Evaluate) and this header is unused.
//
// The loop follows Ceres' published TrustRegionMinimizer + LevenbergMarquardt
// strategy: Jacobi column scaling 1/(1+||J_j||), damping diag(J'J)/radius
// clamped to [1e-6, 1e32], step quality rho = cost change / model cost change
// (optionally the non-monotonic variant), radius /= max(1/3, 1-(2 rho-1)^3)
// on success, radius /= 2^k on failure, and Ceres' default function /
// gradient / parameter tolerances. The linear step solves the damped normal
// equations by Cholesky instead of QR of the augmented Jacobian (same minimiser).
//
// Templated on the cost-function type so that the same solver drives the GPU
// cost functions of this library and (in tests) the CPU oracle's.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <limits>
#include <vector>

namespace hg {

enum class LocalParameterization {
  kIdentity = 0,
  // ceres::QuaternionParameterization: x_plus_delta = exp(delta) * x, (w,x,y,z).
  kQuaternion = 1,
  // cartographer/mapping/internal/3d/rotation_parameterization.h:27-39.
  kYawOnlyQuaternion = 2,
};

struct LmOptions {
  int max_num_iterations = 10;
  bool use_nonmonotonic_steps = false;
  int max_consecutive_nonmonotonic_steps = 5;
  double initial_trust_region_radius = 1e4;
  double max_trust_region_radius = 1e16;
  double min_trust_region_radius = 1e-32;
  double min_relative_decrease = 1e-3;
  double min_lm_diagonal = 1e-6;
  double max_lm_diagonal = 1e32;
  double function_tolerance = 1e-6;
  double gradient_tolerance = 1e-10;
  double parameter_tolerance = 1e-8;
};

struct LmSummary {
  double initial_cost = 0.0;
  double final_cost = 0.0;
  int num_iterations = 0;        // linear solves attempted
  int num_successful_steps = 0;
  int num_cost_evaluations = 0;  // residual-only + residual+Jacobian
  int num_jacobian_evaluations = 0;
  int termination = 0;  // 0 max iterations, 1 function tol, 2 gradient tol, 3 parameter tol, 4 failure
};

template <typename CostFunctionT>
class DenseLmProblem {
 public:
  // Returns the block id. 'values' stays owned by the caller and is updated in place.
  int AddParameterBlock(double* values, int size, LocalParameterization p) {
    Block b;
    b.values = values;
    b.size = size;
    b.param = p;
    b.local_size = (p == LocalParameterization::kIdentity)
                       ? size
                       : (p == LocalParameterization::kQuaternion ? 3 : 1);
    b.constant = false;
    blocks_.push_back(b);
    return static_cast<int>(blocks_.size()) - 1;
  }
  void SetParameterBlockConstant(int id) { blocks_[id].constant = true; }
  // The cost function is not owned.
  void AddResidualBlock(const CostFunctionT* f, const std::vector<int>& block_ids) {
    Residual r;
    r.f = f;
    r.ids = block_ids;
    residuals_.push_back(r);
  }

  int NumResiduals() const {
    int m = 0;
    for (const Residual& r : residuals_) m += r.f->num_residuals();
    return m;
  }

  // Evaluates cost = 1/2 |r|^2; optionally residuals and the dense row-major
  // Jacobian w.r.t. the local (tangent) parameters of the non-constant blocks.
  bool Evaluate(double* cost, std::vector<double>* residuals, std::vector<double>* jacobian) {
    Layout();
    const int m = NumResiduals();
    std::vector<double> r(m);
    std::vector<double> J;
    if (jacobian != nullptr) J.assign(static_cast<size_t>(m) * num_local_, 0.0);
    if (!EvaluateInternal(r.data(), jacobian != nullptr ? J.data() : nullptr)) return false;
    double c = 0.0;
    for (int i = 0; i < m; ++i) c += r[i] * r[i];
    *cost = 0.5 * c;
    if (residuals != nullptr) *residuals = r;
    if (jacobian != nullptr) *jacobian = J;
    return true;
  }
  int NumLocalParameters() {
    Layout();
    return num_local_;
  }

  LmSummary Solve(const LmOptions& options) {
    LmSummary summary;
    Layout();
    const int m = NumResiduals();
    const int n = num_local_;
    std::vector<double> r(m), J(static_cast<size_t>(m) * n), r_new(m);
    std::vector<double> g(n), scale(n), JtJ(static_cast<size_t>(n) * n), A(static_cast<size_t>(n) * n);
    std::vector<double> delta(n), x_backup;

    if (!EvaluateInternal(r.data(), J.data())) {
      summary.termination = 4;
      return summary;
    }
    ++summary.num_cost_evaluations;
    ++summary.num_jacobian_evaluations;
    double cost = HalfSquaredNorm(r);
    summary.initial_cost = cost;
    summary.final_cost = cost;

    // Jacobi scaling, computed once from the initial Jacobian (Ceres).
    for (int j = 0; j < n; ++j) {
      double s = 0.0;
      for (int i = 0; i < m; ++i) s += J[static_cast<size_t>(i) * n + j] * J[static_cast<size_t>(i) * n + j];
      scale[j] = 1.0 / (1.0 + std::sqrt(s));
    }
    auto build_normal = [&]() {
      for (int i = 0; i < m; ++i)
        for (int j = 0; j < n; ++j) J[static_cast<size_t>(i) * n + j] *= scale[j];
      std::fill(JtJ.begin(), JtJ.end(), 0.0);
      std::fill(g.begin(), g.end(), 0.0);
      for (int i = 0; i < m; ++i) {
        const double* row = &J[static_cast<size_t>(i) * n];
        for (int a = 0; a < n; ++a) {
          g[a] += row[a] * r[i];
          for (int b = a; b < n; ++b) JtJ[static_cast<size_t>(a) * n + b] += row[a] * row[b];
        }
      }
      for (int a = 0; a < n; ++a)
        for (int b = 0; b < a; ++b) JtJ[static_cast<size_t>(a) * n + b] = JtJ[static_cast<size_t>(b) * n + a];
    };
    build_normal();
    auto max_abs_unscaled_gradient = [&]() {
      double v = 0.0;
      for (int a = 0; a < n; ++a) v = std::max(v, std::fabs(g[a] / scale[a]));
      return v;
    };
    if (max_abs_unscaled_gradient() <= options.gradient_tolerance) {
      summary.termination = 2;
      return summary;
    }

    double radius = options.initial_trust_region_radius;
    double decrease_factor = 2.0;
    // Non-monotonic bookkeeping (Ceres trust_region_minimizer).
    double reference_cost = cost, candidate_cost = cost, minimum_cost = cost;
    double accumulated_reference_model_cost_change = 0.0;
    double accumulated_candidate_model_cost_change = 0.0;
    int num_consecutive_nonmonotonic_steps = 0;
    std::vector<double> x_min;
    SaveState(&x_min);

    for (int iter = 0; iter < options.max_num_iterations; ++iter) {
      ++summary.num_iterations;
      // Damped normal equations.
      A = JtJ;
      for (int a = 0; a < n; ++a) {
        const double d = std::min(std::max(JtJ[static_cast<size_t>(a) * n + a], options.min_lm_diagonal),
                                  options.max_lm_diagonal);
        A[static_cast<size_t>(a) * n + a] += d / radius;
      }
      bool step_ok = CholeskySolve(A, g, n, &delta);
      double model_cost_change = 0.0;
      if (step_ok) {
        for (int a = 0; a < n; ++a) delta[a] = -delta[a];
        // -(J d)'(r + J d / 2) = -(g'd + d'JtJ d / 2)
        double gd = 0.0, dJd = 0.0;
        for (int a = 0; a < n; ++a) {
          gd += g[a] * delta[a];
          double t = 0.0;
          for (int b = 0; b < n; ++b) t += JtJ[static_cast<size_t>(a) * n + b] * delta[b];
          dJd += delta[a] * t;
        }
        model_cost_change = -(gd + 0.5 * dJd);
        if (!(model_cost_change > 0.0)) step_ok = false;
      }
      if (!step_ok) {
        radius /= decrease_factor;
        decrease_factor *= 2.0;
        if (radius < options.min_trust_region_radius) {
          summary.termination = 4;
          break;
        }
        continue;
      }
      // Candidate.
      SaveState(&x_backup);
      double step_norm2 = 0.0, x_norm2 = 0.0;
      for (int a = 0; a < n; ++a) step_norm2 += (delta[a] * scale[a]) * (delta[a] * scale[a]);
      for (const Block& b : blocks_)
        for (int k = 0; k < b.size; ++k) x_norm2 += b.values[k] * b.values[k];
      ApplyDelta(delta, scale);
      const bool eval_ok = EvaluateInternal(r_new.data(), nullptr);
      ++summary.num_cost_evaluations;
      const double new_cost = eval_ok ? HalfSquaredNorm(r_new) : std::numeric_limits<double>::max();
      double relative_decrease = (cost - new_cost) / model_cost_change;
      if (options.use_nonmonotonic_steps) {
        const double historical =
            (reference_cost - new_cost) / (accumulated_reference_model_cost_change + model_cost_change);
        relative_decrease = std::max(relative_decrease, historical);
      }
      if (eval_ok && relative_decrease > options.min_relative_decrease) {
        // Accept.
        ++summary.num_successful_steps;
        const double cost_change = cost - new_cost;
        const double old_cost = cost;
        cost = new_cost;
        if (!EvaluateInternal(r.data(), J.data())) {
          summary.termination = 4;
          break;
        }
        ++summary.num_jacobian_evaluations;
        build_normal();
        radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * relative_decrease - 1.0, 3));
        radius = std::min(options.max_trust_region_radius, radius);
        decrease_factor = 2.0;
        if (options.use_nonmonotonic_steps) {
          accumulated_candidate_model_cost_change += model_cost_change;
          accumulated_reference_model_cost_change += model_cost_change;
          if (cost < minimum_cost) {
            minimum_cost = cost;
            SaveState(&x_min);
            num_consecutive_nonmonotonic_steps = 0;
            candidate_cost = cost;
            accumulated_candidate_model_cost_change = 0.0;
          } else {
            ++num_consecutive_nonmonotonic_steps;
            if (cost > candidate_cost) {
              candidate_cost = cost;
              accumulated_candidate_model_cost_change = 0.0;
            }
          }
          if (num_consecutive_nonmonotonic_steps == options.max_consecutive_nonmonotonic_steps) {
            reference_cost = candidate_cost;
            accumulated_reference_model_cost_change = accumulated_candidate_model_cost_change;
          }
        } else {
          minimum_cost = cost;
        }
        if (max_abs_unscaled_gradient() <= options.gradient_tolerance) {
          summary.termination = 2;
          break;
        }
        if (std::fabs(cost_change) <= options.function_tolerance * old_cost) {
          summary.termination = 1;
          break;
        }
        if (std::sqrt(step_norm2) <=
            options.parameter_tolerance * (std::sqrt(x_norm2) + options.parameter_tolerance)) {
          summary.termination = 3;
          break;
        }
      } else {
        RestoreState(x_backup);
        radius /= decrease_factor;
        decrease_factor *= 2.0;
        if (radius < options.min_trust_region_radius) {
          summary.termination = 4;
          break;
        }
        if (std::sqrt(step_norm2) <=
            options.parameter_tolerance * (std::sqrt(x_norm2) + options.parameter_tolerance)) {
          summary.termination = 3;
          break;
        }
      }
    }
    if (options.use_nonmonotonic_steps && minimum_cost < cost) {
      // Ceres reports the best point seen.
      RestoreState(x_min);
      cost = minimum_cost;
    }
    summary.final_cost = cost;
    return summary;
  }

 private:
  struct Block {
    double* values;
    int size;
    int local_size;
    int local_offset;
    LocalParameterization param;
    bool constant;
  };
  struct Residual {
    const CostFunctionT* f;
    std::vector<int> ids;
  };

  void Layout() {
    int off = 0;
    for (Block& b : blocks_) {
      b.local_offset = b.constant ? -1 : off;
      if (!b.constant) off += b.local_size;
    }
    num_local_ = off;
  }

  static double HalfSquaredNorm(const std::vector<double>& r) {
    double c = 0.0;
    for (double v : r) c += v * v;
    return 0.5 * c;
  }

  void SaveState(std::vector<double>* x) const {
    x->clear();
    for (const Block& b : blocks_)
      for (int k = 0; k < b.size; ++k) x->push_back(b.values[k]);
  }
  void RestoreState(const std::vector<double>& x) {
    size_t i = 0;
    for (Block& b : blocks_)
      for (int k = 0; k < b.size; ++k) b.values[k] = x[i++];
  }

  // (w,x,y,z) Hamilton product z * w.
  static void QuatProduct(const double* z, const double* w, double* zw) {
    zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
    zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
    zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
    zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
  }

  void ApplyDelta(const std::vector<double>& delta, const std::vector<double>& scale) {
    for (Block& b : blocks_) {
      if (b.constant) continue;
      const double* d = &delta[b.local_offset];
      const double* s = &scale[b.local_offset];
      if (b.param == LocalParameterization::kIdentity) {
        for (int k = 0; k < b.size; ++k) b.values[k] += d[k] * s[k];
      } else if (b.param == LocalParameterization::kQuaternion) {
        const double dx = d[0] * s[0], dy = d[1] * s[1], dz = d[2] * s[2];
        const double norm = std::sqrt(dx * dx + dy * dy + dz * dz);
        if (norm > 0.0) {
          const double sin_by_norm = std::sin(norm) / norm;
          const double q_delta[4] = {std::cos(norm), sin_by_norm * dx, sin_by_norm * dy, sin_by_norm * dz};
          double out[4];
          QuatProduct(q_delta, b.values, out);
          for (int k = 0; k < 4; ++k) b.values[k] = out[k];
        }
      } else {
        double c = d[0] * s[0];
        c = c > 0.5 ? 0.5 : (c < -0.5 ? -0.5 : c);
        const double q_delta[4] = {std::sqrt(1. - c * c), 0., 0., c};
        double out[4];
        QuatProduct(q_delta, b.values, out);
        for (int k = 0; k < 4; ++k) b.values[k] = out[k];
      }
    }
  }

  // d(x (+) delta)/d delta at delta = 0, row-major size x local_size.
  static void PlusJacobian(const Block& b, double* P) {
    const double* x = b.values;
    if (b.param == LocalParameterization::kQuaternion) {
      // ceres::QuaternionParameterization::ComputeJacobian
      P[0] = -x[1]; P[1] = -x[2]; P[2] = -x[3];
      P[3] = x[0];  P[4] = x[3];  P[5] = -x[2];
      P[6] = -x[3]; P[7] = x[0];  P[8] = x[1];
      P[9] = x[2];  P[10] = -x[1]; P[11] = x[0];
    } else {
      // yaw-only: q_delta = (sqrt(1-c^2), 0, 0, c); d/dc at 0 of q_delta * x.
      P[0] = -x[3];
      P[1] = -x[2];
      P[2] = x[1];
      P[3] = x[0];
    }
  }

  bool EvaluateInternal(double* r, double* J) {
    const int n = num_local_;
    int row0 = 0;
    if (J != nullptr) std::fill(J, J + static_cast<size_t>(NumResiduals()) * n, 0.0);
    std::vector<std::vector<double>> jac_store;
    std::vector<double*> jac_ptrs;
    std::vector<const double*> params;
    for (const Residual& res : residuals_) {
      const int nr = res.f->num_residuals();
      params.clear();
      for (int id : res.ids) params.push_back(blocks_[id].values);
      if (J == nullptr) {
        if (!res.f->Evaluate(params.data(), r + row0, nullptr)) return false;
      } else {
        jac_store.resize(res.ids.size());
        jac_ptrs.resize(res.ids.size());
        for (size_t k = 0; k < res.ids.size(); ++k) {
          const Block& b = blocks_[res.ids[k]];
          if (b.constant) {
            jac_ptrs[k] = nullptr;
          } else {
            jac_store[k].assign(static_cast<size_t>(nr) * b.size, 0.0);
            jac_ptrs[k] = jac_store[k].data();
          }
        }
        if (!res.f->Evaluate(params.data(), r + row0, jac_ptrs.data())) return false;
        for (size_t k = 0; k < res.ids.size(); ++k) {
          const Block& b = blocks_[res.ids[k]];
          if (b.constant) continue;
          const double* Jb = jac_store[k].data();
          if (b.param == LocalParameterization::kIdentity) {
            for (int i = 0; i < nr; ++i)
              for (int c = 0; c < b.size; ++c)
                J[static_cast<size_t>(row0 + i) * n + b.local_offset + c] += Jb[static_cast<size_t>(i) * b.size + c];
          } else {
            double P[12];
            PlusJacobian(b, P);
            for (int i = 0; i < nr; ++i)
              for (int c = 0; c < b.local_size; ++c) {
                double s = 0.0;
                for (int a = 0; a < b.size; ++a) s += Jb[static_cast<size_t>(i) * b.size + a] * P[a * b.local_size + c];
                J[static_cast<size_t>(row0 + i) * n + b.local_offset + c] += s;
              }
          }
        }
      }
      row0 += nr;
    }
    return true;
  }

  static bool CholeskySolve(std::vector<double> A, const std::vector<double>& b, int n,
                            std::vector<double>* x) {
    // In-place lower Cholesky.
    for (int j = 0; j < n; ++j) {
      double d = A[static_cast<size_t>(j) * n + j];
      for (int k = 0; k < j; ++k) d -= A[static_cast<size_t>(j) * n + k] * A[static_cast<size_t>(j) * n + k];
      if (!(d > 0.0) || !std::isfinite(d)) return false;
      const double l = std::sqrt(d);
      A[static_cast<size_t>(j) * n + j] = l;
      for (int i = j + 1; i < n; ++i) {
        double s = A[static_cast<size_t>(i) * n + j];
        for (int k = 0; k < j; ++k) s -= A[static_cast<size_t>(i) * n + k] * A[static_cast<size_t>(j) * n + k];
        A[static_cast<size_t>(i) * n + j] = s / l;
      }
    }
    std::vector<double> y(n);
    for (int i = 0; i < n; ++i) {
      double s = b[i];
      for (int k = 0; k < i; ++k) s -= A[static_cast<size_t>(i) * n + k] * y[k];
      y[i] = s / A[static_cast<size_t>(i) * n + i];
    }
    x->assign(n, 0.0);
    for (int i = n - 1; i >= 0; --i) {
      double s = y[i];
      for (int k = i + 1; k < n; ++k) s -= A[static_cast<size_t>(k) * n + i] * (*x)[k];
      (*x)[i] = s / A[static_cast<size_t>(i) * n + i];
    }
    return true;
  }

  std::vector<Block> blocks_;
  std::vector<Residual> residuals_;
  int num_local_ = 0;
};

}  // namespace hg
