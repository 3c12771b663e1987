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
// Only J'J, J'r and the cost are ever needed, so a cost function that can produce
// those directly (member NormalEquations — the GPU blocks do, without
// materialising their Jacobian) is asked for them instead of for rows.
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
    const int n = num_local_;
    std::vector<double> H(static_cast<size_t>(n) * n), g(n), scale(n);
    std::vector<double> Hs(static_cast<size_t>(n) * n), gs(n), A(static_cast<size_t>(n) * n);
    std::vector<double> delta(n), x_backup;

    double cost = 0.0;
    if (!EvaluateNormal(&cost, H.data(), g.data())) {
      summary.termination = 4;
      return summary;
    }
    ++summary.num_cost_evaluations;
    ++summary.num_jacobian_evaluations;
    summary.initial_cost = cost;
    summary.final_cost = cost;

    // Jacobi scaling, computed once from the initial Jacobian's column norms (Ceres).
    for (int j = 0; j < n; ++j) scale[j] = 1.0 / (1.0 + std::sqrt(std::max(0.0, H[static_cast<size_t>(j) * n + j])));
    auto apply_scaling = [&]() {
      for (int a = 0; a < n; ++a) {
        gs[a] = g[a] * scale[a];
        for (int b = 0; b < n; ++b) Hs[static_cast<size_t>(a) * n + b] = H[static_cast<size_t>(a) * n + b] * scale[a] * scale[b];
      }
    };
    apply_scaling();
    auto max_abs_gradient = [&]() {
      double v = 0.0;
      for (int a = 0; a < n; ++a) v = std::max(v, std::fabs(g[a]));
      return v;
    };
    if (max_abs_gradient() <= options.gradient_tolerance) {
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
      A = Hs;
      for (int a = 0; a < n; ++a) {
        const double d = std::min(std::max(Hs[static_cast<size_t>(a) * n + a], options.min_lm_diagonal),
                                  options.max_lm_diagonal);
        A[static_cast<size_t>(a) * n + a] += d / radius;
      }
      bool step_ok = CholeskySolve(A, gs, n, &delta);
      double model_cost_change = 0.0;
      if (step_ok) {
        for (int a = 0; a < n; ++a) delta[a] = -delta[a];
        // -(J d)'(r + J d / 2) = -(g'd + d'J'J d / 2)
        double gd = 0.0, dJd = 0.0;
        for (int a = 0; a < n; ++a) {
          gd += gs[a] * delta[a];
          double t = 0.0;
          for (int b = 0; b < n; ++b) t += Hs[static_cast<size_t>(a) * n + b] * delta[b];
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
      double new_cost = 0.0;
      const bool eval_ok = EvaluateCost(&new_cost);
      ++summary.num_cost_evaluations;
      if (!eval_ok) new_cost = std::numeric_limits<double>::max();
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
        if (!EvaluateNormal(&cost, H.data(), g.data())) {
          summary.termination = 4;
          break;
        }
        ++summary.num_jacobian_evaluations;
        apply_scaling();
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
        if (max_abs_gradient() <= options.gradient_tolerance) {
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

  // A cost function may offer its Gauss-Newton terms directly (ambient
  // parameters, block order): bool NormalEquations(parameters, JtJ, Jtr, cost)
  // returning false when it does not. Detected at compile time.
  template <typename F>
  static auto TryNormalEquations(const F* f, double const* const* p, double* JtJ, double* Jtr, double* c, int)
      -> decltype(f->NormalEquations(p, JtJ, Jtr, c), bool()) {
    return f->NormalEquations(p, JtJ, Jtr, c);
  }
  template <typename F>
  static bool TryNormalEquations(const F*, double const* const*, double*, double*, double*, long) {
    return false;
  }

  // cost = 1/2 |r|^2, H = J'J and g = J'r in the local (tangent) coordinates of
  // the non-constant blocks (row-major n x n, n).
  bool EvaluateNormal(double* cost, double* H, double* g) {
    const int n = num_local_;
    std::fill(H, H + static_cast<size_t>(n) * n, 0.0);
    std::fill(g, g + n, 0.0);
    *cost = 0.0;
    std::vector<const double*> params;
    std::vector<double> JtJ, Jtr, r;
    std::vector<std::vector<double>> jac_store;
    std::vector<double*> jac_ptrs;
    for (const Residual& res : residuals_) {
      const int nb = static_cast<int>(res.ids.size());
      params.clear();
      std::vector<int> off(nb + 1, 0);
      for (int k = 0; k < nb; ++k) {
        params.push_back(blocks_[res.ids[k]].values);
        off[k + 1] = off[k] + blocks_[res.ids[k]].size;
      }
      const int na = off[nb];
      JtJ.assign(static_cast<size_t>(na) * na, 0.0);
      Jtr.assign(na, 0.0);
      double c = 0.0;
      if (!TryNormalEquations(res.f, params.data(), JtJ.data(), Jtr.data(), &c, 0)) {
        const int nr = res.f->num_residuals();
        r.assign(nr, 0.0);
        jac_store.resize(nb);
        jac_ptrs.resize(nb);
        for (int k = 0; k < nb; ++k) {
          jac_store[k].assign(static_cast<size_t>(nr) * blocks_[res.ids[k]].size, 0.0);
          jac_ptrs[k] = blocks_[res.ids[k]].constant ? nullptr : jac_store[k].data();
        }
        if (!res.f->Evaluate(params.data(), r.data(), jac_ptrs.data())) return false;
        for (int i = 0; i < nr; ++i) {
          c += r[i] * r[i];
          for (int ka = 0; ka < nb; ++ka) {
            if (jac_ptrs[ka] == nullptr) continue;
            const int sa = blocks_[res.ids[ka]].size;
            const double* ra = jac_store[ka].data() + static_cast<size_t>(i) * sa;
            for (int a = 0; a < sa; ++a) {
              Jtr[off[ka] + a] += ra[a] * r[i];
              for (int kb = ka; kb < nb; ++kb) {
                if (jac_ptrs[kb] == nullptr) continue;
                const int sb = blocks_[res.ids[kb]].size;
                const double* rb = jac_store[kb].data() + static_cast<size_t>(i) * sb;
                for (int b = 0; b < sb; ++b) JtJ[static_cast<size_t>(off[ka] + a) * na + off[kb] + b] += ra[a] * rb[b];
              }
            }
          }
        }
        c *= 0.5;
        for (int a = 0; a < na; ++a)
          for (int b = 0; b < a; ++b) JtJ[static_cast<size_t>(a) * na + b] = JtJ[static_cast<size_t>(b) * na + a];
      }
      *cost += c;
      // Project onto the tangent spaces: H += P_a' JtJ[a,b] P_b, g += P_a' Jtr[a].
      for (int ka = 0; ka < nb; ++ka) {
        const Block& A = blocks_[res.ids[ka]];
        if (A.constant) continue;
        double Pa[12];
        LocalBasis(A, Pa);
        for (int la = 0; la < A.local_size; ++la) {
          double s = 0.0;
          for (int a = 0; a < A.size; ++a) s += Pa[a * A.local_size + la] * Jtr[off[ka] + a];
          g[A.local_offset + la] += s;
        }
        for (int kb = 0; kb < nb; ++kb) {
          const Block& B = blocks_[res.ids[kb]];
          if (B.constant) continue;
          double Pb[12];
          LocalBasis(B, Pb);
          for (int la = 0; la < A.local_size; ++la)
            for (int lb = 0; lb < B.local_size; ++lb) {
              double s = 0.0;
              for (int a = 0; a < A.size; ++a) {
                double t = 0.0;
                for (int b = 0; b < B.size; ++b) t += JtJ[static_cast<size_t>(off[ka] + a) * na + off[kb] + b] * Pb[b * B.local_size + lb];
                s += Pa[a * A.local_size + la] * t;
              }
              H[static_cast<size_t>(A.local_offset + la) * n + B.local_offset + lb] += s;
            }
        }
      }
    }
    return true;
  }

  // d(x (+) delta)/d delta at delta = 0, row-major size x local_size (identity for plain blocks).
  static void LocalBasis(const Block& b, double* P) {
    if (b.param == LocalParameterization::kIdentity) {
      for (int a = 0; a < b.size; ++a)
        for (int l = 0; l < b.size; ++l) P[a * b.size + l] = a == l ? 1.0 : 0.0;
    } else {
      PlusJacobian(b, P);
    }
  }

  bool EvaluateCost(double* cost) {
    *cost = 0.0;
    std::vector<const double*> params;
    std::vector<double> r, JtJ, Jtr;
    for (const Residual& res : residuals_) {
      params.clear();
      int na = 0;
      for (int id : res.ids) {
        params.push_back(blocks_[id].values);
        na += blocks_[id].size;
      }
      JtJ.assign(static_cast<size_t>(na) * na, 0.0);
      Jtr.assign(na, 0.0);
      double c = 0.0;
      if (!TryNormalEquations(res.f, params.data(), JtJ.data(), Jtr.data(), &c, 0)) {
        r.assign(res.f->num_residuals(), 0.0);
        if (!res.f->Evaluate(params.data(), r.data(), nullptr)) return false;
        for (double v : r) c += v * v;
        c *= 0.5;
      }
      *cost += c;
    }
    return true;
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
