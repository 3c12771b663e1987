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
// The solve is the trust-region loop of include/hg_trust_region.h — Ceres 1.13's
// TrustRegionMinimizer / LevenbergMarquardtStrategy / TrustRegionStepEvaluator restated
// function by function — driven here with host-side evaluations; the device-side
// solve of csrc/scan_match.cu drives the same state machine. This class adds what
// Ceres' Problem / Evaluator do around it: parameter blocks with local
// parameterizations (Plus and its Jacobian), the projection of every residual
// block's Jacobian onto the tangent spaces, and the projected gradient norm
// |Plus(x, -g) - x|. Only J'J, J'r and the cost are ever needed, so a cost function
// that can produce those directly (member NormalEquations — the GPU blocks do,
// without materialising their Jacobian) is asked for them instead of for rows.
//
// Templated on the cost-function type so that the same solver drives the GPU
// cost functions of this library and (in tests) the CPU oracle's.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <limits>
#include <vector>

#include "hg_trust_region.h"

namespace hg {

enum class LocalParameterization {
  kIdentity = 0,
  // ceres::QuaternionParameterization: x_plus_delta = exp(delta) * x, (w,x,y,z).
  kQuaternion = 1,
  // cartographer/mapping/internal/3d/rotation_parameterization.h:27-39.
  kYawOnlyQuaternion = 2,
};

// One entry of ceres::Solver::Summary::iterations (the fields the tests look at).
struct LmIteration {
  int iteration = 0;
  double cost = 0.0;               // cost at the current iterate x
  double candidate_cost = 0.0;     // cost at the candidate of this iteration
  double model_cost_change = 0.0;
  double relative_decrease = 0.0;  // step quality
  double radius = 0.0;             // after this iteration's update
  double gradient_max_norm = 0.0;
  bool successful = false;
};

struct LmOptions : TrustRegionOptions {
  LmOptions() { max_num_iterations = 10; }
  // Optional per-iteration log (Ceres' Solver::Summary::iterations), for tests.
  std::vector<LmIteration>* trace = nullptr;
};

struct LmSummary {
  double initial_cost = 0.0;
  double final_cost = 0.0;
  int num_iterations = 0;        // linear solves attempted
  int num_successful_steps = 0;
  int num_cost_evaluations = 0;  // residual-only + residual+Jacobian
  int num_jacobian_evaluations = 0;
  // 0 max iterations, 1 function tolerance, 2 gradient tolerance, 3 parameter tolerance,
  // 4 failure (evaluation failed, or 5 invalid steps in a row), 5 minimum trust-region radius
  int termination = 0;
};

template <typename CostFunctionT>
class DenseLmProblem {
 public:
  // Tangent-space size limit of Solve (the fixed-size arrays of TrustRegionCore).
  static constexpr int kMaxLocalParameters = 24;

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

  // ceres::Solve for this problem: the trust-region loop of include/hg_trust_region.h (Ceres
  // 1.13's TrustRegionMinimizer, restated) driven with host-side evaluations.
  LmSummary Solve(const LmOptions& options) {
    LmSummary summary;
    Layout();
    const int n = num_local_;
    if (n > kMaxLocalParameters) {
      summary.termination = kTerminationFailure;
      return summary;
    }
    std::vector<double> H(static_cast<size_t>(n) * n), g(n), delta(n), x, candidate_x, best, tmp;
    SaveState(&x);
    best = x;
    double cost = 0.0;
    // |Plus(x, -g) - x|_inf: the gradient norm Ceres tests (EvaluateGradientAndJacobian).
    auto gradient_max_norm = [&]() {
      for (int a = 0; a < n; ++a) delta[a] = -g[a];
      Plus(x, delta, &tmp);
      double v = 0.0;
      for (size_t k = 0; k < x.size(); ++k) v = std::max(v, std::fabs(x[k] - tmp[k]));
      return v;
    };
    auto log_iteration = [&](const TrustRegionCore<kMaxLocalParameters>& core) {
      if (options.trace == nullptr) return;
      LmIteration it;
      it.iteration = core.iteration;
      it.cost = core.x_cost;
      it.candidate_cost = core.candidate_cost;
      it.model_cost_change = core.model_cost_change;
      it.relative_decrease = core.step_quality;
      it.radius = core.radius;
      it.gradient_max_norm = core.gradient_max_norm;
      it.successful = core.step_is_successful;
      options.trace->push_back(it);
    };
    TrustRegionCore<kMaxLocalParameters> core;
    if (!EvaluateNormal(&cost, H.data(), g.data())) {
      summary.termination = kTerminationFailure;
      return summary;
    }
    core.Start(n, options, cost, H.data(), g.data(), gradient_max_norm());
    log_iteration(core);
    while (core.NextStep(delta.data()) == kEvaluate) {
      Plus(x, delta, &candidate_x);
      RestoreState(candidate_x);
      double candidate_cost = 0.0;
      if (!EvaluateCost(&candidate_cost)) candidate_cost = std::numeric_limits<double>::max();
      double step_norm2 = 0.0;
      for (size_t k = 0; k < x.size(); ++k) step_norm2 += (x[k] - candidate_x[k]) * (x[k] - candidate_x[k]);
      const int action = core.CandidateEvaluated(candidate_cost, std::sqrt(step_norm2), Norm(x));
      if (action == kAccepted) {
        x = candidate_x;
        if (!EvaluateNormal(&cost, H.data(), g.data())) {
          core.Fail();
          break;
        }
        if (core.Accepted(cost, H.data(), g.data(), gradient_max_norm())) best = x;
      }
      log_iteration(core);
      if (action == kDone) break;
    }
    // Ceres hands back the lowest-cost iterate that was accepted.
    RestoreState(best);
    summary.initial_cost = core.initial_cost;
    summary.final_cost = core.minimum_cost;
    summary.num_iterations = core.iteration;
    summary.num_successful_steps = core.num_successful_steps;
    summary.num_cost_evaluations = core.num_cost_evaluations;
    summary.num_jacobian_evaluations = core.num_jacobian_evaluations;
    summary.termination = core.termination;
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

  // Evaluator::Plus: ambient x (all blocks, SaveState order) moved by the tangent-space delta
  // of the non-constant blocks. ceres::QuaternionParameterization::Plus /
  // YawOnlyQuaternionPlus (rotation_parameterization.h:27-39).
  void Plus(const std::vector<double>& x, const std::vector<double>& delta, std::vector<double>* out) const {
    *out = x;
    size_t at = 0;
    for (const Block& b : blocks_) {
      const double* xb = &x[at];
      double* ob = &(*out)[at];
      at += b.size;
      if (b.constant) continue;
      const double* d = &delta[b.local_offset];
      if (b.param == LocalParameterization::kIdentity) {
        for (int k = 0; k < b.size; ++k) ob[k] = xb[k] + d[k];
      } else if (b.param == LocalParameterization::kQuaternion) {
        const double norm = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        if (norm > 0.0) {
          const double sin_by_norm = std::sin(norm) / norm;
          const double q_delta[4] = {std::cos(norm), sin_by_norm * d[0], sin_by_norm * d[1], sin_by_norm * d[2]};
          QuatProduct(q_delta, xb, ob);
        }
      } else {
        double c = d[0];
        c = c > 0.5 ? 0.5 : (c < -0.5 ? -0.5 : c);
        const double q_delta[4] = {std::sqrt(1. - c * c), 0., 0., c};
        QuatProduct(q_delta, xb, ob);
      }
    }
  }

  // 2-norm over the non-constant blocks (Ceres' reduced program holds only those).
  double Norm(const std::vector<double>& x) const {
    double s = 0.0;
    size_t at = 0;
    for (const Block& b : blocks_) {
      if (!b.constant)
        for (int k = 0; k < b.size; ++k) s += x[at + k] * x[at + k];
      at += b.size;
    }
    return std::sqrt(s);
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

  std::vector<Block> blocks_;
  std::vector<Residual> residuals_;
  int num_local_ = 0;
};

}  // namespace hg
