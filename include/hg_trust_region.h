// hg_trust_region.h — the trust-region loop of Ceres 1.13, as a state machine.
//
// The reference solves every scan match with ceres::Solve (Levenberg-Marquardt, DENSE_QR):
// cartographer/mapping/internal/3d/scan_matching/ceres_scan_matcher_3d.cc:65-95. Ceres itself
// is a third-party dependency that is not vendored in the reference (pinned 1.13.0,
// scripts/install_ceres.sh:20); this header restates, step by step, the published algorithm of
//   internal/ceres/trust_region_minimizer.cc      TrustRegionMinimizer::Minimize and helpers
//   internal/ceres/levenberg_marquardt_strategy.cc  ComputeStep / StepAccepted / StepRejected
//   internal/ceres/trust_region_step_evaluator.cc   StepQuality / StepAccepted
// for a small dense problem given by its normal equations in tangent (local) coordinates. Every
// member function below names the Ceres function it follows. What is NOT Ceres: the linear step
// solves the damped normal equations (J'J + D'D) y = J'r by Cholesky, where DENSE_QR factors
// the stacked matrix [J; D] — the same minimiser up to rounding; the cost functions of this
// library deliver J'J and J'r without ever forming J.
//
// It is a state machine because the evaluations happen elsewhere: on the host between C-ABI
// calls (include/hg_dense_lm.h) or inside one kernel launch by a whole thread block
// (hectorgrapher_b200/csrc/scan_match.cu). Both drive this one implementation:
//
//   core.Start(n, options, cost, H, g, gradient_max_norm)       evaluation at the start point
//   while (core.NextStep(delta) == kEvaluate) {                 delta: tangent step, unscaled
//     candidate = Plus(x, delta); evaluate its cost
//     switch (core.CandidateEvaluated(cost, |candidate - x|, |x|)) {
//       case kAccepted: x = candidate; evaluate H, g there;
//                       if (core.Accepted(cost, H, g, gradient_max_norm)) best = x;  break;
//       case kRejected: break;
//       case kDone:     leave the loop
//     }
//   }
//   result: 'best' (the lowest-cost accepted iterate, which is what Ceres writes back).
#pragma once

#include <cmath>

#if defined(__CUDACC__)
#define HG_HD __host__ __device__
#else
#define HG_HD
#endif

namespace hg {

struct TrustRegionOptions {
  // ceres::Solver::Options defaults (include/ceres/solver.h); the reference sets only
  // max_num_iterations, use_nonmonotonic_steps and num_threads
  // (cartographer/common/ceres_solver_options.cc:36-43).
  int max_num_iterations = 50;
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
  int max_num_consecutive_invalid_steps = 5;
};

enum TrustRegionTermination {
  kTerminationMaxIterations = 0,
  kTerminationFunctionTolerance = 1,
  kTerminationGradientTolerance = 2,
  kTerminationParameterTolerance = 3,
  kTerminationFailure = 4,  // an evaluation failed, or too many invalid steps in a row
  kTerminationMinRadius = 5,
};

enum TrustRegionAction { kEvaluate = 0, kAccepted = 1, kRejected = 2, kDone = 3 };

template <int MAXN>
struct TrustRegionCore {
  TrustRegionOptions options;
  int n;
  // Scaled system at the current iterate: Hs = S H S, gs = S g, S = diag(scale).
  double Hs[MAXN * MAXN], gs[MAXN], scale[MAXN], diagonal[MAXN];
  double x_cost, minimum_cost, gradient_max_norm;
  // LevenbergMarquardtStrategy
  double radius, decrease_factor;
  bool reuse_diagonal;
  // TrustRegionStepEvaluator
  int max_consecutive_nonmonotonic_steps, num_consecutive_nonmonotonic_steps;
  double ev_minimum_cost, ev_current_cost, ev_reference_cost, ev_candidate_cost;
  double accumulated_reference_model_cost_change, accumulated_candidate_model_cost_change;
  // iteration state
  int iteration, num_consecutive_invalid_steps, termination;
  bool step_is_successful, finished;
  double model_cost_change, candidate_cost, step_quality;
  // summary
  double initial_cost;
  int num_successful_steps, num_cost_evaluations, num_jacobian_evaluations;

  // TrustRegionMinimizer::Init + IterationZero + the step evaluator's constructor.
  // (cost, H = J'J, g = J'r) at the start point; gradient_max_norm = max |Plus(x, -g) - x|.
  HG_HD void Start(const int n_in, const TrustRegionOptions& o, const double cost, const double* H,
                   const double* g, const double gradient_max_norm_in) {
    options = o;
    n = n_in;
    minimum_cost = 1.7976931348623157e308;
    num_consecutive_invalid_steps = 0;
    radius = o.initial_trust_region_radius;
    decrease_factor = 2.0;
    reuse_diagonal = false;
    iteration = 0;
    termination = kTerminationMaxIterations;
    finished = false;
    num_successful_steps = 0;
    num_cost_evaluations = 0;
    num_jacobian_evaluations = 0;
    model_cost_change = 0.0;
    candidate_cost = 0.0;
    step_quality = 0.0;
    // EvaluateGradientAndJacobian at iteration 0: Jacobi scaling 1 / (1 + |J_j|).
    for (int j = 0; j < n; ++j) scale[j] = 1.0 / (1.0 + sqrt(fmax(0.0, H[j * n + j])));
    TakeOver(cost, H, g, gradient_max_norm_in);
    initial_cost = cost;
    step_is_successful = true;
    max_consecutive_nonmonotonic_steps = o.use_nonmonotonic_steps ? o.max_consecutive_nonmonotonic_steps : 0;
    ev_minimum_cost = ev_current_cost = ev_reference_cost = ev_candidate_cost = cost;
    accumulated_reference_model_cost_change = 0.0;
    accumulated_candidate_model_cost_change = 0.0;
    num_consecutive_nonmonotonic_steps = 0;
    // FinalizeIteration of iteration 0: the start point is the first minimum.
    minimum_cost = cost;
    ++num_successful_steps;
  }

  // Loop head + ComputeTrustRegionStep (+ HandleInvalidStep): returns kEvaluate with the
  // unscaled tangent step in 'delta', or kDone (see 'termination').
  HG_HD int NextStep(double* delta) {
    double A[MAXN * MAXN], step[MAXN], y[MAXN];
    for (;;) {
      // FinalizeIterationAndCheckIfMinimizerCanContinue: MaxSolverIterationsReached,
      // GradientToleranceReached, MinTrustRegionRadiusReached, in this order.
      if (finished) return kDone;
      if (iteration >= options.max_num_iterations) return Finish(kTerminationMaxIterations);
      if (step_is_successful && gradient_max_norm <= options.gradient_tolerance)
        return Finish(kTerminationGradientTolerance);
      if (radius <= options.min_trust_region_radius) return Finish(kTerminationMinRadius);
      ++iteration;
      step_is_successful = false;
      // LevenbergMarquardtStrategy::ComputeStep: D^2 = clamp(diag(J'J)) / radius, the diagonal
      // kept from the last accepted iterate while steps are being rejected.
      if (!reuse_diagonal) {
        for (int a = 0; a < n; ++a)
          diagonal[a] = fmin(fmax(Hs[a * n + a], options.min_lm_diagonal), options.max_lm_diagonal);
      }
      reuse_diagonal = true;
      for (int i = 0; i < n * n; ++i) A[i] = Hs[i];
      for (int a = 0; a < n; ++a) A[a * n + a] += diagonal[a] / radius;
      bool step_is_valid = CholeskySolve(A, gs, y, step);
      model_cost_change = 0.0;
      if (step_is_valid) {
        for (int a = 0; a < n; ++a) step[a] = -step[a];
        // model_cost_change = -(J step)'(r + J step / 2) = -(g'step + step' J'J step / 2)
        double gd = 0.0, dJd = 0.0;
        for (int a = 0; a < n; ++a) {
          gd += gs[a] * step[a];
          double t = 0.0;
          for (int b = 0; b < n; ++b) t += Hs[a * n + b] * step[b];
          dJd += step[a] * t;
        }
        model_cost_change = -(gd + 0.5 * dJd);
        step_is_valid = model_cost_change > 0.0;
      }
      if (!step_is_valid) {
        // HandleInvalidStep: an iteration with no progress; StepIsInvalid = StepRejected(0).
        if (++num_consecutive_invalid_steps >= options.max_num_consecutive_invalid_steps)
          return Finish(kTerminationFailure);
        radius /= decrease_factor;
        decrease_factor *= 2.0;
        step_quality = 0.0;
        candidate_cost = x_cost;
        continue;
      }
      num_consecutive_invalid_steps = 0;
      for (int a = 0; a < n; ++a) delta[a] = step[a] * scale[a];  // undo the column scaling
      return kEvaluate;
    }
  }

  // ParameterToleranceReached, FunctionToleranceReached, IsStepSuccessful and — for a rejected
  // step — HandleUnsuccessfulStep. step_norm = |candidate - x| and x_norm = |x| in the ambient
  // space. A failed evaluation is reported as cost = max double.
  HG_HD int CandidateEvaluated(const double cost, const double step_norm, const double x_norm) {
    ++num_cost_evaluations;
    candidate_cost = cost;
    if (step_norm <= options.parameter_tolerance * (x_norm + options.parameter_tolerance))
      return Finish(kTerminationParameterTolerance);
    // Ceres returns here without taking the candidate over.
    if (fabs(x_cost - cost) <= options.function_tolerance * x_cost) return Finish(kTerminationFunctionTolerance);
    // TrustRegionStepEvaluator::StepQuality
    const double relative_decrease = (ev_current_cost - cost) / model_cost_change;
    const double historical_relative_decrease =
        (ev_reference_cost - cost) / (accumulated_reference_model_cost_change + model_cost_change);
    step_quality = fmax(relative_decrease, historical_relative_decrease);
    if (step_quality > options.min_relative_decrease) return kAccepted;
    // LevenbergMarquardtStrategy::StepRejected
    radius /= decrease_factor;
    decrease_factor *= 2.0;
    reuse_diagonal = true;
    return kRejected;
  }

  // HandleSuccessfulStep with the evaluation at the accepted point. Returns true when that
  // point is the lowest-cost iterate so far (the caller keeps it as the result).
  HG_HD bool Accepted(const double cost, const double* H, const double* g, const double gradient_max_norm_in) {
    TakeOver(cost, H, g, gradient_max_norm_in);
    step_is_successful = true;
    // LevenbergMarquardtStrategy::StepAccepted
    const double t = 2.0 * step_quality - 1.0;
    radius = radius / fmax(1.0 / 3.0, 1.0 - t * t * t);
    radius = fmin(options.max_trust_region_radius, radius);
    decrease_factor = 2.0;
    reuse_diagonal = false;
    // TrustRegionStepEvaluator::StepAccepted(candidate_cost, model_cost_change)
    ev_current_cost = candidate_cost;
    accumulated_candidate_model_cost_change += model_cost_change;
    accumulated_reference_model_cost_change += model_cost_change;
    if (ev_current_cost < ev_minimum_cost) {
      ev_minimum_cost = ev_current_cost;
      num_consecutive_nonmonotonic_steps = 0;
      ev_candidate_cost = ev_current_cost;
      accumulated_candidate_model_cost_change = 0.0;
    } else {
      ++num_consecutive_nonmonotonic_steps;
      if (ev_current_cost > ev_candidate_cost) {
        ev_candidate_cost = ev_current_cost;
        accumulated_candidate_model_cost_change = 0.0;
      }
    }
    if (num_consecutive_nonmonotonic_steps == max_consecutive_nonmonotonic_steps) {
      ev_reference_cost = ev_candidate_cost;
      accumulated_reference_model_cost_change = accumulated_candidate_model_cost_change;
    }
    // FinalizeIteration: parameters_ = x only when the cost is a new minimum.
    ++num_successful_steps;
    if (x_cost < minimum_cost) {
      minimum_cost = x_cost;
      return true;
    }
    return false;
  }

  // An evaluation failed at an accepted point (Ceres: FAILURE).
  HG_HD void Fail() { Finish(kTerminationFailure); }

 private:
  HG_HD int Finish(const int why) {
    termination = why;
    finished = true;
    return kDone;
  }

  HG_HD void TakeOver(const double cost, const double* H, const double* g, const double gradient_max_norm_in) {
    x_cost = cost;
    gradient_max_norm = gradient_max_norm_in;
    ++num_cost_evaluations;
    ++num_jacobian_evaluations;
    for (int a = 0; a < n; ++a) {
      gs[a] = g[a] * scale[a];
      for (int b = 0; b < n; ++b) Hs[a * n + b] = H[a * n + b] * scale[a] * scale[b];
    }
  }

  // In-place lower Cholesky of A (row-major n x n), then A x = b.
  HG_HD bool CholeskySolve(double* A, const double* b, double* y, double* x) const {
    for (int j = 0; j < n; ++j) {
      double d = A[j * n + j];
      for (int k = 0; k < j; ++k) d -= A[j * n + k] * A[j * n + k];
      if (!(d > 0.0) || !((d - d) == 0.0)) return false;  // not positive, or not finite
      const double l = sqrt(d);
      A[j * n + j] = l;
      for (int i = j + 1; i < n; ++i) {
        double s = A[i * n + j];
        for (int k = 0; k < j; ++k) s -= A[i * n + k] * A[j * n + k];
        A[i * n + j] = s / l;
      }
    }
    for (int i = 0; i < n; ++i) {
      double s = b[i];
      for (int k = 0; k < i; ++k) s -= A[i * n + k] * y[k];
      y[i] = s / A[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
      double s = y[i];
      for (int k = i + 1; k < n; ++k) s -= A[k * n + i] * x[k];
      x[i] = s / A[i * n + i];
    }
    return true;
  }
};

}  // namespace hg
