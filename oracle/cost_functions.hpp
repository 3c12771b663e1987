// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of the eight scan-to-grid cost functors and of what
// ceres::AutoDiffCostFunction does with them (evaluate the functor once on
// Jet<double, 7> / Jet<double, 14> and read residuals and Jacobian rows off
// the dual parts; evaluate on plain doubles when no Jacobian is requested).
//
// Follows (under /root/reference/cartographer/mapping/internal/3d/scan_matching):
//   occupied_space_cost_function_3d.h:36-80
//   interpolated_occupied_space_cost_function_3d.h:37-92
//   tsdf_space_cost_function_3d.h:37-79
//   interpolated_tsdf_space_cost_function_3d.h:38-96
//   interpolated_tsdf_per_point_space_cost_function_3d.h:38-74
//   multi_resolution_tsdf_space_cost_function_3d.h:37-82
//   interpolated_multi_resolution_tsdf_space_cost_function_3d.h:38-98
//   interpolated_multi_resolution_tsdf_per_point_space_cost_function_3d.h:38-74
//   translation_delta_cost_functor_3d.h:31-45, rotation_delta_cost_functor_3d.h:35-54
//   common/math.h:74-81 (QuaternionProduct)
// Parameter layout: translation[3], rotation[4] = (w, x, y, z)
// (mapping/internal/optimization/ceres_pose.h:49-50).
#pragma once

#include <memory>
#include <vector>

#include "geometry.hpp"
#include "interpolation.hpp"

namespace oracle {

// Shape of ceres::CostFunction (Ceres 1.13 include/ceres/cost_function.h).
class CostFunction {
 public:
  virtual ~CostFunction() {}
  virtual bool Evaluate(double const* const* parameters, double* residuals,
                        double** jacobians) const = 0;
  const std::vector<int>& parameter_block_sizes() const { return sizes_; }
  int num_residuals() const { return num_residuals_; }

 protected:
  std::vector<int> sizes_;
  int num_residuals_ = 0;
};

enum CostKind {
  kOccupiedSpace = 0,
  kInterpolatedOccupiedSpace = 1,
  kTSDFSpace = 2,
  kInterpolatedTSDFSpace = 3,
  kInterpolatedTSDFPerPointSpace = 4,
  kMultiResolutionTSDFSpace = 5,
  kInterpolatedMultiResolutionTSDFSpace = 6,
  kInterpolatedMultiResolutionTSDFPerPointSpace = 7,
};

inline bool IsTwoPose(const int kind) {
  return kind == kInterpolatedOccupiedSpace || kind == kInterpolatedTSDFSpace ||
         kind == kInterpolatedTSDFPerPointSpace ||
         kind == kInterpolatedMultiResolutionTSDFSpace ||
         kind == kInterpolatedMultiResolutionTSDFPerPointSpace;
}

// A point cloud view: 'stride' floats per point (3 = sensor::PointCloud,
// 4 = sensor::TimedPointCloud, sensor/rangefinder_point.h:31-40).
struct CloudView {
  const float* data = nullptr;
  int n = 0;
  int stride = 3;
};

// One scan-to-grid residual block. The grid lookup is selected by 'kind'.
class ScanCostFunction : public CostFunction {
 public:
  ScanCostFunction(const int kind, const double scaling_factor, const CloudView cloud,
                   const HybridGrid* prob, const std::vector<const HybridGridTSDF*>& tsdfs,
                   const double interpolation_ratio)
      : kind_(kind),
        scaling_factor_(scaling_factor),
        ratio_(interpolation_ratio),
        stride_(cloud.stride),
        n_(cloud.n),
        // the one-pose functors keep a reference to the caller's cloud, the
        // interpolated ones a copy; a copy is equivalent for evaluation.
        points_(cloud.data, cloud.data + static_cast<size_t>(cloud.n) * cloud.stride) {
    num_residuals_ = n_;
    if (IsTwoPose(kind)) {
      sizes_ = {3, 4, 3, 4};
    } else {
      sizes_ = {3, 4};
    }
    if (kind == kOccupiedSpace || kind == kInterpolatedOccupiedSpace) {
      occupancy_.reset(new InterpolatedGrid(*prob));
    } else if (kind == kTSDFSpace || kind == kInterpolatedTSDFSpace ||
               kind == kInterpolatedTSDFPerPointSpace) {
      tsdf_.reset(new InterpolatedTSDF(*tsdfs.front()));
    } else {
      multires_.reset(new InterpolatedMultiResolutionTSDF(tsdfs));
    }
  }

  bool Evaluate(double const* const* parameters, double* residuals,
                double** jacobians) const override {
    const int nblocks = static_cast<int>(sizes_.size());
    if (jacobians == nullptr) {
      std::vector<double> r(n_);
      if (nblocks == 2) {
        Run<double>(parameters[0], parameters[1], nullptr, nullptr, r.data());
      } else {
        Run<double>(parameters[0], parameters[1], parameters[2], parameters[3], r.data());
      }
      for (int i = 0; i < n_; ++i) residuals[i] = r[i];
      return true;
    }
    if (nblocks == 2) {
      EvaluateWithJets<7>(parameters, residuals, jacobians);
    } else {
      EvaluateWithJets<14>(parameters, residuals, jacobians);
    }
    return true;
  }

 private:
  template <int N>
  void EvaluateWithJets(double const* const* parameters, double* residuals,
                        double** jacobians) const {
    using J = Jet<N>;
    J p[N];
    int k = 0;
    for (size_t b = 0; b < sizes_.size(); ++b) {
      for (int j = 0; j < sizes_[b]; ++j, ++k) p[k] = J(parameters[b][j], k);
    }
    std::vector<J> r(n_);
    if (N == 7) {
      Run<J>(p, p + 3, nullptr, nullptr, r.data());
    } else {
      Run<J>(p, p + 3, p + 7, p + 10, r.data());
    }
    for (int i = 0; i < n_; ++i) residuals[i] = r[i].a;
    int offset = 0;
    for (size_t b = 0; b < sizes_.size(); ++b) {
      if (jacobians[b] != nullptr) {
        for (int i = 0; i < n_; ++i) {
          for (int j = 0; j < sizes_[b]; ++j) {
            jacobians[b][static_cast<size_t>(i) * sizes_[b] + j] = r[i].v[offset + j];
          }
        }
      }
      offset += sizes_[b];
    }
  }

  template <typename T>
  void Run(const T* t0, const T* q0, const T* t1, const T* q1, T* residual) const {
    const Rigid3<T> first(Vec3<T>(t0[0], t0[1], t0[2]), Quat<T>(q0[0], q0[1], q0[2], q0[3]));
    if (t1 == nullptr) {
      Residuals(first, residual);
    } else {
      const Rigid3<T> second(Vec3<T>(t1[0], t1[1], t1[2]), Quat<T>(q1[0], q1[1], q1[2], q1[3]));
      Residuals(InterpolateTransform(first, second, ratio_), residual);
    }
  }

  template <typename T>
  void Residuals(const Rigid3<T>& transform, T* residual) const {
    for (int i = 0; i < n_; ++i) {
      const float* p = &points_[static_cast<size_t>(i) * stride_];
      const Vec3<T> world = Apply(
          transform, Vec3<T>(T(static_cast<double>(p[0])), T(static_cast<double>(p[1])),
                             T(static_cast<double>(p[2]))));
      if (occupancy_) {
        const T probability = occupancy_->GetProbability(world.x, world.y, world.z);
        residual[i] = scaling_factor_ * (1. - probability);
      } else if (tsdf_) {
        const T tsd = tsdf_->GetTSD(world.x, world.y, world.z);
        residual[i] = scaling_factor_ * tsd;
      } else {
        const T tsd = multires_->GetTSD(world.x, world.y, world.z);
        residual[i] = scaling_factor_ * tsd;
      }
    }
  }

  const int kind_;
  const double scaling_factor_;
  const double ratio_;
  const int stride_;
  const int n_;
  const std::vector<float> points_;
  std::unique_ptr<InterpolatedGrid> occupancy_;
  std::unique_ptr<InterpolatedTSDF> tsdf_;
  std::unique_ptr<InterpolatedMultiResolutionTSDF> multires_;
};

// translation_delta_cost_functor_3d.h:31-45
class TranslationDeltaCostFunction : public CostFunction {
 public:
  TranslationDeltaCostFunction(const double scaling_factor, const double* target)
      : s_(scaling_factor), x_(target[0]), y_(target[1]), z_(target[2]) {
    num_residuals_ = 3;
    sizes_ = {3};
  }
  bool Evaluate(double const* const* parameters, double* residuals,
                double** jacobians) const override {
    const double* t = parameters[0];
    residuals[0] = s_ * (t[0] - x_);
    residuals[1] = s_ * (t[1] - y_);
    residuals[2] = s_ * (t[2] - z_);
    if (jacobians != nullptr && jacobians[0] != nullptr) {
      for (int i = 0; i < 9; ++i) jacobians[0][i] = 0.0;
      jacobians[0][0] = s_;
      jacobians[0][4] = s_;
      jacobians[0][8] = s_;
    }
    return true;
  }

 private:
  const double s_, x_, y_, z_;
};

// rotation_delta_cost_functor_3d.h:35-54 + common/math.h:74-81
class RotationDeltaCostFunction : public CostFunction {
 public:
  // target = (w, x, y, z)
  RotationDeltaCostFunction(const double scaling_factor, const double* target) : s_(scaling_factor) {
    inv_[0] = target[0];
    inv_[1] = -target[1];
    inv_[2] = -target[2];
    inv_[3] = -target[3];
    num_residuals_ = 3;
    sizes_ = {4};
  }
  bool Evaluate(double const* const* parameters, double* residuals,
                double** jacobians) const override {
    using J = Jet<4>;
    J w[4];
    for (int i = 0; i < 4; ++i) w[i] = J(parameters[0][i], i);
    const double* z = inv_;
    J zw[4];
    zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
    zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
    zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
    zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
    for (int r = 0; r < 3; ++r) {
      const J res = s_ * zw[r + 1];
      residuals[r] = res.a;
      if (jacobians != nullptr && jacobians[0] != nullptr) {
        for (int c = 0; c < 4; ++c) jacobians[0][r * 4 + c] = res.v[c];
      }
    }
    return true;
  }

 private:
  const double s_;
  double inv_[4];
};

}  // namespace oracle
