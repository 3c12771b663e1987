// CeresScanMatcher3D::Match as ONE kernel launch: the whole
// Levenberg-Marquardt solve of the 7-DoF refine runs on the device, no host round trip per
// iteration.
//
// Reference behaviour reproduced (paths under /root/reference/cartographer):
//   mapping/internal/3d/scan_matching/ceres_scan_matcher_3d.cc:72-95, 120-172   problem set-up
//   .../occupied_space_cost_function_3d.h, tsdf_space_cost_function_3d.h         scan blocks
//   .../translation_delta_cost_functor_3d.h:31-45, rotation_delta_cost_functor_3d.h:35-54
//   mapping/internal/3d/rotation_parameterization.h:27-39                        yaw-only plus
// The solver is the state machine of include/hg_trust_region.h (Ceres 1.13's trust-region loop,
// restated) — the same code the host loop of include/hg_dense_lm.h drives; hg_scan_match keeps
// that host loop for HG_SCAN_MATCH_HOST=1 and the tests compare the two.
//
// One CTA of 512 threads per problem (a thread-block cluster of eight for clouds above 4096
// points, reduced through distributed shared memory): every evaluation strides the points of all
// clouds over the threads (the Jacobian row of a one-pose block is the 7-vector u of cost.cu, so a thread
// adds 36 products per point into registers), warps reduce by shuffles, 36 threads add the
// per-warp sums in warp order, and thread 0 runs the LM bookkeeping and the 6x6 Cholesky.
// A clouds-worth of points (a few thousand) is a few microseconds per evaluation; the host
// loop it replaces paid two launches, two synchronisations and two PCIe round trips per
// evaluation.
#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>

#include "../../include/hg_trust_region.h"
#include "interpolation.cuh"

namespace cg = cooperative_groups;

namespace hg {
namespace {

constexpr int kLmThreads = 512;
constexpr int kLmWarps = kLmThreads / 32;
constexpr int kMaxClouds = 4;
constexpr int kTerms = 36;  // upper triangle of u u' (28), u r (7), r r (1)

struct ScanMatchArgs {
  GridDev grids[kMaxClouds];
  const float* points[kMaxClouds];
  int n[kMaxClouds];
  double scale[kMaxClouds];
  int num_clouds;
  int only_optimize_yaw;
  int max_num_iterations;
  int use_nonmonotonic_steps;
  double translation_weight, rotation_weight;
  double target_t[3];
  double target_q[4];
  double pose[7];
  double* out;  // pose[7], summary[6], termination
};

__device__ void QuatProduct(const double* z, const double* w, double* zw) {
  zw[0] = z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3];
  zw[1] = z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2];
  zw[2] = z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1];
  zw[3] = z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0];
}

// State of the solve; lives in shared memory, written by thread 0 only. 'eval_x' is the pose
// every thread evaluates next.
struct LmState {
  TrustRegionCore<6> core;
  double x[7], candidate_x[7], best[7], eval_x[7];
  int n;
  int phase;  // 0 = first evaluation pending, 1 = candidate evaluation pending
  int done;
};

// Shared scratch of one evaluation: ambient normal equations, tangent basis, projected system.
struct EvalShared {
  double JtJ[49], Jtr[7], P[42], H[36], g[6], A[36], y[6];
  double M[12];  // rotation-delta Jacobian / s (constant over the solve)
  double cost;
};

// ceres::QuaternionParameterization::ComputeJacobian as (sign, component of q) per entry.
__constant__ signed char kPlusSign[12] = {-1, -1, -1, 1, 1, -1, -1, 1, 1, 1, -1, 1};
__constant__ signed char kPlusIndex[12] = {1, 2, 3, 0, 3, 2, 3, 0, 1, 2, 1, 0};
// Yaw-only plus (rotation_parameterization.h:27-39): d/dc of (sqrt(1 - c^2), 0, 0, c) * q at 0.
__constant__ signed char kYawSign[4] = {-1, -1, 1, 1};
__constant__ signed char kYawIndex[4] = {3, 2, 1, 0};

// From the summed terms S at pose x: cost, H = J'J and g = J'r in tangent coordinates, computed
// by the first ~110 threads of the CTA (one matrix entry each) instead of one thread's loops.
// Contains two __syncthreads(): every thread of the CTA must call it.
__device__ void AssembleNormal(const ScanMatchArgs& a, const double* S, const double* x, const int n,
                               const int tid, EvalShared* ev) {
  // rotation_delta_cost_functor_3d.h:35-54: r = s imag(target^-1 q) = s M q, J = s M (3 x 4).
  const double* M = ev->M;
  const double st = a.translation_weight, sr = a.rotation_weight;
  if (tid < 49) {
    const int i = tid / 7, j = tid - 7 * i;
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    double v = S[lo * 7 - lo * (lo - 1) / 2 + (hi - lo)];
    if (i < 3 && i == j) v += st * st;  // translation_delta_cost_functor_3d.h:31-45: J = s I
    if (i >= 3 && j >= 3) {
      double t = 0.0;
      for (int k = 0; k < 3; ++k) t += (sr * M[k * 4 + i - 3]) * (sr * M[k * 4 + j - 3]);
      v += t;
    }
    ev->JtJ[tid] = v;
  } else if (tid < 56) {
    const int i = tid - 49;
    double v = S[28 + i];
    if (i < 3) {
      v += st * (st * (x[i] - a.target_t[i]));
    } else {
      double jr = 0.0;
      for (int k = 0; k < 3; ++k) {
        const double r = sr * (M[k * 4] * x[3] + M[k * 4 + 1] * x[4] + M[k * 4 + 2] * x[5] + M[k * 4 + 3] * x[6]);
        jr += (sr * M[k * 4 + i - 3]) * r;
      }
      v += jr;
    }
    ev->Jtr[i] = v;
  } else if (tid == 56) {
    double c = 0.5 * S[35];
    double rr = 0.0;
    for (int k = 0; k < 3; ++k) {
      const double r = st * (x[k] - a.target_t[k]);
      rr += r * r;
    }
    c += 0.5 * rr;
    rr = 0.0;
    for (int k = 0; k < 3; ++k) {
      const double r = sr * (M[k * 4] * x[3] + M[k * 4 + 1] * x[4] + M[k * 4 + 2] * x[5] + M[k * 4 + 3] * x[6]);
      rr += r * r;
    }
    ev->cost = c + 0.5 * rr;
  } else if (tid >= 64 && tid < 64 + 7 * n) {
    // Tangent basis P (7 x n): identity for t; ceres::QuaternionParameterization::ComputeJacobian
    // or the yaw-only plus (rotation_parameterization.h:27-39) for q.
    const int e = tid - 64, i = e / n, l = e - i * n;
    const double* q = x + 3;
    double v = 0.0;
    if (i < 3) {
      v = (l == i) ? 1.0 : 0.0;
    } else if (l >= 3) {
      if (n == 6) {
        const int k = (i - 3) * 3 + (l - 3);
        v = kPlusSign[k] * q[kPlusIndex[k]];
      } else {
        v = kYawSign[i - 3] * q[kYawIndex[i - 3]];
      }
    }
    ev->P[e] = v;
  }
  __syncthreads();
  if (tid < n * n) {
    const int la = tid / n, lb = tid - la * n;
    double h = 0.0;
    for (int i = 0; i < 7; ++i) {
      double t = 0.0;
      for (int j = 0; j < 7; ++j) t += ev->JtJ[i * 7 + j] * ev->P[j * n + lb];
      h += ev->P[i * n + la] * t;
    }
    ev->H[tid] = h;
  } else if (tid >= 64 && tid < 64 + n) {
    const int la = tid - 64;
    double g = 0.0;
    for (int i = 0; i < 7; ++i) g += ev->P[i * n + la] * ev->Jtr[i];
    ev->g[la] = g;
  }
  __syncthreads();
}

// Evaluator::Plus for the pose (t, q): identity for t; ceres::QuaternionParameterization::Plus
// (exp(delta) * q) or the yaw-only plus (rotation_parameterization.h:27-39) for q.
__device__ void Plus(const double* x, const double* delta, const int yaw_only, double* out) {
  for (int k = 0; k < 3; ++k) out[k] = x[k] + delta[k];
  for (int k = 3; k < 7; ++k) out[k] = x[k];
  if (!yaw_only) {
    const double dx = delta[3], dy = delta[4], dz = delta[5];
    const double norm = sqrt(dx * dx + dy * dy + dz * dz);
    if (norm > 0.0) {
      const double sin_by_norm = sin(norm) / norm;
      const double q_delta[4] = {cos(norm), sin_by_norm * dx, sin_by_norm * dy, sin_by_norm * dz};
      QuatProduct(q_delta, x + 3, out + 3);
    }
  } else {
    double c = delta[3];
    c = c > 0.5 ? 0.5 : (c < -0.5 ? -0.5 : c);
    const double q_delta[4] = {sqrt(1. - c * c), 0., 0., c};
    QuatProduct(q_delta, x + 3, out + 3);
  }
}

// |Plus(x, -g) - x|_inf (TrustRegionMinimizer::EvaluateGradientAndJacobian).
__device__ double GradientMaxNorm(const double* x, const double* g, const int n, const int yaw_only) {
  double neg[6], moved[7];
  for (int i = 0; i < n; ++i) neg[i] = -g[i];
  Plus(x, neg, yaw_only, moved);
  double v = 0.0;
  for (int k = 0; k < 7; ++k) v = fmax(v, fabs(x[k] - moved[k]));
  return v;
}

// Asks the core for the next step; publishes the candidate pose or ends the solve.
__device__ void NextCandidate(LmState* s, const ScanMatchArgs& a) {
  double delta[6];
  if (s->core.NextStep(delta) != kEvaluate) {
    s->done = 1;
    return;
  }
  Plus(s->x, delta, a.only_optimize_yaw, s->candidate_x);
  for (int k = 0; k < 7; ++k) s->eval_x[k] = s->candidate_x[k];
  s->phase = 1;
}

// The evaluation at s->eval_x is in ev (cost, H, g). DenseLmProblem::Solve of
// include/hg_dense_lm.h, the part between two evaluations. The candidate's normal equations are
// only used if the step is accepted (the host loop re-evaluates at the accepted point, which
// gives these same numbers).
__device__ void Advance(LmState* s, const ScanMatchArgs& a, const EvalShared* ev) {
  const int n = s->n;
  if (s->phase == 0) {
    TrustRegionOptions options;
    options.max_num_iterations = a.max_num_iterations;
    options.use_nonmonotonic_steps = a.use_nonmonotonic_steps != 0;
    s->core.Start(n, options, ev->cost, ev->H, ev->g, GradientMaxNorm(s->x, ev->g, n, a.only_optimize_yaw));
    NextCandidate(s, a);
    return;
  }
  const double cost = isfinite(ev->cost) ? ev->cost : 1.7976931348623157e308;
  double step_norm2 = 0.0, x_norm2 = 0.0;
  for (int k = 0; k < 7; ++k) {
    const double d = s->x[k] - s->candidate_x[k];
    step_norm2 += d * d;
    x_norm2 += s->x[k] * s->x[k];
  }
  const int action = s->core.CandidateEvaluated(cost, sqrt(step_norm2), sqrt(x_norm2));
  if (action == kDone) {
    s->done = 1;
    return;
  }
  if (action == kAccepted) {
    for (int k = 0; k < 7; ++k) s->x[k] = s->candidate_x[k];
    if (s->core.Accepted(cost, ev->H, ev->g, GradientMaxNorm(s->x, ev->g, n, a.only_optimize_yaw))) {
      for (int k = 0; k < 7; ++k) s->best[k] = s->x[k];
    }
  }
  NextCandidate(s, a);
}

// CLUSTER = 1: one CTA. CLUSTER = 8: a thread-block cluster; every CTA evaluates its share of the
// points and reduces it to 36 totals in its own shared memory, CTA 0 adds the eight vectors in rank
// order through distributed shared memory and runs the solver, the others read the next pose (and
// the stop flag) from CTA 0's shared memory. Two cluster barriers per evaluation.
template <int CLUSTER>
__device__ __forceinline__ void ScanMatchBody(const ScanMatchArgs& a) {
  __shared__ LmState st;
  __shared__ double warp_sums[kLmWarps][kTerms];
  __shared__ double totals[kTerms];
  __shared__ double grand[kTerms];
  __shared__ EvalShared ev;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = CLUSTER == 1 ? 0 : static_cast<int>(cluster.block_rank());
  const LmState* st0 = CLUSTER == 1 ? &st : cluster.map_shared_rank(&st, 0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (rank == 0 && tid == 0) {
    for (int k = 0; k < 7; ++k) st.x[k] = st.best[k] = st.eval_x[k] = a.pose[k];
    st.n = a.only_optimize_yaw ? 4 : 6;
    st.phase = 0;
    st.done = 0;
    const double z0 = a.target_q[0], z1 = -a.target_q[1], z2 = -a.target_q[2], z3 = -a.target_q[3];
    ev.M[0] = z1; ev.M[1] = z0; ev.M[2] = -z3; ev.M[3] = z2;
    ev.M[4] = z2; ev.M[5] = z3; ev.M[6] = z0; ev.M[7] = -z1;
    ev.M[8] = z3; ev.M[9] = -z2; ev.M[10] = z1; ev.M[11] = z0;
  }
  if (CLUSTER == 1) __syncthreads(); else cluster.sync();
  for (;;) {
    // ---- evaluate this CTA's share of all scan blocks at the current pose -----------------
    const double tx = st0->eval_x[0], ty = st0->eval_x[1], tz = st0->eval_x[2];
    const double qw = st0->eval_x[3], qx = st0->eval_x[4], qy = st0->eval_x[5], qz = st0->eval_x[6];
    double S[kTerms];
#pragma unroll
    for (int e = 0; e < kTerms; ++e) S[e] = 0.0;
    for (int c = 0; c < a.num_clouds; ++c) {
      const GridDev& grid = a.grids[c];
      const float* pts = a.points[c];
      const double scale = a.scale[c];
      for (int i = rank * kLmThreads + tid; i < a.n[c]; i += kLmThreads * CLUSTER) {
        const double px = static_cast<double>(pts[3 * i]), py = static_cast<double>(pts[3 * i + 1]),
                     pz = static_cast<double>(pts[3 * i + 2]);
        const WorldPoint wp = TransformPoint(qw, qx, qy, qz, tx, ty, tz, px, py, pz);
        ValueGrad v;
        if (grid.kind == kGridProb) {
          const ValueGrad p = LookupOccupancy(grid, wp.x, wp.y, wp.z);
          v.v = scale * (1. - p.v);  // occupied_space_cost_function_3d.h:77
          v.gx = -scale * p.gx;
          v.gy = -scale * p.gy;
          v.gz = -scale * p.gz;
        } else {
          ValueGrad t;
          LookupTsdf(grid, wp.x, wp.y, wp.z, &t);
          v.v = scale * t.v;  // tsdf_space_cost_function_3d.h:76
          v.gx = scale * t.gx;
          v.gy = scale * t.gy;
          v.gz = scale * t.gz;
        }
        double u[7];
        u[0] = v.gx; u[1] = v.gy; u[2] = v.gz;
        RotationTerms(qw, qx, qy, qz, px, py, pz, wp, v, &u[3], &u[4], &u[5], &u[6]);
        int e = 0;
#pragma unroll
        for (int ca = 0; ca < 7; ++ca)
#pragma unroll
          for (int cb = ca; cb < 7; ++cb) S[e++] += u[ca] * u[cb];
#pragma unroll
        for (int ca = 0; ca < 7; ++ca) S[28 + ca] += u[ca] * v.v;
        S[35] += v.v * v.v;
      }
    }
    // ---- fixed-order reduction: lanes (shuffle tree), then warps in order -------------------
#pragma unroll
    for (int e = 0; e < kTerms; ++e) {
      double s = S[e];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
      if (lane == 0) warp_sums[warp][e] = s;
    }
    __syncthreads();
    if (tid < kTerms) {
      double s = 0.0;
      for (int w = 0; w < kLmWarps; ++w) s += warp_sums[w][tid];
      totals[tid] = s;
    }
    __syncthreads();
    if (CLUSTER > 1) cluster.sync();  // every CTA's totals are in place
    if (rank == 0) {
      if (CLUSTER > 1) {
        if (tid < kTerms) {
          double s = 0.0;
          for (int r = 0; r < CLUSTER; ++r) s += cluster.map_shared_rank(totals, r)[tid];
          grand[tid] = s;
        }
        __syncthreads();
      }
      AssembleNormal(a, CLUSTER > 1 ? grand : totals, st.eval_x, st.n, tid, &ev);
      if (tid == 0) Advance(&st, a, &ev);
    }
    if (CLUSTER == 1) __syncthreads(); else cluster.sync();  // the next pose / the stop flag is published
    if (st0->done) break;
  }
  if (CLUSTER > 1) cluster.sync();  // nobody leaves while CTA 0's shared memory may still be read
  if (rank == 0 && tid == 0) {
    // Ceres hands back the lowest-cost accepted iterate.
    for (int k = 0; k < 7; ++k) a.out[k] = st.best[k];
    a.out[7] = st.core.initial_cost;
    a.out[8] = st.core.minimum_cost;
    a.out[9] = st.core.iteration;
    a.out[10] = st.core.num_successful_steps;
    a.out[11] = st.core.num_cost_evaluations;
    a.out[12] = st.core.num_jacobian_evaluations;
    a.out[13] = st.core.termination;
  }
}

template <int CLUSTER>
__global__ void __cluster_dims__(CLUSTER, 1, 1) __launch_bounds__(kLmThreads, 1)
    ScanMatchKernel(const ScanMatchArgs a) {
  ScanMatchBody<CLUSTER>(a);
}

// One CTA per problem of a batch (the refines of all the constraints found for a submap): the
// arguments come from device memory.
__global__ void __launch_bounds__(kLmThreads, 1) ScanMatchBatchKernel(const ScanMatchArgs* __restrict__ args) {
  __shared__ ScanMatchArgs a;
  {
    const int words = sizeof(ScanMatchArgs) / 4;
    const int* src = reinterpret_cast<const int*>(args + blockIdx.x);
    int* dst = reinterpret_cast<int*>(&a);
    for (int k = threadIdx.x; k < words; k += kLmThreads) dst[k] = src[k];
  }
  __syncthreads();
  ScanMatchBody<1>(a);
}

}  // namespace

// Device-side CeresScanMatcher3D::Match. Returns HG_ERR_LIMIT when the problem does not fit the
// kernel (more than kMaxClouds clouds); the caller then uses the host loop.
int ScanMatchOnDevice(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                      const hg_grid* const* grids, const double* occupied_space_weights,
                      double translation_weight, double rotation_weight, int only_optimize_yaw,
                      int max_num_iterations, int use_nonmonotonic_steps, const double* target_translation,
                      const double* initial_pose, double* pose_out, double* summary_out) {
  if (num_clouds > kMaxClouds) return HG_ERR_LIMIT;
  ScanMatchArgs a;
  std::memset(&a, 0, sizeof(a));
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "hg_scan_match: cannot create a stream");
  size_t total_floats = 0;
  for (int i = 0; i < num_clouds; ++i) total_floats += static_cast<size_t>(cloud_sizes[i]) * 3;
  float* d_points = static_cast<float*>(CachedDeviceAlloc(total_floats * sizeof(float)));
  double* d_out = static_cast<double*>(CachedDeviceAlloc(14 * sizeof(double)));
  double* h_out = static_cast<double*>(CachedPinnedAlloc(14 * sizeof(double)));
  struct Release {
    void* p[3];
    ~Release() {
      for (void* q : p) CachedFree(q);
    }
  } release{{d_points, d_out, h_out}};
  if (d_points == nullptr || d_out == nullptr || h_out == nullptr) return HG_ERR_CUDA;
  size_t offset = 0;
  for (int i = 0; i < num_clouds; ++i) {
    a.grids[i] = grids[i]->dev;
    a.points[i] = d_points + offset;
    a.n[i] = cloud_sizes[i];
    // ceres_scan_matcher_3d.cc:138-139 / :149-150
    a.scale[i] = occupied_space_weights[i] / std::sqrt(static_cast<double>(cloud_sizes[i]));
    HG_CUDA(cudaMemcpyAsync(d_points + offset, clouds[i], static_cast<size_t>(cloud_sizes[i]) * 3 * sizeof(float),
                            cudaMemcpyHostToDevice, stream));
    offset += static_cast<size_t>(cloud_sizes[i]) * 3;
  }
  a.num_clouds = num_clouds;
  a.only_optimize_yaw = only_optimize_yaw;
  a.max_num_iterations = max_num_iterations;
  a.use_nonmonotonic_steps = use_nonmonotonic_steps;
  a.translation_weight = translation_weight;
  a.rotation_weight = rotation_weight;
  for (int k = 0; k < 3; ++k) a.target_t[k] = target_translation[k];
  for (int k = 0; k < 4; ++k) a.target_q[k] = initial_pose[3 + k];
  for (int k = 0; k < 7; ++k) a.pose[k] = initial_pose[k];
  a.out = d_out;
  // A few thousand points are one CTA's worth; a full-resolution scan gets a cluster of eight.
  if (total_floats / 3 > 4096) {
    ScanMatchKernel<8><<<8, kLmThreads, 0, stream>>>(a);
  } else {
    ScanMatchKernel<1><<<1, kLmThreads, 0, stream>>>(a);
  }
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpyAsync(h_out, d_out, 14 * sizeof(double), cudaMemcpyDeviceToHost, stream));
  HG_CUDA(cudaStreamSynchronize(stream));
  // A solve that Ceres would end as FAILURE (five invalid steps in a row) still returns the best
  // pose seen: the reference does not look at the summary's termination type.
  for (int k = 0; k < 7; ++k) pose_out[k] = h_out[k];
  if (summary_out) {
    for (int k = 0; k < 6; ++k) summary_out[k] = h_out[7 + k];
  }
  return HG_OK;
}

// CeresScanMatcher3D::Match for a batch of problems in ONE launch (a CTA each). Every problem has
// at most kMaxClouds clouds and at most 4096 points in total; HG_ERR_LIMIT otherwise (the caller
// then solves that batch problem by problem).
int ScanMatchBatchOnDevice(int num_problems, const ScanMatchBatchItem* items, const double* occupied_space_weights,
                           double translation_weight, double rotation_weight, int only_optimize_yaw,
                           int max_num_iterations, int use_nonmonotonic_steps, double* poses_out,
                           double* summaries_out) {
  if (num_problems == 0) return HG_OK;
  static_assert(sizeof(ScanMatchArgs) % 4 == 0, "ScanMatchArgs is copied word by word");
  size_t total_floats = 0;
  for (int p = 0; p < num_problems; ++p) {
    if (items[p].num_clouds > kMaxClouds || items[p].num_clouds < 1) return HG_ERR_LIMIT;
    size_t floats = 0;
    for (int i = 0; i < items[p].num_clouds; ++i) floats += static_cast<size_t>(items[p].sizes[i]) * 3;
    if (floats / 3 > 4096) return HG_ERR_LIMIT;
    total_floats += floats;
  }
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "hg_scan_match: cannot create a stream");
  const size_t off_args = (total_floats * sizeof(float) + 15) & ~size_t(15);
  const size_t upload_bytes = off_args + num_problems * sizeof(ScanMatchArgs);
  char* h_up = static_cast<char*>(CachedPinnedAlloc(upload_bytes));
  char* d_up = static_cast<char*>(CachedDeviceAlloc(upload_bytes));
  double* d_out = static_cast<double*>(CachedDeviceAlloc(num_problems * 14 * sizeof(double)));
  double* h_out = static_cast<double*>(CachedPinnedAlloc(num_problems * 14 * sizeof(double)));
  struct Release {
    void* p[4];
    ~Release() {
      for (void* q : p) CachedFree(q);
    }
  } release{{h_up, d_up, d_out, h_out}};
  if (h_up == nullptr || d_up == nullptr || d_out == nullptr || h_out == nullptr) return HG_ERR_CUDA;
  ScanMatchArgs* h_args = reinterpret_cast<ScanMatchArgs*>(h_up + off_args);
  size_t offset = 0;
  for (int p = 0; p < num_problems; ++p) {
    const ScanMatchBatchItem& it = items[p];
    ScanMatchArgs& a = h_args[p];
    std::memset(&a, 0, sizeof(a));
    for (int i = 0; i < it.num_clouds; ++i) {
      a.grids[i] = it.grids[i]->dev;
      a.points[i] = reinterpret_cast<const float*>(d_up) + offset;
      a.n[i] = it.sizes[i];
      a.scale[i] = occupied_space_weights[i] / std::sqrt(static_cast<double>(it.sizes[i]));  // :138-139 / :149-150
      std::memcpy(reinterpret_cast<float*>(h_up) + offset, it.clouds[i], static_cast<size_t>(it.sizes[i]) * 3 * sizeof(float));
      offset += static_cast<size_t>(it.sizes[i]) * 3;
    }
    a.num_clouds = it.num_clouds;
    a.only_optimize_yaw = only_optimize_yaw;
    a.max_num_iterations = max_num_iterations;
    a.use_nonmonotonic_steps = use_nonmonotonic_steps;
    a.translation_weight = translation_weight;
    a.rotation_weight = rotation_weight;
    for (int k = 0; k < 3; ++k) a.target_t[k] = it.target_translation[k];
    for (int k = 0; k < 4; ++k) a.target_q[k] = it.initial_pose[3 + k];
    for (int k = 0; k < 7; ++k) a.pose[k] = it.initial_pose[k];
    a.out = d_out + 14 * static_cast<size_t>(p);
  }
  HG_CUDA(cudaMemcpyAsync(d_up, h_up, upload_bytes, cudaMemcpyHostToDevice, stream));
  ScanMatchBatchKernel<<<num_problems, kLmThreads, 0, stream>>>(reinterpret_cast<const ScanMatchArgs*>(d_up + off_args));
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpyAsync(h_out, d_out, num_problems * 14 * sizeof(double), cudaMemcpyDeviceToHost, stream));
  HG_CUDA(cudaStreamSynchronize(stream));
  for (int p = 0; p < num_problems; ++p) {
    for (int k = 0; k < 7; ++k) poses_out[7 * p + k] = h_out[14 * p + k];
    if (summaries_out) {
      for (int k = 0; k < 6; ++k) summaries_out[6 * p + k] = h_out[14 * p + 7 + k];
    }
  }
  return HG_OK;
}

}  // namespace hg
