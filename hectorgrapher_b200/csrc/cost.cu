// Scan-to-grid residual + pose Jacobian kernels (the ceres::CostFunction::Evaluate
// hot loop of CeresScanMatcher3D / OptimizingLocalTrajectoryBuilder).
//
// Reference behaviour reproduced (paths under
// /root/reference/cartographer/mapping/internal/3d/scan_matching unless noted):
//   tsdf_space_cost_function_3d.h:68-79 and the seven sibling functors  (hot loop)
//   interpolated_grid.h:50-102            tricubic-Hermite occupancy lookup
//   interpolated_tsdf.h:29-46, 71-116     weight-aware trilinear TSDF lookup
//   interpolated_multi_resolution_tsdf.h:82-137  fine-to-coarse fall-through
//   interpolated_tsdf.h:159-192           lower-voxel selection (float/double mix)
//   transform/rigid_transform.h:191-196 + Eigen _transformVector  (point transform)
//   transform/timestamped_transform.h:39-48  (two-pose interpolation; host side, pose_prep.h)
//
// Instead of carrying Jet<double,7|14> through every operation, each point
// computes the value and its gradient with respect to the world point (three
// partials), then applies the chain rule through d world / d q (3x4) and, for
// two-pose functors, the per-block 4x4 matrices d q / d q0, d q / d q1. This is
// the same derivative (forward mode is exact), evaluated with ~4x fewer FP64
// operations. Interval / voxel selection replicates the reference's float steps
// operation by operation (no FMA contraction) so that the same 2x2x2 stencil is
// chosen.
//
// Layout: one thread per point, 256-thread CTAs, one CTA per tile of 1024 points of
// one residual block. Voxels come from the pre-decoded apron bricks (hg_internal.cuh,
// interpolation.cuh): one 4-byte brick-index load + eight float loads at fixed offsets
// per pyramid level. Outputs are written in Ceres' layout with streaming stores; the
// normal-equation variant sums the 36 terms of the per-point 7-vector in registers.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>

#include "../../include/hg_dense_lm.h"
#include "interpolation.cuh"
#include "host/pose_prep.h"

namespace hg {

struct BlockDev {
  double t[3];
  double q[4];
  double ct0, ct1;
  double m0[16];
  double m1[16];
  double s0, s1;  // m0 = s0 I + c q1', m1 = s1 I + c q0' (pose_prep.h)
  double c[4];
  double q0[4], q1[4];
  double scale;
  long long point_offset;  // first point of the block in the concatenated cloud / outputs
  int n;
  int pad;
};

struct alignas(16) TileDev {
  int block;
  int start;        // first point of the tile inside its block
  int count;        // points in the tile
  int pad;
  long long first;  // the tile's first point in the concatenated cloud / outputs (block offset + start)
  long long pad2;
};

struct CostArgs {
  GridDev grids[HG_MAX_PYRAMID_LEVELS];
  int num_grids;
  const float* points;
  int stride;
  const BlockDev* blocks;
  const TileDev* tiles;
  double* r;
  double* jt0;
  double* jq0;
  double* jt1;
  double* jq1;
  double* partials;  // normal-equation mode: [tile][P]
  int vec_store;     // jq0 / jq1 rows: 2 = 32-byte aligned (one 256-bit store), 1 = 16-byte aligned, 0 = neither
};

constexpr int kThreads = 256;

template <int LOOKUP>
__device__ __forceinline__ ValueGrad Residual(const CostArgs& a, const double scale, const double x,
                                              const double y, const double z) {
  ValueGrad r;
  if (LOOKUP == 0) {
    const ValueGrad p = LookupOccupancy(a.grids[0], x, y, z);
    r.v = scale * (1. - p.v);  // occupied_space_cost_function_3d.h:77
    r.gx = -scale * p.gx;
    r.gy = -scale * p.gy;
    r.gz = -scale * p.gz;
    return r;
  }
  ValueGrad t;
  if (LOOKUP == 1) {
    LookupTsdf(a.grids[0], x, y, z, &t);
  } else {
    bool found = false;
#pragma unroll
    for (int l = 0; l < HG_MAX_PYRAMID_LEVELS; ++l) {
      if (l < a.num_grids && !found) found = LookupTsdfLevelAllObserved(a.grids[l], x, y, z, &t);
    }
    if (!found) {  // interpolated_multi_resolution_tsdf.h:136
      t.v = static_cast<double>(a.grids[0].min_tsd);
      t.gx = t.gy = t.gz = 0.0;
    }
  }
  r.v = scale * t.v;  // tsdf_space_cost_function_3d.h:76
  r.gx = scale * t.gx;
  r.gy = scale * t.gy;
  r.gz = scale * t.gz;
  return r;
}

// Number of Gauss-Newton outputs for n ambient parameters: upper triangle of
// J'J, J'r, and sum r^2.
__host__ __device__ constexpr int NumNormalTerms(int n) { return n * (n + 1) / 2 + n + 1; }
// Terms the kernel accumulates per tile: those of the 7-vector u (see CostKernel).
constexpr int kNormalTermsU = NumNormalTerms(7);

// One 4-wide Jacobian row. vec: 2 = one 256-bit streaming store (32-byte aligned rows), 1 = two
// 128-bit ones, 0 = scalar stores.
__device__ __forceinline__ void StoreRow4(double* base, const long long gi, const int vec, const double j0,
                                          const double j1, const double j2, const double j3) {
  double* o = base + gi * 4;
  if (vec == 2) {
    asm volatile("st.global.cs.v4.b64 [%0], {%1, %2, %3, %4};" ::"l"(o), "d"(j0), "d"(j1), "d"(j2), "d"(j3) : "memory");
  } else if (vec == 1) {
    __stcs(reinterpret_cast<double2*>(o), make_double2(j0, j1));
    __stcs(reinterpret_cast<double2*>(o) + 1, make_double2(j2, j3));
  } else {
    o[0] = j0; o[1] = j1; o[2] = j2; o[3] = j3;
  }
}

// One point of a residual block: residual and Jacobian row, written straight to the outputs.
// (aw ax ay az) = grad . d world / d q of the block's effective pose; for two poses the row of
// pose k is that times d q / d q_k = s_k I + c q_other' (pose_prep.h).
template <int LOOKUP, bool TWO_POSE>
__device__ __forceinline__ void PointJacobian(const CostArgs& a, const BlockDev& sb, const double px,
                                              const double py, const double pz, const long long gi) {
  const double qw = sb.q[0], qx = sb.q[1], qy = sb.q[2], qz = sb.q[3];
  const WorldPoint wp = TransformPoint(qw, qx, qy, qz, sb.t[0], sb.t[1], sb.t[2], px, py, pz);
  const ValueGrad v = Residual<LOOKUP>(a, sb.scale, wp.x, wp.y, wp.z);
  double A[4];
  RotationTerms(qw, qx, qy, qz, px, py, pz, wp, v, &A[0], &A[1], &A[2], &A[3]);
  // Outputs are written once and never re-read by the kernel: streaming stores keep them from
  // evicting the voxel bricks out of L2.
  if (a.r) __stcs(a.r + gi, v.v);
  if (!TWO_POSE) {
    if (a.jt0) {
      double* o = a.jt0 + gi * 3;
      __stcs(o, v.gx); __stcs(o + 1, v.gy); __stcs(o + 2, v.gz);
    }
    if (a.jq0) StoreRow4(a.jq0, gi, a.vec_store, A[0], A[1], A[2], A[3]);
    return;
  }
  const double ac = A[0] * sb.c[0] + A[1] * sb.c[1] + A[2] * sb.c[2] + A[3] * sb.c[3];
  if (a.jt0) {
    double* o = a.jt0 + gi * 3;
    __stcs(o, v.gx * sb.ct0); __stcs(o + 1, v.gy * sb.ct0); __stcs(o + 2, v.gz * sb.ct0);
  }
  if (a.jq0) {
    double J[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) J[j] = sb.s0 * A[j] + ac * sb.q1[j];
    StoreRow4(a.jq0, gi, a.vec_store, J[0], J[1], J[2], J[3]);
  }
  if (a.jt1) {
    double* o = a.jt1 + gi * 3;
    __stcs(o, v.gx * sb.ct1); __stcs(o + 1, v.gy * sb.ct1); __stcs(o + 2, v.gz * sb.ct1);
  }
  if (a.jq1) {
    double J[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) J[j] = sb.s1 * A[j] + ac * sb.q0[j];
    StoreRow4(a.jq1, gi, a.vec_store, J[0], J[1], J[2], J[3]);
  }
}

__device__ __forceinline__ void LoadPoint(const CostArgs& a, const long long index, float* x, float* y, float* z) {
  const float* pp = a.points + index * a.stride;
  *x = __ldcs(pp); *y = __ldcs(pp + 1); *z = __ldcs(pp + 2);
}

// Residuals and Jacobians of a whole batch: persistent CTAs (4 per SM), CTA c takes tiles c,
// c + gridDim.x, ... A CTA that starts a tile waits ~1.5 us on the chain tile -> block descriptor
// -> first points before it has anything to compute (measured: 256 / 512 / 1024-point tiles of
// one-tile CTAs ran the C2 batch in 0.243 / 0.214 / 0.199 ms); here that chain runs one tile
// ahead: the next tile's block descriptor is copied into the other half of a double buffer with
// cp.async (no registers, no stall) during the first round of this tile, and the next tile's
// first points are the prefetch of this tile's last round.
template <int LOOKUP, bool TWO_POSE>
__global__ void __launch_bounds__(kThreads, 4) CostKernelPersistent(const CostArgs a, const int num_tiles) {
  constexpr int kWords = sizeof(BlockDev) / 8;
  // One private double buffer per warp: the warps of a CTA walk the same tiles (each its 32-point
  // slice of every 256-point round) but never wait for each other — with one shared buffer the
  // barrier at each tile boundary cost 1.7 stall cycles per issued instruction (ncu).
  __shared__ BlockDev descriptors[kThreads / 32][2];
  const int lane = static_cast<int>(threadIdx.x);
  const int warp = lane >> 5, lane32 = lane & 31;
  const int stride = gridDim.x;
  int t = blockIdx.x;
  if (t >= num_tiles) return;
  int count = a.tiles[t].count;
  long long first = a.tiles[t].first;
  {
    const double* src = reinterpret_cast<const double*>(a.blocks + a.tiles[t].block);
    double* dst = reinterpret_cast<double*>(&descriptors[warp][0]);
    for (int w = lane32; w < kWords; w += 32) dst[w] = src[w];
  }
  float nx = 0.f, ny = 0.f, nz = 0.f;
  if (lane < count) LoadPoint(a, first + lane, &nx, &ny, &nz);
  __syncwarp();
  int cur = 0;
  for (;;) {
    const int t_next = t + stride;
    const bool has_next = t_next < num_tiles;
    const BlockDev& sb = descriptors[warp][cur];
    int next_count = 0;
    long long next_first = 0;
    for (int base = 0; base < count; base += kThreads) {
      const int i = base + lane;
      const double px = static_cast<double>(nx), py = static_cast<double>(ny), pz = static_cast<double>(nz);
      if (base == 0 && has_next) {
        const double* src = reinterpret_cast<const double*>(a.blocks + __ldg(&a.tiles[t_next].block));
        double* dst = reinterpret_cast<double*>(&descriptors[warp][cur ^ 1]);
        for (int w = lane32; w < kWords; w += 32) {
          const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(dst + w));
          asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(src + w) : "memory");
        }
      }
      // The next round's point is loaded while this round's stencil gathers are in flight.
      if (base + kThreads < count) {
        if (i + kThreads < count) LoadPoint(a, first + i + kThreads, &nx, &ny, &nz);
      } else if (has_next) {
        next_count = __ldg(&a.tiles[t_next].count);
        next_first = __ldg(&a.tiles[t_next].first);
        if (lane < next_count) LoadPoint(a, next_first + lane, &nx, &ny, &nz);
      }
      if (i < count) PointJacobian<LOOKUP, TWO_POSE>(a, sb, px, py, pz, first + i);
    }
    if (!has_next) break;
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    t = t_next;
    count = next_count;
    first = next_first;
    cur ^= 1;
  }
}

// Gauss-Newton sums of a batch (NORMAL_EQ), one CTA per tile.
template <int LOOKUP, bool TWO_POSE, bool NORMAL_EQ>
__global__ void __launch_bounds__(kThreads, 2) CostKernel(const CostArgs a) {
  static_assert(NORMAL_EQ, "the Jacobian outputs are CostKernelPersistent's");
  __shared__ BlockDev sb;
  extern __shared__ double red[];  // kNormalTermsU x kThreads

  const TileDev tile = a.tiles[blockIdx.x];
  {
    // Cooperative copy of the block descriptor.
    const int words = sizeof(BlockDev) / 8;
    const double* src = reinterpret_cast<const double*>(a.blocks + tile.block);
    double* dst = reinterpret_cast<double*>(&sb);
    if (threadIdx.x < words) dst[threadIdx.x] = src[threadIdx.x];
  }
  __syncthreads();

  // Every row of [J | r] is u M with u = (grad value, grad . d world / d q) — 7 numbers per point
  // — and a block-constant 7 x NC matrix M (identity for one pose; the interpolation weights and
  // d q / d q0, d q / d q1 for two). So J'J = M' (sum u u') M: the kernel only sums the 36 terms
  // of u and r (upper triangle of u u', u r, r r) per thread in registers, and
  // ReducePartialsKernel applies M once per block.
  double S[kNormalTermsU];
#pragma unroll
  for (int e = 0; e < kNormalTermsU; ++e) S[e] = 0.0;

  const int tile_end = tile.start + tile.count;
  for (int base = tile.start; base < tile_end; base += kThreads) {
    const int i = base + threadIdx.x;
    if (i < tile_end) {
      float fx, fy, fz;
      LoadPoint(a, sb.point_offset + i, &fx, &fy, &fz);
      const double px = static_cast<double>(fx), py = static_cast<double>(fy), pz = static_cast<double>(fz);
      const double qw = sb.q[0], qx = sb.q[1], qy = sb.q[2], qz = sb.q[3];
      const WorldPoint wp = TransformPoint(qw, qx, qy, qz, sb.t[0], sb.t[1], sb.t[2], px, py, pz);
      const ValueGrad v = Residual<LOOKUP>(a, sb.scale, wp.x, wp.y, wp.z);
      double aw, ax, ay, az;
      RotationTerms(qw, qx, qy, qz, px, py, pz, wp, v, &aw, &ax, &ay, &az);
      const double u[7] = {v.gx, v.gy, v.gz, aw, ax, ay, az};
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
  if (NORMAL_EQ) {
    // Tile total of every term in a fixed order: thread-private sums -> shared memory
    // [term][thread] (conflict-free both ways) -> one warp per term: 8 values per lane, then
    // a shuffle tree.
#pragma unroll
    for (int e = 0; e < kNormalTermsU; ++e) red[e * kThreads + threadIdx.x] = S[e];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = warp; e < kNormalTermsU; e += kThreads / 32) {
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < kThreads / 32; ++k) s += red[e * kThreads + k * 32 + lane];
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
      if (lane == 0) a.partials[static_cast<size_t>(blockIdx.x) * kNormalTermsU + e] = s;
    }
  }
}

// Sums the tile partials of each residual block in tile order (deterministic), then maps the
// 36 terms of u to the block's parameters: one pose — u is the Jacobian row, the terms are the
// output; two poses — row = u M with M = [ct0 I3, 0 | 0, m0 | ct1 I3, 0 | 0, m1] (7 x 14), so
// J'J = M' S M, J'r = M' (sum u r). Output per block: upper triangle of J'J, J'r, r'r.
__global__ void __launch_bounds__(128) ReducePartialsKernel(const double* __restrict__ partials,
                                                            const int* __restrict__ tile_begin,
                                                            const BlockDev* __restrict__ blocks,
                                                            const int two_pose, double* __restrict__ out) {
  __shared__ double S[kNormalTermsU];
  const int b = blockIdx.x;
  const int e = threadIdx.x;
  if (e < kNormalTermsU) {
    double s = 0.0;
    for (int t = tile_begin[b]; t < tile_begin[b + 1]; ++t) s += partials[static_cast<size_t>(t) * kNormalTermsU + e];
    S[e] = s;
  }
  __syncthreads();
  if (!two_pose) {
    if (e < kNormalTermsU) out[static_cast<size_t>(b) * kNormalTermsU + e] = S[e];
    return;
  }
  constexpr int NC = 14, P = NumNormalTerms(NC);
  if (e >= P) return;
  const BlockDev& blk = blocks[b];
  // Column c of M as a 7-vector.
  const auto column = [&blk](const int c, double* m) {
#pragma unroll
    for (int i = 0; i < 7; ++i) m[i] = 0.0;
    if (c < 3) {
      m[c] = blk.ct0;
    } else if (c < 7) {
      for (int k = 0; k < 4; ++k) m[3 + k] = blk.m0[4 * k + (c - 3)];
    } else if (c < 10) {
      m[c - 7] = blk.ct1;
    } else {
      for (int k = 0; k < 4; ++k) m[3 + k] = blk.m1[4 * k + (c - 10)];
    }
  };
  const auto sym = [](const int i, const int j) {  // index of (i, j) in the upper triangle of the 7 x 7 sums
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    return lo * 7 - lo * (lo - 1) / 2 + (hi - lo);
  };
  double value;
  if (e < NC * (NC + 1) / 2) {
    int ca = 0, r = e;
    while (r >= NC - ca) { r -= NC - ca; ++ca; }
    const int cb = ca + r;
    double ma[7], mb[7];
    column(ca, ma);
    column(cb, mb);
    value = 0.0;
    for (int i = 0; i < 7; ++i) {
      double row = 0.0;
      for (int j = 0; j < 7; ++j) row += S[sym(i, j)] * mb[j];
      value += ma[i] * row;
    }
  } else if (e < NC * (NC + 1) / 2 + NC) {
    double m[7];
    column(e - NC * (NC + 1) / 2, m);
    value = 0.0;
    for (int i = 0; i < 7; ++i) value += m[i] * S[28 + i];
  } else {
    value = S[35];
  }
  out[static_cast<size_t>(b) * P + e] = value;
}

// Copies 8-byte words from pinned host memory (read over PCIe in place) to device memory.
__global__ void UploadWordsKernel(const double* __restrict__ src, double* __restrict__ dst, const int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = src[i];
}

// Which pyramid level answers each point of a multi-resolution TSDF batch
// (interpolated_multi_resolution_tsdf.h:82-137: the first level whose eight corners are all
// observed): counts[l] for level l, counts[num_grids] for "none". Settles how many stencil
// gathers a point really costs (SURVEY 8d counts 32 B per level touched).
__global__ void __launch_bounds__(kThreads) LevelUsageKernel(const CostArgs a, unsigned long long* __restrict__ counts) {
  __shared__ BlockDev sb;
  const TileDev tile = a.tiles[blockIdx.x];
  {
    const int words = sizeof(BlockDev) / 8;
    const double* src = reinterpret_cast<const double*>(a.blocks + tile.block);
    double* dst = reinterpret_cast<double*>(&sb);
    if (threadIdx.x < words) dst[threadIdx.x] = src[threadIdx.x];
  }
  __syncthreads();
  const int tile_end = tile.start + tile.count;
  for (int i = tile.start + threadIdx.x; i < tile_end; i += kThreads) {
    const float* pp = a.points + (sb.point_offset + i) * a.stride;
    const WorldPoint wp = TransformPoint(sb.q[0], sb.q[1], sb.q[2], sb.q[3], sb.t[0], sb.t[1], sb.t[2],
                                         static_cast<double>(pp[0]), static_cast<double>(pp[1]),
                                         static_cast<double>(pp[2]));
    int level = a.num_grids;
    ValueGrad t;
    for (int l = a.num_grids - 1; l >= 0; --l)
      if (LookupTsdfLevelAllObserved(a.grids[l], wp.x, wp.y, wp.z, &t)) level = l;
    atomicAdd(&counts[level], 1ull);
  }
}

}  // namespace hg

using namespace hg;

// ---------------------------------------------------------------------------
// Host side.

struct hg_batch {
  int device = 0;
  int kind = 0;
  int lookup = 0;  // 0 occupancy, 1 tsdf, 2 multires
  bool two_pose = false;
  int num_blocks = 0;
  int stride = 3;
  int64_t num_points = 0;
  std::vector<int64_t> offsets;  // num_blocks + 1
  std::vector<double> scales;
  std::vector<double> ratios;
  // The grids are read through their handles at every call, so that a cost function created
  // before a device-side insertion (insert.cu) sees the grid's current buffers afterwards.
  const hg_grid* grids[HG_MAX_PYRAMID_LEVELS];
  int num_grids = 0;
  float* d_points = nullptr;
  // Tile tables for 256-point (output) and 1024-point (normal equations) tiles.
  TileDev* d_tiles_out = nullptr;
  int num_tiles_out = 0;
  int tile_out = 256;
  bool wants_stencils = false;  // large batch: keep the grids' packed stencil copies built
  TileDev* d_tiles_ne = nullptr;
  int num_tiles_ne = 0;
  int* d_ne_tile_begin = nullptr;  // num_blocks + 1
  // Per-call scratch, guarded by 'mu' (calls on one handle serialise on it).
  mutable std::mutex mu;
  // Stream of the host-pointer entry points when the caller passes none: the legacy default
  // stream, or the owning host thread's stream for the cost functions of hg_scan_match.
  cudaStream_t stream = nullptr;
  // Block descriptors travel through a ring of (pinned, device) slots: an evaluation with device
  // outputs returns while its descriptor copy and its kernel are still queued, so the next call
  // must not write the same staging memory. A slot is reused only after the event recorded behind
  // the work that reads it has completed.
  static constexpr int kBlockSlots = 4;
  mutable BlockDev* d_blocks[kBlockSlots] = {nullptr, nullptr, nullptr, nullptr};
  mutable BlockDev* h_blocks[kBlockSlots] = {nullptr, nullptr, nullptr, nullptr};  // pinned
  mutable cudaEvent_t slot_done[kBlockSlots] = {nullptr, nullptr, nullptr, nullptr};
  mutable int next_slot = 0;
  // End-to-end pipeline of the host-pointer surface (large batches): the batch is cut into chunks
  // of whole blocks; hg_batch_update_points queues the chunks' points on 'copy_in', evaluate runs a
  // chunk's tiles as soon as its points are there and hands its outputs to 'copy_out', so that
  // host-to-device copies, kernels and device-to-host copies of different chunks overlap.
  static constexpr int kMaxChunks = 8;
  std::vector<int> chunk_block_begin;  // num_chunks + 1
  std::vector<int> out_tile_begin;     // num_blocks + 1: first output tile of every block
  mutable cudaStream_t copy_in = nullptr, copy_out = nullptr;
  mutable cudaEvent_t chunk_in[kMaxChunks] = {nullptr};
  mutable cudaEvent_t chunk_done[kMaxChunks] = {nullptr};
  mutable bool points_pending = false;  // chunk_in events guard d_points
  mutable double* d_partials = nullptr;
  mutable double* d_ne_out = nullptr;
  mutable double* h_ne_out = nullptr;  // pinned
  mutable double* d_out = nullptr;     // staging for host-pointer evaluate
  mutable size_t d_out_doubles = 0;
  mutable double* h_out = nullptr;     // pinned staging for small host-pointer evaluates
  mutable size_t h_out_doubles = 0;
};

struct hg_cost {
  hg_batch* batch = nullptr;
};

namespace {

constexpr int kTileOut = 1024;
constexpr int kTileNe = 1024;

int LookupOfKind(int kind) {
  if (kind == HG_COST_OCCUPIED_SPACE || kind == HG_COST_INTERPOLATED_OCCUPIED_SPACE) return 0;
  if (kind == HG_COST_TSDF_SPACE || kind == HG_COST_INTERPOLATED_TSDF_SPACE ||
      kind == HG_COST_INTERPOLATED_TSDF_PER_POINT_SPACE)
    return 1;
  return 2;
}
bool TwoPoseKind(int kind) {
  return kind == HG_COST_INTERPOLATED_OCCUPIED_SPACE || kind == HG_COST_INTERPOLATED_TSDF_SPACE ||
         kind == HG_COST_INTERPOLATED_TSDF_PER_POINT_SPACE ||
         kind == HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_SPACE ||
         kind == HG_COST_INTERPOLATED_MULTI_RESOLUTION_TSDF_PER_POINT_SPACE;
}

void BuildTiles(const std::vector<int64_t>& offsets, int tile_points, std::vector<TileDev>* tiles,
                std::vector<int>* tile_begin) {
  tiles->clear();
  if (tile_begin) tile_begin->clear();
  for (size_t b = 0; b + 1 < offsets.size(); ++b) {
    if (tile_begin) tile_begin->push_back(static_cast<int>(tiles->size()));
    const int64_t n = offsets[b + 1] - offsets[b];
    for (int64_t s = 0; s < n; s += tile_points) {
      TileDev t;
      t.block = static_cast<int>(b);
      t.start = static_cast<int>(s);
      t.count = static_cast<int>(std::min<int64_t>(tile_points, n - s));
      t.pad = 0;
      t.first = offsets[b] + s;
      t.pad2 = 0;
      tiles->push_back(t);
    }
  }
  if (tile_begin) tile_begin->push_back(static_cast<int>(tiles->size()));
}

// Persistent grid of the Jacobian kernel: 4 CTAs per SM (its launch bound), never more than tiles.
int PersistentGrid(const int device, const int num_tiles) {
  static std::atomic<int> sms[64];
  const int slot = device & 63;
  int n = sms[slot].load(std::memory_order_acquire);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || n <= 0) n = 148;
    sms[slot].store(n, std::memory_order_release);
  }
  int per_sm = 4;
  if (const char* env = std::getenv("HG_COST_CTAS_PER_SM")) per_sm = std::max(1, std::atoi(env));
  return std::min(num_tiles, n * per_sm);
}

template <bool NE>
int LaunchCost(const hg_batch* B, const CostArgs& args, int num_tiles, cudaStream_t stream) {
  if (num_tiles == 0) return HG_OK;
  const size_t smem = NE ? static_cast<size_t>(kThreads) * kNormalTermsU * sizeof(double) : 0;
  const int grid = PersistentGrid(B->device, num_tiles);
#define HG_LAUNCH(L, T)                                                                         \
  do {                                                                                          \
    if (!NE) {                                                                                  \
      CostKernelPersistent<L, T><<<grid, kThreads, 0, stream>>>(args, num_tiles);              \
      break;                                                                                    \
    }                                                                                           \
    if (smem > 48 * 1024) {                                                                     \
      /* once per device and kernel: the attribute sticks */                                    \
      static std::atomic<unsigned long long> configured{0};                                     \
      const unsigned long long bit = 1ull << (B->device & 63);                                  \
      if ((configured.load(std::memory_order_acquire) & bit) == 0) {                            \
        HG_CUDA(cudaFuncSetAttribute(CostKernel<L, T, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                     static_cast<int>(smem)));                                  \
        configured.fetch_or(bit, std::memory_order_release);                                    \
      }                                                                                         \
    }                                                                                           \
    CostKernel<L, T, true><<<num_tiles, kThreads, smem, stream>>>(args);                        \
  } while (0)
  if (B->lookup == 0) {
    if (B->two_pose) HG_LAUNCH(0, true); else HG_LAUNCH(0, false);
  } else if (B->lookup == 1) {
    if (B->two_pose) HG_LAUNCH(1, true); else HG_LAUNCH(1, false);
  } else {
    if (B->two_pose) HG_LAUNCH(2, true); else HG_LAUNCH(2, false);
  }
#undef HG_LAUNCH
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  return HG_OK;
}

int LazyEvent(cudaEvent_t* e) {
  if (*e == nullptr) HG_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  return HG_OK;
}

// Waits until nothing queued by earlier calls still reads the batch's device buffers.
int WaitForSlots(const hg_batch* B) {
  for (int k = 0; k < hg_batch::kBlockSlots; ++k)
    if (B->slot_done[k] != nullptr) HG_CUDA(cudaEventSynchronize(B->slot_done[k]));
  return HG_OK;
}

// Fills a free slot's pinned block descriptors from host poses and queues their copy to the
// device on 'stream'. The caller records B->slot_done[*slot] behind the last launch that reads them.
int PrepareBlocks(const hg_batch* B, const double* poses, int num_poses, const int32_t* i0,
                  const int32_t* i1, cudaStream_t stream, int* slot_out) {
  const int slot = B->next_slot;
  B->next_slot = (slot + 1) % hg_batch::kBlockSlots;
  *slot_out = slot;
  if (B->d_blocks[slot] == nullptr) {
    BlockDev* dev = nullptr;
    HG_ALLOC(dev, CachedDeviceAlloc(sizeof(BlockDev) * B->num_blocks));
    B->d_blocks[slot] = dev;
  }
  if (B->h_blocks[slot] == nullptr) {
    BlockDev* pinned = nullptr;
    HG_ALLOC(pinned, CachedPinnedAlloc(sizeof(BlockDev) * B->num_blocks));
    B->h_blocks[slot] = pinned;
  }
  int rc = LazyEvent(&B->slot_done[slot]);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaEventSynchronize(B->slot_done[slot]));  // never recorded = complete
  for (int b = 0; b < B->num_blocks; ++b) {
    const int a0 = i0 ? i0[b] : 0;
    HG_REQUIRE(a0 >= 0 && a0 < num_poses, "pose_index0 out of range");
    BlockPose p;
    if (B->two_pose) {
      HG_REQUIRE(i1 != nullptr, "pose_index1 is required for interpolated cost functions");
      const int a1 = i1[b];
      HG_REQUIRE(a1 >= 0 && a1 < num_poses, "pose_index1 out of range");
      p = TwoPose(poses + 7 * a0, poses + 7 * a0 + 3, poses + 7 * a1, poses + 7 * a1 + 3, B->ratios[b]);
    } else {
      p = OnePose(poses + 7 * a0, poses + 7 * a0 + 3);
    }
    BlockDev& d = B->h_blocks[slot][b];
    std::memcpy(d.t, p.t, sizeof(d.t));
    std::memcpy(d.q, p.q, sizeof(d.q));
    d.ct0 = p.ct0;
    d.ct1 = p.ct1;
    std::memcpy(d.m0, p.m0, sizeof(d.m0));
    std::memcpy(d.m1, p.m1, sizeof(d.m1));
    d.s0 = p.s0;
    d.s1 = p.s1;
    std::memcpy(d.c, p.c, sizeof(d.c));
    std::memcpy(d.q0, p.q0, sizeof(d.q0));
    std::memcpy(d.q1, p.q1, sizeof(d.q1));
    d.scale = B->scales[b];
    d.point_offset = B->offsets[b];
    d.n = static_cast<int>(B->offsets[b + 1] - B->offsets[b]);
    d.pad = 0;
  }
  // The descriptors travel by a kernel that reads the pinned array in place (unified addressing),
  // not by a copy: a cudaMemcpyAsync would queue on the host-to-device copy engine behind the
  // point chunks hg_batch_update_points has just queued (77 MB on C2), and the first chunk's
  // kernel would wait ~1.4 ms for 30 KB.
  const int words = static_cast<int>(sizeof(BlockDev) / 8) * B->num_blocks;
  UploadWordsKernel<<<std::min((words + 255) / 256, 64), 256, 0, stream>>>(
      reinterpret_cast<const double*>(B->h_blocks[slot]), reinterpret_cast<double*>(B->d_blocks[slot]), words);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  return HG_OK;
}

// Points queued by hg_batch_update_points: work on 'stream' that reads all of them waits for the
// last chunk (the chunks arrive in order on one stream).
int WaitForAllPoints(const hg_batch* B, cudaStream_t stream) {
  if (B->points_pending) {
    const int last = static_cast<int>(B->chunk_block_begin.size()) - 2;
    HG_CUDA(cudaStreamWaitEvent(stream, B->chunk_in[last], 0));
  }
  return HG_OK;
}

// build_stencils: make sure every grid has its packed stencil copy (built when the batch was
// created; built again here after a grid was written in between). The no-allocation entry point
// passes false and gathers from the apron bricks in that case.
CostArgs BaseArgs(const hg_batch* B, const bool build_stencils = true) {
  CostArgs a;
  std::memset(&a, 0, sizeof(a));
  const char* env = std::getenv("HG_NO_STENCILS");  // diagnostic: gather from the apron bricks instead
  const bool packed = env == nullptr || std::atoi(env) == 0;
  for (int i = 0; i < B->num_grids; ++i) {
    if (packed && build_stencils && B->wants_stencils) EnsureStencils(B->grids[i]);  // no-op once built
    a.grids[i] = B->grids[i]->dev;
    if (!packed) a.grids[i].stencils = nullptr;
  }
  a.num_grids = B->num_grids;
  a.points = B->d_points;
  a.stride = B->stride;
  return a;  // a.blocks: the slot PrepareBlocks filled

}

}  // namespace

extern "C" {

int hg_batch_create(int kind, int num_blocks, const float* points, int stride_floats,
                    const int64_t* cloud_offsets, const double* scaling_factors,
                    const double* interpolation_ratios, const hg_grid* const* grids, int num_grids,
                    hg_batch** out) {
  HG_REQUIRE(out != nullptr && kind >= 0 && kind <= 7, "hg_batch_create: bad kind");
  HG_REQUIRE(num_blocks >= 1 && cloud_offsets != nullptr && scaling_factors != nullptr,
             "hg_batch_create: bad block arrays");
  HG_REQUIRE(stride_floats >= 3, "hg_batch_create: stride must be >= 3 floats");
  HG_REQUIRE(num_grids >= 1 && num_grids <= HG_MAX_PYRAMID_LEVELS && grids != nullptr,
             "hg_batch_create: bad grid list");
  const int lookup = LookupOfKind(kind);
  HG_REQUIRE(lookup == 2 || num_grids == 1, "hg_batch_create: this functor takes exactly one grid");
  float last_resolution = 0.f;
  for (int i = 0; i < num_grids; ++i) {
    HG_REQUIRE(grids[i] != nullptr, "hg_batch_create: null grid");
    HG_REQUIRE((grids[i]->dev.kind == kGridProb) == (lookup == 0),
               "hg_batch_create: grid type does not match the functor");
    // interpolated_multi_resolution_tsdf.h:62-67 CHECKs ascending resolution.
    HG_REQUIRE(last_resolution < grids[i]->dev.resolution,
               "hg_batch_create: TSDF pyramid must be sorted by ascending resolution");
    last_resolution = grids[i]->dev.resolution;
  }
  const bool two = TwoPoseKind(kind);
  std::unique_ptr<hg_batch> B(new hg_batch);
  HG_CUDA(cudaGetDevice(&B->device));
  B->kind = kind;
  B->lookup = lookup;
  B->two_pose = two;
  B->num_blocks = num_blocks;
  B->stride = stride_floats;
  B->offsets.assign(cloud_offsets, cloud_offsets + num_blocks + 1);
  HG_REQUIRE(B->offsets[0] == 0, "hg_batch_create: cloud_offsets[0] must be 0");
  for (int b = 0; b < num_blocks; ++b) {
    HG_REQUIRE(B->offsets[b + 1] >= B->offsets[b], "hg_batch_create: cloud_offsets must be non-decreasing");
    HG_REQUIRE(B->offsets[b + 1] - B->offsets[b] < (int64_t(1) << 31), "hg_batch_create: block too large");
  }
  B->num_points = B->offsets[num_blocks];
  HG_REQUIRE(B->num_points == 0 || points != nullptr, "hg_batch_create: null points");
  B->scales.assign(scaling_factors, scaling_factors + num_blocks);
  if (two) {
    HG_REQUIRE(interpolation_ratios != nullptr, "hg_batch_create: interpolation ratios required");
    B->ratios.assign(interpolation_ratios, interpolation_ratios + num_blocks);
  } else {
    B->ratios.assign(num_blocks, 0.0);
  }
  for (int i = 0; i < num_grids; ++i) B->grids[i] = grids[i];
  B->num_grids = num_grids;
  {
    // Packed stencil copies of the grids (GridDev::stencils), built once per grid — for batches large
    // enough to gain from them: a scan matcher's few thousand points against the submap that is being
    // written every scan would pay a rebuild per scan for a kernel that is latency-bound anyway.
    int64_t min_points = 32768;
    if (const char* env = std::getenv("HG_STENCIL_MIN_POINTS")) min_points = std::atoll(env);
    const char* off = std::getenv("HG_NO_STENCILS");
    B->wants_stencils = B->num_points >= min_points;
    if (B->wants_stencils && (off == nullptr || std::atoi(off) == 0)) {
      for (int i = 0; i < num_grids; ++i) {
        const int rc = EnsureStencils(grids[i]);
        if (rc != HG_OK) return rc;
      }
    }
  }

  const size_t point_bytes = static_cast<size_t>(std::max<int64_t>(B->num_points, 1)) * stride_floats * sizeof(float);
  HG_ALLOC(B->d_points, CachedDeviceAlloc(point_bytes));
  if (B->num_points > 0) {
    HG_CUDA(cudaMemcpy(B->d_points, points, static_cast<size_t>(B->num_points) * stride_floats * sizeof(float),
                       cudaMemcpyHostToDevice));
  }
  std::vector<TileDev> tiles;
  std::vector<int> tile_begin;
  B->tile_out = kTileOut;
  if (const char* env = std::getenv("HG_TILE_OUT")) B->tile_out = std::max(256, (std::atoi(env) / 256) * 256);
  BuildTiles(B->offsets, B->tile_out, &tiles, &B->out_tile_begin);
  B->num_tiles_out = static_cast<int>(tiles.size());
  {
    // Chunks of whole blocks for the end-to-end pipeline: at most kMaxChunks, each of at least
    // ~256k points.
    int max_chunks = hg_batch::kMaxChunks;
    if (const char* env = std::getenv("HG_E2E_CHUNKS")) max_chunks = std::min(max_chunks, std::max(1, std::atoi(env)));
    const int64_t target = std::max<int64_t>(int64_t(1) << 18, (B->num_points + max_chunks - 1) / max_chunks);
    B->chunk_block_begin.assign(1, 0);
    int64_t in_chunk = 0;
    for (int b = 0; b < num_blocks; ++b) {
      in_chunk += B->offsets[b + 1] - B->offsets[b];
      const int chunks_so_far = static_cast<int>(B->chunk_block_begin.size());
      if (b + 1 == num_blocks || (in_chunk >= target && chunks_so_far < max_chunks)) {
        B->chunk_block_begin.push_back(b + 1);
        in_chunk = 0;
      }
    }
  }
  HG_ALLOC(B->d_tiles_out, CachedDeviceAlloc(sizeof(TileDev) * std::max<size_t>(tiles.size(), 1)));
  if (!tiles.empty())
    HG_CUDA(cudaMemcpy(B->d_tiles_out, tiles.data(), sizeof(TileDev) * tiles.size(), cudaMemcpyHostToDevice));
  BuildTiles(B->offsets, kTileNe, &tiles, &tile_begin);
  B->num_tiles_ne = static_cast<int>(tiles.size());
  HG_ALLOC(B->d_tiles_ne, CachedDeviceAlloc(sizeof(TileDev) * std::max<size_t>(tiles.size(), 1)));
  if (!tiles.empty())
    HG_CUDA(cudaMemcpy(B->d_tiles_ne, tiles.data(), sizeof(TileDev) * tiles.size(), cudaMemcpyHostToDevice));
  HG_ALLOC(B->d_ne_tile_begin, CachedDeviceAlloc(sizeof(int) * tile_begin.size()));
  HG_CUDA(cudaMemcpy(B->d_ne_tile_begin, tile_begin.data(), sizeof(int) * tile_begin.size(), cudaMemcpyHostToDevice));

  const int P = NumNormalTerms(two ? 14 : 7);
  HG_ALLOC(B->d_partials, CachedDeviceAlloc(sizeof(double) * kNormalTermsU * std::max(B->num_tiles_ne, 1)));
  HG_ALLOC(B->d_ne_out, CachedDeviceAlloc(sizeof(double) * P * num_blocks));
  HG_ALLOC(B->h_ne_out, CachedPinnedAlloc(sizeof(double) * P * num_blocks));
  *out = B.release();
  return HG_OK;
}

void hg_batch_destroy(hg_batch* B) {
  if (B == nullptr) return;
  CachedFree(B->d_points);
  CachedFree(B->d_tiles_out);
  CachedFree(B->d_tiles_ne);
  CachedFree(B->d_ne_tile_begin);
  // Work queued by evaluations that returned early may still read these buffers.
  WaitForSlots(B);
  if (B->copy_in) cudaStreamSynchronize(B->copy_in);
  if (B->copy_out) cudaStreamSynchronize(B->copy_out);
  for (int k = 0; k < hg_batch::kBlockSlots; ++k) {
    CachedFree(B->d_blocks[k]);
    CachedFree(B->h_blocks[k]);
    if (B->slot_done[k]) cudaEventDestroy(B->slot_done[k]);
  }
  for (int k = 0; k < hg_batch::kMaxChunks; ++k) {
    if (B->chunk_in[k]) cudaEventDestroy(B->chunk_in[k]);
    if (B->chunk_done[k]) cudaEventDestroy(B->chunk_done[k]);
  }
  if (B->copy_in) cudaStreamDestroy(B->copy_in);
  if (B->copy_out) cudaStreamDestroy(B->copy_out);
  CachedFree(B->d_partials);
  CachedFree(B->d_ne_out);
  CachedFree(B->h_ne_out);
  CachedFree(B->d_out);
  CachedFree(B->h_out);
  delete B;
}

int64_t hg_batch_num_residuals(const hg_batch* B) { return B ? B->num_points : 0; }

int hg_batch_update_points(hg_batch* B, const float* points, int stride_floats) {
  HG_REQUIRE(B != nullptr && points != nullptr && stride_floats == B->stride,
             "hg_batch_update_points: bad arguments");
  HG_ON_DEVICE(B->device);
  std::lock_guard<std::mutex> lock(B->mu);
  // Evaluations that returned early (device outputs) may still be reading the old points.
  int rc = WaitForSlots(B);
  if (rc != HG_OK) return rc;
  if (B->copy_in == nullptr) HG_CUDA(cudaStreamCreateWithFlags(&B->copy_in, cudaStreamNonBlocking));
  // One copy per chunk of blocks, each followed by an event: the next evaluation starts on a chunk
  // as soon as its points have arrived. The call returns when the copies are queued (pageable
  // memory: when the runtime has staged them).
  const int num_chunks = static_cast<int>(B->chunk_block_begin.size()) - 1;
  for (int c = 0; c < num_chunks; ++c) {
    const int64_t p0 = B->offsets[B->chunk_block_begin[c]], p1 = B->offsets[B->chunk_block_begin[c + 1]];
    rc = LazyEvent(&B->chunk_in[c]);
    if (rc != HG_OK) return rc;
    if (p1 > p0) {
      HG_CUDA(cudaMemcpyAsync(B->d_points + p0 * B->stride, points + p0 * B->stride,
                              static_cast<size_t>(p1 - p0) * B->stride * sizeof(float), cudaMemcpyHostToDevice,
                              B->copy_in));
    }
    HG_CUDA(cudaEventRecord(B->chunk_in[c], B->copy_in));
  }
  B->points_pending = true;
  return HG_OK;
}

int hg_batch_evaluate(const hg_batch* B, const double* poses, int num_poses, const int32_t* pose_index0,
                      const int32_t* pose_index1, int device_outputs, double* residuals, double* jt0,
                      double* jq0, double* jt1, double* jq1, void* cuda_stream) {
  HG_REQUIRE(B != nullptr && poses != nullptr && num_poses >= 1, "hg_batch_evaluate: bad arguments");
  HG_ON_DEVICE(B->device);
  cudaStream_t stream = cuda_stream != nullptr ? static_cast<cudaStream_t>(cuda_stream) : B->stream;
  std::lock_guard<std::mutex> lock(B->mu);
  int slot = 0;
  int rc = PrepareBlocks(B, poses, num_poses, pose_index0, pose_index1, stream, &slot);
  if (rc != HG_OK) return rc;
  CostArgs a = BaseArgs(B);
  a.blocks = B->d_blocks[slot];
  a.tiles = B->d_tiles_out;
  const int64_t N = B->num_points;
  if (!B->two_pose) {
    jt1 = nullptr;
    jq1 = nullptr;
  }
  if (device_outputs) {
    rc = WaitForAllPoints(B, stream);
    if (rc != HG_OK) return rc;
    a.r = residuals; a.jt0 = jt0; a.jq0 = jq0; a.jt1 = jt1; a.jq1 = jq1;
    const uintptr_t rows = reinterpret_cast<uintptr_t>(jq0) | reinterpret_cast<uintptr_t>(jq1);
    a.vec_store = rows % 32 == 0 ? 2 : (rows % 16 == 0 ? 1 : 0);
    if (const char* env = std::getenv("HG_VEC_STORE_MAX")) a.vec_store = std::min(a.vec_store, std::atoi(env));
    rc = LaunchCost<false>(B, a, B->num_tiles_out, stream);
    if (rc != HG_OK) return rc;
    HG_CUDA(cudaEventRecord(B->slot_done[slot], stream));  // returns without waiting: see the slot ring
    return HG_OK;
  }
  // Host outputs: stage in device memory, then one copy per requested array.
  const size_t need = static_cast<size_t>(N) * 15;
  if (B->d_out_doubles < need) {
    rc = WaitForSlots(B);  // nothing queued may still write the old staging area
    if (rc != HG_OK) return rc;
    if (B->copy_out) HG_CUDA(cudaStreamSynchronize(B->copy_out));
    CachedFree(B->d_out);
    B->d_out = nullptr;
    B->d_out_doubles = 0;
    HG_ALLOC(B->d_out, CachedDeviceAlloc(std::max<size_t>(need, 1) * sizeof(double)));
    B->d_out_doubles = need;
  }
  // 4-wide arrays first so that they stay 32-byte aligned for any N.
  double* d = B->d_out;
  a.jq0 = jq0 ? d : nullptr;
  a.jq1 = jq1 ? d + 4 * N : nullptr;
  a.jt0 = jt0 ? d + 8 * N : nullptr;
  a.jt1 = jt1 ? d + 11 * N : nullptr;
  a.r = residuals ? d + 14 * N : nullptr;
  a.vec_store = 2;  // the staging area is 256-byte aligned and the 4-wide arrays come first
  // Small blocks (a scan matcher's Evaluate: ~1 MB into Ceres' pageable arrays): one DMA of the
  // whole staging area into pinned memory and host copies from there cost less than up to five
  // separate pageable device-to-host copies.
  constexpr size_t kPinnedStageDoubles = size_t(1) << 20;  // 8 MB
  if (need <= kPinnedStageDoubles) {
    rc = WaitForAllPoints(B, stream);
    if (rc != HG_OK) return rc;
    rc = LaunchCost<false>(B, a, B->num_tiles_out, stream);
    if (rc != HG_OK) return rc;
    if (B->h_out_doubles < need) {
      CachedFree(B->h_out);
      B->h_out = nullptr;
      B->h_out_doubles = 0;
      HG_ALLOC(B->h_out, CachedPinnedAlloc(std::max<size_t>(need, 1) * sizeof(double)));
      B->h_out_doubles = need;
    }
    HG_CUDA(cudaMemcpyAsync(B->h_out, d, need * sizeof(double), cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaEventRecord(B->slot_done[slot], stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    B->points_pending = false;
    const double* h = B->h_out;
    if (jq0) std::memcpy(jq0, h, sizeof(double) * N * 4);
    if (jq1) std::memcpy(jq1, h + 4 * N, sizeof(double) * N * 4);
    if (jt0) std::memcpy(jt0, h + 8 * N, sizeof(double) * N * 3);
    if (jt1) std::memcpy(jt1, h + 11 * N, sizeof(double) * N * 3);
    if (residuals) std::memcpy(residuals, h + 14 * N, sizeof(double) * N);
    return HG_OK;
  }
  // Large batches: chunk by chunk. A chunk's kernel waits only for that chunk's points (when
  // hg_batch_update_points queued them), and its outputs start their way to the host while the
  // next chunk computes and later chunks' points are still arriving: PCIe carries both directions
  // at once, so the step costs about what the device-to-host copies alone cost.
  if (B->copy_out == nullptr) HG_CUDA(cudaStreamCreateWithFlags(&B->copy_out, cudaStreamNonBlocking));
  const int num_chunks = static_cast<int>(B->chunk_block_begin.size()) - 1;
  const bool trace = std::getenv("HG_TRACE_E2E") != nullptr;
  std::vector<cudaEvent_t> tk, td;
  cudaEvent_t t_begin = nullptr;
  if (trace) {
    cudaEventCreate(&t_begin);
    cudaEventRecord(t_begin, stream);
  }
  for (int c = 0; c < num_chunks; ++c) {
    const int b0 = B->chunk_block_begin[c], b1 = B->chunk_block_begin[c + 1];
    const int t0 = B->out_tile_begin[b0], t1 = B->out_tile_begin[b1];
    const int64_t p0 = B->offsets[b0], n = B->offsets[b1] - p0;
    if (B->points_pending) HG_CUDA(cudaStreamWaitEvent(stream, B->chunk_in[c], 0));
    CostArgs ac = a;
    ac.tiles = B->d_tiles_out + t0;
    rc = LaunchCost<false>(B, ac, t1 - t0, stream);
    if (rc != HG_OK) return rc;
    rc = LazyEvent(&B->chunk_done[c]);
    if (rc != HG_OK) return rc;
    HG_CUDA(cudaEventRecord(B->chunk_done[c], stream));
    if (trace) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      cudaEventRecord(e, stream);
      tk.push_back(e);
    }
    HG_CUDA(cudaStreamWaitEvent(B->copy_out, B->chunk_done[c], 0));
    if (n == 0) continue;
    if (residuals)
      HG_CUDA(cudaMemcpyAsync(residuals + p0, a.r + p0, sizeof(double) * n, cudaMemcpyDeviceToHost, B->copy_out));
    if (jt0)
      HG_CUDA(cudaMemcpyAsync(jt0 + 3 * p0, a.jt0 + 3 * p0, sizeof(double) * n * 3, cudaMemcpyDeviceToHost, B->copy_out));
    if (jq0)
      HG_CUDA(cudaMemcpyAsync(jq0 + 4 * p0, a.jq0 + 4 * p0, sizeof(double) * n * 4, cudaMemcpyDeviceToHost, B->copy_out));
    if (jt1)
      HG_CUDA(cudaMemcpyAsync(jt1 + 3 * p0, a.jt1 + 3 * p0, sizeof(double) * n * 3, cudaMemcpyDeviceToHost, B->copy_out));
    if (jq1)
      HG_CUDA(cudaMemcpyAsync(jq1 + 4 * p0, a.jq1 + 4 * p0, sizeof(double) * n * 4, cudaMemcpyDeviceToHost, B->copy_out));
    if (trace) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      cudaEventRecord(e, B->copy_out);
      td.push_back(e);
    }
  }
  HG_CUDA(cudaEventRecord(B->slot_done[slot], stream));
  HG_CUDA(cudaStreamSynchronize(B->copy_out));
  HG_CUDA(cudaStreamSynchronize(stream));
  if (trace) {
    std::fprintf(stderr, "[hg trace] host-pointer Evaluate, ms after the first launch was queued: kernels done");
    for (cudaEvent_t e : tk) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, t_begin, e);
      std::fprintf(stderr, " %.2f", ms);
      cudaEventDestroy(e);
    }
    std::fprintf(stderr, "; copies out done");
    for (cudaEvent_t e : td) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, t_begin, e);
      std::fprintf(stderr, " %.2f", ms);
      cudaEventDestroy(e);
    }
    std::fprintf(stderr, "\n");
    cudaEventDestroy(t_begin);
  }
  B->points_pending = false;
  return HG_OK;
}

int hg_batch_normal_equations(const hg_batch* B, const double* poses, int num_poses,
                              const int32_t* pose_index0, const int32_t* pose_index1, double* JtJ,
                              double* Jtr, double* cost_out) {
  HG_REQUIRE(B != nullptr && poses != nullptr && num_poses >= 1, "hg_batch_normal_equations: bad arguments");
  HG_ON_DEVICE(B->device);
  std::lock_guard<std::mutex> lock(B->mu);
  cudaStream_t stream = B->stream;
  int slot = 0;
  int rc = PrepareBlocks(B, poses, num_poses, pose_index0, pose_index1, stream, &slot);
  if (rc != HG_OK) return rc;
  CostArgs a = BaseArgs(B);
  a.blocks = B->d_blocks[slot];
  a.tiles = B->d_tiles_ne;
  a.partials = B->d_partials;
  rc = WaitForAllPoints(B, stream);
  if (rc != HG_OK) return rc;
  rc = LaunchCost<true>(B, a, B->num_tiles_ne, stream);
  if (rc != HG_OK) return rc;
  const int n = B->two_pose ? 14 : 7;
  const int P = NumNormalTerms(n);
  ReducePartialsKernel<<<B->num_blocks, 128, 0, stream>>>(B->d_partials, B->d_ne_tile_begin, B->d_blocks[slot],
                                                          B->two_pose ? 1 : 0, B->d_ne_out);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpyAsync(B->h_ne_out, B->d_ne_out, sizeof(double) * P * B->num_blocks,
                          cudaMemcpyDeviceToHost, stream));
  HG_CUDA(cudaEventRecord(B->slot_done[slot], stream));
  HG_CUDA(cudaStreamSynchronize(stream));
  B->points_pending = false;
  for (int b = 0; b < B->num_blocks; ++b) {
    const double* s = B->h_ne_out + static_cast<size_t>(b) * P;
    int e = 0;
    for (int ca = 0; ca < n; ++ca)
      for (int cb = ca; cb < n; ++cb, ++e) {
        if (JtJ) {
          JtJ[(static_cast<size_t>(b) * n + ca) * n + cb] = s[e];
          JtJ[(static_cast<size_t>(b) * n + cb) * n + ca] = s[e];
        }
      }
    for (int c = 0; c < n; ++c, ++e)
      if (Jtr) Jtr[static_cast<size_t>(b) * n + c] = s[e];
    if (cost_out) cost_out[b] = 0.5 * s[e];
  }
  return HG_OK;
}

int hg_batch_level_usage(const hg_batch* B, const double* poses, int num_poses, const int32_t* pose_index0,
                         const int32_t* pose_index1, int64_t* counts) {
  HG_REQUIRE(B != nullptr && poses != nullptr && num_poses >= 1 && counts != nullptr,
             "hg_batch_level_usage: bad arguments");
  HG_REQUIRE(B->lookup == 2, "hg_batch_level_usage: only multi-resolution TSDF cost functions have levels");
  HG_ON_DEVICE(B->device);
  std::lock_guard<std::mutex> lock(B->mu);
  cudaStream_t stream = B->stream;
  int slot = 0;
  int rc = PrepareBlocks(B, poses, num_poses, pose_index0, pose_index1, stream, &slot);
  if (rc != HG_OK) return rc;
  CostArgs a = BaseArgs(B);
  a.blocks = B->d_blocks[slot];
  a.tiles = B->d_tiles_out;
  rc = WaitForAllPoints(B, stream);
  if (rc != HG_OK) return rc;
  unsigned long long* d_counts = static_cast<unsigned long long*>(
      CachedDeviceAlloc(sizeof(unsigned long long) * (HG_MAX_PYRAMID_LEVELS + 1)));
  HG_REQUIRE(d_counts != nullptr, "hg_batch_level_usage: out of device memory");
  struct Release {
    void* p;
    ~Release() { CachedFree(p); }
  } release{d_counts};
  HG_CUDA(cudaMemsetAsync(d_counts, 0, sizeof(unsigned long long) * (HG_MAX_PYRAMID_LEVELS + 1), stream));
  if (B->num_tiles_out > 0) {
    LevelUsageKernel<<<B->num_tiles_out, kThreads, 0, stream>>>(a, d_counts);
    CountLaunch();
    HG_CUDA(cudaGetLastError());
  }
  unsigned long long h[HG_MAX_PYRAMID_LEVELS + 1];
  HG_CUDA(cudaMemcpyAsync(h, d_counts, sizeof(h), cudaMemcpyDeviceToHost, stream));
  HG_CUDA(cudaStreamSynchronize(stream));
  for (int l = 0; l <= B->num_grids; ++l) counts[l] = static_cast<int64_t>(h[l]);
  return HG_OK;
}

// ---- single residual block = batch of one ---------------------------------

int hg_cost_create(int kind, double scaling_factor, const float* points, int stride_floats,
                   int num_points, const hg_grid* const* grids, int num_grids,
                   double interpolation_ratio, hg_cost** out) {
  HG_REQUIRE(out != nullptr && num_points >= 0, "hg_cost_create: bad arguments");
  const int64_t offsets[2] = {0, num_points};
  hg_batch* B = nullptr;
  const int rc = hg_batch_create(kind, 1, points, stride_floats, offsets, &scaling_factor,
                                 &interpolation_ratio, grids, num_grids, &B);
  if (rc != HG_OK) return rc;
  hg_cost* c = new hg_cost;
  c->batch = B;
  *out = c;
  return HG_OK;
}

void hg_cost_destroy(hg_cost* cost) {
  if (cost == nullptr) return;
  hg_batch_destroy(cost->batch);
  delete cost;
}

int hg_cost_num_residuals(const hg_cost* cost) { return cost ? static_cast<int>(cost->batch->num_points) : 0; }
int hg_cost_num_parameter_blocks(const hg_cost* cost) { return cost ? (cost->batch->two_pose ? 4 : 2) : 0; }

static void GatherPoses(const hg_cost* cost, const double* const* parameters, double* poses) {
  std::memcpy(poses, parameters[0], 3 * sizeof(double));
  std::memcpy(poses + 3, parameters[1], 4 * sizeof(double));
  if (cost->batch->two_pose) {
    std::memcpy(poses + 7, parameters[2], 3 * sizeof(double));
    std::memcpy(poses + 10, parameters[3], 4 * sizeof(double));
  }
}

int hg_cost_evaluate(const hg_cost* cost, const double* const* parameters, double* residuals,
                     double** jacobians) {
  HG_REQUIRE(cost != nullptr && parameters != nullptr, "hg_cost_evaluate: bad arguments");
  double poses[14];
  GatherPoses(cost, parameters, poses);
  const int32_t i0 = 0, i1 = 1;
  const bool two = cost->batch->two_pose;
  return hg_batch_evaluate(cost->batch, poses, two ? 2 : 1, &i0, &i1, 0, residuals,
                           jacobians ? jacobians[0] : nullptr, jacobians ? jacobians[1] : nullptr,
                           (jacobians && two) ? jacobians[2] : nullptr,
                           (jacobians && two) ? jacobians[3] : nullptr, nullptr);
}

int hg_cost_evaluate_device(const hg_cost* cost, const double* const* parameters, double* d_residuals,
                            double* const* d_jacobians, void* cuda_stream) {
  HG_REQUIRE(cost != nullptr && parameters != nullptr, "hg_cost_evaluate_device: bad arguments");
  double poses[14];
  GatherPoses(cost, parameters, poses);
  const int32_t i0 = 0, i1 = 1;
  const bool two = cost->batch->two_pose;
  return hg_batch_evaluate(cost->batch, poses, two ? 2 : 1, &i0, &i1, 1, d_residuals,
                           d_jacobians ? d_jacobians[0] : nullptr, d_jacobians ? d_jacobians[1] : nullptr,
                           (d_jacobians && two) ? d_jacobians[2] : nullptr,
                           (d_jacobians && two) ? d_jacobians[3] : nullptr, cuda_stream);
}

int hg_cost_normal_equations(const hg_cost* cost, const double* const* parameters, double* JtJ,
                             double* Jtr, double* cost_out) {
  HG_REQUIRE(cost != nullptr && parameters != nullptr, "hg_cost_normal_equations: bad arguments");
  double poses[14];
  GatherPoses(cost, parameters, poses);
  const int32_t i0 = 0, i1 = 1;
  return hg_batch_normal_equations(cost->batch, poses, cost->batch->two_pose ? 2 : 1, &i0, &i1, JtJ, Jtr,
                                   cost_out);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// CeresScanMatcher3D::Match on the dense LM.

namespace {

// Minimal cost-function shape for hg::DenseLmProblem.
class LmCost {
 public:
  virtual ~LmCost() {}
  virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
  virtual bool NormalEquations(double const* const*, double*, double*, double*) const { return false; }
  virtual int num_residuals() const = 0;
};

class GpuScanCost : public LmCost {
 public:
  explicit GpuScanCost(hg_cost* c) : c_(c) {}
  ~GpuScanCost() override { hg_cost_destroy(c_); }
  bool Evaluate(double const* const* p, double* r, double** J) const override {
    const bool ok = hg_cost_evaluate(c_, p, r, J) == HG_OK;
    if (!ok) failed_ = true;
    return ok;
  }
  // Fused J'J / J'r / cost on the device: the solver never needs the rows.
  bool NormalEquations(double const* const* p, double* JtJ, double* Jtr, double* cost) const override {
    const bool ok = hg_cost_normal_equations(c_, p, JtJ, Jtr, cost) == HG_OK;
    if (!ok) failed_ = true;
    return ok;
  }
  bool failed() const { return failed_; }
  int num_residuals() const override { return hg_cost_num_residuals(c_); }

 private:
  hg_cost* c_;
  mutable bool failed_ = false;  // a device evaluation returned an error
};

// translation_delta_cost_functor_3d.h:31-45
class TranslationDelta : public LmCost {
 public:
  TranslationDelta(double s, const double* target) : s_(s) { std::memcpy(t_, target, sizeof(t_)); }
  bool Evaluate(double const* const* p, double* r, double** J) const override {
    for (int k = 0; k < 3; ++k) r[k] = s_ * (p[0][k] - t_[k]);
    if (J && J[0]) {
      for (int k = 0; k < 9; ++k) J[0][k] = 0.0;
      J[0][0] = J[0][4] = J[0][8] = s_;
    }
    return true;
  }
  int num_residuals() const override { return 3; }

 private:
  double s_, t_[3];
};

// rotation_delta_cost_functor_3d.h:35-54 with common/math.h:74-81: residual =
// s * imag(target^-1 * q); linear in q, so the Jacobian is the constant matrix below.
class RotationDelta : public LmCost {
 public:
  RotationDelta(double s, const double* target_wxyz) : s_(s) {
    z_[0] = target_wxyz[0];
    z_[1] = -target_wxyz[1];
    z_[2] = -target_wxyz[2];
    z_[3] = -target_wxyz[3];
  }
  bool Evaluate(double const* const* p, double* r, double** J) const override {
    const double* w = p[0];
    const double* z = z_;
    r[0] = s_ * (z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2]);
    r[1] = s_ * (z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1]);
    r[2] = s_ * (z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0]);
    if (J && J[0]) {
      const double M[12] = {z[1], z[0], -z[3], z[2], z[2], z[3], z[0], -z[1], z[3], -z[2], z[1], z[0]};
      for (int k = 0; k < 12; ++k) J[0][k] = s_ * M[k];
    }
    return true;
  }
  int num_residuals() const override { return 3; }

 private:
  double s_, z_[4];
};

}  // namespace

extern "C" int hg_scan_match(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                             const hg_grid* const* grids, const double* occupied_space_weights,
                             double translation_weight, double rotation_weight, int only_optimize_yaw,
                             int max_num_iterations, int use_nonmonotonic_steps,
                             const double* target_translation, const double* initial_pose,
                             double* pose_out, double* summary_out) {
  HG_REQUIRE(num_clouds >= 1 && clouds && cloud_sizes && grids && occupied_space_weights && initial_pose &&
                 pose_out && target_translation,
             "hg_scan_match: bad arguments");
  for (int i = 0; i < num_clouds; ++i) {
    HG_REQUIRE(occupied_space_weights[i] > 0., "hg_scan_match: occupied_space_weight must be > 0");
    HG_REQUIRE(cloud_sizes[i] > 0 && clouds[i] != nullptr && grids[i] != nullptr, "hg_scan_match: empty point cloud");
  }
  // The whole solve in one launch (scan_match.cu) unless the problem does not fit the kernel or
  // HG_SCAN_MATCH_HOST=1 asks for the host loop over the fused normal equations (same solver).
  const bool host_loop = std::getenv("HG_SCAN_MATCH_HOST") != nullptr;  // read per call: tests switch it
  if (!host_loop) {
    const int rc = ScanMatchOnDevice(num_clouds, clouds, cloud_sizes, grids, occupied_space_weights,
                                     translation_weight, rotation_weight, only_optimize_yaw, max_num_iterations,
                                     use_nonmonotonic_steps, target_translation, initial_pose, pose_out,
                                     summary_out);
    if (rc != HG_ERR_LIMIT) return rc;
  }
  double t[3] = {initial_pose[0], initial_pose[1], initial_pose[2]};
  double q[4] = {initial_pose[3], initial_pose[4], initial_pose[5], initial_pose[6]};
  hg::DenseLmProblem<LmCost> problem;
  const int tb = problem.AddParameterBlock(t, 3, hg::LocalParameterization::kIdentity);
  const int qb = problem.AddParameterBlock(q, 4, only_optimize_yaw ? hg::LocalParameterization::kYawOnlyQuaternion
                                                                   : hg::LocalParameterization::kQuaternion);
  std::vector<std::unique_ptr<LmCost>> owned;
  std::vector<const GpuScanCost*> scan_costs;
  for (int i = 0; i < num_clouds; ++i) {
    // ceres_scan_matcher_3d.cc:138-139 / :149-150
    const double scaling = occupied_space_weights[i] / std::sqrt(static_cast<double>(cloud_sizes[i]));
    hg_cost* c = nullptr;
    const int kind = grids[i]->dev.kind == kGridProb ? HG_COST_OCCUPIED_SPACE : HG_COST_TSDF_SPACE;
    const int rc = hg_cost_create(kind, scaling, clouds[i], 3, cloud_sizes[i], &grids[i], 1, 0.0, &c);
    if (rc == HG_OK) c->batch->stream = ThreadStream();  // concurrent matches do not share the legacy stream
    if (rc != HG_OK) return rc;
    scan_costs.push_back(new GpuScanCost(c));
    owned.emplace_back(const_cast<GpuScanCost*>(scan_costs.back()));
    problem.AddResidualBlock(owned.back().get(), {tb, qb});
  }
  owned.emplace_back(new TranslationDelta(translation_weight, target_translation));
  problem.AddResidualBlock(owned.back().get(), {tb});
  owned.emplace_back(new RotationDelta(rotation_weight, initial_pose + 3));
  problem.AddResidualBlock(owned.back().get(), {qb});
  hg::LmOptions options;
  options.max_num_iterations = max_num_iterations;
  options.use_nonmonotonic_steps = use_nonmonotonic_steps != 0;
  const hg::LmSummary s = problem.Solve(options);
  // Ceres' own FAILURE (five invalid steps in a row) still yields the best pose seen, and the
  // reference does not look at the termination type; only a failed device call is an error here.
  for (const GpuScanCost* c : scan_costs) {
    if (c->failed()) {
      SetError(std::string("hg_scan_match: evaluation failed: ") + hg_last_error());
      return HG_ERR_CUDA;
    }
  }
  for (int k = 0; k < 3; ++k) pose_out[k] = t[k];
  for (int k = 0; k < 4; ++k) pose_out[3 + k] = q[k];
  if (summary_out) {
    summary_out[0] = s.initial_cost;
    summary_out[1] = s.final_cost;
    summary_out[2] = s.num_iterations;
    summary_out[3] = s.num_successful_steps;
    summary_out[4] = s.num_cost_evaluations;
    summary_out[5] = s.num_jacobian_evaluations;
  }
  return HG_OK;
}

// CeresScanMatcher3D::Evaluate (ceres_scan_matcher_3d.cc:97-118): the same problem as Match,
// evaluated once at the initial pose — cost, all residuals (clouds in order, then the translation
// and rotation deltas), and what the reference asks ceres::Problem::Evaluate for in its fourth
// argument, which is the GRADIENT of the cost in the tangent space of the pose.
extern "C" int hg_scan_match_evaluate(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                                      const hg_grid* const* grids, const double* occupied_space_weights,
                                      double translation_weight, double rotation_weight, int only_optimize_yaw,
                                      const double* target_translation, const double* initial_pose, double* cost,
                                      double* residuals, double* gradient) {
  HG_REQUIRE(num_clouds >= 1 && clouds && cloud_sizes && grids && occupied_space_weights && initial_pose &&
                 target_translation,
             "hg_scan_match_evaluate: bad arguments");
  double t[3] = {initial_pose[0], initial_pose[1], initial_pose[2]};
  double q[4] = {initial_pose[3], initial_pose[4], initial_pose[5], initial_pose[6]};
  hg::DenseLmProblem<LmCost> problem;
  const int tb = problem.AddParameterBlock(t, 3, hg::LocalParameterization::kIdentity);
  const int qb = problem.AddParameterBlock(q, 4, only_optimize_yaw ? hg::LocalParameterization::kYawOnlyQuaternion
                                                                   : hg::LocalParameterization::kQuaternion);
  std::vector<std::unique_ptr<LmCost>> owned;
  for (int i = 0; i < num_clouds; ++i) {
    HG_REQUIRE(occupied_space_weights[i] > 0., "hg_scan_match_evaluate: occupied_space_weight must be > 0");
    HG_REQUIRE(cloud_sizes[i] > 0 && clouds[i] != nullptr && grids[i] != nullptr,
               "hg_scan_match_evaluate: empty point cloud");
    const double scaling = occupied_space_weights[i] / std::sqrt(static_cast<double>(cloud_sizes[i]));
    hg_cost* c = nullptr;
    const int kind = grids[i]->dev.kind == kGridProb ? HG_COST_OCCUPIED_SPACE : HG_COST_TSDF_SPACE;
    const int rc = hg_cost_create(kind, scaling, clouds[i], 3, cloud_sizes[i], &grids[i], 1, 0.0, &c);
    if (rc != HG_OK) return rc;
    c->batch->stream = ThreadStream();
    owned.emplace_back(new GpuScanCost(c));
    problem.AddResidualBlock(owned.back().get(), {tb, qb});
  }
  owned.emplace_back(new TranslationDelta(translation_weight, target_translation));
  problem.AddResidualBlock(owned.back().get(), {tb});
  owned.emplace_back(new RotationDelta(rotation_weight, initial_pose + 3));
  problem.AddResidualBlock(owned.back().get(), {qb});
  double c = 0.0;
  std::vector<double> r, J;
  if (!problem.Evaluate(&c, &r, &J)) {
    SetError(std::string("hg_scan_match_evaluate: evaluation failed: ") + hg_last_error());
    return HG_ERR_CUDA;
  }
  if (cost != nullptr) *cost = c;
  if (residuals != nullptr) std::copy(r.begin(), r.end(), residuals);
  if (gradient != nullptr) {
    const int n = problem.NumLocalParameters();
    for (int k = 0; k < n; ++k) gradient[k] = 0.0;
    for (size_t i = 0; i < r.size(); ++i)
      for (int k = 0; k < n; ++k) gradient[k] += J[i * n + k] * r[i];
  }
  return HG_OK;
}

// ---------------------------------------------------------------------------
// hg_cost_normal_equations_static — the Gauss-Newton terms of one residual
// block computed with NO allocation of any kind at call time.
//
// Who needs this: callers that are not allowed to touch the CUDA allocator once
// they are running — a SLAM front end inside its real-time loop (cudaMalloc and
// cudaFree synchronise the whole device and take an unbounded time), or a host
// program that has already handed all free device memory to its own arena, so
// that a cudaMalloc issued by a library would simply fail. For them the scratch
// memory the kernels need (per-tile partial sums, the reduced result and the
// block descriptor) is not allocated at all: it is the file-scope device array
// below, which every GPU that loads this library carries from load to exit.
//
// Limit: one residual block of at most kStaticMaxPoints = 1,048,576 points
// (the array is sized for it: 984 KB per GPU). Larger blocks are rejected with
// HG_ERR_LIMIT; use hg_cost_normal_equations for those.
//
// Sharing: there is ONE such array per GPU. Every call uses it and leaves its
// last contents behind. Nothing in it has to start at zero — each call writes
// every word it later reads (all terms of every tile it launches, then the
// reduced vector) — so no clearing pass is needed. Calls that overlapped on the
// GPU would race: they would overwrite each other's partial sums. Within one
// process that cannot happen, because a call holds a process-wide lock and
// returns only after the device has finished its work.
//
// Set-up: the first call on a device looks up the array's device address with
// cudaGetSymbolAddress (the array's name is only a host-side stand-in in host
// code, never a pointer a kernel may be given) while that device is current,
// and caches it. That first call should therefore happen before any stream
// capture is begun on the device.
namespace {

constexpr int kStaticMaxPoints = 1 << 20;
constexpr int kStaticMaxTiles = kStaticMaxPoints / kTileNe;
constexpr int kStaticTerms = NumNormalTerms(14);
constexpr int kStaticBlockDoubles = (sizeof(BlockDev) + 7) / 8;
constexpr int kStaticDoubles = kStaticMaxTiles * kStaticTerms + kStaticTerms + kStaticBlockDoubles;

static __device__ __align__(256) double g_static_scratch[kStaticDoubles];

constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;
double* g_static_address[kMaxDevices] = {nullptr};

// Device address of g_static_scratch on the current device (cached per device).
int StaticScratchAddress(int device, double** out) {
  HG_REQUIRE(device >= 0 && device < kMaxDevices, "hg_cost_normal_equations_static: device index out of range");
  if (g_static_address[device] == nullptr) {
    double* address = nullptr;
    HG_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&address), g_static_scratch));
    g_static_address[device] = address;
  }
  *out = g_static_address[device];
  return HG_OK;
}

}  // namespace

extern "C" int hg_cost_normal_equations_static(const hg_cost* cost, const double* const* parameters,
                                               double* JtJ, double* Jtr, double* cost_out) {
  HG_REQUIRE(cost != nullptr && parameters != nullptr, "hg_cost_normal_equations_static: bad arguments");
  const hg_batch* B = cost->batch;
  if (B->num_points > kStaticMaxPoints) {
    SetError("hg_cost_normal_equations_static: the block has " + std::to_string(B->num_points) +
             " points; the static scratch array holds at most " + std::to_string(kStaticMaxPoints) +
             " (use hg_cost_normal_equations)");
    return HG_ERR_LIMIT;
  }
  int device = -1;
  HG_CUDA(cudaGetDevice(&device));
  HG_REQUIRE(device == B->device,
             "hg_cost_normal_equations_static: the cost function's device must be the current device");
  std::lock_guard<std::mutex> lock(g_static_mutex);
  double* scratch = nullptr;
  int rc = StaticScratchAddress(device, &scratch);
  if (rc != HG_OK) return rc;
  double* d_partials = scratch;
  double* d_reduced = scratch + static_cast<size_t>(kStaticMaxTiles) * kStaticTerms;
  BlockDev* d_block = reinterpret_cast<BlockDev*>(d_reduced + kStaticTerms);

  // Block descriptor on the stack, copied straight into the static array.
  const BlockPose p = B->two_pose ? TwoPose(parameters[0], parameters[1], parameters[2], parameters[3], B->ratios[0])
                                  : OnePose(parameters[0], parameters[1]);
  BlockDev block;
  std::memcpy(block.t, p.t, sizeof(block.t));
  std::memcpy(block.q, p.q, sizeof(block.q));
  block.ct0 = p.ct0;
  block.ct1 = p.ct1;
  std::memcpy(block.m0, p.m0, sizeof(block.m0));
  std::memcpy(block.m1, p.m1, sizeof(block.m1));
  block.s0 = p.s0;
  block.s1 = p.s1;
  std::memcpy(block.c, p.c, sizeof(block.c));
  std::memcpy(block.q0, p.q0, sizeof(block.q0));
  std::memcpy(block.q1, p.q1, sizeof(block.q1));
  block.scale = B->scales[0];
  block.point_offset = 0;
  block.n = static_cast<int>(B->num_points);
  block.pad = 0;
  cudaStream_t stream = nullptr;
  HG_CUDA(cudaMemcpyAsync(d_block, &block, sizeof(block), cudaMemcpyHostToDevice, stream));

  CostArgs a = BaseArgs(B, /*build_stencils=*/false);  // this entry point never allocates
  a.blocks = d_block;
  a.tiles = B->d_tiles_ne;
  a.partials = d_partials;
  rc = WaitForAllPoints(B, stream);  // points queued by hg_batch_update_points (an event wait: no allocation)
  if (rc != HG_OK) return rc;
  rc = LaunchCost<true>(B, a, B->num_tiles_ne, stream);
  if (rc != HG_OK) return rc;
  const int n = B->two_pose ? 14 : 7;
  const int P = NumNormalTerms(n);
  ReducePartialsKernel<<<1, 128, 0, stream>>>(d_partials, B->d_ne_tile_begin, d_block, B->two_pose ? 1 : 0,
                                              d_reduced);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  double reduced[kStaticTerms];
  HG_CUDA(cudaMemcpyAsync(reduced, d_reduced, sizeof(double) * P, cudaMemcpyDeviceToHost, stream));
  HG_CUDA(cudaStreamSynchronize(stream));
  int e = 0;
  for (int ca = 0; ca < n; ++ca)
    for (int cb = ca; cb < n; ++cb, ++e) {
      if (JtJ) {
        JtJ[ca * n + cb] = reduced[e];
        JtJ[cb * n + ca] = reduced[e];
      }
    }
  for (int c = 0; c < n; ++c, ++e)
    if (Jtr) Jtr[c] = reduced[e];
  if (cost_out) *cost_out = 0.5 * reduced[e];
  return HG_OK;
}
