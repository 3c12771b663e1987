// FastCorrelativeScanMatcher3D on the device.
//
// Reference behaviour reproduced (paths under
// /root/reference/cartographer/mapping/internal/3d/scan_matching unless noted):
//   fast_correlative_scan_matcher_3d.cc:157-228  Match / MatchFullSubmap / MatchWithSearchParameters
//   ... :230-274   DiscretizeScan (float point transform, GetCellIndex, low-resolution index fold)
//   ... :276-326   GenerateDiscreteScans (angular step, yaw list, rotational prefilter, per-yaw pose)
//   ... :328-398   lowest-resolution candidates, ScoreCandidates (integer gather sum, float score, sort)
//   ... :400-471   GetPoseFromCandidate, BranchAndBound (DFS order, pruning, leaf gate)
//   low_resolution_matcher.cc:24-69               leaf gate on the low-resolution grid
//   rotational_scan_matcher.cc:121-185            RotateHistogram / MatchHistograms / Match
//
// Division of labour. The data-parallel parts run as CUDA kernels: rotating and
// discretising every scan for every yaw (bit-exact float steps), scoring
// candidates (integer gather sums over the uint8 pyramid), and the
// low-resolution gate. The branch-and-bound itself is replayed on the host in
// EXACTLY the reference's depth-first order over scores the GPU produced in
// large speculative waves: whenever the replay needs the children of a node
// that has not been scored yet, the pending part of the search frontier is
// expanded level by level on the GPU (one launch per level over all pending
// parents whose score still beats the current bound), so the replay consumes a
// superset of what the reference's DFS scores and returns the same candidate,
// including its tie-breaking. Integer sums are exact; float scores are computed
// from them with the reference's expression, unfused.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <memory>

#include <chrono>
#include <cstdlib>

#include "device_math.cuh"
#include "host/float_geometry.h"

namespace hg {
namespace {

using fgeom::Q4;
using fgeom::Rigid;
using fgeom::V3;

// HG_TRACE=1 prints phase timings of every search to stderr.
inline bool TraceEnabled() {
  static const bool on = std::getenv("HG_TRACE") != nullptr;
  return on;
}
struct PhaseTimer {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  double Lap() {
    const auto t1 = std::chrono::steady_clock::now();
    const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    t0 = t1;
    return ms;
  }
};

// PrecomputationGrid3D::ToProbability (precomputation_grid_3d.h:33-36) applied to
// sum / float(N) (fast_correlative_scan_matcher_3d.cc:380-382).
inline float ScoreFromSum(const int sum, const int num_points) {
  const float kMin = 0.1f;
  const float kMax = 1.f - kMin;
  const float value = sum / static_cast<float>(num_points);
  return kMin + value * ((kMax - kMin) / 255.f);
}

// ---------------------------------------------------------------- kernels --

// DiscretizeScan for every (scan, point): cells of the unique levels.
// Layout: cells[(level * S + s) * N + i] as int3; level 0 = full resolution
// (shared by depths 0..full_resolution_depth-1), level l >= 1 = reduction
// exponent l.
__global__ void __launch_bounds__(256) DiscretizeKernel(const float* __restrict__ points, const int N,
                                                        const float* __restrict__ poses, const int S,
                                                        const float resolution, const float inv_resolution, const int num_low_levels,
                                                        const int start_xy, const int start_z,
                                                        int* __restrict__ cells) {
  const int64_t gid = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (gid >= static_cast<int64_t>(S) * N) return;
  const int s = static_cast<int>(gid / N);
  const int i = static_cast<int>(gid - static_cast<int64_t>(s) * N);
  const float* P = poses + 7 * s;
  float wx, wy, wz;
  RotateF(P[3], P[4], P[5], P[6], points[3 * i], points[3 * i + 1], points[3 * i + 2], &wx, &wy, &wz);
  wx = __fadd_rn(wx, P[0]);
  wy = __fadd_rn(wy, P[1]);
  wz = __fadd_rn(wz, P[2]);
  const int cx = CellIndexFast(wx, resolution, inv_resolution), cy = CellIndexFast(wy, resolution, inv_resolution),
            cz = CellIndexFast(wz, resolution, inv_resolution);
  int* o = cells + gid * 3;
  o[0] = cx; o[1] = cy; o[2] = cz;
  const int64_t level_stride = static_cast<int64_t>(S) * N * 3;
  for (int l = 1; l <= num_low_levels; ++l) {
    o += level_stride;
    o[0] = ((cx + start_xy) >> l) - (start_xy >> l);
    o[1] = ((cy + start_xy) >> l) - (start_xy >> l);
    o[2] = ((cz + start_z) >> l) - (start_z >> l);
  }
}

// ScoreCandidates' inner loop: one warp per candidate, lanes stride over the
// scan's points, integer gather from the uint8 level, warp-shuffle sum.
__global__ void __launch_bounds__(256) ScoreKernel(const BrickTable level, const int* __restrict__ cells,
                                                   const int N, const int reduction_exponent,
                                                   const int4* __restrict__ candidates,
                                                   const int64_t num_candidates, int* __restrict__ sums) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= num_candidates) return;
  const int4 c = candidates[warp];
  const int ox = c.y >> reduction_exponent, oy = c.z >> reduction_exponent, oz = c.w >> reduction_exponent;
  const int* base = cells + static_cast<int64_t>(c.x) * N * 3;
  int sum = 0;
  for (int i = lane; i < N; i += 32) {
    const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
    sum += LoadU8(level, x + ox, y + oy, z + oz);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  if (lane == 0) sums[warp] = sum;
}

// The lowest-resolution candidates of GenerateLowestResolutionCandidates
// (:328-361) are a regular lattice (scan, z, y, x with step 2^max_depth), so the
// kernel derives each candidate from its index in generation order instead of
// reading a candidate list: nothing but the integer sums crosses PCIe.
__global__ void __launch_bounds__(256) ScoreLatticeKernel(const BrickTable level, const int* __restrict__ cells,
                                                          const int N, const int reduction_exponent,
                                                          const int nxy, const int nz, const int wxy,
                                                          const int wz, const int step,
                                                          const int64_t num_candidates, int* __restrict__ sums) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= num_candidates) return;
  const int per_scan = nxy * nxy * nz;
  const int scan = static_cast<int>(warp / per_scan);
  int rem = static_cast<int>(warp - static_cast<int64_t>(scan) * per_scan);
  const int iz = rem / (nxy * nxy);
  rem -= iz * nxy * nxy;
  const int iy = rem / nxy;
  const int ix = rem - iy * nxy;
  const int ox = (-wxy + ix * step) >> reduction_exponent, oy = (-wxy + iy * step) >> reduction_exponent,
            oz = (-wz + iz * step) >> reduction_exponent;
  const int* base = cells + static_cast<int64_t>(scan) * N * 3;
  int sum = 0;
  for (int i = lane; i < N; i += 32) {
    const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
    sum += LoadU8(level, x + ox, y + oy, z + oz);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  if (lane == 0) sums[warp] = sum;
}

// Same sums as ScoreLatticeKernel for a level that has a dense copy: one warp
// scores KX x-adjacent lattice candidates at once, so that a point's cell is
// loaded and range-checked in y and z once for KX gathers, and a gather is one
// byte load without the brick-index indirection.
template <int KX>
__global__ void __launch_bounds__(256) ScoreLatticeDenseKernel(const DenseLevel level, const int* __restrict__ cells,
                                                               const int N, const int reduction_exponent,
                                                               const int nxy, const int nz, const int wxy,
                                                               const int wz, const int step,
                                                               const int64_t num_groups, int* __restrict__ sums) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= num_groups) return;
  const int gx = (nxy + KX - 1) / KX;  // groups per lattice row
  const int rows_per_scan = nxy * nz;
  const int64_t row = warp / gx;  // (scan, iz, iy)
  const int g = static_cast<int>(warp - row * gx);
  const int scan = static_cast<int>(row / rows_per_scan);
  const int rem = static_cast<int>(row - static_cast<int64_t>(scan) * rows_per_scan);
  const int iz = rem / nxy, iy = rem - iz * nxy;
  const int oy = ((-wxy + iy * step) >> reduction_exponent) - level.min[1];
  const int oz = ((-wz + iz * step) >> reduction_exponent) - level.min[2];
  int ox[KX];
#pragma unroll
  for (int k = 0; k < KX; ++k) ox[k] = ((-wxy + (g * KX + k) * step) >> reduction_exponent) - level.min[0];
  const int* base = cells + static_cast<int64_t>(scan) * N * 3;
  int sum[KX];
#pragma unroll
  for (int k = 0; k < KX; ++k) sum[k] = 0;
  for (int i = lane; i < N; i += 32) {
    const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1) + oy, z = __ldg(base + 3 * i + 2) + oz;
    if (static_cast<unsigned>(y) < static_cast<unsigned>(level.dim[1]) &&
        static_cast<unsigned>(z) < static_cast<unsigned>(level.dim[2])) {
      const unsigned char* rowp = level.data + (static_cast<size_t>(z) * level.dim[1] + y) * level.dim[0];
#pragma unroll
      for (int k = 0; k < KX; ++k) {
        const int xx = x + ox[k];
        if (static_cast<unsigned>(xx) < static_cast<unsigned>(level.dim[0])) sum[k] += __ldg(rowp + xx);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < KX; ++k) {
    int v = sum[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    const int ix = g * KX + k;
    if (lane == 0 && ix < nxy) sums[row * nxy + ix] = v;
  }
}

// Lowest-resolution level re-laid out for the lattice: lattice candidates that are adjacent in
// x are 'cs' = step >> reduction_exponent level cells apart, so the voxels one scan point reads
// for a whole lattice row are every cs-th cell of one x-row of the level. Each x-row is
// therefore stored as cs residue rows (cells x with x mod cs = r, in order of x / cs): the
// voxels of consecutive x-candidates become CONSECUTIVE BYTES, four per 32-bit load. Rows
// carry kLatticePadLeft zero bytes before and kLatticePadRight after their payload, so that any
// window of up to 32 candidates that touches the payload can be read without bounds checks.
// One pyramid level as the kernels read it; the matcher keeps an array of them on the device so that
// one launch can walk several levels.
struct LevelDev {
  BrickTable table;
  DenseLevel dense;
};

struct LatticeRows {
  const unsigned* data = nullptr;  // [(z * dim1 + y) * cs + r][row_words]
  int min[3] = {0, 0, 0};
  int dim[3] = {0, 0, 0};
  int cs_log2 = 0;
  int row_words = 0;
  int payload = 0;  // bytes of payload per residue row: ceil(dim0 / cs)
};
constexpr int kLatticePadLeft = 32, kLatticePadRight = 40;

__global__ void __launch_bounds__(256) BuildLatticeRowsKernel(const DenseLevel dense, const LatticeRows rows,
                                                              const int64_t total_bytes,
                                                              unsigned char* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= total_bytes) return;
  const int row_bytes = rows.row_words * 4;
  const int64_t row = i / row_bytes;
  const int j = static_cast<int>(i - row * row_bytes);
  const int r = static_cast<int>(row & ((1 << rows.cs_log2) - 1));
  const int64_t zy = row >> rows.cs_log2;
  const int q = j - kLatticePadLeft;
  const int x = (q << rows.cs_log2) + r;
  unsigned char v = 0;
  if (q >= 0 && x < dense.dim[0]) v = dense.data[zy * dense.dim[0] + x];
  out[i] = v;
}

// Same sums as ScoreLatticeKernel, from the residue rows: one warp scores KX = 4 * KXW
// x-adjacent lattice candidates; a lane reads, per scan point, KXW + 1 aligned words that hold
// the KX voxels, realigns them with funnel shifts and adds them four at a time as two packed
// pairs of 16-bit sums (flushed to 32-bit accumulators before they can overflow).
template <int KXW, bool ONE_CHUNK>
__global__ void __launch_bounds__(256, ONE_CHUNK ? 8 : 4)
    ScoreLatticePackedKernel(const LatticeRows rows, const int* __restrict__ cells, const int N,
                             const int reduction_exponent, const int nxy, const int nz, const int wxy, const int wz,
                             const int step, const int64_t num_groups, int* __restrict__ sums) {
  constexpr int KX = 4 * KXW;
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= num_groups) return;
  const int gx = (nxy + KX - 1) / KX;  // groups per lattice row
  const int rows_per_scan = nxy * nz;
  const int64_t row = warp / gx;  // (scan, iz, iy)
  const int g = static_cast<int>(warp - row * gx);
  const int scan = static_cast<int>(row / rows_per_scan);
  const int rem = static_cast<int>(row - static_cast<int64_t>(scan) * rows_per_scan);
  const int iz = rem / nxy, iy = rem - iz * nxy;
  const int oy = ((-wxy + iy * step) >> reduction_exponent) - rows.min[1];
  const int oz = ((-wz + iz * step) >> reduction_exponent) - rows.min[2];
  const int ox0 = ((-wxy + (g * KX) * step) >> reduction_exponent) - rows.min[0];
  const int cs_mask = (1 << rows.cs_log2) - 1;
  const int* base = cells + static_cast<int64_t>(scan) * N * 3;
  // ONE_CHUNK (N <= 8192): a lane adds at most 256 voxels per candidate, 256 * 255 < 65536, so the
  // packed 16-bit sums never overflow and no 32-bit accumulators are live in the loop (fewer
  // registers, 6 CTAs/SM). Otherwise the packed sums are flushed every 256 points per lane.
  int sum[ONE_CHUNK ? 1 : KX];
  if (!ONE_CHUNK) {
#pragma unroll
    for (int k = 0; k < KX; ++k) sum[k] = 0;
  }
  unsigned even[KXW], odd[KXW];
#pragma unroll
  for (int j = 0; j < KXW; ++j) even[j] = odd[j] = 0u;  // lanes beyond a cloud of fewer than 32 points add nothing
  for (int chunk = lane; chunk < N; chunk += 32 * 256) {
#pragma unroll
    for (int j = 0; j < KXW; ++j) even[j] = odd[j] = 0u;
    const int chunk_end = min(N, chunk + 32 * 256);
    for (int i = chunk; i < chunk_end; i += 32) {
      const int x = __ldg(base + 3 * i) + ox0, y = __ldg(base + 3 * i + 1) + oy, z = __ldg(base + 3 * i + 2) + oz;
      const int q0 = x >> rows.cs_log2;
      if (static_cast<unsigned>(y) >= static_cast<unsigned>(rows.dim[1]) ||
          static_cast<unsigned>(z) >= static_cast<unsigned>(rows.dim[2]) || q0 < -kLatticePadLeft ||
          q0 >= rows.payload) {
        continue;
      }
      const int b = kLatticePadLeft + q0;
      const unsigned* rowp = rows.data +
                             ((static_cast<size_t>(z) * rows.dim[1] + y) * (cs_mask + 1) + (x & cs_mask)) * rows.row_words +
                             (b >> 2);
      const unsigned shift = (b & 3) * 8;
      unsigned w[KXW + 1];
#pragma unroll
      for (int j = 0; j <= KXW; ++j) w[j] = __ldg(rowp + j);
#pragma unroll
      for (int j = 0; j < KXW; ++j) {
        const unsigned v = __funnelshift_r(w[j], w[j + 1], shift);
        even[j] += v & 0x00FF00FFu;
        odd[j] += (v >> 8) & 0x00FF00FFu;
      }
    }
    if (!ONE_CHUNK) {
#pragma unroll
      for (int j = 0; j < KXW; ++j) {
        sum[4 * j] += even[j] & 0xFFFFu;
        sum[4 * j + 1] += odd[j] & 0xFFFFu;
        sum[4 * j + 2] += even[j] >> 16;
        sum[4 * j + 3] += odd[j] >> 16;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < KX; ++k) {
    int v;
    if (ONE_CHUNK) {
      const unsigned packed = (k & 1) ? odd[k >> 2] : even[k >> 2];
      v = static_cast<int>((k & 2) ? (packed >> 16) : (packed & 0xFFFFu));
    } else {
      v = sum[k];
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    const int ix = g * KX + k;
    if (lane == 0 && ix < nxy) sums[row * nxy + ix] = v;
  }
}

// Children of a wave of parents at 'depth' (the loop nest at :441-461): warp w
// scores child slot w % 8 = (z, y, x) in {0, half}^3 of parent w / 8, so the
// host uploads one record per parent and reads eight sums back. A child beyond
// the search window (the reference's 'break') gets -1.
__device__ __forceinline__ int LoadDenseU8(const DenseLevel& d, const int x, const int y, const int z) {
  const unsigned ux = static_cast<unsigned>(x - d.min[0]), uy = static_cast<unsigned>(y - d.min[1]),
                 uz = static_cast<unsigned>(z - d.min[2]);
  if (ux >= static_cast<unsigned>(d.dim[0]) || uy >= static_cast<unsigned>(d.dim[1]) ||
      uz >= static_cast<unsigned>(d.dim[2])) {
    return 0;
  }
  return __ldg(d.data + (static_cast<size_t>(uz) * d.dim[1] + uy) * d.dim[0] + ux);
}

__global__ void __launch_bounds__(256) ScoreChildrenKernel(const BrickTable level, const DenseLevel dense,
                                                           const int* __restrict__ cells,
                                                           const int N, const int reduction_exponent,
                                                           const int4* __restrict__ parents,
                                                           const int64_t num_parents, const int half,
                                                           const int wxy, const int wz, int* __restrict__ sums) {
  const int64_t warp = (blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= num_parents * 8) return;
  const int4 p = parents[warp >> 3];
  const int slot = static_cast<int>(warp & 7);
  const int cx = p.y + ((slot & 1) ? half : 0), cy = p.z + ((slot & 2) ? half : 0), cz = p.w + ((slot & 4) ? half : 0);
  if (cx > wxy || cy > wxy || cz > wz) {
    if (lane == 0) sums[warp] = -1;
    return;
  }
  const int ox = cx >> reduction_exponent, oy = cy >> reduction_exponent, oz = cz >> reduction_exponent;
  const int* base = cells + static_cast<int64_t>(p.x) * N * 3;
  int sum = 0;
  if (dense.data != nullptr) {  // one byte load per gather instead of brick index + voxel
    for (int i = lane; i < N; i += 32) {
      const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
      sum += LoadDenseU8(dense, x + ox, y + oy, z + oz);
    }
  } else {
    for (int i = lane; i < N; i += 32) {
      const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
      sum += LoadU8(level, x + ox, y + oy, z + oz);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  if (lane == 0) sums[warp] = sum;
}

// Order-preserving compaction of the lattice sums that reach 'threshold':
// count per 1024-element chunk, scan the chunk counts, then write (index, sum)
// pairs — so only the candidates the reference's loop can reach cross PCIe.
constexpr int kChunk = 1024;
__global__ void __launch_bounds__(256) CountKeptKernel(const int* __restrict__ sums, const int64_t n,
                                                       const int threshold, int* __restrict__ chunk_counts) {
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kChunk;
  int c = 0;
  for (int k = threadIdx.x; k < kChunk; k += 256) {
    const int64_t i = base + k;
    c += (i < n && sums[i] >= threshold) ? 1 : 0;
  }
  __shared__ int s[8];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < 8; ++w) t += s[w];
    chunk_counts[blockIdx.x] = t;
  }
}
// Single-CTA exclusive scan of the chunk counts; chunk_counts[num_chunks] receives the total.
__global__ void __launch_bounds__(1024) ScanChunksKernel(int* __restrict__ chunk_counts, const int num_chunks) {
  __shared__ int partial[1024];
  const int per = (num_chunks + 1023) / 1024;
  const int begin = threadIdx.x * per, end = min(num_chunks, begin + per);
  int local = 0;
  for (int i = begin; i < end; ++i) local += chunk_counts[i];
  partial[threadIdx.x] = local;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const int v = threadIdx.x >= d ? partial[threadIdx.x - d] : 0;
    __syncthreads();
    partial[threadIdx.x] += v;
    __syncthreads();
  }
  int run = partial[threadIdx.x] - local;
  for (int i = begin; i < end; ++i) {
    const int c = chunk_counts[i];
    chunk_counts[i] = run;
    run += c;
  }
  if (threadIdx.x == 1023) chunk_counts[num_chunks] = partial[1023];
}
__global__ void __launch_bounds__(256) WriteKeptKernel(const int* __restrict__ sums, const int64_t n,
                                                       const int threshold, const int* __restrict__ chunk_offsets,
                                                       int2* __restrict__ kept) {
  // One warp per 32 consecutive elements inside the chunk, chunks of 1024 = 32 warps' worth handled
  // by 8 warps in 4 rounds; ballot + popc keep the order.
  __shared__ int warp_counts[32];
  const int64_t base = static_cast<int64_t>(blockIdx.x) * kChunk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int my_flag[4];
  int my_sum[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int seg = r * 8 + warp;  // segment of 32 elements, in order
    const int64_t i = base + seg * 32 + lane;
    const int v = i < n ? sums[i] : 0;
    my_sum[r] = v;
    my_flag[r] = (i < n && v >= threshold) ? 1 : 0;
    const unsigned b = __ballot_sync(0xffffffffu, my_flag[r]);
    if (lane == 0) warp_counts[seg] = __popc(b);
  }
  __syncthreads();
  const int out_base = chunk_offsets[blockIdx.x];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int seg = r * 8 + warp;
    int before = 0;
    for (int k = 0; k < seg; ++k) before += warp_counts[k];
    const unsigned b = __ballot_sync(0xffffffffu, my_flag[r]);
    if (my_flag[r]) {
      const int pos = out_base + before + __popc(b & ((1u << lane) - 1u));
      kept[pos] = make_int2(static_cast<int>(base + seg * 32 + lane), my_sum[r]);
    }
  }
}

// Histogram of the lattice sums >= tmin in kHistogramBins bins of width 1 << shift.
constexpr int kHistogramBins = 2048;
__global__ void __launch_bounds__(256) HistogramKernel(const int* __restrict__ sums, const int64_t n, const int tmin,
                                                       const int shift, int* __restrict__ bins) {
  __shared__ int local[kHistogramBins];
  for (int k = threadIdx.x; k < kHistogramBins; k += 256) local[k] = 0;
  __syncthreads();
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int v = sums[i];
    if (v >= tmin) atomicAdd(&local[min(kHistogramBins - 1, (v - tmin) >> shift)], 1);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < kHistogramBins; k += 256)
    if (local[k] != 0) atomicAdd(&bins[k], local[k]);
}

// LowResolutionMatcher closure; the float accumulation is sequential over the points in cloud
// order, exactly like the reference loop.
// One warp per pose: the lanes evaluate the points' values in parallel (the lookups are the
// latency), then lane 0 adds them in cloud order with the same float additions as the loop.
constexpr int kLowChunk = 1024;  // values staged per warp and round
__global__ void __launch_bounds__(128) LowResolutionKernel(const GridDev g, const float* __restrict__ points,
                                                           const int N, const float* __restrict__ poses,
                                                           const int64_t num_poses, float* __restrict__ scores) {
  __shared__ float staged[4][kLowChunk];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t p = blockIdx.x * static_cast<int64_t>(4) + warp;
  if (p >= num_poses) return;
  const float* P = poses + 7 * p;
  const float tx = P[0], ty = P[1], tz = P[2], qw = P[3], qx = P[4], qy = P[5], qz = P[6];
  float* mine = staged[warp];
  float score = 0.f;
  for (int base = 0; base < N; base += kLowChunk) {
    const int count = min(kLowChunk, N - base);
    for (int k = lane; k < count; k += 32) {
      const int i = base + k;
      float wx, wy, wz;
      RotateF(qw, qx, qy, qz, points[3 * i], points[3 * i + 1], points[3 * i + 2], &wx, &wy, &wz);
      wx = __fadd_rn(wx, tx);
      wy = __fadd_rn(wy, ty);
      wz = __fadd_rn(wz, tz);
      const int cx = CellIndexFast(wx, g.resolution, g.inv_resolution),
                cy = CellIndexFast(wy, g.resolution, g.inv_resolution),
                cz = CellIndexFast(wz, g.resolution, g.inv_resolution);
      float v;
      if (g.kind == kGridProb) {
        v = Decode(g.a, LoadU16(g.table, cx, cy, cz));
      } else {
        const float tsd = Decode(g.a, LoadTsdf(g.table, cx, cy, cz).x);
        v = __fsub_rn(1.f, fabsf(__fdiv_rn(tsd, g.max_tsd)));
      }
      mine[k] = v;
    }
    __syncwarp();
    if (lane == 0) {
      for (int k = 0; k < count; ++k) score = __fadd_rn(score, mine[k]);
    }
    __syncwarp();
  }
  if (lane == 0) scores[p] = __fdiv_rn(score, static_cast<float>(N));
}

// ------------------------------------------------------------- host parts --

// rotational_scan_matcher.cc:134-155, written into 'out' (size of the histogram). The indices
// (i + full_buckets) % size are stepped instead of divided; the float operations are the
// reference's: fraction * h1 + (1 - fraction) * h0, unfused (-ffp-contract=off).
void RotateHistogram(const std::vector<float>& histogram, const float angle, float* out) {
  const long size = static_cast<long>(histogram.size());
  if (size == 0) return;
  const float rotate_by_buckets = static_cast<float>((-angle * size) / M_PI);
  int full_buckets = static_cast<int>(std::lround(rotate_by_buckets - 0.5f));
  const float fraction = rotate_by_buckets - full_buckets;
  while (full_buckets < 0) full_buckets += static_cast<int>(size);
  long i0 = full_buckets % size, i1 = (full_buckets + 1) % size;
  const float one_minus = 1.f - fraction;
  for (long i = 0; i != size; ++i) {
    out[i] = fraction * histogram[i1] + one_minus * histogram[i0];
    if (++i0 == size) i0 = 0;
    if (++i1 == size) i1 = 0;
  }
}

// Eigen's sum of n floats over a VectorXf expression (Eigen 3.3 Core/Redux.h; Eigen is not
// vendored in the reference, restated from memory): flavor 1 = the SSE2 build of an x86-64
// reference — packets of 4 from the aligned start, two packet accumulators taking alternate
// packets, their sum, one more packet when their number is odd, predux (a0 + a2) + (a1 + a3),
// scalar tail; flavor 0 = the generic scalar code, left to right.
template <typename Term>
float EigenSum(const long n, const int flavor, const Term& term) {
  if (flavor != 1 || n < 4) {
    float res = term(0);
    for (long i = 1; i < n; ++i) res = res + term(i);
    return res;
  }
  const long aligned_end2 = (n / 8) * 8, aligned_end = (n / 4) * 4;
  float p0[4], p1[4];
  for (int k = 0; k < 4; ++k) p0[k] = term(k);
  if (aligned_end > 4) {
    for (int k = 0; k < 4; ++k) p1[k] = term(4 + k);
    for (long index = 8; index < aligned_end2; index += 8) {
      for (int k = 0; k < 4; ++k) p0[k] = p0[k] + term(index + k);
      for (int k = 0; k < 4; ++k) p1[k] = p1[k] + term(index + 4 + k);
    }
    for (int k = 0; k < 4; ++k) p0[k] = p0[k] + p1[k];
    if (aligned_end > aligned_end2) {
      for (int k = 0; k < 4; ++k) p0[k] = p0[k] + term(aligned_end2 + k);
    }
  }
  float res = (p0[0] + p0[2]) + (p0[1] + p0[3]);
  for (long i = aligned_end; i < n; ++i) res = res + term(i);
  return res;
}

// rotational_scan_matcher.cc:121-132 (VectorXf::norm and ::dot). 'submap_norm' is the submap
// histogram's norm, which does not depend on the yaw.
float MatchHistograms(const std::vector<float>& submap, const float submap_norm, const float* scan, const int flavor) {
  const long n = static_cast<long>(submap.size());
  if (n == 0) return 1.f;
  const float scan_norm = std::sqrt(EigenSum(n, flavor, [&](long i) { return scan[i] * scan[i]; }));
  const float normalization = scan_norm * submap_norm;
  if (normalization < 1e-3f) return 1.f;
  return EigenSum(n, flavor, [&](long i) { return submap[i] * scan[i]; }) / normalization;
}

struct SearchParameters {
  int wxy;
  int wz;
  double angular;
};

struct DeviceBuffer {
  void* p = nullptr;
  ~DeviceBuffer() {
    if (p) cudaFree(p);
  }
  int Alloc(size_t bytes) {
    if (p) cudaFree(p);
    p = nullptr;
    HG_CUDA(cudaMalloc(&p, std::max<size_t>(bytes, 16)));
    return HG_OK;
  }
  size_t cap = 0;
  // Grow-only (with headroom); contents are not preserved.
  int Reserve(size_t bytes) {
    if (bytes <= cap) return HG_OK;
    const size_t want = bytes + bytes / 2 + 4096;
    const int rc = Alloc(want);
    cap = rc == HG_OK ? want : 0;
    return rc;
  }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};
struct PinnedBuffer {
  void* p = nullptr;
  size_t cap = 0;
  ~PinnedBuffer() {
    if (p) cudaFreeHost(p);
  }
  int Reserve(size_t bytes) {
    if (bytes <= cap) return HG_OK;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    const size_t want = std::max<size_t>(bytes * 2, 1 << 16);
    HG_CUDA(cudaMallocHost(&p, want));
    cap = want;
    return HG_OK;
  }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

// Scratch memory of the searches running on this host thread, per device.
// Grow-only and reused from search to search: the allocations (pinned host
// memory in particular) cost more than the kernels they serve.
struct Workspace {
  DeviceBuffer d_points, d_poses, d_cells;                     // discrete scans
  DeviceBuffer d_cand, d_sums;                                 // waves
  DeviceBuffer d_lattice, d_chunks, d_kept;                    // lowest-resolution lattice
  DeviceBuffer d_low_points, d_low_poses, d_low_scores;        // leaf gate
  PinnedBuffer h_cand, h_sums, h_kept, h_poses, h_low, h_count;
  // device-resident search (DeviceSearch)
  DeviceBuffer d_upload, d_starts, d_lists, d_ctrl;
  DeviceBuffer d_exchange;  // dive bounds offered to / received from the other shards of a search
  PinnedBuffer h_upload, h_ctrl;
};
Workspace& GetWorkspace(int device) {
  thread_local std::map<int, std::unique_ptr<Workspace>> spaces;
  std::unique_ptr<Workspace>& w = spaces[device];
  if (!w) w.reset(new Workspace);
  return *w;
}

}  // namespace
}  // namespace hg

using namespace hg;

struct hg_fcsm {
  int device = 0;
  hg_fcsm_options options;
  const hg_grid* grid = nullptr;
  const hg_grid* low_grid = nullptr;
  hg_stack* stack = nullptr;
  std::vector<float> histogram;
  float resolution = 0.f;
  int width_in_voxels = 0;
  // Residue-row copy of the lowest-resolution level for the lattice kernel (see LatticeRows).
  LatticeRows lattice_rows;
  void* d_lattice_rows = nullptr;
  LevelDev* d_levels = nullptr;  // [branch_and_bound_depth]
};

namespace {

// The discrete scans of one search, resident on the device.
struct ScanSet {
  int S = 0;               // scans kept after the rotational prefilter
  int N = 0;               // points per scan
  int num_low_levels = 0;  // branch_and_bound_depth - full_resolution_depth (>= 0)
  int frd = 0;             // min(full_resolution_depth, branch_and_bound_depth)
  std::vector<Rigid> poses;
  std::vector<float> rotational_scores;
  std::vector<int> global_scan_index;  // index in the unsharded scan list
  Workspace* ws = nullptr;  // owns d_points / d_poses / d_cells
  // Cells of depth d live at level max(0, d - frd + 1).
  const int* CellsOfDepth(int depth) const {
    const int level = std::max(0, depth - frd + 1);
    return ws->d_cells.as<int>() + static_cast<int64_t>(level) * S * N * 3;
  }
};

// Host half of GenerateDiscreteScans (:276-326): yaw list, rotational prefilter, per-yaw poses of
// the scans this shard keeps. Fills out->{poses, rotational_scores, global_scan_index, S, N, frd,
// num_low_levels}; no device work.
int ComputeScanPoses(const hg_fcsm* m, const SearchParameters& sp, const hg_node_data* data,
                     const Rigid& node_pose, const Rigid& submap_pose, int shard_index, int num_shards,
                     ScanSet* out) {
  const int flavor = m->options.float_flavor;
  const int N = data->num_high_resolution_points;
  HG_REQUIRE(N > 0 && data->high_resolution_point_cloud != nullptr, "empty high resolution point cloud");
  const float* pts = data->high_resolution_point_cloud;
  const float resolution = m->resolution;
  // :281-294
  float max_scan_range = 3.f * resolution;
  for (int i = 0; i < N; ++i) {
    const float range = fgeom::Norm3(V3{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]});
    max_scan_range = std::max(range, max_scan_range);
  }
  const float kSafetyMargin = 1.f - 1e-2f;
  const float angular_step_size =
      kSafetyMargin * std::acos(1.f - (resolution * resolution) / (2.f * (max_scan_range * max_scan_range)));
  const int angular_window_size = static_cast<int>(std::lround(sp.angular / angular_step_size));
  std::vector<float> angles;
  for (int rz = -angular_window_size; rz <= angular_window_size; ++rz) angles.push_back(rz * angular_step_size);
  const Rigid node_to_submap = fgeom::RigidProduct(fgeom::RigidInverse(submap_pose), node_pose, flavor);
  std::vector<float> scores;
  if (m->options.use_rotational_scan_matcher) {
    HG_REQUIRE(data->histogram_size == static_cast<int>(m->histogram.size()),
               "rotational scan matcher histograms differ in size");
    const double* ga = data->gravity_alignment;
    const double n2 = (ga[1] * ga[1] + ga[2] * ga[2]) + (ga[3] * ga[3] + ga[0] * ga[0]);
    const Q4 gravity_inverse{static_cast<float>(ga[0] / n2), static_cast<float>(-ga[1] / n2),
                             static_cast<float>(-ga[2] / n2), static_cast<float>(-ga[3] / n2)};
    const float initial_angle = fgeom::GetYaw(fgeom::Product(node_to_submap.q, gravity_inverse, flavor));
    const std::vector<float> scan_hist(data->rotational_scan_matcher_histogram,
                                       data->rotational_scan_matcher_histogram + data->histogram_size);
    scores.reserve(angles.size());
    const long hn = static_cast<long>(m->histogram.size());
    const float submap_norm =
        hn == 0 ? 0.f : std::sqrt(EigenSum(hn, flavor, [&](long i) { return m->histogram[i] * m->histogram[i]; }));
    std::vector<float> rotated(scan_hist.size());
    for (const float angle : angles) {
      RotateHistogram(scan_hist, initial_angle + angle, rotated.data());
      scores.push_back(MatchHistograms(m->histogram, submap_norm, rotated.data(), flavor));
    }
  }
  out->poses.clear();
  out->rotational_scores.clear();
  out->global_scan_index.clear();
  int kept = 0;
  for (size_t i = 0; i != angles.size(); ++i) {
    if (m->options.use_rotational_scan_matcher &&
        static_cast<double>(scores[i]) < m->options.min_rotational_score) {
      continue;
    }
    const int global_index = kept++;
    if (num_shards > 1 && (global_index % num_shards) != shard_index) continue;
    const V3 angle_axis{0.f, 0.f, angles[i]};
    const Q4 q = fgeom::Product(
        fgeom::Product(fgeom::Inverse(submap_pose.q, flavor),
                       fgeom::AngleAxisVectorToRotationQuaternion(angle_axis), flavor),
        node_pose.q, flavor);
    out->poses.push_back(Rigid{node_to_submap.t, q});
    out->rotational_scores.push_back(m->options.use_rotational_scan_matcher ? scores[i] : 1.0f);
    out->global_scan_index.push_back(global_index);
  }
  out->S = static_cast<int>(out->poses.size());
  out->N = N;
  const int D = m->options.branch_and_bound_depth;
  out->frd = std::min(m->options.full_resolution_depth, D);
  out->num_low_levels = D - out->frd;
  return HG_OK;
}

int GenerateDiscreteScans(const hg_fcsm* m, const SearchParameters& sp, const hg_node_data* data,
                          const Rigid& node_pose, const Rigid& submap_pose, int shard_index, int num_shards,
                          cudaStream_t stream, ScanSet* out) {
  int rc = ComputeScanPoses(m, sp, data, node_pose, submap_pose, shard_index, num_shards, out);
  if (rc != HG_OK) return rc;
  const int S = out->S, N = out->N;
  const float* pts = data->high_resolution_point_cloud;
  const float resolution = m->resolution;
  if (S == 0) return HG_OK;

  Workspace& ws = GetWorkspace(m->device);
  out->ws = &ws;
  rc = ws.d_points.Reserve(sizeof(float) * 3 * N);
  if (rc != HG_OK) return rc;
  rc = ws.d_poses.Reserve(sizeof(float) * 7 * S);
  if (rc != HG_OK) return rc;
  const size_t cell_ints = static_cast<size_t>(1 + out->num_low_levels) * S * N * 3;
  rc = ws.d_cells.Reserve(sizeof(int) * cell_ints);
  if (rc != HG_OK) return rc;
  std::vector<float> flat(7 * static_cast<size_t>(S));
  for (int s = 0; s < S; ++s) {
    const Rigid& r = out->poses[s];
    float* o = &flat[7 * s];
    o[0] = r.t.x; o[1] = r.t.y; o[2] = r.t.z; o[3] = r.q.w; o[4] = r.q.x; o[5] = r.q.y; o[6] = r.q.z;
  }
  HG_CUDA(cudaMemcpyAsync(ws.d_points.p, pts, sizeof(float) * 3 * N, cudaMemcpyHostToDevice, stream));
  HG_CUDA(cudaMemcpyAsync(ws.d_poses.p, flat.data(), sizeof(float) * flat.size(), cudaMemcpyHostToDevice, stream));
  const int64_t total = static_cast<int64_t>(S) * N;
  DiscretizeKernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream>>>(
      ws.d_points.as<float>(), N, ws.d_poses.as<float>(), S, resolution, 1.f / resolution, out->num_low_levels, -sp.wxy,
      -sp.wz, ws.d_cells.as<int>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaStreamSynchronize(stream));  // 'flat' goes out of scope
  return HG_OK;
}

inline int ReduceExponentOf(const hg_fcsm* m, int depth) {
  return std::max(0, depth - m->options.full_resolution_depth + 1);
}

// Launches the scoring of one search's whole lowest-resolution lattice (S scans x nz x nxy x nxy
// candidates in generation order) into 'out'. Picks the residue-row kernel, the dense-level
// kernel or the brick-table kernel, whichever the level offers. *packed_out tells which.
int LaunchLattice(const hg_fcsm* m, const int* cells, int S, int N, int depth, int nxy, int nz, int wxy, int wz,
                  int* out, cudaStream_t stream, bool* packed_out) {
  const size_t n = static_cast<size_t>(S) * nxy * nxy * nz;
  if (n == 0) return HG_OK;
  const int64_t threads = static_cast<int64_t>(n) * 32;
  const hg_stack_level& L = m->stack->levels[depth];
  const LatticeRows& R = m->lattice_rows;
  const int e = ReduceExponentOf(m, depth);
  const bool packed = R.data != nullptr && depth == static_cast<int>(m->stack->levels.size()) - 1 && nxy >= 5 &&
                      ((1 << depth) >> e) == (1 << R.cs_log2);
  if (packed_out) *packed_out = packed;
  if (packed) {
#define HG_PACKED(KXW)                                                                                    \
  do {                                                                                                    \
    const int64_t groups = static_cast<int64_t>(S) * nz * nxy * ((nxy + 4 * KXW - 1) / (4 * KXW));        \
    if (N <= 8192) {                                                                                      \
      ScoreLatticePackedKernel<KXW, true><<<static_cast<unsigned>((groups * 32 + 255) / 256), 256, 0, stream>>>( \
          R, cells, N, e, nxy, nz, wxy, wz, 1 << depth, groups, out);                                     \
    } else {                                                                                              \
      ScoreLatticePackedKernel<KXW, false><<<static_cast<unsigned>((groups * 32 + 255) / 256), 256, 0, stream>>>( \
          R, cells, N, e, nxy, nz, wxy, wz, 1 << depth, groups, out);                                     \
    }                                                                                                     \
  } while (0)
    if (nxy <= 8) HG_PACKED(2); else if (nxy <= 20) HG_PACKED(5); else HG_PACKED(8);
#undef HG_PACKED
  } else if (L.dense.data != nullptr) {
    constexpr int KX = 4;
    const int64_t groups = static_cast<int64_t>(S) * nz * nxy * ((nxy + KX - 1) / KX);
    ScoreLatticeDenseKernel<KX><<<static_cast<unsigned>((groups * 32 + 255) / 256), 256, 0, stream>>>(
        L.dense, cells, N, e, nxy, nz, wxy, wz, 1 << depth, groups, out);
  } else {
    ScoreLatticeKernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
        L.table, cells, N, e, nxy, nz, wxy, wz, 1 << depth, static_cast<int64_t>(n), out);
  }
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  return HG_OK;
}

// Scores candidates on the device; synchronous. Buffers live in the thread's workspace.
struct Scorer {
  const hg_fcsm* m;
  const ScanSet* scans;
  cudaStream_t stream;
  Workspace* ws;
  int64_t total_scored = 0;
  int launches = 0;
  float lattice_ms = 0.f;
  int64_t lattice_candidates = 0;

  int ReduceExponent(int depth) const { return std::max(0, depth - m->options.full_resolution_depth + 1); }
  static constexpr size_t kZeroCopyParents = 512;

  int Reserve(size_t n) {
    int rc = ws->h_cand.Reserve(n * sizeof(int4));
    if (rc != HG_OK) return rc;
    rc = ws->h_sums.Reserve(n * 8 * sizeof(int));
    if (rc != HG_OK) return rc;
    rc = ws->d_cand.Reserve(n * sizeof(int4));
    if (rc != HG_OK) return rc;
    return ws->d_sums.Reserve(n * 8 * sizeof(int));
  }
  int4* HostCandidates() const { return ws->h_cand.as<int4>(); }
  const int* HostSums() const { return ws->h_sums.as<int>(); }

  // Explicit candidates (already in HostCandidates(), n records) -> HostSums()[n].
  int Run(int depth, size_t n, float* kernel_ms) {
    if (n == 0) return HG_OK;
    HG_CUDA(cudaMemcpyAsync(ws->d_cand.p, ws->h_cand.p, n * sizeof(int4), cudaMemcpyHostToDevice, stream));
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (kernel_ms) {
      HG_CUDA(cudaEventCreate(&e0));
      HG_CUDA(cudaEventCreate(&e1));
      HG_CUDA(cudaEventRecord(e0, stream));
    }
    const int64_t threads = static_cast<int64_t>(n) * 32;
    ScoreKernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
        m->stack->levels[depth].table, scans->CellsOfDepth(depth), scans->N, ReduceExponent(depth),
        ws->d_cand.as<int4>(), static_cast<int64_t>(n), ws->d_sums.as<int>());
    CountLaunch();
    HG_CUDA(cudaGetLastError());
    if (kernel_ms) HG_CUDA(cudaEventRecord(e1, stream));
    HG_CUDA(cudaMemcpyAsync(ws->h_sums.p, ws->d_sums.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    if (kernel_ms) {
      float ms = 0.f;
      HG_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      *kernel_ms += ms;
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
    }
    total_scored += static_cast<int64_t>(n);
    ++launches;
    return HG_OK;
  }

  // Children of n parents (records in HostCandidates()) scored at 'child_depth':
  // HostSums()[8 n], -1 where the child lies beyond the window.
  int RunChildren(int child_depth, size_t n, int half, int wxy, int wz) {
    if (n == 0) return HG_OK;
    const int64_t threads = static_cast<int64_t>(n) * 8 * 32;
    if (n <= kZeroCopyParents) {
      // A small wave (a local search has dozens of them): the kernel reads the parent records
      // from, and writes the sums to, the pinned host buffers directly (they are mapped into the
      // device's address space), so the wave is one launch and one synchronisation instead of
      // copy + launch + copy + synchronisation.
      ScoreChildrenKernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
          m->stack->levels[child_depth].table, m->stack->levels[child_depth].dense,
          scans->CellsOfDepth(child_depth), scans->N, ReduceExponent(child_depth), ws->h_cand.as<int4>(),
          static_cast<int64_t>(n), half, wxy, wz, ws->h_sums.as<int>());
      CountLaunch();
      HG_CUDA(cudaGetLastError());
      HG_CUDA(cudaStreamSynchronize(stream));
      ++launches;
      return HG_OK;
    }
    HG_CUDA(cudaMemcpyAsync(ws->d_cand.p, ws->h_cand.p, n * sizeof(int4), cudaMemcpyHostToDevice, stream));
    ScoreChildrenKernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
        m->stack->levels[child_depth].table, m->stack->levels[child_depth].dense, scans->CellsOfDepth(child_depth),
        scans->N, ReduceExponent(child_depth), ws->d_cand.as<int4>(), static_cast<int64_t>(n), half, wxy, wz,
        ws->d_sums.as<int>());
    CountLaunch();
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws->h_sums.p, ws->d_sums.p, n * 8 * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    ++launches;
    return HG_OK;
  }

  // Scores the whole lowest-resolution lattice into ws->d_lattice (generation order).
  int RunLattice(int depth, int nxy, int nz, int wxy, int wz, size_t n) {
    if (n == 0) return HG_OK;
    int rc = ws->d_lattice.Reserve(n * sizeof(int));
    if (rc != HG_OK) return rc;
    const int64_t threads = static_cast<int64_t>(n) * 32;
    // The kernel time is reported for full-submap searches (bench.py's roofline); a local search
    // is a few microseconds of kernel and skips the event calls.
    const bool timed = n >= (size_t(1) << 20);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (timed) {
      HG_CUDA(cudaEventCreate(&e0));
      HG_CUDA(cudaEventCreate(&e1));
      HG_CUDA(cudaEventRecord(e0, stream));
    }
    const hg_stack_level& L = m->stack->levels[depth];
    bool packed = false;
    rc = LaunchLattice(m, scans->CellsOfDepth(depth), scans->S, scans->N, depth, nxy, nz, wxy, wz,
                       ws->d_lattice.as<int>(), stream, &packed);
    if (rc != HG_OK) return rc;
    if (timed) {
      HG_CUDA(cudaEventRecord(e1, stream));
      HG_CUDA(cudaEventSynchronize(e1));
      HG_CUDA(cudaEventElapsedTime(&lattice_ms, e0, e1));
      cudaEventDestroy(e0);
      cudaEventDestroy(e1);
    }
    lattice_candidates = static_cast<int64_t>(n);
    total_scored += static_cast<int64_t>(n);
    ++launches;
    if (packed && std::getenv("HG_LATTICE_CHECK") != nullptr) {
      // Development / test aid: the packed kernel against the brick-table kernel, every sum.
      std::vector<int> got(n), want(n);
      HG_CUDA(cudaMemcpyAsync(got.data(), ws->d_lattice.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
      DeviceBuffer check;
      rc = check.Reserve(n * sizeof(int));
      if (rc != HG_OK) return rc;
      ScoreLatticeKernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, stream>>>(
          L.table, scans->CellsOfDepth(depth), scans->N, ReduceExponent(depth), nxy, nz, wxy, wz, 1 << depth,
          static_cast<int64_t>(n), check.as<int>());
      HG_CUDA(cudaMemcpyAsync(want.data(), check.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
      HG_CUDA(cudaStreamSynchronize(stream));
      for (size_t i = 0; i < n; ++i) {
        if (got[i] != want[i]) {
          SetError("HG_LATTICE_CHECK: packed lattice sum " + std::to_string(i) + " is " + std::to_string(got[i]) +
                   ", the brick-table kernel gives " + std::to_string(want[i]));
          return HG_ERR_CUDA;
        }
      }
    }
    return HG_OK;
  }
  // All lattice sums to the host (strict order needs every one).
  int DownloadLattice(size_t n, std::vector<int>* out) {
    out->resize(n);
    if (n == 0) return HG_OK;
    HG_CUDA(cudaMemcpyAsync(out->data(), ws->d_lattice.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    return HG_OK;
  }
  // Number of lattice candidates with sum >= threshold.
  int CountKept(size_t n, int threshold, int* count) {
    *count = 0;
    if (n == 0) return HG_OK;
    const int chunks = static_cast<int>((n + kChunk - 1) / kChunk);
    int rc = ws->d_chunks.Reserve((chunks + 1) * sizeof(int));
    if (rc != HG_OK) return rc;
    rc = ws->h_count.Reserve(sizeof(int));
    if (rc != HG_OK) return rc;
    CountKeptKernel<<<chunks, 256, 0, stream>>>(ws->d_lattice.as<int>(), static_cast<int64_t>(n), threshold,
                                                ws->d_chunks.as<int>());
    ScanChunksKernel<<<1, 1024, 0, stream>>>(ws->d_chunks.as<int>(), chunks);
    CountLaunch(2);
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws->h_count.p, ws->d_chunks.as<int>() + chunks, sizeof(int), cudaMemcpyDeviceToHost,
                            stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    *count = *ws->h_count.as<int>();
    return HG_OK;
  }
  // A threshold T >= tmin such that at least 'want' lattice candidates have sum >= T (as large as
  // a 2048-bin histogram of the sums can tell), or tmin if fewer than 'want' reach tmin.
  int TopThreshold(size_t n, int tmin, int want, int* threshold) {
    *threshold = tmin;
    const int64_t range = static_cast<int64_t>(255) * scans->N - tmin + 1;
    if (range <= 0 || n == 0) return HG_OK;
    int shift = 0;
    while ((range >> shift) > kHistogramBins) ++shift;
    int rc = ws->d_sums.Reserve(kHistogramBins * sizeof(int));
    if (rc != HG_OK) return rc;
    rc = ws->h_sums.Reserve(kHistogramBins * sizeof(int));
    if (rc != HG_OK) return rc;
    HG_CUDA(cudaMemsetAsync(ws->d_sums.p, 0, kHistogramBins * sizeof(int), stream));
    HistogramKernel<<<592, 256, 0, stream>>>(ws->d_lattice.as<int>(), static_cast<int64_t>(n), tmin, shift,
                                             ws->d_sums.as<int>());
    CountLaunch();
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws->h_sums.p, ws->d_sums.p, kHistogramBins * sizeof(int), cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    const int* bins = ws->h_sums.as<int>();
    int64_t above = 0;
    for (int b = kHistogramBins - 1; b >= 0; --b) {
      above += bins[b];
      if (above >= want) {
        *threshold = tmin + (b << shift);
        return HG_OK;
      }
    }
    return HG_OK;
  }
  // (generation index, sum) of the lattice candidates with sum >= threshold, in generation order.
  int DownloadKept(size_t n, int threshold, std::vector<int2>* out) {
    out->clear();
    if (n == 0) return HG_OK;
    const int chunks = static_cast<int>((n + kChunk - 1) / kChunk);
    int rc = ws->d_chunks.Reserve((chunks + 1) * sizeof(int));
    if (rc != HG_OK) return rc;
    rc = ws->h_count.Reserve(sizeof(int));
    if (rc != HG_OK) return rc;
    CountKeptKernel<<<chunks, 256, 0, stream>>>(ws->d_lattice.as<int>(), static_cast<int64_t>(n), threshold,
                                                ws->d_chunks.as<int>());
    ScanChunksKernel<<<1, 1024, 0, stream>>>(ws->d_chunks.as<int>(), chunks);
    CountLaunch(2);
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws->h_count.p, ws->d_chunks.as<int>() + chunks, sizeof(int), cudaMemcpyDeviceToHost,
                            stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    const int kept = *ws->h_count.as<int>();
    if (kept == 0) return HG_OK;
    rc = ws->d_kept.Reserve(static_cast<size_t>(kept) * sizeof(int2));
    if (rc != HG_OK) return rc;
    rc = ws->h_kept.Reserve(static_cast<size_t>(kept) * sizeof(int2));
    if (rc != HG_OK) return rc;
    WriteKeptKernel<<<chunks, 256, 0, stream>>>(ws->d_lattice.as<int>(), static_cast<int64_t>(n), threshold,
                                                ws->d_chunks.as<int>(), ws->d_kept.as<int2>());
    CountLaunch();
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws->h_kept.p, ws->d_kept.p, static_cast<size_t>(kept) * sizeof(int2),
                            cudaMemcpyDeviceToHost, stream));
    HG_CUDA(cudaStreamSynchronize(stream));
    out->assign(ws->h_kept.as<int2>(), ws->h_kept.as<int2>() + kept);
    return HG_OK;
  }
};

// ---- branch and bound --------------------------------------------------------

struct Node {
  int scan;
  int ox, oy, oz;
  float score;
  float low_resolution_score;  // leaves only; NaN = not evaluated yet
  int first_child;             // -1 = not expanded
  int num_children;
};

constexpr int kChildrenPruned = -2;  // expanded; no child beat the bound, so none was materialised

struct BestCandidate {
  int node = -1;
  float score = -std::numeric_limits<float>::infinity();
  float low_resolution_score = 0.f;
};

class BranchAndBoundSearch {
 public:
  BranchAndBoundSearch(const hg_fcsm* m, const SearchParameters& sp, const hg_node_data* data,
                       const ScanSet* scans, cudaStream_t stream)
      : m_(m), sp_(sp), data_(data), scans_(scans), stream_(stream) {
    ws_ = &GetWorkspace(m->device);
    scorer_.m = m;
    scorer_.scans = scans;
    scorer_.stream = stream;
    scorer_.ws = ws_;
  }

  // strict = reproduce the reference's visiting order candidate by candidate.
  // Otherwise the search runs on the reachable candidates only and returns the
  // same candidate: the answer is the best valid leaf, ties going to the leaf
  // the reference visits first; the only tie the fast order cannot settle by
  // itself is one between top-level subtrees of equal score, which is detected
  // afterwards (a probe of the winner's equal-score twins) and, if real,
  // settled by rerunning in strict order.
  int Run(float min_score, bool strict, BestCandidate* best_out) {
    const int max_depth = m_->options.branch_and_bound_depth - 1;
    strict_ = strict;
    global_bound_ = -std::numeric_limits<float>::infinity();
    int rc = UploadLowResolutionCloud();
    if (rc != HG_OK) return rc;
    std::vector<Key> keys;
    allow_partial_keys_ = !strict && max_depth >= 1;
    rc = CollectKeys(max_depth, min_score, strict, nullptr, &keys);
    allow_partial_keys_ = false;
    if (rc != HG_OK) return rc;
    float dive_bound = -std::numeric_limits<float>::infinity();
    if (!strict && max_depth >= 1 && keys.size() >= 16) {
      size_t dive_width = kDiveWidth;
      if (const char* env = std::getenv("HG_DIVE_WIDTH")) dive_width = std::max(16, std::atoi(env));  // experiments
      const size_t width = std::min(dive_width, keys.size());
      if (keys.size() > width) {
        std::nth_element(keys.begin(), keys.begin() + width, keys.end(), [](const Key& a, const Key& b) {
          return a.score > b.score || (a.score == b.score && a.index < b.index);
        });
      }
      rc = GreedyDive(keys, width, max_depth, min_score, &dive_bound);
      if (rc != HG_OK) return rc;
      if (dive_bound > min_score) {
        global_bound_ = dive_bound;
        wave_budget_ = 4096;  // the bound keeps the waves small: expand everything that survives it at once
        if (partial_keys_) {
          rc = CollectKeysReaching(dive_bound, &keys);  // only the top candidates are here so far
          if (rc != HG_OK) return rc;
        } else {
          size_t kept = 0;
          for (const Key& k : keys)
            if (k.score >= dive_bound) keys[kept++] = k;
          keys.resize(kept);
        }
      } else if (partial_keys_) {
        rc = CollectKeys(max_depth, min_score, false, nullptr, &keys);  // no bound: every candidate > min_score
        if (rc != HG_OK) return rc;
      }
    }
    BuildTopLevel(&keys, strict);
    failed_ = HG_OK;
    BestCandidate best = Search(0, static_cast<int>(top_count_), max_depth, min_score);
    if (failed_ != HG_OK) return failed_;
    if (!strict && global_bound_ > min_score && !(best.node >= 0 && best.score >= global_bound_)) {
      // The dive's leaf was not reached again: a parent scored below its child
      // somewhere (the pyramid is not a strict upper bound there). Redo without
      // the global bound, which is correct by construction.
      if (TraceEnabled()) std::fprintf(stderr, "[hg trace] global bound not confirmed: rerun without it\n");
      global_bound_ = -std::numeric_limits<float>::infinity();
      wave_budget_ = 8;
      rc = CollectKeys(max_depth, min_score, false, nullptr, &keys);
      if (rc != HG_OK) return rc;
      BuildTopLevel(&keys, false);
      best = Search(0, static_cast<int>(top_count_), max_depth, min_score);
      if (failed_ != HG_OK) return failed_;
    }
    if (!strict && best.node >= 0 && best.score > min_score) {
      // Top-level ancestor of the winner and its equal-score twins.
      const int ancestor = TopAncestor(best.node);
      std::vector<int> twins;
      for (size_t i = 0; i < top_count_; ++i)
        if (static_cast<int>(i) != ancestor && nodes_[i].score == nodes_[ancestor].score) twins.push_back(top_index_[i]);
      if (!twins.empty()) {
        // Does a twin subtree hold a valid leaf with the winning score?
        const Node winner = nodes_[best.node];
        const float winner_low = best.low_resolution_score;
        const float below = std::nextafter(best.score, -std::numeric_limits<float>::infinity());
        strict_ = true;
        global_bound_ = -std::numeric_limits<float>::infinity();
        wave_budget_ = 8;
        keys.clear();
        for (int t : twins) keys.push_back(Key{nodes_[ancestor].score, static_cast<uint32_t>(t)});
        BuildTopLevel(&keys, true);
        const BestCandidate probe = Search(0, static_cast<int>(top_count_), max_depth, below);
        if (failed_ != HG_OK) return failed_;
        if (probe.node >= 0 && probe.score >= best.score) {
          // A real tie across equal-score subtrees: only the strict order settles it.
          if (TraceEnabled()) std::fprintf(stderr, "[hg trace] tie across twin subtrees: strict rerun\n");
          wave_budget_ = 8;
          rc = CollectKeys(max_depth, min_score, true, nullptr, &keys);
          if (rc != HG_OK) return rc;
          BuildTopLevel(&keys, true);
          best = Search(0, static_cast<int>(top_count_), max_depth, min_score);
          if (failed_ != HG_OK) return failed_;
        } else {
          // Keep the winner; re-home it in the (rebuilt) node table.
          nodes_.clear();
          nodes_.push_back(winner);
          top_count_ = 0;
          best.node = 0;
          best.low_resolution_score = winner_low;
        }
      }
    }
    *best_out = best;
    return HG_OK;
  }

  const Node& node(int i) const { return nodes_[i]; }
  int64_t scored() const { return scorer_.total_scored; }
  int64_t visited() const { return visited_; }
  int launches() const { return scorer_.launches + low_launches_; }
  float lattice_ms() const { return scorer_.lattice_ms; }
  int64_t lattice_candidates() const { return scorer_.lattice_candidates; }

  // GetPoseFromCandidate (:400-406): Translation(resolution * offset) * scan pose.
  Rigid PoseOf(const Node& n) const {
    const Rigid translation{V3{m_->resolution * static_cast<float>(n.ox), m_->resolution * static_cast<float>(n.oy),
                               m_->resolution * static_cast<float>(n.oz)},
                            Q4{1.f, 0.f, 0.f, 0.f}};
    return fgeom::RigidProduct(translation, scans_->poses[n.scan], m_->options.float_flavor);
  }

 private:
  struct Key {
    float score;
    uint32_t index;  // position in generation order
  };

  // GenerateLowestResolutionCandidates + ScoreCandidates(max_depth) (:328-398),
  // without the sort: (score, generation index) of the lattice candidates in
  // generation order. strict: all of them. Otherwise only those the reference's
  // loop can reach (score > min_score, :437-440), compacted on the device.
  // 'only' restricts the list to the given generation indices.
  static constexpr size_t kSmallLattice = 1 << 16;
  // Starts of the greedy dive. On C4 256 ... 4096 starts all find the same bound (the optimum);
  // the dive costs 0.8 / 0.9 / 1.2 / 2.1 ms at 256 / 1024 / 2048 / 4096. HG_DIVE_WIDTH overrides.
  static constexpr size_t kDiveWidth = 1024;
  bool allow_partial_keys_ = false;  // CollectKeys may return only the best candidates ...
  bool partial_keys_ = false;        // ... and did
  int CollectKeys(int max_depth, float min_score, bool strict, const std::vector<int>* only,
                  std::vector<Key>* keys) {
    PhaseTimer pt;
    const int step = 1 << max_depth;
    nxy_ = (2 * sp_.wxy + step) / step;
    nz_ = (2 * sp_.wz + step) / step;
    per_scan_ = nxy_ * nxy_ * nz_;
    step_ = step;
    const size_t n = static_cast<size_t>(scans_->S) * per_scan_;
    if (!lattice_scored_) {
      int rc = scorer_.RunLattice(max_depth, nxy_, nz_, sp_.wxy, sp_.wz, n);
      if (rc != HG_OK) return rc;
      lattice_scored_ = true;
      visited_ += static_cast<int64_t>(n);
    }
    keys->clear();
    if (only != nullptr || strict) {
      // Every sum is needed on the host.
      if (h_lattice_sums_.size() != n) {
        const int rc = scorer_.DownloadLattice(n, &h_lattice_sums_);
        if (rc != HG_OK) return rc;
      }
      const int* sums = h_lattice_sums_.data();
      if (only != nullptr) {
        for (int i : *only) keys->push_back(Key{ScoreFromSum(sums[i], scans_->N), static_cast<uint32_t>(i)});
      } else {
        keys->reserve(n);
        for (size_t i = 0; i < n; ++i) keys->push_back(Key{ScoreFromSum(sums[i], scans_->N), static_cast<uint32_t>(i)});
      }
    } else if (n <= kSmallLattice) {
      // A local search: a few thousand sums. One copy and a host filter cost less than the three
      // compaction kernels and their two synchronisations.
      if (h_lattice_sums_.size() != n) {
        const int rc = scorer_.DownloadLattice(n, &h_lattice_sums_);
        if (rc != HG_OK) return rc;
      }
      const int threshold = SmallestSumBeating(min_score);
      for (size_t i = 0; i < n; ++i) {
        if (h_lattice_sums_[i] >= threshold)
          keys->push_back(Key{ScoreFromSum(h_lattice_sums_[i], scans_->N), static_cast<uint32_t>(i)});
      }
    } else {
      // A full-submap search keeps hundreds of thousands of candidates above min_score, of which the
      // greedy dive needs the best few thousand and the search afterwards only those that reach
      // the dive's bound. When allowed (first call of Run), fetch only the top ones now; Run
      // fetches the rest by threshold once the bound is known.
      int threshold = SmallestSumBeating(min_score);
      partial_keys_ = false;
      if (allow_partial_keys_) {
        int total = 0;
        int rc = scorer_.CountKept(n, threshold, &total);
        if (rc != HG_OK) return rc;
        if (static_cast<size_t>(total) > 8 * kDiveWidth) {
          rc = scorer_.TopThreshold(n, threshold, static_cast<int>(kDiveWidth), &threshold);
          if (rc != HG_OK) return rc;
          partial_keys_ = true;
        }
      }
      std::vector<int2> kept;
      const int rc = scorer_.DownloadKept(n, threshold, &kept);
      if (rc != HG_OK) return rc;
      keys->reserve(kept.size());
      for (const int2& k : kept) keys->push_back(Key{ScoreFromSum(k.y, scans_->N), static_cast<uint32_t>(k.x)});
    }
    if (TraceEnabled())
      std::fprintf(stderr, "[hg trace] lowest level: %zu candidates (%zu kept), score+select %.1f ms\n", n,
                   keys->size(), pt.Lap());
    return HG_OK;
  }

  // All lattice candidates with score >= bound (sum >= the smallest sum that reaches it).
  int CollectKeysReaching(float bound, std::vector<Key>* keys) {
    const size_t n = static_cast<size_t>(scans_->S) * per_scan_;
    int64_t lo = 0, hi = static_cast<int64_t>(255) * scans_->N + 1;
    while (lo < hi) {
      const int64_t mid = (lo + hi) / 2;
      if (ScoreFromSum(static_cast<int>(mid), scans_->N) >= bound) hi = mid; else lo = mid + 1;
    }
    std::vector<int2> kept;
    const int rc = scorer_.DownloadKept(n, static_cast<int>(lo), &kept);
    if (rc != HG_OK) return rc;
    keys->clear();
    keys->reserve(kept.size());
    for (const int2& k : kept) keys->push_back(Key{ScoreFromSum(k.y, scans_->N), static_cast<uint32_t>(k.x)});
    return HG_OK;
  }

  // score(sum) is non-decreasing in sum: the smallest sum whose score is > s.
  int SmallestSumBeating(float s) const {
    int64_t lo = 0, hi = static_cast<int64_t>(255) * scans_->N + 1;
    while (lo < hi) {
      const int64_t mid = (lo + hi) / 2;
      if (ScoreFromSum(static_cast<int>(mid), scans_->N) > s) hi = mid; else lo = mid + 1;
    }
    return static_cast<int>(lo);
  }

  Node NodeOfKey(const Key& k) const {
    const int scan = static_cast<int>(k.index / per_scan_);
    int rem = static_cast<int>(k.index - static_cast<size_t>(scan) * per_scan_);
    const int iz = rem / (nxy_ * nxy_);
    rem -= iz * nxy_ * nxy_;
    const int iy = rem / nxy_;
    const int ix = rem - iy * nxy_;
    return Node{scan, -sp_.wxy + ix * step_, -sp_.wxy + iy * step_, -sp_.wz + iz * step_, k.score,
                std::numeric_limits<float>::quiet_NaN(), -1, 0};
  }

  // Sorts the keys and makes them the top-level node list. strict: std::sort with
  // std::greater<Candidate3D>'s comparison over the generation order — the
  // comparator only looks at the score, so the same algorithm on (score, index)
  // keys yields the reference's permutation, ties included. Otherwise a total
  // order (score, then generation index).
  void BuildTopLevel(std::vector<Key>* keys, bool strict) {
    PhaseTimer pt;
    if (strict) {
      std::sort(keys->begin(), keys->end(), [](const Key& a, const Key& b) { return a.score > b.score; });
    } else {
      std::sort(keys->begin(), keys->end(), [](const Key& a, const Key& b) {
        return a.score > b.score || (a.score == b.score && a.index < b.index);
      });
    }
    nodes_.clear();
    nodes_.reserve(keys->size() * 2);
    top_index_.clear();
    for (const Key& k : *keys) {
      nodes_.push_back(NodeOfKey(k));
      top_index_.push_back(static_cast<int>(k.index));
    }
    top_count_ = keys->size();
    parent_of_.assign(nodes_.size(), -1);
    leaves_pending_.clear();
    if (TraceEnabled()) std::fprintf(stderr, "[hg trace] top level: %zu nodes, sort+build %.1f ms\n", top_count_, pt.Lap());
  }

  // A quick lower bound on the answer: the 'width' best lowest-resolution
  // candidates are each followed greedily (best child per level, one launch per
  // level for all of them) down to a leaf; the best such leaf that passes the
  // low-resolution gate gives a score that the final answer cannot be below.
  // Subtrees whose score is strictly below it can then be dropped without
  // touching either the answer or its tie-breaking.
  int GreedyDive(const std::vector<Key>& keys, size_t width, int max_depth, float min_score, float* bound) {
    *bound = -std::numeric_limits<float>::infinity();
    PhaseTimer pt;
    std::vector<Node> cur;
    cur.reserve(width);
    for (size_t i = 0; i < width && i < keys.size(); ++i) cur.push_back(NodeOfKey(keys[i]));
    for (int d = max_depth; d >= 1 && !cur.empty(); --d) {
      const size_t np = cur.size();
      int rc = scorer_.Reserve(np);
      if (rc != HG_OK) return rc;
      int4* hc = scorer_.HostCandidates();
      for (size_t i = 0; i < np; ++i) hc[i] = make_int4(cur[i].scan, cur[i].ox, cur[i].oy, cur[i].oz);
      const int half_width = 1 << (d - 1);
      rc = scorer_.RunChildren(d - 1, np, half_width, sp_.wxy, sp_.wz);
      if (rc != HG_OK) return rc;
      const int* sums = scorer_.HostSums();
      for (size_t i = 0; i < np; ++i) {
        int best_slot = 0, best_sum = -1;
        for (int slot = 0; slot < 8; ++slot) {
          if (sums[8 * i + slot] < 0) continue;
          ++scorer_.total_scored;
          if (sums[8 * i + slot] > best_sum) {
            best_sum = sums[8 * i + slot];
            best_slot = slot;
          }
        }
        Node& c = cur[i];
        c.ox += (best_slot & 1) ? half_width : 0;
        c.oy += (best_slot & 2) ? half_width : 0;
        c.oz += (best_slot & 4) ? half_width : 0;
        c.score = ScoreFromSum(std::max(best_sum, 0), scans_->N);
      }
    }
    std::sort(cur.begin(), cur.end(), [](const Node& a, const Node& b) { return a.score > b.score; });
    // Leaf gate in descending score, in batches, until one passes.
    for (size_t at = 0; at < cur.size(); at += 256) {
      const size_t end = std::min(cur.size(), at + 256);
      if (cur[at].score <= min_score) break;
      std::vector<Node> batch(cur.begin() + at, cur.begin() + end);
      std::vector<float> low;
      const int rc = LowResolutionScores(batch, &low);
      if (rc != HG_OK) return rc;
      for (size_t k = 0; k < batch.size(); ++k) {
        if (batch[k].score <= min_score) break;
        if (static_cast<double>(low[k]) >= m_->options.min_low_resolution_score) {
          *bound = batch[k].score;
          break;
        }
      }
      if (*bound > -std::numeric_limits<float>::infinity()) break;
    }
    if (TraceEnabled())
      std::fprintf(stderr, "[hg trace] greedy dive: %zu starts, bound %.6f, %.1f ms\n", cur.size(), *bound, pt.Lap());
    return HG_OK;
  }

  int UploadLowResolutionCloud() {
    const int n = data_->num_low_resolution_points;
    if (n <= 0) return HG_OK;
    int rc = ws_->d_low_points.Reserve(sizeof(float) * 3 * n);
    if (rc != HG_OK) return rc;
    HG_CUDA(cudaMemcpyAsync(ws_->d_low_points.p, data_->low_resolution_point_cloud, sizeof(float) * 3 * n,
                            cudaMemcpyHostToDevice, stream_));
    return HG_OK;
  }

  // BranchAndBound (:408-471) over nodes_[begin, begin + count) at 'depth'.
  BestCandidate Search(int begin, int count, int depth, float min_score) {
    BestCandidate unsuccessful;  // score -inf
    if (failed_ != HG_OK) return unsuccessful;
    if (depth == 0) {
      for (int i = begin; i < begin + count; ++i) {
        if (nodes_[i].score <= min_score) return unsuccessful;
        if (std::isnan(nodes_[i].low_resolution_score)) {
          failed_ = EvaluateLowResolution(begin, count, min_score);
          if (failed_ != HG_OK) return unsuccessful;
        }
        const float low = nodes_[i].low_resolution_score;
        if (static_cast<double>(low) >= m_->options.min_low_resolution_score) {
          BestCandidate b;
          b.node = i;
          b.score = nodes_[i].score;
          b.low_resolution_score = low;
          return b;
        }
      }
      return unsuccessful;
    }
    BestCandidate best;
    best.score = min_score;
    for (int i = begin; i < begin + count; ++i) {
      // The reference compares against its min_score argument (:437-440); a
      // candidate in (min_score, running best] still gets its children scored,
      // but the recursion then stops at the first child and changes nothing.
      // Outside strict order those candidates (and, the list being sorted, all
      // that follow) are skipped.
      if (nodes_[i].score <= (strict_ ? min_score : best.score)) break;
      if (nodes_[i].score < global_bound_) break;  // cannot hold the answer or a tie with it
      if (nodes_[i].first_child == -1) {
        failed_ = Speculate(begin, count, i, depth, strict_ ? min_score : best.score, best.score);
        if (failed_ != HG_OK) return unsuccessful;
      }
      visited_ += nodes_[i].num_children;
      if (nodes_[i].first_child == kChildrenPruned) continue;  // every child <= a bound that best.score has reached
      const BestCandidate sub = Search(nodes_[i].first_child, nodes_[i].num_children, depth - 1, best.score);
      if (best.score < sub.score) best = sub;  // std::max(best, sub): the earlier one wins ties
    }
    return best;
  }

  // Expands nodes_[trigger ..) of the sibling list [begin, begin+count) — every
  // sibling the reference's loop will reach, i.e. with score > 'list_min_score'
  // (the loop compares against its min_score argument, not the running best) —
  // and then their descendants that still beat 'bound' (the running best),
  // level by level, one scoring launch per level. The wave is truncated to a
  // prefix in DFS order; the budget grows geometrically from call to call so
  // that the first calls act as a quick dive that tightens the bound.
  int Speculate(int begin, int count, int trigger, int depth, float list_min_score, float bound) {
    std::vector<int> wave;
    for (int i = trigger; i < begin + count; ++i) {
      if (nodes_[i].score <= list_min_score || nodes_[i].score < global_bound_) break;
      if (nodes_[i].first_child == -1) wave.push_back(i);
      if (static_cast<int64_t>(wave.size()) >= wave_budget_) break;
    }
    const int64_t budget = wave_budget_;
    wave_budget_ = std::min<int64_t>(wave_budget_ * 8, int64_t(1) << 17);
    Node kids[8];
    for (int d = depth; d >= 1 && !wave.empty(); --d) {
      const size_t np = wave.size();
      int rc = scorer_.Reserve(np);
      if (rc != HG_OK) return rc;
      int4* hc = scorer_.HostCandidates();
      for (size_t i = 0; i < np; ++i) {
        const Node& p = nodes_[wave[i]];
        hc[i] = make_int4(p.scan, p.ox, p.oy, p.oz);
      }
      const int half_width = 1 << (d - 1);
      PhaseTimer wt;
      rc = scorer_.RunChildren(d - 1, np, half_width, sp_.wxy, sp_.wz);
      if (rc != HG_OK) return rc;
      wave_gpu_ms_ += wt.Lap();
      wave_parents_ += static_cast<int64_t>(np);
      const int* sums = scorer_.HostSums();
      // Smallest child sum that beats the bound and is not below the global
      // bound (score is non-decreasing in the sum).
      int beats_bound = SmallestSumBeating(bound);
      if (global_bound_ > -std::numeric_limits<float>::infinity()) {
        const float just_below = std::nextafter(global_bound_, -std::numeric_limits<float>::infinity());
        beats_bound = std::max(beats_bound, SmallestSumBeating(just_below));
      }
      std::vector<int> next;
      for (size_t w = 0; w < np; ++w) {
        const int* cs = sums + 8 * w;
        int nc = 0, best_sum = -1;
        for (int slot = 0; slot < 8; ++slot) {
          if (cs[slot] < 0) continue;
          ++nc;
          best_sum = std::max(best_sum, cs[slot]);
        }
        scorer_.total_scored += nc;
        Node& parent = nodes_[wave[w]];
        if (!strict_ && best_sum < beats_bound) {
          // The recursion into these children would stop at the first one
          // (:437-440 / :414-419 with min_score = running best >= bound).
          parent.first_child = kChildrenPruned;
          parent.num_children = nc;
          continue;
        }
        // Materialise in generation order (z outer, y, x inner), then the
        // reference's sort: std::sort on <= 16 elements is an insertion sort in
        // libstdc++, stable in generation order for equal scores.
        int k = 0;
        for (int slot = 0; slot < 8; ++slot) {
          if (cs[slot] < 0) continue;
          kids[k++] = Node{parent.scan, parent.ox + ((slot & 1) ? half_width : 0),
                           parent.oy + ((slot & 2) ? half_width : 0), parent.oz + ((slot & 4) ? half_width : 0),
                           ScoreFromSum(cs[slot], scans_->N), std::numeric_limits<float>::quiet_NaN(), -1, 0};
        }
        for (int a = 1; a < nc; ++a) {
          const Node v = kids[a];
          int b = a - 1;
          while (b >= 0 && v.score > kids[b].score) {
            kids[b + 1] = kids[b];
            --b;
          }
          kids[b + 1] = v;
        }
        ++materialised_;
        const int at = static_cast<int>(nodes_.size());
        const int parent_index = wave[w];
        nodes_.insert(nodes_.end(), kids, kids + nc);  // may reallocate: 'parent' is dead from here
        parent_of_.resize(nodes_.size(), -1);
        for (int c = 0; c < nc; ++c) parent_of_[at + c] = parent_index;
        nodes_[parent_index].first_child = at;
        nodes_[parent_index].num_children = nc;
        if (d - 1 >= 1) {
          for (int c = 0; c < nc; ++c) {
            if (nodes_[at + c].score <= bound || nodes_[at + c].score < global_bound_) break;
            if (static_cast<int64_t>(next.size()) < budget) next.push_back(at + c);
          }
        } else {
          leaves_pending_.push_back(at);
          leaves_pending_.push_back(nc);
        }
      }
      wave.swap(next);
    }
    // Leaf gate for every freshly created leaf list.
    int rc = HG_OK;
    if (!leaves_pending_.empty()) {
      std::vector<int> leaves;
      for (size_t k = 0; k < leaves_pending_.size(); k += 2) {
        for (int c = 0; c < leaves_pending_[k + 1]; ++c) {
          const int idx = leaves_pending_[k] + c;
          if (nodes_[idx].score <= bound) break;
          leaves.push_back(idx);
        }
      }
      leaves_pending_.clear();
      rc = LowResolutionFor(leaves);
    }
    return rc;
  }

  int EvaluateLowResolution(int begin, int count, float min_score) {
    std::vector<int> leaves;
    for (int i = begin; i < begin + count; ++i) {
      if (nodes_[i].score <= min_score) break;
      if (std::isnan(nodes_[i].low_resolution_score)) leaves.push_back(i);
    }
    return LowResolutionFor(leaves);
  }

  int LowResolutionFor(const std::vector<int>& leaves) {
    if (leaves.empty()) return HG_OK;
    std::vector<Node> list;
    list.reserve(leaves.size());
    for (int idx : leaves) list.push_back(nodes_[idx]);
    std::vector<float> scores;
    const int rc = LowResolutionScores(list, &scores);
    if (rc != HG_OK) return rc;
    for (size_t k = 0; k < leaves.size(); ++k) nodes_[leaves[k]].low_resolution_score = scores[k];
    return HG_OK;
  }

  // CreateLowResolutionMatcher's closure for the candidate poses of 'list'.
  int LowResolutionScores(const std::vector<Node>& list, std::vector<float>* out) {
    const size_t n = list.size();
    out->assign(n, -std::numeric_limits<float>::infinity());
    if (n == 0) return HG_OK;
    PhaseTimer lt;
    struct Acc { double* t; PhaseTimer* p; ~Acc() { *t += p->Lap(); } } acc{&low_ms_, &lt};
    const int np = data_->num_low_resolution_points;
    if (np <= 0) return HG_OK;  // score / points->size() with an empty cloud is 0 / 0 in the reference: never passes
    int rc = ws_->h_poses.Reserve(n * 7 * sizeof(float));
    if (rc != HG_OK) return rc;
    rc = ws_->h_low.Reserve(n * sizeof(float));
    if (rc != HG_OK) return rc;
    rc = ws_->d_low_poses.Reserve(n * 7 * sizeof(float));
    if (rc != HG_OK) return rc;
    rc = ws_->d_low_scores.Reserve(n * sizeof(float));
    if (rc != HG_OK) return rc;
    float* hp = ws_->h_poses.as<float>();
    for (size_t k = 0; k < n; ++k) {
      const Rigid r = PoseOf(list[k]);
      float* o = hp + 7 * k;
      o[0] = r.t.x; o[1] = r.t.y; o[2] = r.t.z; o[3] = r.q.w; o[4] = r.q.x; o[5] = r.q.y; o[6] = r.q.z;
    }
    if (n <= 256) {  // poses and scores through the mapped pinned buffers: one launch, one synchronisation
      LowResolutionKernel<<<static_cast<unsigned>((n + 3) / 4), 128, 0, stream_>>>(
          m_->low_grid->dev, ws_->d_low_points.as<float>(), np, hp, static_cast<int64_t>(n), ws_->h_low.as<float>());
      CountLaunch();
      ++low_launches_;
      HG_CUDA(cudaGetLastError());
      HG_CUDA(cudaStreamSynchronize(stream_));
    } else {
      HG_CUDA(cudaMemcpyAsync(ws_->d_low_poses.p, hp, n * 7 * sizeof(float), cudaMemcpyHostToDevice, stream_));
      LowResolutionKernel<<<static_cast<unsigned>((n + 3) / 4), 128, 0, stream_>>>(
          m_->low_grid->dev, ws_->d_low_points.as<float>(), np, ws_->d_low_poses.as<float>(),
          static_cast<int64_t>(n), ws_->d_low_scores.as<float>());
      CountLaunch();
      ++low_launches_;
      HG_CUDA(cudaGetLastError());
      HG_CUDA(cudaMemcpyAsync(ws_->h_low.p, ws_->d_low_scores.p, n * sizeof(float), cudaMemcpyDeviceToHost, stream_));
      HG_CUDA(cudaStreamSynchronize(stream_));
    }
    const float* hs = ws_->h_low.as<float>();
    for (size_t k = 0; k < n; ++k) (*out)[k] = hs[k];
    return HG_OK;
  }

  int TopAncestor(int node) const {
    while (node >= static_cast<int>(top_count_)) node = parent_of_[node];
    return node;
  }

  const hg_fcsm* m_;
  const SearchParameters sp_;
  const hg_node_data* data_;
  const ScanSet* scans_;
  cudaStream_t stream_;
  Scorer scorer_;
  bool strict_ = false;
  Workspace* ws_ = nullptr;
  bool lattice_scored_ = false;
  float global_bound_ = -std::numeric_limits<float>::infinity();
  int nxy_ = 0, nz_ = 0, per_scan_ = 0, step_ = 1;
 public:
  double wave_gpu_ms_ = 0.0, low_ms_ = 0.0;
  int64_t wave_parents_ = 0, materialised_ = 0;
 private:
  std::vector<int> h_lattice_sums_;  // lowest-resolution sums in generation order (strict / probe only)
  std::vector<int> top_index_;       // generation index of nodes_[0 .. top_count_)
  std::vector<int> parent_of_;
  std::vector<Node> nodes_;
  size_t top_count_ = 0;
  int64_t visited_ = 0;
  int64_t wave_budget_ = 8;
  int failed_ = HG_OK;
  int low_launches_ = 0;
  std::vector<int> leaves_pending_;
};

Rigid ToRigidF(const double* t, const double* q) {
  return Rigid{V3{static_cast<float>(t[0]), static_cast<float>(t[1]), static_cast<float>(t[2])},
               Q4{static_cast<float>(q[0]), static_cast<float>(q[1]), static_cast<float>(q[2]),
                  static_cast<float>(q[3])}};
}

// MatchWithSearchParameters (:202-228) with the branch-and-bound replayed on the host.
int HostReplayMatch(const hg_fcsm* m, const SearchParameters& sp, const Rigid& node_pose,
                              const Rigid& submap_pose, const hg_node_data* data, float min_score,
                              int shard_index, int num_shards, hg_match_result* result) {
  std::memset(result, 0, sizeof(*result));
  result->pose_rotation[0] = 1.0;
  HG_ON_DEVICE(m->device);
  // The calling host thread's own non-blocking stream: searches driven by different host
  // threads overlap on the device.
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "hg_fcsm: cannot create a stream");
  ScanSet scans;
  PhaseTimer pt;
  int rc = GenerateDiscreteScans(m, sp, data, node_pose, submap_pose, shard_index, num_shards, stream, &scans);
  if (rc != HG_OK) return rc;
  const double t_scans = pt.Lap();
  result->num_discrete_scans = scans.S;
  BranchAndBoundSearch search(m, sp, data, &scans, stream);
  BestCandidate best;
  rc = search.Run(min_score, m->options.strict_visit_order != 0, &best);
  if (rc != HG_OK) return rc;
  if (TraceEnabled())
    std::fprintf(stderr, "[hg trace] search: %d scans x %d pts, window %d/%d; discretise %.1f ms, branch-and-bound %.1f ms (waves on GPU %.1f ms, %lld parents, %lld materialised; leaf gate %.1f ms), scored %lld visited %lld launches %d\n",
                 scans.S, scans.N, sp.wxy, sp.wz, t_scans, pt.Lap(), search.wave_gpu_ms_,
                 static_cast<long long>(search.wave_parents_), static_cast<long long>(search.materialised_),
                 search.low_ms_, static_cast<long long>(search.scored()),
                 static_cast<long long>(search.visited()), search.launches());
  result->num_candidates_scored = search.scored();
  result->num_candidates_visited = search.visited();
  result->num_score_launches = search.launches();
  result->num_lowest_resolution_candidates = search.lattice_candidates();
  result->lowest_resolution_kernel_ms = search.lattice_ms();
  if (best.node >= 0 && best.score > min_score) {
    const Node& n = search.node(best.node);
    const Rigid pose = search.PoseOf(n);
    result->found = 1;
    result->score = best.score;
    result->pose_translation[0] = pose.t.x;
    result->pose_translation[1] = pose.t.y;
    result->pose_translation[2] = pose.t.z;
    result->pose_rotation[0] = pose.q.w;
    result->pose_rotation[1] = pose.q.x;
    result->pose_rotation[2] = pose.q.y;
    result->pose_rotation[3] = pose.q.z;
    result->rotational_score = scans.rotational_scores[n.scan];
    result->low_resolution_score = best.low_resolution_score;
    result->best_scan_index = scans.global_scan_index[n.scan];
    result->best_offset[0] = n.ox;
    result->best_offset[1] = n.oy;
    result->best_offset[2] = n.oz;
  }
  return HG_OK;
}

}  // namespace

namespace hg {
namespace {

// ------------------------------------------------ device-resident search --
//
// The same search without a host decision per level, for one problem or a batch of problems that
// share a matcher (all the (node, submap) pairs of one submap). Order-independent form of the
// reference's depth-first search: the answer is the best-scoring leaf with
// score > min_score that passes the low-resolution gate, and — every level of the pyramid being a
// maximum over the finer one — a subtree whose root scores below a known valid leaf cannot hold it.
//   phase A  lattice (one launch per problem), then a GREEDY DIVE from the best lattice candidate
//            of every discrete scan: best child per level, one launch per level for all starts of
//            all problems; the low-resolution gate on the dive's leaves gives every problem a
//            bound B = score of its best valid dive leaf;
//   phase B  every lattice candidate with score >= B seeds a frontier that is expanded level by
//            level on the device (children with score >= B are appended to the next level's list
//            with an atomic counter; the launches are persistent grids that read the counts from
//            device memory), the gate runs on all surviving leaves, and the leaves that reach the
//            best valid score are returned.
// One synchronisation at the end (two when the caller exchanges the bound between the phases, as
// the sharded full-submap search does). Without a bound (no dive leaf passed the gate) B is
// "score > min_score". Exactly one winner is the answer, ties included in the sense that equal
// scores were all kept; several winners (an exact tie), a bound that phase B cannot reach again
// (a parent scored below its child: the pyramid is designed, not proven, to be an upper bound)
// or a full list send that problem through the host replay above, which settles them in the
// reference's visiting order.

constexpr float kScoreMin = 0.1f;
constexpr float kScoreMax = 1.f - kScoreMin;
constexpr float kScoreScale = (kScoreMax - kScoreMin) / 255.f;

// ScoreFromSum on the device: the same three float operations, unfused.
__device__ __forceinline__ float DeviceScore(const int sum, const int num_points) {
  const float value = __fdiv_rn(static_cast<float>(sum), static_cast<float>(num_points));
  return __fadd_rn(kScoreMin, __fmul_rn(value, kScoreScale));
}

struct ProblemDev {
  int scan_begin, S;       // this problem's discrete scans in the batch-wide scan list
  int N, NL;               // high- and low-resolution points
  long long cells_offset;  // ints into the batch's cell array: [level][scan][point][3]
  long long level_stride;  // ints per level = S * N * 3
  int low_offset;          // points into the batch's low-resolution cloud array
  int wxy, wz;
  float min_score;
  long long point_offset;  // points into the batch's high-resolution cloud array
};

// One discrete scan: owning problem and the pieces of GetPoseFromCandidate (:400-406) that do not
// depend on the offset — translation of the scan pose and the renormalised rotation.
struct ScanDev {
  int problem;
  float tx, ty, tz;
  float qw, qx, qy, qz;
};

struct WinnerRec {
  int scan, ox, oy, oz;  // scan = index in the batch-wide scan list
  int sum;
  float low;
  int ancestor_sum;  // lattice sum of the leaf's lowest-resolution ancestor
  int problem;
};

struct SearchDev {
  const ProblemDev* problems;
  const ScanDev* scans;
  const int* cells;
  const float* low_points;
  const int* lattice;  // [total_scans * per_scan]
  int4* starts;        // [total_scans] dive state
  int* start_sums;     // -1 = the scan's best lattice candidate does not beat min_score
  int* start_dead;     // 1 = not among the scans the dive starts from
  int* bound_bits;     // [P] score bits of the best valid dive leaf, 0 = none
  int* thresh_bits;    // [P] phase B keeps score bits >= this
  int* best_bits;      // [P] best valid leaf of phase B
  int* win_count;      // [P]
  int4* list[2];       // frontier of depth d in list[d & 1]
  int* list_sum[2];
  float* leaf_low;
  int* counts;         // [depth] entries of the depth's list (may exceed list_cap: overflow)
  int list_cap;
  WinnerRec* winners;
  int* winners_total;
  int winners_cap;
  unsigned long long* scored;  // [P] candidates scored below the lattice
  int total_scans, per_scan, nxy, nz, step, max_depth, frd;
  double min_low_resolution_score;
};

__device__ __forceinline__ const int* CellsOf(const SearchDev& s, const ProblemDev& P, const int local_scan,
                                              const int depth) {
  const int level = max(0, depth - s.frd + 1);
  return s.cells + P.cells_offset + level * P.level_stride + static_cast<long long>(local_scan) * P.N * 3;
}

// DiscretizeKernel for every (scan, point) of a batch in one launch: blockIdx.y = scan of the
// batch-wide list, blockIdx.x = chunk of 256 points.
__global__ void __launch_bounds__(256) DiscretizeBatchKernel(const ProblemDev* __restrict__ problems,
                                                             const ScanDev* __restrict__ scans,
                                                             const float* __restrict__ points,
                                                             const float* __restrict__ poses, const float resolution,
                                                             const float inv_resolution, const int num_low_levels,
                                                             int* __restrict__ cells) {
  const int sg = blockIdx.y;
  const ProblemDev& P = problems[scans[sg].problem];
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P.N) return;
  const float* pt = points + 3 * (P.point_offset + i);
  const float* Q = poses + 7 * static_cast<long long>(sg);
  float wx, wy, wz;
  RotateF(Q[3], Q[4], Q[5], Q[6], pt[0], pt[1], pt[2], &wx, &wy, &wz);
  wx = __fadd_rn(wx, Q[0]);
  wy = __fadd_rn(wy, Q[1]);
  wz = __fadd_rn(wz, Q[2]);
  const int cx = CellIndexFast(wx, resolution, inv_resolution), cy = CellIndexFast(wy, resolution, inv_resolution),
            cz = CellIndexFast(wz, resolution, inv_resolution);
  int* o = cells + P.cells_offset + (static_cast<long long>(sg - P.scan_begin) * P.N + i) * 3;
  o[0] = cx; o[1] = cy; o[2] = cz;
  const int start_xy = -P.wxy, start_z = -P.wz;
  for (int l = 1; l <= num_low_levels; ++l) {
    o += P.level_stride;
    o[0] = ((cx + start_xy) >> l) - (start_xy >> l);
    o[1] = ((cy + start_xy) >> l) - (start_xy >> l);
    o[2] = ((cz + start_z) >> l) - (start_z >> l);
  }
}

// Lowest-resolution lattice of a batch of small searches (a local window has a handful of
// lattice candidates per scan): a CTA per (scan, eight candidates) loads every point's cell once
// for the eight gathers. Same sums as ScoreLatticeKernel, generation order.
__global__ void __launch_bounds__(256) LatticeBatchKernel(const SearchDev s, const BrickTable level,
                                                          const DenseLevel dense, int* __restrict__ lattice) {
  __shared__ int warp_sums[8][8];
  const int sg = blockIdx.y;
  const int first = blockIdx.x * 8;
  const ProblemDev& P = s.problems[s.scans[sg].problem];
  const int e = max(0, s.max_depth - s.frd + 1);
  const int* base = CellsOf(s, P, sg - P.scan_begin, s.max_depth);
  int ox[8], oy[8], oz[8];
  unsigned valid = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int at = first + k;
    ox[k] = oy[k] = oz[k] = 0;
    if (at < s.per_scan) {
      const int iz = at / (s.nxy * s.nxy), rem = at - iz * s.nxy * s.nxy, iy = rem / s.nxy, ix = rem - iy * s.nxy;
      ox[k] = (-P.wxy + ix * s.step) >> e;
      oy[k] = (-P.wxy + iy * s.step) >> e;
      oz[k] = (-P.wz + iz * s.step) >> e;
      valid |= 1u << k;
    }
  }
  int sums[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) sums[k] = 0;
#pragma unroll 2
  for (int i = threadIdx.x; i < P.N; i += 256) {
    const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if ((valid >> k) & 1u) {
        sums[k] += dense.data != nullptr ? LoadDenseU8(dense, x + ox[k], y + oy[k], z + oz[k])
                                         : LoadU8(level, x + ox[k], y + oy[k], z + oz[k]);
      }
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], d);
    if (lane == 0) warp_sums[warp][k] = sums[k];
  }
  __syncthreads();
  if (threadIdx.x < 8 && first + threadIdx.x < s.per_scan) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += warp_sums[w][threadIdx.x];
    lattice[static_cast<long long>(sg) * s.per_scan + first + threadIdx.x] = t;
  }
}

// Keeps the dive starts of the 'keep' best-scoring scans of every problem (rank by counting:
// blockIdx.x = problem, blockIdx.y = chunk of 256 of its scans); the others are marked in
// start_dead. The dive only has to produce a good bound.
__global__ void __launch_bounds__(256) SelectStartsKernel(const SearchDev s, const int keep) {
  __shared__ int tile[1024];
  const ProblemDev& P = s.problems[blockIdx.x];
  const int t = blockIdx.y * 256 + threadIdx.x;
  if (blockIdx.y * 256 >= P.S) return;  // whole CTAs leave; the others stay together for the barriers
  const int* sums = s.start_sums + P.scan_begin;
  const bool active = t < P.S;
  int dead = 0;
  if (P.S > keep) {
    const int mine = active ? sums[t] : 0;
    int rank = 0;
    // the problem's sums pass through shared memory 1 024 at a time: every thread compares its own
    // with all of them (read straight from global memory this loop was 100 us for 1 633 scans)
    for (int base = 0; base < P.S; base += 1024) {
      const int n = min(1024, P.S - base);
      __syncthreads();
      for (int j = threadIdx.x; j < n; j += 256) tile[j] = __ldg(sums + base + j);
      __syncthreads();
#pragma unroll 8
      for (int j = 0; j < n; ++j) {
        const int other = tile[j];
        rank += (other > mine || (other == mine && base + j < t)) ? 1 : 0;
      }
    }
    dead = rank >= keep ? 1 : 0;
  }
  if (active) s.start_dead[P.scan_begin + t] = dead;
}

// Dive start of every discrete scan: its best lattice candidate (first one on equal sums).
// WARPS warps per scan: one for the small lattices of local windows, eight for a full-submap
// lattice (4 913 sums per scan: a single warp walks them in 154 dependent rounds, ~100 us for the
// whole launch however many scans there are).
template <int WARPS>
__global__ void __launch_bounds__(256) ArgmaxPerScanKernel(const SearchDev s) {
  __shared__ int s_best[8], s_at[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sg = WARPS == 8 ? static_cast<int>(blockIdx.x) : static_cast<int>((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (sg >= s.total_scans) return;  // whole warps (WARPS == 1) or whole CTAs (WARPS == 8)
  const int* sums = s.lattice + static_cast<long long>(sg) * s.per_scan;
  int best = -1, at = 0x7fffffff;
  const int first = WARPS == 8 ? static_cast<int>(threadIdx.x) : lane;
#pragma unroll 4
  for (int i = first; i < s.per_scan; i += 32 * WARPS) {
    const int v = __ldg(sums + i);
    if (v > best) {
      best = v;
      at = i;
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const int ob = __shfl_xor_sync(0xffffffffu, best, d), oa = __shfl_xor_sync(0xffffffffu, at, d);
    if (ob > best || (ob == best && oa < at)) {
      best = ob;
      at = oa;
    }
  }
  if (WARPS == 8) {
    if (lane == 0) {
      s_best[warp] = best;
      s_at[warp] = at;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    for (int w = 1; w < 8; ++w) {
      if (s_best[w] > best || (s_best[w] == best && s_at[w] < at)) {
        best = s_best[w];
        at = s_at[w];
      }
    }
  } else if (lane != 0) {
    return;
  }
  const ProblemDev& P = s.problems[s.scans[sg].problem];
  const int iz = at / (s.nxy * s.nxy), rem = at - iz * s.nxy * s.nxy, iy = rem / s.nxy, ix = rem - iy * s.nxy;
  s.starts[sg] = make_int4(sg, -P.wxy + ix * s.step, -P.wxy + iy * s.step, -P.wz + iz * s.step);
  s.start_sums[sg] = DeviceScore(best, P.N) > P.min_score ? best : -1;
}

// Sums of the eight children of one parent (slot = 4 z + 2 y + x, offsets {0, half}^3, the loop
// nest at :441-461), computed by a whole CTA of 256 threads: every point's cell is loaded once and
// feeds eight independent gathers. 'valid' masks the children beyond the search window (the
// reference's 'break'); their sums stay 0. The totals land in totals[0..7] (shared memory); the
// function ends with a __syncthreads().
__device__ __forceinline__ void BlockChildrenSums(const BrickTable& level, const DenseLevel& dense, const int* base,
                                                  const int N, const int px, const int py, const int pz,
                                                  const int half, const int e, const unsigned valid,
                                                  int (*warp_sums)[8], int* totals) {
  int sums[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ox[2] = {px >> e, (px + half) >> e}, oy[2] = {py >> e, (py + half) >> e},
            oz[2] = {pz >> e, (pz + half) >> e};
#pragma unroll
  for (int k = 0; k < 8; ++k) sums[k] = 0;
  if (dense.data != nullptr) {
    const int dx = dense.dim[0], dy = dense.dim[1], dz = dense.dim[2];
#pragma unroll 2
    for (int i = threadIdx.x; i < N; i += 256) {
      const int x = __ldg(base + 3 * i) - dense.min[0], y = __ldg(base + 3 * i + 1) - dense.min[1],
                z = __ldg(base + 3 * i + 2) - dense.min[2];
      int xs[2], rows[4];
      bool xin[2], rin[4];
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        xs[a] = x + ox[a];
        xin[a] = static_cast<unsigned>(xs[a]) < static_cast<unsigned>(dx);
      }
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int yy = y + oy[b], zz = z + oz[c];
          rin[2 * c + b] = static_cast<unsigned>(yy) < static_cast<unsigned>(dy) &&
                           static_cast<unsigned>(zz) < static_cast<unsigned>(dz);
          rows[2 * c + b] = (zz * dy + yy) * dx;
        }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (((valid >> k) & 1u) && rin[k >> 1] && xin[k & 1]) sums[k] += __ldg(dense.data + rows[k >> 1] + xs[k & 1]);
      }
    }
  } else {
#pragma unroll 2
    for (int i = threadIdx.x; i < N; i += 256) {
      const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if ((valid >> k) & 1u) sums[k] += LoadU8(level, x + ox[k & 1], y + oy[(k >> 1) & 1], z + oz[k >> 2]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], d);
    if (lane == 0) warp_sums[warp][k] = sums[k];
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += warp_sums[w][threadIdx.x];
    totals[threadIdx.x] = t;
  }
  __syncthreads();
}

// Children of (px, py, pz) that lie inside the window, as a bit mask over the slots.
__device__ __forceinline__ unsigned ValidChildren(const int px, const int py, const int pz, const int half,
                                                  const int wxy, const int wz) {
  unsigned valid = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int cx = px + ((k & 1) ? half : 0), cy = py + ((k & 2) ? half : 0), cz = pz + ((k & 4) ? half : 0);
    if (cx <= wxy && cy <= wxy && cz <= wz) valid |= 1u << k;
  }
  return valid;
}

// The greedy dive: a CTA per start walks from the lowest resolution to a leaf; at every level it
// scores the eight children and the best one (the first on equal sums) becomes the next parent.
__global__ void __launch_bounds__(256) DiveKernel(const SearchDev s, const LevelDev* __restrict__ levels) {
  __shared__ int warp_sums[8][8];
  __shared__ int totals[8];
  const int sg = blockIdx.x;
  if (s.start_sums[sg] < 0 || s.start_dead[sg]) return;  // CTA-uniform
  int4 p = s.starts[sg];
  const ProblemDev& P = s.problems[s.scans[sg].problem];
  const int* cells0 = CellsOf(s, P, sg - P.scan_begin, 0);
  int best_sum = s.start_sums[sg];
  unsigned long long scored = 0;
  for (int child_depth = s.max_depth - 1; child_depth >= 0; --child_depth) {
    const int half = 1 << child_depth;
    const int e = max(0, child_depth - s.frd + 1);
    const unsigned valid = ValidChildren(p.y, p.z, p.w, half, P.wxy, P.wz);
    const LevelDev& L = levels[child_depth];
    BlockChildrenSums(L.table, L.dense, cells0 + e * P.level_stride, P.N, p.y, p.z, p.w, half, e, valid, warp_sums,
                      totals);
    int best = 0;
    best_sum = -1;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (((valid >> k) & 1u) && totals[k] > best_sum) {
        best_sum = totals[k];
        best = k;
      }
    }
    p = make_int4(sg, p.y + ((best & 1) ? half : 0), p.z + ((best & 2) ? half : 0), p.w + ((best & 4) ? half : 0));
    scored += __popc(valid);
    __syncthreads();  // 'totals' is rewritten by the next level
  }
  if (threadIdx.x != 0) return;
  s.starts[sg] = p;
  s.start_sums[sg] = best_sum;
  atomicAdd(&s.scored[s.scans[sg].problem], scored);
}

// LowResolutionMatcher closure (low_resolution_matcher.cc:24-69) for candidates: a warp per
// candidate evaluates the points' values in parallel and lane 0 adds them in cloud order — the
// reference's float sum, bit for bit (see LowResolutionKernel). DIVE: the candidates are the dive
// leaves and a valid one raises its problem's bound; otherwise they are phase B's leaves, the
// score is kept and a valid one raises its problem's best.
struct GateArgs {
  GridDev grid;                // the low-resolution grid
  float candidate_resolution;  // resolution the candidate offsets count in (the matcher's grid)
};
template <bool DIVE>
__global__ void __launch_bounds__(128) GateKernel(const SearchDev s, const GateArgs g) {
  __shared__ float staged[4][kLowChunk];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int count = DIVE ? s.total_scans : min(s.counts[0], s.list_cap);
  float* mine = staged[warp];
  for (int item = blockIdx.x * 4 + warp; item < count; item += gridDim.x * 4) {
    const int4 c = DIVE ? s.starts[item] : s.list[0][item];
    const int sum = DIVE ? s.start_sums[item] : s.list_sum[0][item];
    if (sum < 0 || (DIVE && s.start_dead[item])) continue;  // warp-uniform
    const ScanDev sc = s.scans[c.x];
    const ProblemDev& P = s.problems[sc.problem];
    const float score_value = DeviceScore(sum, P.N);
    // GetPoseFromCandidate (:400-406): Translation(resolution * offset) * scan pose, in float.
    const float tx = __fadd_rn(sc.tx, __fmul_rn(g.candidate_resolution, static_cast<float>(c.y)));
    const float ty = __fadd_rn(sc.ty, __fmul_rn(g.candidate_resolution, static_cast<float>(c.z)));
    const float tz = __fadd_rn(sc.tz, __fmul_rn(g.candidate_resolution, static_cast<float>(c.w)));
    float low = -__int_as_float(0x7f800000);  // an empty cloud is 0 / 0 in the reference: never passes
    if (P.NL > 0) {
      const float* points = s.low_points + 3 * static_cast<long long>(P.low_offset);
      float acc = 0.f;
      for (int base = 0; base < P.NL; base += kLowChunk) {
        const int n = min(kLowChunk, P.NL - base);
        for (int k = lane; k < n; k += 32) {
          const int i = base + k;
          float wx, wy, wz;
          RotateF(sc.qw, sc.qx, sc.qy, sc.qz, points[3 * i], points[3 * i + 1], points[3 * i + 2], &wx, &wy, &wz);
          wx = __fadd_rn(wx, tx);
          wy = __fadd_rn(wy, ty);
          wz = __fadd_rn(wz, tz);
          const int cx = CellIndexFast(wx, g.grid.resolution, g.grid.inv_resolution),
                    cy = CellIndexFast(wy, g.grid.resolution, g.grid.inv_resolution),
                    cz = CellIndexFast(wz, g.grid.resolution, g.grid.inv_resolution);
          float v;
          if (g.grid.kind == kGridProb) {
            v = Decode(g.grid.a, LoadU16(g.grid.table, cx, cy, cz));
          } else {
            const float tsd = Decode(g.grid.a, LoadTsdf(g.grid.table, cx, cy, cz).x);
            v = __fsub_rn(1.f, fabsf(__fdiv_rn(tsd, g.grid.max_tsd)));
          }
          mine[k] = v;
        }
        __syncwarp();
        if (lane == 0) {
          for (int k = 0; k < n; ++k) acc = __fadd_rn(acc, mine[k]);
        }
        __syncwarp();
      }
      low = __fdiv_rn(acc, static_cast<float>(P.NL));
    }
    if (lane == 0) {
      const bool valid = score_value > P.min_score && static_cast<double>(low) >= s.min_low_resolution_score;
      if (DIVE) {
        if (valid) atomicMax(&s.bound_bits[sc.problem], __float_as_int(score_value));
      } else {
        s.leaf_low[item] = low;
        if (valid) atomicMax(&s.best_bits[sc.problem], __float_as_int(score_value));
      }
    }
  }
}

// Phase B keeps candidates whose score bits are >= thresh_bits: the problem's bound (the larger of
// its own dive's and, for a sharded search, the one the other shards found), or the smallest
// score above min_score when there is none. Scores are >= 0.1, so their bit patterns order like
// the floats.
// exchange_bits: null, or one float per problem that the caller's collective left in device memory.
__global__ void ThresholdKernel(const SearchDev s, const int num_problems, const float external_bound,
                                const int* __restrict__ exchange_bits) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= num_problems) return;
  float t = nextafterf(s.problems[p].min_score, __int_as_float(0x7f800000));
  if (!(t > 0.f)) t = 0.f;
  int bits = __float_as_int(t);
  bits = max(bits, s.bound_bits[p]);
  if (external_bound > 0.f) bits = max(bits, __float_as_int(external_bound));
  if (exchange_bits != nullptr) {
    const float e = __int_as_float(exchange_bits[p]);
    if (e > 0.f && e <= 1.f) bits = max(bits, exchange_bits[p]);  // a score; anything else is ignored
  }
  s.thresh_bits[p] = bits;
}

// Warp-aggregated append of the lanes with 'keep' set; returns each lane's position (or -1).
__device__ __forceinline__ int WarpAppend(int* counter, const bool keep) {
  const unsigned mask = __ballot_sync(0xffffffffu, keep);
  if (mask == 0) return -1;
  const int lane = threadIdx.x & 31, leader = __ffs(mask) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(mask));
  base = __shfl_sync(0xffffffffu, base, leader);
  return keep ? base + __popc(mask & ((1u << lane) - 1u)) : -1;
}

// Lattice candidates that reach their problem's threshold -> the frontier of the lowest
// resolution.
__global__ void __launch_bounds__(256) SeedKernel(const SearchDev s) {
  const int n = s.total_scans * s.per_scan;  // < 2^31 (checked on the host)
  const int d = s.max_depth;
  for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
    const int i = base + threadIdx.x;
    bool keep = false;
    int sg = 0, at = 0, sum = 0;
    if (i < n) {
      sg = i / s.per_scan;
      at = i - sg * s.per_scan;
      sum = __ldg(s.lattice + i);
      const int p = s.scans[sg].problem;
      keep = __float_as_int(DeviceScore(sum, s.problems[p].N)) >= s.thresh_bits[p];
    }
    const int pos = WarpAppend(&s.counts[d], keep);
    if (keep) {
      if (pos < s.list_cap) {
        const ProblemDev& P = s.problems[s.scans[sg].problem];
        const int iz = at / (s.nxy * s.nxy), rem = at - iz * s.nxy * s.nxy, iy = rem / s.nxy, ix = rem - iy * s.nxy;
        s.list[d & 1][pos] = make_int4(sg, -P.wxy + ix * s.step, -P.wxy + iy * s.step, -P.wz + iz * s.step);
        s.list_sum[d & 1][pos] = sum;
      }
    }
  }
}

// Children of the frontier of 'child_depth + 1' that reach the threshold -> frontier of
// child_depth. A CTA per parent; persistent grid, the parent count is read from the device.
__global__ void __launch_bounds__(256) ExpandKernel(const SearchDev s, const BrickTable level, const DenseLevel dense,
                                                    const int child_depth) {
  __shared__ int warp_sums[8][8];
  __shared__ int totals[8];
  const int d = child_depth + 1;
  const int parents = min(s.counts[d], s.list_cap);
  const int half = 1 << child_depth;
  const int e = max(0, child_depth - s.frd + 1);
  for (int w = blockIdx.x; w < parents; w += gridDim.x) {
    const int4 p = s.list[d & 1][w];
    const int problem = s.scans[p.x].problem;
    const ProblemDev& P = s.problems[problem];
    const unsigned valid = ValidChildren(p.y, p.z, p.w, half, P.wxy, P.wz);
    BlockChildrenSums(level, dense, CellsOf(s, P, p.x - P.scan_begin, child_depth), P.N, p.y, p.z, p.w, half, e, valid,
                      warp_sums, totals);
    if (threadIdx.x < 32) {  // lanes 0..7 of the first warp each decide one child
      const int lane = threadIdx.x;
      const int mine = lane < 8 ? totals[lane] : 0;
      const bool keep = lane < 8 && ((valid >> lane) & 1u) &&
                        __float_as_int(DeviceScore(mine, P.N)) >= s.thresh_bits[problem];
      const int pos = WarpAppend(&s.counts[child_depth], keep);
      if (keep && pos < s.list_cap) {
        s.list[child_depth & 1][pos] = make_int4(p.x, p.y + ((lane & 1) ? half : 0), p.z + ((lane & 2) ? half : 0),
                                                 p.w + ((lane & 4) ? half : 0));
        s.list_sum[child_depth & 1][pos] = mine;
      }
      if (lane == 0) atomicAdd(&s.scored[problem], static_cast<unsigned long long>(__popc(valid)));
    }
    __syncthreads();  // 'totals' is rewritten by the next parent
  }
}

// The valid leaves that reach their problem's best valid score.
__global__ void __launch_bounds__(256) WinnersKernel(const SearchDev s) {
  const int count = min(s.counts[0], s.list_cap);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const int4 c = s.list[0][i];
    const int sum = s.list_sum[0][i];
    const int problem = s.scans[c.x].problem;
    const ProblemDev& P = s.problems[problem];
    const float score_value = DeviceScore(sum, P.N);
    const float low = s.leaf_low[i];
    if (!(score_value > P.min_score) || !(static_cast<double>(low) >= s.min_low_resolution_score)) continue;
    if (__float_as_int(score_value) != s.best_bits[problem]) continue;
    atomicAdd(&s.win_count[problem], 1);
    const int at = atomicAdd(s.winners_total, 1);
    if (at >= s.winners_cap) continue;
    const int ix = (c.y + P.wxy) >> s.max_depth, iy = (c.z + P.wxy) >> s.max_depth, iz = (c.w + P.wz) >> s.max_depth;
    WinnerRec r;
    r.scan = c.x; r.ox = c.y; r.oy = c.z; r.oz = c.w;
    r.sum = sum;
    r.low = low;
    r.ancestor_sum = s.lattice[static_cast<long long>(c.x) * s.per_scan + (iz * s.nxy + iy) * s.nxy + ix];
    r.problem = problem;
    s.winners[at] = r;
  }
}

// Sums of explicit candidates (scan of the batch-wide list, offset) at one depth: a warp each. Used
// to order exactly tied winners (the scores of their ancestors decide).
__global__ void __launch_bounds__(256) ScoreListKernel(const SearchDev s, const BrickTable level, const DenseLevel dense,
                                                       const int depth, const int4* __restrict__ candidates,
                                                       const int n, int* __restrict__ sums) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const int4 c = candidates[w];
  const ProblemDev& P = s.problems[s.scans[c.x].problem];
  const int e = max(0, depth - s.frd + 1);
  const int ox = c.y >> e, oy = c.z >> e, oz = c.w >> e;
  const int* base = CellsOf(s, P, c.x - P.scan_begin, depth);
  int sum = 0;
  for (int i = lane; i < P.N; i += 32) {
    const int x = __ldg(base + 3 * i), y = __ldg(base + 3 * i + 1), z = __ldg(base + 3 * i + 2);
    sum += dense.data != nullptr ? LoadDenseU8(dense, x + ox, y + oy, z + oz) : LoadU8(level, x + ox, y + oy, z + oz);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  if (lane == 0) sums[w] = sum;
}

}  // namespace
}  // namespace hg

namespace {

struct CtrlDev {
  int counts[32];
  int winners_total;
  int pad;
};

// Host driver of the device-resident search: problems are added (host half of
// GenerateDiscreteScans), then phase A and phase B are queued on the calling thread's stream.
class DeviceSearch {
 public:
  struct Outcome {
    bool found = false;
    bool needs_host_replay = false;  // exact tie, unconfirmed bound or a full list
    int scan = 0, ox = 0, oy = 0, oz = 0;  // scan = local index in the problem's scan list
    float score = 0.f, low = 0.f;
    float ancestor_score = 0.f;
    int64_t ancestor_index = 0;  // generation index of the lowest-resolution ancestor, unsharded scan numbering
    int64_t scored = 0;
    float local_bound = 0.f;  // this problem's own dive bound (0 = none)
    bool order_unresolved = false;  // tied winners below equal-scoring ancestors inside one shard of a sweep
  };

  DeviceSearch(const hg_fcsm* m, cudaStream_t stream) : m_(m), stream_(stream), ws_(&GetWorkspace(m->device)) {}
  ~DeviceSearch() {
    if (e0_) cudaEventDestroy(e0_);
    if (e1_) cudaEventDestroy(e1_);
    for (auto& m : marks_) cudaEventDestroy(m.second);
  }

  int AddProblem(const SearchParameters& sp, const Rigid& node_pose, const Rigid& submap_pose,
                 const hg_node_data* data, float min_score, int shard_index, int num_shards) {
    if (!problems_.empty()) {
      HG_REQUIRE(sp.wxy == problems_[0].sp.wxy && sp.wz == problems_[0].sp.wz,
                 "device search: the problems of one batch must share their search window");
    }
    problems_.emplace_back();
    Problem& p = problems_.back();
    p.sp = sp;
    p.data = data;
    p.min_score = min_score;
    return ComputeScanPoses(m_, sp, data, node_pose, submap_pose, shard_index, num_shards, &p.scans);
  }
  int num_problems() const { return static_cast<int>(problems_.size()); }
  const ScanSet& scans(int p) const { return problems_[p].scans; }
  float lattice_ms() const { return lattice_ms_; }
  int64_t lattice_candidates(int p) const { return static_cast<int64_t>(problems_[p].scans.S) * per_scan_; }
  int launches() const { return launches_; }

  // Uploads, discretises, scores the lattices, dives and gates the dive's leaves.
  int RunPhaseA(bool time_lattice) {
    const int P = num_problems();
    const int D = m_->options.branch_and_bound_depth;
    max_depth_ = D - 1;
    HG_REQUIRE(D >= 1 && D <= 32, "device search: bad branch_and_bound_depth");
    const SearchParameters& sp = problems_[0].sp;
    step_ = 1 << max_depth_;
    nxy_ = (2 * sp.wxy + step_) / step_;
    nz_ = (2 * sp.wz + step_) / step_;
    per_scan_ = nxy_ * nxy_ * nz_;
    // ---- layout of the upload arena: points | low points | scan poses (7 floats) | ScanDev | ProblemDev
    size_t n_points = 0, n_low = 0, n_scans = 0, cell_ints = 0;
    std::vector<ProblemDev> pd(P);
    for (int p = 0; p < P; ++p) {
      const ScanSet& sc = problems_[p].scans;
      ProblemDev& d = pd[p];
      d.scan_begin = static_cast<int>(n_scans);
      d.S = sc.S;
      d.N = sc.N;
      d.NL = std::max(0, problems_[p].data->num_low_resolution_points);
      d.cells_offset = static_cast<long long>(cell_ints);
      d.level_stride = static_cast<long long>(sc.S) * sc.N * 3;
      d.low_offset = static_cast<int>(n_low);
      d.wxy = sp.wxy;
      d.wz = sp.wz;
      d.min_score = problems_[p].min_score;
      d.point_offset = static_cast<long long>(n_points);
      problems_[p].point_offset = n_points;
      n_points += static_cast<size_t>(sc.N);
      n_low += static_cast<size_t>(d.NL);
      n_scans += static_cast<size_t>(sc.S);
      cell_ints += static_cast<size_t>(1 + sc.num_low_levels) * sc.S * sc.N * 3;
    }
    total_scans_ = static_cast<int>(n_scans);
    HG_REQUIRE(n_scans * static_cast<size_t>(per_scan_) < (size_t(1) << 31), "device search: lattice too large");
    if (total_scans_ == 0) return HG_OK;
    const size_t off_low = n_points * 3 * sizeof(float);
    const size_t off_poses = off_low + n_low * 3 * sizeof(float);
    const size_t off_scans = (off_poses + n_scans * 7 * sizeof(float) + 15) & ~size_t(15);
    const size_t off_problems = (off_scans + n_scans * sizeof(ScanDev) + 15) & ~size_t(15);
    const size_t upload_bytes = off_problems + P * sizeof(ProblemDev);
    int rc = ws_->h_upload.Reserve(upload_bytes);
    if (rc != HG_OK) return rc;
    rc = ws_->d_upload.Reserve(upload_bytes);
    if (rc != HG_OK) return rc;
    char* h = ws_->h_upload.as<char>();
    const int flavor = m_->options.float_flavor;
    for (int p = 0; p < P; ++p) {
      const Problem& pr = problems_[p];
      std::memcpy(h + pr.point_offset * 3 * sizeof(float), pr.data->high_resolution_point_cloud,
                  static_cast<size_t>(pr.scans.N) * 3 * sizeof(float));
      if (pd[p].NL > 0) {
        std::memcpy(h + off_low + static_cast<size_t>(pd[p].low_offset) * 3 * sizeof(float),
                    pr.data->low_resolution_point_cloud, static_cast<size_t>(pd[p].NL) * 3 * sizeof(float));
      }
      float* poses = reinterpret_cast<float*>(h + off_poses) + 7 * static_cast<size_t>(pd[p].scan_begin);
      ScanDev* sd = reinterpret_cast<ScanDev*>(h + off_scans) + pd[p].scan_begin;
      for (int s = 0; s < pr.scans.S; ++s) {
        const Rigid& r = pr.scans.poses[s];
        float* o = poses + 7 * s;
        o[0] = r.t.x; o[1] = r.t.y; o[2] = r.t.z; o[3] = r.q.w; o[4] = r.q.x; o[5] = r.q.y; o[6] = r.q.z;
        // GetPoseFromCandidate with a zero offset: what every candidate of this scan shares.
        const Rigid leaf = fgeom::RigidProduct(Rigid{V3{0.f, 0.f, 0.f}, Q4{1.f, 0.f, 0.f, 0.f}}, r, flavor);
        sd[s] = ScanDev{p, leaf.t.x, leaf.t.y, leaf.t.z, leaf.q.w, leaf.q.x, leaf.q.y, leaf.q.z};
      }
    }
    std::memcpy(h + off_problems, pd.data(), P * sizeof(ProblemDev));
    Mark("start");
    HG_CUDA(cudaMemcpyAsync(ws_->d_upload.p, h, upload_bytes, cudaMemcpyHostToDevice, stream_));
    Mark("upload");
    char* d = ws_->d_upload.as<char>();
    // ---- device state
    const size_t lattice_n = n_scans * per_scan_;
    list_cap_ = kListCapacity;
    winners_cap_ = std::max(4096, 8 * P);
    if ((rc = ws_->d_cells.Reserve(cell_ints * sizeof(int))) != HG_OK) return rc;
    if ((rc = ws_->d_lattice.Reserve(lattice_n * sizeof(int))) != HG_OK) return rc;
    if ((rc = ws_->d_starts.Reserve(n_scans * (sizeof(int4) + 2 * sizeof(int)))) != HG_OK) return rc;
    if ((rc = ws_->d_lists.Reserve(static_cast<size_t>(list_cap_) * (2 * sizeof(int4) + 2 * sizeof(int) + sizeof(float)))) != HG_OK)
      return rc;
    // control block: CtrlDev | per problem: bound, thresh, best, win_count (ints) | scored (u64) | winners
    const size_t off_pp = sizeof(CtrlDev);
    const size_t off_scored = (off_pp + 4 * P * sizeof(int) + 7) & ~size_t(7);
    const size_t off_win = (off_scored + P * sizeof(unsigned long long) + 15) & ~size_t(15);
    ctrl_bytes_ = off_win + winners_cap_ * sizeof(WinnerRec);
    off_pp_ = off_pp;
    off_scored_ = off_scored;
    off_win_ = off_win;
    if ((rc = ws_->d_ctrl.Reserve(ctrl_bytes_)) != HG_OK) return rc;
    if ((rc = ws_->h_ctrl.Reserve(ctrl_bytes_)) != HG_OK) return rc;
    HG_CUDA(cudaMemsetAsync(ws_->d_ctrl.p, 0, off_win, stream_));
    char* c = ws_->d_ctrl.as<char>();
    SearchDev& s = dev_;
    s.problems = reinterpret_cast<const ProblemDev*>(d + off_problems);
    s.scans = reinterpret_cast<const ScanDev*>(d + off_scans);
    s.cells = ws_->d_cells.as<int>();
    s.low_points = reinterpret_cast<const float*>(d + off_low);
    s.lattice = ws_->d_lattice.as<int>();
    s.starts = ws_->d_starts.as<int4>();
    s.start_sums = reinterpret_cast<int*>(ws_->d_starts.as<int4>() + n_scans);
    s.start_dead = s.start_sums + n_scans;
    int* pp = reinterpret_cast<int*>(c + off_pp);
    s.bound_bits = pp;
    s.thresh_bits = pp + P;
    s.best_bits = pp + 2 * P;
    s.win_count = pp + 3 * P;
    char* l = ws_->d_lists.as<char>();
    s.list[0] = reinterpret_cast<int4*>(l);
    s.list[1] = s.list[0] + list_cap_;
    s.list_sum[0] = reinterpret_cast<int*>(s.list[1] + list_cap_);
    s.list_sum[1] = s.list_sum[0] + list_cap_;
    s.leaf_low = reinterpret_cast<float*>(s.list_sum[1] + list_cap_);
    CtrlDev* ctrl = reinterpret_cast<CtrlDev*>(c);
    s.counts = ctrl->counts;
    s.list_cap = list_cap_;
    s.winners = reinterpret_cast<WinnerRec*>(c + off_win);
    s.winners_total = &ctrl->winners_total;
    s.winners_cap = winners_cap_;
    s.scored = reinterpret_cast<unsigned long long*>(c + off_scored);
    s.total_scans = total_scans_;
    s.per_scan = per_scan_;
    s.nxy = nxy_;
    s.nz = nz_;
    s.step = step_;
    s.max_depth = max_depth_;
    s.frd = problems_[0].scans.frd;
    s.min_low_resolution_score = m_->options.min_low_resolution_score;
    // ---- discretise: one launch for the batch (per problem when the grid would not fit)
    const float* d_points = reinterpret_cast<const float*>(d);
    const float* d_poses = reinterpret_cast<const float*>(d + off_poses);
    const float resolution = m_->resolution;
    int max_n = 0;
    for (int p = 0; p < P; ++p) max_n = std::max(max_n, problems_[p].scans.N);
    if (total_scans_ <= 65535) {
      DiscretizeBatchKernel<<<dim3((max_n + 255) / 256, total_scans_), 256, 0, stream_>>>(
          s.problems, s.scans, d_points, d_poses, resolution, 1.f / resolution, problems_[0].scans.num_low_levels,
          ws_->d_cells.as<int>());
      CountLaunch();
      ++launches_;
    } else {
      for (int p = 0; p < P; ++p) {
        const ScanSet& sc = problems_[p].scans;
        if (sc.S == 0) continue;
        const int64_t total = static_cast<int64_t>(sc.S) * sc.N;
        DiscretizeKernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, stream_>>>(
            d_points + 3 * problems_[p].point_offset, sc.N, d_poses + 7 * static_cast<size_t>(pd[p].scan_begin), sc.S,
            resolution, 1.f / resolution, sc.num_low_levels, -sp.wxy, -sp.wz, ws_->d_cells.as<int>() + pd[p].cells_offset);
        CountLaunch();
        ++launches_;
      }
    }
    HG_CUDA(cudaGetLastError());
    Mark("discretise");
    if (time_lattice) {
      HG_CUDA(cudaEventCreate(&e0_));
      HG_CUDA(cudaEventCreate(&e1_));
      HG_CUDA(cudaEventRecord(e0_, stream_));
    }
    // ---- lattice: small lattices (local windows) in one launch for the batch; a large one
    // (full submap) per problem through the residue-row kernel
    const int lattice_level = std::max(0, max_depth_ - s.frd + 1);
    if (nxy_ < 5 && total_scans_ <= 65535) {
      const hg_stack_level& L = m_->stack->levels[max_depth_];
      LatticeBatchKernel<<<dim3((per_scan_ + 7) / 8, total_scans_), 256, 0, stream_>>>(s, L.table, L.dense,
                                                                                      ws_->d_lattice.as<int>());
      CountLaunch();
      ++launches_;
    } else {
      for (int p = 0; p < P; ++p) {
        const ScanSet& sc = problems_[p].scans;
        if (sc.S == 0) continue;
        const int* cells = ws_->d_cells.as<int>() + pd[p].cells_offset + lattice_level * pd[p].level_stride;
        rc = LaunchLattice(m_, cells, sc.S, sc.N, max_depth_, nxy_, nz_, sp.wxy, sp.wz,
                           ws_->d_lattice.as<int>() + static_cast<size_t>(pd[p].scan_begin) * per_scan_, stream_, nullptr);
        if (rc != HG_OK) return rc;
        ++launches_;
      }
    }
    if (time_lattice) HG_CUDA(cudaEventRecord(e1_, stream_));
    Mark("lattice");
    if (std::getenv("HG_LATTICE_CHECK") != nullptr) {
      // Development / test aid: whatever kernel scored the lattice, against the brick-table kernel, every sum.
      const hg_stack_level& L = m_->stack->levels[max_depth_];
      for (int p = 0; p < P; ++p) {
        const ScanSet& sc = problems_[p].scans;
        const size_t n = static_cast<size_t>(sc.S) * per_scan_;
        if (n == 0) continue;
        DeviceBuffer check;
        if ((rc = check.Reserve(n * sizeof(int))) != HG_OK) return rc;
        const int* cells = ws_->d_cells.as<int>() + pd[p].cells_offset + lattice_level * pd[p].level_stride;
        ScoreLatticeKernel<<<static_cast<unsigned>((n * 32 + 255) / 256), 256, 0, stream_>>>(
            L.table, cells, sc.N, ReduceExponentOf(m_, max_depth_), nxy_, nz_, sp.wxy, sp.wz, 1 << max_depth_,
            static_cast<int64_t>(n), check.as<int>());
        CountLaunch();
        std::vector<int> got(n), want(n);
        HG_CUDA(cudaMemcpyAsync(got.data(), ws_->d_lattice.as<int>() + static_cast<size_t>(pd[p].scan_begin) * per_scan_,
                                n * sizeof(int), cudaMemcpyDeviceToHost, stream_));
        HG_CUDA(cudaMemcpyAsync(want.data(), check.p, n * sizeof(int), cudaMemcpyDeviceToHost, stream_));
        HG_CUDA(cudaStreamSynchronize(stream_));
        for (size_t i = 0; i < n; ++i) {
          if (got[i] != want[i]) {
            SetError("HG_LATTICE_CHECK: lattice sum " + std::to_string(i) + " of problem " + std::to_string(p) + " is " +
                     std::to_string(got[i]) + ", the brick-table kernel gives " + std::to_string(want[i]));
            return HG_ERR_CUDA;
          }
        }
      }
    }
    // ---- dive
    if (per_scan_ >= 1024) {
      ArgmaxPerScanKernel<8><<<total_scans_, 256, 0, stream_>>>(s);
    } else {
      ArgmaxPerScanKernel<1><<<(total_scans_ * 32 + 255) / 256, 256, 0, stream_>>>(s);
    }
    CountLaunch();
    ++launches_;
    {
      // The dive only has to produce a good bound: a few dozen starts per problem of a batch of local
      // searches, a few hundred for a single (full-submap) search — neighbouring yaws are near copies.
      int max_s = 1;
      for (int p = 0; p < P; ++p) max_s = std::max(max_s, problems_[p].scans.S);
      SelectStartsKernel<<<dim3(P, (max_s + 255) / 256), 256, 0, stream_>>>(
          s, P > 1 ? kDiveStartsPerProblem : kDiveStartsSingle);
      CountLaunch();
      ++launches_;
    }
    Mark("argmax+starts");
    if (max_depth_ >= 1) {
      DiveKernel<<<total_scans_, 256, 0, stream_>>>(s, m_->d_levels);
      CountLaunch();
      ++launches_;
    }
    Mark("dive");
    GateKernel<true><<<std::min(PersistentBlocks(16), (total_scans_ + 3) / 4), 128, 0, stream_>>>(s, GateGrid());
    CountLaunch();
    ++launches_;
    Mark("dive-gate");
    HG_CUDA(cudaGetLastError());
    phase_a_done_ = true;
    return HG_OK;
  }

  // Queues a copy of every problem's dive bound into a separate device array and returns it: the
  // caller's collective (an all-reduce-max over the shards of the search) works on that array in
  // place, on this search's stream or one ordered after it, and RunPhaseB reads it — no host round
  // trip between the phases. The problem's own bound stays where it is (phase B must meet that leaf again).
  int OfferBounds(float** exchange) {
    const int P = num_problems();
    int rc;
    if ((rc = ws_->d_exchange.Reserve(std::max(P, 1) * sizeof(int))) != HG_OK) return rc;
    exchange_ = ws_->d_exchange.as<int>();
    if (total_scans_ == 0) {
      HG_CUDA(cudaMemsetAsync(exchange_, 0, std::max(P, 1) * sizeof(int), stream_));
    } else {
      HG_CUDA(cudaMemcpyAsync(exchange_, dev_.bound_bits, P * sizeof(int), cudaMemcpyDeviceToDevice, stream_));
    }
    *exchange = reinterpret_cast<float*>(exchange_);
    return HG_OK;
  }

  // Synchronises and returns every problem's own dive bound (0 = none).
  int ReadBounds(std::vector<float>* bounds) {
    const int P = num_problems();
    bounds->assign(P, 0.f);
    if (total_scans_ == 0) return HG_OK;
    HG_CUDA(cudaMemcpyAsync(ws_->h_ctrl.as<char>() + off_pp_, ws_->d_ctrl.as<char>() + off_pp_, P * sizeof(int),
                            cudaMemcpyDeviceToHost, stream_));
    HG_CUDA(cudaStreamSynchronize(stream_));
    const int* bits = reinterpret_cast<const int*>(ws_->h_ctrl.as<char>() + off_pp_);
    for (int p = 0; p < P; ++p) std::memcpy(&(*bounds)[p], &bits[p], sizeof(float));
    return HG_OK;
  }

  // Frontier expansion under the bound, leaf gate, winners; one synchronisation at the end.
  // external_bound > 0: a bound found elsewhere (other shards of the same search).
  int RunPhaseB(float external_bound, std::vector<Outcome>* outcomes) {
    const int P = num_problems();
    outcomes->assign(P, Outcome());
    if (total_scans_ == 0) return HG_OK;
    const SearchDev& s = dev_;
    Mark("(host: bound exchange)");
    ThresholdKernel<<<(P + 127) / 128, 128, 0, stream_>>>(s, P, external_bound, exchange_);
    SeedKernel<<<PersistentBlocks(8), 256, 0, stream_>>>(s);
    CountLaunch(2);
    launches_ += 2;
    Mark("threshold+seed");
    for (int depth = max_depth_; depth >= 1; --depth) {
      const hg_stack_level& L = m_->stack->levels[depth - 1];
      ExpandKernel<<<PersistentBlocks(8), 256, 0, stream_>>>(s, L.table, L.dense, depth - 1);
      CountLaunch();
      ++launches_;
      Mark("expand");
    }
    GateKernel<false><<<PersistentBlocks(16), 128, 0, stream_>>>(s, GateGrid());
    Mark("gate");
    WinnersKernel<<<PersistentBlocks(4), 256, 0, stream_>>>(s);
    CountLaunch(2);
    launches_ += 2;
    Mark("winners");
    HG_CUDA(cudaGetLastError());
    HG_CUDA(cudaMemcpyAsync(ws_->h_ctrl.p, ws_->d_ctrl.p, ctrl_bytes_, cudaMemcpyDeviceToHost, stream_));
    Mark("readback");
    HG_CUDA(cudaStreamSynchronize(stream_));
    PrintMarks();
    if (e0_ != nullptr) HG_CUDA(cudaEventElapsedTime(&lattice_ms_, e0_, e1_));
    const char* h = ws_->h_ctrl.as<char>();
    const CtrlDev* ctrl = reinterpret_cast<const CtrlDev*>(h);
    const int* pp = reinterpret_cast<const int*>(h + off_pp_);
    const unsigned long long* scored = reinterpret_cast<const unsigned long long*>(h + off_scored_);
    const WinnerRec* winners = reinterpret_cast<const WinnerRec*>(h + off_win_);
    bool overflow = false;
    for (int dd = 0; dd <= max_depth_; ++dd) overflow = overflow || ctrl->counts[dd] > list_cap_;
    if (TraceEnabled()) {
      std::fprintf(stderr, "[hg trace] device search: %d problems, %d scans, frontier sizes", P, total_scans_);
      for (int dd = max_depth_; dd >= 0; --dd) std::fprintf(stderr, " %d", ctrl->counts[dd]);
      std::fprintf(stderr, ", winners %d\n", ctrl->winners_total);
    }
    const int recorded = std::min(ctrl->winners_total, winners_cap_);
    for (int p = 0; p < P; ++p) {
      Outcome& o = (*outcomes)[p];
      const ScanSet& sc = problems_[p].scans;
      o.scored = static_cast<int64_t>(scored[p]) + static_cast<int64_t>(sc.S) * per_scan_;
      std::memcpy(&o.local_bound, &pp[p], sizeof(float));
      const int thresh_bits = pp[P + p], wins = pp[3 * P + p];
      if (overflow || ctrl->winners_total > winners_cap_) {
        o.needs_host_replay = sc.S > 0;
        continue;
      }
      if (wins == 0) {
        // The problem's own bound is a leaf phase B must meet again; if it did not, a parent
        // scored below its child somewhere.
        o.needs_host_replay = pp[p] != 0 && thresh_bits == pp[p];
        continue;
      }
      int chosen = -1;
      if (wins > 1) {
        // exact tie: the leaf the reference's search meets first wins
        std::vector<WinnerRec> tied;
        for (int k = 0; k < recorded; ++k)
          if (winners[k].problem == p) tied.push_back(winners[k]);
        int first = -1;
        const int rc = FirstVisited(p, tied, &first, &o.order_unresolved);
        if (rc != HG_OK) return rc;
        if (first < 0) {
          o.needs_host_replay = true;
          continue;
        }
        for (int k = 0; k < recorded; ++k) {
          const WinnerRec& w = winners[k];
          if (w.problem == p && w.scan == tied[first].scan && w.ox == tied[first].ox && w.oy == tied[first].oy &&
              w.oz == tied[first].oz) {
            chosen = k;
          }
        }
      }
      for (int k = 0; k < recorded; ++k) {
        const WinnerRec& w = winners[k];
        if (w.problem != p || (chosen >= 0 && k != chosen)) continue;
        o.found = true;
        o.scan = w.scan - scan_begin(p);
        o.ox = w.ox; o.oy = w.oy; o.oz = w.oz;
        o.score = ScoreFromSum(w.sum, sc.N);
        o.low = w.low;
        o.ancestor_score = ScoreFromSum(w.ancestor_sum, sc.N);
        const int ix = (w.ox + problems_[p].sp.wxy) >> max_depth_, iy = (w.oy + problems_[p].sp.wxy) >> max_depth_,
                  iz = (w.oz + problems_[p].sp.wz) >> max_depth_;
        o.ancestor_index = static_cast<int64_t>(sc.global_scan_index[o.scan]) * per_scan_ + (iz * nxy_ + iy) * nxy_ + ix;
        break;
      }
    }
    return HG_OK;
  }

  // Among exactly tied winners of problem p, the one the reference's depth-first search meets first
  // (:408-471): subtrees are visited in the order of the sorted candidate lists — the lowest-resolution
  // list by std::sort on the score (:384), every list of children by score with equal scores in
  // generation order (std::sort on <= 16 elements is an insertion sort in libstdc++: stable) — and
  // std::max keeps the earlier of equal leaves. So the winners are compared by the scores of their
  // ancestors, level by level from the top. *first = -1 when this cannot be settled here.
  int FirstVisited(int p, const std::vector<WinnerRec>& tied, int* first, bool* order_unresolved) {
    *first = -1;
    *order_unresolved = false;
    const int n = static_cast<int>(tied.size());
    if (n < 2 || n > 4096) return HG_OK;
    const SearchParameters& sp = problems_[p].sp;
    const ScanSet& sc = problems_[p].scans;
    // ancestors at depth d = 1 .. max_depth - 1 (the leaf's sum and the top ancestor's are known)
    const auto ancestor = [&](const WinnerRec& w, int d, int* o) {
      o[0] = -sp.wxy + (((w.ox + sp.wxy) >> d) << d);
      o[1] = -sp.wxy + (((w.oy + sp.wxy) >> d) << d);
      o[2] = -sp.wz + (((w.oz + sp.wz) >> d) << d);
    };
    // scores[d][k]: score of winner k's ancestor at depth d (the reference compares float scores)
    std::vector<std::vector<float>> scores(max_depth_ + 1, std::vector<float>(n, 0.f));
    for (int k = 0; k < n; ++k) {
      scores[0][k] = ScoreFromSum(tied[k].sum, sc.N);
      scores[max_depth_][k] = ScoreFromSum(tied[k].ancestor_sum, sc.N);
    }
    if (max_depth_ >= 2) {
      const size_t bytes = static_cast<size_t>(n) * (sizeof(int4) + sizeof(int));
      int rc = ws_->h_cand.Reserve(bytes);
      if (rc != HG_OK) return rc;
      if ((rc = ws_->d_cand.Reserve(bytes)) != HG_OK) return rc;
      int4* hc = ws_->h_cand.as<int4>();
      int* hs = reinterpret_cast<int*>(hc + n);
      int4* dc = ws_->d_cand.as<int4>();
      int* ds = reinterpret_cast<int*>(dc + n);
      for (int d = 1; d < max_depth_; ++d) {
        for (int k = 0; k < n; ++k) {
          int o[3];
          ancestor(tied[k], d, o);
          hc[k] = make_int4(tied[k].scan, o[0], o[1], o[2]);
        }
        const hg_stack_level& L = m_->stack->levels[d];
        HG_CUDA(cudaMemcpyAsync(dc, hc, n * sizeof(int4), cudaMemcpyHostToDevice, stream_));
        ScoreListKernel<<<(n * 32 + 255) / 256, 256, 0, stream_>>>(dev_, L.table, L.dense, d, dc, n, ds);
        CountLaunch();
        ++launches_;
        HG_CUDA(cudaGetLastError());
        HG_CUDA(cudaMemcpyAsync(hs, ds, n * sizeof(int), cudaMemcpyDeviceToHost, stream_));
        HG_CUDA(cudaStreamSynchronize(stream_));
        for (int k = 0; k < n; ++k) scores[d][k] = ScoreFromSum(hs[k], sc.N);
      }
    }
    // generation index of the top ancestor in this problem's lattice (local scans)
    const auto top_index = [&](const WinnerRec& w) {
      const int ix = (w.ox + sp.wxy) >> max_depth_, iy = (w.oy + sp.wxy) >> max_depth_, iz = (w.oz + sp.wz) >> max_depth_;
      return static_cast<int64_t>(w.scan - scan_begin(p)) * per_scan_ + (iz * nxy_ + iy) * nxy_ + ix;
    };
    // Position of the distinct top ancestors in the sorted lowest-resolution list. Different scores:
    // descending. Equal scores: the permutation std::sort gives the whole list (introsort is not stable, but
    // it is a function of the comparison results alone), reproduced by sorting (score, index) keys of the
    // whole lattice in generation order with the same comparison.
    std::vector<int64_t> tops(n);
    for (int k = 0; k < n; ++k) tops[k] = top_index(tied[k]);
    bool equal_tops_differ = false;
    for (int a = 0; a < n && !equal_tops_differ; ++a)
      for (int b = a + 1; b < n; ++b)
        if (tops[a] != tops[b] && scores[max_depth_][a] == scores[max_depth_][b]) {
          equal_tops_differ = true;
          break;
        }
    std::vector<int64_t> position(n, 0);
    const bool whole_sweep = UnshardedScans(sc);
    if (equal_tops_differ && whole_sweep) {
      const size_t ln = static_cast<size_t>(sc.S) * per_scan_;
      std::vector<int> lattice(ln);
      HG_CUDA(cudaMemcpyAsync(lattice.data(), dev_.lattice + static_cast<size_t>(scan_begin(p)) * per_scan_,
                              ln * sizeof(int), cudaMemcpyDeviceToHost, stream_));
      HG_CUDA(cudaStreamSynchronize(stream_));
      struct Key {
        float score;
        uint32_t index;
      };
      std::vector<Key> keys(ln);
      for (size_t i = 0; i < ln; ++i) keys[i] = Key{ScoreFromSum(lattice[i], sc.N), static_cast<uint32_t>(i)};
      std::sort(keys.begin(), keys.end(), [](const Key& a, const Key& b) { return a.score > b.score; });
      std::vector<uint32_t> where(ln);
      for (size_t i = 0; i < ln; ++i) where[keys[i].index] = static_cast<uint32_t>(i);
      for (int k = 0; k < n; ++k) position[k] = where[tops[k]];
    } else {
      // Distinct scores order the tops. Equal scores of different tops inside ONE SHARD of a sweep: the
      // permutation of the whole list is not known here; the earlier generation index stands in, as in
      // hg_select_best_record.
      *order_unresolved = equal_tops_differ;
      std::vector<int> order(n);
      for (int k = 0; k < n; ++k) order[k] = k;
      std::sort(order.begin(), order.end(), [&](int a, int b) {
        if (scores[max_depth_][a] != scores[max_depth_][b]) return scores[max_depth_][a] > scores[max_depth_][b];
        return tops[a] < tops[b];
      });
      for (int r = 0; r < n; ++r) position[order[r]] = r;
      for (int r = 1; r < n; ++r)  // the same node gets the same position
        if (tops[order[r]] == tops[order[r - 1]]) position[order[r]] = position[order[r - 1]];
    }
    const auto slot_at = [&](const WinnerRec& w, int d) {  // generation slot of the depth-d ancestor among its siblings
      int o[3], q[3];
      ancestor(w, d, o);
      ancestor(w, d + 1, q);
      return ((o[2] != q[2]) ? 4 : 0) | ((o[1] != q[1]) ? 2 : 0) | ((o[0] != q[0]) ? 1 : 0);
    };
    const auto before = [&](int a, int b) {  // is winner a met before winner b?
      if (tops[a] != tops[b]) return position[a] < position[b];
      for (int d = max_depth_ - 1; d >= 0; --d) {
        int oa[3], ob[3];
        ancestor(tied[a], d, oa);
        ancestor(tied[b], d, ob);
        if (oa[0] == ob[0] && oa[1] == ob[1] && oa[2] == ob[2]) continue;
        if (scores[d][a] != scores[d][b]) return scores[d][a] > scores[d][b];
        return slot_at(tied[a], d) < slot_at(tied[b], d);
      }
      return false;
    };
    int best = 0;
    for (int k = 1; k < n; ++k)
      if (before(k, best)) best = k;
    if (TraceEnabled()) {
      for (int k = 0; k < n; ++k) {
        std::fprintf(stderr, "[hg trace] tied winner %d: scan %d offset (%d, %d, %d) top %lld position %lld scores", k,
                     tied[k].scan - scan_begin(p), tied[k].ox, tied[k].oy, tied[k].oz, static_cast<long long>(tops[k]),
                     static_cast<long long>(position[k]));
        for (int d = max_depth_; d >= 0; --d) std::fprintf(stderr, " %.6f", scores[d][k]);
        std::fprintf(stderr, "%s\n", k == best ? "  <- first" : "");
      }
    }
    *first = best;
    return HG_OK;
  }
  // The whole yaw sweep is in this problem (not a shard of it): the global scan numbering is 0 .. S-1.
  static bool UnshardedScans(const ScanSet& sc) {
    for (int k = 0; k < sc.S; ++k)
      if (sc.global_scan_index[k] != k) return false;
    return true;
  }

  // GetPoseFromCandidate (:400-406).
  Rigid PoseOf(int p, const Outcome& o) const {
    const Rigid translation{V3{m_->resolution * static_cast<float>(o.ox), m_->resolution * static_cast<float>(o.oy),
                               m_->resolution * static_cast<float>(o.oz)},
                            Q4{1.f, 0.f, 0.f, 0.f}};
    return fgeom::RigidProduct(translation, problems_[p].scans.poses[o.scan], m_->options.float_flavor);
  }

 private:
  static constexpr int kListCapacity = 1 << 20;
  static constexpr int kDiveStartsPerProblem = 32;
  static constexpr int kDiveStartsSingle = 256;
  struct Problem {
    SearchParameters sp;
    const hg_node_data* data = nullptr;
    float min_score = 0.f;
    ScanSet scans;
    size_t point_offset = 0;
  };
  int scan_begin(int p) const {
    int b = 0;
    for (int k = 0; k < p; ++k) b += problems_[k].scans.S;
    return b;
  }
  GateArgs GateGrid() const { return GateArgs{m_->low_grid->dev, m_->resolution}; }
  static int PersistentBlocks(int per_sm) {
    static const int sms = [] {
      int dev = 0, n = 148;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
      return n;
    }();
    return sms * per_sm;
  }

  const hg_fcsm* m_;
  cudaStream_t stream_;
  Workspace* ws_;
  std::vector<Problem> problems_;
  SearchDev dev_;
  int max_depth_ = 0, step_ = 1, nxy_ = 0, nz_ = 0, per_scan_ = 0, total_scans_ = 0;
  int list_cap_ = 0, winners_cap_ = 0;
  size_t ctrl_bytes_ = 0, off_pp_ = 0, off_scored_ = 0, off_win_ = 0;
  cudaEvent_t e0_ = nullptr, e1_ = nullptr;
  int* exchange_ = nullptr;
  // HG_TRACE: events between the phases of a search, printed by RunPhaseB.
  std::vector<std::pair<const char*, cudaEvent_t>> marks_;
  void Mark(const char* label) {
    if (!TraceEnabled()) return;
    cudaEvent_t e = nullptr;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, stream_);
    marks_.emplace_back(label, e);
  }
  void PrintMarks() {
    if (marks_.size() > 1) {
      std::fprintf(stderr, "[hg trace] device phases (us):");
      for (size_t i = 1; i < marks_.size(); ++i) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, marks_[i - 1].second, marks_[i].second);
        std::fprintf(stderr, " %s %.0f", marks_[i].first, 1e3f * ms);
      }
      float total = 0.f;
      cudaEventElapsedTime(&total, marks_.front().second, marks_.back().second);
      std::fprintf(stderr, " | first to last mark %.0f (includes the host's wait between the phases)\n", 1e3f * total);
    }
    for (auto& m : marks_) cudaEventDestroy(m.second);
    marks_.clear();
  }
  float lattice_ms_ = 0.f;
  int launches_ = 0;
  bool phase_a_done_ = false;
};

void FillResult(const DeviceSearch& search, int p, const DeviceSearch::Outcome& o, hg_match_result* result) {
  const ScanSet& scans = search.scans(p);
  std::memset(result, 0, sizeof(*result));
  result->pose_rotation[0] = 1.0;
  result->num_discrete_scans = scans.S;
  result->num_candidates_scored = o.scored;
  result->num_candidates_visited = o.scored;
  result->num_score_launches = search.launches();
  result->num_lowest_resolution_candidates = search.lattice_candidates(p);
  result->lowest_resolution_kernel_ms = search.lattice_ms();
  result->dive_bound = o.local_bound;
  if (!o.found) return;
  const Rigid pose = search.PoseOf(p, o);
  result->found = 1;
  result->score = o.score;
  result->pose_translation[0] = pose.t.x;
  result->pose_translation[1] = pose.t.y;
  result->pose_translation[2] = pose.t.z;
  result->pose_rotation[0] = pose.q.w;
  result->pose_rotation[1] = pose.q.x;
  result->pose_rotation[2] = pose.q.y;
  result->pose_rotation[3] = pose.q.z;
  result->rotational_score = scans.rotational_scores[o.scan];
  result->low_resolution_score = o.low;
  result->best_scan_index = scans.global_scan_index[o.scan];
  result->best_offset[0] = o.ox;
  result->best_offset[1] = o.oy;
  result->best_offset[2] = o.oz;
  result->best_ancestor_score = o.ancestor_score;
  result->best_ancestor_index = o.ancestor_index;
  result->order_unresolved = o.order_unresolved ? 1 : 0;
}

// HG_SEARCH_HOST_REPLAY=1 sends every search through the host replay (tests compare the two).
bool ForceHostReplay() { return std::getenv("HG_SEARCH_HOST_REPLAY") != nullptr; }

// MatchWithSearchParameters (:202-228): device-resident search; strict visiting order, exact ties
// and the other cases DeviceSearch hands back go through the host replay.
int MatchWithSearchParameters(const hg_fcsm* m, const SearchParameters& sp, const Rigid& node_pose,
                              const Rigid& submap_pose, const hg_node_data* data, float min_score,
                              int shard_index, int num_shards, hg_match_result* result) {
  if (m->options.strict_visit_order != 0 || ForceHostReplay()) {
    return HostReplayMatch(m, sp, node_pose, submap_pose, data, min_score, shard_index, num_shards, result);
  }
  HG_ON_DEVICE(m->device);
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "hg_fcsm: cannot create a stream");
  PhaseTimer pt;
  DeviceSearch search(m, stream);
  int rc = search.AddProblem(sp, node_pose, submap_pose, data, min_score, shard_index, num_shards);
  if (rc != HG_OK) return rc;
  const bool time_lattice = static_cast<int64_t>(search.scans(0).S) * search.scans(0).N >= (int64_t(1) << 20);
  rc = search.RunPhaseA(time_lattice);
  if (rc != HG_OK) return rc;
  std::vector<DeviceSearch::Outcome> outcomes;
  rc = search.RunPhaseB(0.f, &outcomes);
  if (rc != HG_OK) return rc;
  if (TraceEnabled()) std::fprintf(stderr, "[hg trace] device search: %.2f ms\n", pt.Lap());
  if (outcomes[0].needs_host_replay) {
    if (TraceEnabled()) std::fprintf(stderr, "[hg trace] device search handed the problem to the host replay\n");
    return HostReplayMatch(m, sp, node_pose, submap_pose, data, min_score, shard_index, num_shards, result);
  }
  FillResult(search, 0, outcomes[0], result);
  return HG_OK;
}

int CheckNodeData(const hg_node_data* d) {
  HG_REQUIRE(d != nullptr, "null node data");
  HG_REQUIRE(d->num_high_resolution_points > 0 && d->high_resolution_point_cloud != nullptr,
             "node data: empty high resolution point cloud");
  HG_REQUIRE(d->num_low_resolution_points >= 0, "node data: negative low resolution point count");
  HG_REQUIRE(d->num_low_resolution_points == 0 || d->low_resolution_point_cloud != nullptr,
             "node data: null low resolution point cloud");
  return HG_OK;
}

// MatchFullSubmap's window (:181-192).
SearchParameters FullSubmapWindow(const hg_fcsm* m, const hg_node_data* data) {
  float max_point_distance = 0.f;
  const float* pts = data->high_resolution_point_cloud;
  for (int i = 0; i < data->num_high_resolution_points; ++i) {
    max_point_distance = std::max(max_point_distance, fgeom::Norm3(V3{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}));
  }
  const int linear_window_size =
      (m->width_in_voxels + 1) / 2 + static_cast<int>(std::lround(max_point_distance / m->resolution + 0.5f));
  return SearchParameters{linear_window_size, linear_window_size, M_PI};
}

SearchParameters LocalWindow(const hg_fcsm* m) {
  // :163-166 RoundToInt(double window / float resolution)
  return SearchParameters{
      static_cast<int>(std::lround(m->options.linear_xy_search_window / m->resolution)),
      static_cast<int>(std::lround(m->options.linear_z_search_window / m->resolution)),
      m->options.angular_search_window};
}

}  // namespace

extern "C" {

int hg_fcsm_create(const hg_grid* grid, const hg_grid* low_resolution_grid,
                   const float* rotational_scan_matcher_histogram, int histogram_size,
                   const hg_fcsm_options* options, hg_fcsm** out) {
  HG_REQUIRE(grid && low_resolution_grid && options && out, "hg_fcsm_create: bad arguments");
  HG_REQUIRE(histogram_size >= 0 && (histogram_size == 0 || rotational_scan_matcher_histogram != nullptr),
             "hg_fcsm_create: bad histogram");
  std::unique_ptr<hg_fcsm> m(new hg_fcsm);
  m->device = grid->device;
  m->options = *options;
  m->grid = grid;
  m->low_grid = low_resolution_grid;
  m->resolution = grid->dev.resolution;
  m->width_in_voxels = grid->grid_size;
  if (histogram_size > 0)
    m->histogram.assign(rotational_scan_matcher_histogram, rotational_scan_matcher_histogram + histogram_size);
  const int rc = hg_stack_build(grid, options->branch_and_bound_depth, options->full_resolution_depth, &m->stack);
  if (rc != HG_OK) return rc;
  {
    std::vector<LevelDev> levels;
    for (const hg_stack_level& L : m->stack->levels) levels.push_back(LevelDev{L.table, L.dense});
    HG_CUDA(cudaMalloc(&m->d_levels, levels.size() * sizeof(LevelDev)));
    HG_CUDA(cudaMemcpy(m->d_levels, levels.data(), levels.size() * sizeof(LevelDev), cudaMemcpyHostToDevice));
  }
  // Residue rows of the lowest-resolution level for the lattice kernel.
  const int depth = static_cast<int>(m->stack->levels.size()) - 1;
  const hg_stack_level& L = m->stack->levels[depth];
  if (depth >= 1 && L.dense.data != nullptr) {
    const int e = std::max(0, depth - options->full_resolution_depth + 1);
    LatticeRows& R = m->lattice_rows;
    R.cs_log2 = depth - e;
    for (int k = 0; k < 3; ++k) {
      R.min[k] = L.dense.min[k];
      R.dim[k] = L.dense.dim[k];
    }
    R.payload = (R.dim[0] + (1 << R.cs_log2) - 1) >> R.cs_log2;
    R.row_words = (kLatticePadLeft + R.payload + kLatticePadRight + 3) / 4;
    const int64_t total = static_cast<int64_t>(R.dim[2]) * R.dim[1] * (int64_t(1) << R.cs_log2) * R.row_words * 4;
    if (total <= (int64_t(256) << 20)) {
      if (cudaMalloc(&m->d_lattice_rows, static_cast<size_t>(total)) != cudaSuccess) {
        cudaGetLastError();  // optional copy: the brick / dense kernels still work
        m->d_lattice_rows = nullptr;
      } else {
        BuildLatticeRowsKernel<<<static_cast<unsigned>((total + 255) / 256), 256>>>(
            L.dense, R, total, static_cast<unsigned char*>(m->d_lattice_rows));
        CountLaunch();
        HG_CUDA(cudaGetLastError());
        HG_CUDA(cudaDeviceSynchronize());
        R.data = static_cast<const unsigned*>(m->d_lattice_rows);
      }
    }
  }
  *out = m.release();
  return HG_OK;
}

void hg_fcsm_destroy(hg_fcsm* m) {
  if (m == nullptr) return;
  if (m->d_lattice_rows) cudaFree(m->d_lattice_rows);
  if (m->d_levels) cudaFree(m->d_levels);
  hg_stack_destroy(m->stack);
  delete m;
}

const hg_stack* hg_fcsm_stack(const hg_fcsm* m) { return m ? m->stack : nullptr; }

int hg_fcsm_match(const hg_fcsm* m, const double* global_node_pose, const double* global_submap_pose,
                  const hg_node_data* data, float min_score, hg_match_result* result) {
  HG_REQUIRE(m && global_node_pose && global_submap_pose && result, "hg_fcsm_match: bad arguments");
  const int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  return MatchWithSearchParameters(m, LocalWindow(m), ToRigidF(global_node_pose, global_node_pose + 3),
                                   ToRigidF(global_submap_pose, global_submap_pose + 3), data, min_score, 0, 1,
                                   result);
}

int hg_fcsm_match_full_submap(const hg_fcsm* m, const double* global_node_rotation,
                              const double* global_submap_rotation, const hg_node_data* data, float min_score,
                              int shard_index, int num_shards, hg_match_result* result) {
  HG_REQUIRE(m && global_node_rotation && global_submap_rotation && result,
             "hg_fcsm_match_full_submap: bad arguments");
  HG_REQUIRE(num_shards >= 1 && shard_index >= 0 && shard_index < num_shards,
             "hg_fcsm_match_full_submap: bad shard");
  const int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  const SearchParameters sp = FullSubmapWindow(m, data);
  const double zero[3] = {0., 0., 0.};
  return MatchWithSearchParameters(m, sp, ToRigidF(zero, global_node_rotation),
                                   ToRigidF(zero, global_submap_rotation), data, min_score, shard_index,
                                   num_shards, result);
}

int hg_fcsm_match_batch(const hg_fcsm* m, int num_problems, const double* global_node_poses,
                        const double* global_submap_poses, const hg_node_data* data, float min_score,
                        hg_match_result* results) {
  HG_REQUIRE(m && num_problems >= 0 && (num_problems == 0 || (global_node_poses && global_submap_poses && data && results)),
             "hg_fcsm_match_batch: bad arguments");
  for (int p = 0; p < num_problems; ++p) {
    const int rc = CheckNodeData(&data[p]);
    if (rc != HG_OK) return rc;
  }
  const SearchParameters sp = LocalWindow(m);
  if (m->options.strict_visit_order != 0 || ForceHostReplay()) {
    for (int p = 0; p < num_problems; ++p) {
      const int rc = HostReplayMatch(m, sp, ToRigidF(global_node_poses + 7 * p, global_node_poses + 7 * p + 3),
                                     ToRigidF(global_submap_poses + 7 * p, global_submap_poses + 7 * p + 3), &data[p],
                                     min_score, 0, 1, &results[p]);
      if (rc != HG_OK) return rc;
    }
    return HG_OK;
  }
  HG_ON_DEVICE(m->device);
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "hg_fcsm: cannot create a stream");
  // Sub-batches bounded by the memory of the discretised scans (12 bytes per scan point and level).
  constexpr int64_t kMaxCellInts = int64_t(1) << 28;  // 1 GiB
  int begin = 0;
  while (begin < num_problems) {
    PhaseTimer pt;
    DeviceSearch search(m, stream);
    int64_t cell_ints = 0;
    int end = begin;
    while (end < num_problems && end - begin < 1024) {
      int rc = search.AddProblem(sp, ToRigidF(global_node_poses + 7 * end, global_node_poses + 7 * end + 3),
                                 ToRigidF(global_submap_poses + 7 * end, global_submap_poses + 7 * end + 3), &data[end],
                                 min_score, 0, 1);
      if (rc != HG_OK) return rc;
      const ScanSet& sc = search.scans(end - begin);
      cell_ints += static_cast<int64_t>(1 + sc.num_low_levels) * sc.S * sc.N * 3;
      ++end;
      if (cell_ints >= kMaxCellInts) break;
    }
    const double t_add = pt.Lap();
    int rc = search.RunPhaseA(false);
    if (rc != HG_OK) return rc;
    const double t_a = pt.Lap();
    std::vector<DeviceSearch::Outcome> outcomes;
    rc = search.RunPhaseB(0.f, &outcomes);
    if (rc != HG_OK) return rc;
    if (TraceEnabled())
      std::fprintf(stderr, "[hg trace] batch of %d: scan poses %.2f ms, phase A queued in %.2f ms, phase B + wait %.2f ms\n",
                   end - begin, t_add, t_a, pt.Lap());
    for (int p = begin; p < end; ++p) {
      if (outcomes[p - begin].needs_host_replay) {
        rc = HostReplayMatch(m, sp, ToRigidF(global_node_poses + 7 * p, global_node_poses + 7 * p + 3),
                             ToRigidF(global_submap_poses + 7 * p, global_submap_poses + 7 * p + 3), &data[p], min_score,
                             0, 1, &results[p]);
        if (rc != HG_OK) return rc;
      } else {
        FillResult(search, p - begin, outcomes[p - begin], &results[p]);
      }
    }
    begin = end;
  }
  return HG_OK;
}

}  // extern "C"

// A sharded MatchFullSubmap between its two phases.
struct hg_fcsm_search {
  const hg_fcsm* m = nullptr;
  SearchParameters sp{0, 0, 0.};
  Rigid node_pose, submap_pose;
  const hg_node_data* data = nullptr;
  float min_score = 0.f;
  int shard_index = 0, num_shards = 1;
  std::unique_ptr<DeviceSearch> search;  // null: the host replay does everything in 'finish'
};

extern "C" {

int hg_fcsm_full_submap_begin(const hg_fcsm* m, const double* global_node_rotation,
                              const double* global_submap_rotation, const hg_node_data* data, float min_score,
                              int shard_index, int num_shards, hg_fcsm_search** search_out, float* dive_bound) {
  HG_REQUIRE(m && global_node_rotation && global_submap_rotation && search_out && dive_bound,
             "hg_fcsm_full_submap_begin: bad arguments");
  HG_REQUIRE(num_shards >= 1 && shard_index >= 0 && shard_index < num_shards, "hg_fcsm_full_submap_begin: bad shard");
  int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  *dive_bound = 0.f;
  std::unique_ptr<hg_fcsm_search> h(new hg_fcsm_search);
  h->m = m;
  h->sp = FullSubmapWindow(m, data);
  const double zero[3] = {0., 0., 0.};
  h->node_pose = ToRigidF(zero, global_node_rotation);
  h->submap_pose = ToRigidF(zero, global_submap_rotation);
  h->data = data;
  h->min_score = min_score;
  h->shard_index = shard_index;
  h->num_shards = num_shards;
  if (m->options.strict_visit_order == 0 && !ForceHostReplay()) {
    HG_ON_DEVICE(m->device);
    cudaStream_t stream = ThreadStream();
    HG_REQUIRE(stream != nullptr, "hg_fcsm: cannot create a stream");
    h->search.reset(new DeviceSearch(m, stream));
    rc = h->search->AddProblem(h->sp, h->node_pose, h->submap_pose, data, min_score, shard_index, num_shards);
    if (rc != HG_OK) return rc;
    rc = h->search->RunPhaseA(true);
    if (rc != HG_OK) return rc;
    std::vector<float> bounds;
    rc = h->search->ReadBounds(&bounds);
    if (rc != HG_OK) return rc;
    *dive_bound = bounds[0];
  }
  *search_out = h.release();
  return HG_OK;
}

int hg_fcsm_full_submap_begin_async(const hg_fcsm* m, const double* global_node_rotation,
                                    const double* global_submap_rotation, const hg_node_data* data, float min_score,
                                    int shard_index, int num_shards, hg_fcsm_search** search_out, float** d_bound,
                                    void** stream_out) {
  HG_REQUIRE(m && global_node_rotation && global_submap_rotation && search_out && d_bound && stream_out,
             "hg_fcsm_full_submap_begin_async: bad arguments");
  HG_REQUIRE(num_shards >= 1 && shard_index >= 0 && shard_index < num_shards,
             "hg_fcsm_full_submap_begin_async: bad shard");
  int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  *d_bound = nullptr;
  *stream_out = nullptr;
  std::unique_ptr<hg_fcsm_search> h(new hg_fcsm_search);
  h->m = m;
  h->sp = FullSubmapWindow(m, data);
  const double zero[3] = {0., 0., 0.};
  h->node_pose = ToRigidF(zero, global_node_rotation);
  h->submap_pose = ToRigidF(zero, global_submap_rotation);
  h->data = data;
  h->min_score = min_score;
  h->shard_index = shard_index;
  h->num_shards = num_shards;
  if (m->options.strict_visit_order == 0 && !ForceHostReplay()) {
    HG_ON_DEVICE(m->device);
    cudaStream_t stream = ThreadStream();
    HG_REQUIRE(stream != nullptr, "hg_fcsm: cannot create a stream");
    h->search.reset(new DeviceSearch(m, stream));
    rc = h->search->AddProblem(h->sp, h->node_pose, h->submap_pose, data, min_score, shard_index, num_shards);
    if (rc != HG_OK) return rc;
    rc = h->search->RunPhaseA(true);
    if (rc != HG_OK) return rc;
    rc = h->search->OfferBounds(d_bound);
    if (rc != HG_OK) return rc;
    *stream_out = static_cast<void*>(stream);
  }
  *search_out = h.release();
  return HG_OK;
}

int hg_fcsm_full_submap_finish(hg_fcsm_search* search, float bound, hg_match_result* result) {
  HG_REQUIRE(search != nullptr && result != nullptr, "hg_fcsm_full_submap_finish: bad arguments");
  std::unique_ptr<hg_fcsm_search> h(search);
  HG_ON_DEVICE(h->m->device);
  if (h->search) {
    std::vector<DeviceSearch::Outcome> outcomes;
    const int rc = h->search->RunPhaseB(bound, &outcomes);
    if (rc != HG_OK) return rc;
    if (!outcomes[0].needs_host_replay) {
      FillResult(*h->search, 0, outcomes[0], result);
      return HG_OK;
    }
  }
  return HostReplayMatch(h->m, h->sp, h->node_pose, h->submap_pose, h->data, h->min_score, h->shard_index,
                         h->num_shards, result);
}

void hg_fcsm_search_abandon(hg_fcsm_search* search) { delete search; }

// ---- best-record exchange between the shards of one search --------------------------------
// Layout of a record (HG_BEST_RECORD_DOUBLES doubles): 0 found, 1 score, 2 ancestor score,
// 3 ancestor generation index, 4 scan index, 5..7 offset, 8..14 pose (t3 q4), 15 rotational score,
// 16 low-resolution score, 17 candidates scored, 18 order unresolved inside the shard, 19 reserved.
int hg_pack_best_record(const hg_match_result* r, double* record) {
  HG_REQUIRE(r != nullptr && record != nullptr, "hg_pack_best_record: bad arguments");
  for (int k = 0; k < HG_BEST_RECORD_DOUBLES; ++k) record[k] = 0.0;
  record[0] = r->found ? 1.0 : 0.0;
  record[1] = r->score;
  record[2] = r->best_ancestor_score;
  record[3] = static_cast<double>(r->best_ancestor_index);
  record[4] = r->best_scan_index;
  for (int k = 0; k < 3; ++k) record[5 + k] = r->best_offset[k];
  for (int k = 0; k < 3; ++k) record[8 + k] = r->pose_translation[k];
  for (int k = 0; k < 4; ++k) record[11 + k] = r->pose_rotation[k];
  record[15] = r->rotational_score;
  record[16] = r->low_resolution_score;
  record[17] = static_cast<double>(r->num_candidates_scored);
  record[18] = r->order_unresolved;
  return HG_OK;
}

int hg_unpack_best_record(const double* record, hg_match_result* r) {
  HG_REQUIRE(r != nullptr && record != nullptr, "hg_unpack_best_record: bad arguments");
  std::memset(r, 0, sizeof(*r));
  r->found = record[0] > 0.0 ? 1 : 0;
  r->score = static_cast<float>(record[1]);
  r->best_ancestor_score = static_cast<float>(record[2]);
  r->best_ancestor_index = static_cast<int64_t>(record[3]);
  r->best_scan_index = static_cast<int32_t>(record[4]);
  for (int k = 0; k < 3; ++k) r->best_offset[k] = static_cast<int32_t>(record[5 + k]);
  for (int k = 0; k < 3; ++k) r->pose_translation[k] = record[8 + k];
  for (int k = 0; k < 4; ++k) r->pose_rotation[k] = record[11 + k];
  r->rotational_score = static_cast<float>(record[15]);
  r->low_resolution_score = static_cast<float>(record[16]);
  r->num_candidates_scored = static_cast<int64_t>(record[17]);
  r->order_unresolved = static_cast<int32_t>(record[18]);
  return HG_OK;
}

// The reference's search keeps the first best leaf it visits (std::max keeps the earlier of equal
// scores, :465-466) and visits the lowest-resolution candidates in descending order of their
// score (:384, :437). Across shards that is: the higher leaf score; on equal leaf scores the leaf
// whose lowest-resolution ancestor scores higher; ancestors of equal score are ordered by the
// reference's std::sort, which no shard can reproduce alone — the earlier generation index is
// taken and *order_unresolved is set so that the caller can rerun unsharded in strict order.
int hg_select_best_record(const double* records, int num_records, int* winner, int* order_unresolved) {
  HG_REQUIRE(records != nullptr && num_records >= 0 && winner != nullptr, "hg_select_best_record: bad arguments");
  *winner = -1;
  if (order_unresolved) *order_unresolved = 0;
  for (int k = 0; k < num_records; ++k) {
    const double* r = records + static_cast<size_t>(k) * HG_BEST_RECORD_DOUBLES;
    if (!(r[0] > 0.0)) continue;
    if (*winner < 0) {
      *winner = k;
      continue;
    }
    const double* b = records + static_cast<size_t>(*winner) * HG_BEST_RECORD_DOUBLES;
    if (r[1] != b[1]) {
      if (r[1] > b[1]) {
        *winner = k;
        if (order_unresolved) *order_unresolved = 0;
      }
      continue;
    }
    if (r[2] != b[2]) {
      if (r[2] > b[2]) {
        *winner = k;
        if (order_unresolved) *order_unresolved = 0;
      }
      continue;
    }
    if (order_unresolved) *order_unresolved = 1;
    if (r[3] < b[3]) *winner = k;
  }
  // a shard that could not order its own tied leaves passes that on
  if (*winner >= 0 && order_unresolved && records[static_cast<size_t>(*winner) * HG_BEST_RECORD_DOUBLES + 18] > 0.0)
    *order_unresolved = 1;
  return HG_OK;
}

int hg_fcsm_discrete_scans(const hg_fcsm* m, int wxy, int wz, double angular_search_window,
                           const double* global_node_pose, const double* global_submap_pose,
                           const hg_node_data* data, int max_scans, int32_t* num_scans, float* poses,
                           float* rotational_scores, int32_t* cells) {
  HG_REQUIRE(m && global_node_pose && global_submap_pose && num_scans, "hg_fcsm_discrete_scans: bad arguments");
  int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  HG_ON_DEVICE(m->device);
  ScanSet scans;
  const SearchParameters sp{wxy, wz, angular_search_window};
  rc = GenerateDiscreteScans(m, sp, data, ToRigidF(global_node_pose, global_node_pose + 3),
                             ToRigidF(global_submap_pose, global_submap_pose + 3), 0, 1, nullptr, &scans);
  if (rc != HG_OK) return rc;
  *num_scans = scans.S;
  if (scans.S > max_scans || scans.S == 0) return HG_OK;
  for (int s = 0; s < scans.S; ++s) {
    if (poses) {
      const Rigid& r = scans.poses[s];
      float* o = poses + 7 * s;
      o[0] = r.t.x; o[1] = r.t.y; o[2] = r.t.z; o[3] = r.q.w; o[4] = r.q.x; o[5] = r.q.y; o[6] = r.q.z;
    }
    if (rotational_scores) rotational_scores[s] = scans.rotational_scores[s];
  }
  if (cells) {
    const int D = m->options.branch_and_bound_depth;
    const size_t per = static_cast<size_t>(scans.N) * 3;
    for (int d = 0; d < D; ++d) {
      const int* src = scans.CellsOfDepth(d);
      for (int s = 0; s < scans.S; ++s) {
        HG_CUDA(cudaMemcpy(cells + (static_cast<size_t>(s) * D + d) * per, src + static_cast<size_t>(s) * per,
                           per * sizeof(int), cudaMemcpyDeviceToHost));
      }
    }
  }
  return HG_OK;
}

int hg_fcsm_score_candidates(const hg_fcsm* m, int wxy, int wz, double angular_search_window,
                             const double* global_node_pose, const double* global_submap_pose,
                             const hg_node_data* data, int depth, int64_t num_candidates,
                             const int32_t* candidates, int32_t* sums, float* scores, float* kernel_ms) {
  HG_REQUIRE(m && global_node_pose && global_submap_pose && candidates && num_candidates >= 0,
             "hg_fcsm_score_candidates: bad arguments");
  HG_REQUIRE(depth >= 0 && depth < m->options.branch_and_bound_depth, "hg_fcsm_score_candidates: bad depth");
  int rc = CheckNodeData(data);
  if (rc != HG_OK) return rc;
  HG_ON_DEVICE(m->device);
  ScanSet scans;
  const SearchParameters sp{wxy, wz, angular_search_window};
  rc = GenerateDiscreteScans(m, sp, data, ToRigidF(global_node_pose, global_node_pose + 3),
                             ToRigidF(global_submap_pose, global_submap_pose + 3), 0, 1, nullptr, &scans);
  if (rc != HG_OK) return rc;
  for (int64_t i = 0; i < num_candidates; ++i) {
    HG_REQUIRE(candidates[4 * i] >= 0 && candidates[4 * i] < scans.S, "hg_fcsm_score_candidates: scan index out of range");
  }
  Scorer scorer;
  scorer.m = m;
  scorer.scans = &scans;
  scorer.stream = nullptr;
  scorer.ws = &GetWorkspace(m->device);
  rc = scorer.Reserve(static_cast<size_t>(num_candidates));
  if (rc != HG_OK) return rc;
  std::memcpy(scorer.HostCandidates(), candidates, static_cast<size_t>(num_candidates) * sizeof(int4));
  if (kernel_ms) *kernel_ms = 0.f;
  rc = scorer.Run(depth, static_cast<size_t>(num_candidates), kernel_ms);
  if (rc != HG_OK) return rc;
  const int* hs = scorer.HostSums();
  for (int64_t i = 0; i < num_candidates; ++i) {
    if (sums) sums[i] = hs[i];
    if (scores) scores[i] = ScoreFromSum(hs[i], scans.N);
  }
  return HG_OK;
}

int hg_low_resolution_scores(const hg_grid* low_resolution_grid, const float* points, int num_points,
                             int64_t num_poses, const float* poses, float* scores) {
  HG_REQUIRE(low_resolution_grid && points && num_points > 0 && num_poses >= 0 && poses && scores,
             "hg_low_resolution_scores: bad arguments");
  if (num_poses == 0) return HG_OK;
  DeviceBuffer d_points, d_poses, d_scores;
  int rc = d_points.Alloc(sizeof(float) * 3 * num_points);
  if (rc != HG_OK) return rc;
  rc = d_poses.Alloc(sizeof(float) * 7 * num_poses);
  if (rc != HG_OK) return rc;
  rc = d_scores.Alloc(sizeof(float) * num_poses);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemcpy(d_points.p, points, sizeof(float) * 3 * num_points, cudaMemcpyHostToDevice));
  HG_CUDA(cudaMemcpy(d_poses.p, poses, sizeof(float) * 7 * num_poses, cudaMemcpyHostToDevice));
  LowResolutionKernel<<<static_cast<unsigned>((num_poses + 3) / 4), 128>>>(
      low_resolution_grid->dev, d_points.as<float>(), num_points, d_poses.as<float>(), num_poses,
      d_scores.as<float>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(scores, d_scores.p, sizeof(float) * num_poses, cudaMemcpyDeviceToHost));
  return HG_OK;
}

}  // extern "C"
