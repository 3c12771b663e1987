// Voxel filters on the device (the step right before the matcher).
//
// Reference behaviour reproduced (bit-exact output: the same points, in input order):
//   sensor/internal/voxel_filter.cc:25-73       VoxelFilter::Filter — the first point of every voxel,
//                                               GetCellIndex = lround(p / size), 96-bit key
//   sensor/internal/adaptive_voxel_filter.h:34-84  FilterByMaxRange + AdaptivelyVoxelFiltered
//
// "First point per voxel in input order" is the minimum point index per voxel:
//   1. every point computes its cell (exact float division semantics, CellIndexFast);
//   2. points claim slots of an open-addressing hash set. A slot's key is the cell
//      of the point whose index sits in 'owner', so a key is never read half
//      written: other points compare against cells[owner];
//   3. atomicMin of the point index per slot; a point is kept iff it is its voxel's minimum;
//   4. order-preserving compaction of the kept flags (count per chunk, scan, write).
#include <algorithm>
#include <cmath>
#include <map>
#include <memory>

#include "device_math.cuh"

namespace hg {
namespace {

struct VfBuffer {
  void* p = nullptr;
  size_t cap = 0;
  ~VfBuffer() {
    if (p) cudaFree(p);
  }
  int Reserve(size_t bytes) {
    if (bytes <= cap) return HG_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t want = bytes + bytes / 2 + 4096;
    HG_CUDA(cudaMalloc(&p, want));
    cap = want;
    return HG_OK;
  }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

struct VfWorkspace {
  VfBuffer d_points, d_cells, d_active, d_owner, d_min, d_slot, d_flags, d_chunks, d_kept;
};
VfWorkspace& GetVfWorkspace(int device) {
  thread_local std::map<int, std::unique_ptr<VfWorkspace>> spaces;
  std::unique_ptr<VfWorkspace>& w = spaces[device];
  if (!w) w.reset(new VfWorkspace);
  return *w;
}

__device__ __forceinline__ unsigned HashCell(const int x, const int y, const int z) {
  unsigned h = static_cast<unsigned>(x) * 73856093u ^ static_cast<unsigned>(y) * 19349663u ^
               static_cast<unsigned>(z) * 83492791u;
  h ^= h >> 15;
  h *= 0x2c1b3c6du;
  h ^= h >> 13;
  return h;
}

// FilterByMaxRange (adaptive_voxel_filter.h:34-43): norm() <= max_range, Eigen's
// unrolled 3-vector reduction c0 + (c1 + c2), unfused.
__global__ void __launch_bounds__(256) RangeKernel(const float* __restrict__ pts, const int stride, const int n,
                                                   const float max_range, int* __restrict__ active) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float x = pts[static_cast<size_t>(i) * stride], y = pts[static_cast<size_t>(i) * stride + 1],
              z = pts[static_cast<size_t>(i) * stride + 2];
  const float sq = __fadd_rn(__fmul_rn(x, x), __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z)));
  active[i] = __fsqrt_rn(sq) <= max_range ? 1 : 0;
}

__global__ void __launch_bounds__(256) CellsKernel(const float* __restrict__ pts, const int stride, const int n,
                                                   const float size, const float inv_size,
                                                   int* __restrict__ cells) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = pts + static_cast<size_t>(i) * stride;
  cells[3 * i] = CellIndexFast(p[0], size, inv_size);
  cells[3 * i + 1] = CellIndexFast(p[1], size, inv_size);
  cells[3 * i + 2] = CellIndexFast(p[2], size, inv_size);
}

__global__ void __launch_bounds__(256) InsertKernel(const int* __restrict__ cells, const int* __restrict__ active,
                                                    const int n, const unsigned mask, int* __restrict__ owner,
                                                    int* __restrict__ min_index, int* __restrict__ slot_of) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (active != nullptr && !active[i]) {
    slot_of[i] = -1;
    return;
  }
  const int x = cells[3 * i], y = cells[3 * i + 1], z = cells[3 * i + 2];
  unsigned h = HashCell(x, y, z) & mask;
  for (;;) {
    const int o = atomicCAS(owner + h, -1, i);
    if (o == -1 || (cells[3 * o] == x && cells[3 * o + 1] == y && cells[3 * o + 2] == z)) break;
    h = (h + 1) & mask;
  }
  atomicMin(min_index + h, i);
  slot_of[i] = static_cast<int>(h);
}

__global__ void __launch_bounds__(256) KeepKernel(const int* __restrict__ slot_of, const int* __restrict__ min_index,
                                                  const int n, int* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = slot_of[i];
  flags[i] = (s >= 0 && min_index[s] == i) ? 1 : 0;
}

constexpr int kVfChunk = 1024;
__global__ void __launch_bounds__(256) VfCountKernel(const int* __restrict__ flags, const int n,
                                                     int* __restrict__ chunk_counts) {
  const int base = blockIdx.x * kVfChunk;
  int c = 0;
  for (int k = threadIdx.x; k < kVfChunk; k += 256) c += (base + k < n) ? flags[base + k] : 0;
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
__global__ void __launch_bounds__(1024) VfScanKernel(int* __restrict__ chunk_counts, const int num_chunks) {
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
__global__ void __launch_bounds__(256) VfWriteKernel(const int* __restrict__ flags, const int n,
                                                     const int* __restrict__ chunk_offsets,
                                                     int* __restrict__ kept) {
  __shared__ int warp_counts[32];
  const int base = blockIdx.x * kVfChunk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int my_flag[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int seg = r * 8 + warp;
    const int i = base + seg * 32 + lane;
    my_flag[r] = (i < n) ? flags[i] : 0;
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
    if (my_flag[r]) kept[out_base + before + __popc(b & ((1u << lane) - 1u))] = base + seg * 32 + lane;
  }
}

// One VoxelFilter(size).Filter pass over the points already on the device.
// Leaves the kept indices in ws.d_kept and returns their number.
int FilterPass(VfWorkspace& ws, const int stride, const int n, const float size, const int* d_active, int* kept) {
  const int blocks = (n + 255) / 256;
  unsigned cap = 64;
  while (cap < 2u * static_cast<unsigned>(n)) cap <<= 1;
  int rc = ws.d_cells.Reserve(sizeof(int) * 3 * n);
  if (rc != HG_OK) return rc;
  rc = ws.d_owner.Reserve(sizeof(int) * cap);
  if (rc != HG_OK) return rc;
  rc = ws.d_min.Reserve(sizeof(int) * cap);
  if (rc != HG_OK) return rc;
  rc = ws.d_slot.Reserve(sizeof(int) * n);
  if (rc != HG_OK) return rc;
  rc = ws.d_flags.Reserve(sizeof(int) * n);
  if (rc != HG_OK) return rc;
  const int chunks = (n + kVfChunk - 1) / kVfChunk;
  rc = ws.d_chunks.Reserve(sizeof(int) * (chunks + 1));
  if (rc != HG_OK) return rc;
  rc = ws.d_kept.Reserve(sizeof(int) * n);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemsetAsync(ws.d_owner.p, 0xff, sizeof(int) * cap));
  HG_CUDA(cudaMemsetAsync(ws.d_min.p, 0x7f, sizeof(int) * cap));
  CellsKernel<<<blocks, 256>>>(ws.d_points.as<float>(), stride, n, size, 1.f / size, ws.d_cells.as<int>());
  InsertKernel<<<blocks, 256>>>(ws.d_cells.as<int>(), d_active, n, cap - 1, ws.d_owner.as<int>(),
                                ws.d_min.as<int>(), ws.d_slot.as<int>());
  KeepKernel<<<blocks, 256>>>(ws.d_slot.as<int>(), ws.d_min.as<int>(), n, ws.d_flags.as<int>());
  VfCountKernel<<<chunks, 256>>>(ws.d_flags.as<int>(), n, ws.d_chunks.as<int>());
  VfScanKernel<<<1, 1024>>>(ws.d_chunks.as<int>(), chunks);
  VfWriteKernel<<<chunks, 256>>>(ws.d_flags.as<int>(), n, ws.d_chunks.as<int>(), ws.d_kept.as<int>());
  CountLaunch(6);
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(kept, ws.d_chunks.as<int>() + chunks, sizeof(int), cudaMemcpyDeviceToHost));
  return HG_OK;
}

}  // namespace
}  // namespace hg

using namespace hg;

extern "C" {

int hg_voxel_filter(float size, const float* points, int stride_floats, int64_t num_points, int64_t* num_kept,
                    int32_t* kept_indices) {
  HG_REQUIRE(size > 0.f && stride_floats >= 3 && num_points >= 0 && num_kept != nullptr,
             "hg_voxel_filter: bad arguments");
  HG_REQUIRE(num_points < (int64_t(1) << 30), "hg_voxel_filter: too many points");
  *num_kept = 0;
  if (num_points == 0) return HG_OK;
  HG_REQUIRE(points != nullptr && kept_indices != nullptr, "hg_voxel_filter: null arrays");
  int device = 0;
  HG_CUDA(cudaGetDevice(&device));
  VfWorkspace& ws = GetVfWorkspace(device);
  const int n = static_cast<int>(num_points);
  int rc = ws.d_points.Reserve(sizeof(float) * stride_floats * n);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemcpy(ws.d_points.p, points, sizeof(float) * stride_floats * n, cudaMemcpyHostToDevice));
  int kept = 0;
  rc = FilterPass(ws, stride_floats, n, size, nullptr, &kept);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemcpy(kept_indices, ws.d_kept.p, sizeof(int) * kept, cudaMemcpyDeviceToHost));
  *num_kept = kept;
  return HG_OK;
}

int hg_adaptive_voxel_filter(float max_length, int min_num_points, float max_range, const float* points,
                             int stride_floats, int64_t num_points, int64_t* num_kept, int32_t* kept_indices) {
  HG_REQUIRE(max_length > 0.f && min_num_points >= 0 && stride_floats >= 3 && num_points >= 0 && num_kept != nullptr,
             "hg_adaptive_voxel_filter: bad arguments");
  HG_REQUIRE(num_points < (int64_t(1) << 30), "hg_adaptive_voxel_filter: too many points");
  *num_kept = 0;
  if (num_points == 0) return HG_OK;
  HG_REQUIRE(points != nullptr && kept_indices != nullptr, "hg_adaptive_voxel_filter: null arrays");
  int device = 0;
  HG_CUDA(cudaGetDevice(&device));
  VfWorkspace& ws = GetVfWorkspace(device);
  const int n = static_cast<int>(num_points);
  int rc = ws.d_points.Reserve(sizeof(float) * stride_floats * n);
  if (rc != HG_OK) return rc;
  rc = ws.d_active.Reserve(sizeof(int) * n);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemcpy(ws.d_points.p, points, sizeof(float) * stride_floats * n, cudaMemcpyHostToDevice));
  const int blocks = (n + 255) / 256;
  RangeKernel<<<blocks, 256>>>(ws.d_points.as<float>(), stride_floats, n, max_range, ws.d_active.as<int>());
  CountLaunch();
  // Points in range, in input order (= the cloud the reference hands to AdaptivelyVoxelFiltered).
  const int chunks = (n + kVfChunk - 1) / kVfChunk;
  rc = ws.d_chunks.Reserve(sizeof(int) * (chunks + 1));
  if (rc != HG_OK) return rc;
  rc = ws.d_kept.Reserve(sizeof(int) * n);
  if (rc != HG_OK) return rc;
  auto download = [&](int count) {
    HG_CUDA(cudaMemcpy(kept_indices, ws.d_kept.p, sizeof(int) * count, cudaMemcpyDeviceToHost));
    *num_kept = count;
    return HG_OK;
  };
  VfCountKernel<<<chunks, 256>>>(ws.d_active.as<int>(), n, ws.d_chunks.as<int>());
  VfScanKernel<<<1, 1024>>>(ws.d_chunks.as<int>(), chunks);
  VfWriteKernel<<<chunks, 256>>>(ws.d_active.as<int>(), n, ws.d_chunks.as<int>(), ws.d_kept.as<int>());
  CountLaunch(3);
  HG_CUDA(cudaGetLastError());
  int in_range = 0;
  HG_CUDA(cudaMemcpy(&in_range, ws.d_chunks.as<int>() + chunks, sizeof(int), cudaMemcpyDeviceToHost));
  if (in_range <= min_num_points) return download(in_range);  // already sparse enough (:49-52)

  int count = 0;
  rc = FilterPass(ws, stride_floats, n, max_length, ws.d_active.as<int>(), &count);
  if (rc != HG_OK) return rc;
  if (count >= min_num_points) return download(count);  // :53-57
  // :58-83 — halve the edge length until dense enough, then bisect to within 10 %.
  for (float high_length = max_length; high_length > 1e-2f * max_length; high_length /= 2.f) {
    float low_length = high_length / 2.f;
    rc = FilterPass(ws, stride_floats, n, low_length, ws.d_active.as<int>(), &count);
    if (rc != HG_OK) return rc;
    if (count >= min_num_points) {
      float result_length = low_length;
      while ((high_length - low_length) / low_length > 1e-1f) {
        const float mid_length = (low_length + high_length) / 2.f;
        int candidate = 0;
        rc = FilterPass(ws, stride_floats, n, mid_length, ws.d_active.as<int>(), &candidate);
        if (rc != HG_OK) return rc;
        if (candidate >= min_num_points) {
          low_length = mid_length;
          result_length = mid_length;
        } else {
          high_length = mid_length;
        }
      }
      // d_kept holds the last pass; redo the winning length if the last pass was a rejected candidate.
      rc = FilterPass(ws, stride_floats, n, result_length, ws.d_active.as<int>(), &count);
      if (rc != HG_OK) return rc;
      return download(count);
    }
  }
  return download(count);  // gave up: the result of the last (smallest) length, as the reference returns
}

}  // extern "C"
