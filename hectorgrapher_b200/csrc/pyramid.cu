// PrecomputationGridStack3D on the device: uint8 conversion of the submap grid
// and the sliding-maximum pyramid.
//
// Reference behaviour reproduced (bit-exact integer outputs):
//   mapping/internal/3d/scan_matching/precomputation_grid_3d.cc:49-61   HybridGrid -> uint8
//   ... :62-77                                                          HybridGridTSDF -> uint8
//   ... :79-97   PrecomputeGrid: every non-zero source cell c is max-scattered to the
//                8 cells c - shift * octant, optionally folded by >> 1
//   mapping/internal/3d/scan_matching/fast_correlative_scan_matcher_3d.cc:60-94  level loop / shift rule
//
// The scatter is evaluated in gather form, one thread per output voxel:
//   full resolution : out[c] = max_{o in {0,1}^3} in[c + shift o]
//   half resolution : out[h] = max_{o, b in {0,1}^3} in[2h + b + shift o]
// which reads each source voxel 8 (64) times from L1/L2 instead of doing 8
// atomic max-scatters per source voxel, and makes the result independent of
// execution order. The set of output bricks is derived on the host from the
// source level's brick list (a few 10^4 entries); all voxel work is on the GPU.
#include <algorithm>
#include <unordered_set>

#include "device_math.cuh"

namespace hg {
namespace {

__global__ void __launch_bounds__(256) ConvertKernel(const GridDev g, const int64_t num_voxels,
                                                     unsigned char* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= num_voxels) return;
  // i-th core voxel (brick i / 512, z-major 8^3) inside the 9^3 apron brick.
  const int v = static_cast<int>(i & 511);
  const int64_t src = (i >> 9) * kApronBrickVoxels + (((v >> 6) * 9 + ((v >> 3) & 7)) * 9 + (v & 7));
  int cell_value = 0;
  if (g.kind == kGridProb) {
    const unsigned short raw = static_cast<const unsigned short*>(g.table.bricks)[src];
    if (raw != 0) {  // the reference iterates non-default voxels only
      // (ValueToProbability(v) - kMinProbability) * (255.f / (kMaxProbability - kMinProbability))
      const float kMin = 0.1f;
      const float kMax = __fsub_rn(1.f, kMin);
      const float p = Decode(g.a, raw);
      cell_value = RoundToIntF(__fmul_rn(__fsub_rn(p, kMin), __fdiv_rn(255.f, __fsub_rn(kMax, kMin))));
    }
  } else {
    const ushort2 raw = static_cast<const ushort2*>(g.table.bricks)[src];
    if (raw.x != 0 || raw.y != 0) {
      const float tsd = Decode(g.a, raw.x);
      const float normalized = __fsub_rn(1.f, __fdiv_rn(fabsf(tsd), g.max_tsd));
      cell_value = RoundToIntF(__fmul_rn(normalized, 255.f));
    }
  }
  out[i] = static_cast<unsigned char>(cell_value);
}

// One CTA per output brick, two voxels per thread.
__global__ void __launch_bounds__(256) MaxPoolKernel(const BrickTable src, const BrickTable dst,
                                                     const int32_t* __restrict__ dst_brick_coords,
                                                     const int shift, const int half,
                                                     unsigned char* __restrict__ out) {
  const int b = blockIdx.x;
  const int bx = dst_brick_coords[3 * b], by = dst_brick_coords[3 * b + 1], bz = dst_brick_coords[3 * b + 2];
  for (int v = threadIdx.x; v < 512; v += blockDim.x) {
    const int x = (bx << 3) + (v & 7), y = (by << 3) + ((v >> 3) & 7), z = (bz << 3) + (v >> 6);
    int m = 0;
    if (!half) {
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        m = max(m, static_cast<int>(LoadU8(src, x + ((o & 1) ? shift : 0), y + ((o & 2) ? shift : 0),
                                          z + ((o & 4) ? shift : 0))));
      }
    } else {
      for (int o = 0; o < 8; ++o) {
        const int sx = 2 * x + ((o & 1) ? shift : 0), sy = 2 * y + ((o & 2) ? shift : 0),
                  sz = 2 * z + ((o & 4) ? shift : 0);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          m = max(m, static_cast<int>(LoadU8(src, sx + (c & 1), sy + ((c >> 1) & 1), sz + (c >> 2))));
        }
      }
    }
    out[static_cast<size_t>(b) * 512 + v] = static_cast<unsigned char>(m);
  }
}

// Expands the bricks of a level into its dense array (which starts zeroed).
__global__ void __launch_bounds__(256) DensifyKernel(const unsigned char* __restrict__ bricks,
                                                     const int32_t* __restrict__ brick_coords,
                                                     const DenseLevel dense, unsigned char* __restrict__ out) {
  const int b = blockIdx.x;
  const int bx = brick_coords[3 * b], by = brick_coords[3 * b + 1], bz = brick_coords[3 * b + 2];
  for (int v = threadIdx.x; v < 512; v += blockDim.x) {
    const int x = (bx << 3) + (v & 7) - dense.min[0], y = (by << 3) + ((v >> 3) & 7) - dense.min[1],
              z = (bz << 3) + (v >> 6) - dense.min[2];
    out[(static_cast<size_t>(z) * dense.dim[1] + y) * dense.dim[0] + x] = bricks[static_cast<size_t>(b) * 512 + v];
  }
}

__global__ void LookupU8Kernel(const BrickTable t, const int64_t n, const int32_t* __restrict__ xyz,
                               unsigned char* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  out[i] = LoadU8(t, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
}

struct Key {
  int x, y, z;
  bool operator==(const Key& o) const { return x == o.x && y == o.y && z == o.z; }
};
struct KeyHash {
  size_t operator()(const Key& k) const {
    uint64_t h = static_cast<uint32_t>(k.x) * 0x9E3779B97F4A7C15ull;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.y)) + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.z)) + 0x165667B1ull) * 0xD6E8FEB86659FD93ull;
    return static_cast<size_t>(h ^ (h >> 29));
  }
};

// Bricks of the next level that can receive a non-zero value from the bricks
// of the source level: targets of source cell c are (c - shift o) [>> 1].
std::vector<int32_t> TargetBricks(const std::vector<int32_t>& src, const int shift, const bool half) {
  std::unordered_set<Key, KeyHash> seen;
  std::vector<int32_t> out;
  const int64_t nb = static_cast<int64_t>(src.size() / 3);
  for (int64_t i = 0; i < nb; ++i) {
    int lo[3], hi[3];
    for (int k = 0; k < 3; ++k) {
      int a = 8 * src[3 * i + k] - shift;
      int b = 8 * src[3 * i + k] + 7;
      if (half) {
        a >>= 1;
        b >>= 1;
      }
      lo[k] = a >> 3;
      hi[k] = b >> 3;
    }
    for (int z = lo[2]; z <= hi[2]; ++z)
      for (int y = lo[1]; y <= hi[1]; ++y)
        for (int x = lo[0]; x <= hi[0]; ++x) {
          if (seen.insert(Key{x, y, z}).second) {
            out.push_back(x);
            out.push_back(y);
            out.push_back(z);
          }
        }
  }
  return out;
}

}  // namespace
}  // namespace hg

using namespace hg;

extern "C" {

int hg_stack_build(const hg_grid* grid, int branch_and_bound_depth, int full_resolution_depth,
                   hg_stack** out) {
  HG_REQUIRE(grid != nullptr && out != nullptr, "hg_stack_build: bad arguments");
  HG_ON_DEVICE(grid->device);
  // fast_correlative_scan_matcher_3d.cc:63-64 CHECK_GE(..., 1)
  HG_REQUIRE(branch_and_bound_depth >= 1 && full_resolution_depth >= 1,
             "hg_stack_build: branch_and_bound_depth and full_resolution_depth must be >= 1");
  hg_stack* s = new hg_stack;
  s->device = grid->device;
  s->resolution = grid->dev.resolution;
  s->full_resolution_depth = full_resolution_depth;
  s->levels.resize(branch_and_bound_depth);
  s->brick_coords.resize(branch_and_bound_depth);
  int rc = HG_OK;
  auto fail = [&](int code) {
    hg_stack_destroy(s);
    return code;
  };

  // Level 0: same brick set as the grid, uint8 voxels.
  {
    hg_stack_level& L = s->levels[0];
    // core bricks only: apron-only bricks hold no voxel of their own
    s->brick_coords[0].assign(grid->brick_coords.begin(), grid->brick_coords.begin() + 3 * grid->num_bricks);
    rc = UploadBrickIndex(s->brick_coords[0], kCoreBrickVoxels, &L.table, &L.d_index, &L.d_bricks);
    if (rc != HG_OK) return fail(rc);
    L.num_bricks = grid->num_bricks;
    const int64_t nv = grid->num_bricks * 512;
    if (nv > 0) {
      ConvertKernel<<<static_cast<unsigned>((nv + 255) / 256), 256>>>(
          grid->dev, nv, static_cast<unsigned char*>(L.d_bricks));
      CountLaunch();
      if (cudaGetLastError() != cudaSuccess) {
        SetError("ConvertKernel launch failed");
        return fail(HG_ERR_CUDA);
      }
    }
  }
  int last_width = 1;
  for (int depth = 1; depth != branch_and_bound_depth; ++depth) {
    const bool half = depth >= full_resolution_depth;
    const int next_width = 1 << depth;
    const int full_voxels_per_high_resolution_voxel = 1 << std::max(0, depth - full_resolution_depth);
    const int shift = (next_width - last_width + (full_voxels_per_high_resolution_voxel - 1)) /
                      full_voxels_per_high_resolution_voxel;
    last_width = next_width;
    hg_stack_level& L = s->levels[depth];
    const hg_stack_level& S = s->levels[depth - 1];
    s->brick_coords[depth] = TargetBricks(s->brick_coords[depth - 1], shift, half);
    rc = UploadBrickIndex(s->brick_coords[depth], kCoreBrickVoxels, &L.table, &L.d_index, &L.d_bricks);
    if (rc != HG_OK) return fail(rc);
    L.num_bricks = static_cast<int64_t>(s->brick_coords[depth].size() / 3);
    if (L.num_bricks > 0) {
      int32_t* d_coords = nullptr;
      if (cudaMalloc(&d_coords, s->brick_coords[depth].size() * sizeof(int32_t)) != cudaSuccess ||
          cudaMemcpy(d_coords, s->brick_coords[depth].data(),
                     s->brick_coords[depth].size() * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
        SetError("hg_stack_build: brick coordinate upload failed");
        return fail(HG_ERR_CUDA);
      }
      MaxPoolKernel<<<static_cast<unsigned>(L.num_bricks), 256>>>(
          S.table, L.table, d_coords, shift, half ? 1 : 0, static_cast<unsigned char*>(L.d_bricks));
      CountLaunch();
      const cudaError_t e = cudaDeviceSynchronize();
      cudaFree(d_coords);
      if (e != cudaSuccess) {
        SetError(std::string("MaxPoolKernel: ") + cudaGetErrorString(e));
        return fail(HG_ERR_CUDA);
      }
    }
  }
  // Dense copies of the small levels (<= 32 MiB each).
  for (int depth = 0; depth != branch_and_bound_depth; ++depth) {
    hg_stack_level& L = s->levels[depth];
    const std::vector<int32_t>& coords = s->brick_coords[depth];
    if (L.num_bricks == 0) continue;
    int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
    for (int64_t i = 0; i < L.num_bricks; ++i)
      for (int k = 0; k < 3; ++k) {
        const int c = coords[3 * i + k];
        if (i == 0 || c < lo[k]) lo[k] = c;
        if (i == 0 || c > hi[k]) hi[k] = c;
      }
    int64_t cells = 1;
    for (int k = 0; k < 3; ++k) cells *= static_cast<int64_t>(hi[k] - lo[k] + 1) * 8;
    if (cells > (int64_t(32) << 20)) continue;
    int32_t* d_coords = nullptr;
    if (cudaMalloc(&L.d_dense, static_cast<size_t>(cells)) != cudaSuccess ||
        cudaMemset(L.d_dense, 0, static_cast<size_t>(cells)) != cudaSuccess ||
        cudaMalloc(&d_coords, coords.size() * sizeof(int32_t)) != cudaSuccess ||
        cudaMemcpy(d_coords, coords.data(), coords.size() * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess) {
      SetError("hg_stack_build: dense level allocation failed");
      if (d_coords) cudaFree(d_coords);
      return fail(HG_ERR_CUDA);
    }
    for (int k = 0; k < 3; ++k) {
      L.dense.min[k] = lo[k] * 8;
      L.dense.dim[k] = (hi[k] - lo[k] + 1) * 8;
    }
    DensifyKernel<<<static_cast<unsigned>(L.num_bricks), 256>>>(static_cast<const unsigned char*>(L.d_bricks),
                                                               d_coords, L.dense,
                                                               static_cast<unsigned char*>(L.d_dense));
    CountLaunch();
    const cudaError_t e = cudaDeviceSynchronize();
    cudaFree(d_coords);
    if (e != cudaSuccess) {
      SetError(std::string("DensifyKernel: ") + cudaGetErrorString(e));
      return fail(HG_ERR_CUDA);
    }
    L.dense.data = static_cast<const unsigned char*>(L.d_dense);
  }
  if (cudaDeviceSynchronize() != cudaSuccess) {
    SetError("hg_stack_build: device error");
    return fail(HG_ERR_CUDA);
  }
  *out = s;
  return HG_OK;
}

void hg_stack_destroy(hg_stack* stack) {
  if (stack == nullptr) return;
  for (hg_stack_level& L : stack->levels) {
    if (L.d_index) cudaFree(L.d_index);
    if (L.d_bricks) cudaFree(L.d_bricks);
    if (L.d_dense) cudaFree(L.d_dense);
  }
  delete stack;
}

int hg_stack_depth(const hg_stack* stack) { return stack ? static_cast<int>(stack->levels.size()) : 0; }

int hg_stack_lookup(const hg_stack* stack, int depth, int64_t n, const int32_t* xyz, uint8_t* out) {
  HG_REQUIRE(stack != nullptr && depth >= 0 && depth < static_cast<int>(stack->levels.size()) && n >= 0,
             "hg_stack_lookup: bad arguments");
  if (n == 0) return HG_OK;
  HG_ON_DEVICE(stack->device);
  int32_t* d_xyz = nullptr;
  unsigned char* d_out = nullptr;
  HG_CUDA(cudaMalloc(&d_xyz, n * 3 * sizeof(int32_t)));
  HG_CUDA(cudaMalloc(&d_out, n));
  HG_CUDA(cudaMemcpy(d_xyz, xyz, n * 3 * sizeof(int32_t), cudaMemcpyHostToDevice));
  LookupU8Kernel<<<static_cast<unsigned>((n + 255) / 256), 256>>>(stack->levels[depth].table, n, d_xyz, d_out);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(out, d_out, n, cudaMemcpyDeviceToHost));
  cudaFree(d_xyz);
  cudaFree(d_out);
  return HG_OK;
}

static int DownloadLevel(const hg_stack* stack, int depth, std::vector<unsigned char>* bricks) {
  const hg_stack_level& L = stack->levels[depth];
  bricks->resize(static_cast<size_t>(L.num_bricks) * 512);
  if (L.num_bricks > 0) {
    HG_CUDA(cudaMemcpy(bricks->data(), L.d_bricks, bricks->size(), cudaMemcpyDeviceToHost));
  }
  return HG_OK;
}

int64_t hg_stack_num_nonzero(const hg_stack* stack, int depth) {
  if (stack == nullptr || depth < 0 || depth >= static_cast<int>(stack->levels.size())) return -1;
  std::vector<unsigned char> bricks;
  if (DownloadLevel(stack, depth, &bricks) != HG_OK) return -1;
  int64_t n = 0;
  for (unsigned char v : bricks) n += v != 0;
  return n;
}

int hg_stack_export(const hg_stack* stack, int depth, int64_t capacity, int32_t* xyz, uint8_t* values,
                    int64_t* num_written) {
  HG_REQUIRE(stack != nullptr && depth >= 0 && depth < static_cast<int>(stack->levels.size()),
             "hg_stack_export: bad arguments");
  std::vector<unsigned char> bricks;
  const int rc = DownloadLevel(stack, depth, &bricks);
  if (rc != HG_OK) return rc;
  const std::vector<int32_t>& coords = stack->brick_coords[depth];
  int64_t n = 0;
  for (size_t b = 0; b < coords.size() / 3; ++b) {
    for (int v = 0; v < 512; ++v) {
      const unsigned char val = bricks[b * 512 + v];
      if (val == 0) continue;
      if (n < capacity) {
        xyz[3 * n] = 8 * coords[3 * b] + (v & 7);
        xyz[3 * n + 1] = 8 * coords[3 * b + 1] + ((v >> 3) & 7);
        xyz[3 * n + 2] = 8 * coords[3 * b + 2] + (v >> 6);
        values[n] = val;
      }
      ++n;
    }
  }
  if (num_written) *num_written = n;
  return HG_OK;
}

}  // extern "C"
