// Internal declarations shared by the .cu files of libhg_b200.so.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/hg_b200.h"

namespace hg {

// ---------------------------------------------------------------- errors --
void SetError(const std::string& msg);
extern std::atomic<uint64_t> g_kernel_launches;
inline void CountLaunch(uint64_t n = 1) { g_kernel_launches.fetch_add(n, std::memory_order_relaxed); }

#define HG_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::hg::SetError(std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ + \
                     ":" + std::to_string(__LINE__) + ")");                               \
      return HG_ERR_CUDA;                                                                 \
    }                                                                                     \
  } while (0)

#define HG_REQUIRE(cond, msg)          \
  do {                                 \
    if (!(cond)) {                     \
      ::hg::SetError(msg);             \
      return HG_ERR_INVALID;           \
    }                                  \
  } while (0)

// ---------------------------------------------------------------- devices --
// One policy for every entry point that takes a handle: the call runs on the handle's device and
// the caller's current device is what it was when the call returns.
class DeviceGuard {
 public:
  explicit DeviceGuard(const int device) {
    if (cudaGetDevice(&previous_) != cudaSuccess) previous_ = -1;
    if (previous_ != device) {
      ok_ = cudaSetDevice(device) == cudaSuccess;
      switched_ = ok_;
    }
  }
  ~DeviceGuard() {
    if (switched_ && previous_ >= 0) cudaSetDevice(previous_);
  }
  bool ok() const { return ok_; }
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;

 private:
  int previous_ = -1;
  bool switched_ = false;
  bool ok_ = true;
};
#define HG_ON_DEVICE(device)                                             \
  ::hg::DeviceGuard _hg_device_guard(device);                            \
  do {                                                                   \
    if (!_hg_device_guard.ok()) {                                        \
      ::hg::SetError("cannot switch to the handle's CUDA device");      \
      return HG_ERR_CUDA;                                                \
    }                                                                    \
  } while (0)

// ------------------------------------------------------------ allocation --
// alloc.cu: blocks released by a host thread are cached by that thread and reused (no
// cudaFree on the per-pair path); a block may be released by another thread than the
// one that allocated it. ThreadStream() is a non-blocking stream owned by the calling
// host thread (one per device), created on first use.
void* CachedDeviceAlloc(size_t bytes);  // nullptr + SetError on failure
void* CachedPinnedAlloc(size_t bytes);
void CachedFree(void* p);
cudaStream_t ThreadStream();

#define HG_ALLOC(var, expr)                          \
  do {                                               \
    void* _p = (expr);                               \
    if (_p == nullptr) return HG_ERR_CUDA;           \
    (var) = static_cast<decltype(var)>(_p);          \
  } while (0)

// ----------------------------------------------------------- brick table --
// A sparse voxel grid flattened for the device:
//   * 'index' is a dense int32 array over the bounding box of occupied 8x8x8
//     bricks (brick coordinates = voxel index >> 3, arithmetic shift), holding
//     the brick id or -1;
//   * 'bricks' holds the voxels, z-major inside a brick like the reference's
//     FlatGrid (hybrid_grid_base.h:40-43). Precomputation levels (uint8) store
//     the plain 8x8x8 = 512 voxels per brick: ((z*8)+y)*8+x.
//     Submap grids (uint16 probability, ushort2 {tsd, weight} TSDF) store 9x9x9 =
//     729 voxels per brick: the 8^3 core plus a one-voxel apron on the high side
//     of each axis that repeats the neighbouring bricks' first planes, index
//     ((z*9)+y)*9+x with x,y,z in [0,8]. A 2x2x2 interpolation stencil whose
//     lower corner lies in a brick's core therefore never leaves that brick: one
//     index load and eight voxel loads at fixed offsets. Bricks that hold only
//     apron copies (no core voxel) are created where needed and listed after the
//     core bricks.
//   * when that bounding box would be too large (a few voxels at opposite
//     corners of the +-8192 index range), 'index' is instead an open-addressing
//     hash table of int4 {bx, by, bz, id} entries (id -1 = empty), selected by
//     hash_mask != 0.
struct BrickTable {
  const int32_t* index;  // device
  const void* bricks;    // device
  int bmin[3];           // brick-coordinate origin of 'index'
  int bdim[3];           // extent of 'index' in bricks
  unsigned hash_mask;    // 0 = dense index; otherwise capacity - 1 of the hash table
};

__host__ __device__ inline unsigned BrickHash(int bx, int by, int bz) {
  unsigned h = static_cast<unsigned>(bx) * 73856093u ^ static_cast<unsigned>(by) * 19349663u ^
               static_cast<unsigned>(bz) * 83492791u;
  h ^= h >> 15;
  return h * 0x2c1b3c6du;
}

// Value decoding constants: value -> float is v' = v & 0x7FFF; v' == 0 ? unknown
// : v' * scale + offset with offset = lo - scale (value_conversion_tables.cc:29-51),
// evaluated with separate round-to-nearest multiply and add.
constexpr int kCoreBrickVoxels = 512;   // precomputation levels
constexpr int kApronBrickVoxels = 729;  // submap grids

struct ValueDecode {
  float scale;
  float offset;
  float unknown;
};

enum GridKind { kGridProb = 0, kGridTsdf = 1 };

struct GridDev {
  BrickTable table;
  // Same bricks (apron layout), pre-decoded for the cost kernels: one float per
  // voxel. Probability grids: the decoded probability. TSDF grids: the decoded
  // TSD of an observed voxel, NaN for an unobserved one (weight value 0 or 1) —
  // the cost functors never read the TSD of an unobserved voxel.
  const float* decoded;
  // Optional third copy for the Jacobian kernels (EnsureStencils, grid.cu): for every voxel of
  // every brick the 2x2x2 stencil whose lower corner it is, eight decoded floats in the
  // reference's corner order (111,112,121,122,211,212,221,222) = one 32-byte sector, so an
  // interpolated lookup is two 16-byte loads from one sector instead of eight 4-byte loads from
  // four to eight. Null until a cost function asks for it; dropped when the grid is written.
  const float4* stencils;
  float resolution;
  float inv_resolution;     // fl(1 / resolution): fast path of the exact cell-index computation
  double inv_resolution_d;  // 1.0 / double(resolution): seed for the reciprocal of the cell width
  ValueDecode a;  // probability, or tsd
  ValueDecode b;  // weight (TSDF only)
  float min_tsd;  // -max_tsd
  float max_tsd;
  int kind;
};

struct GridMutable;  // insert.cu: state of a grid that is written on the device

}  // namespace hg

// Host-side handle definitions (opaque in the public header).
struct hg_grid {
  hg::GridDev dev;
  int device = 0;
  int grid_size = 0;
  int64_t num_bricks = 0;       // bricks with at least one core voxel (listed first)
  int64_t num_all_bricks = 0;   // + apron-only bricks
  int64_t num_voxels = 0;
  float relative_truncation_distance = 0.f;
  float max_weight = 0.f;
  void* d_index = nullptr;
  void* d_bricks = nullptr;
  void* d_decoded = nullptr;
  void* d_stencils = nullptr;
  int64_t stencil_capacity = 0;  // voxels d_stencils has room for (32 bytes each)
  // Host copy of the brick coordinates (brick id -> brick coords), used to
  // derive the precomputation levels' brick sets.
  std::vector<int32_t> brick_coords;
  hg::GridMutable* mut = nullptr;  // created by the first device-side insertion
};

namespace hg {
void DestroyGridMutable(hg_grid* grid);  // insert.cu
// grid.cu: builds GridDev::stencils if it is missing (thread-safe; the grid is logically const).
int EnsureStencils(const hg_grid* grid);
// Marks GridDev::stencils stale (the grid's voxels are about to change); the buffer is kept.
void DropStencils(hg_grid* grid);

// scan_match.cu: CeresScanMatcher3D::Match solved entirely on the device (arguments of
// hg_scan_match). HG_ERR_LIMIT when the problem does not fit the kernel.
int ScanMatchOnDevice(int num_clouds, const float* const* clouds, const int32_t* cloud_sizes,
                      const hg_grid* const* grids, const double* occupied_space_weights,
                      double translation_weight, double rotation_weight, int only_optimize_yaw,
                      int max_num_iterations, int use_nonmonotonic_steps, const double* target_translation,
                      const double* initial_pose, double* pose_out, double* summary_out);

// scan_match.cu: the same for a batch of problems in one launch (one CTA each; every problem at most
// 4 clouds and 4096 points in total, HG_ERR_LIMIT otherwise). poses_out: n x 7, summaries_out: n x 6 or null.
struct ScanMatchBatchItem {
  int num_clouds;
  const float* clouds[4];
  int32_t sizes[4];
  const hg_grid* grids[4];
  double target_translation[3];
  double initial_pose[7];
};
int ScanMatchBatchOnDevice(int num_problems, const ScanMatchBatchItem* items, const double* occupied_space_weights,
                           double translation_weight, double rotation_weight, int only_optimize_yaw,
                           int max_num_iterations, int use_nonmonotonic_steps, double* poses_out,
                           double* summaries_out);

// Optional dense copy of a (small) precomputation level: a plain uint8 3-D array
// over the bounding box of its bricks, x fastest. Coarse levels of the pyramid
// are a few hundred KB; reading them densely saves the brick-index load.
struct DenseLevel {
  const unsigned char* data;  // nullptr = not available
  int min[3];                 // cell coordinate of data[0]
  int dim[3];                 // extent in cells
};
}  // namespace hg

struct hg_stack_level {
  hg::BrickTable table;
  hg::DenseLevel dense = {nullptr, {0, 0, 0}, {0, 0, 0}};
  void* d_dense = nullptr;
  void* d_index = nullptr;
  void* d_bricks = nullptr;
  int64_t num_bricks = 0;
  bool owns_index = true;
};

struct hg_stack {
  int device = 0;
  float resolution = 0.f;
  int full_resolution_depth = 0;
  std::vector<hg_stack_level> levels;
  std::vector<std::vector<int32_t>> brick_coords;  // per level: brick id -> brick coordinates
};

namespace hg {
// Builds a device brick table (index + zeroed bricks of 'voxel_bytes' each)
// from a list of brick coordinates. Used by grid upload and the pyramid.
int UploadBrickIndex(const std::vector<int32_t>& brick_coords, int brick_bytes, BrickTable* table,
                     void** d_index, void** d_bricks);
// Host half of UploadBrickIndex: fills table->{bmin, bdim, hash_mask} and the index
// entries (dense int32 array, or int4 hash entries when hash_mask != 0).
int BuildBrickIndexHost(const std::vector<int32_t>& brick_coords, BrickTable* table,
                        std::vector<int32_t>* entries);
}  // namespace hg
