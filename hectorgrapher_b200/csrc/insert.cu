// Range-data insertion into device-resident submap grids.
//
// Reference behaviour reproduced here (paths under /root/reference/cartographer):
//   mapping/3d/range_data_inserter_3d.cc:27-101       OccupancyGridRangeDataInserter3D::Insert
//   mapping/3d/hybrid_grid.h:76-82, 119-130            FinishUpdate / ApplyLookupTable (update marker)
//   mapping/probability_values.cc:76-87                ComputeLookupTableToApplyOdds
//   mapping/3d/tsdf_range_data_inserter_3d.cc:294-343  InsertHit, default ray cast
//   mapping/3d/tsdf_range_data_inserter_3d.cc:395-401, 700-737  Insert (no normal projection), UpdateCell
//   mapping/3d/hybrid_grid_tsdf.h:87-93                SetCell (TSD value carries bit 15)
//   mapping/2d/tsd_value_converter.h:39-55             TSDToValue / WeightToValue
//   mapping/3d/hybrid_grid_base.h:385-403              DynamicGrid::Grow
//
// How it is laid out for the GPU. The reference walks the returns one by one and
// read-modify-writes voxels through the tree; a TSDF voxel's new value depends on
// every earlier update of that voxel (running weighted mean, re-quantised to uint16
// after each update), so the order of updates PER VOXEL is part of the result. Here:
//   1. one thread per return counts its ray samples; an exclusive scan gives every
//      (return, sample) pair its sequence number — the reference's visiting order;
//   2. one thread per (return, sample) computes the cell and the update (tsd, weight)
//      with the reference's float operations and emits a 42-bit cell key;
//   3. a stable radix sort by key (the update pair rides along as the value) groups the
//      updates of a cell and keeps them in sequence order inside the group;
//   4. new bricks are created (host bookkeeping over the few thousand touched bricks);
//   5. one thread per touched cell replays that cell's updates in order on the stored
//      uint16 pair — bit-identical to the sequential loop, no atomics on voxel data;
//   6. the read-optimised copies (apron bricks, pre-decoded floats; hg_internal.cuh)
//      are regenerated from the core bricks.
// The occupancy inserter is a set operation (each cell at most once per scan, hits win):
// the same sort, keyed (cell, hit-or-miss), and one table lookup per touched cell.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <thrust/iterator/transform_iterator.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <unordered_map>

#include "device_math.cuh"

namespace hg {

namespace {

constexpr int kKeyBias = 8192;  // cells live in [-8192, 8191] (grid_size <= 64 << 8)
constexpr int kCellKeyBits = 42;

// (bz:11 | by:11 | bx:11 | z:3 | y:3 | x:3): cells of one brick are contiguous in key order.
__host__ __device__ inline uint64_t CellKey(const int x, const int y, const int z) {
  const unsigned ux = static_cast<unsigned>(x + kKeyBias), uy = static_cast<unsigned>(y + kKeyBias),
                 uz = static_cast<unsigned>(z + kKeyBias);
  const uint64_t brick = (static_cast<uint64_t>(uz >> 3) << 22) | (static_cast<uint64_t>(uy >> 3) << 11) |
                         static_cast<uint64_t>(ux >> 3);
  return (brick << 9) | ((uz & 7u) << 6) | ((uy & 7u) << 3) | (ux & 7u);
}
__host__ __device__ inline void CellOfKey(const uint64_t key, int* x, int* y, int* z) {
  const uint64_t brick = key >> 9;
  *x = static_cast<int>(((brick & 2047u) << 3) | (key & 7u)) - kKeyBias;
  *y = static_cast<int>((((brick >> 11) & 2047u) << 3) | ((key >> 3) & 7u)) - kKeyBias;
  *z = static_cast<int>((((brick >> 22) & 2047u) << 3) | ((key >> 6) & 7u)) - kKeyBias;
}
inline uint64_t BrickKeyOfCoords(const int bx, const int by, const int bz) {
  return (static_cast<uint64_t>(bz + 1024) << 22) | (static_cast<uint64_t>(by + 1024) << 11) |
         static_cast<uint64_t>(bx + 1024);
}
inline void CoordsOfBrickKey(const uint64_t k, int* bx, int* by, int* bz) {
  *bx = static_cast<int>(k & 2047u) - 1024;
  *by = static_cast<int>((k >> 11) & 2047u) - 1024;
  *bz = static_cast<int>((k >> 22) & 2047u) - 1024;
}

// HG_TRACE=1 prints phase timings of every insertion to stderr (synchronises at phase ends).
struct PhaseTrace {
  bool on;
  std::chrono::steady_clock::time_point t;
  std::string text;
  PhaseTrace() : on(std::getenv("HG_TRACE") != nullptr), t(std::chrono::steady_clock::now()) {}
  void Mark(const char* what) {
    if (!on) return;
    cudaDeviceSynchronize();
    const auto now = std::chrono::steady_clock::now();
    char buf[96];
    std::snprintf(buf, sizeof(buf), " %s %.3f", what, std::chrono::duration<double, std::milli>(now - t).count());
    text += buf;
    t = now;
  }
  ~PhaseTrace() {
    if (on) std::fprintf(stderr, "[hg insert]%s ms\n", text.c_str());
  }
};

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int Ensure(const size_t bytes) {
    if (bytes <= cap) return HG_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    const size_t want = bytes + bytes / 2 + 256;
    HG_CUDA(cudaMalloc(&p, want));
    cap = want;
    return HG_OK;
  }
  void Free() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const {
    return static_cast<T*>(p);
  }
};

// Device-side counters of one insertion.
struct InsertState {
  int error;        // a cell fell outside +-8192
  int need;         // max over written cells of (c >= 0 ? c + 1 : -c): half the grid size that contains them
  int new_voxels;   // voxels that went from default to written
  int pad;
};

}  // namespace

struct GridMutable {
  // Core bricks: 512 voxels per brick, z-major, ids = the first num_bricks entries of
  // hg_grid::brick_coords. The truth the apron / decoded copies are regenerated from.
  void* d_core = nullptr;
  int64_t cap_core = 0;  // bricks
  std::unordered_map<uint64_t, int32_t> core_ids;  // brick key -> core id
  std::vector<int32_t> core_coords;                // 3 per core brick
  std::unordered_map<uint64_t, int32_t> apron_pos;  // apron-only brick key -> position in apron_keys
  std::vector<uint64_t> apron_keys;
  int64_t cap_all = 0;       // bricks that fit d_bricks / d_decoded
  size_t cap_index = 0;      // bytes of d_index
  DevBuf coords, points, mask, counts, offsets, keys[2], vals[2], flags, starts, brick_heads, brick_keys, temp,
      state, num;
  uint16_t* d_tables = nullptr;  // hit table, miss table (32768 entries each)
  float table_hit = -1.f, table_miss = -1.f;
  void* h_pinned = nullptr;  // small pinned staging area
};

void DestroyGridMutable(hg_grid* grid) {
  GridMutable* m = grid->mut;
  if (m == nullptr) return;
  if (m->d_core) cudaFree(m->d_core);
  for (DevBuf* b : {&m->coords, &m->points, &m->mask, &m->counts, &m->offsets, &m->keys[0], &m->keys[1],
                    &m->vals[0], &m->vals[1], &m->flags, &m->starts, &m->brick_heads,
                    &m->brick_keys, &m->temp, &m->state, &m->num}) {
    b->Free();
  }
  if (m->d_tables) cudaFree(m->d_tables);
  if (m->h_pinned) cudaFreeHost(m->h_pinned);
  delete m;
  grid->mut = nullptr;
}

namespace {

// ------------------------------------------------------------------ kernels --

template <typename Voxel>
__global__ void __launch_bounds__(256) ExtractCoreKernel(const Voxel* __restrict__ apron, const int64_t num_core,
                                                         Voxel* __restrict__ core) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= num_core * 512) return;
  const int64_t b = i >> 9;
  const int v = static_cast<int>(i & 511);
  core[i] = apron[b * kApronBrickVoxels + (((v >> 6) * 9 + ((v >> 3) & 7)) * 9 + (v & 7))];
}

// Regenerates apron bricks and their decoded copy from the core bricks: voxel (lx, ly, lz)
// in [0, 8]^3 of brick b is the core voxel (l & 7) of brick b + (l == 8).
template <typename Voxel>
__global__ void __launch_bounds__(256) BuildApronKernel(const GridDev g, const int32_t* __restrict__ coords,
                                                        const Voxel* __restrict__ core, const int num_core,
                                                        const int64_t num_all, Voxel* __restrict__ apron,
                                                        float* __restrict__ decoded) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= num_all * kApronBrickVoxels) return;
  const int b = static_cast<int>(i / kApronBrickVoxels);
  const int v = static_cast<int>(i - static_cast<int64_t>(b) * kApronBrickVoxels);
  const int lz = v / 81, ly = (v / 9) % 9, lx = v % 9;
  int src = b;
  if (lx == 8 || ly == 8 || lz == 8) {
    src = BrickId(g.table, (coords[3 * b] + (lx >> 3)) << 3, (coords[3 * b + 1] + (ly >> 3)) << 3,
                  (coords[3 * b + 2] + (lz >> 3)) << 3);
  }
  Voxel raw;
  memset(&raw, 0, sizeof(Voxel));
  if (src >= 0 && src < num_core) {
    raw = core[static_cast<int64_t>(src) * 512 + (((lz & 7) << 6) | ((ly & 7) << 3) | (lx & 7))];
  }
  apron[i] = raw;
  if constexpr (sizeof(Voxel) == 2) {
    decoded[i] = Decode(g.a, raw);
  } else {
    decoded[i] = (raw.y & 0x7FFFu) > 1u ? Decode(g.a, raw.x) : __int_as_float(0x7fc00000);
  }
}

struct RayParams {
  int stride;  // floats between consecutive returns
  float ox, oy, oz;
  float resolution, inv_resolution;
  // TSDF
  float truncation_distance;
  int update_free_space;
  float epsilon, sigma;
  double min_range, max_range;
  // occupancy
  int num_free_space_voxels;
};

struct Ray {
  int bx, by, bz;  // begin cell (occupancy: origin cell)
  int dx, dy, dz;
  int num_samples;
  float range;
};

__device__ __forceinline__ float Norm3(const float x, const float y, const float z) {
  // Eigen's Vector3f::norm(): sqrt(x^2 + (y^2 + z^2)), scalar path.
  return __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z))));
}

// InsertHit up to num_samples (tsdf_range_data_inserter_3d.cc:294-318) with the gates of
// Insert (:709-716). Returns false when the return contributes nothing.
__device__ __forceinline__ bool SetupTsdfRay(const RayParams& p, const float hx, const float hy, const float hz,
                                             Ray* r) {
  if (isnan(hx) || isnan(hy) || isnan(hz)) return false;
  const float rx = __fsub_rn(hx, p.ox), ry = __fsub_rn(hy, p.oy), rz = __fsub_rn(hz, p.oz);
  const float range = Norm3(rx, ry, rz);
  if (static_cast<double>(range) < p.min_range) return false;
  if (static_cast<double>(range) > p.max_range) return false;
  if (range < p.truncation_distance) return false;
  const float ratio = __fdiv_rn(p.truncation_distance, range);
  float bxf = p.ox, byf = p.oy, bzf = p.oz;
  if (!p.update_free_space) {
    const float fb = __fsub_rn(1.0f, ratio);
    bxf = __fadd_rn(p.ox, __fmul_rn(fb, rx));
    byf = __fadd_rn(p.oy, __fmul_rn(fb, ry));
    bzf = __fadd_rn(p.oz, __fmul_rn(fb, rz));
  }
  const float fe = __fadd_rn(1.0f, ratio);
  const float exf = __fadd_rn(p.ox, __fmul_rn(fe, rx));
  const float eyf = __fadd_rn(p.oy, __fmul_rn(fe, ry));
  const float ezf = __fadd_rn(p.oz, __fmul_rn(fe, rz));
  r->bx = CellIndexFast(bxf, p.resolution, p.inv_resolution);
  r->by = CellIndexFast(byf, p.resolution, p.inv_resolution);
  r->bz = CellIndexFast(bzf, p.resolution, p.inv_resolution);
  r->dx = CellIndexFast(exf, p.resolution, p.inv_resolution) - r->bx;
  r->dy = CellIndexFast(eyf, p.resolution, p.inv_resolution) - r->by;
  r->dz = CellIndexFast(ezf, p.resolution, p.inv_resolution) - r->bz;
  r->num_samples = max(abs(r->dx), max(abs(r->dy), abs(r->dz)));
  r->range = range;
  return true;
}

__global__ void __launch_bounds__(256) TsdfCountKernel(const RayParams p, const float* __restrict__ pts,
                                                       const unsigned char* __restrict__ mask, const int n,
                                                       int* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int c = 0;
  if (mask == nullptr || mask[i]) {
    Ray r;
    if (SetupTsdfRay(p, pts[static_cast<size_t>(p.stride) * i], pts[static_cast<size_t>(p.stride) * i + 1], pts[static_cast<size_t>(p.stride) * i + 2], &r)) c = r.num_samples + 1;
  }
  counts[i] = c;
}

// Largest i with offsets[i] <= u (offsets is an exclusive scan with n + 1 entries).
__device__ __forceinline__ int OwnerOf(const int* __restrict__ offsets, const int n, const int u) {
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(offsets + mid) <= u) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ bool CellInRange(const int x, const int y, const int z) {
  return x >= -kKeyBias && x < kKeyBias && y >= -kKeyBias && y < kKeyBias && z >= -kKeyBias && z < kKeyBias;
}

// One thread per (return, sample): the cell (:321-323), the clamped distance update and its
// weight (:325-338).
__global__ void __launch_bounds__(256) TsdfEmitKernel(const RayParams p, const float* __restrict__ pts,
                                                      const int* __restrict__ offsets, const int n,
                                                      const int total, uint64_t* __restrict__ keys,
                                                      float2* __restrict__ updates,
                                                      InsertState* __restrict__ state) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= total) return;
  const int i = OwnerOf(offsets, n, u);
  const int position = u - offsets[i];
  Ray r;
  SetupTsdfRay(p, pts[static_cast<size_t>(p.stride) * i], pts[static_cast<size_t>(p.stride) * i + 1], pts[static_cast<size_t>(p.stride) * i + 2], &r);
  const float fp = static_cast<float>(position), fn = static_cast<float>(r.num_samples);
  const int cx = r.bx + RoundToIntF(__fdiv_rn(__fmul_rn(static_cast<float>(r.dx), fp), fn));
  const int cy = r.by + RoundToIntF(__fdiv_rn(__fmul_rn(static_cast<float>(r.dy), fp), fn));
  const int cz = r.bz + RoundToIntF(__fdiv_rn(__fmul_rn(static_cast<float>(r.dz), fp), fn));
  if (!CellInRange(cx, cy, cz)) {
    atomicOr(&state->error, 1);
    keys[u] = 0;
    updates[u] = make_float2(0.f, 0.f);
    return;
  }
  keys[u] = CellKey(cx, cy, cz);
  const float dx = __fsub_rn(__fmul_rn(static_cast<float>(cx), p.resolution), p.ox);
  const float dy = __fsub_rn(__fmul_rn(static_cast<float>(cy), p.resolution), p.oy);
  const float dz = __fsub_rn(__fmul_rn(static_cast<float>(cz), p.resolution), p.oz);
  float tsd = __fsub_rn(r.range, Norm3(dx, dy, dz));
  if (tsd > p.truncation_distance) tsd = p.truncation_distance;
  if (tsd < -p.truncation_distance) tsd = -p.truncation_distance;
  float weight = 1.0f;
  const float normalized = __fdiv_rn(tsd, p.truncation_distance);
  if (normalized < -p.epsilon) {
    // float(std::exp(-sigma * std::pow(-normalized - epsilon, 2))): the subtraction is a float
    // operation, pow / the product / exp are double (std::pow(float, int) promotes).
    const double t = static_cast<double>(__fsub_rn(-normalized, p.epsilon));
    weight = static_cast<float>(exp(__dmul_rn(static_cast<double>(-p.sigma), __dmul_rn(t, t))));
  }
  updates[u] = make_float2(tsd, weight);
}

template <int SHIFT>
__global__ void __launch_bounds__(256) HeadFlagsKernel(const uint64_t* __restrict__ keys, const int n,
                                                       unsigned char* __restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  flags[i] = (i == 0 || (keys[i] >> SHIFT) != (keys[i - 1] >> SHIFT)) ? 1 : 0;
}

// Heads of brick runs among the unique cells; SHIFT = bits below the brick key.
template <int SHIFT>
__global__ void __launch_bounds__(256) BrickFlagsKernel(const uint64_t* __restrict__ keys,
                                                        const int* __restrict__ starts, const int num_cells,
                                                        unsigned char* __restrict__ flags) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_cells) return;
  flags[j] = (j == 0 || (keys[starts[j]] >> SHIFT) != (keys[starts[j - 1]] >> SHIFT)) ? 1 : 0;
}

template <int SHIFT>
__global__ void __launch_bounds__(256) GatherBrickKeysKernel(const uint64_t* __restrict__ keys,
                                                             const int* __restrict__ starts,
                                                             const int* __restrict__ brick_heads,
                                                             const int num_bricks, uint64_t* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_bricks) return;
  out[j] = keys[starts[brick_heads[j]]] >> SHIFT;
}

__device__ __forceinline__ void NoteWritten(InsertState* state, const int x, const int y, const int z,
                                            const bool was_default) {
  const int need = max(x >= 0 ? x + 1 : -x, max(y >= 0 ? y + 1 : -y, z >= 0 ? z + 1 : -z));
  atomicMax(&state->need, need);
  if (was_default) atomicAdd(&state->new_voxels, 1);
}

struct TsdfApplyParams {
  ValueDecode tsd, weight;
  float min_tsd, max_tsd, tsd_resolution;
  float max_weight, weight_resolution;  // the grid's converter
  float maximum_weight;                 // the inserter's cap
};

// UpdateCell (:725-737) + SetCell (hybrid_grid_tsdf.h:87-93), one thread per touched cell,
// that cell's updates in sequence order.
__global__ void __launch_bounds__(128) TsdfApplyKernel(const BrickTable table, const TsdfApplyParams p,
                                                       const uint64_t* __restrict__ keys,
                                                       const float2* __restrict__ updates,
                                                       const int* __restrict__ starts, const int num_cells,
                                                       const int total, ushort2* __restrict__ core,
                                                       InsertState* __restrict__ state) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_cells) return;
  const int s = starts[j];
  const int e = j + 1 < num_cells ? starts[j + 1] : total;
  int x, y, z;
  CellOfKey(keys[s], &x, &y, &z);
  const int id = BrickId(table, x, y, z);
  if (id < 0) {
    atomicOr(&state->error, 2);
    return;
  }
  ushort2* const voxel = core + static_cast<int64_t>(id) * 512 + InBrick(x, y, z);
  ushort2 v = *voxel;
  const bool was_default = v.x == 0 && v.y == 0;
  bool wrote = false;
  // The chain is sequential by definition; its updates are fetched a chunk ahead so that only
  // the arithmetic is on the critical path. Padding entries have weight 0 and are skipped like
  // the reference skips zero-weight updates.
  constexpr int kChunk = 8;
  float2 ahead[kChunk];
#pragma unroll
  for (int i = 0; i < kChunk; ++i) ahead[i] = s + i < e ? __ldg(updates + s + i) : make_float2(0.f, 0.f);
  for (int k = s; k < e; k += kChunk) {
    float2 current[kChunk];
#pragma unroll
    for (int i = 0; i < kChunk; ++i) current[i] = ahead[i];
#pragma unroll
    for (int i = 0; i < kChunk; ++i) {
      ahead[i] = k + kChunk + i < e ? __ldg(updates + k + kChunk + i) : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < kChunk; ++i) {
      const float update_weight = current[i].y;
      if (update_weight == 0.f) continue;
      const float update_sdf = current[i].x;
      const float old_weight = Decode(p.weight, v.y);
      const float old_sdf = Decode(p.tsd, v.x);
      float updated_weight = __fadd_rn(old_weight, update_weight);
      const float updated_sdf = __fdiv_rn(
          __fadd_rn(__fmul_rn(old_sdf, old_weight), __fmul_rn(update_sdf, update_weight)), updated_weight);
      updated_weight = updated_weight < p.maximum_weight ? updated_weight : p.maximum_weight;  // std::min
      float c = updated_sdf;
      if (c > p.max_tsd) c = p.max_tsd;
      if (c < p.min_tsd) c = p.min_tsd;
      float w = updated_weight;
      if (w > p.max_weight) w = p.max_weight;
      if (w < 0.f) w = 0.f;
      v.x = static_cast<unsigned short>(
          (RoundToIntF(__fmul_rn(__fsub_rn(c, p.min_tsd), p.tsd_resolution)) + 1 + 0x8000) & 0xFFFF);
      v.y = static_cast<unsigned short>(RoundToIntF(__fmul_rn(__fsub_rn(w, 0.f), p.weight_resolution)) + 1);
      wrote = true;
    }
  }
  if (wrote) {
    *voxel = v;
    NoteWritten(state, x, y, z, was_default);
  }
}

// ---- occupancy ---------------------------------------------------------------

__device__ __forceinline__ void SetupOccupancyRay(const RayParams& p, const float hx, const float hy,
                                                  const float hz, Ray* r, int* hit_x, int* hit_y, int* hit_z) {
  r->bx = CellIndexFast(p.ox, p.resolution, p.inv_resolution);
  r->by = CellIndexFast(p.oy, p.resolution, p.inv_resolution);
  r->bz = CellIndexFast(p.oz, p.resolution, p.inv_resolution);
  *hit_x = CellIndexFast(hx, p.resolution, p.inv_resolution);
  *hit_y = CellIndexFast(hy, p.resolution, p.inv_resolution);
  *hit_z = CellIndexFast(hz, p.resolution, p.inv_resolution);
  r->dx = *hit_x - r->bx;
  r->dy = *hit_y - r->by;
  r->dz = *hit_z - r->bz;
  r->num_samples = max(abs(r->dx), max(abs(r->dy), abs(r->dz)));
}

__global__ void __launch_bounds__(256) OccupancyCountKernel(const RayParams p, const float* __restrict__ pts,
                                                            const int n, int* __restrict__ counts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray r;
  int hx, hy, hz;
  SetupOccupancyRay(p, pts[static_cast<size_t>(p.stride) * i], pts[static_cast<size_t>(p.stride) * i + 1], pts[static_cast<size_t>(p.stride) * i + 2], &r, &hx, &hy, &hz);
  counts[i] = 1 + min(r.num_samples, max(p.num_free_space_voxels, 0));
}

// Sample 0 of a return is its hit cell; samples 1.. are the last num_free_space_voxels
// positions of InsertMissesIntoGrid (range_data_inserter_3d.cc:34-51). Key bit 0: 0 hit, 1 miss,
// so that a hit sorts before a miss of the same cell.
__global__ void __launch_bounds__(256) OccupancyEmitKernel(const RayParams p, const float* __restrict__ pts,
                                                           const int* __restrict__ offsets, const int n,
                                                           const int total, uint64_t* __restrict__ keys,
                                                           InsertState* __restrict__ state) {
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= total) return;
  const int i = OwnerOf(offsets, n, u);
  const int k = u - offsets[i];
  Ray r;
  int cx, cy, cz;
  SetupOccupancyRay(p, pts[static_cast<size_t>(p.stride) * i], pts[static_cast<size_t>(p.stride) * i + 1], pts[static_cast<size_t>(p.stride) * i + 2], &r, &cx, &cy, &cz);
  unsigned miss = 0;
  if (k > 0) {
    const int position = max(0, r.num_samples - p.num_free_space_voxels) + (k - 1);
    cx = r.bx + r.dx * position / r.num_samples;  // integer division truncates like Eigen's Array3i
    cy = r.by + r.dy * position / r.num_samples;
    cz = r.bz + r.dz * position / r.num_samples;
    miss = 1;
  }
  if (!CellInRange(cx, cy, cz)) {
    atomicOr(&state->error, 1);
    keys[u] = 0;
    return;
  }
  keys[u] = (CellKey(cx, cy, cz) << 1) | miss;
}

__global__ void __launch_bounds__(128) OccupancyApplyKernel(const BrickTable table,
                                                            const uint64_t* __restrict__ keys,
                                                            const int* __restrict__ starts, const int num_cells,
                                                            const uint16_t* __restrict__ tables,
                                                            unsigned short* __restrict__ core,
                                                            InsertState* __restrict__ state) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= num_cells) return;
  const uint64_t key = keys[starts[j]];
  int x, y, z;
  CellOfKey(key >> 1, &x, &y, &z);
  const int id = BrickId(table, x, y, z);
  if (id < 0) {
    atomicOr(&state->error, 2);
    return;
  }
  unsigned short* const voxel = core + static_cast<int64_t>(id) * 512 + InBrick(x, y, z);
  const unsigned short old = *voxel;
  // table[old] carries the update marker; FinishUpdate removes it again.
  *voxel = static_cast<unsigned short>(tables[(key & 1u) * 32768u + (old & 0x7FFFu)] & 0x7FFFu);
  NoteWritten(state, x, y, z, old == 0);
}

// --------------------------------------------------------------- host side --

unsigned Blocks(const int64_t n, const int threads) { return static_cast<unsigned>((n + threads - 1) / threads); }

size_t VoxelBytes(const hg_grid* g) { return g->dev.kind == kGridProb ? sizeof(uint16_t) : sizeof(ushort2); }

// First insertion: take over the buffers of an uploaded grid and extract its core bricks.
int MakeMutable(hg_grid* g) {
  if (g->mut != nullptr) return HG_OK;
  GridMutable* m = new GridMutable;
  g->mut = m;
  m->core_coords.assign(g->brick_coords.begin(), g->brick_coords.begin() + 3 * g->num_bricks);
  for (int64_t b = 0; b < g->num_bricks; ++b) {
    m->core_ids.emplace(BrickKeyOfCoords(g->brick_coords[3 * b], g->brick_coords[3 * b + 1],
                                         g->brick_coords[3 * b + 2]),
                        static_cast<int32_t>(b));
  }
  for (int64_t b = g->num_bricks; b < g->num_all_bricks; ++b) {
    const uint64_t k =
        BrickKeyOfCoords(g->brick_coords[3 * b], g->brick_coords[3 * b + 1], g->brick_coords[3 * b + 2]);
    m->apron_pos.emplace(k, static_cast<int32_t>(m->apron_keys.size()));
    m->apron_keys.push_back(k);
  }
  m->cap_all = std::max<int64_t>(g->num_all_bricks, 1);
  m->cap_index = 0;  // unknown: forces a reallocation when the index is first rebuilt
  m->cap_core = std::max<int64_t>(2 * g->num_bricks, 1024);
  HG_CUDA(cudaMalloc(&m->d_core, static_cast<size_t>(m->cap_core) * 512 * VoxelBytes(g)));
  HG_CUDA(cudaMemset(m->d_core, 0, static_cast<size_t>(m->cap_core) * 512 * VoxelBytes(g)));
  if (g->num_bricks > 0) {
    const int64_t nv = g->num_bricks * 512;
    if (g->dev.kind == kGridProb) {
      ExtractCoreKernel<unsigned short><<<Blocks(nv, 256), 256>>>(
          static_cast<const unsigned short*>(g->d_bricks), g->num_bricks, static_cast<unsigned short*>(m->d_core));
    } else {
      ExtractCoreKernel<ushort2><<<Blocks(nv, 256), 256>>>(static_cast<const ushort2*>(g->d_bricks),
                                                            g->num_bricks, static_cast<ushort2*>(m->d_core));
    }
    CountLaunch();
    HG_CUDA(cudaGetLastError());
  }
  HG_CUDA(cudaMallocHost(&m->h_pinned, 64));
  if (m->state.Ensure(sizeof(InsertState)) != HG_OK) return HG_ERR_CUDA;
  if (m->num.Ensure(2 * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  return HG_OK;
}

// Creates the core bricks named by 'keys' that do not exist yet and keeps the apron-only
// set consistent (a core brick needs its seven lower neighbours to exist in the all-brick
// list so that their aprons can show its low faces). Returns the number of new bricks.
int AddBricks(hg_grid* g, const std::vector<uint64_t>& keys, int64_t* added) {
  GridMutable* m = g->mut;
  *added = 0;
  for (const uint64_t k : keys) {
    if (m->core_ids.count(k)) continue;
    const int32_t id = static_cast<int32_t>(m->core_coords.size() / 3);
    int bx, by, bz;
    CoordsOfBrickKey(k, &bx, &by, &bz);
    m->core_coords.push_back(bx);
    m->core_coords.push_back(by);
    m->core_coords.push_back(bz);
    m->core_ids.emplace(k, id);
    ++*added;
    auto it = m->apron_pos.find(k);
    if (it != m->apron_pos.end()) {  // promoted from apron-only to core: swap-remove
      const int32_t pos = it->second;
      const uint64_t last = m->apron_keys.back();
      m->apron_keys[pos] = last;
      m->apron_pos[last] = pos;
      m->apron_keys.pop_back();
      m->apron_pos.erase(k);
    }
  }
  if (*added == 0) return HG_OK;
  // Lower neighbours of the new bricks (ids assigned above are the tail of core_coords).
  const int64_t nc = static_cast<int64_t>(m->core_coords.size() / 3);
  for (int64_t b = nc - *added; b < nc; ++b) {
    const int bx = m->core_coords[3 * b], by = m->core_coords[3 * b + 1], bz = m->core_coords[3 * b + 2];
    for (int d = 1; d < 8; ++d) {
      const int nx = bx - (d & 1), ny = by - ((d >> 1) & 1), nz = bz - ((d >> 2) & 1);
      if (nx < -1024 || ny < -1024 || nz < -1024) continue;
      const uint64_t k = BrickKeyOfCoords(nx, ny, nz);
      if (m->core_ids.count(k) || m->apron_pos.count(k)) continue;
      m->apron_pos.emplace(k, static_cast<int32_t>(m->apron_keys.size()));
      m->apron_keys.push_back(k);
    }
  }
  // Grow the core store (contents kept).
  if (nc > m->cap_core) {
    const int64_t new_cap = std::max<int64_t>(2 * m->cap_core, nc + nc / 2);
    void* p = nullptr;
    const size_t vb = VoxelBytes(g) * 512;
    HG_CUDA(cudaMalloc(&p, static_cast<size_t>(new_cap) * vb));
    HG_CUDA(cudaMemcpy(p, m->d_core, static_cast<size_t>(nc - *added) * vb, cudaMemcpyDeviceToDevice));
    HG_CUDA(cudaMemset(static_cast<char*>(p) + static_cast<size_t>(nc - *added) * vb, 0,
                       static_cast<size_t>(new_cap - (nc - *added)) * vb));
    cudaFree(m->d_core);
    m->d_core = p;
    m->cap_core = new_cap;
  }
  // All-brick list: core bricks first (ids stable), apron-only bricks after them.
  g->brick_coords = m->core_coords;
  g->brick_coords.reserve(m->core_coords.size() + 3 * m->apron_keys.size());
  for (const uint64_t k : m->apron_keys) {
    int bx, by, bz;
    CoordsOfBrickKey(k, &bx, &by, &bz);
    g->brick_coords.push_back(bx);
    g->brick_coords.push_back(by);
    g->brick_coords.push_back(bz);
  }
  g->num_bricks = nc;
  g->num_all_bricks = static_cast<int64_t>(g->brick_coords.size() / 3);
  // Index and read-side buffers.
  std::vector<int32_t> entries;
  BuildBrickIndexHost(g->brick_coords, &g->dev.table, &entries);
  const size_t index_bytes = entries.size() * sizeof(int32_t);
  if (index_bytes > m->cap_index) {
    if (g->d_index) cudaFree(g->d_index);
    g->d_index = nullptr;
    m->cap_index = index_bytes + index_bytes / 2;
    HG_CUDA(cudaMalloc(&g->d_index, m->cap_index));
  }
  HG_CUDA(cudaMemcpy(g->d_index, entries.data(), index_bytes, cudaMemcpyHostToDevice));
  if (g->num_all_bricks > m->cap_all) {
    m->cap_all = g->num_all_bricks + g->num_all_bricks / 2;
    if (g->d_bricks) cudaFree(g->d_bricks);
    if (g->d_decoded) cudaFree(g->d_decoded);
    g->d_bricks = g->d_decoded = nullptr;
    HG_CUDA(cudaMalloc(&g->d_bricks, static_cast<size_t>(m->cap_all) * kApronBrickVoxels * VoxelBytes(g)));
    HG_CUDA(cudaMalloc(&g->d_decoded, static_cast<size_t>(m->cap_all) * kApronBrickVoxels * sizeof(float)));
  }
  g->dev.table.index = static_cast<const int32_t*>(g->d_index);
  g->dev.table.bricks = g->d_bricks;
  g->dev.decoded = static_cast<const float*>(g->d_decoded);
  if (m->coords.Ensure(g->brick_coords.size() * sizeof(int32_t)) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cudaMemcpy(m->coords.p, g->brick_coords.data(), g->brick_coords.size() * sizeof(int32_t),
                     cudaMemcpyHostToDevice));
  return HG_OK;
}

int RebuildReadCopies(hg_grid* g) {
  DropStencils(g);  // rebuilt from the new decoded bricks by the next cost function that wants them
  GridMutable* m = g->mut;
  if (g->num_all_bricks == 0) return HG_OK;
  if (m->coords.cap == 0) {  // an uploaded grid written to without any new brick
    if (m->coords.Ensure(g->brick_coords.size() * sizeof(int32_t)) != HG_OK) return HG_ERR_CUDA;
    HG_CUDA(cudaMemcpy(m->coords.p, g->brick_coords.data(), g->brick_coords.size() * sizeof(int32_t),
                       cudaMemcpyHostToDevice));
  }
  const int64_t nv = g->num_all_bricks * kApronBrickVoxels;
  if (g->dev.kind == kGridProb) {
    BuildApronKernel<unsigned short><<<Blocks(nv, 256), 256>>>(
        g->dev, m->coords.as<int32_t>(), static_cast<const unsigned short*>(m->d_core),
        static_cast<int>(g->num_bricks), g->num_all_bricks, static_cast<unsigned short*>(g->d_bricks),
        static_cast<float*>(g->d_decoded));
  } else {
    BuildApronKernel<ushort2><<<Blocks(nv, 256), 256>>>(
        g->dev, m->coords.as<int32_t>(), static_cast<const ushort2*>(m->d_core), static_cast<int>(g->num_bricks),
        g->num_all_bricks, static_cast<ushort2*>(g->d_bricks), static_cast<float*>(g->d_decoded));
  }
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  return HG_OK;
}

// Uploads the returns as they are (one contiguous copy); the kernels read them with the stride.
int UploadReturns(GridMutable* m, const float* returns, const int stride, const int64_t n) {
  const size_t bytes = (static_cast<size_t>(n - 1) * stride + 3) * sizeof(float);
  if (m->points.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cudaMemcpy(m->points.p, returns, bytes, cudaMemcpyHostToDevice));
  return HG_OK;
}

struct WidenCount {
  __host__ __device__ unsigned long long operator()(const int v) const { return static_cast<unsigned long long>(v); }
};

// Exclusive scan of counts[0..n) into offsets[0..n]; returns the total.
int ScanCounts(GridMutable* m, const int64_t n, int* total) {
  // counts has n + 1 entries with counts[n] = 0, so that offsets[n] is the total.
  HG_CUDA(cudaMemset(m->counts.as<int>() + n, 0, sizeof(int)));
  size_t bytes = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, bytes, m->counts.as<int>(), m->offsets.as<int>(), static_cast<int>(n + 1));
  if (m->temp.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cub::DeviceScan::ExclusiveSum(m->temp.p, bytes, m->counts.as<int>(), m->offsets.as<int>(),
                                        static_cast<int>(n + 1)));
  CountLaunch();
  HG_CUDA(cudaMemcpy(total, m->offsets.as<int>() + n, sizeof(int), cudaMemcpyDeviceToHost));
  // The offsets are 32-bit: a scan whose updates do not fit is refused instead of wrapping around
  // (the sum is taken again in 64 bits).
  if (m->num.Ensure(sizeof(unsigned long long)) != HG_OK) return HG_ERR_CUDA;
  thrust::transform_iterator<WidenCount, const int*> wide(m->counts.as<int>(), WidenCount());
  unsigned long long* d_sum = reinterpret_cast<unsigned long long*>(m->num.p);
  bytes = 0;
  cub::DeviceReduce::Sum(nullptr, bytes, wide, d_sum, static_cast<int>(n));
  if (m->temp.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cub::DeviceReduce::Sum(m->temp.p, bytes, wide, d_sum, static_cast<int>(n)));
  CountLaunch();
  unsigned long long sum = 0;
  HG_CUDA(cudaMemcpy(&sum, d_sum, sizeof(sum), cudaMemcpyDeviceToHost));
  if (sum >= (1ull << 31)) {
    SetError("range-data insertion: " + std::to_string(sum) + " cell updates in one scan exceed the 2^31 - 1 this path indexes");
    return HG_ERR_LIMIT;
  }
  return HG_OK;
}

// Positions of run heads (flags != 0) -> out, count -> *num.
int SelectHeads(GridMutable* m, const int n, int* out, int* num) {
  size_t bytes = 0;
  thrust::counting_iterator<int> it(0);
  cub::DeviceSelect::Flagged(nullptr, bytes, it, m->flags.as<unsigned char>(), out, m->num.as<int>(), n);
  if (m->temp.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cub::DeviceSelect::Flagged(m->temp.p, bytes, it, m->flags.as<unsigned char>(), out, m->num.as<int>(), n));
  CountLaunch();
  HG_CUDA(cudaMemcpy(num, m->num.p, sizeof(int), cudaMemcpyDeviceToHost));
  return HG_OK;
}

// After the sort: unique cells (starts), touched bricks -> created on the host.
// KEY_SHIFT = bits below the cell key (1 for occupancy keys, 0 for TSDF keys).
template <int KEY_SHIFT>
int GroupAndCreateBricks(hg_grid* g, const uint64_t* keys, const int total, int* num_cells) {
  GridMutable* m = g->mut;
  if (m->flags.Ensure(static_cast<size_t>(total)) != HG_OK) return HG_ERR_CUDA;
  if (m->starts.Ensure(static_cast<size_t>(total) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  HeadFlagsKernel<KEY_SHIFT><<<Blocks(total, 256), 256>>>(keys, total, m->flags.as<unsigned char>());
  CountLaunch();
  int rc = SelectHeads(m, total, m->starts.as<int>(), num_cells);
  if (rc != HG_OK) return rc;
  if (m->brick_heads.Ensure(static_cast<size_t>(*num_cells) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  BrickFlagsKernel<KEY_SHIFT + 9><<<Blocks(*num_cells, 256), 256>>>(keys, m->starts.as<int>(), *num_cells,
                                                                     m->flags.as<unsigned char>());
  CountLaunch();
  int num_touched = 0;
  rc = SelectHeads(m, *num_cells, m->brick_heads.as<int>(), &num_touched);
  if (rc != HG_OK) return rc;
  if (m->brick_keys.Ensure(static_cast<size_t>(num_touched) * sizeof(uint64_t)) != HG_OK) return HG_ERR_CUDA;
  GatherBrickKeysKernel<KEY_SHIFT + 9><<<Blocks(num_touched, 256), 256>>>(
      keys, m->starts.as<int>(), m->brick_heads.as<int>(), num_touched, m->brick_keys.as<uint64_t>());
  CountLaunch();
  std::vector<uint64_t> touched(static_cast<size_t>(num_touched));
  HG_CUDA(cudaMemcpy(touched.data(), m->brick_keys.p, touched.size() * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  int64_t added = 0;
  return AddBricks(g, touched, &added);
}

// Reads the counters back, grows grid_size like DynamicGrid::Grow and refreshes the read copies.
int FinishInsertion(hg_grid* g, const char* who) {
  GridMutable* m = g->mut;
  InsertState st;
  HG_CUDA(cudaMemcpy(&st, m->state.p, sizeof(st), cudaMemcpyDeviceToHost));
  if (st.error & 2) {
    SetError(std::string(who) + ": internal error, a touched brick was not created");
    return HG_ERR_CUDA;
  }
  g->num_voxels += st.new_voxels;
  while (g->grid_size / 2 < st.need && g->grid_size < (64 << 8)) g->grid_size *= 2;
  return RebuildReadCopies(g);
}

}  // namespace
}  // namespace hg

using namespace hg;

extern "C" {

int hg_grid_insert_tsdf(hg_grid* grid, const hg_tsdf_inserter_options* options, const float origin[3],
                        const float* returns, int stride_floats, int64_t num_returns, int64_t* num_updates) {
  HG_REQUIRE(grid != nullptr && options != nullptr && origin != nullptr && num_returns >= 0 && stride_floats >= 3,
             "hg_grid_insert_tsdf: bad arguments");
  HG_REQUIRE(grid->dev.kind == kGridTsdf, "hg_grid_insert_tsdf: not a TSDF grid");
  HG_REQUIRE(num_returns == 0 || returns != nullptr, "hg_grid_insert_tsdf: null returns");
  HG_REQUIRE(num_returns < (int64_t(1) << 27), "hg_grid_insert_tsdf: too many returns for one call");
  if (num_updates) *num_updates = 0;
  if (num_returns == 0) return HG_OK;
  HG_ON_DEVICE(grid->device);
  int rc = MakeMutable(grid);
  if (rc != HG_OK) return rc;
  GridMutable* m = grid->mut;
  const int n = static_cast<int>(num_returns);

  RayParams p;
  memset(&p, 0, sizeof(p));
  p.stride = stride_floats;
  p.ox = origin[0];
  p.oy = origin[1];
  p.oz = origin[2];
  p.resolution = grid->dev.resolution;
  p.inv_resolution = grid->dev.inv_resolution;
  // options_.relative_truncation_distance() * tsdf->resolution(): double * float, narrowed to float.
  p.truncation_distance =
      static_cast<float>(options->relative_truncation_distance * static_cast<double>(grid->dev.resolution));
  p.update_free_space = options->num_free_space_voxels > 0 ? 1 : 0;
  p.epsilon = static_cast<float>(options->weight_function_epsilon);
  p.sigma = static_cast<float>(options->weight_function_sigma);
  p.min_range = options->min_range;
  p.max_range = options->max_range;

  PhaseTrace trace;
  rc = UploadReturns(m, returns, stride_floats, n);
  if (rc != HG_OK) return rc;
  trace.Mark("upload");
  // The insertion-ratio counter (:701-708) depends only on the position in the cloud.
  const unsigned char* d_mask = nullptr;
  if (options->insertion_ratio < 1.0) {
    std::vector<unsigned char> mask(static_cast<size_t>(n));
    size_t inserted = 0, omitted = 0;
    for (int i = 0; i < n; ++i) {
      if (static_cast<double>(inserted) <= options->insertion_ratio * static_cast<double>(inserted + omitted)) {
        ++inserted;
        mask[i] = 1;
      } else {
        ++omitted;
        mask[i] = 0;
      }
    }
    if (m->mask.Ensure(mask.size()) != HG_OK) return HG_ERR_CUDA;
    HG_CUDA(cudaMemcpy(m->mask.p, mask.data(), mask.size(), cudaMemcpyHostToDevice));
    d_mask = m->mask.as<unsigned char>();
  }
  if (m->counts.Ensure(static_cast<size_t>(n + 1) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  if (m->offsets.Ensure(static_cast<size_t>(n + 1) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cudaMemset(m->state.p, 0, sizeof(InsertState)));
  TsdfCountKernel<<<Blocks(n, 256), 256>>>(p, m->points.as<float>(), d_mask, n, m->counts.as<int>());
  CountLaunch();
  int total = 0;
  rc = ScanCounts(m, n, &total);
  if (rc != HG_OK) return rc;
  if (total == 0) return HG_OK;
  HG_REQUIRE(total > 0, "hg_grid_insert_tsdf: more than 2^31 cell updates in one call");
  trace.Mark("count+scan");

  for (int k = 0; k < 2; ++k) {
    if (m->keys[k].Ensure(static_cast<size_t>(total) * sizeof(uint64_t)) != HG_OK) return HG_ERR_CUDA;
    if (m->vals[k].Ensure(static_cast<size_t>(total) * sizeof(uint64_t)) != HG_OK) return HG_ERR_CUDA;
  }
  TsdfEmitKernel<<<Blocks(total, 256), 256>>>(p, m->points.as<float>(), m->offsets.as<int>(), n, total,
                                               m->keys[0].as<uint64_t>(), m->vals[0].as<float2>(),
                                               m->state.as<InsertState>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  int err = 0;
  HG_CUDA(cudaMemcpy(&err, m->state.p, sizeof(int), cudaMemcpyDeviceToHost));
  HG_REQUIRE(err == 0, "hg_grid_insert_tsdf: a ray leaves the largest grid (cells beyond +-8192); the reference "
                       "CHECK-fails in DynamicGrid::Grow");
  trace.Mark("emit");

  cub::DoubleBuffer<uint64_t> keys(m->keys[0].as<uint64_t>(), m->keys[1].as<uint64_t>());
  // The (tsd, weight) pair travels through the sort as one 64-bit value; the sort is stable, so
  // the updates of a cell stay in sequence order.
  cub::DoubleBuffer<uint64_t> vals(m->vals[0].as<uint64_t>(), m->vals[1].as<uint64_t>());
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys, vals, total, 0, kCellKeyBits);
  if (m->temp.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cub::DeviceRadixSort::SortPairs(m->temp.p, bytes, keys, vals, total, 0, kCellKeyBits));
  CountLaunch(4);
  trace.Mark("sort");

  int num_cells = 0;
  rc = GroupAndCreateBricks<0>(grid, keys.Current(), total, &num_cells);
  if (rc != HG_OK) return rc;
  trace.Mark("group+bricks");

  TsdfApplyParams ap;
  ap.tsd = grid->dev.a;
  ap.weight = grid->dev.b;
  ap.min_tsd = grid->dev.min_tsd;
  ap.max_tsd = grid->dev.max_tsd;
  ap.tsd_resolution = 32766.f / (grid->dev.max_tsd - grid->dev.min_tsd);  // tsd_value_converter.cc:27
  ap.max_weight = grid->max_weight;
  ap.weight_resolution = 32766.f / (grid->max_weight - 0.f);              // :28
  ap.maximum_weight = static_cast<float>(options->maximum_weight);
  TsdfApplyKernel<<<Blocks(num_cells, 128), 128>>>(grid->dev.table, ap, keys.Current(),
                                                    reinterpret_cast<const float2*>(vals.Current()),
                                                    m->starts.as<int>(), num_cells, total,
                                                    static_cast<ushort2*>(m->d_core), m->state.as<InsertState>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  trace.Mark("apply");
  rc = FinishInsertion(grid, "hg_grid_insert_tsdf");
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaDeviceSynchronize());
  trace.Mark("rebuild");
  if (num_updates) *num_updates = total;
  return HG_OK;
}

int hg_grid_insert_occupancy(hg_grid* grid, const hg_occupancy_inserter_options* options,
                             const float origin[3], const float* returns, int stride_floats,
                             int64_t num_returns) {
  HG_REQUIRE(grid != nullptr && options != nullptr && origin != nullptr && num_returns >= 0 && stride_floats >= 3,
             "hg_grid_insert_occupancy: bad arguments");
  HG_REQUIRE(grid->dev.kind == kGridProb, "hg_grid_insert_occupancy: not a probability grid");
  HG_REQUIRE(num_returns == 0 || returns != nullptr, "hg_grid_insert_occupancy: null returns");
  HG_REQUIRE(num_returns < (int64_t(1) << 27), "hg_grid_insert_occupancy: too many returns for one call");
  // CreateProbabilityGridRangeDataInserterOptions3D CHECKs these (range_data_inserter_3d.cc:64-65).
  HG_REQUIRE(options->hit_probability > 0.5 && options->miss_probability < 0.5 &&
                 options->hit_probability < 1.0 && options->miss_probability > 0.0,
             "hg_grid_insert_occupancy: need miss_probability < 0.5 < hit_probability");
  if (num_returns == 0) return HG_OK;
  HG_ON_DEVICE(grid->device);
  int rc = MakeMutable(grid);
  if (rc != HG_OK) return rc;
  GridMutable* m = grid->mut;
  const int n = static_cast<int>(num_returns);

  // ComputeLookupTableToApplyOdds (probability_values.cc:76-87) for both odds, in float.
  const float hit_p = static_cast<float>(options->hit_probability);
  const float miss_p = static_cast<float>(options->miss_probability);
  if (m->d_tables == nullptr || m->table_hit != hit_p || m->table_miss != miss_p) {
    const float kMin = 0.1f, kMax = 1.f - kMin;
    auto to_value = [&](float prob) {
      if (prob > kMax) prob = kMax;
      if (prob < kMin) prob = kMin;
      return static_cast<uint16_t>(std::lround((prob - kMin) * (32766.f / (kMax - kMin))) + 1);
    };
    std::vector<uint16_t> tables(2 * 32768);
    for (int t = 0; t < 2; ++t) {
      const float pr = t == 0 ? hit_p : miss_p;
      const float odds = pr / (1.f - pr);
      const float o0 = odds;
      tables[t * 32768] = static_cast<uint16_t>(to_value(o0 / (o0 + 1.f)) + 0x8000);
      const float scale = (kMax - kMin) / 32766.f;
      for (int cell = 1; cell < 32768; ++cell) {
        const float cell_p = cell * scale + (kMin - scale);
        const float o = odds * (cell_p / (1.f - cell_p));
        tables[t * 32768 + cell] = static_cast<uint16_t>(to_value(o / (o + 1.f)) + 0x8000);
      }
    }
    if (m->d_tables == nullptr) HG_CUDA(cudaMalloc(&m->d_tables, tables.size() * sizeof(uint16_t)));
    HG_CUDA(cudaMemcpy(m->d_tables, tables.data(), tables.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    m->table_hit = hit_p;
    m->table_miss = miss_p;
  }

  RayParams p;
  memset(&p, 0, sizeof(p));
  p.stride = stride_floats;
  p.ox = origin[0];
  p.oy = origin[1];
  p.oz = origin[2];
  p.resolution = grid->dev.resolution;
  p.inv_resolution = grid->dev.inv_resolution;
  p.num_free_space_voxels = options->num_free_space_voxels;

  rc = UploadReturns(m, returns, stride_floats, n);
  if (rc != HG_OK) return rc;
  if (m->counts.Ensure(static_cast<size_t>(n + 1) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  if (m->offsets.Ensure(static_cast<size_t>(n + 1) * sizeof(int)) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cudaMemset(m->state.p, 0, sizeof(InsertState)));
  OccupancyCountKernel<<<Blocks(n, 256), 256>>>(p, m->points.as<float>(), n, m->counts.as<int>());
  CountLaunch();
  int total = 0;
  rc = ScanCounts(m, n, &total);
  if (rc != HG_OK) return rc;
  HG_REQUIRE(total > 0, "hg_grid_insert_occupancy: more than 2^31 cell updates in one call");
  for (int k = 0; k < 2; ++k) {
    if (m->keys[k].Ensure(static_cast<size_t>(total) * sizeof(uint64_t)) != HG_OK) return HG_ERR_CUDA;
  }
  OccupancyEmitKernel<<<Blocks(total, 256), 256>>>(p, m->points.as<float>(), m->offsets.as<int>(), n, total,
                                                    m->keys[0].as<uint64_t>(), m->state.as<InsertState>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  int err = 0;
  HG_CUDA(cudaMemcpy(&err, m->state.p, sizeof(int), cudaMemcpyDeviceToHost));
  HG_REQUIRE(err == 0, "hg_grid_insert_occupancy: a return leaves the largest grid (cells beyond +-8192); the "
                       "reference CHECK-fails in DynamicGrid::Grow");
  cub::DoubleBuffer<uint64_t> keys(m->keys[0].as<uint64_t>(), m->keys[1].as<uint64_t>());
  size_t bytes = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, bytes, keys, total, 0, kCellKeyBits + 1);
  if (m->temp.Ensure(bytes) != HG_OK) return HG_ERR_CUDA;
  HG_CUDA(cub::DeviceRadixSort::SortKeys(m->temp.p, bytes, keys, total, 0, kCellKeyBits + 1));
  CountLaunch(4);
  int num_cells = 0;
  rc = GroupAndCreateBricks<1>(grid, keys.Current(), total, &num_cells);
  if (rc != HG_OK) return rc;
  OccupancyApplyKernel<<<Blocks(num_cells, 128), 128>>>(grid->dev.table, keys.Current(), m->starts.as<int>(),
                                                         num_cells, m->d_tables,
                                                         static_cast<unsigned short*>(m->d_core),
                                                         m->state.as<InsertState>());
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  rc = FinishInsertion(grid, "hg_grid_insert_occupancy");
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaDeviceSynchronize());
  return HG_OK;
}

int hg_grid_export(const hg_grid* grid, int64_t max_voxels, int64_t* num_voxels, int32_t* xyz,
                   uint16_t* values_a, uint16_t* values_b) {
  HG_REQUIRE(grid != nullptr && num_voxels != nullptr && max_voxels >= 0, "hg_grid_export: bad arguments");
  HG_ON_DEVICE(grid->device);
  const bool tsdf = grid->dev.kind == kGridTsdf;
  const size_t vb = tsdf ? sizeof(ushort2) : sizeof(uint16_t);
  const int64_t nb = grid->num_bricks;
  std::vector<unsigned char> host(static_cast<size_t>(nb) * kApronBrickVoxels * vb);
  if (nb > 0) HG_CUDA(cudaMemcpy(host.data(), grid->d_bricks, host.size(), cudaMemcpyDeviceToHost));
  int64_t count = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int64_t k = 0;
    for (int64_t b = 0; b < nb; ++b) {
      const int bx = grid->brick_coords[3 * b] * 8, by = grid->brick_coords[3 * b + 1] * 8,
                bz = grid->brick_coords[3 * b + 2] * 8;
      for (int z = 0; z < 8; ++z)
        for (int y = 0; y < 8; ++y)
          for (int x = 0; x < 8; ++x) {
            const size_t at = static_cast<size_t>(b) * kApronBrickVoxels + ((z * 9 + y) * 9 + x);
            uint16_t a, w = 0;
            if (tsdf) {
              a = reinterpret_cast<const ushort2*>(host.data())[at].x;
              w = reinterpret_cast<const ushort2*>(host.data())[at].y;
            } else {
              a = reinterpret_cast<const uint16_t*>(host.data())[at];
            }
            if (a == 0 && w == 0) continue;
            if (pass == 1) {
              xyz[3 * k] = bx + x;
              xyz[3 * k + 1] = by + y;
              xyz[3 * k + 2] = bz + z;
              values_a[k] = a;
              if (tsdf && values_b) values_b[k] = w;
            }
            ++k;
          }
    }
    count = k;
    if (pass == 0) {
      *num_voxels = count;
      if (count > max_voxels || count == 0) return HG_OK;
      HG_REQUIRE(xyz != nullptr && values_a != nullptr, "hg_grid_export: null output arrays");
    }
  }
  return HG_OK;
}

}  // extern "C"
