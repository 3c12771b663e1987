// Submap storage on the device: HybridGrid / HybridGridTSDF flattened to a
// brick table (see hg_internal.cuh), plus library-wide plumbing.
//
// Reference behaviour reproduced here:
//   mapping/3d/hybrid_grid_base.h:263-279  value(): default outside the grid / absent bricks
//   mapping/3d/hybrid_grid_base.h:40-43    z-major voxel order inside an 8^3 block
//   mapping/3d/hybrid_grid.h:133-135, hybrid_grid_tsdf.h:105-113  value -> float
//   mapping/value_conversion_tables.cc:29-51, 2d/tsd_value_converter.cc:22-32  decoding constants
#include <algorithm>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "device_math.cuh"
#include "interpolation.cuh"

namespace hg {

static thread_local std::string t_error;
std::atomic<uint64_t> g_kernel_launches{0};

void SetError(const std::string& msg) { t_error = msg; }

int BuildBrickIndexHost(const std::vector<int32_t>& brick_coords, BrickTable* table,
                        std::vector<int32_t>* entries) {
  const int64_t nb = static_cast<int64_t>(brick_coords.size() / 3);
  int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
  for (int64_t i = 0; i < nb; ++i) {
    for (int k = 0; k < 3; ++k) {
      const int c = brick_coords[3 * i + k];
      if (i == 0 || c < lo[k]) lo[k] = c;
      if (i == 0 || c > hi[k]) hi[k] = c;
    }
  }
  int64_t cells = 1;
  for (int k = 0; k < 3; ++k) {
    table->bmin[k] = lo[k];
    table->bdim[k] = nb == 0 ? 1 : hi[k] - lo[k] + 1;
    cells *= table->bdim[k];
  }
  table->hash_mask = 0;
  // Dense index while it stays within 64 Mi entries (256 MB) and within 64x the
  // brick count + slack; otherwise an open-addressing hash table.
  if (cells > (int64_t(1) << 26) || cells > 64 * nb + (int64_t(1) << 20)) {
    unsigned cap = 16;
    while (cap < 2 * static_cast<uint64_t>(nb) + 1) cap <<= 1;
    entries->assign(static_cast<size_t>(cap) * 4, -1);
    for (int64_t i = 0; i < nb; ++i) {
      const int bx = brick_coords[3 * i], by = brick_coords[3 * i + 1], bz = brick_coords[3 * i + 2];
      unsigned h = BrickHash(bx, by, bz) & (cap - 1);
      while ((*entries)[4 * static_cast<size_t>(h) + 3] >= 0) h = (h + 1) & (cap - 1);
      (*entries)[4 * static_cast<size_t>(h)] = bx;
      (*entries)[4 * static_cast<size_t>(h) + 1] = by;
      (*entries)[4 * static_cast<size_t>(h) + 2] = bz;
      (*entries)[4 * static_cast<size_t>(h) + 3] = static_cast<int32_t>(i);
    }
    table->hash_mask = cap - 1;
    return HG_OK;
  }
  entries->assign(static_cast<size_t>(cells), -1);
  for (int64_t i = 0; i < nb; ++i) {
    const int64_t bx = brick_coords[3 * i] - lo[0];
    const int64_t by = brick_coords[3 * i + 1] - lo[1];
    const int64_t bz = brick_coords[3 * i + 2] - lo[2];
    (*entries)[(bz * table->bdim[1] + by) * table->bdim[0] + bx] = static_cast<int32_t>(i);
  }
  return HG_OK;
}

int UploadBrickIndex(const std::vector<int32_t>& brick_coords, int brick_bytes, BrickTable* table,
                     void** d_index, void** d_bricks) {
  const int64_t nb = static_cast<int64_t>(brick_coords.size() / 3);
  std::vector<int32_t> entries;
  BuildBrickIndexHost(brick_coords, table, &entries);
  HG_CUDA(cudaMalloc(d_index, entries.size() * sizeof(int32_t)));
  HG_CUDA(cudaMemcpy(*d_index, entries.data(), entries.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
  const size_t all_bytes = static_cast<size_t>(std::max<int64_t>(nb, 1)) * brick_bytes;
  HG_CUDA(cudaMalloc(d_bricks, all_bytes));
  HG_CUDA(cudaMemset(*d_bricks, 0, all_bytes));
  table->index = static_cast<const int32_t*>(*d_index);
  table->bricks = *d_bricks;
  return HG_OK;
}

namespace {

struct BrickKey {
  int x, y, z;
  bool operator==(const BrickKey& o) const { return x == o.x && y == o.y && z == o.z; }
};
struct BrickKeyHash {
  size_t operator()(const BrickKey& k) const {
    uint64_t h = static_cast<uint32_t>(k.x) * 0x9E3779B97F4A7C15ull;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.y)) + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
    h ^= (static_cast<uint64_t>(static_cast<uint32_t>(k.z)) + 0x165667B1ull) * 0xD6E8FEB86659FD93ull;
    return static_cast<size_t>(h ^ (h >> 29));
  }
};

// Smallest reference grid_size (64 << bits, bits >= 1) whose symmetric index
// range contains every voxel (DynamicGrid::Grow, hybrid_grid_base.h:385-403).
int DeriveGridSize(int64_t n, const int32_t* xyz) {
  int bits = 1;
  for (int64_t i = 0; i < 3 * n; ++i) {
    while (bits < 8) {
      const int half = (64 << bits) >> 1;
      if (xyz[i] >= -half && xyz[i] < half) break;
      ++bits;
    }
  }
  return 64 << bits;
}

template <typename Voxel, typename MakeVoxel>
int Flatten(int64_t n, const int32_t* xyz, MakeVoxel make, hg_grid* g) {
  std::unordered_map<BrickKey, int32_t, BrickKeyHash> ids;
  ids.reserve(static_cast<size_t>(n / 16 + 16));
  std::vector<int32_t>& coords = g->brick_coords;
  auto brick_id = [&](const BrickKey& k) {
    auto it = ids.find(k);
    if (it == ids.end()) {
      const int32_t id = static_cast<int32_t>(coords.size() / 3);
      coords.push_back(k.x);
      coords.push_back(k.y);
      coords.push_back(k.z);
      it = ids.emplace(k, id).first;
    }
    return it->second;
  };
  // Pass 1: core bricks, in order of first appearance (these come first in the list).
  std::vector<int32_t> brick_of(static_cast<size_t>(n));
  for (int64_t i = 0; i < n; ++i) {
    brick_of[static_cast<size_t>(i)] = brick_id(BrickKey{xyz[3 * i] >> 3, xyz[3 * i + 1] >> 3, xyz[3 * i + 2] >> 3});
  }
  const int64_t num_core = static_cast<int64_t>(coords.size() / 3);
  // Pass 2: a voxel on a low face of its brick is also the apron of the brick(s) before it.
  for (int64_t i = 0; i < n; ++i) {
    const int x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if ((x & 7) && (y & 7) && (z & 7)) continue;
    for (int dz = 0; dz <= ((z & 7) == 0 ? 1 : 0); ++dz)
      for (int dy = 0; dy <= ((y & 7) == 0 ? 1 : 0); ++dy)
        for (int dx = 0; dx <= ((x & 7) == 0 ? 1 : 0); ++dx)
          if (dx | dy | dz) brick_id(BrickKey{(x >> 3) - dx, (y >> 3) - dy, (z >> 3) - dz});
  }
  const int64_t nb = static_cast<int64_t>(coords.size() / 3);
  std::vector<Voxel> bricks(static_cast<size_t>(std::max<int64_t>(nb, 1)) * kApronBrickVoxels);
  std::memset(bricks.data(), 0, bricks.size() * sizeof(Voxel));
  for (int64_t i = 0; i < n; ++i) {
    const int x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    const Voxel v = make(i);
    bricks[static_cast<size_t>(brick_of[static_cast<size_t>(i)]) * kApronBrickVoxels +
           (((z & 7) * 9 + (y & 7)) * 9 + (x & 7))] = v;
    if ((x & 7) && (y & 7) && (z & 7)) continue;
    for (int dz = 0; dz <= ((z & 7) == 0 ? 1 : 0); ++dz)
      for (int dy = 0; dy <= ((y & 7) == 0 ? 1 : 0); ++dy)
        for (int dx = 0; dx <= ((x & 7) == 0 ? 1 : 0); ++dx) {
          if (!(dx | dy | dz)) continue;
          const int32_t id = ids.find(BrickKey{(x >> 3) - dx, (y >> 3) - dy, (z >> 3) - dz})->second;
          const int lx = dx ? 8 : (x & 7), ly = dy ? 8 : (y & 7), lz = dz ? 8 : (z & 7);
          bricks[static_cast<size_t>(id) * kApronBrickVoxels + ((lz * 9 + ly) * 9 + lx)] = v;
        }
  }
  const int rc = UploadBrickIndex(coords, static_cast<int>(sizeof(Voxel)) * kApronBrickVoxels, &g->dev.table,
                                  &g->d_index, &g->d_bricks);
  if (rc != HG_OK) return rc;
  HG_CUDA(cudaMemcpy(g->d_bricks, bricks.data(), bricks.size() * sizeof(Voxel), cudaMemcpyHostToDevice));
  g->num_bricks = num_core;
  g->num_all_bricks = nb;
  g->num_voxels = n;
  return HG_OK;
}

// Decoding constants computed exactly as SlowValueToBoundedFloat does.
ValueDecode MakeDecode(float unknown, float lo, float hi) {
  ValueDecode d;
  const float scale = (hi - lo) / 32766.f;
  d.scale = scale;
  d.offset = lo - scale;
  d.unknown = unknown;
  return d;
}

// Fills GridDev::decoded from the raw bricks (same float operations as Decode).
__global__ void __launch_bounds__(256) DecodeBricksKernel(const GridDev g, const int64_t num_voxels,
                                                          float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= num_voxels) return;
  if (g.kind == kGridProb) {
    out[i] = Decode(g.a, static_cast<const unsigned short*>(g.table.bricks)[i]);
  } else {
    const ushort2 raw = static_cast<const ushort2*>(g.table.bricks)[i];
    out[i] = (raw.y & 0x7FFFu) > 1u ? Decode(g.a, raw.x) : __int_as_float(0x7fc00000);
  }
}

__global__ void LookupKernel(const GridDev g, const int what, const int64_t n,
                             const int32_t* __restrict__ xyz, float* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const int x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
  float v;
  if (g.kind == kGridProb) {
    v = Decode(g.a, LoadU16(g.table, x, y, z));
  } else {
    const ushort2 t = LoadTsdf(g.table, x, y, z);
    v = what == 1 ? Decode(g.a, t.x) : Decode(g.b, t.y);
  }
  out[i] = v;
}

// InterpolatedTSDF::GetWeight (interpolated_tsdf.h:129-156): plain trilinear blend of the eight
// corner weights around the lower voxel, z then y then x, each step (b - a) * n + a in double with
// no fused multiply-add; the normalised coordinates are true divisions (the reference's double
// instantiation). InterpolatedMultiResolutionTSDF::GetWeight
// (interpolated_multi_resolution_tsdf.h:139-171) returns from the first iteration of its loop, so
// it is this lookup on the finest grid of the pyramid.
__global__ void InterpolatedWeightKernel(const GridDev g, const int64_t n, const double* __restrict__ p,
                                         double* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const double x = p[3 * i], y = p[3 * i + 1], z = p[3 * i + 2];
  double x1, y1, z1, x2, y2, z2;
  int ix, iy, iz;
  LowerVoxel(g, x, &x1, &x2, &ix);
  LowerVoxel(g, y, &y1, &y2, &iy);
  LowerVoxel(g, z, &z1, &z2, &iz);
  double q[8];  // corner order 111,112,121,122,211,212,221,222
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const ushort2 t = LoadTsdf(g.table, ix + (k >> 2), iy + ((k >> 1) & 1), iz + (k & 1));
    q[k] = static_cast<double>(Decode(g.b, t.y));
  }
  const double nx = __ddiv_rn(__dsub_rn(x, x1), __dsub_rn(x2, x1));
  const double ny = __ddiv_rn(__dsub_rn(y, y1), __dsub_rn(y2, y1));
  const double nz = __ddiv_rn(__dsub_rn(z, z1), __dsub_rn(z2, z1));
  const double q11 = __dadd_rn(__dmul_rn(__dsub_rn(q[1], q[0]), nz), q[0]);
  const double q12 = __dadd_rn(__dmul_rn(__dsub_rn(q[3], q[2]), nz), q[2]);
  const double q21 = __dadd_rn(__dmul_rn(__dsub_rn(q[5], q[4]), nz), q[4]);
  const double q22 = __dadd_rn(__dmul_rn(__dsub_rn(q[7], q[6]), nz), q[6]);
  const double q1 = __dadd_rn(__dmul_rn(__dsub_rn(q12, q11), ny), q11);
  const double q2 = __dadd_rn(__dmul_rn(__dsub_rn(q22, q21), ny), q21);
  out[i] = __dadd_rn(__dmul_rn(__dsub_rn(q2, q1), nx), q1);
}

__global__ void PackStencilsKernel(const float* __restrict__ decoded, const int64_t n, float4* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  const int v = static_cast<int>(i & 511);
  const float* base = decoded + (i >> 9) * kApronBrickVoxels + (((v >> 6) * 9 + ((v >> 3) & 7)) * 9 + (v & 7));
  out[2 * i] = make_float4(base[0], base[81], base[9], base[90]);
  out[2 * i + 1] = make_float4(base[1], base[82], base[10], base[91]);
}

__global__ void CellIndexKernel(const float resolution, const float inv_resolution, const int64_t n,
                                const float* __restrict__ p, int32_t* __restrict__ out) {
  const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
  if (i >= 3 * n) return;
  out[i] = CellIndexFast(p[i], resolution, inv_resolution);
}

}  // namespace
}  // namespace hg

using namespace hg;

static int BuildDecodedBricks(hg_grid* g) {
  const int64_t nv = std::max<int64_t>(g->num_all_bricks, 1) * kApronBrickVoxels;
  HG_CUDA(cudaMalloc(&g->d_decoded, static_cast<size_t>(nv) * sizeof(float)));
  g->dev.decoded = static_cast<const float*>(g->d_decoded);
  DecodeBricksKernel<<<static_cast<unsigned>((nv + 255) / 256), 256>>>(g->dev, nv, static_cast<float*>(g->d_decoded));
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaDeviceSynchronize());
  return HG_OK;
}


namespace hg {

static std::mutex g_stencil_mutex;

int EnsureStencils(const hg_grid* grid_const) {
  hg_grid* g = const_cast<hg_grid*>(grid_const);
  if (__atomic_load_n(&g->dev.stencils, __ATOMIC_ACQUIRE) != nullptr) return HG_OK;
  std::lock_guard<std::mutex> lock(g_stencil_mutex);
  if (g->dev.stencils != nullptr || g->num_all_bricks == 0 || g->d_decoded == nullptr) return HG_OK;
  HG_ON_DEVICE(g->device);
  const int64_t n = g->num_all_bricks * 512;
  if (g->stencil_capacity < n) {  // grow-only: a grid that is written between uses keeps its buffer
    if (g->d_stencils) cudaFree(g->d_stencils);
    g->d_stencils = nullptr;
    g->stencil_capacity = 0;
    const int64_t cap = n + n / 2;
    HG_CUDA(cudaMalloc(&g->d_stencils, static_cast<size_t>(cap) * 32));
    g->stencil_capacity = cap;
  }
  cudaStream_t stream = ThreadStream();
  HG_REQUIRE(stream != nullptr, "EnsureStencils: cannot create a stream");
  // the decoded bricks were written on the legacy default stream (upload / insertion), which a
  // non-blocking stream does not wait for
  HG_CUDA(cudaStreamSynchronize(nullptr));
  PackStencilsKernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(
      static_cast<const float*>(g->d_decoded), n, static_cast<float4*>(g->d_stencils));
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaStreamSynchronize(stream));
  __atomic_store_n(&g->dev.stencils, static_cast<const float4*>(g->d_stencils), __ATOMIC_RELEASE);
  return HG_OK;
}

// The grid's voxels are about to change: the packed copy is stale (its buffer is kept).
void DropStencils(hg_grid* g) {
  std::lock_guard<std::mutex> lock(g_stencil_mutex);
  g->dev.stencils = nullptr;
}

}  // namespace hg

extern "C" {

const char* hg_last_error(void) { return hg::t_error.c_str(); }
const char* hg_version(void) { return "hectorgrapher_b200 0.1 (sm_100a)"; }
uint64_t hg_kernel_launch_count(void) { return g_kernel_launches.load(); }

int hg_set_device(int device) {
  HG_CUDA(cudaSetDevice(device));
  return HG_OK;
}

int hg_grid_upload_prob(float resolution, int grid_size, int64_t num_voxels, const int32_t* xyz,
                        const uint16_t* values, hg_grid** out) {
  HG_REQUIRE(out != nullptr && resolution > 0.f && num_voxels >= 0, "hg_grid_upload_prob: bad arguments");
  HG_REQUIRE(num_voxels == 0 || (xyz != nullptr && values != nullptr), "hg_grid_upload_prob: null voxel arrays");
  hg_grid* g = new hg_grid();
  HG_CUDA(cudaGetDevice(&g->device));
  g->dev.kind = kGridProb;
  g->dev.resolution = resolution;
  g->dev.inv_resolution = 1.f / resolution;
  g->dev.inv_resolution_d = 1.0 / static_cast<double>(resolution);
  // probability_values.cc:56-60: unknown -> kMinProbability, range [0.1, 0.9].
  const float kMin = 0.1f, kMax = 1.f - kMin;
  g->dev.a = MakeDecode(kMin, kMin, kMax);
  g->dev.b = MakeDecode(0.f, 0.f, 1.f);
  g->dev.min_tsd = 0.f;
  g->dev.max_tsd = 0.f;
  g->grid_size = grid_size > 0 ? grid_size : DeriveGridSize(num_voxels, xyz);
  int rc = Flatten<uint16_t>(num_voxels, xyz, [values](int64_t i) { return values[i]; }, g);
  if (rc == HG_OK) rc = BuildDecodedBricks(g);
  if (rc != HG_OK) {
    hg_grid_destroy(g);
    return rc;
  }
  *out = g;
  return HG_OK;
}

int hg_grid_upload_tsdf(float resolution, float relative_truncation_distance, float max_weight,
                        int grid_size, int64_t num_voxels, const int32_t* xyz,
                        const uint16_t* tsd_values, const uint16_t* weight_values, hg_grid** out) {
  HG_REQUIRE(out != nullptr && resolution > 0.f && num_voxels >= 0 && max_weight > 0.f &&
                 relative_truncation_distance > 0.f,
             "hg_grid_upload_tsdf: bad arguments");
  HG_REQUIRE(num_voxels == 0 || (xyz != nullptr && tsd_values != nullptr && weight_values != nullptr),
             "hg_grid_upload_tsdf: null voxel arrays");
  hg_grid* g = new hg_grid();
  HG_CUDA(cudaGetDevice(&g->device));
  g->dev.kind = kGridTsdf;
  g->dev.resolution = resolution;
  g->dev.inv_resolution = 1.f / resolution;
  g->dev.inv_resolution_d = 1.0 / static_cast<double>(resolution);
  g->relative_truncation_distance = relative_truncation_distance;
  g->max_weight = max_weight;
  // hybrid_grid_tsdf.h:65-67, tsd_value_converter.cc:22-32.
  const float max_tsd = relative_truncation_distance * resolution;
  const float min_tsd = -max_tsd;
  g->dev.max_tsd = max_tsd;
  g->dev.min_tsd = min_tsd;
  g->dev.a = MakeDecode(min_tsd, min_tsd, max_tsd);
  g->dev.b = MakeDecode(0.f, 0.f, max_weight);
  g->grid_size = grid_size > 0 ? grid_size : DeriveGridSize(num_voxels, xyz);
  int rc = Flatten<ushort2>(
      num_voxels, xyz,
      [tsd_values, weight_values](int64_t i) { return make_ushort2(tsd_values[i], weight_values[i]); }, g);
  if (rc == HG_OK) rc = BuildDecodedBricks(g);
  if (rc != HG_OK) {
    hg_grid_destroy(g);
    return rc;
  }
  *out = g;
  return HG_OK;
}

void hg_grid_destroy(hg_grid* grid) {
  if (grid == nullptr) return;
  if (grid->d_index) cudaFree(grid->d_index);
  if (grid->d_bricks) cudaFree(grid->d_bricks);
  if (grid->d_decoded) cudaFree(grid->d_decoded);
  if (grid->d_stencils) cudaFree(grid->d_stencils);
  DestroyGridMutable(grid);
  delete grid;
}

int64_t hg_grid_stencil_bytes(const hg_grid* grid) {
  if (grid == nullptr || grid->dev.stencils == nullptr) return 0;
  return grid->num_all_bricks * 512 * 32;
}

int hg_grid_size(const hg_grid* grid) { return grid ? grid->grid_size : 0; }
int64_t hg_grid_num_bricks(const hg_grid* grid) { return grid ? grid->num_bricks : 0; }

int hg_grid_lookup(const hg_grid* grid, int what, int64_t n, const int32_t* xyz, float* out) {
  HG_REQUIRE(grid != nullptr && n >= 0 && what >= 0 && what <= 2, "hg_grid_lookup: bad arguments");
  HG_REQUIRE((grid->dev.kind == kGridProb) == (what == 0), "hg_grid_lookup: 'what' does not match the grid type");
  if (n == 0) return HG_OK;
  HG_ON_DEVICE(grid->device);
  int32_t* d_xyz = nullptr;
  float* d_out = nullptr;
  HG_CUDA(cudaMalloc(&d_xyz, n * 3 * sizeof(int32_t)));
  HG_CUDA(cudaMalloc(&d_out, n * sizeof(float)));
  HG_CUDA(cudaMemcpy(d_xyz, xyz, n * 3 * sizeof(int32_t), cudaMemcpyHostToDevice));
  LookupKernel<<<static_cast<unsigned>((n + 255) / 256), 256>>>(grid->dev, what, n, d_xyz, d_out);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(out, d_out, n * sizeof(float), cudaMemcpyDeviceToHost));
  cudaFree(d_xyz);
  cudaFree(d_out);
  return HG_OK;
}

int hg_grid_interpolated_weight(const hg_grid* grid, int64_t n, const double* points_xyz, double* out) {
  HG_REQUIRE(grid != nullptr && n >= 0, "hg_grid_interpolated_weight: bad arguments");
  HG_REQUIRE(grid->dev.kind != kGridProb, "hg_grid_interpolated_weight: needs a TSDF grid");
  if (n == 0) return HG_OK;
  HG_REQUIRE(points_xyz != nullptr && out != nullptr, "hg_grid_interpolated_weight: null buffer");
  HG_ON_DEVICE(grid->device);
  double* d_p = nullptr;
  double* d_out = nullptr;
  HG_CUDA(cudaMalloc(&d_p, n * 3 * sizeof(double)));
  HG_CUDA(cudaMalloc(&d_out, n * sizeof(double)));
  HG_CUDA(cudaMemcpy(d_p, points_xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice));
  InterpolatedWeightKernel<<<static_cast<unsigned>((n + 255) / 256), 256>>>(grid->dev, n, d_p, d_out);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(out, d_out, n * sizeof(double), cudaMemcpyDeviceToHost));
  cudaFree(d_p);
  cudaFree(d_out);
  return HG_OK;
}

int hg_get_cell_index(float resolution, int64_t n, const float* points_xyz, int32_t* out_xyz) {
  HG_REQUIRE(n >= 0 && resolution > 0.f, "hg_get_cell_index: bad arguments");
  if (n == 0) return HG_OK;
  float* d_p = nullptr;
  int32_t* d_o = nullptr;
  HG_CUDA(cudaMalloc(&d_p, n * 3 * sizeof(float)));
  HG_CUDA(cudaMalloc(&d_o, n * 3 * sizeof(int32_t)));
  HG_CUDA(cudaMemcpy(d_p, points_xyz, n * 3 * sizeof(float), cudaMemcpyHostToDevice));
  CellIndexKernel<<<static_cast<unsigned>((3 * n + 255) / 256), 256>>>(resolution, 1.f / resolution, n, d_p, d_o);
  CountLaunch();
  HG_CUDA(cudaGetLastError());
  HG_CUDA(cudaMemcpy(out_xyz, d_o, n * 3 * sizeof(int32_t), cudaMemcpyDeviceToHost));
  cudaFree(d_p);
  cudaFree(d_o);
  return HG_OK;
}

}  // extern "C"
