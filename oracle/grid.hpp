// ORACLE — TEST INFRASTRUCTURE ONLY. Not part of the product path.
//
// CPU restatement of HectorGrapher's sparse voxel storage and value encodings.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may use anything under oracle/.
//
// Follows (reference paths relative to /root/reference/cartographer):
//   mapping/3d/hybrid_grid_base.h:40-43    flat z-major index
//   mapping/3d/hybrid_grid_base.h:66-135   FlatGrid (8^3 values)
//   mapping/3d/hybrid_grid_base.h:140-245  NestedGrid (8^3 lazily allocated FlatGrids)
//   mapping/3d/hybrid_grid_base.h:250-405  DynamicGrid (2^bits per axis of NestedGrids, grows x2)
//   mapping/3d/hybrid_grid_base.h:412-460  HybridGridBase (GetCellIndex / GetCenterOfCell)
//   mapping/3d/hybrid_grid.h:46-140        HybridGrid (uint16 probability + update marker)
//   mapping/3d/hybrid_grid_tsdf.h:41-150   TSDFVoxel / HybridGridTSDF
//   mapping/probability_values.{h,cc}      u16 <-> probability, odds tables
//   mapping/value_conversion_tables.cc:29-51  u16 -> float table law
//   mapping/2d/tsd_value_converter.{h,cc}  TSD / weight <-> u16
//   common/port.h:40-46 (RoundToInt = lround), common/math.h:31-40 (Clamp)
//
// The three-level pointer tree and the 65536-entry float tables are kept on
// purpose: this file is also the timed CPU baseline and should cost what the
// reference costs per lookup.
//
// Build with -O3 -ffp-contract=off (reference Release flags have no FMA).
#pragma once

#include <array>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <map>
#include <memory>
#include <tuple>
#include <vector>

namespace oracle {

using uint8 = uint8_t;
using uint16 = uint16_t;

// common/port.h:40-42 — std::lround, i.e. half away from zero.
inline int RoundToInt(const float x) { return static_cast<int>(std::lround(x)); }
inline int RoundToInt(const double x) { return static_cast<int>(std::lround(x)); }

// common/math.h:31-40
template <typename T>
inline T Clamp(const T value, const T lo, const T hi) {
  if (value > hi) return hi;
  if (value < lo) return lo;
  return value;
}

struct Array3i {
  int x = 0, y = 0, z = 0;
  Array3i() = default;
  Array3i(int x_, int y_, int z_) : x(x_), y(y_), z(z_) {}
  int operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline Array3i operator+(const Array3i& a, const Array3i& b) {
  return Array3i(a.x + b.x, a.y + b.y, a.z + b.z);
}
inline Array3i operator-(const Array3i& a, const Array3i& b) {
  return Array3i(a.x - b.x, a.y - b.y, a.z - b.z);
}
inline bool operator==(const Array3i& a, const Array3i& b) {
  return a.x == b.x && a.y == b.y && a.z == b.z;
}

struct Vector3f {
  float x = 0.f, y = 0.f, z = 0.f;
  Vector3f() = default;
  Vector3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};

// hybrid_grid_base.h:40-43
inline int FlatIndex(const Array3i& i, const int bits) {
  return (((i.z << bits) + i.y) << bits) + i.x;
}
// hybrid_grid_base.h:47-52
inline Array3i UnflattenIndex(const int index, const int bits) {
  const int mask = (1 << bits) - 1;
  return Array3i(index & mask, (index >> bits) & mask, (index >> bits) >> bits);
}

template <typename V>
inline bool IsDefault(const V& v) {
  return v == V();
}

// Leaf: 8x8x8 values, contiguous.
template <typename V, int kBits>
class FlatGrid {
 public:
  using ValueType = V;
  FlatGrid() {
    for (V& c : cells_) c = V();
  }
  static int grid_size() { return 1 << kBits; }
  V value(const Array3i& i) const { return cells_[FlatIndex(i, kBits)]; }
  V* mutable_value(const Array3i& i) { return &cells_[FlatIndex(i, kBits)]; }
  // Visits non-default values in storage order.
  template <typename F>
  void ForEach(const Array3i& base, F&& f) const {
    for (int k = 0; k != (1 << (3 * kBits)); ++k) {
      if (!IsDefault(cells_[k])) {
        f(base + UnflattenIndex(k, kBits), cells_[k]);
      }
    }
  }

 private:
  std::array<V, 1 << (3 * kBits)> cells_;
};

// Middle level: 8x8x8 lazily created leaves.
template <typename Wrapped, int kBits>
class NestedGrid {
 public:
  using ValueType = typename Wrapped::ValueType;
  static int grid_size() { return Wrapped::grid_size() << kBits; }
  ValueType value(const Array3i& i) const {
    const Array3i meta(i.x / Wrapped::grid_size(), i.y / Wrapped::grid_size(),
                       i.z / Wrapped::grid_size());
    const Wrapped* const cell = meta_[FlatIndex(meta, kBits)].get();
    if (cell == nullptr) return ValueType();
    return cell->value(Array3i(i.x - meta.x * Wrapped::grid_size(),
                               i.y - meta.y * Wrapped::grid_size(),
                               i.z - meta.z * Wrapped::grid_size()));
  }
  ValueType* mutable_value(const Array3i& i) {
    const Array3i meta(i.x / Wrapped::grid_size(), i.y / Wrapped::grid_size(),
                       i.z / Wrapped::grid_size());
    std::unique_ptr<Wrapped>& cell = meta_[FlatIndex(meta, kBits)];
    if (cell == nullptr) cell.reset(new Wrapped());
    return cell->mutable_value(Array3i(i.x - meta.x * Wrapped::grid_size(),
                                       i.y - meta.y * Wrapped::grid_size(),
                                       i.z - meta.z * Wrapped::grid_size()));
  }
  template <typename F>
  void ForEach(const Array3i& base, F&& f) const {
    for (int k = 0; k != (1 << (3 * kBits)); ++k) {
      if (meta_[k] != nullptr) {
        const Array3i m = UnflattenIndex(k, kBits);
        meta_[k]->ForEach(
            base + Array3i(m.x * Wrapped::grid_size(), m.y * Wrapped::grid_size(),
                           m.z * Wrapped::grid_size()),
            f);
      }
    }
  }

 private:
  std::array<std::unique_ptr<Wrapped>, 1 << (3 * kBits)> meta_;
};

// Top level: starts 2x2x2, doubles per axis on demand (at most 2^8 per axis),
// indices symmetric around the origin.
template <typename Wrapped>
class DynamicGrid {
 public:
  using ValueType = typename Wrapped::ValueType;
  DynamicGrid() : bits_(1), meta_(8) {}
  DynamicGrid(DynamicGrid&&) = default;
  DynamicGrid& operator=(DynamicGrid&&) = default;

  int grid_size() const { return Wrapped::grid_size() << bits_; }

  // hybrid_grid_base.h:263-279
  ValueType value(const Array3i& index) const {
    const int half = grid_size() >> 1;
    const Array3i s(index.x + half, index.y + half, index.z + half);
    const unsigned gs = static_cast<unsigned>(grid_size());
    if (static_cast<unsigned>(s.x) >= gs || static_cast<unsigned>(s.y) >= gs ||
        static_cast<unsigned>(s.z) >= gs) {
      return ValueType();
    }
    const Array3i meta(s.x / Wrapped::grid_size(), s.y / Wrapped::grid_size(),
                       s.z / Wrapped::grid_size());
    const Wrapped* const cell = meta_[FlatIndex(meta, bits_)].get();
    if (cell == nullptr) return ValueType();
    return cell->value(Array3i(s.x - meta.x * Wrapped::grid_size(),
                               s.y - meta.y * Wrapped::grid_size(),
                               s.z - meta.z * Wrapped::grid_size()));
  }

  // hybrid_grid_base.h:283-301
  ValueType* mutable_value(const Array3i& index) {
    for (;;) {
      const int half = grid_size() >> 1;
      const Array3i s(index.x + half, index.y + half, index.z + half);
      const unsigned gs = static_cast<unsigned>(grid_size());
      if (static_cast<unsigned>(s.x) >= gs || static_cast<unsigned>(s.y) >= gs ||
          static_cast<unsigned>(s.z) >= gs) {
        Grow();
        continue;
      }
      const Array3i meta(s.x / Wrapped::grid_size(), s.y / Wrapped::grid_size(),
                         s.z / Wrapped::grid_size());
      std::unique_ptr<Wrapped>& cell = meta_[FlatIndex(meta, bits_)];
      if (cell == nullptr) cell.reset(new Wrapped());
      return cell->mutable_value(Array3i(s.x - meta.x * Wrapped::grid_size(),
                                         s.y - meta.y * Wrapped::grid_size(),
                                         s.z - meta.z * Wrapped::grid_size()));
    }
  }

  // Visits every non-default voxel (index, value) in the reference iterator's
  // order (hybrid_grid_base.h:304-372).
  template <typename F>
  void ForEach(F&& f) const {
    const int half = grid_size() >> 1;
    for (int k = 0; k != (1 << (3 * bits_)); ++k) {
      if (meta_[k] != nullptr) {
        const Array3i m = UnflattenIndex(k, bits_);
        meta_[k]->ForEach(Array3i(m.x * Wrapped::grid_size() - half,
                                  m.y * Wrapped::grid_size() - half,
                                  m.z * Wrapped::grid_size() - half),
                          f);
      }
    }
  }

 private:
  // hybrid_grid_base.h:385-403
  void Grow() {
    const int new_bits = bits_ + 1;
    if (new_bits > 8) {
      std::abort();  // reference: CHECK_LE(new_bits, 8)
    }
    std::vector<std::unique_ptr<Wrapped>> grown(8 * meta_.size());
    const int n = 1 << bits_;
    const int off = 1 << (bits_ - 1);
    for (int z = 0; z != n; ++z)
      for (int y = 0; y != n; ++y)
        for (int x = 0; x != n; ++x)
          grown[FlatIndex(Array3i(x + off, y + off, z + off), new_bits)] =
              std::move(meta_[FlatIndex(Array3i(x, y, z), bits_)]);
    meta_ = std::move(grown);
    bits_ = new_bits;
  }

  int bits_;
  std::vector<std::unique_ptr<Wrapped>> meta_;
};

template <typename V>
using GridBase = DynamicGrid<NestedGrid<FlatGrid<V, 3>, 3>>;

// hybrid_grid_base.h:412-460
template <typename V>
class HybridGridBase : public GridBase<V> {
 public:
  explicit HybridGridBase(const float resolution) : resolution_(resolution) {}
  float resolution() const { return resolution_; }
  // float division per component, then lround.
  Array3i GetCellIndex(const Vector3f& p) const {
    const float ix = p.x / resolution_;
    const float iy = p.y / resolution_;
    const float iz = p.z / resolution_;
    return Array3i(RoundToInt(ix), RoundToInt(iy), RoundToInt(iz));
  }
  Vector3f GetCenterOfCell(const Array3i& i) const {
    return Vector3f(static_cast<float>(i.x) * resolution_,
                    static_cast<float>(i.y) * resolution_,
                    static_cast<float>(i.z) * resolution_);
  }

 private:
  const float resolution_;
};

// ---------------------------------------------------------------------------
// Value encodings.

constexpr float kMinProbability = 0.1f;
constexpr float kMaxProbability = 1.f - kMinProbability;
constexpr uint16 kUpdateMarker = 1u << 15;

// probability_values.h:32-44
inline uint16 BoundedFloatToValue(const float v, const float lo, const float hi) {
  const int value = RoundToInt((Clamp(v, lo, hi) - lo) * (32766.f / (hi - lo))) + 1;
  return static_cast<uint16>(value);
}
inline uint16 ProbabilityToValue(const float p) {
  return BoundedFloatToValue(p, kMinProbability, kMaxProbability);
}
inline float Odds(const float p) { return p / (1.f - p); }
inline float ProbabilityFromOdds(const float odds) { return odds / (odds + 1.f); }

// value_conversion_tables.cc:29-51 (and probability_values.cc:29-54, which is
// the same law with kValueCount - 2.f == 32766.f): entry for a u16 with the
// update-marker bit masked off; 0 -> unknown_result.
inline std::vector<float> MakeValueToFloatTable(const float unknown_result,
                                                const float lo, const float hi) {
  std::vector<float> table;
  table.reserve(65536);
  for (int raw = 0; raw != 65536; ++raw) {
    const uint16 value = static_cast<uint16>(raw) & static_cast<uint16>(~kUpdateMarker);
    if (value == 0) {
      table.push_back(unknown_result);
    } else {
      const float scale = (hi - lo) / 32766.f;
      table.push_back(value * scale + (lo - scale));
    }
  }
  return table;
}

inline const std::vector<float>& ValueToProbabilityTable() {
  static const std::vector<float> table =
      MakeValueToFloatTable(kMinProbability, kMinProbability, kMaxProbability);
  return table;
}
inline float ValueToProbability(const uint16 v) { return ValueToProbabilityTable()[v]; }

// probability_values.cc:76-87
inline std::vector<uint16> ComputeLookupTableToApplyOdds(const float odds) {
  std::vector<uint16> result;
  result.reserve(32768);
  result.push_back(ProbabilityToValue(ProbabilityFromOdds(odds)) + kUpdateMarker);
  for (int cell = 1; cell != 32768; ++cell) {
    result.push_back(
        ProbabilityToValue(ProbabilityFromOdds(odds * Odds(ValueToProbabilityTable()[cell]))) +
        kUpdateMarker);
  }
  return result;
}

enum class GridType { PROBABILITY_GRID, TSDF };

struct GridInterface {
  virtual ~GridInterface() {}
  virtual GridType GetGridType() const = 0;
};

// hybrid_grid.h:46-140
class HybridGrid : public GridInterface, public HybridGridBase<uint16> {
 public:
  explicit HybridGrid(const float resolution) : HybridGridBase<uint16>(resolution) {}
  GridType GetGridType() const override { return GridType::PROBABILITY_GRID; }
  void FinishUpdate() {
    while (!update_indices_.empty()) {
      *update_indices_.back() -= kUpdateMarker;
      update_indices_.pop_back();
    }
  }
  bool IsKnown(const Array3i& i) const { return value(i) != 0; }
  void SetProbability(const Array3i& i, const float p) {
    *mutable_value(i) = ProbabilityToValue(p);
  }
  bool ApplyLookupTable(const Array3i& i, const std::vector<uint16>& table) {
    uint16* const cell = mutable_value(i);
    if (*cell >= kUpdateMarker) return false;
    update_indices_.push_back(cell);
    *cell = table[*cell];
    return true;
  }
  float GetProbability(const Array3i& i) const { return ValueToProbability(value(i)); }

 private:
  std::vector<uint16*> update_indices_;
};

// hybrid_grid_tsdf.h:41-50
struct TSDFVoxel {
  uint16 discrete_tsd = 0;
  uint16 discrete_weight = 0;
  bool operator==(const TSDFVoxel& o) const {
    return discrete_tsd == o.discrete_tsd && discrete_weight == o.discrete_weight;
  }
};

// 2d/tsd_value_converter.{h,cc}
class TSDValueConverter {
 public:
  TSDValueConverter(const float max_tsd, const float max_weight)
      : max_tsd_(max_tsd),
        min_tsd_(-max_tsd),
        max_weight_(max_weight),
        tsd_resolution_(32766.f / (max_tsd_ - min_tsd_)),
        weight_resolution_(32766.f / (max_weight_ - 0.f)),
        value_to_tsd_(MakeValueToFloatTable(min_tsd_, min_tsd_, max_tsd_)),
        value_to_weight_(MakeValueToFloatTable(0.f, 0.f, max_weight)) {}
  uint16 TSDToValue(const float tsd) const {
    return static_cast<uint16>(
        RoundToInt((Clamp(tsd, min_tsd_, max_tsd_) - min_tsd_) * tsd_resolution_) + 1);
  }
  uint16 WeightToValue(const float w) const {
    return static_cast<uint16>(
        RoundToInt((Clamp(w, 0.f, max_weight_) - 0.f) * weight_resolution_) + 1);
  }
  float ValueToTSD(const uint16 v) const { return value_to_tsd_[v]; }
  float ValueToWeight(const uint16 v) const { return value_to_weight_[v]; }
  float getMaxTSD() const { return max_tsd_; }
  float getMinTSD() const { return min_tsd_; }
  float getMaxWeight() const { return max_weight_; }

 private:
  float max_tsd_, min_tsd_, max_weight_, tsd_resolution_, weight_resolution_;
  std::vector<float> value_to_tsd_, value_to_weight_;
};

// hybrid_grid_tsdf.h:57-150
class HybridGridTSDF : public GridInterface, public HybridGridBase<TSDFVoxel> {
 public:
  HybridGridTSDF(const float resolution, const float relative_truncation_distance,
                 const float max_weight)
      : HybridGridBase<TSDFVoxel>(resolution),
        converter_(relative_truncation_distance * resolution, max_weight) {}
  GridType GetGridType() const override { return GridType::TSDF; }
  // Note (hybrid_grid_tsdf.h:87-93): the stored TSD value carries the update
  // marker bit and SetCell never registers the cell for FinishUpdate.
  void SetCell(const Array3i& i, const float tsd, const float weight) {
    TSDFVoxel v;
    v.discrete_tsd = static_cast<uint16>(converter_.TSDToValue(tsd) + kUpdateMarker);
    v.discrete_weight = converter_.WeightToValue(weight);
    *mutable_value(i) = v;
  }
  float GetTSD(const Array3i& i) const { return converter_.ValueToTSD(value(i).discrete_tsd); }
  float GetWeight(const Array3i& i) const {
    return converter_.ValueToWeight(value(i).discrete_weight);
  }
  bool IsKnown(const Array3i& i) const { return value(i).discrete_weight != 0; }
  const TSDValueConverter& ValueConverter() const { return converter_; }

 private:
  TSDValueConverter converter_;
};

// internal/3d/scan_matching/precomputation_grid_3d.h:27-37
class PrecomputationGrid3D : public HybridGridBase<uint8> {
 public:
  explicit PrecomputationGrid3D(const float resolution) : HybridGridBase<uint8>(resolution) {}
  static float ToProbability(const float value) {
    return kMinProbability + value * ((kMaxProbability - kMinProbability) / 255.f);
  }
};

}  // namespace oracle
