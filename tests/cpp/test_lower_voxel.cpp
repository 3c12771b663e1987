// CPU-only check of the arithmetic behind LowerVoxelNoBranch (hectorgrapher_b200/csrc/interpolation.cuh):
// outside the band in which it asks for the exact path, rounding q = fl(xf * fl(1 / res)) to the
// nearest integer gives the cell index the reference computes with an IEEE division and lround
// (HybridGridBase::GetCellIndex, hybrid_grid_base.h:428-433; common/port.h:40-42), and the lower-voxel
// selection that follows (interpolated_tsdf.h:159-192) is the same. The device code uses the _rn
// intrinsics; this restatement uses the same IEEE single-precision operations (no contraction:
// compile with -ffp-contract=off). Prints "LOWER VOXEL OK" and exits 0 on success.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <random>

struct Lower {
  double x1, x2;
  int index;
  bool exact_path;
};

static Lower NoBranch(const float resolution, const float inv_resolution, const double x) {
  const float xf = static_cast<float>(x);
  const float q = xf * inv_resolution;
  const int i = static_cast<int>(std::nearbyintf(q));  // round to nearest even, like __float2int_rn
  const float fi = static_cast<float>(i);
  const float d = q - fi;
  Lower r;
  r.exact_path = !((0.5f - std::fabs(d)) > std::fabs(q) * 4.8e-7f);
  const float c = fi * resolution;
  const bool above = static_cast<double>(c) > x;
  const float lower = above ? c - resolution : c;
  r.x1 = static_cast<double>(lower);
  r.x2 = static_cast<double>(lower + resolution);
  r.index = above ? i - 1 : i;
  return r;
}

// The reference, operation by operation.
static Lower Reference(const float resolution, const double x) {
  const float xf = static_cast<float>(x);
  int i = static_cast<int>(std::lround(xf / resolution));
  float c = static_cast<float>(i) * resolution;  // GetCenterOfCell: index * resolution
  if (static_cast<double>(c) > x) {
    c -= resolution;
    i = static_cast<int>(std::lround(c / resolution));  // the reference's second GetCellIndex
  }
  Lower r;
  r.x1 = static_cast<double>(c);
  r.x2 = static_cast<double>(c + resolution);
  r.index = i;
  r.exact_path = false;
  return r;
}

int main() {
  std::mt19937_64 rng(12345);
  const float resolutions[] = {0.05f, 0.1f, 0.45f, 2.0f, 0.3f};
  long long checked = 0, exact = 0, exact_on_random = 0, random_values = 0;
  for (const float res : resolutions) {
    const float inv = 1.f / res;
    std::uniform_real_distribution<double> wide(-400.0, 400.0);
    std::uniform_int_distribution<int> cell(-8000, 8000);
    std::uniform_real_distribution<double> tiny(-1e-5, 1e-5);
    for (int k = 0; k < 4000000; ++k) {
      double x;
      switch (k & 3) {
        case 0: x = wide(rng); break;
        case 1: x = (cell(rng) + 0.5) * static_cast<double>(res) + tiny(rng); break;  // around a cell boundary
        case 2: x = cell(rng) * static_cast<double>(res) + tiny(rng); break;           // around a cell centre
        default: x = static_cast<double>(static_cast<float>((cell(rng) + 0.5f) * res)); break;  // float half cells
      }
      const Lower a = NoBranch(res, inv, x);
      random_values += (k & 3) == 0;
      if (a.exact_path) {
        ++exact;
        exact_on_random += (k & 3) == 0;
        continue;  // the kernel takes LowerVoxel (exact division) for these
      }
      const Lower b = Reference(res, x);
      ++checked;
      if (a.index != b.index || a.x1 != b.x1 || a.x2 != b.x2) {
        std::fprintf(stderr, "FAILED: res %.9g x %.17g: index %d vs %d, x1 %.17g vs %.17g\n", res, x, a.index, b.index,
                     a.x1, b.x1);
        return 1;
      }
    }
  }
  // the exact path is for the neighbourhood of cell boundaries (a band of 4.8e-7 |q| around the half-way
  // points: 0.14 % of uniform values at |q| up to 8 000): taken there, rare on random input
  if (exact == 0 || exact_on_random * 100 > random_values) {
    std::fprintf(stderr, "FAILED: exact path taken %lld times, %lld of %lld random values\n", exact, exact_on_random,
                 random_values);
    return 1;
  }
  std::printf("LOWER VOXEL OK (%lld values, %lld through the exact path)\n", checked, exact);
  return 0;
}
