// Grid wire format: proto::HybridGrid / proto::HybridGridTSDF messages
// and the .pbstream container, decoded without libprotobuf, and uploaded to brick tables.
//
// Reference behaviour reproduced here (paths under /root/reference/cartographer):
//   mapping/proto/3d/hybrid_grid.proto:19-29, hybrid_grid_tsdf.proto:19-32   message layouts
//   mapping/3d/hybrid_grid.h:61-73          HybridGrid(proto): SetProbability(ValueToProbability(v))
//   mapping/3d/hybrid_grid_tsdf.h:69-83     HybridGridTSDF(proto): SetCell(ValueToTSD, ValueToWeight)
//   mapping/3d/hybrid_grid.h:88-99, hybrid_grid_tsdf.h:119-134   ToProto (incl. the max_tsd quirk)
//   io/proto_stream.cc:27-106               magic + little-endian size + gzip chunks
//   mapping/proto/serialization.proto:26-30, 72-88; mapping/proto/submap.proto:33-47
//   transform/proto/transform.proto:33-75   Rigid3d = Vector3d translation, Quaterniond (x, y, z, w)
// The protobuf wire format itself (varints, zig-zag, packed repeated fields, length-delimited
// sub-messages) is the published encoding of Protocol Buffers (proto3).
#include <zlib.h>

#include <cmath>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "hg_internal.cuh"

namespace hg {
namespace {

struct Reader {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  Reader(const void* data, size_t size)
      : p(static_cast<const uint8_t*>(data)), end(static_cast<const uint8_t*>(data) + size) {}
  bool done() const { return p >= end || !ok; }
  uint64_t Varint() {
    uint64_t v = 0;
    for (int shift = 0; shift < 64; shift += 7) {
      if (p >= end) { ok = false; return 0; }
      const uint8_t b = *p++;
      v |= static_cast<uint64_t>(b & 0x7F) << shift;
      if (!(b & 0x80)) return v;
    }
    ok = false;
    return 0;
  }
  uint32_t Fixed32() {
    if (end - p < 4) { ok = false; return 0; }
    uint32_t v;
    std::memcpy(&v, p, 4);
    p += 4;
    return v;
  }
  uint64_t Fixed64() {
    if (end - p < 8) { ok = false; return 0; }
    uint64_t v;
    std::memcpy(&v, p, 8);
    p += 8;
    return v;
  }
  float Float() {
    const uint32_t v = Fixed32();
    float f;
    std::memcpy(&f, &v, 4);
    return f;
  }
  double Double() {
    const uint64_t v = Fixed64();
    double d;
    std::memcpy(&d, &v, 8);
    return d;
  }
  Reader Sub() {  // length-delimited payload
    const uint64_t n = Varint();
    if (!ok || n > static_cast<uint64_t>(end - p)) { ok = false; return Reader(p, 0); }
    Reader r(p, static_cast<size_t>(n));
    p += n;
    return r;
  }
  void Skip(const int wire_type) {
    switch (wire_type) {
      case 0: Varint(); break;
      case 1: Fixed64(); break;
      case 2: Sub(); break;
      case 5: Fixed32(); break;
      default: ok = false;
    }
  }
};

inline int32_t ZigZag32(const uint64_t v) {
  const uint32_t n = static_cast<uint32_t>(v);
  return static_cast<int32_t>((n >> 1) ^ (~(n & 1) + 1));
}

// repeated sint32 / int32, packed (wire type 2) or not (wire type 0).
bool ReadRepeatedInt(Reader* r, const int wire_type, const bool zigzag, std::vector<int32_t>* out) {
  if (wire_type == 2) {
    Reader s = r->Sub();
    while (!s.done()) {
      const uint64_t v = s.Varint();
      out->push_back(zigzag ? ZigZag32(v) : static_cast<int32_t>(v));
    }
    if (!s.ok) r->ok = false;
    return r->ok;
  }
  if (wire_type == 0) {
    const uint64_t v = r->Varint();
    out->push_back(zigzag ? ZigZag32(v) : static_cast<int32_t>(v));
    return r->ok;
  }
  r->ok = false;
  return false;
}

struct GridMessage {
  float resolution = 0.f;
  float relative_truncation_distance = 0.f;
  float max_weight = 0.f;
  std::vector<int32_t> x, y, z, a, b;
};

int ParseGrid(const void* bytes, const size_t size, const bool tsdf, GridMessage* m) {
  Reader r(bytes, size);
  while (!r.done()) {
    const uint64_t tag = r.Varint();
    const int field = static_cast<int>(tag >> 3), wt = static_cast<int>(tag & 7);
    if (field == 1 && wt == 5) m->resolution = r.Float();
    else if (field == 3) ReadRepeatedInt(&r, wt, true, &m->x);
    else if (field == 4) ReadRepeatedInt(&r, wt, true, &m->y);
    else if (field == 5) ReadRepeatedInt(&r, wt, true, &m->z);
    else if (field == 6) ReadRepeatedInt(&r, wt, false, &m->a);
    else if (tsdf && field == 7) ReadRepeatedInt(&r, wt, false, &m->b);
    else if (tsdf && field == 8 && wt == 5) m->relative_truncation_distance = r.Float();
    else if (tsdf && field == 9 && wt == 5) m->max_weight = r.Float();
    else r.Skip(wt);
  }
  HG_REQUIRE(r.ok, "grid proto: malformed message");
  // The CHECK_EQs of the constructors.
  HG_REQUIRE(m->a.size() == m->x.size() && m->a.size() == m->y.size() && m->a.size() == m->z.size(),
             "grid proto: index and value arrays differ in length");
  HG_REQUIRE(!tsdf || m->b.size() == m->a.size(), "grid proto: values_tsd and values_weight differ in length");
  HG_REQUIRE(m->resolution > 0.f, "grid proto: resolution missing");
  return HG_OK;
}

inline float ClampF(const float v, const float lo, const float hi) { return v > hi ? hi : (v < lo ? lo : v); }

// value -> float -> value, the way the proto constructors store a voxel.
struct RoundTrip {
  float lo, hi;
  float unknown;
  uint16_t operator()(const int32_t raw) const {
    const uint16_t v = static_cast<uint16_t>(raw) & 0x7FFF;
    const float scale = (hi - lo) / 32766.f;
    const float f = v == 0 ? unknown : v * scale + (lo - scale);  // value_conversion_tables.cc:29-51
    return static_cast<uint16_t>(std::lround((ClampF(f, lo, hi) - lo) * (32766.f / (hi - lo))) + 1);
  }
};

int UploadMessage(const GridMessage& m, const bool tsdf, hg_grid** out) {
  const int64_t n = static_cast<int64_t>(m.a.size());
  std::vector<int32_t> xyz(static_cast<size_t>(3 * n));
  std::vector<uint16_t> a(static_cast<size_t>(n)), b(static_cast<size_t>(tsdf ? n : 0));
  for (int64_t i = 0; i < n; ++i) {
    xyz[3 * i] = m.x[i];
    xyz[3 * i + 1] = m.y[i];
    xyz[3 * i + 2] = m.z[i];
  }
  if (!tsdf) {
    const RoundTrip prob{0.1f, 1.f - 0.1f, 0.1f};
    for (int64_t i = 0; i < n; ++i) a[i] = prob(m.a[i]);
    return hg_grid_upload_prob(m.resolution, 0, n, xyz.data(), a.data(), out);
  }
  HG_REQUIRE(m.relative_truncation_distance > 0.f && m.max_weight > 0.f,
             "grid proto: relative_truncation_distance / max_weight missing");
  const float max_tsd = m.relative_truncation_distance * m.resolution;  // hybrid_grid_tsdf.h:65-67
  const RoundTrip tsd{-max_tsd, max_tsd, -max_tsd};
  const RoundTrip weight{0.f, m.max_weight, 0.f};
  for (int64_t i = 0; i < n; ++i) {
    a[i] = static_cast<uint16_t>(tsd(m.a[i]) + 0x8000);  // SetCell adds the update marker
    b[i] = weight(m.b[i]);
  }
  return hg_grid_upload_tsdf(m.resolution, m.relative_truncation_distance, m.max_weight, 0, n, xyz.data(),
                             a.data(), b.data(), out);
}

// ---- encoder -------------------------------------------------------------------

void PutVarint(std::string* s, uint64_t v) {
  while (v >= 0x80) {
    s->push_back(static_cast<char>((v & 0x7F) | 0x80));
    v >>= 7;
  }
  s->push_back(static_cast<char>(v));
}
void PutFloatField(std::string* s, const int field, const float f) {
  uint32_t bits;
  std::memcpy(&bits, &f, 4);
  if (bits == 0) return;  // proto3 omits default values
  PutVarint(s, static_cast<uint64_t>(field) << 3 | 5);
  s->append(reinterpret_cast<const char*>(&bits), 4);
}
template <typename T, typename Encode>
void PutPacked(std::string* s, const int field, const T* values, const int64_t n, const int64_t stride,
               Encode encode) {
  if (n == 0) return;
  std::string body;
  body.reserve(static_cast<size_t>(n) * 2);
  for (int64_t i = 0; i < n; ++i) PutVarint(&body, encode(values[i * stride]));
  PutVarint(s, static_cast<uint64_t>(field) << 3 | 2);
  PutVarint(s, body.size());
  s->append(body);
}

// ---- pbstream ------------------------------------------------------------------

constexpr uint64_t kMagic = 0x7b1d1f7b5bf501dbull;  // io/proto_stream.cc:27

bool Gunzip(const uint8_t* data, const size_t size, std::string* out) {
  z_stream zs;
  std::memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, 16 + MAX_WBITS) != Z_OK) return false;
  zs.next_in = const_cast<Bytef*>(data);
  zs.avail_in = static_cast<uInt>(size);
  out->clear();
  char buffer[1 << 16];
  int rc;
  do {
    zs.next_out = reinterpret_cast<Bytef*>(buffer);
    zs.avail_out = sizeof(buffer);
    rc = inflate(&zs, Z_NO_FLUSH);
    if (rc != Z_OK && rc != Z_STREAM_END) {
      inflateEnd(&zs);
      return false;
    }
    out->append(buffer, sizeof(buffer) - zs.avail_out);
  } while (rc != Z_STREAM_END);
  inflateEnd(&zs);
  return true;
}

struct Slice {
  const uint8_t* p = nullptr;
  size_t n = 0;
};

}  // namespace
}  // namespace hg

struct hg_pbstream {
  std::string file;
  size_t offset = 8;
  bool header_read = false;
  std::string chunk;  // current decompressed SerializedData
  hg::Slice grids[2];
  int is_tsdf[2] = {-1, -1};
  std::vector<float> histogram;
};

using namespace hg;

extern "C" {

int hg_grid_proto_decode(const void* bytes, int64_t size, int is_tsdf, hg_grid_proto_info* info,
                         int64_t capacity, int32_t* xyz, uint16_t* values_a, uint16_t* values_b) {
  HG_REQUIRE(bytes != nullptr && size >= 0 && info != nullptr && capacity >= 0, "hg_grid_proto_decode: bad arguments");
  GridMessage m;
  const int rc = ParseGrid(bytes, static_cast<size_t>(size), is_tsdf != 0, &m);
  if (rc != HG_OK) return rc;
  info->resolution = m.resolution;
  info->relative_truncation_distance = m.relative_truncation_distance;
  info->max_weight = m.max_weight;
  info->is_tsdf = is_tsdf != 0;
  info->num_voxels = static_cast<int64_t>(m.a.size());
  if (capacity < info->num_voxels || info->num_voxels == 0) return HG_OK;
  HG_REQUIRE(xyz != nullptr && values_a != nullptr && (!is_tsdf || values_b != nullptr),
             "hg_grid_proto_decode: null output arrays");
  for (size_t i = 0; i < m.a.size(); ++i) {
    xyz[3 * i] = m.x[i];
    xyz[3 * i + 1] = m.y[i];
    xyz[3 * i + 2] = m.z[i];
    values_a[i] = static_cast<uint16_t>(m.a[i]);
    if (is_tsdf) values_b[i] = static_cast<uint16_t>(m.b[i]);
  }
  return HG_OK;
}

int hg_grid_from_proto(const void* bytes, int64_t size, int is_tsdf, hg_grid** out) {
  HG_REQUIRE(bytes != nullptr && size >= 0 && out != nullptr, "hg_grid_from_proto: bad arguments");
  GridMessage m;
  const int rc = ParseGrid(bytes, static_cast<size_t>(size), is_tsdf != 0, &m);
  if (rc != HG_OK) return rc;
  return UploadMessage(m, is_tsdf != 0, out);
}

int hg_grid_to_proto(const hg_grid* grid, void* buffer, int64_t capacity, int64_t* size) {
  HG_REQUIRE(grid != nullptr && size != nullptr && capacity >= 0, "hg_grid_to_proto: bad arguments");
  int64_t n = 0;
  int rc = hg_grid_export(grid, 0, &n, nullptr, nullptr, nullptr);
  if (rc != HG_OK) return rc;
  const bool tsdf = grid->dev.kind == kGridTsdf;
  std::vector<int32_t> xyz(static_cast<size_t>(3 * n));
  std::vector<uint16_t> a(static_cast<size_t>(n)), b(static_cast<size_t>(n));
  if (n > 0) {
    rc = hg_grid_export(grid, n, &n, xyz.data(), a.data(), b.data());
    if (rc != HG_OK) return rc;
  }
  std::string s;
  PutFloatField(&s, 1, grid->dev.resolution);
  const auto zigzag = [](const int32_t v) {
    return static_cast<uint64_t>((static_cast<uint32_t>(v) << 1) ^ static_cast<uint32_t>(v >> 31));
  };
  const auto plain = [](const uint16_t v) { return static_cast<uint64_t>(v); };
  PutPacked(&s, 3, xyz.data(), n, 3, zigzag);
  PutPacked(&s, 4, xyz.data() + 1, n, 3, zigzag);
  PutPacked(&s, 5, xyz.data() + 2, n, 3, zigzag);
  PutPacked(&s, 6, a.data(), n, 1, plain);
  if (tsdf) {
    PutPacked(&s, 7, b.data(), n, 1, plain);
    PutFloatField(&s, 8, grid->dev.max_tsd);  // hybrid_grid_tsdf.h:131 stores getMaxTSD() here
    PutFloatField(&s, 9, grid->max_weight);
  }
  *size = static_cast<int64_t>(s.size());
  if (*size > capacity) return HG_OK;
  HG_REQUIRE(buffer != nullptr || s.empty(), "hg_grid_to_proto: null buffer");
  if (!s.empty()) std::memcpy(buffer, s.data(), s.size());
  return HG_OK;
}

int hg_pbstream_open(const char* path, hg_pbstream** out) {
  HG_REQUIRE(path != nullptr && out != nullptr, "hg_pbstream_open: bad arguments");
  std::ifstream in(path, std::ios::in | std::ios::binary);
  HG_REQUIRE(in.good(), std::string("hg_pbstream_open: cannot open ") + path);
  hg_pbstream* s = new hg_pbstream;
  s->file.assign(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
  uint64_t magic = 0;
  if (s->file.size() >= 8) std::memcpy(&magic, s->file.data(), 8);  // little-endian host
  if (magic != kMagic) {
    delete s;
    SetError(std::string("hg_pbstream_open: ") + path + " is not a proto stream (bad magic)");
    return HG_ERR_INVALID;
  }
  *out = s;
  return HG_OK;
}

int hg_pbstream_next_submap(hg_pbstream* s, int* has_submap, hg_submap_info* info) {
  HG_REQUIRE(s != nullptr && has_submap != nullptr && info != nullptr, "hg_pbstream_next_submap: bad arguments");
  *has_submap = 0;
  for (;;) {
    if (s->offset + 8 > s->file.size()) return HG_OK;  // end of stream
    uint64_t compressed = 0;
    std::memcpy(&compressed, s->file.data() + s->offset, 8);
    s->offset += 8;
    HG_REQUIRE(compressed <= s->file.size() - s->offset, "hg_pbstream: truncated chunk");
    const uint8_t* data = reinterpret_cast<const uint8_t*>(s->file.data()) + s->offset;
    s->offset += compressed;
    HG_REQUIRE(Gunzip(data, static_cast<size_t>(compressed), &s->chunk), "hg_pbstream: gzip error");
    if (!s->header_read) {  // SerializationHeader
      s->header_read = true;
      continue;
    }
    // SerializedData: field 3 = Submap.
    Reader r(s->chunk.data(), s->chunk.size());
    Slice submap3d;
    hg_submap_info found;
    std::memset(&found, 0, sizeof(found));
    found.local_pose[3] = 1.0;
    while (!r.done()) {
      const uint64_t tag = r.Varint();
      const int field = static_cast<int>(tag >> 3), wt = static_cast<int>(tag & 7);
      if (field == 3 && wt == 2) {
        Reader sub = r.Sub();
        while (!sub.done()) {
          const uint64_t t2 = sub.Varint();
          const int f2 = static_cast<int>(t2 >> 3), w2 = static_cast<int>(t2 & 7);
          if (f2 == 1 && w2 == 2) {  // SubmapId
            Reader id = sub.Sub();
            while (!id.done()) {
              const uint64_t t3 = id.Varint();
              if ((t3 >> 3) == 1 && (t3 & 7) == 0) found.trajectory_id = static_cast<int32_t>(id.Varint());
              else if ((t3 >> 3) == 2 && (t3 & 7) == 0) found.submap_index = static_cast<int32_t>(id.Varint());
              else id.Skip(static_cast<int>(t3 & 7));
            }
            if (!id.ok) sub.ok = false;
          } else if (f2 == 3 && w2 == 2) {  // Submap3D
            Reader body = sub.Sub();
            submap3d.p = body.p;
            submap3d.n = static_cast<size_t>(body.end - body.p);
          } else {
            sub.Skip(w2);
          }
        }
        if (!sub.ok) r.ok = false;
      } else {
        r.Skip(wt);
      }
    }
    HG_REQUIRE(r.ok, "hg_pbstream: malformed SerializedData");
    if (submap3d.p == nullptr) continue;  // some other kind of data
    // Submap3D (submap.proto:33-47).
    s->grids[0] = s->grids[1] = Slice();
    s->is_tsdf[0] = s->is_tsdf[1] = -1;
    s->histogram.clear();
    Reader b(submap3d.p, submap3d.n);
    while (!b.done()) {
      const uint64_t tag = b.Varint();
      const int field = static_cast<int>(tag >> 3), wt = static_cast<int>(tag & 7);
      if (field == 1 && wt == 2) {  // Rigid3d
        Reader pose = b.Sub();
        while (!pose.done()) {
          const uint64_t t2 = pose.Varint();
          const int f2 = static_cast<int>(t2 >> 3), w2 = static_cast<int>(t2 & 7);
          if ((f2 == 1 || f2 == 2) && w2 == 2) {
            Reader v = pose.Sub();
            if (f2 == 2) found.local_pose[3] = 0.0;  // a present Quaterniond has all four fields (0 if omitted)
            while (!v.done()) {
              const uint64_t t3 = v.Varint();
              const int f3 = static_cast<int>(t3 >> 3);
              if ((t3 & 7) != 1 || f3 < 1 || f3 > 4) { v.Skip(static_cast<int>(t3 & 7)); continue; }
              const double d = v.Double();
              if (f2 == 1 && f3 <= 3) found.local_pose[f3 - 1] = d;          // Vector3d x, y, z
              if (f2 == 2) found.local_pose[f3 == 4 ? 3 : 3 + f3] = d;        // Quaterniond x, y, z, w -> w, x, y, z
            }
            if (!v.ok) pose.ok = false;
          } else {
            pose.Skip(w2);
          }
        }
        if (!pose.ok) b.ok = false;
      } else if (field == 2 && wt == 0) {
        found.num_range_data = static_cast<int32_t>(b.Varint());
      } else if (field == 3 && wt == 0) {
        found.finished = b.Varint() != 0;
      } else if ((field == 4 || field == 7 || field == 5 || field == 8) && wt == 2) {
        Reader g = b.Sub();
        const int which = (field == 4 || field == 7) ? 0 : 1;
        s->grids[which].p = g.p;
        s->grids[which].n = static_cast<size_t>(g.end - g.p);
        s->is_tsdf[which] = field >= 7 ? 1 : 0;
      } else if (field == 6 && wt == 2) {
        Reader h = b.Sub();
        while (!h.done()) s->histogram.push_back(h.Float());
        if (!h.ok) b.ok = false;
      } else if (field == 6 && wt == 5) {
        s->histogram.push_back(b.Float());
      } else {
        b.Skip(wt);
      }
    }
    HG_REQUIRE(b.ok, "hg_pbstream: malformed Submap3D");
    found.high_is_tsdf = s->is_tsdf[0];
    found.low_is_tsdf = s->is_tsdf[1];
    found.histogram_size = static_cast<int32_t>(s->histogram.size());
    *info = found;
    *has_submap = 1;
    return HG_OK;
  }
}

int hg_pbstream_submap_grid(hg_pbstream* s, int which, hg_grid** out) {
  HG_REQUIRE(s != nullptr && out != nullptr && (which == 0 || which == 1), "hg_pbstream_submap_grid: bad arguments");
  HG_REQUIRE(s->is_tsdf[which] >= 0, "hg_pbstream_submap_grid: the current submap has no such grid");
  return hg_grid_from_proto(s->grids[which].p, static_cast<int64_t>(s->grids[which].n), s->is_tsdf[which], out);
}

int hg_pbstream_submap_histogram(hg_pbstream* s, int capacity, float* histogram) {
  HG_REQUIRE(s != nullptr && capacity >= 0, "hg_pbstream_submap_histogram: bad arguments");
  HG_REQUIRE(static_cast<size_t>(capacity) >= s->histogram.size() && (histogram != nullptr || s->histogram.empty()),
             "hg_pbstream_submap_histogram: buffer too small");
  if (!s->histogram.empty()) std::memcpy(histogram, s->histogram.data(), s->histogram.size() * sizeof(float));
  return HG_OK;
}

void hg_pbstream_close(hg_pbstream* s) { delete s; }

}  // extern "C"
