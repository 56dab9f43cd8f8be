// weight_cache.cu — the pre-swizzled weight cache file: everything load_network_from_dir
// (src/lib.rs:108-174) plus the tile builders produce for one scene, in one file, so that a context starts from a
// single sequential read instead of 48 small files and three re-swizzles per network.  Host-only code.
//
// Layout (little-endian, no padding):
//   char magic[8] = "NRFB200W" | u32 version | u32 n_nets (2: coarse, fine)
//   per net: u32 n_tensors | per tensor: u16 name_len, name, u32 ndim, i64 dims[ndim], f32 data[prod(dims)]
//            per operand format (bf16, fp16, fp16 hi/lo split): u64 bytes, chunk stream
//   u64 FNV-1a of every byte before it
// The f32 tensors are the file format of lego_rust/<net>/*.bin (lego_rust/README.md:23-36), names per shapes.txt
// (weights.rs:11-95); the chunk streams are exactly what build_tc_image returns, and a loader checks that.
#include <cstdio>
#include <cstring>

#include "weight_cache.h"

namespace nerf {
namespace {

constexpr char kMagic[8] = {'N', 'R', 'F', 'B', '2', '0', '0', 'W'};
// File version = format revision (high 8 bits) | a hash of kTcLayoutTag (mlp_tc.h): the chunk streams are shared-memory
// images of one kernel generation, and a file written for another chunk order or swizzle has the right size and a
// valid checksum - it must be refused, not rendered.
constexpr uint32_t fnv1a24(const char* s) {
  uint32_t h = 2166136261u;
  while (*s) h = (h ^ (uint32_t)(unsigned char)*s++) * 16777619u;
  return (h ^ (h >> 24)) & 0xffffffu;
}
constexpr uint32_t kVersion = (2u << 24) | fnv1a24(kTcLayoutTag);

uint64_t fnv1a(const uint8_t* p, size_t n) {
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) {
    h ^= p[i];
    h *= 1099511628211ull;
  }
  return h;
}

template <class T>
void put(std::vector<uint8_t>& b, const T& v) {
  const uint8_t* p = reinterpret_cast<const uint8_t*>(&v);
  b.insert(b.end(), p, p + sizeof(T));
}

struct Reader {
  const uint8_t* p;
  size_t n, off = 0;
  bool ok = true;
  template <class T>
  T get() {
    T v{};
    if (off + sizeof(T) > n) {
      ok = false;
      return v;
    }
    std::memcpy(&v, p + off, sizeof(T));
    off += sizeof(T);
    return v;
  }
  const uint8_t* bytes(size_t k) {
    if (!ok || k > n - off) {
      ok = false;
      return nullptr;
    }
    const uint8_t* q = p + off;
    off += k;
    return q;
  }
};

}  // namespace

int serialize_weight_cache(const TensorMap nets[2], std::vector<uint8_t>* out) {
  std::vector<uint8_t>& b = *out;
  b.clear();
  b.reserve(20u << 20);
  b.insert(b.end(), kMagic, kMagic + 8);
  put<uint32_t>(b, kVersion);
  put<uint32_t>(b, 2u);
  for (int n = 0; n < 2; ++n) {
    int rc = validate_network(nets[n], n == 0 ? "coarse" : "fine");
    if (rc != NERF_OK) return rc;
    put<uint32_t>(b, (uint32_t)nets[n].size());
    for (const auto& kv : nets[n]) {
      put<uint16_t>(b, (uint16_t)kv.first.size());
      b.insert(b.end(), kv.first.begin(), kv.first.end());
      put<uint32_t>(b, (uint32_t)kv.second.dims.size());
      for (int64_t d : kv.second.dims) put<int64_t>(b, d);
      const uint8_t* p = reinterpret_cast<const uint8_t*>(kv.second.data.data());
      b.insert(b.end(), p, p + kv.second.data.size() * 4);
    }
    for (int fmt = 0; fmt < kNumTcFormats; ++fmt) {
      std::vector<uint8_t> img;
      build_tc_image(nets[n], fmt, &img);
      put<uint64_t>(b, (uint64_t)img.size());
      b.insert(b.end(), img.begin(), img.end());
    }
  }
  put<uint64_t>(b, fnv1a(b.data(), b.size()));
  return NERF_OK;
}

int parse_weight_cache(const uint8_t* data, size_t size, CachedWeights* out) {
  if (size < 24 || std::memcmp(data, kMagic, 8) != 0) {
    set_error("weight cache: bad magic");
    return NERF_E_IO;
  }
  uint64_t want;
  std::memcpy(&want, data + size - 8, 8);
  if (fnv1a(data, size - 8) != want) {
    set_error("weight cache: checksum mismatch (truncated or corrupted file)");
    return NERF_E_IO;
  }
  Reader r{data, size - 8};
  r.off = 8;
  if (r.get<uint32_t>() != kVersion || r.get<uint32_t>() != 2u) {
    set_error("weight cache: unsupported version");
    return NERF_E_UNSUPPORTED;
  }
  for (int n = 0; n < 2 && r.ok; ++n) {
    const uint32_t nt = r.get<uint32_t>();
    for (uint32_t i = 0; i < nt && r.ok; ++i) {
      const uint16_t len = r.get<uint16_t>();
      const uint8_t* nm = r.bytes(len);
      if (!nm) break;
      HostTensor t;
      const uint32_t nd = r.get<uint32_t>();
      if (nd < 1 || nd > 2) {
        r.ok = false;
        break;
      }
      size_t cnt = 1;
      for (uint32_t d = 0; d < nd; ++d) {
        const int64_t v = r.get<int64_t>();
        if (v <= 0 || v > (1 << 20)) r.ok = false;
        t.dims.push_back(v);
        cnt *= (size_t)(v > 0 ? v : 1);
      }
      const uint8_t* p = r.ok ? r.bytes(cnt * 4) : nullptr;
      if (!p) break;
      t.data.resize(cnt);
      std::memcpy(t.data.data(), p, cnt * 4);
      out->nets[n][std::string(reinterpret_cast<const char*>(nm), len)] = std::move(t);
    }
    for (int fmt = 0; fmt < kNumTcFormats && r.ok; ++fmt) {
      const uint64_t bytes = r.get<uint64_t>();
      const uint8_t* p = r.bytes((size_t)bytes);
      if (!p) break;
      out->images[n][fmt].assign(p, p + bytes);
    }
  }
  if (!r.ok || r.off != size - 8) {
    set_error("weight cache: malformed body");
    return NERF_E_IO;
  }
  for (int n = 0; n < 2; ++n) {
    int rc = validate_network(out->nets[n], n == 0 ? "coarse" : "fine");
    if (rc != NERF_OK) return rc;
  }
  return NERF_OK;
}

int write_file(const std::string& path, const std::vector<uint8_t>& b) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) {
    set_error("write weight cache: " + path);
    return NERF_E_IO;
  }
  const size_t got = std::fwrite(b.data(), 1, b.size(), f);
  if (std::fclose(f) != 0 || got != b.size()) {
    set_error("short write: " + path);
    return NERF_E_IO;
  }
  return NERF_OK;
}

int read_file(const std::string& path, std::vector<uint8_t>* b) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) {
    set_error("read weight cache: " + path);
    return NERF_E_IO;
  }
  std::fseek(f, 0, SEEK_END);
  const long n = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  b->resize(n > 0 ? (size_t)n : 0);
  const size_t got = std::fread(b->data(), 1, b->size(), f);
  std::fclose(f);
  if (got != b->size()) {
    set_error("short read: " + path);
    return NERF_E_IO;
  }
  return NERF_OK;
}

}  // namespace nerf
