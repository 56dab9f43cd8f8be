// host_util.cpp — host side of the boundary (no CUDA here): error slot, weight files, camera JSON, PPM.
// Compiled with -ffp-contract=off so the f32 camera maths rounds like the reference's Rust.
#include "host_util.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace nerf {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }
const char* get_error() { return g_error.c_str(); }

int load_tensor_dir(const std::string& dir, TensorMap* out) {
  std::ifstream sf(dir + "/shapes.txt");
  if (!sf) {
    set_error("read shapes: " + dir + "/shapes.txt");
    return NERF_E_IO;
  }
  std::string line;
  while (std::getline(sf, line)) {  // "name dim0 [dim1]" (src/lib.rs:67-72)
    std::istringstream ls(line);
    std::string name;
    if (!(ls >> name)) continue;
    HostTensor t;
    int64_t d;
    while (ls >> d) t.dims.push_back(d);
    const std::string path = dir + "/" + name + ".bin";
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) {
      set_error("read tensor: " + path);
      return NERF_E_IO;
    }
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    t.data.resize((size_t)bytes / 4);  // chunks_exact(4) of little-endian f32 (src/lib.rs:38-40)
    const size_t got = std::fread(t.data.data(), 4, t.data.size(), f);
    std::fclose(f);
    if (got != t.data.size()) {
      set_error("short read: " + path);
      return NERF_E_IO;
    }
    (*out)[name] = std::move(t);
  }
  return NERF_OK;
}

int validate_network(const TensorMap& t, const char* which) {
  struct Spec {
    const char* name;
    int64_t d0, d1;
  };
  static const Spec specs[] = {
      {"dense0", 63, 256},  {"dense1", 256, 256}, {"dense2", 256, 256},     {"dense3", 256, 256},
      {"dense4", 256, 256}, {"dense5", 319, 256}, {"dense6", 256, 256},     {"dense7", 256, 256},
      {"bottleneck", 256, 256}, {"viewdirs", 283, 128}, {"rgb", 128, 3},    {"alpha", 256, 1}};
  for (const Spec& s : specs) {
    const std::string kn = std::string(s.name) + "_kernel", bn = std::string(s.name) + "_bias";
    auto k = t.find(kn), b = t.find(bn);
    if (k == t.end()) {
      set_error(std::string(which) + ": missing matrix parameter: " + kn);  // src/lib.rs:118
      return NERF_E_SHAPE;
    }
    if (b == t.end()) {
      set_error(std::string(which) + ": missing bias parameter: " + bn);  // src/lib.rs:127
      return NERF_E_SHAPE;
    }
    const HostTensor& kt = k->second;
    const HostTensor& bt = b->second;
    if (kt.dims.size() != 2 || kt.dims[0] != s.d0 || kt.dims[1] != s.d1 || (int64_t)kt.data.size() != s.d0 * s.d1) {
      set_error(std::string(which) + ": matrix dims mismatch for " + kn);  // src/lib.rs:119-120
      return NERF_E_SHAPE;
    }
    if (bt.dims.size() != 1 || bt.dims[0] != s.d1 || (int64_t)bt.data.size() != s.d1) {
      set_error(std::string(which) + ": bias dims mismatch for " + bn);  // src/lib.rs:128-129
      return NERF_E_SHAPE;
    }
  }
  return NERF_OK;
}

void normalize3(const float in[3], float out[3]) {
  const float len = std::sqrt(in[0] * in[0] + in[1] * in[1] + in[2] * in[2]);
  out[0] = in[0] / len;
  out[1] = in[1] / len;
  out[2] = in[2] / len;
}
void cross3(const float a[3], const float b[3], float out[3]) {
  const float x = a[1] * b[2] - a[2] * b[1];
  const float y = a[2] * b[0] - a[0] * b[2];
  const float z = a[0] * b[1] - a[1] * b[0];
  out[0] = x; out[1] = y; out[2] = z;
}

namespace {

// Minimal JSON reader for the flat camera file: finds "key" at any depth-1 position and parses the number
// or the array of numbers after the colon.  Enough for tf_reference_samples.json and orbit pose files.
bool find_key(const std::string& s, const std::string& key, size_t* pos) {
  const std::string pat = "\"" + key + "\"";
  size_t p = 0;
  while ((p = s.find(pat, p)) != std::string::npos) {
    size_t q = p + pat.size();
    while (q < s.size() && std::isspace((unsigned char)s[q])) ++q;
    if (q < s.size() && s[q] == ':') {
      *pos = q + 1;
      return true;
    }
    p = q;
  }
  return false;
}
bool parse_number(const std::string& s, size_t* pos, double* out) {
  while (*pos < s.size() && std::isspace((unsigned char)s[*pos])) ++*pos;
  const char* b = s.c_str() + *pos;
  char* e = nullptr;
  const double v = std::strtod(b, &e);
  if (e == b) return false;
  *pos += (size_t)(e - b);
  *out = v;
  return true;
}
bool get_number(const std::string& s, const std::string& key, double* out) {
  size_t p;
  return find_key(s, key, &p) && parse_number(s, &p, out);
}
bool get_array3(const std::string& s, const std::string& key, double out[3]) {
  size_t p;
  if (!find_key(s, key, &p)) return false;
  while (p < s.size() && std::isspace((unsigned char)s[p])) ++p;
  if (p >= s.size() || s[p] != '[') return false;
  ++p;
  for (int i = 0; i < 3; ++i) {
    if (!parse_number(s, &p, &out[i])) return false;
    while (p < s.size() && (std::isspace((unsigned char)s[p]) || s[p] == ',')) ++p;
  }
  return true;
}

inline uint8_t quantize_one(float x) {  // (clamp(v,0,1)*255+0.5) as u8; NaN -> 0 (src/lib.rs:574-576)
  if (x != x) return 0;
  x = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
  const float q = x * 255.0f + 0.5f;
  return (uint8_t)(q >= 255.0f ? 255 : (int)q);
}

}  // namespace
}  // namespace nerf

using namespace nerf;

extern "C" {

const char* nerf_last_error(void) { return get_error(); }
const char* nerf_version(void) { return "nerf_b200 0.1 (sm_100a)"; }

int nerf_camera_from_values(double near_, double far_, const double origin[3], const double forward[3],
                            const double up[3], const double hwf[3], nerf_camera* out) {
  if (!origin || !forward || !up || !hwf || !out) {
    set_error("nerf_camera_from_values: null argument");
    return NERF_E_ARG;
  }
  out->near_ = (float)near_;  // f64 -> f32 casts first (src/lib.rs:620-629)
  out->far_ = (float)far_;
  const float fwd[3] = {(float)forward[0], (float)forward[1], (float)forward[2]};
  const float upv[3] = {(float)up[0], (float)up[1], (float)up[2]};
  for (int i = 0; i < 3; ++i) out->origin[i] = (float)origin[i];
  normalize3(fwd, out->forward);  // src/lib.rs:623
  normalize3(upv, out->up);       // src/lib.rs:624
  const float hw = (float)hwf[1], hh = (float)hwf[0], focal = (float)hwf[2];
  out->alpha_w = std::atan((0.5f * hw) / focal);  // src/lib.rs:630
  out->alpha_h = std::atan((0.5f * hh) / focal);  // src/lib.rs:631
  return NERF_OK;
}

int nerf_camera_from_json(const char* path, nerf_camera* out) {
  if (!path || !out) {
    set_error("nerf_camera_from_json: null argument");
    return NERF_E_ARG;
  }
  std::ifstream f(path);
  if (!f) {
    set_error(std::string("read camera json: ") + path);
    return NERF_E_IO;
  }
  std::stringstream ss;
  ss << f.rdbuf();
  const std::string s = ss.str();
  double near_, far_, o[3], fw[3], up[3], hwf[3];
  if (!get_number(s, "near", &near_) || !get_number(s, "far", &far_) || !get_array3(s, "camera_origin", o) ||
      !get_array3(s, "camera_forward", fw) || !get_array3(s, "camera_up", up) || !get_array3(s, "hwf", hwf)) {
    set_error(std::string("camera json lacks near/far/camera_origin/camera_forward/camera_up/hwf: ") + path);
    return NERF_E_IO;
  }
  return nerf_camera_from_values(near_, far_, o, fw, up, hwf, out);
}

int nerf_quantize_rgb8(const float* rgb, int64_t n, uint8_t* out) {
  if ((!rgb || !out) && n > 0) {
    set_error("nerf_quantize_rgb8: null argument");
    return NERF_E_ARG;
  }
  for (int64_t i = 0; i < n; ++i) out[i] = quantize_one(rgb[i]);
  return NERF_OK;
}

int nerf_save_ppm(const char* path, int W, int H, const float* rgb) {
  if (!path || !rgb || W <= 0 || H <= 0) {
    set_error("nerf_save_ppm: bad argument");
    return NERF_E_ARG;
  }
  FILE* f = std::fopen(path, "wb");
  if (!f) {
    set_error(std::string("create ") + path);
    return NERF_E_IO;
  }
  std::fprintf(f, "P6\n%d %d\n255\n", W, H);  // src/lib.rs:571
  std::vector<uint8_t> buf((size_t)W * H * 3);
  for (size_t i = 0; i < buf.size(); ++i) buf[i] = quantize_one(rgb[i]);
  const size_t wr = std::fwrite(buf.data(), 1, buf.size(), f);
  std::fclose(f);
  if (wr != buf.size()) {
    set_error(std::string("short write ") + path);
    return NERF_E_IO;
  }
  return NERF_OK;
}

}  // extern "C"
