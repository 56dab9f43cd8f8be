// oracle.cpp — TEST INFRASTRUCTURE ONLY (see oracle.h).  CPU restatement of the nerf-rs hot path.
//
// Build: g++ -O3 -mavx2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/Makefile).
// -ffp-contract=off matters: rustc never contracts a*b+c into an FMA, so every mul/add below
// rounds exactly like the reference's f32 arithmetic.  libm calls (sinf/cosf/expf/tanf/atanf)
// are glibc's, the same functions Rust's f32::sin/cos/exp/tan/atan lower to on Linux.
//
// All file:line citations are into the reference crate (elisabeth96/nerf-rs).
#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

thread_local std::string g_err;

// ---------------------------------------------------------------- Vec3 (src/vec3.rs:3-47)
struct V3 {
  float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }   // vec3.rs:49-54
inline V3 muls(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }     // Mul<f32>
inline V3 cross(V3 a, V3 b) {                                               // vec3.rs:19-25
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float length(V3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }  // vec3.rs:27-29
inline V3 normalize(V3 a) {                                                          // vec3.rs:31-34
  float l = length(a);
  return v3(a.x / l, a.y / l, a.z / l);
}

// ---------------------------------------------------------------- operand rounding emulation
inline float round_bf16(float v) {
  uint32_t b;
  std::memcpy(&b, &v, 4);
  if ((b & 0x7f800000u) == 0x7f800000u) return v;
  b += 0x7fffu + ((b >> 16) & 1u);
  b &= 0xffff0000u;
  std::memcpy(&v, &b, 4);
  return v;
}
inline float round_tf32(float v) {
  uint32_t b;
  std::memcpy(&b, &v, 4);
  if ((b & 0x7f800000u) == 0x7f800000u) return v;
  b += 0xfffu + ((b >> 13) & 1u);
  b &= 0xffffe000u;
  std::memcpy(&v, &b, 4);
  return v;
}
inline float round_fp16(float v) {
  // RNE to IEEE binary16 (with subnormals and overflow to inf), returned as f32.
  return (float)(_Float16)v;
}
inline float round_mode_apply(float v, int mode) {
  switch (mode) {
    case ORACLE_ROUND_BF16: return round_bf16(v);
    case ORACLE_ROUND_FP16: return round_fp16(v);
    case ORACLE_ROUND_TF32: return round_tf32(v);
    default: return v;
  }
}

// ---------------------------------------------------------------- Layer / Network (src/network.rs:62-238)
struct Layer {
  int in = 0, out = 0;
  std::vector<float> w;  // [in][out] row-major, exactly the .bin contents (network.rs:63, lib.rs:121)
  std::vector<float> b;  // [out]
  mutable std::vector<float> w_rounded[4];  // built by prepare() before any parallel region
  void prepare(int mode) const {
    if (mode == ORACLE_ROUND_NONE || !w_rounded[mode].empty()) return;
    std::vector<float> tmp(w.size());
    for (size_t i = 0; i < w.size(); ++i) tmp[i] = round_mode_apply(w[i], mode);
    w_rounded[mode].swap(tmp);
  }
  const float* weights(int mode) const { return mode == ORACLE_ROUND_NONE ? w.data() : w_rounded[mode].data(); }
  // The tensor-core kernel preloads the f32 accumulator with the bias, so the bias is never rounded.
  const float* bias(int) const { return b.data(); }
};

enum Act { RELU, SIGMOID, NONE };  // network.rs:67-72

inline float activate(float v, Act a) {  // network.rs:161-169
  switch (a) {
    case RELU: return v > 0.0f ? v : 0.0f;  // f32::max(v, 0.0); NaN -> 0 as in Rust
    case SIGMOID: return 1.0f / (1.0f + std::exp(-v));
    default: return v;
  }
}

// out[o][c] = act(b[o] + sum_{i ascending} W[i][o] * h[i][c])      (network.rs:125-147, fill_with_bias :149-159)
// Reordered variant: per block of 16 columns, i outer / o inner.  Per output element the operation
// sequence is bias, then +W[0][o]*h[0], +W[1][o]*h[1], ... exactly as the reference's in->col->out
// loops produce it (each element's partial sums are independent of the others), so it is bit-identical.
void dense_reordered(const Layer& L, const float* W, const float* bias, const float* h, int64_t B, Act act,
                     float* out) {
  const int in = L.in, no = L.out;
  constexpr int CB = 16;  // one cache line of activations per input row
  std::vector<float> acc((size_t)CB * no), hb((size_t)in * CB);
  for (int64_t c0 = 0; c0 < B; c0 += CB) {
    const int cb = (int)std::min<int64_t>(CB, B - c0);
    for (int i = 0; i < in; ++i)
      for (int k = 0; k < cb; ++k) hb[(size_t)i * CB + k] = h[(size_t)i * B + c0 + k];
    for (int k = 0; k < cb; ++k)
      for (int o = 0; o < no; ++o) acc[(size_t)k * no + o] = bias[o];
    for (int i = 0; i < in; ++i) {
      const float* __restrict wr = W + (size_t)i * no;
      for (int k = 0; k < cb; ++k) {
        const float hv = hb[(size_t)i * CB + k];
        float* __restrict a = acc.data() + (size_t)k * no;
#pragma omp simd
        for (int o = 0; o < no; ++o) a[o] = a[o] + wr[o] * hv;
      }
    }
    for (int k = 0; k < cb; ++k)
      for (int o = 0; o < no; ++o) out[(size_t)o * B + c0 + k] = activate(acc[(size_t)k * no + o], act);
  }
}

// The reference's literal loop nest (network.rs:134-143): in -> col -> out with strided get/set.
void dense_literal(const Layer& L, const float* W, const float* bias, const float* h, int64_t B, Act act,
                   float* out) {
  const int in = L.in, no = L.out;
  for (int o = 0; o < no; ++o)
    for (int64_t c = 0; c < B; ++c) out[(size_t)o * B + c] = bias[o];
  for (int i = 0; i < in; ++i) {
    const float* wr = W + (size_t)i * no;
    for (int64_t c = 0; c < B; ++c) {
      const float hv = h[(size_t)i * B + c];
      for (int o = 0; o < no; ++o) {
        float cur = out[(size_t)o * B + c] + wr[o] * hv;
        out[(size_t)o * B + c] = cur;
      }
    }
  }
  for (size_t k = 0; k < (size_t)no * B; ++k) out[k] = activate(out[k], act);
}

// rm selects the operand set: ORACLE_ROUND_NONE is the reference's f32 arithmetic.
void dense(const Layer& L, int rm, const float* h, int64_t B, Act act, int order, float* out) {
  if (order == ORACLE_GEMM_LITERAL)
    dense_literal(L, L.weights(rm), L.bias(rm), h, B, act, out);
  else
    dense_reordered(L, L.weights(rm), L.bias(rm), h, B, act, out);
}

// positional_encoding_batch / _dirs (network.rs:263-330): rows [v, sin(f v)x3, cos(f v)x3, ...], f doubled.
void positional_encoding(const float* x, const float* y, const float* z, int64_t stride, int64_t B, int n,
                         float* enc /*[3+6n][B]*/) {
  for (int64_t c = 0; c < B; ++c) {
    const float p[3] = {x[c * stride], y[c * stride], z[c * stride]};
    enc[0 * B + c] = p[0];
    enc[1 * B + c] = p[1];
    enc[2 * B + c] = p[2];
    float f = 1.0f;
    int row = 3;
    for (int k = 0; k < n; ++k) {
      for (int a = 0; a < 3; ++a) enc[(size_t)(row++) * B + c] = std::sin(f * p[a]);
      for (int a = 0; a < 3; ++a) enc[(size_t)(row++) * B + c] = std::cos(f * p[a]);
      f *= 2.0f;
    }
  }
}

void round_inplace(std::vector<float>& v, int mode) {
  if (mode == ORACLE_ROUND_NONE) return;
  for (auto& e : v) e = round_mode_apply(e, mode);
}

}  // namespace

struct oracle_net {
  Layer layers[8];
  Layer bottleneck, viewdirs, rgb, alpha;
  std::map<std::string, std::pair<std::vector<int64_t>, std::vector<float>>> raw;
  void prepare(int mode) const {
    for (auto& l : layers) l.prepare(mode);
    bottleneck.prepare(mode);
    viewdirs.prepare(mode);
  }
};

namespace {

// Network::forward_batch (network.rs:197-237).
// round_mode != 0 emulates the tensor-core kernel's numerics (operands of the 256-wide GEMMs rounded to
// the 16/19-bit format, f32 accumulation; sigma and rgb heads evaluated in f32 on unrounded inputs).
void forward_batch(const oracle_net& N, const float* points, const float* dirs, int64_t B, int order,
                   int rm, float* rgb, float* sigma) {
  if (B == 0) return;  // network.rs:199-201
  std::vector<float> h0((size_t)63 * B);
  positional_encoding(points, points + B, points + 2 * B, 1, B, 10, h0.data());  // network.rs:204
  round_inplace(h0, rm);
  std::vector<float> h = h0, nxt((size_t)256 * B);
  for (int i = 0; i < 5; ++i) {  // network.rs:206-208
    dense(N.layers[i], rm, h.data(), B, RELU, order, nxt.data());
    h = nxt;
    round_inplace(h, rm);
  }
  std::vector<float> cat((size_t)319 * B);  // concat_rows(h_0, h4): h_0 FIRST (network.rs:210, 240-261)
  std::memcpy(cat.data(), h0.data(), sizeof(float) * 63 * B);
  std::memcpy(cat.data() + (size_t)63 * B, h.data(), sizeof(float) * 256 * B);
  h = cat;
  std::vector<float> h8_f32;
  for (int i = 5; i < 8; ++i) {  // network.rs:211-213
    dense(N.layers[i], rm, h.data(), B, RELU, order, nxt.data());
    h = nxt;
    if (i == 7) h8_f32 = h;
    round_inplace(h, rm);
  }
  // sigma = relu(alpha(h8))  (network.rs:216): f32 head on the unrounded h8 in every mode.
  std::vector<float> sig((size_t)B);
  dense(N.alpha, ORACLE_ROUND_NONE, h8_f32.data(), B, RELU, order, sig.data());
  std::memcpy(sigma, sig.data(), sizeof(float) * B);  // row_slice(0) (network.rs:234)

  std::vector<float> bn((size_t)256 * B);  // network.rs:218, ActivationType::None
  dense(N.bottleneck, rm, h.data(), B, NONE, order, bn.data());
  round_inplace(bn, rm);
  std::vector<float> q((size_t)283 * B);  // concat_rows(bottleneck, PE4(dirs)): bottleneck FIRST (network.rs:219-220)
  std::memcpy(q.data(), bn.data(), sizeof(float) * 256 * B);
  {
    std::vector<float> dpe((size_t)27 * B);
    positional_encoding(dirs, dirs + 1, dirs + 2, 3, B, 4, dpe.data());
    round_inplace(dpe, rm);
    std::memcpy(q.data() + (size_t)256 * B, dpe.data(), sizeof(float) * 27 * B);
  }
  std::vector<float> ch((size_t)128 * B), cm((size_t)3 * B);
  dense(N.viewdirs, rm, q.data(), B, RELU, order, ch.data());  // network.rs:222
  dense(N.rgb, ORACLE_ROUND_NONE, ch.data(), B, SIGMOID, order, cm.data());           // network.rs:223 (f32 head)
  for (int64_t c = 0; c < B; ++c) {  // network.rs:225-232
    rgb[c * 3 + 0] = cm[0 * B + c];
    rgb[c * 3 + 1] = cm[1 * B + c];
    rgb[c * 3 + 2] = cm[2 * B + c];
  }
}

// Camera::get_ray_dir (lib.rs:214-230); `offset` is 0.5 in the reference.
V3 get_ray_dir(const oracle_camera& cam, int nx, int ny, int i, int j, float offset) {
  V3 f = normalize(v3(cam.dir[0], cam.dir[1], cam.dir[2]));
  V3 r = normalize(cross(f, v3(cam.up[0], cam.up[1], cam.up[2])));
  V3 u = normalize(cross(r, f));
  float x = (((float)j + offset) / (float)nx) * 2.0f - 1.0f;
  float y = 1.0f - (((float)i + offset) / (float)ny) * 2.0f;
  float sx = std::tan(cam.alpha_w);
  float sy = std::tan(cam.alpha_h);
  return add(add(muls(r, x * sx), muls(u, y * sy)), f);
}

// Philox4x32-10 (Salmon et al. 2011), shared with the CUDA path.
inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

}  // namespace

extern "C" {

const char* oracle_last_error(void) { return g_err.c_str(); }

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// load_shapes (lib.rs:63-74) + load_tensor (lib.rs:35-42) + load_network_from_dir (lib.rs:108-174).
oracle_net* oracle_load_network(const char* dir) {
  std::string root(dir);
  std::ifstream sf(root + "/shapes.txt");
  if (!sf) {
    g_err = "read shapes: " + root + "/shapes.txt";
    return nullptr;
  }
  auto* net = new oracle_net();
  std::string line;
  while (std::getline(sf, line)) {
    std::istringstream ls(line);
    std::string name;
    if (!(ls >> name)) continue;
    std::vector<int64_t> dims;
    int64_t d;
    while (ls >> d) dims.push_back(d);
    std::ifstream bf(root + "/" + name + ".bin", std::ios::binary);
    if (!bf) {
      g_err = "read tensor: " + name;
      delete net;
      return nullptr;
    }
    std::vector<char> bytes((std::istreambuf_iterator<char>(bf)), std::istreambuf_iterator<char>());
    std::vector<float> data(bytes.size() / 4);  // chunks_exact(4), little-endian f32
    std::memcpy(data.data(), bytes.data(), data.size() * 4);
    net->raw[name] = {dims, data};
  }
  auto take = [&](const char* kn, const char* bn, Layer& L) -> bool {
    auto k = net->raw.find(kn);
    auto b = net->raw.find(bn);
    if (k == net->raw.end()) { g_err = std::string("missing matrix parameter: ") + kn; return false; }
    if (b == net->raw.end()) { g_err = std::string("missing bias parameter: ") + bn; return false; }
    if (k->second.first.size() != 2 || b->second.first.size() != 1) { g_err = "dims mismatch"; return false; }
    L.in = (int)k->second.first[0];
    L.out = (int)k->second.first[1];
    L.w = k->second.second;
    L.b = b->second.second;
    if ((int64_t)L.w.size() != (int64_t)L.in * L.out || (int)L.b.size() != L.out) { g_err = "size mismatch"; return false; }
    return true;
  };
  bool ok = true;
  for (int i = 0; i < 8 && ok; ++i) {
    std::string kn = "dense" + std::to_string(i) + "_kernel", bn = "dense" + std::to_string(i) + "_bias";
    ok = take(kn.c_str(), bn.c_str(), net->layers[i]);
  }
  ok = ok && take("bottleneck_kernel", "bottleneck_bias", net->bottleneck);
  ok = ok && take("viewdirs_kernel", "viewdirs_bias", net->viewdirs);
  ok = ok && take("rgb_kernel", "rgb_bias", net->rgb);
  ok = ok && take("alpha_kernel", "alpha_bias", net->alpha);
  if (!ok) {
    delete net;
    return nullptr;
  }
  return net;
}

void oracle_free_network(oracle_net* n) { delete n; }

int64_t oracle_get_tensor(const oracle_net* n, const char* name, float* out, int64_t cap) {
  auto it = n->raw.find(name);
  if (it == n->raw.end()) return -1;
  int64_t cnt = (int64_t)it->second.second.size();
  if (out && cap >= cnt) std::memcpy(out, it->second.second.data(), cnt * 4);
  return cnt;
}

int oracle_forward_batch(const oracle_net* net, const float* points, const float* dirs, int64_t B,
                         int gemm_order, int round_mode, float* rgb, float* sigma) {
  if (!net || B < 0) { g_err = "bad argument"; return -1; }
  net->prepare(round_mode);
  forward_batch(*net, points, dirs, B, gemm_order, round_mode, rgb, sigma);
  return 0;
}

// camera_from_samples (lib.rs:614-645) + get_vec3 (lib.rs:594-601): f64 JSON numbers cast to f32 first.
void oracle_camera_from_values(double near_, double far_, const double origin[3], const double forward[3],
                               const double up[3], const double hwf[3], oracle_camera* out) {
  out->near_ = (float)near_;
  out->far_ = (float)far_;
  V3 pos = v3((float)origin[0], (float)origin[1], (float)origin[2]);
  V3 dir = normalize(v3((float)forward[0], (float)forward[1], (float)forward[2]));
  V3 upv = normalize(v3((float)up[0], (float)up[1], (float)up[2]));
  float hw = (float)hwf[1], hh = (float)hwf[0], focal = (float)hwf[2];
  out->alpha_w = std::atan((0.5f * hw) / focal);
  out->alpha_h = std::atan((0.5f * hh) / focal);
  out->pos[0] = pos.x; out->pos[1] = pos.y; out->pos[2] = pos.z;
  out->dir[0] = dir.x; out->dir[1] = dir.y; out->dir[2] = dir.z;
  out->up[0] = upv.x; out->up[1] = upv.y; out->up[2] = upv.z;
}

void oracle_ray_dirs(const oracle_camera* cam, int nx, int ny, float offset, int do_norm, int64_t ray_begin,
                     int64_t ray_end, float* out) {
  for (int64_t p = ray_begin; p < ray_end; ++p) {
    int i = (int)(p / nx), j = (int)(p % nx);  // pixel_index = i*nx + j (lib.rs:371)
    V3 d = get_ray_dir(*cam, nx, ny, i, j, offset);
    if (do_norm) d = normalize(d);  // lib.rs:371
    float* o = out + (p - ray_begin) * 3;
    o[0] = d.x; o[1] = d.y; o[2] = d.z;
  }
}

// stratified_samples (lib.rs:233-248).  `upper - lower` is recomputed as the reference does.
void oracle_stratified(float near_, float far_, int count, const float* jitter, float* t) {
  if (count == 0) return;
  float interval = (far_ - near_) / (float)count;
  for (int i = 0; i < count; ++i) {
    float lower = near_ + (float)i * interval;
    float upper = lower + interval;
    t[i] = lower + (upper - lower) * jitter[i];
  }
}

// compute_weights (lib.rs:250-283).
void oracle_compute_weights(const float* sigma, const float* t, int n, float far_, float* w) {
  float transmittance = 1.0f;
  int i = 0;
  for (; i < n; ++i) {
    float delta = (i + 1 < n) ? (t[i + 1] - t[i]) : (far_ - t[i]);
    if (delta < 0.0f) delta = 0.0f;
    float alpha = 1.0f - std::exp(-sigma[i] * delta);
    w[i] = transmittance * alpha;
    transmittance *= 1.0f - alpha;
    if (transmittance < 1e-4f) {  // lib.rs:276-279: the rest are exactly zero
      for (int k = i + 1; k < n; ++k) w[k] = 0.0f;
      break;
    }
  }
}

// midpoints (lib.rs:285-287) + sample_importance (lib.rs:289-351).
int oracle_sample_importance(const float* t, const float* w, int n, const float* u, int count, float* out) {
  if (count == 0 || n < 3) return 0;  // lib.rs:295-297
  const int m = n - 2;                // pdf_weights = weights[1..n-1]
  if (m <= 0) return 0;
  std::vector<float> bins(n - 1);
  for (int i = 0; i + 1 < n; ++i) bins[i] = 0.5f * (t[i] + t[i + 1]);
  if ((int)bins.size() < 2 || (int)bins.size() != m + 1) return 0;
  std::vector<float> adj(m);
  for (int i = 0; i < m; ++i) {
    float v = w[i + 1];
    v = (v > 0.0f) ? v : 0.0f;  // f32::max(w, 0.0)
    adj[i] = v + 1e-5f;
  }
  float sum = 0.0f;
  for (int i = 0; i < m; ++i) sum += adj[i];
  if (sum <= 0.0f) return 0;
  for (int i = 0; i < m; ++i) adj[i] /= sum;
  std::vector<float> cdf(m + 1);
  cdf[0] = 0.0f;
  float cumulative = 0.0f;
  for (int i = 0; i < m; ++i) {
    cumulative += adj[i];
    cdf[i + 1] = cumulative;
  }
  cdf[m] = 1.0f;  // lib.rs:326-328
  for (int k = 0; k < count; ++k) {
    float uu = u[k];
    int idx = m - 1;
    for (int j = 0; j < m; ++j) {
      if (uu >= cdf[j] && uu < cdf[j + 1]) {
        idx = j;
        break;
      }
    }
    float lo = cdf[idx], hi = cdf[idx + 1];
    float denom = hi - lo;
    if (!(denom > 1e-6f)) denom = 1e-6f;  // f32::max(denom, 1e-6)
    float bl = bins[idx], bu = bins[idx + 1];
    float tt = (uu - lo) / denom;
    out[k] = bl + (bu - bl) * tt;
  }
  return count;
}

// Depth and accumulated opacity of a ray.  The reference has neither output; the definition used here is the usual
// one over the weights of compute_weights (lib.rs:250-283) that integrate_ray (lib.rs:176-195) also uses:
// acc = sum_i w_i (so that integrate_ray's background term is 1 - acc), depth = sum_i w_i * t_i, f32, ascending i.
void oracle_integrate_aux(const float* sigma, const float* t, int n, float far_, float out[2]) {
  out[0] = out[1] = 0.0f;
  if (n == 0) return;
  std::vector<float> w(n);
  oracle_compute_weights(sigma, t, n, far_, w.data());
  float depth = 0.0f, acc = 0.0f;
  for (int i = 0; i < n; ++i) {
    depth += w[i] * t[i];
    acc += w[i];
  }
  out[0] = depth;
  out[1] = acc;
}

// integrate_ray (lib.rs:176-195).
void oracle_integrate_ray(const float* rgb, const float* sigma, const float* t, int n, float far_, float out[3]) {
  if (n == 0) {
    out[0] = out[1] = out[2] = 0.0f;
    return;
  }
  std::vector<float> w(n);
  oracle_compute_weights(sigma, t, n, far_, w.data());
  float r = 0.0f, g = 0.0f, b = 0.0f, acc = 0.0f;
  for (int i = 0; i < n; ++i) {
    r += rgb[i * 3 + 0] * w[i];
    g += rgb[i * 3 + 1] * w[i];
    b += rgb[i * 3 + 2] * w[i];
    acc += w[i];
  }
  float bg = 1.0f * (1.0f - acc);
  out[0] = r + bg;
  out[1] = g + bg;
  out[2] = b + bg;
}

void oracle_philox_uniforms(uint64_t seed, uint32_t pixel_index, uint32_t stream, int count, float* out) {
  for (int k0 = 0; k0 < count; k0 += 4) {
    uint32_t c[4] = {pixel_index, (uint32_t)(k0 / 4), stream, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    for (int e = 0; e < 4 && k0 + e < count; ++e) out[k0 + e] = (float)(c[e] >> 8) * (1.0f / 16777216.0f);
  }
}

// render_image (lib.rs:474-565) driving render_block (lib.rs:353-472) over 8x8 pixel blocks.
int oracle_render(const oracle_net* coarse, const oracle_net* fine, const oracle_camera* cam, int W, int H,
                  int nc, int nf, const oracle_sampling* smp, int64_t ray_begin, int64_t ray_end, int order,
                  int rm, int n_threads, float* rgb_out, const oracle_stages* st) {
  if (nc <= 0) { g_err = "coarse samples per ray must be greater than 0"; return -1; }   // lib.rs:483-486
  const int bs = 8;
  if (W % bs != 0) { g_err = "image width must be multiple of block size"; return -1; }   // lib.rs:492-496
  if (H % bs != 0) { g_err = "image height must be multiple of block size"; return -1; }  // lib.rs:497-501
  if (ray_begin < 0 || ray_end > (int64_t)W * H || ray_begin > ray_end) { g_err = "bad ray range"; return -1; }
  coarse->prepare(rm);
  fine->prepare(rm);
  const V3 origin = v3(cam->pos[0], cam->pos[1], cam->pos[2]);
  const int bxn = W / bs, byn = H / bs;
  const int mode = smp ? smp->mode : ORACLE_SAMPLING_DET;
#ifdef _OPENMP
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#else
  n_threads = 1;
#endif
#pragma omp parallel for schedule(dynamic, 1) num_threads(n_threads)
  for (int blk = 0; blk < bxn * byn; ++blk) {
    const int by = (blk / bxn) * bs, bx = (blk % bxn) * bs;  // row-major block order (lib.rs:503-510)
    // rays of this block that fall inside [ray_begin, ray_end)
    std::vector<int64_t> pix;
    std::vector<V3> dirs;
    for (int i = by; i < by + bs; ++i)
      for (int j = bx; j < bx + bs; ++j) {
        int64_t p = (int64_t)i * W + j;
        if (p < ray_begin || p >= ray_end) continue;
        pix.push_back(p);
        dirs.push_back(normalize(get_ray_dir(*cam, W, H, i, j, 0.5f)));  // lib.rs:370-371
      }
    const int nr = (int)pix.size();
    if (nr == 0) continue;
    // coarse t (lib.rs:375-386)
    std::vector<float> tc((size_t)nr * nc), jit(nc);
    for (int r = 0; r < nr; ++r) {
      const int64_t li = pix[r] - ray_begin;
      if (mode == ORACLE_SAMPLING_DET) {
        for (int s = 0; s < nc; ++s) jit[s] = 0.5f;
      } else if (mode == ORACLE_SAMPLING_PHILOX) {
        oracle_philox_uniforms(smp->seed, (uint32_t)pix[r], 0u, nc, jit.data());
      } else {
        std::memcpy(jit.data(), smp->jitter + li * nc, sizeof(float) * nc);
      }
      oracle_stratified(cam->near_, cam->far_, nc, jit.data(), &tc[(size_t)r * nc]);
    }
    // pack (lib.rs:388-402): p = origin + dir_hat * t, column = ray*nc + s
    const int64_t Bc = (int64_t)nr * nc;
    std::vector<float> pts((size_t)3 * Bc), vd((size_t)3 * Bc);
    for (int r = 0; r < nr; ++r)
      for (int s = 0; s < nc; ++s) {
        int64_t col = (int64_t)r * nc + s;
        V3 p = add(origin, muls(dirs[r], tc[col]));
        pts[0 * Bc + col] = p.x; pts[1 * Bc + col] = p.y; pts[2 * Bc + col] = p.z;
        vd[col * 3 + 0] = dirs[r].x; vd[col * 3 + 1] = dirs[r].y; vd[col * 3 + 2] = dirs[r].z;
      }
    std::vector<float> crgb((size_t)3 * Bc), csig((size_t)Bc);
    forward_batch(*coarse, pts.data(), vd.data(), Bc, order, rm, crgb.data(), csig.data());  // lib.rs:404
    // per ray: weights, importance samples, merge + sort (lib.rs:409-421)
    const int nm = nc + nf;
    std::vector<int> fcount(nr);
    std::vector<float> tf((size_t)nr * nm), wc(nc), uu(std::max(nf, 1));
    for (int r = 0; r < nr; ++r) {
      const int64_t li = pix[r] - ray_begin;
      oracle_compute_weights(&csig[(size_t)r * nc], &tc[(size_t)r * nc], nc, cam->far_, wc.data());
      if (mode == ORACLE_SAMPLING_DET) {
        for (int k = 0; k < nf; ++k) uu[k] = ((float)k + 0.5f) / (float)nf;
      } else if (mode == ORACLE_SAMPLING_PHILOX) {
        oracle_philox_uniforms(smp->seed, (uint32_t)pix[r], 1u, nf, uu.data());
      } else if (nf > 0) {
        std::memcpy(uu.data(), smp->u + li * nf, sizeof(float) * nf);
      }
      float* m = &tf[(size_t)r * nm];
      std::memcpy(m, &tc[(size_t)r * nc], sizeof(float) * nc);
      int got = oracle_sample_importance(&tc[(size_t)r * nc], wc.data(), nc, uu.data(), nf, m + nc);
      fcount[r] = nc + got;
      std::stable_sort(m, m + fcount[r]);  // sort_by(partial_cmp) is a stable sort (lib.rs:419)
      if (st) {
        if (st->dirs) { st->dirs[li * 3] = dirs[r].x; st->dirs[li * 3 + 1] = dirs[r].y; st->dirs[li * 3 + 2] = dirs[r].z; }
        if (st->t_c) std::memcpy(st->t_c + li * nc, &tc[(size_t)r * nc], sizeof(float) * nc);
        if (st->sigma_c) std::memcpy(st->sigma_c + li * nc, &csig[(size_t)r * nc], sizeof(float) * nc);
        if (st->w_c) std::memcpy(st->w_c + li * nc, wc.data(), sizeof(float) * nc);
        if (st->t_f) std::memcpy(st->t_f + li * nm, m, sizeof(float) * fcount[r]);
      }
    }
    // fine pack (lib.rs:423-443), ragged-capable through offsets
    int64_t Bf = 0;
    std::vector<int64_t> off(nr);
    for (int r = 0; r < nr; ++r) { off[r] = Bf; Bf += fcount[r]; }
    std::vector<float> fp((size_t)3 * Bf), fvd((size_t)3 * Bf);
    for (int r = 0; r < nr; ++r)
      for (int s = 0; s < fcount[r]; ++s) {
        int64_t col = off[r] + s;
        V3 p = add(origin, muls(dirs[r], tf[(size_t)r * nm + s]));
        fp[0 * Bf + col] = p.x; fp[1 * Bf + col] = p.y; fp[2 * Bf + col] = p.z;
        fvd[col * 3 + 0] = dirs[r].x; fvd[col * 3 + 1] = dirs[r].y; fvd[col * 3 + 2] = dirs[r].z;
      }
    std::vector<float> frgb((size_t)3 * Bf), fsig((size_t)Bf);
    forward_batch(*fine, fp.data(), fvd.data(), Bf, order, rm, frgb.data(), fsig.data());  // lib.rs:445
    for (int r = 0; r < nr; ++r) {  // lib.rs:447-459
      const int64_t li = pix[r] - ray_begin;
      oracle_integrate_ray(&frgb[(size_t)off[r] * 3], &fsig[off[r]], &tf[(size_t)r * nm], fcount[r], cam->far_,
                           rgb_out + li * 3);
      if (st) {
        if (st->sigma_f) std::memcpy(st->sigma_f + li * nm, &fsig[off[r]], sizeof(float) * fcount[r]);
        if (st->rgb_f) std::memcpy(st->rgb_f + li * nm * 3, &frgb[(size_t)off[r] * 3], sizeof(float) * 3 * fcount[r]);
      }
    }
  }
  return 0;
}

// save_ppm quantiser (lib.rs:574-576): (clamp(v,0,1)*255+0.5) as u8; Rust's `as u8` saturates, NaN -> 0.
void oracle_quantize(const float* v, int64_t n, uint8_t* out) {
  for (int64_t i = 0; i < n; ++i) {
    float x = v[i];
    if (x != x) { out[i] = 0; continue; }
    x = x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x);
    float q = x * 255.0f + 0.5f;
    out[i] = (uint8_t)(q >= 255.0f ? 255 : (q <= 0.0f ? 0 : (int)q));
  }
}

// save_ppm (lib.rs:567-580).
int oracle_save_ppm(const char* path, int W, int H, const float* rgb) {
  FILE* f = std::fopen(path, "wb");
  if (!f) { g_err = std::string("create ") + path; return -1; }
  std::fprintf(f, "P6\n%d %d\n255\n", W, H);
  std::vector<uint8_t> buf((size_t)W * H * 3);
  oracle_quantize(rgb, (int64_t)W * H * 3, buf.data());
  std::fwrite(buf.data(), 1, buf.size(), f);
  std::fclose(f);
  return 0;
}

}  // extern "C"
