// nerf_rs.hpp — the reference crate's own surface (elisabeth96/nerf-rs: src/lib.rs, src/network.rs, src/vec3.rs) as a
// header-only C++17 layer over the C ABI of libnerf_b200.so (nerf_b200.h).
//
// The reference is Rust and no Rust toolchain exists in this environment, so this is the compiled-language mirror of
// its API: same names, same argument meaning, same error behaviour (the reference panics - here a nerf_rs::Panic is
// thrown with the library's message), so that code and tests written against the crate read the same:
//
//   reference (Rust)                                          here (C++)
//   nerf_rs::vec3::Vec3 {x, y, z}            vec3.rs:3-47     nerf_rs::Vec3
//   nerf_rs::network::Matrix                 network.rs:3-60  nerf_rs::Matrix      (row-major f32, dims i32)
//   nerf_rs::network::Layer::new             network.rs:75    nerf_rs::Layer
//   nerf_rs::network::Network::new           network.rs:181   nerf_rs::Network::create
//   nerf_rs::network::Network::forward_batch network.rs:197   nerf_rs::Network::forward_batch
//   load_network_from_dir(dir)               lib.rs:108-174   nerf_rs::load_network_from_dir
//   Camera / camera_from_samples             lib.rs:197-211,614-645   nerf_rs::Camera / camera_from_samples
//   render_image(&coarse, &fine, &cam, n_f)  lib.rs:474-565   nerf_rs::render_image
//   save_ppm(path, w, h, pixels)             lib.rs:567-580   nerf_rs::save_ppm
//   render_cli_image()                       lib.rs:647-677   nerf_rs::render_cli_image
//
// There is no CPU arithmetic in this file: every forward pass and every render goes to the GPU through the C ABI, and
// without an sm_100 device every entry point throws.
#ifndef NERF_RS_HPP
#define NERF_RS_HPP

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "nerf_b200.h"

namespace nerf_rs {

// What a Rust panic becomes on this side of the ABI.
struct Panic : std::runtime_error {
  using std::runtime_error::runtime_error;
};
inline void check(int rc, const char* what) {
  if (rc != NERF_OK) throw Panic(std::string(what) + ": " + nerf_last_error());
}

// ---------------------------------------------------------------------------------------------- vec3.rs
struct Vec3 {  // #[repr(C)]-compatible: three consecutive f32, as the ABI's [B][3] arrays expect
  float x, y, z;
  Vec3() : x(0), y(0), z(0) {}
  Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
  float dot(const Vec3& o) const { return x * o.x + y * o.y + z * o.z; }
  Vec3 cross(const Vec3& o) const { return Vec3(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x); }
  float length() const { return std::sqrt(dot(*this)); }
  Vec3 normalize() const {  // three divisions by the length, like vec3.rs:31-34
    const float l = length();
    return Vec3(x / l, y / l, z / l);
  }
};
inline Vec3 operator+(const Vec3& a, const Vec3& b) { return Vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return Vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vec3 operator*(const Vec3& a, float s) { return Vec3(a.x * s, a.y * s, a.z * s); }
static_assert(sizeof(Vec3) == 12, "Vec3 must be three packed floats");

// ---------------------------------------------------------------------------------------------- network.rs
class Matrix {  // row-major; activations are [features x batch], a column is a sample (network.rs:3-60)
 public:
  Matrix(int32_t rows, int32_t cols, std::vector<float> data) : rows_(rows), cols_(cols), data_(std::move(data)) {
    if (rows < 0 || cols < 0 || data_.size() != (size_t)rows * (size_t)cols) throw Panic("Matrix::new: data length");
  }
  static Matrix zeros(int32_t rows, int32_t cols) { return Matrix(rows, cols, std::vector<float>((size_t)rows * cols)); }
  int32_t rows() const { return rows_; }
  int32_t cols() const { return cols_; }
  const std::vector<float>& data() const { return data_; }
  std::vector<float>& data_mut() { return data_; }
  float get(int32_t r, int32_t c) const { return data_.at((size_t)r * cols_ + c); }
  void set(int32_t r, int32_t c, float v) { data_.at((size_t)r * cols_ + c) = v; }
  const float* row_slice(int32_t r) const { return data_.data() + (size_t)r * cols_; }

 private:
  int32_t rows_, cols_;
  std::vector<float> data_;
};

// network.rs:62-78: a dense layer as the caller holds it - weight is [in x out] row-major (the file layout of
// lego_rust/*/_kernel.bin), bias has `out` entries.  Host data only; the arithmetic happens on the GPU once the layers
// are handed to Network::create.
struct Layer {
  Matrix weight;
  std::vector<float> bias;
  Layer(Matrix weight_, std::vector<float> bias_) : weight(std::move(weight_)), bias(std::move(bias_)) {  // Layer::new
    if ((size_t)weight.cols() != bias.size()) throw Panic("Layer::new: bias length != weight.cols()");
  }
};

// The tensors of one network in the order and under the names of shapes.txt (lib.rs:108-174).
struct HostNet {
  std::vector<Layer> layers;  // dense0..7, bottleneck, viewdirs, rgb, alpha
  std::vector<std::string> names;
  std::vector<nerf_tensor> tensors() const {
    std::vector<nerf_tensor> t;
    for (size_t i = 0; i < layers.size(); ++i) {
      nerf_tensor k{}, b{};
      k.name = names[2 * i].c_str();
      k.ndim = 2;
      k.dims[0] = layers[i].weight.rows();
      k.dims[1] = layers[i].weight.cols();
      k.data = layers[i].weight.data().data();
      b.name = names[2 * i + 1].c_str();
      b.ndim = 1;
      b.dims[0] = (int64_t)layers[i].bias.size();
      b.data = layers[i].bias.data();
      t.push_back(k);
      t.push_back(b);
    }
    return t;
  }
};

// One GPU context per weight root (it holds both networks of a scene: a render needs both).
class Context {
 public:
  Context(const std::string& root, int device) {
    check(nerf_create(root.c_str(), device, &ctx_), "load_network_from_dir");
  }
  Context(const HostNet& coarse, const HostNet& fine, int device) {  // Network::new (network.rs:181-195)
    const std::vector<nerf_tensor> c = coarse.tensors(), f = fine.tensors();
    check(nerf_create_from_tensors(c.data(), (int)c.size(), f.data(), (int)f.size(), device, &ctx_), "Network::new");
  }
  ~Context() { nerf_destroy(ctx_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  nerf_ctx* get() const { return ctx_; }

 private:
  nerf_ctx* ctx_ = nullptr;
};

class Network {
 public:
  Network(std::shared_ptr<Context> ctx, int net_id) : ctx_(std::move(ctx)), net_(net_id) {}
  // Network::new (network.rs:181-195): eight trunk layers, then bottleneck, viewdirs, rgb, alpha.  A GPU context
  // always holds the two networks of a scene, so a network built on its own sits in both slots of its context; hand
  // two of them to render_image and it builds (and keeps) the coarse + fine context of the pair.
  static Network create(std::vector<Layer> layers, Layer bottleneck, Layer viewdirs, Layer rgb, Layer alpha,
                        int device = 0) {
    if (layers.size() != 8) throw Panic("Network::new: the trunk has eight layers (network.rs:206-213)");
    auto host = std::make_shared<HostNet>();
    host->layers = std::move(layers);
    host->layers.push_back(std::move(bottleneck));
    host->layers.push_back(std::move(viewdirs));
    host->layers.push_back(std::move(rgb));
    host->layers.push_back(std::move(alpha));
    static const char* const stem[12] = {"dense0", "dense1", "dense2", "dense3", "dense4", "dense5",
                                         "dense6", "dense7", "bottleneck", "viewdirs", "rgb", "alpha"};
    for (const char* n : stem) {
      host->names.push_back(std::string(n) + "_kernel");
      host->names.push_back(std::string(n) + "_bias");
    }
    Network net(std::make_shared<Context>(*host, *host, device), NERF_NET_FINE);
    net.host_ = std::move(host);
    net.device_ = device;
    return net;
  }
  const std::shared_ptr<HostNet>& host() const { return host_; }
  int device() const { return device_; }
  // Arithmetic of the 256-wide layers (nerf_precision).  Unset, forward_batch runs NERF_PREC_FP16X3 - the tensor-core
  // mode that passes the reference's own unit test (1e-2 abs on sigma up to 213, lib.rs:732-742) with margin - and
  // render_image runs NERF_PREC_MIXED, where the image bound is the criterion.
  void set_precision(int p) { precision_ = p; }
  int precision() const { return precision_ >= 0 ? precision_ : NERF_PREC_FP16X3; }
  int render_precision() const { return precision_ >= 0 ? precision_ : NERF_PREC_MIXED; }
  const std::shared_ptr<Context>& context() const { return ctx_; }
  int net_id() const { return net_; }

  // network.rs:197-237: points is 3 x B (a column per sample), view_dirs has B entries; returns (rgb, sigma).
  std::pair<std::vector<Vec3>, std::vector<float>> forward_batch(const Matrix& points,
                                                                 const std::vector<Vec3>& view_dirs) const {
    const int64_t batch = points.cols();
    if (batch == 0) return {};  // network.rs:199-201
    if (points.rows() != 3 || (int64_t)view_dirs.size() != batch) throw Panic("forward_batch: shape mismatch");
    std::vector<Vec3> rgb((size_t)batch);
    std::vector<float> sigma((size_t)batch);
    check(nerf_forward_batch(ctx_->get(), net_, points.data().data(), reinterpret_cast<const float*>(view_dirs.data()),
                             batch, precision(), reinterpret_cast<float*>(rgb.data()), sigma.data()),
          "forward_batch");
    return {std::move(rgb), std::move(sigma)};
  }

 private:
  std::shared_ptr<Context> ctx_;
  int net_;
  int precision_ = -1;  // unset: see set_precision
  std::shared_ptr<HostNet> host_;  // only for networks built from caller-held layers
  int device_ = 0;
};

// lib.rs:108-174.  `dir` is <root>/coarse or <root>/fine, as the reference calls it (lib.rs:651-652); both calls share
// one context per root.
inline Network load_network_from_dir(const std::string& dir, int device = 0) {
  std::string d = dir;
  while (!d.empty() && d.back() == '/') d.pop_back();
  const size_t cut = d.find_last_of('/');
  const std::string leaf = cut == std::string::npos ? d : d.substr(cut + 1);
  const std::string root = cut == std::string::npos ? "." : d.substr(0, cut);
  int net;
  if (leaf == "coarse") net = NERF_NET_COARSE;
  else if (leaf == "fine") net = NERF_NET_FINE;
  else throw Panic("load_network_from_dir: expected <root>/coarse or <root>/fine, got " + dir);
  static std::mutex mu;
  static std::map<std::string, std::weak_ptr<Context>> cache;
  std::lock_guard<std::mutex> lk(mu);
  const std::string key = root + "#" + std::to_string(device);
  std::shared_ptr<Context> ctx = cache[key].lock();
  if (!ctx) {
    ctx = std::make_shared<Context>(root, device);
    cache[key] = ctx;
  }
  return Network(ctx, net);
}

// ---------------------------------------------------------------------------------------------- lib.rs
struct Camera {  // lib.rs:197-211
  nerf_camera c{};
  size_t nx = 0, ny = 0, samples_per_ray = 0;
};
// lib.rs:614-645: the JSON's near, far, camera_origin, camera_forward, camera_up, hwf.
inline Camera camera_from_samples(const std::string& json_path, size_t nx, size_t ny, size_t samples_per_ray) {
  Camera cam;
  check(nerf_camera_from_json(json_path.c_str(), &cam.c), "camera_from_samples");
  cam.nx = nx;
  cam.ny = ny;
  cam.samples_per_ray = samples_per_ray;
  return cam;
}

// lib.rs:474-565.  The reference jitters with an OS-seeded thread_rng; `sampling` picks the reproducible substitute
// (default: Philox seed 0; NERF_SAMPLING_DET for the parity tests).
inline std::vector<Vec3> render_image(const Network& coarse, const Network& fine, const Camera& cam,
                                      size_t fine_samples_per_ray, const nerf_sampling* sampling = nullptr) {
  std::shared_ptr<Context> ctx = coarse.context();
  if (coarse.host() && fine.host()) {
    // two networks built with Network::create: the pair's own context, built once
    static std::mutex mu;
    static std::map<std::pair<const HostNet*, const HostNet*>, std::shared_ptr<Context>> pairs;
    std::lock_guard<std::mutex> lk(mu);
    std::shared_ptr<Context>& slot = pairs[{coarse.host().get(), fine.host().get()}];
    if (!slot) slot = std::make_shared<Context>(*coarse.host(), *fine.host(), coarse.device());
    ctx = slot;
  } else if (coarse.context() != fine.context() || coarse.net_id() != NERF_NET_COARSE || fine.net_id() != NERF_NET_FINE) {
    throw Panic("render_image: coarse and fine must come from the same weight root");
  }
  nerf_sampling smp{};
  smp.mode = NERF_SAMPLING_PHILOX;
  if (sampling) smp = *sampling;
  std::vector<Vec3> image(cam.nx * cam.ny);
  check(nerf_render(ctx->get(), &cam.c, (int)cam.nx, (int)cam.ny, (int)cam.samples_per_ray,
                    (int)fine_samples_per_ray, fine.render_precision(), &smp, 0, (int64_t)(cam.nx * cam.ny),
                    reinterpret_cast<float*>(image.data())),
        "render_image");  // W % 8, H % 8, samples == 0 -> the asserts of lib.rs:483-501
  return image;
}

// lib.rs:567-580.
inline void save_ppm(const std::string& path, size_t width, size_t height, const std::vector<Vec3>& pixels) {
  if (pixels.size() != width * height) throw Panic("pixels.len() != width * height");  // lib.rs:569
  check(nerf_save_ppm(path.c_str(), (int)width, (int)height, reinterpret_cast<const float*>(pixels.data())), "save_ppm");
}

// lib.rs:603-612 (native target).
inline std::pair<size_t, size_t> default_sample_counts() { return {64, 128}; }

// lib.rs:647-677: both networks and the camera from `weights_root` (the reference bakes in CARGO_MANIFEST_DIR/lego_rust),
// 256 x 256, output.ppm in the working directory, the same progress lines.
inline void render_cli_image(const std::string& weights_root = "lego_rust") {
  const Network coarse = load_network_from_dir(weights_root + "/coarse");
  const Network fine = load_network_from_dir(weights_root + "/fine");
  const auto counts = default_sample_counts();
  const size_t image_width = 256, image_height = 256;
  std::printf("Rendering with %zu coarse samples and %zu fine samples per ray\n", counts.first, counts.second);
  const Camera camera = camera_from_samples(weights_root + "/tf_reference_samples.json", image_width, image_height, counts.first);
  std::printf("Starting image rendering...\n");
  const auto start = std::chrono::steady_clock::now();
  const std::vector<Vec3> pixels = render_image(coarse, fine, camera, counts.second);
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - start).count();
  std::printf("Rendering completed in %.2f seconds\n", secs);
  save_ppm("output.ppm", image_width, image_height, pixels);
}

}  // namespace nerf_rs
#endif  // NERF_RS_HPP
