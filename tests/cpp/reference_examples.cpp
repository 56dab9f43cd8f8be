// reference_examples.cpp — the reference crate's only unit test, `coarse_and_fine_match_reference_examples`
// (src/lib.rs:753-916), written against the C++ mirror of the crate's API (include/nerf_rs.hpp) so that it runs on
// the GPU path.  Like the original it loads both networks with load_network_from_dir, builds for each of the three
// example rays a 3 x 5 Matrix of points `origin + ray_dir * t` (the un-normalised TensorFlow ray direction,
// lib.rs:855) with five copies of the unit view direction (lib.rs:854,859), calls forward_batch on the coarse and the
// fine network and requires every sigma and every rgb channel within 1e-2 (assert_close, lib.rs:732-742).  The
// original hard-codes the expected numbers as literals (lib.rs:759-847); this reads the same numbers from
// lego_rust/tf_reference_samples.json, where the reference keeps them in f64-printed form.
//
//   usage: reference_examples <lego_rust dir> [precision: default | fp32 | fp16x3 | fp16 | bf16 | mixed] [out.ppm]
// With a third argument it also goes through the render half of the API the way render_cli_image does (lib.rs:647-677):
// camera_from_samples at 32 x 24 with 64 coarse samples, render_image with 128 fine samples (deterministic sampling
// so that the caller can compare bytes), save_ppm.
// exit 0 = all assertions hold; 1 = assertion failed (printed); 2 = panic (no GPU, missing files, ...).
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/nerf_rs.hpp"

namespace {

// All numbers of the JSON array (nested or flat) that follows the `occurrence`-th `"key":` in `text`.
std::vector<double> numbers_after(const std::string& text, const std::string& key, int occurrence) {
  size_t pos = 0;
  const std::string pat = "\"" + key + "\"";
  for (int k = 0; k <= occurrence; ++k) {
    pos = text.find(pat, pos);
    if (pos == std::string::npos) throw nerf_rs::Panic("tf_reference_samples.json: missing key " + key);
    pos += pat.size();
  }
  pos = text.find('[', pos);
  std::vector<double> out;
  int depth = 0;
  for (size_t i = pos; i < text.size(); ++i) {
    const char ch = text[i];
    if (ch == '[') ++depth;
    else if (ch == ']') {
      if (--depth == 0) break;
    } else if (ch == '-' || ch == '+' || std::isdigit((unsigned char)ch)) {
      char* end = nullptr;
      out.push_back(std::strtod(text.c_str() + i, &end));
      i = (size_t)(end - text.c_str()) - 1;
    }
  }
  return out;
}

int failures = 0;
void assert_close(float got, double want, const char* what, int example, int sample) {  // lib.rs:732-742
  const double tol = 1e-2;
  if (!(std::fabs((double)got - want) < tol)) {
    std::printf("FAIL example %d sample %d %s: got %.6f, expected %.6f (|diff| %.3e >= %.0e)\n", example, sample, what,
                got, want, std::fabs((double)got - want), tol);
    ++failures;
  }
}

// The tensors of <dir> the way load_network_from_dir reads them (lib.rs:35-42,63-74,108-174), but kept by the caller:
// little-endian f32 files, kernels [in x out] row-major.  Feeds Network::create, the mirror of Network::new.
nerf_rs::Network network_from_layers(const std::string& dir, int precision) {
  auto layer = [&](const std::string& stem, int32_t in, int32_t out) {
    auto read = [&](const std::string& name, size_t count) {
      std::ifstream f(dir + "/" + name + ".bin", std::ios::binary);
      std::vector<float> v(count);
      if (!f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(count * 4))) throw nerf_rs::Panic("read " + name);
      return v;
    };
    return nerf_rs::Layer(nerf_rs::Matrix(in, out, read(stem + "_kernel", (size_t)in * out)), read(stem + "_bias", out));
  };
  std::vector<nerf_rs::Layer> trunk;
  for (int i = 0; i < 8; ++i) trunk.push_back(layer("dense" + std::to_string(i), i == 0 ? 63 : (i == 5 ? 319 : 256), 256));
  nerf_rs::Network n = nerf_rs::Network::create(std::move(trunk), layer("bottleneck", 256, 256), layer("viewdirs", 283, 128),
                                                layer("rgb", 128, 3), layer("alpha", 256, 1));
  if (precision >= 0) n.set_precision(precision);
  return n;
}

int precision_from(const char* s) {
  if (!std::strcmp(s, "fp32")) return NERF_PREC_FP32;
  if (!std::strcmp(s, "bf16")) return NERF_PREC_BF16;
  if (!std::strcmp(s, "fp16")) return NERF_PREC_FP16;
  if (!std::strcmp(s, "fp16x3")) return NERF_PREC_FP16X3;
  if (!std::strcmp(s, "mixed")) return NERF_PREC_MIXED;
  throw nerf_rs::Panic(std::string("unknown precision ") + s);
}

}  // namespace

int main(int argc, char** argv) {
  using namespace nerf_rs;
  try {
    const std::string root = argc > 1 ? argv[1] : "lego_rust";
    const char* prec_name = argc > 2 ? argv[2] : "default";
    Network coarse = load_network_from_dir(root + "/coarse");  // lib.rs:756
    Network fine = load_network_from_dir(root + "/fine");      // lib.rs:757
    if (std::strcmp(prec_name, "default") != 0) {  // "default": exactly what a caller of the crate's API gets
      const int precision = precision_from(prec_name);
      coarse.set_precision(precision);
      fine.set_precision(precision);
    }

    std::ifstream f(root + "/tf_reference_samples.json");
    if (!f) throw Panic("read " + root + "/tf_reference_samples.json");
    std::stringstream ss;
    ss << f.rdbuf();
    const std::string text = ss.str();
    const std::vector<double> o = numbers_after(text, "camera_origin", 0);
    const std::vector<double> z = numbers_after(text, "z_vals", 0);
    if (o.size() != 3 || z.size() != 5) throw Panic("tf_reference_samples.json: camera_origin / z_vals");
    const Vec3 origin((float)o[0], (float)o[1], (float)o[2]);

    for (int ex = 0; ex < 3; ++ex) {
      const std::vector<double> rd = numbers_after(text, "ray_d", ex), vd = numbers_after(text, "viewdir_unit", ex);
      const Vec3 ray_dir((float)rd[0], (float)rd[1], (float)rd[2]), view_dir((float)vd[0], (float)vd[1], (float)vd[2]);
      Matrix points = Matrix::zeros(3, (int32_t)z.size());
      std::vector<Vec3> dirs;
      for (size_t s = 0; s < z.size(); ++s) {
        const Vec3 p = origin + ray_dir * (float)z[s];  // lib.rs:855
        points.set(0, (int32_t)s, p.x);
        points.set(1, (int32_t)s, p.y);
        points.set(2, (int32_t)s, p.z);
        dirs.push_back(view_dir);
      }
      const struct { const Network* net; const char* sigma; const char* rgb; const char* name; } nets[2] = {
          {&coarse, "coarse_sigma", "coarse_rgb", "coarse"}, {&fine, "fine_sigma", "fine_rgb", "fine"}};
      for (const auto& n : nets) {
        const auto out = n.net->forward_batch(points, dirs);  // lib.rs:862,889
        const std::vector<double> want_s = numbers_after(text, n.sigma, ex), want_c = numbers_after(text, n.rgb, ex);
        if (out.second.size() != 5 || want_s.size() != 5 || want_c.size() != 15) throw Panic("fixture shape");
        for (int s = 0; s < 5; ++s) {
          char what[64];
          std::snprintf(what, sizeof what, "%s sigma", n.name);
          assert_close(out.second[s], want_s[s], what, ex, s);
          std::snprintf(what, sizeof what, "%s rgb", n.name);
          assert_close(out.first[s].x, want_c[3 * s + 0], what, ex, s);
          assert_close(out.first[s].y, want_c[3 * s + 1], what, ex, s);
          assert_close(out.first[s].z, want_c[3 * s + 2], what, ex, s);
        }
      }
    }
    // network.rs:199-201: an empty batch gives empty vectors
    const auto empty = coarse.forward_batch(Matrix::zeros(3, 0), {});
    if (!empty.first.empty() || !empty.second.empty()) {
      std::printf("FAIL empty batch\n");
      ++failures;
    }
    if (argc > 3) {
      const Camera camera = camera_from_samples(root + "/tf_reference_samples.json", 32, 24, default_sample_counts().first);
      nerf_sampling det{};
      det.mode = NERF_SAMPLING_DET;
      const std::vector<Vec3> pixels = render_image(coarse, fine, camera, default_sample_counts().second, &det);
      save_ppm(argv[3], 32, 24, pixels);
      // the asserts of lib.rs:483-501 surface as panics
      bool panicked = false;
      try {
        render_image(coarse, fine, camera_from_samples(root + "/tf_reference_samples.json", 30, 24, 64), 128, &det);
      } catch (const Panic&) {
        panicked = true;
      }
      if (!panicked) {
        std::printf("FAIL width not a multiple of 8 must panic\n");
        ++failures;
      }
      // Layer::new / Network::new (network.rs:75,181-195): the same networks from caller-held layers evaluate and
      // render bit-identically to the ones load_network_from_dir built
      const int p = std::strcmp(prec_name, "default") != 0 ? precision_from(prec_name) : -1;
      const Network coarse2 = network_from_layers(root + "/coarse", p), fine2 = network_from_layers(root + "/fine", p);
      Matrix pts = Matrix::zeros(3, 7);
      std::vector<Vec3> dirs;
      for (int32_t s = 0; s < 7; ++s) {
        pts.set(0, s, 0.1f * (float)s - 0.3f);
        pts.set(1, s, 0.05f * (float)s);
        pts.set(2, s, 0.2f - 0.07f * (float)s);
        dirs.push_back(Vec3(0.6f, 0.0f, -0.8f));
      }
      const struct { const Network *a, *b; } same[2] = {{&coarse, &coarse2}, {&fine, &fine2}};
      for (const auto& n : same) {
        const auto x = n.a->forward_batch(pts, dirs), y = n.b->forward_batch(pts, dirs);
        if (std::memcmp(x.first.data(), y.first.data(), 7 * sizeof(Vec3)) != 0 ||
            std::memcmp(x.second.data(), y.second.data(), 7 * sizeof(float)) != 0) {
          std::printf("FAIL Network::new differs from load_network_from_dir in forward_batch\n");
          ++failures;
        }
      }
      const std::vector<Vec3> pixels2 = render_image(coarse2, fine2, camera, default_sample_counts().second, &det);
      if (std::memcmp(pixels.data(), pixels2.data(), pixels.size() * sizeof(Vec3)) != 0) {
        std::printf("FAIL Network::new pair renders a different image\n");
        ++failures;
      }
    }
    std::printf("%s: %d assertion(s) failed\n", failures ? "FAILED" : "ok", failures);
    return failures ? 1 : 0;
  } catch (const nerf_rs::Panic& e) {
    std::fprintf(stderr, "panic: %s\n", e.what());
    return 2;
  }
}
