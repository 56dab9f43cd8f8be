// nerf_cli — the reference's command line (src/main.rs:1-3 -> render_cli_image, src/lib.rs:647-677) over the C ABI.
//
// With no arguments it does what `cargo run --release` does: loads lego_rust/{coarse,fine}, takes the camera from
// lego_rust/tf_reference_samples.json, renders 256x256 with 64 coarse + 128 fine samples per ray
// (default_sample_counts, lib.rs:603-612; width/height lib.rs:657-658), prints the same progress lines and writes
// output.ppm (save_ppm, lib.rs:567-580).  The flags only expose what the reference hard-codes.
//
// --gpus N spreads the frame over N devices the way the reference spreads it over cores (rayon par_iter over pixel
// blocks, lib.rs:533-550): one context and one host thread per GPU, each rendering its share of the frame straight
// into the host frame.  The shares are interleaved one-row stripes (row k -> GPU k % N), not contiguous ranges: rows
// through empty space are cheaper than rows through the object, and rayon's work stealing balances the same thing in
// the reference.  --cache FILE starts from the pre-swizzled weight cache (written on first use).
//
// The reference jitters samples with an OS-seeded thread_rng (lib.rs:375,407); the default here is the seeded
// Philox stream (--seed), --det selects the deterministic mid-point sampling used by the parity tests.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/nerf_b200.h"

namespace {

void usage() {
  std::printf(
      "usage: nerf_cli [--weights DIR] [--camera JSON] [--width W] [--height H] [--nc N] [--nf N]\n"
      "                [--precision fp32|bf16|fp16|tf32|fp16x3|mixed] [--det | --seed S] [--device D] [--gpus N]\n"
      "                [--cache FILE] [--out FILE]\n"
      "defaults reproduce the reference CLI: 256x256, 64 coarse + 128 fine samples, output.ppm\n");
}

int fail(const char* what) {
  std::fprintf(stderr, "nerf_cli: %s: %s\n", what, nerf_last_error());
  return 1;
}

}  // namespace

int main(int argc, char** argv) {
  std::string root = "lego_rust", cam_path, out = "output.ppm", cache;
  int gpus = 1;
  if (const char* e = std::getenv("NERF_B200_WEIGHTS")) root = e;
  int W = 256, H = 256, nc = 64, nf = 128, precision = NERF_PREC_MIXED, device = 0;
  nerf_sampling smp{};
  smp.mode = NERF_SAMPLING_PHILOX;
  smp.seed = 0;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&](const char* name) -> const char* {
      if (i + 1 >= argc) {
        std::fprintf(stderr, "nerf_cli: %s needs a value\n", name);
        std::exit(2);
      }
      return argv[++i];
    };
    if (a == "--weights") root = next("--weights");
    else if (a == "--camera") cam_path = next("--camera");
    else if (a == "--width") W = std::atoi(next("--width"));
    else if (a == "--height") H = std::atoi(next("--height"));
    else if (a == "--nc") nc = std::atoi(next("--nc"));
    else if (a == "--nf") nf = std::atoi(next("--nf"));
    else if (a == "--device") device = std::atoi(next("--device"));
    else if (a == "--out") out = next("--out");
    else if (a == "--gpus") gpus = std::atoi(next("--gpus"));
    else if (a == "--cache") cache = next("--cache");
    else if (a == "--det") smp.mode = NERF_SAMPLING_DET;
    else if (a == "--seed") { smp.mode = NERF_SAMPLING_PHILOX; smp.seed = std::strtoull(next("--seed"), nullptr, 10); }
    else if (a == "--precision") {
      const std::string p = next("--precision");
      if (p == "fp32") precision = NERF_PREC_FP32;
      else if (p == "bf16") precision = NERF_PREC_BF16;
      else if (p == "fp16") precision = NERF_PREC_FP16;
      else if (p == "tf32") precision = NERF_PREC_TF32;
      else if (p == "fp16x3") precision = NERF_PREC_FP16X3;
      else if (p == "mixed") precision = NERF_PREC_MIXED;
      else { std::fprintf(stderr, "nerf_cli: unknown precision %s\n", p.c_str()); return 2; }
    } else if (a == "--help" || a == "-h") { usage(); return 0; }
    else { std::fprintf(stderr, "nerf_cli: unknown argument %s\n", a.c_str()); usage(); return 2; }
  }
  if (cam_path.empty()) cam_path = root + "/tf_reference_samples.json";  // lib.rs:653-654

  nerf_camera cam;
  if (nerf_camera_from_json(cam_path.c_str(), &cam) != NERF_OK) return fail("camera");  // lib.rs:665
  if (gpus < 1) {
    std::fprintf(stderr, "nerf_cli: --gpus needs a positive count\n");
    return 2;
  }
  if (!cache.empty()) {
    if (FILE* f = std::fopen(cache.c_str(), "rb")) std::fclose(f);
    else if (nerf_build_weight_cache(root.c_str(), cache.c_str()) != NERF_OK) return fail("build weight cache");
  }
  std::vector<nerf_ctx*> ctx((size_t)gpus, nullptr);
  for (int g = 0; g < gpus; ++g) {  // lib.rs:651-652, once per device
    const int rc = cache.empty() ? nerf_create(root.c_str(), device + g, &ctx[g])
                                 : nerf_create_from_cache(cache.c_str(), device + g, &ctx[g]);
    if (rc != NERF_OK) {
      for (nerf_ctx* c : ctx) nerf_destroy(c);
      return fail("load networks");
    }
  }

  std::printf("Rendering with %d coarse samples and %d fine samples per ray\n", nc, nf);  // lib.rs:660-663
  std::printf("Starting image rendering...\n");                                            // lib.rs:667
  std::vector<float> rgb((size_t)W * H * 3);
  const int64_t n_rays = (int64_t)W * H;
  std::vector<int> status((size_t)gpus, NERF_OK);
  std::vector<std::string> errors((size_t)gpus);
  const auto t0 = std::chrono::steady_clock::now();  // the reference times render_image only (lib.rs:668-670)
  {
    std::vector<std::thread> workers;
    for (int g = 0; g < gpus; ++g) {
      workers.emplace_back([&, g] {
        status[g] = gpus == 1 ? nerf_render(ctx[g], &cam, W, H, nc, nf, precision, &smp, 0, n_rays, rgb.data())
                              : nerf_render_striped(ctx[g], &cam, W, H, nc, nf, precision, &smp, W, gpus, g, rgb.data());
        if (status[g] != NERF_OK) errors[g] = nerf_last_error();  // the message is thread-local
      });
    }
    for (std::thread& t : workers) t.join();
  }
  for (int g = 0; g < gpus; ++g) {
    if (status[g] != NERF_OK) {
      std::fprintf(stderr, "nerf_cli: render (gpu %d): %s\n", device + g, errors[g].c_str());
      for (nerf_ctx* c : ctx) nerf_destroy(c);
      return 1;
    }
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::printf("Rendering complete: %lld/%lld pixels (100.0%%)\n", (long long)W * H, (long long)W * H);  // lib.rs:559-562
  std::printf("Rendering completed in %.2f seconds\n", secs);  // lib.rs:671-675
  nerf_timing tm{};
  for (int g = 0; g < gpus; ++g)
    if (nerf_get_timing(ctx[g], &tm) == NERF_OK)
      std::printf("gpu %d: device time %.3f ms (coarse MLP %.3f, fine MLP %.3f, sampling %.3f, compositing %.3f)\n",
                  device + g, tm.total_ms, tm.mlp_coarse_ms, tm.mlp_fine_ms, tm.sampling_ms, tm.composite_ms);
  for (nerf_ctx* c : ctx) nerf_destroy(c);
  if (nerf_save_ppm(out.c_str(), W, H, rgb.data()) != NERF_OK) return fail("save ppm");  // lib.rs:676
  std::printf("Wrote %s\n", out.c_str());
  return 0;
}
