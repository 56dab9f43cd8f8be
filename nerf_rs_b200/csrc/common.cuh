// common.cuh — shared device/host declarations of libnerf_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/nerf_b200.h"

namespace nerf {

// ------------------------------------------------------------------ network dimensions (shapes.txt)
constexpr int kPosOctaves = 10;          // network.rs:204
constexpr int kDirOctaves = 4;           // network.rs:219
constexpr int kPosEnc = 3 + 6 * kPosOctaves;  // 63
constexpr int kDirEnc = 3 + 6 * kDirOctaves;  // 27
constexpr int kHidden = 256;
constexpr int kViewHidden = 128;
constexpr int kSkipIn = kPosEnc + kHidden;    // 319, [h0; h4] (network.rs:210)
constexpr int kViewIn = kHidden + kDirEnc;    // 283, [bottleneck; PE4(dir)] (network.rs:220)

// f32 weights of one network on the device, in the file layout ([in][out] row-major).
struct NetF32 {
  const float* w[8];   // dense0..7
  const float* b[8];
  const float* w_bott; // [256][256]
  const float* b_bott;
  const float* w_view; // [283][128]
  const float* b_view;
  const float* w_rgb;  // [128][3]
  const float* b_rgb;
  const float* w_alpha;  // [256][1]
  const float* b_alpha;
};

// Where the samples of an MLP launch come from.
//   mode 0: forward_batch — points [3][B] feature-major + dirs [B][3]          (network.rs:197)
//   mode 1: render        — sample g belongs to ray g / S; p = origin + dir_hat * t   (lib.rs:396,436)
//   mode 1 + index        — refinement pass over a compacted list: launch slot i evaluates sample g = index[i] and
//                           stores its outputs at g; the number of slots is read from *count on the device
//                           (the list is built by coarse_select_kernel in the same stream), B is ignored
struct SampleSrc {
  int mode;
  int S;                // samples per ray (mode 1)
  int64_t B;            // number of samples
  const float* points;  // mode 0
  const float* dirs;    // mode 0: [B][3]; mode 1: [n_rays][3]
  const float* t;       // mode 1: [n_rays][S]
  float origin[3];
  const int32_t* index;   // mode 1 only, nullable: sample ids of the launch slots
  const uint32_t* count;  // with index: device-resident number of slots
  // Staged coarse pass (mlp_tc_kernel's raw-sigma instantiation only): the launch covers samples
  // [sub_s0, sub_s0 + sub_n) of every ray (sub_n = 0: the whole ray); with `index` the rays are the listed ones
  // (`index` holds ray ids, `count` their number) and slot i is sample sub_s0 + i % sub_n of ray index[i / sub_n].
  int sub_s0, sub_n;
};

struct SampleDst {
  float* sigma;  // [B]
  float* rgb;    // [B][3] or nullptr: skip bottleneck/viewdirs/rgb (the reference discards coarse rgb, lib.rs:404)
  // sigma = max(alpha(h8), sigma_floor): 0 is the reference's ReLU (network.rs:216, fmaxf also maps NaN to 0);
  // -inf stores the pre-activation itself (first pass of the two-pass coarse densities)
  float sigma_floor;
  // 1: samples whose density is 0 get rgb = 0 instead of the network's colour (their compositing weight is exactly 0,
  // lib.rs:176-195); lets the fused kernel skip the colour branch where a whole tile is empty.  Render launches only.
  int skip_empty_rgb;
  // Two-phase fine pass (mlp_tc.cu): the launch stops after the density head; a sample with density gets
  // its dense7 activations (256 16-bit values, the colour branch's input) appended to `h8` and {ray direction, sample
  // id as bits} to `h8_meta`, their number accumulates in *h8_count (zeroed by the caller); a sample without density
  // gets rgb = 0.  The colour branch then runs over those rows (launch_mlp_colour).  Null: one fused launch.
  uint16_t* h8;
  float4* h8_meta;
  uint32_t* h8_count;
};

// Camera constants hoisted out of get_ray_dir (lib.rs:216-226): they do not depend on the pixel, so
// evaluating them once on the host (glibc tanf, IEEE sqrt/div) is bit-identical to the reference.
struct RayBasis {
  float f[3], r[3], u[3];
  float sx, sy;
  float origin[3];
  float near_, far_;
  int nx, ny;
};

// Which pixel a launch's ray r is.  A contiguous range (parts <= 1): pixel = first + r.  A striped view - the share of
// part `part` when the frame is dealt out in stripes of `stripe` pixel indices to `parts` renderers in turn, which
// balances the data-dependent work (empty rows are cheaper than rows through the object) across GPUs: view index
// q = first + r lies in the view's stripe q / stripe, which is the frame's stripe (q / stripe) * parts + part.
struct RayMap {
  int64_t first;
  int64_t stripe;
  int32_t parts, part;
  __host__ __device__ int64_t pixel(int64_t r) const {
    const int64_t q = first + r;
    if (parts <= 1) return q;
    const int64_t k = q / stripe;
    return (k * parts + part) * stripe + (q - k * stripe);
  }
};
// Number of pixels of a striped view of an n-pixel frame.
inline int64_t view_size(int64_t n, int64_t stripe, int parts, int part) {
  const int64_t round = stripe * parts, full = n / round, rest = n - full * round;
  const int64_t tail = rest - (int64_t)part * stripe;
  return full * stripe + (tail <= 0 ? 0 : (tail < stripe ? tail : stripe));
}

// ------------------------------------------------------------------ Philox4x32-10 (shared with oracle)
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
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
// k-th uniform of (seed, pixel, stream): block k/4, element k%4, float = (x >> 8) * 2^-24 in [0,1).
__host__ __device__ inline float philox_uniform(uint64_t seed, uint32_t pixel, uint32_t stream, int k) {
  uint32_t c[4] = {pixel, (uint32_t)(k >> 2), stream, 0u};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  return (float)(c[k & 3] >> 8) * (1.0f / 16777216.0f);
}

// ------------------------------------------------------------------ kernel launchers (one per .cu)
// mlp_fp32.cu
cudaError_t launch_mlp_fp32(const NetF32& net, const SampleSrc& src, const SampleDst& dst, cudaStream_t s);
// sampling.cu
cudaError_t init_sampling();
cudaError_t launch_raygen_stratified(const RayBasis& rb, int nc, int mode, uint64_t seed, const float* d_jitter,
                                     RayMap map, int64_t n_rays, float* d_dirs, float* d_tc, cudaStream_t s);
cudaError_t launch_ray_dirs(const RayBasis& rb, int64_t ray_begin, int64_t n_rays, float* d_dirs, cudaStream_t s);
// weights come from d_sigc (compute_weights) unless d_w_in is given (stage entry point).
cudaError_t launch_importance(const float* d_tc, const float* d_sigc, const float* d_w_in, int nc, int nf, float far_,
                              int mode,
                              uint64_t seed, const float* d_u, RayMap map, int64_t n_rays, float* d_tf,
                              float* d_wc_or_null, cudaStream_t s, const int32_t* d_ray_ids = nullptr);
// d_ray_ids (nullable, skip-empty-rays mode): launch slot i works on ray d_ray_ids[i] - the coarse inputs, the
// uniforms, the Philox key and d_wc are indexed by ray, the merged samples d_tf by slot.
// Two-pass coarse densities: d_sig holds the single-pass fp16 pre-activations on entry and
// relu(pre-activation) on return; the ids of the samples whose density can reach the coarse weights are appended to
// d_list, their number to *d_count (zeroed by the caller).
cudaError_t launch_coarse_select(float* d_sig, const float* d_t, int64_t n_rays, int nc, float far_, float margin,
                                 float tau_cut, int32_t* d_list, uint32_t* d_count, cudaStream_t s);
// Staged coarse pass: appends to d_list the rays (all n_rays of them, or those of d_in_list) whose optical depth over
// samples [0, s_end) - from the pre-activations in d_sig - is still below tau_cut; the number goes to *d_count.
cudaError_t launch_coarse_alive(const float* d_sig, const float* d_t, int64_t n_rays, int nc, int s_end, float far_,
                                float tau_cut, const int32_t* d_in_list, const uint32_t* d_in_count, int32_t* d_list,
                                uint32_t* d_count, cudaStream_t s);
cudaError_t launch_weights(const float* d_sigma, const float* d_t, int64_t n_rays, int n, float far_, float* d_w,
                           cudaStream_t s);
cudaError_t launch_quantize(const float* d_rgb, int64_t n_pixels, int channels, uint8_t* d_out, cudaStream_t s);
// ray r's outputs go to element omap.pixel(r) of d_out / d_depth / d_acc
cudaError_t launch_composite(const float* d_rgb, const float* d_sigma, const float* d_t, int64_t n_rays, int n,
                             float far_, RayMap omap, float* d_out, float* d_depth /*nullable*/, float* d_acc /*nullable*/,
                             cudaStream_t s, const int32_t* d_ray_ids = nullptr);
// d_ray_ids (nullable): the inputs are indexed by launch slot i and the outputs go to omap.pixel(d_ray_ids[i]).
// Skip-empty-rays mode (opt-in, not the reference's image): rays all of whose coarse weights are exactly
// zero get the background (rgb 1, depth 0, opacity 0) at omap.pixel(ray) straight away; the others are appended to
// d_list (their number to *d_count, zeroed by the caller) and their directions copied to d_dirs_out in list order.
cudaError_t launch_ray_pick(const float* d_sigc, const float* d_tc, const float* d_dirs, int64_t n_rays, int nc,
                            float far_, RayMap omap, int32_t* d_list, uint32_t* d_count, float* d_dirs_out, float* d_out,
                            float* d_depth /*nullable*/, float* d_acc /*nullable*/, cudaStream_t s);

}  // namespace nerf
