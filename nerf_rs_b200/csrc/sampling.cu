// sampling.cu — ray generation, stratified sampling, hierarchical importance resampling and alpha
// compositing as sm_100a kernels.  These are HBM-bound stages (K1 268 B/ray, K3 1 280 B/ray,
// K4 3 856 B/ray); one warp owns one ray, loads are lane-strided (coalesced), transmittance is a warp-level
// exclusive product scan.  Compiled with -fmad=false: every mul/add rounds separately, as in the
// reference's f32 Rust arithmetic (rustc never contracts).
//
// Reference functions restated here (file:line into elisabeth96/nerf-rs):
//   get_ray_dir + normalize   src/lib.rs:214-230, :371      stratified_samples  src/lib.rs:233-248
//   compute_weights           src/lib.rs:250-283            midpoints           src/lib.rs:285-287
//   sample_importance         src/lib.rs:289-351            merge + sort        src/lib.rs:416-420
//   integrate_ray             src/lib.rs:176-195
#include <cstdlib>

#include "common.cuh"

namespace nerf {

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kWarpsPerCta = 8;

__device__ __forceinline__ void ray_dir_normalized(const RayBasis& rb, int64_t pixel, float out[3]) {
  const int i = (int)(pixel / rb.nx), j = (int)(pixel % rb.nx);  // pixel_index = i*nx + j (lib.rs:371)
  const float x = (((float)j + 0.5f) / (float)rb.nx) * 2.0f - 1.0f;  // lib.rs:221
  const float y = 1.0f - (((float)i + 0.5f) / (float)rb.ny) * 2.0f;  // lib.rs:222
  const float a = x * rb.sx, b = y * rb.sy;
  const float dx = (rb.r[0] * a + rb.u[0] * b) + rb.f[0];  // r*(x*sx) + u*(y*sy) + f (lib.rs:229)
  const float dy = (rb.r[1] * a + rb.u[1] * b) + rb.f[1];
  const float dz = (rb.r[2] * a + rb.u[2] * b) + rb.f[2];
  const float len = sqrtf(dx * dx + dy * dy + dz * dz);  // vec3.rs:27-34: three divisions, not an rsqrt
  out[0] = dx / len;
  out[1] = dy / len;
  out[2] = dz / len;
}

// K1: one thread per (ray, coarse sample); thread s==0 of a ray also writes the direction.
__global__ void __launch_bounds__(256) raygen_stratified_kernel(RayBasis rb, int nc, int mode, uint64_t seed,
                                                               const float* __restrict__ jitter,
                                                               RayMap map, int64_t n_rays,
                                                               float* __restrict__ dirs, float* __restrict__ tc) {
  const int64_t total = n_rays * nc;
  const float interval = (rb.far_ - rb.near_) / (float)nc;  // lib.rs:239
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    const int64_t ray = e / nc;
    const int s = (int)(e - ray * nc);
    const int64_t pixel = map.pixel(ray);
    float jit;
    if (mode == NERF_SAMPLING_DET) jit = 0.5f;
    else if (mode == NERF_SAMPLING_PHILOX) jit = philox_uniform(seed, (uint32_t)pixel, 0u, s);
    else jit = jitter[e];
    const float lower = rb.near_ + (float)s * interval;  // lib.rs:241
    const float upper = lower + interval;                // lib.rs:242
    tc[e] = lower + (upper - lower) * jit;               // lib.rs:244
    if (s == 0) {
      float d[3];
      ray_dir_normalized(rb, pixel, d);
      dirs[ray * 3 + 0] = d[0];
      dirs[ray * 3 + 1] = d[1];
      dirs[ray * 3 + 2] = d[2];
    }
  }
}

// K1 for sample counts that are a multiple of four: one thread per four consecutive samples of a ray - one Philox block
// (its four uniforms are exactly those samples'), one 128-bit store, one division by nc / 4 instead of four by nc (ncu,
// round 2: the per-sample form was issue-bound at six times its HBM time).  Same expressions per sample as above.
__global__ void __launch_bounds__(256) raygen_stratified_quad_kernel(RayBasis rb, int nc, int mode, uint64_t seed,
                                                                    const float* __restrict__ jitter, RayMap map,
                                                                    int64_t n_rays, float* __restrict__ dirs,
                                                                    float* __restrict__ tc) {
  const int qpr = nc >> 2;  // quads per ray
  const int64_t total = n_rays * qpr;
  const float interval = (rb.far_ - rb.near_) / (float)nc;  // lib.rs:239
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
    int64_t ray;
    if (total <= 0x7fffffff) ray = (int64_t)((uint32_t)e / (uint32_t)qpr);
    else ray = e / qpr;
    const int q = (int)(e - ray * qpr);
    const int64_t pixel = map.pixel(ray);
    float jit[4] = {0.5f, 0.5f, 0.5f, 0.5f};
    if (mode == NERF_SAMPLING_PHILOX) {
      uint32_t c[4] = {(uint32_t)pixel, (uint32_t)q, 0u, 0u};  // philox_uniform(seed, pixel, 0, 4 q + j)
      philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
      for (int j = 0; j < 4; ++j) jit[j] = (float)(c[j] >> 8) * (1.0f / 16777216.0f);
    } else if (mode != NERF_SAMPLING_DET) {
      const float4 v = *reinterpret_cast<const float4*>(jitter + e * 4);  // the launcher checked the alignment
      jit[0] = v.x; jit[1] = v.y; jit[2] = v.z; jit[3] = v.w;
    }
    float t[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float lower = rb.near_ + (float)(4 * q + j) * interval;  // lib.rs:241
      const float upper = lower + interval;                          // lib.rs:242
      t[j] = lower + (upper - lower) * jit[j];                       // lib.rs:244
    }
    *reinterpret_cast<float4*>(tc + e * 4) = make_float4(t[0], t[1], t[2], t[3]);
    if (q == 0) {
      float d[3];
      ray_dir_normalized(rb, pixel, d);
      dirs[ray * 3 + 0] = d[0];
      dirs[ray * 3 + 1] = d[1];
      dirs[ray * 3 + 2] = d[2];
    }
  }
}

__global__ void __launch_bounds__(256) ray_dirs_kernel(RayBasis rb, int64_t ray_begin, int64_t n_rays,
                                                      float* __restrict__ dirs) {
  for (int64_t ray = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; ray < n_rays;
       ray += (int64_t)gridDim.x * blockDim.x) {
    float d[3];
    ray_dir_normalized(rb, ray_begin + ray, d);
    dirs[ray * 3 + 0] = d[0];
    dirs[ray * 3 + 1] = d[1];
    dirs[ray * 3 + 2] = d[2];
  }
}

// compute_weights (lib.rs:250-283) for 32 consecutive samples [base, base+32) of one ray, one per lane.
// `carry` is the transmittance before sample `base`; returns this lane's weight and updates carry.
// T before sample i is an exclusive product scan of (1 - alpha).  The reference zeroes every weight after
// the first sample whose running T drops below 1e-4; T is non-increasing, so that is exactly "T before
// sample i < 1e-4".
// The scan half: this lane's weight from its alpha (0 for lanes past the end of the ray).
__device__ __forceinline__ float warp_weight_from_alpha(float alpha, int lane, float& carry);
// alpha of one sample from values already in registers (the same three operations as below).
__device__ __forceinline__ float sample_alpha(float sig_i, float t_i, float t_next, bool has_next, float far_) {
  float delta = has_next ? (t_next - t_i) : (far_ - t_i);  // lib.rs:262-266
  if (delta < 0.0f) delta = 0.0f;                          // lib.rs:267-269
  return 1.0f - expf(-sig_i * delta);                      // lib.rs:271
}
__device__ __forceinline__ float warp_weights_round(const float* sig, const float* t, int n, float far_, int base,
                                                    int lane, float& carry) {
  const int i = base + lane;
  float alpha = 0.0f;
  if (i < n) alpha = sample_alpha(sig[i], t[i], (i + 1 < n) ? t[i + 1] : 0.0f, i + 1 < n, far_);
  return warp_weight_from_alpha(alpha, lane, carry);
}
__device__ __forceinline__ float warp_weight_from_alpha(float alpha, int lane, float& carry) {
  float p = 1.0f - alpha;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const float q = __shfl_up_sync(kFull, p, d);
    if (lane >= d) p *= q;
  }
  float excl = __shfl_up_sync(kFull, p, 1);
  if (lane == 0) excl = 1.0f;
  const float t_before = carry * excl;
  carry = carry * __shfl_sync(kFull, p, 31);
  return (t_before < 1e-4f) ? 0.0f : t_before * alpha;  // lib.rs:272-279
}

// ---- sorting helpers of K3.  sort_by(partial_cmp) (lib.rs:419) is a stable sort of plain floats: equal floats
// are indistinguishable, so ANY ascending arrangement of the multiset is bit-identical to the reference's.
__host__ __device__ inline int next_pow2(int n) {
  int p = 1;
  while (p < n) p <<= 1;
  return p;
}
// Warp-cooperative bitonic sort of a[0..p) in shared memory, p a power of two (padding = +inf).
__device__ __forceinline__ void warp_bitonic_sort(float* a, int p, int lane) {
  for (int k = 2; k <= p; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = lane; i < (p >> 1); i += 32) {
        const int lo = ((i & ~(j - 1)) << 1) | (i & (j - 1));  // i-th index with bit j clear
        const int hi = lo | j;
        const bool up = (lo & k) == 0;
        const float x = a[lo], y = a[hi];
        if ((x > y) == up) {
          a[lo] = y;
          a[hi] = x;
        }
      }
      __syncwarp();
    }
  }
}
// true iff a[0..n) is non-decreasing (warp-uniform result)
__device__ __forceinline__ bool warp_is_sorted(const float* a, int n, int lane) {
  bool ok = true;
  for (int i = lane; i + 1 < n; i += 32) ok = ok && !(a[i + 1] < a[i]);
  return __all_sync(kFull, ok);
}
// number of elements of the ascending array a[0..n) that are < v (strict) or <= v
__device__ __forceinline__ int count_less(const float* a, int n, float v, bool or_equal) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    const float x = a[mid];
    if (or_equal ? (x <= v) : (x < v)) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// K3: coarse weights -> pdf -> cdf -> inverse-CDF samples -> merge with coarse t -> sort.  One warp per ray.
// Per-warp shared memory: t[nc] w[nc] pdf[nc] cdf[nc] bins[nc] fine[pow2(nf)] (+ coarse copy [pow2(nc)] only if the
// caller's coarse samples are not ascending).
// Rounding-sensitive steps keep the reference's operation order: the pdf sum and the cdf prefix sum are sequential
// f32 folds (lib.rs:310,319-328); every lane evaluates the same fold from broadcast shared-memory reads, which
// costs no more than one lane doing it and needs no broadcast afterwards.
__global__ void __launch_bounds__(kWarpsPerCta * 32) importance_kernel(
    const float* __restrict__ tc, const float* __restrict__ sigc, const float* __restrict__ w_in, int nc, int nf,
    float far_, int mode, uint64_t seed,
    const float* __restrict__ u_in, RayMap map, int64_t n_rays, float* __restrict__ tf,
    float* __restrict__ wc_out, const int32_t* __restrict__ ray_ids) {
  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nm = nc + nf;
  const int pf = next_pow2(nf > 0 ? nf : 1), pc = next_pow2(nc);
  float* s_t = smem + (size_t)warp * (5 * nc + pf + pc);
  float* s_w = s_t + nc;
  float* s_pdf = s_w + nc;
  float* s_cdf = s_pdf + nc;
  float* s_bins = s_cdf + nc;
  float* s_f = s_bins + nc;   // fine samples, padded to pf with +inf
  float* s_c = s_f + pf;      // sorted copy of the coarse samples (only used when they arrive unsorted)
  const int m = nc - 2;  // pdf_weights = weights[1..n-1] (lib.rs:299)

  for (int64_t slot = blockIdx.x * (int64_t)kWarpsPerCta + warp; slot < n_rays;
       slot += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t ray = ray_ids != nullptr ? (int64_t)ray_ids[slot] : slot;  // inputs by ray, samples out by slot
    const float* tr = tc + ray * nc;
    const float* sr = sigc + ray * nc;
    for (int i = lane; i < nc; i += 32) s_t[i] = tr[i];
    __syncwarp();
    if (w_in != nullptr) {
      for (int i = lane; i < nc; i += 32) s_w[i] = w_in[ray * nc + i];
    } else {
      float carry = 1.0f;
      for (int base = 0; base < nc; base += 32) {
        const float w = warp_weights_round(sr, s_t, nc, far_, base, lane, carry);
        if (base + lane < nc) {
          s_w[base + lane] = w;
          if (wc_out) wc_out[ray * nc + base + lane] = w;
        }
      }
    }
    __syncwarp();
    float* out = tf + slot * nm;
    if (nf > 0) {
      for (int i = lane; i < m; i += 32) s_pdf[i] = fmaxf(s_w[i + 1], 0.0f) + 1e-5f;  // lib.rs:309
      for (int i = lane; i < nc - 1; i += 32) s_bins[i] = 0.5f * (s_t[i] + s_t[i + 1]);  // lib.rs:286
      __syncwarp();
      float sum = 0.0f;
      for (int i = 0; i < m; ++i) sum += s_pdf[i];  // sequential fold, as iter().sum() (lib.rs:310)
      // cdf[i+1] = cdf[i] + pdf[i] / sum, sequentially (lib.rs:315-328); lane i % 32 keeps element i + 1
      for (int i = lane; i < m; i += 32) s_pdf[i] = s_pdf[i] / sum;  // lib.rs:315-317
      __syncwarp();
      float cum = 0.0f;
      if (lane == 0) s_cdf[0] = 0.0f;
      for (int i = 0; i < m; ++i) {
        cum += s_pdf[i];
        if ((i & 31) == lane) s_cdf[i + 1] = cum;
      }
      __syncwarp();
      if (lane == 0) s_cdf[m] = 1.0f;  // lib.rs:327
      __syncwarp();
      const int64_t pixel = map.pixel(ray);
      if (mode == NERF_SAMPLING_PHILOX) {
        // one Philox4x32-10 block yields four consecutive uniforms (philox_uniform: block k / 4, element k % 4): evaluate
        // each block once and park the uniforms in s_f, which the loop below overwrites element by element
        for (int b = lane; 4 * b < nf; b += 32) {
          uint32_t c[4] = {(uint32_t)pixel, (uint32_t)b, 1u, 0u};
          philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (4 * b + j < nf) s_f[4 * b + j] = (float)(c[j] >> 8) * (1.0f / 16777216.0f);
        }
        __syncwarp();
      }
      for (int k = lane; k < pf; k += 32) {
        float v = __int_as_float(0x7f800000);  // +inf padding for the sort
        if (k < nf) {
          float uu;
          if (mode == NERF_SAMPLING_DET) uu = ((float)k + 0.5f) / (float)nf;
          else if (mode == NERF_SAMPLING_PHILOX) uu = s_f[k];  // == philox_uniform(seed, pixel, 1, k)
          else uu = u_in[ray * nf + k];
          // lib.rs:333-339 picks the first j with cdf[j] <= u < cdf[j+1], default m - 1.  By induction the first j with
          // u < cdf[j+1] already satisfies cdf[j] <= u, so idx = min{j : u < cdf[j+1]}; cdf[1..m-1] is a running sum of
          // positive terms (non-decreasing), so that is a binary search; j = m - 1 is both the last candidate and
          // the default.  NaN compares false everywhere and falls through to m - 1, as in the reference.
          int lo = 0, hi = m - 1;
          while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (uu < s_cdf[mid + 1]) hi = mid;
            else lo = mid + 1;
          }
          const int idx = lo;
          const float cl = s_cdf[idx], ch = s_cdf[idx + 1];
          const float denom = fmaxf(ch - cl, 1e-6f);  // lib.rs:343
          const float bl = s_bins[idx], bu = s_bins[idx + 1];
          const float tt = (uu - cl) / denom;
          v = bl + (bu - bl) * tt;  // lib.rs:347
        }
        s_f[k] = v;
      }
      __syncwarp();
      // merged = coarse ++ fine; sort (lib.rs:416-420).  Both parts are normally ascending already (stratified
      // samples; inverse CDF of ascending u); otherwise each is sorted on its own, then merged by rank.
      if (!warp_is_sorted(s_f, nf, lane)) warp_bitonic_sort(s_f, pf, lane);
      const float* cs = s_t;
      if (!warp_is_sorted(s_t, nc, lane)) {
        for (int i = lane; i < pc; i += 32) s_c[i] = i < nc ? s_t[i] : __int_as_float(0x7f800000);
        __syncwarp();
        warp_bitonic_sort(s_c, pc, lane);
        cs = s_c;
      }
      // rank of coarse i = i + #{fine < v}; rank of fine k = k + #{coarse <= v}: a permutation of 0..nm-1
      for (int i = lane; i < nc; i += 32) {
        const float v = cs[i];
        out[i + count_less(s_f, nf, v, false)] = v;
      }
      for (int k = lane; k < nf; k += 32) {
        const float v = s_f[k];
        out[k + count_less(cs, nc, v, true)] = v;
      }
    } else {
      if (!warp_is_sorted(s_t, nc, lane)) {
        for (int i = lane; i < pc; i += 32) s_c[i] = i < nc ? s_t[i] : __int_as_float(0x7f800000);
        __syncwarp();
        warp_bitonic_sort(s_c, pc, lane);
        for (int i = lane; i < nc; i += 32) out[i] = s_c[i];
      } else {
        for (int i = lane; i < nc; i += 32) out[i] = s_t[i];
      }
    }
    __syncwarp();
  }
}

// K3 for compile-time sample counts (the reference's 64 + 128): the same arithmetic in the same order as
// importance_kernel, with a third of the instructions (ncu, round 2: the generic kernel is issue-bound at 2 242
// warp-instructions per ray, DRAM at 6 %).
//   - the two sequential f32 folds (pdf sum, cdf prefix sum; lib.rs:310,319-328) are unrolled over 128-bit shared
//     loads; every lane still evaluates the whole chain, lane 0 stores the cdf with 128-bit stores;
//   - the cdf is kept as s_cs[i] = cdf[i + 1] (cdf[0] = 0 is implicit), so that those stores are aligned;
//   - merge: a fine sample drawn from bin idx lies between the midpoints around t[idx + 1], so the number of coarse
//     samples <= it is idx + 1 or idx + 2; the guess is VERIFIED against the sorted coarse samples and the binary
//     search only runs if it fails.  The fine samples mark their output slots; every slot is then written in order
//     (coalesced), unmarked slots taking the coarse samples in sequence - the same permutation as the rank merge.
//   - anything unsorted (jittered u, caller-supplied arrays) takes the generic path: sort each part, rank merge.
template <int NC, int NF>
__global__ void __launch_bounds__(kWarpsPerCta * 32) importance_fixed_kernel(
    const float* __restrict__ tc, const float* __restrict__ sigc, const float* __restrict__ w_in, float far_, int mode,
    uint64_t seed, const float* __restrict__ u_in, RayMap map, int64_t n_rays, float* __restrict__ tf,
    float* __restrict__ wc_out, const int32_t* __restrict__ ray_ids) {
  static_assert(NC % 32 == 0 && NF % 32 == 0 && (NC & (NC - 1)) == 0 && (NF & (NF - 1)) == 0, "power-of-two sample counts");
  constexpr int M = NC - 2, NM = NC + NF;  // pdf_weights = weights[1..n-1] (lib.rs:299)
  constexpr int kPerWarp = 5 * NC + NF + NM;
  extern __shared__ __align__(16) float smem_fixed[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* s_t = smem_fixed + (size_t)warp * kPerWarp;
  float* s_w = s_t + NC;      // coarse weights; later the sorted copy of unsorted coarse samples
  float* s_pdf = s_w + NC;
  float* s_cs = s_pdf + NC;   // s_cs[i] = cdf[i + 1]
  float* s_bins = s_cs + NC;
  float* s_f = s_bins + NC;
  int* s_mark = reinterpret_cast<int*>(s_f + NF);

  for (int64_t slot = blockIdx.x * (int64_t)kWarpsPerCta + warp; slot < n_rays;
       slot += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t ray = ray_ids != nullptr ? (int64_t)ray_ids[slot] : slot;  // inputs by ray, samples out by slot
    const float* tr = tc + ray * NC;
#pragma unroll
    for (int i = lane; i < NC; i += 32) s_t[i] = tr[i];
    __syncwarp();
    if (w_in != nullptr) {
#pragma unroll
      for (int i = lane; i < NC; i += 32) s_w[i] = w_in[ray * NC + i];
    } else {
      const float* sr = sigc + ray * NC;
      float carry = 1.0f;
#pragma unroll
      for (int base = 0; base < NC; base += 32) {
        const float w = warp_weights_round(sr, s_t, NC, far_, base, lane, carry);
        s_w[base + lane] = w;
        if (wc_out) wc_out[ray * NC + base + lane] = w;
      }
    }
    __syncwarp();
#pragma unroll
    for (int i = lane; i < NC; i += 32) {
      s_pdf[i] = i < M ? fmaxf(s_w[i + 1], 0.0f) + 1e-5f : 0.0f;        // lib.rs:309 (entries >= M are never added)
      if (i < NC - 1) s_bins[i] = 0.5f * (s_t[i] + s_t[i + 1]);          // lib.rs:286
    }
    __syncwarp();
    float sum = 0.0f;  // sequential fold, as iter().sum() (lib.rs:310)
#pragma unroll
    for (int q = 0; q < NC / 4; ++q) {
      const float4 v = reinterpret_cast<const float4*>(s_pdf)[q];
      if (4 * q + 0 < M) sum += v.x;
      if (4 * q + 1 < M) sum += v.y;
      if (4 * q + 2 < M) sum += v.z;
      if (4 * q + 3 < M) sum += v.w;
    }
    __syncwarp();  // every lane has read the unnormalised pdf
#pragma unroll
    for (int i = lane; i < NC; i += 32)
      if (i < M) s_pdf[i] = s_pdf[i] / sum;  // lib.rs:315-317
    __syncwarp();
    float cum = 0.0f;  // cdf[i + 1] = cdf[i] + pdf[i], sequentially (lib.rs:319-328); cdf[M] forced to 1 (lib.rs:327)
#pragma unroll
    for (int q = 0; q < NC / 4; ++q) {
      const float4 v = reinterpret_cast<const float4*>(s_pdf)[q];
      float4 o;
      if (4 * q + 0 < M) cum += v.x;
      o.x = 4 * q + 0 < M - 1 ? cum : 1.0f;
      if (4 * q + 1 < M) cum += v.y;
      o.y = 4 * q + 1 < M - 1 ? cum : 1.0f;
      if (4 * q + 2 < M) cum += v.z;
      o.z = 4 * q + 2 < M - 1 ? cum : 1.0f;
      if (4 * q + 3 < M) cum += v.w;
      o.w = 4 * q + 3 < M - 1 ? cum : 1.0f;
      if (lane == 0) reinterpret_cast<float4*>(s_cs)[q] = o;
    }
    __syncwarp();
    const int64_t pixel = map.pixel(ray);
    if (mode == NERF_SAMPLING_PHILOX) {  // one Philox block yields four consecutive uniforms; parked in s_f
      for (int b = lane; 4 * b < NF; b += 32) {
        uint32_t c[4] = {(uint32_t)pixel, (uint32_t)b, 1u, 0u};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
        for (int j = 0; j < 4; ++j) s_f[4 * b + j] = (float)(c[j] >> 8) * (1.0f / 16777216.0f);
      }
      __syncwarp();
    }
    float vals[NF / 32];
    int hint[NF / 32];
#pragma unroll
    for (int r = 0; r < NF / 32; ++r) {
      const int k = lane + 32 * r;
      float uu;
      if (mode == NERF_SAMPLING_DET) uu = ((float)k + 0.5f) / (float)NF;
      else if (mode == NERF_SAMPLING_PHILOX) uu = s_f[k];
      else uu = u_in[ray * NF + k];
      int lo = 0, hi = M - 1;  // idx = min{j : u < cdf[j + 1]}, default M - 1 (see importance_kernel)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (uu < s_cs[mid]) hi = mid;
        else lo = mid + 1;
      }
      const int idx = lo;
      const float cl = idx > 0 ? s_cs[idx - 1] : 0.0f, ch = s_cs[idx];
      const float denom = fmaxf(ch - cl, 1e-6f);  // lib.rs:343
      const float bl = s_bins[idx], bu = s_bins[idx + 1];
      const float tt = (uu - cl) / denom;
      vals[r] = bl + (bu - bl) * tt;  // lib.rs:347
      hint[r] = idx + 1;
    }
    __syncwarp();  // Philox: every lane has taken its uniforms out of s_f
#pragma unroll
    for (int r = 0; r < NF / 32; ++r) s_f[lane + 32 * r] = vals[r];
    __syncwarp();
    float* out = tf + slot * NM;
    const bool fine_sorted = warp_is_sorted(s_f, NF, lane), coarse_sorted = warp_is_sorted(s_t, NC, lane);
    if (fine_sorted && coarse_sorted) {
#pragma unroll
      for (int j = lane; j < NM; j += 32) s_mark[j] = 0;
      __syncwarp();
#pragma unroll
      for (int r = 0; r < NF / 32; ++r) {
        const float v = vals[r];
        const int h = hint[r];  // 1 <= h <= M < NC - 1
        int cnt;                // number of coarse samples <= v
        if (s_t[h - 1] <= v && !(s_t[h] <= v)) cnt = h;
        else if (s_t[h] <= v && !(s_t[h + 1] <= v)) cnt = h + 1;
        else cnt = count_less(s_t, NC, v, true);
        s_mark[lane + 32 * r + cnt] = lane + 32 * r + 1;  // fine k goes to slot k + #{coarse <= v}
      }
      __syncwarp();
      int base = 0;  // fine samples in the slots before this group of 32
#pragma unroll
      for (int g = 0; g < NM / 32; ++g) {
        const int j = g * 32 + lane;
        const int mk = s_mark[j];
        const unsigned bal = __ballot_sync(kFull, mk != 0);
        const int before = base + __popc(bal & ((1u << lane) - 1u));
        const int ci = min(max(j - before, 0), NC - 1);  // in range whenever the marks are a valid merge
        out[j] = mk != 0 ? s_f[mk - 1] : s_t[ci];
        base += __popc(bal);
      }
    } else {
      if (!fine_sorted) warp_bitonic_sort(s_f, NF, lane);
      const float* cs = s_t;
      if (!coarse_sorted) {
#pragma unroll
        for (int i = lane; i < NC; i += 32) s_w[i] = s_t[i];
        __syncwarp();
        warp_bitonic_sort(s_w, NC, lane);
        cs = s_w;
      }
      for (int i = lane; i < NC; i += 32) {
        const float v = cs[i];
        out[i + count_less(s_f, NF, v, false)] = v;
      }
      for (int k = lane; k < NF; k += 32) {
        const float v = s_f[k];
        out[k + count_less(cs, NC, v, true)] = v;
      }
    }
    __syncwarp();
  }
}

// Two-pass coarse densities, selection step.  `sig` arrives as the single-pass fp16 kernel's density
// PRE-activations alpha(h8) and leaves as relu of them; the samples whose density the split-operand kernel has to
// re-evaluate are appended to `list`.  A sample is skipped only if its exact density provably cannot reach a coarse
// weight (compute_weights, lib.rs:250-283):
//   - pre <= -margin: the exact pre-activation is negative too (margin bounds the single-pass error), sigma = 0 both ways;
//   - optical depth in front of it >= tau_cut: transmittance there is far below the 1e-4 at which the reference zeroes
//     every later weight (lib.rs:276-279), whatever this sample's density is.  The optical depth is accumulated from the
//     fp16 densities; tau_cut leaves e^-(tau_cut - 9.21) of slack for their error, and every sample in front of the
//     cut is itself re-evaluated or exactly zero.
// One warp per ray, one list reservation (atomicAdd) per CTA of 8 rays.
__global__ void __launch_bounds__(kWarpsPerCta * 32) coarse_select_kernel(float* __restrict__ sig,
                                                                          const float* __restrict__ t, int64_t n_rays,
                                                                          int nc, float far_, float margin,
                                                                          float tau_cut, int32_t* __restrict__ list,
                                                                          uint32_t* __restrict__ count) {
  __shared__ uint32_t s_cnt[kWarpsPerCta], s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t ray0 = blockIdx.x * (int64_t)kWarpsPerCta; ray0 < n_rays; ray0 += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t ray = ray0 + warp;
    const bool live = ray < n_rays;
    // pass 1: flags (kept as ballots in registers for up to 8 rounds = 256 samples, recomputed otherwise)
    uint32_t total = 0;
    float carry = 0.0f;
    if (live) {
      for (int base = 0; base < nc; base += 32) {
        const int i = base + lane;
        float pre = -1e30f, od = 0.0f;
        if (i < nc) {
          pre = sig[ray * nc + i];
          const float ti = t[ray * nc + i];
          float delta = (i + 1 < nc) ? (t[ray * nc + i + 1] - ti) : (far_ - ti);
          if (delta < 0.0f) delta = 0.0f;
          od = fmaxf(pre, 0.0f) * delta;
        }
        float incl = od;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const float q = __shfl_up_sync(kFull, incl, d);
          if (lane >= d) incl += q;
        }
        const float before = carry + (incl - od);
        carry += __shfl_sync(kFull, incl, 31);
        // NaN pre-activations (broken weights) are selected: both comparisons are written to be true for NaN
        const bool flag = i < nc && !(pre <= -margin) && !(before >= tau_cut);
        total += __popc(__ballot_sync(kFull, flag));
      }
    }
    if (lane == 0) s_cnt[warp] = total;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t sum = 0;
      for (int w = 0; w < kWarpsPerCta; ++w) {
        const uint32_t c = s_cnt[w];
        s_cnt[w] = sum;
        sum += c;
      }
      s_base = sum ? atomicAdd(count, sum) : 0u;
      if (sum) atomicAdd(count + 1, sum);  // running total over the chunks of a render (nerf_timing::coarse_refined)
    }
    __syncthreads();
    // pass 2: same flags again (the loads hit L1), write the ids and the clamped densities
    if (live) {
      uint32_t pos = s_base + s_cnt[warp];
      carry = 0.0f;
      for (int base = 0; base < nc; base += 32) {
        const int i = base + lane;
        float pre = -1e30f, od = 0.0f;
        if (i < nc) {
          pre = sig[ray * nc + i];
          const float ti = t[ray * nc + i];
          float delta = (i + 1 < nc) ? (t[ray * nc + i + 1] - ti) : (far_ - ti);
          if (delta < 0.0f) delta = 0.0f;
          od = fmaxf(pre, 0.0f) * delta;
        }
        float incl = od;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const float q = __shfl_up_sync(kFull, incl, d);
          if (lane >= d) incl += q;
        }
        const float before = carry + (incl - od);
        carry += __shfl_sync(kFull, incl, 31);
        const bool flag = i < nc && !(pre <= -margin) && !(before >= tau_cut);
        const uint32_t mask = __ballot_sync(kFull, flag);
        if (flag) list[pos + __popc(mask & ((1u << lane) - 1u))] = (int32_t)(ray * nc + i);
        pos += __popc(mask);
        __syncwarp();  // every lane has read its neighbours' pre-activations (t only) before anyone overwrites sig
        if (i < nc) sig[ray * nc + i] = fmaxf(pre, 0.0f);
      }
    }
    __syncthreads();  // s_cnt / s_base are reused by the next group of rays
  }
}

// Staged coarse pass: which rays still have light after their first s_end coarse samples?  One warp per
// candidate ray (all rays, or the survivors of the previous stage), optical depth from the fp16 pre-activations as in
// coarse_select_kernel; survivors are appended to `list` (one reservation per CTA of 8 rays).
__global__ void __launch_bounds__(kWarpsPerCta * 32) coarse_alive_kernel(const float* __restrict__ sig,
                                                                         const float* __restrict__ t, int64_t n_rays,
                                                                         int nc, int s_end, float far_, float tau_cut,
                                                                         const int32_t* __restrict__ in_list,
                                                                         const uint32_t* __restrict__ in_count,
                                                                         int32_t* __restrict__ list,
                                                                         uint32_t* __restrict__ count) {
  __shared__ uint32_t s_alive[kWarpsPerCta], s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t n = in_list != nullptr ? (int64_t)*in_count : n_rays;
  for (int64_t c0 = blockIdx.x * (int64_t)kWarpsPerCta; c0 < n; c0 += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t c = c0 + warp;
    int64_t ray = -1;
    bool alive = false;
    if (c < n) {
      ray = in_list != nullptr ? (int64_t)in_list[c] : c;
      float tau = 0.0f;
      for (int i = lane; i < s_end; i += 32) {
        const float ti = t[ray * nc + i];
        float delta = (i + 1 < nc) ? (t[ray * nc + i + 1] - ti) : (far_ - ti);
        if (delta < 0.0f) delta = 0.0f;
        tau += fmaxf(sig[ray * nc + i], 0.0f) * delta;
      }
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) tau += __shfl_xor_sync(kFull, tau, d);
      alive = !(tau >= tau_cut);  // NaN keeps the ray
    }
    if (lane == 0) s_alive[warp] = alive ? 1u : 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t sum = 0;
      for (int w = 0; w < kWarpsPerCta; ++w) {
        const uint32_t a = s_alive[w];
        s_alive[w] = sum;
        sum += a;
      }
      s_base = sum ? atomicAdd(count, sum) : 0u;
    }
    __syncthreads();
    if (alive && lane == 0) list[s_base + s_alive[warp]] = (int32_t)ray;
    __syncthreads();
  }
}

// Skip-empty-rays mode (opt-in): which rays does the coarse pass give any weight at all?  One warp per
// ray evaluates compute_weights (lib.rs:250-283) exactly as K3 will; a ray whose weights are all exactly 0 is written
// out as background here and never reaches the fine pass, the others are appended to `list` (one reservation per CTA
// of 8 rays) with their directions.  NaN weights keep the ray.
__global__ void __launch_bounds__(kWarpsPerCta * 32) ray_pick_kernel(const float* __restrict__ sig,
                                                                     const float* __restrict__ t,
                                                                     const float* __restrict__ dirs, int64_t n_rays,
                                                                     int nc, float far_, RayMap omap,
                                                                     int32_t* __restrict__ list,
                                                                     uint32_t* __restrict__ count,
                                                                     float* __restrict__ dirs_out,
                                                                     float* __restrict__ out,
                                                                     float* __restrict__ depth_out,
                                                                     float* __restrict__ acc_out) {
  __shared__ uint32_t s_alive[kWarpsPerCta], s_base;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t r0 = blockIdx.x * (int64_t)kWarpsPerCta; r0 < n_rays; r0 += (int64_t)gridDim.x * kWarpsPerCta) {
    const int64_t ray = r0 + warp;
    bool alive = false;
    if (ray < n_rays) {
      float carry = 1.0f;
      bool any = false;
      for (int base = 0; base < nc; base += 32) {
        const float w = warp_weights_round(sig + ray * nc, t + ray * nc, nc, far_, base, lane, carry);
        any = any || !(w == 0.0f);
      }
      alive = __any_sync(kFull, any);
      if (!alive) {
        const int64_t opix = omap.pixel(ray);
        if (lane < 3) out[opix * 3 + lane] = 1.0f;  // integrate_ray with no weight: 0 + (1 - 0) * white (lib.rs:194)
        if (lane == 0 && depth_out != nullptr) depth_out[opix] = 0.0f;
        if (lane == 0 && acc_out != nullptr) acc_out[opix] = 0.0f;
      }
    }
    if (lane == 0) s_alive[warp] = alive ? 1u : 0u;
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t sum = 0;
      for (int w = 0; w < kWarpsPerCta; ++w) {
        const uint32_t a = s_alive[w];
        s_alive[w] = sum;
        sum += a;
      }
      s_base = sum ? atomicAdd(count, sum) : 0u;
    }
    __syncthreads();
    if (alive) {
      const uint32_t pos = s_base + s_alive[warp];
      if (lane == 0) list[pos] = (int32_t)ray;
      if (lane < 3) dirs_out[(int64_t)pos * 3 + lane] = dirs[ray * 3 + lane];
    }
    __syncthreads();
  }
}

// compute_weights alone (stage parity entry point).
__global__ void __launch_bounds__(kWarpsPerCta * 32) weights_kernel(const float* __restrict__ sigma,
                                                                    const float* __restrict__ t, int64_t n_rays,
                                                                    int n, float far_, float* __restrict__ w) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerCta + warp; ray < n_rays; ray += (int64_t)gridDim.x * kWarpsPerCta) {
    float carry = 1.0f;
    for (int base = 0; base < n; base += 32) {
      const float wi = warp_weights_round(sigma + ray * n, t + ray * n, n, far_, base, lane, carry);
      if (base + lane < n) w[ray * n + base + lane] = wi;
    }
  }
}

// K4: integrate_ray (lib.rs:176-195): rgb = sum w_i c_i + (1 - sum w_i) * white.
// kRounds > 0: the ray has at most 32 * kRounds samples and ALL of its loads (densities, positions, the colour block)
// are issued before any arithmetic, so a ray pays the memory latency once instead of once per round of 32 samples (ncu,
// round 2: the loop form was latency-bound, DRAM at 37 %).  kRounds = 0: any n, loads inside the loop.
template <int kRounds>
__global__ void __launch_bounds__(kWarpsPerCta * 32) composite_kernel(const float* __restrict__ rgb,
                                                                      const float* __restrict__ sigma,
                                                                      const float* __restrict__ t, int64_t n_rays,
                                                                      int n, float far_, RayMap omap, bool staged,
                                                                      float* __restrict__ out,
                                                                      float* __restrict__ depth_out,
                                                                      float* __restrict__ acc_out,
                                                                      const int32_t* __restrict__ ray_ids) {
  extern __shared__ float4 s_stage[];  // staged launches: one ray's colours per warp, [n * 3] floats
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // A ray's colours are n * 3 consecutive floats; read lane by lane that is three 4-byte loads at a 12-byte stride.
  // When the block is 16-byte aligned (n * 3 a multiple of 4) the warp copies it with 128-bit loads into shared
  // memory first and picks its samples from there (stride 3 words: no bank conflicts).
  const float* s_rgb = reinterpret_cast<const float*>(s_stage) + (size_t)warp * n * 3;
  for (int64_t ray = blockIdx.x * (int64_t)kWarpsPerCta + warp; ray < n_rays; ray += (int64_t)gridDim.x * kWarpsPerCta) {
    const float* sr = sigma + ray * n;
    const float* tr = t + ray * n;
    const float* cr = rgb + ray * n * 3;
    float r = 0.0f, g = 0.0f, b = 0.0f, acc = 0.0f, depth = 0.0f, carry = 1.0f;
    if constexpr (kRounds > 0) {
      constexpr int kQ = (kRounds * 3 + 3) / 4;  // 128-bit colour loads per lane
      float4 st[kQ];
      float sg[kRounds], tv[kRounds], tn[kRounds];
      const float4* src4 = reinterpret_cast<const float4*>(cr);
      if (staged) {
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
          const int k = lane + 32 * q;
          if (k < n * 3 / 4) st[q] = __ldg(src4 + k);
        }
      }
#pragma unroll
      for (int q = 0; q < kRounds; ++q) {
        const int i = q * 32 + lane;
        sg[q] = i < n ? __ldg(sr + i) : 0.0f;
        tv[q] = i < n ? __ldg(tr + i) : 0.0f;
        tn[q] = i + 1 < n ? __ldg(tr + i + 1) : 0.0f;  // the neighbour lane's line: an L1 hit
      }
      if (staged) {
        __syncwarp();  // the previous ray's reads are done
        float4* dst4 = s_stage + (size_t)warp * (n * 3 / 4);
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
          const int k = lane + 32 * q;
          if (k < n * 3 / 4) dst4[k] = st[q];
        }
        __syncwarp();
        cr = s_rgb;
      }
#pragma unroll
      for (int q = 0; q < kRounds; ++q) {
        const int i = q * 32 + lane;
        float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f;
        if (i < n) {
          c0 = cr[i * 3 + 0];
          c1 = cr[i * 3 + 1];
          c2 = cr[i * 3 + 2];
        }
        const float alpha = i < n ? sample_alpha(sg[q], tv[q], tn[q], i + 1 < n, far_) : 0.0f;
        const float wi = warp_weight_from_alpha(alpha, lane, carry);
        r += c0 * wi;
        g += c1 * wi;
        b += c2 * wi;
        acc += wi;
        depth += wi * tv[q];
      }
    } else {
    if (staged) {
      __syncwarp();  // the previous ray's reads are done
      const float4* src4 = reinterpret_cast<const float4*>(cr);
      float4* dst4 = s_stage + (size_t)warp * (n * 3 / 4);
      for (int k = lane; k < n * 3 / 4; k += 32) dst4[k] = __ldg(src4 + k);
      __syncwarp();
      cr = s_rgb;
    }
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      float c0 = 0.0f, c1 = 0.0f, c2 = 0.0f, ti = 0.0f;
      if (i < n) {
        c0 = cr[i * 3 + 0];
        c1 = cr[i * 3 + 1];
        c2 = cr[i * 3 + 2];
        ti = tr[i];
      }
      const float wi = warp_weights_round(sr, tr, n, far_, base, lane, carry);
      r += c0 * wi;
      g += c1 * wi;
      b += c2 * wi;
      acc += wi;
      depth += wi * ti;
    }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      r += __shfl_xor_sync(kFull, r, d);
      g += __shfl_xor_sync(kFull, g, d);
      b += __shfl_xor_sync(kFull, b, d);
      acc += __shfl_xor_sync(kFull, acc, d);
      depth += __shfl_xor_sync(kFull, depth, d);
    }
    // optional outputs the reference does not have: depth = sum w_i t_i, accumulated opacity = sum w_i
    const int64_t opix = omap.pixel(ray_ids != nullptr ? (int64_t)ray_ids[ray] : ray);  // inputs by slot, pixel by ray
    if (lane == 0 && depth_out != nullptr) depth_out[opix] = depth;
    if (lane == 0 && acc_out != nullptr) acc_out[opix] = acc;
    if (n == 0) { r = g = b = 0.0f; acc = 1.0f; }  // lib.rs:178-180: no samples -> black
    const float bg = 1.0f - acc;                   // lib.rs:194
    const float v = lane == 0 ? r + bg : (lane == 1 ? g + bg : b + bg);
    if (lane < 3) out[opix * 3 + lane] = v;  // one 12-byte coalesced store per ray
  }
}

constexpr size_t kImportanceMaxSmem = 200 * 1024;  // check_render_args (api.cu) refuses sample counts beyond it
size_t importance_smem_bytes(int nc, int nf) {
  return (size_t)kWarpsPerCta * (5 * (size_t)nc + next_pow2(nf > 0 ? nf : 1) + next_pow2(nc)) * sizeof(float);
}

inline int grid_for(int64_t work_items, int per_cta) {
  int64_t g = (work_items + per_cta - 1) / per_cta;
  const int64_t cap = 148 * 16;  // 16 resident CTAs of 256 threads would oversubscribe; loops stride the rest
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

cudaError_t init_sampling() {  // once per device, at context creation
  return cudaFuncSetAttribute(importance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kImportanceMaxSmem);
}

cudaError_t launch_raygen_stratified(const RayBasis& rb, int nc, int mode, uint64_t seed, const float* d_jitter,
                                     RayMap map, int64_t n_rays, float* d_dirs, float* d_tc, cudaStream_t s) {
  if (n_rays == 0) return cudaSuccess;
  const bool aligned = (reinterpret_cast<uintptr_t>(d_tc) & 15u) == 0 &&
                       (mode != NERF_SAMPLING_ARRAYS || (reinterpret_cast<uintptr_t>(d_jitter) & 15u) == 0);
  if (nc % 4 == 0 && aligned)
    raygen_stratified_quad_kernel<<<grid_for(n_rays * (nc / 4), 256), 256, 0, s>>>(rb, nc, mode, seed, d_jitter, map,
                                                                                    n_rays, d_dirs, d_tc);
  else
    raygen_stratified_kernel<<<grid_for(n_rays * nc, 256), 256, 0, s>>>(rb, nc, mode, seed, d_jitter, map, n_rays,
                                                                         d_dirs, d_tc);
  return cudaGetLastError();
}

cudaError_t launch_ray_dirs(const RayBasis& rb, int64_t ray_begin, int64_t n_rays, float* d_dirs, cudaStream_t s) {
  if (n_rays == 0) return cudaSuccess;
  ray_dirs_kernel<<<grid_for(n_rays, 256), 256, 0, s>>>(rb, ray_begin, n_rays, d_dirs);
  return cudaGetLastError();
}

cudaError_t launch_importance(const float* d_tc, const float* d_sigc, const float* d_w_in, int nc, int nf, float far_,
                              int mode, uint64_t seed, const float* d_u, RayMap map, int64_t n_rays, float* d_tf,
                              float* d_wc, cudaStream_t s, const int32_t* d_ray_ids) {
  if (n_rays == 0) return cudaSuccess;
  // the reference's sample counts have a kernel of their own (NERF_B200_GENERIC_IMPORTANCE=1 forces the generic one:
  // the two must agree bit for bit, tests/test_gpu_round2.py)
  const char* force = std::getenv("NERF_B200_GENERIC_IMPORTANCE");
  if (nc == 64 && nf == 128 && !(force != nullptr && force[0] == '1')) {
    constexpr size_t kFixedSmem = (size_t)kWarpsPerCta * (5 * 64 + 128 + 192) * sizeof(float);  // 20 KiB
    importance_fixed_kernel<64, 128><<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, kFixedSmem, s>>>(
        d_tc, d_sigc, d_w_in, far_, mode, seed, d_u, map, n_rays, d_tf, d_wc, d_ray_ids);
    return cudaGetLastError();
  }
  const size_t smem = importance_smem_bytes(nc, nf);
  if (smem > kImportanceMaxSmem) return cudaErrorInvalidValue;
  importance_kernel<<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, smem, s>>>(
      d_tc, d_sigc, d_w_in, nc, nf, far_, mode, seed, d_u, map, n_rays, d_tf, d_wc, d_ray_ids);
  return cudaGetLastError();
}

cudaError_t launch_coarse_select(float* d_sig, const float* d_t, int64_t n_rays, int nc, float far_, float margin,
                                 float tau_cut, int32_t* d_list, uint32_t* d_count, cudaStream_t s) {
  if (n_rays == 0) return cudaSuccess;
  coarse_select_kernel<<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, 0, s>>>(d_sig, d_t, n_rays, nc, far_,
                                                                                    margin, tau_cut, d_list, d_count);
  return cudaGetLastError();
}

cudaError_t launch_coarse_alive(const float* d_sig, const float* d_t, int64_t n_rays, int nc, int s_end, float far_,
                                float tau_cut, const int32_t* d_in_list, const uint32_t* d_in_count, int32_t* d_list,
                                uint32_t* d_count, cudaStream_t s) {
  if (n_rays == 0) return cudaSuccess;
  coarse_alive_kernel<<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, 0, s>>>(
      d_sig, d_t, n_rays, nc, s_end, far_, tau_cut, d_in_list, d_in_count, d_list, d_count);
  return cudaGetLastError();
}

cudaError_t launch_ray_pick(const float* d_sigc, const float* d_tc, const float* d_dirs, int64_t n_rays, int nc,
                            float far_, RayMap omap, int32_t* d_list, uint32_t* d_count, float* d_dirs_out, float* d_out,
                            float* d_depth, float* d_acc, cudaStream_t s) {
  if (n_rays == 0) return cudaSuccess;
  ray_pick_kernel<<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, 0, s>>>(
      d_sigc, d_tc, d_dirs, n_rays, nc, far_, omap, d_list, d_count, d_dirs_out, d_out, d_depth, d_acc);
  return cudaGetLastError();
}

cudaError_t launch_weights(const float* d_sigma, const float* d_t, int64_t n_rays, int n, float far_, float* d_w,
                           cudaStream_t s) {
  if (n_rays == 0 || n == 0) return cudaSuccess;
  weights_kernel<<<grid_for(n_rays, kWarpsPerCta), kWarpsPerCta * 32, 0, s>>>(d_sigma, d_t, n_rays, n, far_, d_w);
  return cudaGetLastError();
}

// save_ppm / pixels_to_rgba quantiser (lib.rs:574-576, 586-589): (clamp(v, 0, 1) * 255 + 0.5) as u8.  Rust's `as u8`
// truncates and saturates and maps NaN to 0; fmaxf(NaN, 0) = 0 gives the same.  Multiply and add round separately.
__device__ __forceinline__ uint32_t quantize_one(float v) {
  const float c = fminf(fmaxf(v, 0.0f), 1.0f);
  return __float2uint_rz(__fadd_rn(__fmul_rn(c, 255.0f), 0.5f));
}
// One thread per pixel: 12 B in, 3 B (PPM body) or 4 B (RGBA, alpha 255) out.
__global__ void quantize_kernel(const float* __restrict__ rgb, int64_t n_pixels, int channels, uint8_t* __restrict__ out) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n_pixels) return;
  const uint32_t r = quantize_one(rgb[p * 3 + 0]), g = quantize_one(rgb[p * 3 + 1]), b = quantize_one(rgb[p * 3 + 2]);
  if (channels == 4) {
    reinterpret_cast<uint32_t*>(out)[p] = r | (g << 8) | (b << 16) | 0xff000000u;
  } else {
    out[p * 3 + 0] = (uint8_t)r;
    out[p * 3 + 1] = (uint8_t)g;
    out[p * 3 + 2] = (uint8_t)b;
  }
}

cudaError_t launch_quantize(const float* d_rgb, int64_t n_pixels, int channels, uint8_t* d_out, cudaStream_t s) {
  if (n_pixels == 0) return cudaSuccess;
  quantize_kernel<<<(unsigned)((n_pixels + 255) / 256), 256, 0, s>>>(d_rgb, n_pixels, channels, d_out);
  return cudaGetLastError();
}

cudaError_t launch_composite(const float* d_rgb, const float* d_sigma, const float* d_t, int64_t n_rays, int n,
                             float far_, RayMap omap, float* d_out, float* d_depth, float* d_acc, cudaStream_t s,
                             const int32_t* d_ray_ids) {
  if (n_rays == 0) return cudaSuccess;
  const size_t stage = (size_t)kWarpsPerCta * n * 3 * sizeof(float);
  const bool staged = n > 0 && (n * 3) % 4 == 0 && stage <= 48 * 1024 && (reinterpret_cast<uintptr_t>(d_rgb) & 15u) == 0;
  const int grid = grid_for(n_rays, kWarpsPerCta), threads = kWarpsPerCta * 32;
  const size_t smem = staged ? stage : 0;
#define NERF_COMPOSITE(R) \
  composite_kernel<R><<<grid, threads, smem, s>>>(d_rgb, d_sigma, d_t, n_rays, n, far_, omap, staged, d_out, d_depth, d_acc, \
                                                  d_ray_ids)
  if (n <= 64) NERF_COMPOSITE(2);
  else if (n <= 128) NERF_COMPOSITE(4);
  else if (n <= 192) NERF_COMPOSITE(6);  // the reference's 64 + 128 samples
  else if (n <= 256) NERF_COMPOSITE(8);
  else NERF_COMPOSITE(0);
#undef NERF_COMPOSITE
  return cudaGetLastError();
}

}  // namespace nerf
