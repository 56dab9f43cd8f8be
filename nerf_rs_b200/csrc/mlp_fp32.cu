// mlp_fp32.cu — NERF_PREC_FP32: Network::forward_batch (src/network.rs:197-237) on CUDA cores in f32.
//
// This is the parity mode.  Every output element is accumulated exactly as the reference's
// forward_fallback does (src/network.rs:134-143): acc = bias; acc = acc + W[i][o] * h[i] for i ascending, with
// the multiply and the add rounded separately (-fmad=false), so the GEMMs are bit-identical to the CPU
// reference for identical inputs; only sinf/cosf/expf differ from glibc by <= 2 ulp.
//
// One CTA = 64 samples, 256 threads.  Activations stay in shared memory for all 12 layers
// (feature-major [features][64], like the reference's Matrix); weights stream from L2 in 16-row chunks
// through a double-buffered shared tile.  Thread (ty, tx) owns an 8-output x 8-sample register tile.
#include "common.cuh"

namespace nerf {
namespace {

constexpr int TS = 64;       // samples per CTA
constexpr int NT = 256;      // threads per CTA
constexpr int KC = 16;       // weight rows per chunk
constexpr int kPeRows = 64;  // 63 used
constexpr int kDpeRows = 28; // 27 used
constexpr size_t kSmemFloats = (size_t)(kPeRows + kDpeRows + 2 * kHidden) * TS + 2 * KC * kHidden;

enum Act { ACT_RELU = 0, ACT_NONE = 1 };

// out[o][s] = act(b[o] + sum_k W[k][o] * in[k][s]); input rows are in0[0..rows0) followed by in1[0..rows1)
// (concat_rows order, src/network.rs:240-261).
template <int N>
__device__ __forceinline__ void dense(const float* __restrict__ Wg, const float* __restrict__ bg, const float* in0,
                                      int rows0, const float* in1, int rows1, float* out, int act, float* wt) {
  constexpr int NO = N / 32;  // outputs per thread: 8 (N=256) or 4 (N=128)
  constexpr int LD4 = KC * N / 4 / NT;  // float4 loads per thread per chunk
  const int tid = threadIdx.x, tx = tid & 7, ty = tid >> 3;
  const int K = rows0 + rows1;
  const int nchunks = (K + KC - 1) / KC;

  float acc[NO][8];
#pragma unroll
  for (int a = 0; a < NO; ++a) {
    const int o = (a < 4) ? ty * 4 + a : 128 + ty * 4 + (a - 4);
    const float bv = bg[o];
#pragma unroll
    for (int s = 0; s < 8; ++s) acc[a][s] = bv;  // fill_with_bias (network.rs:149-159)
  }

  float4 pre[LD4];
  auto fetch = [&](int c) {
#pragma unroll
    for (int q = 0; q < LD4; ++q) {
      const int idx = (q * NT + tid) * 4;  // float index inside the chunk
      const int k = c * KC + idx / N;
      pre[q] = (k < K) ? *reinterpret_cast<const float4*>(Wg + (size_t)c * KC * N + idx) : make_float4(0, 0, 0, 0);
    }
  };
  auto stash = [&](int stage) {
#pragma unroll
    for (int q = 0; q < LD4; ++q) *reinterpret_cast<float4*>(wt + stage * KC * N + (q * NT + tid) * 4) = pre[q];
  };

  fetch(0);
  stash(0);
  __syncthreads();
  for (int c = 0; c < nchunks; ++c) {
    if (c + 1 < nchunks) fetch(c + 1);
    const float* w = wt + (c & 1) * KC * N;
    const int rows = min(KC, K - c * KC);
    for (int r = 0; r < rows; ++r) {
      const int k = c * KC + r;
      const float* hrow = (k < rows0) ? in0 + (size_t)k * TS : in1 + (size_t)(k - rows0) * TS;
      const float4 ha = *reinterpret_cast<const float4*>(hrow + tx * 4);
      const float4 hb = *reinterpret_cast<const float4*>(hrow + 32 + tx * 4);
      const float h[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
      float wv[NO];
      {
        const float4 wa = *reinterpret_cast<const float4*>(w + r * N + ty * 4);
        wv[0] = wa.x; wv[1] = wa.y; wv[2] = wa.z; wv[3] = wa.w;
        if constexpr (NO == 8) {
          const float4 wb = *reinterpret_cast<const float4*>(w + r * N + 128 + ty * 4);
          wv[4] = wb.x; wv[5] = wb.y; wv[6] = wb.z; wv[7] = wb.w;
        }
      }
#pragma unroll
      for (int a = 0; a < NO; ++a)
#pragma unroll
        for (int s = 0; s < 8; ++s) acc[a][s] = __fadd_rn(acc[a][s], __fmul_rn(wv[a], h[s]));  // network.rs:139
    }
    if (c + 1 < nchunks) stash((c + 1) & 1);
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < NO; ++a) {
    const int o = (a < 4) ? ty * 4 + a : 128 + ty * 4 + (a - 4);
    float v[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) v[s] = (act == ACT_RELU) ? fmaxf(acc[a][s], 0.0f) : acc[a][s];  // network.rs:164
    *reinterpret_cast<float4*>(out + (size_t)o * TS + tx * 4) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(out + (size_t)o * TS + 32 + tx * 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(NT, 1) mlp_fp32_kernel(NetF32 net, SampleSrc src, SampleDst dst) {
  extern __shared__ __align__(16) float smem[];
  float* pe = smem;                      // [64][TS]   rows 0..62: PE10(p)   (network.rs:263-292)
  float* dpe = pe + kPeRows * TS;        // [28][TS]   rows 0..26: PE4(dir)  (network.rs:294-330)
  float* X = dpe + kDpeRows * TS;        // [256][TS]
  float* Y = X + kHidden * TS;           // [256][TS]
  float* wt = Y + kHidden * TS;          // [2][KC][256]
  const int tid = threadIdx.x;

  for (int64_t tile = blockIdx.x; tile * TS < src.B; tile += gridDim.x) {
    const int64_t g0 = tile * TS;
    if (tid < TS) {
      const int64_t g = g0 + tid;
      float p[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 0.f};
      if (g < src.B) {
        if (src.mode == 0) {
          p[0] = src.points[g];
          p[1] = src.points[src.B + g];
          p[2] = src.points[2 * src.B + g];
          d[0] = src.dirs[g * 3 + 0];
          d[1] = src.dirs[g * 3 + 1];
          d[2] = src.dirs[g * 3 + 2];
        } else {
          const int64_t ray = g / src.S;
          d[0] = src.dirs[ray * 3 + 0];
          d[1] = src.dirs[ray * 3 + 1];
          d[2] = src.dirs[ray * 3 + 2];
          const float t = src.t[g];
          p[0] = src.origin[0] + d[0] * t;  // origin + dir_hat * t, mul then add (lib.rs:396,436)
          p[1] = src.origin[1] + d[1] * t;
          p[2] = src.origin[2] + d[2] * t;
        }
      }
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        pe[a * TS + tid] = p[a];
        dpe[a * TS + tid] = d[a];
      }
    }
    __syncthreads();
    {
      // 30 position (octave, axis) pairs + 12 direction pairs, spread over the 4 thread quarters.
      const int s = tid & (TS - 1), part = tid >> 6;
      for (int q = part; q < 42; q += 4) {
        const bool is_dir = q >= 30;
        const int qq = is_dir ? q - 30 : q;
        const int oct = qq / 3, axis = qq % 3;
        float* enc = is_dir ? dpe : pe;
        const float f = (float)(1 << oct);     // f doubled per octave (network.rs:287): exactly 2^oct
        const float arg = f * enc[axis * TS + s];
        float sv, cv;
        sincosf(arg, &sv, &cv);
        enc[(3 + 6 * oct + axis) * TS + s] = sv;      // rows 3+6k..5+6k: sin
        enc[(3 + 6 * oct + 3 + axis) * TS + s] = cv;  // rows 6+6k..8+6k: cos
      }
    }
    __syncthreads();

    dense<256>(net.w[0], net.b[0], pe, kPosEnc, nullptr, 0, X, ACT_RELU, wt);
    dense<256>(net.w[1], net.b[1], X, kHidden, nullptr, 0, Y, ACT_RELU, wt);
    dense<256>(net.w[2], net.b[2], Y, kHidden, nullptr, 0, X, ACT_RELU, wt);
    dense<256>(net.w[3], net.b[3], X, kHidden, nullptr, 0, Y, ACT_RELU, wt);
    dense<256>(net.w[4], net.b[4], Y, kHidden, nullptr, 0, X, ACT_RELU, wt);
    dense<256>(net.w[5], net.b[5], pe, kPosEnc, X, kHidden, Y, ACT_RELU, wt);  // concat_rows(h_0, h4) (network.rs:210)
    dense<256>(net.w[6], net.b[6], Y, kHidden, nullptr, 0, X, ACT_RELU, wt);
    dense<256>(net.w[7], net.b[7], X, kHidden, nullptr, 0, Y, ACT_RELU, wt);   // h8 in Y

    if (tid < TS) {  // sigma = relu(alpha(h8)) (network.rs:216)
      float acc = net.b_alpha[0];
      for (int k = 0; k < kHidden; ++k) acc = __fadd_rn(acc, __fmul_rn(net.w_alpha[k], Y[k * TS + tid]));
      if (g0 + tid < src.B) dst.sigma[g0 + tid] = fmaxf(acc, 0.0f);
    }
    if (dst.rgb != nullptr) {
      dense<256>(net.w_bott, net.b_bott, Y, kHidden, nullptr, 0, X, ACT_NONE, wt);           // network.rs:218
      dense<128>(net.w_view, net.b_view, X, kHidden, dpe, kDirEnc, Y, ACT_RELU, wt);          // network.rs:219-222
      if (tid < 3 * TS) {  // rgb = sigmoid(rgb(c_hidden)) (network.rs:223)
        const int ch = tid / TS, s = tid % TS;
        float acc = net.b_rgb[ch];
        for (int k = 0; k < kViewHidden; ++k) acc = __fadd_rn(acc, __fmul_rn(net.w_rgb[k * 3 + ch], Y[k * TS + s]));
        const float v = 1.0f / (1.0f + expf(-acc));  // network.rs:165
        if (g0 + s < src.B) dst.rgb[(g0 + s) * 3 + ch] = v;
      }
    }
    __syncthreads();
  }
}

}  // namespace

cudaError_t launch_mlp_fp32(const NetF32& net, const SampleSrc& src, const SampleDst& dst, cudaStream_t s) {
  if (src.B == 0) return cudaSuccess;
  const size_t smem = kSmemFloats * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(mlp_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int64_t tiles = (src.B + TS - 1) / TS;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)(tiles < sms ? tiles : sms);
  mlp_fp32_kernel<<<grid, NT, smem, s>>>(net, src, dst);
  return cudaGetLastError();
}

}  // namespace nerf
