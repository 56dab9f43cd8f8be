// mlp_tc.h — the fused tcgen05/TMEM MLP kernel (mlp_tc.cu): weight tiles, constants and launcher.
#pragma once
#include "common.cuh"
#include <vector>

#include "host_util.h"

namespace nerf {

// Small f32 tables the epilogue needs.  The struct is a __grid_constant__ kernel parameter (12.5 KiB), so every
// entry is a warp-uniform constant-bank operand: no loads, no shared memory.
struct alignas(16) TcConsts {
  float bias[10][256];   // dense0..7, bottleneck, viewdirs (128 used)
  float w_alpha[256];    // sigma head weights (network.rs:216), evaluated in f32 in the epilogue
  float w_rgb[128 * 3];  // rgb head weights [128][3] (network.rs:223)
  float b_rgb[3];
  float b_alpha;
};

// Pre-swizzled 16-bit operand tiles of one network (one operand format), in consumption order.
// Each chunk is the exact shared-memory image of a [N x 64] K-major SWIZZLE_128B tcgen05 B operand
// (N=256: 32 KiB; viewdirs N=128: 16 KiB), so the kernel streams it with a 1-D bulk TMA copy.
struct TcWeights {
  uint8_t* d_chunks = nullptr;
  size_t bytes = 0;
  int fmt = 0;  // 0 = bf16, 1 = fp16, 2 = fp16 hi/lo split (fp16x3), 3 = tf32
  TcConsts consts;
};

// Names the layout of the pre-swizzled chunk streams (both kernels).  CHANGE IT whenever the chunk order, the N split
// between the CTAs of a pair, the K-block size, the swizzle or the hi/lo scaling changes: weight cache files carry its
// hash in their version word (weight_cache.cu) and files of another layout are refused.
constexpr char kTcLayoutTag[] =
    "r2:kmajor-sw128;kblock64;n256(viewdirs128);order=d0:pe|d1-4:4|d5:pe+4|d6-7:4|bott:4|view:4+pe4;x3=hi,lo*2^11 per kblock;"
    "tf32=kblock32,rna,same order";

// Operand formats: 0 = bf16, 1 = fp16 (mlp_tc.cu), 2 = fp16 hi / lo split, a hi and a lo chunk per K-block (mlp_tc_x3.cu),
// 3 = tf32 in 32-bit containers, K-blocks of 32 (mlp_tc_tf32.cu).
constexpr int kNumTcFormats = 4;
// Host side (no GPU needed): the chunk stream of one network in one format, and the epilogue's f32 tables.
void build_tc_image(const TensorMap& t, int fmt, std::vector<uint8_t>* out);
void build_tc_x3_image(const TensorMap& t, std::vector<uint8_t>* out);
void build_tc_tf32_image(const TensorMap& t, std::vector<uint8_t>* out);
size_t tc_tf32_image_bytes();
void fill_tc_consts(const TensorMap& t, TcConsts* c);
size_t tc_image_bytes(int fmt);  // length of that stream: the kernels' bulk copies read exactly this much
// Device upload of a chunk stream (built just now, or read from a weight cache file).
int upload_tc_weights(const uint8_t* img, size_t bytes, int fmt, const TensorMap& t, TcWeights* out);
int build_tc_weights(const TensorMap& t, int fmt, TcWeights* out);
void free_tc_weights(TcWeights* w);
// Opt the kernels into their dynamic shared memory size (once per device, at context creation).
cudaError_t init_mlp_tc();
cudaError_t init_mlp_tc_x3();
cudaError_t init_mlp_tc_tf32();
cudaError_t launch_mlp_tc(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                          cudaStream_t s);
// Second phase of the two-phase fine pass: the colour branch over the *d_count rows of d_h8 that a
// launch_mlp_tc with SampleDst::h8 stored; d_meta[i] = {ray direction, sample id as bits} of row i, colour to d_rgb.
cudaError_t launch_mlp_colour(const TcWeights& w, const uint16_t* d_h8, const float4* d_meta, const uint32_t* d_count,
                              float* d_rgb, int sm_count, cudaStream_t s);
cudaError_t launch_mlp_tc_x3(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                             cudaStream_t s);
cudaError_t launch_mlp_tc_tf32(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                               cudaStream_t s);

}  // namespace nerf
