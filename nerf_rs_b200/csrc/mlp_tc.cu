// mlp_tc.cu — Network::forward_batch (src/network.rs:197-237) as ONE fused persistent tcgen05 kernel.
//
// One CTA per SM processes pairs of 128-sample tiles (slot X, slot Y).  Samples sit on the 128 TMEM lanes
// (M = 128); every 256-wide layer is a D[128 x 256] += A[128 x K] * B[K x 256] chain of tcgen05.mma
// (kind::f16, M128 N256 K16) accumulating f32 in TMEM (slot X: columns 0..255, slot Y: 256..511).
//   - A (activations) never leaves the SM: the epilogue reads D with tcgen05.ld, adds the bias, applies ReLU,
//     rounds to fp16/bf16 and writes the next layer's A operand straight into shared memory in the K-major
//     SWIZZLE_128B layout the MMA descriptor expects.
//   - B (weights) is streamed from L2 as pre-swizzled [N x 64] chunks by 1-D bulk TMA copies through a 2-stage
//     mbarrier ring; both tile slots consume a chunk before it is released, halving L2 -> SMEM traffic.
//   - concat_rows(h0, h4) (network.rs:210) is split-K: the PE tile (K = 64) and h4 (K = 256) accumulate into
//     the same D.  concat_rows(bottleneck, PE4(dir)) (network.rs:220) likewise (K = 256 + 32).
//   - sigma (256 -> 1) and rgb (128 -> 3) heads are f32 dot products in the epilogue on the f32 accumulators.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warp 2 = TMEM allocator,
// warps 4..7 = epilogue of slot X, warps 8..11 = epilogue of slot Y (warp % 4 = its TMEM lane quadrant).
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstring>
#include <string>
#include <vector>

#include "mlp_tc.h"

namespace nerf {
namespace {

constexpr int kTileM = 128;
constexpr int kThreads = 384;
constexpr int kStages = 2;
constexpr uint32_t kChunkK = 64;                 // K elements per weight chunk (one 128-byte swizzle row)
constexpr uint32_t kStageBytes = 256 * 128;      // [256 x 64] 16-bit
constexpr uint32_t kKBlockBytes = kTileM * 128;  // one [128 x 64] K-block of an A tile
constexpr uint32_t kActBytes = 4 * kKBlockBytes; // [128 x 256]
constexpr uint32_t kAuxBytes = kKBlockBytes;     // [128 x 64]: PE10(p) for layers 0/5, later PE4(dir)
constexpr uint32_t kOffAct = 0;
constexpr uint32_t kOffAux = kOffAct + 2 * kActBytes;
constexpr uint32_t kOffStage = kOffAux + 2 * kAuxBytes;
constexpr uint32_t kOffBar = kOffStage + kStages * kStageBytes;
constexpr uint32_t kSmemBytes = kOffBar + 128 + 1024;  // + barriers + slack for manual 1 KiB alignment

// Layer table: chunks per layer in consumption order.
constexpr int kNumLayers = 10;  // dense0..7, bottleneck, viewdirs
__host__ __device__ constexpr int layer_chunks(int l) { return l == 0 ? 1 : ((l == 5 || l == 9) ? 5 : 4); }
__host__ __device__ constexpr int layer_n(int l) { return l == 9 ? 128 : 256; }

// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
//   [0,14) start >> 4   [16,30) LBO >> 4 (unused for swizzled K-major, 1)   [32,46) SBO >> 4 (8 rows = 1024 B)
//   [46,48) version = 1   [61,64) layout = 2 (SWIZZLE_128B)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): c_format F32 [4,6)=1; a/b format [7,10),[10,13):
// 0 = F16, 1 = BF16; a/b major K (0); N>>3 at [17,23); M>>4 at [24,29).
__host__ __device__ constexpr uint32_t make_idesc(int fmt_is_fp16, int n) {
  const uint32_t f = fmt_is_fp16 ? 0u : 1u;
  return (1u << 4) | (f << 7) | (f << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
}

// Byte offset of element (row, k) inside a [rows x 64] 16-bit K-major SWIZZLE_128B tile:
// 8-row atoms of 1024 B; inside an atom row r8 the 16-byte chunk index is XORed with r8.
__host__ __device__ inline uint32_t sw128_offset(uint32_t row, uint32_t k) {
  const uint32_t r8 = row & 7u;
  return (row >> 3) * 1024u + r8 * 128u + ((((k >> 3) ^ r8) & 7u) << 4) + (k & 7u) * 2u;
}

// f32 pair -> packed 16-bit pair (lo = first element), optional ReLU fused into the conversion.
template <bool kFp16, bool kRelu>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  uint32_t r;
  if constexpr (kFp16) {
    if constexpr (kRelu) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    else asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    if constexpr (kRelu) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}
template <bool kFp16>
__device__ __forceinline__ uint16_t cvt1(float v) {
  if constexpr (kFp16) return __half_as_ushort(__float2half_rn(v));
  else return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}


// ------------------------------------------------------------------------------------------ epilogue pieces
// One 32-column block of a layer's accumulator for this thread's row: D -> +bias -> (ReLU) -> 16-bit -> A tile.
// kSigma additionally accumulates the f32 sigma head on the unrounded ReLU outputs.
template <bool kFp16, bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_block(uint32_t taddr, const float* __restrict__ bias, const float* __restrict__ walpha,
                                          float& sig, uint8_t* act_row /* A tile + row offset */, uint32_t r8, int cb) {
  uint32_t v[32];
  tc_ld32(taddr + cb * 32, v);
  tc_wait_ld();
  float x[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]) + bias[cb * 32 + j];
  if constexpr (kSigma) {
#pragma unroll
    for (int j = 0; j < 32; ++j) sig = fmaf(fmaxf(x[j], 0.0f), walpha[cb * 32 + j], sig);
  }
  if constexpr (kStore) {
    uint8_t* kblock = act_row + (cb >> 1) * kKBlockBytes;
    const uint32_t chunk0 = (cb & 1) * 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 q;
      q.x = pack2<kFp16, kRelu>(x[i * 8 + 0], x[i * 8 + 1]);
      q.y = pack2<kFp16, kRelu>(x[i * 8 + 2], x[i * 8 + 3]);
      q.z = pack2<kFp16, kRelu>(x[i * 8 + 4], x[i * 8 + 5]);
      q.w = pack2<kFp16, kRelu>(x[i * 8 + 6], x[i * 8 + 7]);
      *reinterpret_cast<uint4*>(kblock + (((chunk0 + i) ^ r8) << 4)) = q;
    }
  }
}

template <bool kFp16, bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_layer(uint32_t taddr, const float* __restrict__ bias,
                                          const float* __restrict__ walpha, float& sig, uint8_t* act_row, uint32_t r8) {
#pragma unroll 1
  for (int cb = 0; cb < 8; ++cb) epi_block<kFp16, kRelu, kSigma, kStore>(taddr, bias, walpha, sig, act_row, r8, cb);
}

// positional_encoding_* (network.rs:263-330) of one sample into row `row` of a [128 x 64] aux tile.
template <bool kFp16>
__device__ __forceinline__ void write_encoding(uint8_t* aux, uint32_t row, const float v[3], int octaves, int pad_to) {
#pragma unroll
  for (int a = 0; a < 3; ++a) *reinterpret_cast<uint16_t*>(aux + sw128_offset(row, a)) = cvt1<kFp16>(v[a]);
#pragma unroll 1
  for (int o = 0; o < octaves; ++o) {
    const float f = (float)(1 << o);  // f doubles per octave (network.rs:287)
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      float sv, cv;
      sincosf(f * v[a], &sv, &cv);
      *reinterpret_cast<uint16_t*>(aux + sw128_offset(row, 3 + 6 * o + a)) = cvt1<kFp16>(sv);
      *reinterpret_cast<uint16_t*>(aux + sw128_offset(row, 6 + 6 * o + a)) = cvt1<kFp16>(cv);
    }
  }
  for (int k = 3 + 6 * octaves; k < pad_to; ++k) *reinterpret_cast<uint16_t*>(aux + sw128_offset(row, k)) = 0;
}

// ------------------------------------------------------------------------------------------ the kernel
template <bool kFp16>
__global__ void __launch_bounds__(kThreads, 1)
mlp_tc_kernel(const __grid_constant__ TcConsts C, const uint8_t* __restrict__ chunks, const SampleSrc src,
              const SampleDst dst) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1 KiB alignment
  uint8_t* smem = smem_raw + (base - raw);
  const uint32_t bar0 = base + kOffBar;
  const uint32_t bar_full = bar0, bar_empty = bar0 + 16, bar_acc = bar0 + 32, bar_act = bar0 + 48;  // 2 x 8 B each
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffBar + 64);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool want_rgb = dst.rgb != nullptr;
  const int n_layers = want_rgb ? kNumLayers : 8;
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_pairs = (n_tiles + 1) / 2;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_acc + 8 * s, 1);
      mbar_init(bar_act + 8 * s, kTileM);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(base + kOffBar + 64), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================================== TMA producer: stream weight chunks through the ring
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
        const uint8_t* src_ptr = chunks;
        for (int l = 0; l < n_layers; ++l) {
          const uint32_t bytes = (uint32_t)layer_n(l) * 128u;
          for (int c = 0; c < layer_chunks(l); ++c) {
            mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
            mbar_arrive_expect_tx(bar_full + 8 * stage, bytes);
            tma_bulk_g2s(base + kOffStage + stage * kStageBytes, src_ptr, bytes, bar_full + 8 * stage);
            src_ptr += bytes;
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer: one thread drives the tensor core
    if (lane == 0) {
      uint32_t stage = 0, phase = 0, act_par[2] = {0u, 0u};
      for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
        const int n_active = (2 * pair + 1 < n_tiles) ? 2 : 1;
        for (int l = 0; l < n_layers; ++l) {
          const uint32_t idesc = make_idesc(kFp16 ? 1 : 0, layer_n(l));
          for (int s = 0; s < n_active; ++s) {
            mbar_wait(bar_act + 8 * s, act_par[s]);
            act_par[s] ^= 1u;
          }
          tc_fence_after();
          const int nch = layer_chunks(l);
          for (int c = 0; c < nch; ++c) {
            mbar_wait(bar_full + 8 * stage, phase);
            tc_fence_after();
            const bool from_aux = (l == 0) || (l == 5 && c == 0) || (l == 9 && c == 4);
            const int ksteps = (l == 9 && c == 4) ? 2 : 4;  // PE4(dir): 27 rows padded to K = 32
            const int kb = (l == 5) ? c - 1 : c;
            const uint32_t b_addr = base + kOffStage + stage * kStageBytes;
            for (int s = 0; s < n_active; ++s) {
              const uint32_t a_addr =
                  from_aux ? base + kOffAux + s * kAuxBytes : base + kOffAct + s * kActBytes + kb * kKBlockBytes;
              const uint32_t d_tmem = tmem_base + (uint32_t)s * 256u;
              for (int ks = 0; ks < ksteps; ++ks)
                tc_mma_f16(d_tmem, make_desc(a_addr + ks * 32), make_desc(b_addr + ks * 32), idesc,
                           (c | ks) != 0 ? 1u : 0u);
              if (c == nch - 1) tc_commit(bar_acc + 8 * s);  // accumulator of this layer complete -> epilogue
            }
            tc_commit(bar_empty + 8 * stage);  // both slots have consumed the chunk -> producer may refill
            if (++stage == kStages) { stage = 0; phase ^= 1u; }
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================================================== epilogue warps: one thread per sample row
    const int slot = (warp - 4) >> 2;
    const uint32_t quad = (uint32_t)(warp & 3);
    const uint32_t row = quad * 32u + (uint32_t)lane;
    const uint32_t r8 = row & 7u;
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + (uint32_t)slot * 256u;
    uint8_t* act_row = smem + kOffAct + slot * kActBytes + (row >> 3) * 1024u + r8 * 128u;
    uint8_t* aux = smem + kOffAux + slot * kAuxBytes;
    const uint32_t acc_bar = bar_acc + 8 * slot, act_bar = bar_act + 8 * slot;
    uint32_t acc_par = 0;
    for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
      const int64_t tile = 2 * pair + slot;
      if (tile >= n_tiles) continue;
      const int64_t g = tile * kTileM + row;
      const bool valid = g < src.B;
      float p[3] = {0.f, 0.f, 0.f}, d[3] = {0.f, 0.f, 0.f};
      if (valid) {
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
          p[0] = __fadd_rn(src.origin[0], __fmul_rn(d[0], t));  // origin + dir_hat * t (lib.rs:396,436)
          p[1] = __fadd_rn(src.origin[1], __fmul_rn(d[1], t));
          p[2] = __fadd_rn(src.origin[2], __fmul_rn(d[2], t));
        }
      }
      write_encoding<kFp16>(aux, row, p, kPosOctaves, 64);  // layer-0 / skip operand (network.rs:204)
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(act_bar);

      float sig = C.b_alpha;
#pragma unroll 1
      for (int l = 0; l < 7; ++l) {
        mbar_wait(acc_bar, acc_par);
        acc_par ^= 1u;
        tc_fence_after();
        epi_layer<kFp16, true, false, true>(taddr, C.bias[l], C.w_alpha, sig, act_row, r8);
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(act_bar);
      }
      mbar_wait(acc_bar, acc_par);  // dense7 -> h8; sigma = relu(alpha(h8)) (network.rs:216)
      acc_par ^= 1u;
      tc_fence_after();
      if (want_rgb) {
        epi_layer<kFp16, true, true, true>(taddr, C.bias[7], C.w_alpha, sig, act_row, r8);
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(act_bar);
      } else {
        epi_layer<kFp16, true, true, false>(taddr, C.bias[7], C.w_alpha, sig, act_row, r8);
        tc_fence_before();
      }
      if (valid) dst.sigma[g] = fmaxf(sig, 0.0f);
      if (want_rgb) {
        mbar_wait(acc_bar, acc_par);  // bottleneck, ActivationType::None (network.rs:218)
        acc_par ^= 1u;
        tc_fence_after();
        epi_layer<kFp16, false, false, true>(taddr, C.bias[8], C.w_alpha, sig, act_row, r8);
        write_encoding<kFp16>(aux, row, d, kDirOctaves, 32);  // PE4(dir) replaces PE10(p): skip MMA is long done
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(act_bar);

        mbar_wait(acc_bar, acc_par);  // viewdirs (ReLU) then the f32 rgb head + sigmoid (network.rs:222-223)
        acc_par ^= 1u;
        tc_fence_after();
        float r0 = C.b_rgb[0], r1 = C.b_rgb[1], r2 = C.b_rgb[2];
#pragma unroll 1
        for (int cb = 0; cb < 4; ++cb) {
          uint32_t v[32];
          tc_ld32(taddr + cb * 32, v);
          tc_wait_ld();
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float cj = fmaxf(__uint_as_float(v[j]) + C.bias[9][cb * 32 + j], 0.0f);
            r0 = fmaf(cj, C.w_rgb[(cb * 32 + j) * 3 + 0], r0);
            r1 = fmaf(cj, C.w_rgb[(cb * 32 + j) * 3 + 1], r1);
            r2 = fmaf(cj, C.w_rgb[(cb * 32 + j) * 3 + 2], r2);
          }
        }
        tc_fence_before();
        if (valid) {
          dst.rgb[g * 3 + 0] = 1.0f / (1.0f + expf(-r0));  // network.rs:165
          dst.rgb[g * 3 + 1] = 1.0f / (1.0f + expf(-r1));
          dst.rgb[g * 3 + 2] = 1.0f / (1.0f + expf(-r2));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace
}  // namespace nerf

// ============================================================================ host side
namespace nerf {
namespace {

inline uint16_t host_cvt(float v, int fmt) {
  return fmt == 1 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}

// Appends the [n x 64] SWIZZLE_128B image of W rows [k0, k0 + krows) (W is [in][out] row-major, so the B
// operand element (n, k) is W[k0 + k][n]); rows beyond krows are zero padding.
void append_chunk(std::vector<uint8_t>& img, const std::vector<float>& W, int out_dim, int k0, int krows, int n,
                  int fmt) {
  const size_t base = img.size();
  img.resize(base + (size_t)n * 128, 0);
  for (int row = 0; row < n; ++row)
    for (int k = 0; k < krows; ++k) {
      const uint16_t h = host_cvt(W[(size_t)(k0 + k) * out_dim + row], fmt);
      std::memcpy(&img[base + sw128_offset((uint32_t)row, (uint32_t)k)], &h, 2);
    }
}

}  // namespace

int build_tc_weights(const TensorMap& t, int fmt, const NetF32&, TcWeights* out) {
  std::vector<uint8_t> img;
  img.reserve(1200 * 1024);
  static const char* const names[10] = {"dense0", "dense1", "dense2", "dense3", "dense4",
                                        "dense5", "dense6", "dense7", "bottleneck", "viewdirs"};
  std::memset(&out->consts, 0, sizeof(out->consts));
  for (int l = 0; l < kNumLayers; ++l) {
    const std::vector<float>& W = t.at(std::string(names[l]) + "_kernel").data;
    const std::vector<float>& b = t.at(std::string(names[l]) + "_bias").data;
    const int n = layer_n(l);
    for (int o = 0; o < n; ++o) out->consts.bias[l][o] = b[o];
    if (l == 0) {
      append_chunk(img, W, n, 0, kPosEnc, n, fmt);  // PE10 rows 0..62, k = 63 zero
    } else if (l == 5) {
      append_chunk(img, W, n, 0, kPosEnc, n, fmt);  // rows 0..62 multiply h0 (concat_rows(h0, h4), network.rs:210)
      for (int c = 0; c < 4; ++c) append_chunk(img, W, n, kPosEnc + 64 * c, 64, n, fmt);
    } else if (l == 9) {
      for (int c = 0; c < 4; ++c) append_chunk(img, W, n, 64 * c, 64, n, fmt);  // bottleneck FIRST (network.rs:220)
      append_chunk(img, W, n, kHidden, kDirEnc, n, fmt);                        // then PE4(dir) rows 256..282
    } else {
      for (int c = 0; c < 4; ++c) append_chunk(img, W, n, 64 * c, 64, n, fmt);
    }
  }
  const std::vector<float>& wa = t.at("alpha_kernel").data;
  const std::vector<float>& wr = t.at("rgb_kernel").data;
  for (int k = 0; k < 256; ++k) out->consts.w_alpha[k] = wa[k];
  for (int k = 0; k < 128 * 3; ++k) out->consts.w_rgb[k] = wr[k];
  for (int k = 0; k < 3; ++k) out->consts.b_rgb[k] = t.at("rgb_bias").data[k];
  out->consts.b_alpha = t.at("alpha_bias").data[0];
  out->fmt = fmt;
  out->bytes = img.size();
  cudaError_t e = cudaMalloc(&out->d_chunks, img.size());
  if (e == cudaSuccess) e = cudaMemcpy(out->d_chunks, img.data(), img.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error(std::string("upload tensor-core weight tiles: ") + cudaGetErrorString(e));
    return NERF_E_CUDA;
  }
  return NERF_OK;
}

void free_tc_weights(TcWeights* w) {
  if (w->d_chunks) cudaFree(w->d_chunks);
  w->d_chunks = nullptr;
  w->bytes = 0;
}

cudaError_t launch_mlp_tc(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                          cudaStream_t s) {
  if (src.B == 0) return cudaSuccess;
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_pairs = (n_tiles + 1) / 2;
  const int grid = (int)(n_pairs < sm_count ? n_pairs : sm_count);
  cudaError_t e;
  if (w.fmt == 1) {
    e = cudaFuncSetAttribute(mlp_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    mlp_tc_kernel<true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst);
  } else {
    e = cudaFuncSetAttribute(mlp_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    mlp_tc_kernel<false><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst);
  }
  return cudaGetLastError();
}

}  // namespace nerf
