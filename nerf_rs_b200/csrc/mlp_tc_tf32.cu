// mlp_tc_tf32.cu — Network::forward_batch (src/network.rs:197-237) with tcgen05 kind::tf32 MMAs: the "tf32" mode.
//
// The north star names pre-swizzled tf32 tiles and a tf32 mode beside bf16.  Operands here are 32-bit containers of
// which the tensor core uses sign, 8 exponent and 10 mantissa bits - the mantissa of fp16 with the range of f32 - at
// HALF the fp16 rate (K = 8 per instruction instead of 16).  On this network (|activations| <= ~50, |weights| <= 8.4)
// it computes what the fp16 mode computes, at less than half its speed: it exists for completeness of the mode list
// and for the orbit sweep's "tf32 vs bf16" line, not as a mode to choose.
//
// Structure: the split-operand kernel's (mlp_tc_x3.cu), because the shapes coincide.  A [128 x 256] tf32 activation
// tile is 128 KiB - what the hi and lo fp16 images take there - so a CTA pair again holds ONE 128-sample tile per CTA
// and a layer's MMAs and its drain alternate.  K-blocks are [rows x 32] (a 128-byte SWIZZLE_128B row holds 32 tf32), so a
// 256-deep layer is eight ring fills of one [N/2 x 32] chunk per CTA (16 KiB: the same ring stage), four K = 8 MMAs
// each, all into one accumulator (TMEM columns 0..255).
//   shared memory: act [128 x 256] tf32 (8 K-blocks) | aux [128 x 64] tf32 (2 K-blocks) | 4 x 16 KiB ring | barriers
//   aux K-block 0, 1 = PE10(p) columns 0..31, 32..63 (layer 0 / skip operand); later K-block 0 = PE4(dir) and K-block 1
//   = f32 scratch for combining the four column quarters of a row.
// Warp roles as in mlp_tc.cu: warps 0..15 epilogue (warp % 4 = TMEM lane quadrant, warp / 4 = column quarter),
// 16 TMA producer, 17 MMA issuer (leader CTA), 18 TMEM allocator, 19 relay.
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "mlp_tc.h"
#include "tc_ptx.cuh"

namespace nerf {
namespace {

constexpr int kThreads = 640;
constexpr int kEpiWarps = 16, kWarpTma = 16, kWarpMma = 17, kWarpAlloc = 18, kWarpRelay = 19;
constexpr int kStages = 4;
constexpr uint32_t kStageBytes = 128 * 128;      // this CTA's N-half of one [N x 32] tf32 chunk
constexpr uint32_t kKBlockBytes = kTileM * 128;  // one [128 x 32] tf32 K-block of an A tile
constexpr uint32_t kActBytes = 8 * kKBlockBytes; // [128 x 256] tf32
constexpr uint32_t kAuxBytes = 2 * kKBlockBytes; // [128 x 64] tf32
constexpr uint32_t kOffAct = 0, kOffAux = kActBytes, kOffStage = kOffAux + kAuxBytes;
constexpr uint32_t kOffBar = kOffStage + kStages * kStageBytes;
// barrier block, 8 B each: full[4] pfull[4] empty[4] acc act | tmem slot
constexpr uint32_t kBarFull = 0, kBarPFull = 32, kBarEmpty = 64, kBarAcc = 96, kBarAct = 104, kTmemSlotOff = 128;
constexpr uint32_t kSmemBytes = kOffBar + 256;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
constexpr int kEpiRegs = 112, kCtlRegs = 32;
constexpr int kNumLayers = 10;  // dense0..7, bottleneck, viewdirs

// 32-row K-blocks of a layer's input in consumption order: dense0 = PE10(p) (63 -> 64 rows), dense5 = PE10(p) then h4
// (concat_rows(h0, h4), network.rs:210), viewdirs = bottleneck then PE4(dir) (27 -> 32 rows; network.rs:220).
__host__ __device__ constexpr int layer_kblocks(int l) { return l == 0 ? 2 : (l == 5 ? 10 : (l == 9 ? 9 : 8)); }
__host__ __device__ constexpr int layer_n(int l) { return l == 9 ? 128 : 256; }

__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kCtlRegs)); }
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kEpiRegs)); }
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// f32 -> tf32 container, round to nearest (the tensor core would otherwise drop the low 13 bits).
__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
// Instruction descriptor for kind::tf32 (cute::UMMA::InstrDescriptor): c_format F32 [4,6) = 1; a / b format [7,10),
// [10,13) = 2 (TF32); K-major; N >> 3 at [17,23); M >> 4 at [24,29), M = 256 across the CTA pair.
__host__ __device__ constexpr uint32_t make_idesc_tf32(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)((2 * kTileM) >> 4) << 24);
}
__device__ __forceinline__ void tc_mma2_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Byte offset of element (row, k) inside a [rows x 32] tf32 K-major SWIZZLE_128B K-block.
__host__ __device__ inline uint32_t sw128_offset_tf32(uint32_t row, uint32_t k) {
  const uint32_t r8 = row & 7u;
  return (row >> 3) * 1024u + r8 * 128u + ((((k >> 2) ^ r8) & 7u) << 4) + (k & 3u) * 4u;
}

// One 32-column accumulator block of a row = one K-block row of the next layer's A operand: f32 bias, optional sigma
// partial on the unrounded values, optional ReLU, then eight 16-byte chunks of four tf32 each.
template <bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_block(uint32_t (&v)[32], const float* __restrict__ bias,
                                          const float* __restrict__ walpha, float& sig, uint8_t* kblock_row,
                                          const uint32_t (&swz)[8]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float x = __uint_as_float(v[j]) + bias[j];
    if constexpr (kRelu || kSigma) x = fmaxf(x, 0.0f);  // the sigma head reads relu(h8) (network.rs:214-216)
    v[j] = __float_as_uint(x);
  }
  if constexpr (kSigma) {
#pragma unroll
    for (int j = 0; j < 32; ++j) sig = fmaf(__uint_as_float(v[j]), walpha[j], sig);
  }
  if constexpr (kStore) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      uint4 q;
      q.x = to_tf32(__uint_as_float(v[4 * i + 0]));
      q.y = to_tf32(__uint_as_float(v[4 * i + 1]));
      q.z = to_tf32(__uint_as_float(v[4 * i + 2]));
      q.w = to_tf32(__uint_as_float(v[4 * i + 3]));
      *reinterpret_cast<uint4*>(kblock_row + swz[i]) = q;
    }
  }
}

// The warp's 64 columns of one layer = K-blocks 2 * colq and 2 * colq + 1 of the next A tile; `kblock_row` is this row
// of the first of them, bias / walpha already point at the warp's first column.
template <bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_cols(uint32_t taddr, const float* __restrict__ bias,
                                         const float* __restrict__ walpha, float& sig, uint8_t* kblock_row,
                                         const uint32_t (&swz)[8]) {
  uint32_t v[32];
  tc_ld32(taddr, v);
  tc_wait_ld();
  epi_block<kRelu, kSigma, kStore>(v, bias, walpha, sig, kblock_row, swz);
  tc_ld32(taddr + 32, v);
  tc_wait_ld();
  epi_block<kRelu, kSigma, kStore>(v, bias + 32, walpha + 32, sig, kblock_row + kKBlockBytes, swz);
}

// positional_encoding_* (network.rs:263-330): columns [kCol0, kCol0 + 16) of one sample's encoding as four 16-byte
// chunks of its row in aux K-block kCol0 / 32.  Column layout: 0..2 = v, 3 + 6o + a = sin(2^o v_a), 6 + 6o + a =
// cos(2^o v_a), zero padding after 3 + 6 * kOctaves.  Every octave is evaluated directly; 2^o * v is exact, as in the
// reference (f doubles per octave, network.rs:287).
template <int kCol0, int kOctaves>
__device__ __forceinline__ void write_pe16(uint8_t* aux_row, uint32_t r8, const float (&v)[3]) {
  float out[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) out[j] = 0.0f;
#pragma unroll
  for (int o = 0; o < kOctaves; ++o) {
    if (3 + 6 * o + 5 >= kCol0 && 3 + 6 * o <= kCol0 + 15) {  // columns of octave o: [3 + 6o, 9 + 6o)
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        float s, c;
        fast_sincos((float)(1 << o) * v[a], s, c);
        const int js = 3 + 6 * o + a - kCol0, jc = js + 3;
        if (js >= 0 && js < 16) out[js] = s;
        if (jc >= 0 && jc < 16) out[jc] = c;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 16; ++j)
    if (kCol0 + j < 3) out[j] = v[kCol0 + j];
  uint8_t* kb_row = aux_row + (kCol0 / 32) * kKBlockBytes;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 q;
    q.x = to_tf32(out[4 * i + 0]);
    q.y = to_tf32(out[4 * i + 1]);
    q.z = to_tf32(out[4 * i + 2]);
    q.w = to_tf32(out[4 * i + 3]);
    *reinterpret_cast<uint4*>(kb_row + ((((uint32_t)((kCol0 % 32) / 4 + i)) ^ r8) << 4)) = q;
  }
}

// ------------------------------------------------------------------------------------------ the kernel
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_tc_tf32_kernel(const __grid_constant__ TcConsts C, const uint8_t* __restrict__ chunks, const SampleSrc src,
                   const SampleDst dst) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0u) __trap();  // SWIZZLE_128B atoms need 1 KiB alignment
  const uint32_t bar0 = base + kOffBar;
  const uint32_t bar_full = bar0 + kBarFull, bar_pfull = bar0 + kBarPFull, bar_empty = bar0 + kBarEmpty,
                 bar_acc = bar0 + kBarAcc, bar_act = bar0 + kBarAct;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffBar + kTmemSlotOff);

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();  // 0 = leader (issues the pair's MMAs), 1 = peer
  const int64_t pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const bool want_rgb = dst.rgb != nullptr;
  const int n_layers = want_rgb ? kNumLayers : 8;
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_units = (n_tiles + 1) / 2;  // unit u = tiles 2u (leader CTA) and 2u + 1 (peer CTA)

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);   // this CTA's bulk copy (expect_tx)
      mbar_init(bar_pfull + 8 * s, 2);  // leader only: both CTAs' relays
      mbar_init(bar_empty + 8 * s, 1);  // the commit after the chunk's last MMA
    }
    mbar_init(bar_acc, 1);
    mbar_init(bar_act, 32);  // leader only: one elected arrive per epilogue warp, both CTAs
    fence_barrier_init();
  }
  if (warp == kWarpAlloc) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(base + kOffBar + kTmemSlotOff),
                 "r"(256u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // barriers of both CTAs initialised before anyone arrives remotely
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  int n_fills_unit = 0;  // ring-stage fills per unit: one per K-block
  for (int l = 0; l < n_layers; ++l) n_fills_unit += layer_kblocks(l);

  if (warp == kWarpTma) {
    reg_dec();
    // ===================================================== TMA producer: this CTA's half of every weight chunk
    uint32_t stage = 0, phase = 0;
    for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
      const uint8_t* src_ptr = chunks;
      for (int l = 0; l < n_layers; ++l) {
        const uint32_t bytes = (uint32_t)layer_n(l) * 64u;  // [N/2 x 32] tf32
        for (int c = 0; c < layer_kblocks(l); ++c) {
          mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
          if (elect_one()) {
            mbar_arrive_expect_tx(bar_full + 8 * stage, bytes);
            tma_bulk_g2s(base + kOffStage + stage * kStageBytes, src_ptr + rank * bytes, bytes, bar_full + 8 * stage);
          }
          __syncwarp();
          src_ptr += 2 * bytes;
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == kWarpAlloc) {
    reg_dec();  // all four warps of the group must execute it
  } else if (warp == kWarpRelay) {
    reg_dec();
    // ===================================================== relay: tell the leader that this CTA's half has landed
    uint32_t stage = 0, phase = 0;
    for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
      for (int c = 0; c < n_fills_unit; ++c) {
        mbar_wait(bar_full + 8 * stage, phase);
        if (elect_one()) mbar_arrive_cluster(bar_pfull + 8 * stage, 0u);
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == kWarpMma) {
    reg_dec();
    // ===================================================== MMA issuer (leader CTA): one thread drives both tensor cores
    if (rank == 0) {
      const uint64_t act_desc0 = make_desc(base + kOffAct), aux_desc0 = make_desc(base + kOffAux);
      const uint64_t stage_desc0 = make_desc(base + kOffStage);
      uint32_t g = 0;        // ring fill counter
      uint32_t act_par = 0;  // parity of the "layer input ready" barrier
      for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
#pragma unroll 1
        for (int l = 0; l < n_layers; ++l) {
          const int nkb = layer_kblocks(l);
          const uint32_t idesc = make_idesc_tf32(layer_n(l));
          // which K-blocks read the aux tile: dense0 both of its two, dense5 its first two (PE10(p) comes FIRST in
          // concat_rows(h0, h4)), viewdirs its last (PE4(dir) comes after the bottleneck)
          const int aux_first = (l == 0 || l == 5) ? 0 : (l == 9 ? 8 : nkb);
          const int aux_count = (l == 0 || l == 5) ? 2 : (l == 9 ? 1 : 0);
          mbar_wait(bar_act, act_par);  // layer input written and accumulator drained (both CTAs)
          act_par ^= 1u;
#pragma unroll 1
          for (int kb = 0; kb < nkb; ++kb) {
            const bool is_aux = kb >= aux_first && kb < aux_first + aux_count;
            const int a_kb = is_aux ? kb - aux_first : (l == 5 ? kb - 2 : kb);
            const uint64_t a_desc = (is_aux ? aux_desc0 : act_desc0) + (uint32_t)a_kb * (kKBlockBytes >> 4);
            const uint32_t stage = g & (kStages - 1);
            mbar_wait(bar_pfull + 8 * stage, (g / kStages) & 1u);
            tc_fence_after();
            const uint64_t bd = stage_desc0 + stage * (kStageBytes >> 4);
            if (elect_one()) {
#pragma unroll 1
              for (int j = 0; j < 4; ++j)  // four K = 8 steps per 32-deep K-block: 32 bytes = 2 descriptor units each
                tc_mma2_tf32(tmem_base, a_desc + 2u * (uint32_t)j, bd + 2u * (uint32_t)j, idesc, (kb | j) != 0 ? 1u : 0u);
              if (kb == nkb - 1) tc_commit2(bar_acc);  // this layer's accumulator complete -> both epilogues
              tc_commit2(bar_empty + 8 * stage);
            }
            __syncwarp();
            ++g;
          }
        }
      }
    }
  } else if (warp < kEpiWarps) {
    reg_inc();
    // ===================================================== epilogue warps: one thread per (sample row, column quarter)
    const int colq = warp >> 2;
    const uint32_t quad = (uint32_t)(warp & 3);
    const uint32_t row = quad * 32u + (uint32_t)lane;
    const uint32_t r8 = row & 7u;
    const uint32_t row_off = (row >> 3) * 1024u + r8 * 128u;  // this row inside any K-block
    const uint32_t taddr0 = tmem_base + ((quad * 32u) << 16) + (uint32_t)colq * 64u;
    uint8_t* const kblock_row = smem + kOffAct + (uint32_t)(2 * colq) * kKBlockBytes + row_off;  // K-block 2 * colq
    uint8_t* const aux_row = smem + kOffAux + row_off;
    uint32_t swz[8];  // byte offset of 16-byte chunk c inside the row's 128 bytes (SWIZZLE_128B)
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c) swz[c] = (c ^ r8) << 4;
    // f32 scratch for combining the four column quarters of a row: quarter k owns chunk k of this row in aux K-block 1
    // = {rgb partials, sigma partial}; that K-block is free whenever the scratch is used (see the file header)
    auto scratch_of = [&](uint32_t k) { return reinterpret_cast<float*>(smem + kOffAux + kKBlockBytes + row_off + ((k ^ r8) << 4)); };
    float* const scratch = scratch_of((uint32_t)colq);
    const uint32_t row_bar = 1u + quad;  // named barrier of the 4 warps (128 threads) that share a row
    uint32_t acc_par = 0;

    auto acc_wait = [&]() {
      mbar_wait(bar_acc, acc_par);
      acc_par ^= 1u;
      tc_fence_after();
    };
    auto act_ready = [&]() {  // this warp's part of the next A operand is in shared memory, its columns drained
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(bar_act, 0u);
    };

    float4 pre;  // {p.xyz, -} for explicit points, {dir.xyz, t} for samples along rays
    auto prefetch = [&](int64_t unit) {
      const int64_t gi = (2 * unit + (int64_t)rank) * kTileM + row;
      pre = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gi < src.B) {
        if (src.mode == 0) {
          pre = make_float4(src.points[gi], src.points[src.B + gi], src.points[2 * src.B + gi], 0.f);
        } else {
          const int64_t ray = gi / src.S;
          pre = make_float4(src.dirs[ray * 3 + 0], src.dirs[ray * 3 + 1], src.dirs[ray * 3 + 2], src.t[gi]);
        }
      }
    };
    prefetch(pair_id);
    for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
      const int64_t g = (2 * unit + (int64_t)rank) * kTileM + row;
      const bool valid = g < src.B;  // tiles past the end run the same barrier protocol on zeros
      float p[3] = {pre.x, pre.y, pre.z};
      float dir[3] = {0.f, 0.f, 0.f};
      if (src.mode == 0) {
        if (valid) {
          dir[0] = src.dirs[g * 3 + 0];
          dir[1] = src.dirs[g * 3 + 1];
          dir[2] = src.dirs[g * 3 + 2];
        }
      } else {
        dir[0] = pre.x;
        dir[1] = pre.y;
        dir[2] = pre.z;
        if (valid) {
          p[0] = __fadd_rn(src.origin[0], __fmul_rn(dir[0], pre.w));  // origin + dir_hat * t (lib.rs:396,436)
          p[1] = __fadd_rn(src.origin[1], __fmul_rn(dir[1], pre.w));
          p[2] = __fadd_rn(src.origin[2], __fmul_rn(dir[2], pre.w));
        }
      }
      // PE10(p): layer-0 / skip operand (network.rs:204); each column quarter writes its 16 of the 64 columns
      if (colq == 0) write_pe16<0, kPosOctaves>(aux_row, r8, p);
      else if (colq == 1) write_pe16<16, kPosOctaves>(aux_row, r8, p);
      else if (colq == 2) write_pe16<32, kPosOctaves>(aux_row, r8, p);
      else write_pe16<48, kPosOctaves>(aux_row, r8, p);
      act_ready();

      float sig = colq == 0 ? C.b_alpha : 0.0f;
#pragma unroll 1
      for (int l = 0; l < 7; ++l) {
        acc_wait();
        epi_cols<true, false, true>(taddr0, C.bias[l] + colq * 64, C.w_alpha, sig, kblock_row, swz);
        act_ready();
      }
      if (unit + n_pairs < n_units) prefetch(unit + n_pairs);
      // dense7 -> h8; sigma = relu(alpha(h8)) (network.rs:216)
      acc_wait();
      if (want_rgb) {
        epi_cols<true, true, true>(taddr0, C.bias[7] + colq * 64, C.w_alpha + colq * 64, sig, kblock_row, swz);
        act_ready();
      } else {
        epi_cols<true, true, false>(taddr0, C.bias[7] + colq * 64, C.w_alpha + colq * 64, sig, kblock_row, swz);
        tc_fence_before();
      }
      if (colq != 0) scratch[3] = sig;
      named_bar_sync(row_bar, 128);
      if (colq == 0) {
        float v = sig;
#pragma unroll
        for (uint32_t k = 1; k < 4; ++k) v += scratch_of(k)[3];
        if (valid) dst.sigma[g] = fmaxf(v, 0.0f);
      }
      if (want_rgb) {
        // bottleneck, ActivationType::None (network.rs:218)
        acc_wait();
        epi_cols<false, false, true>(taddr0, C.bias[8] + colq * 64, C.w_alpha, sig, kblock_row, swz);
        // PE4(dir) replaces PE10(p) in aux K-block 0: the skip MMAs that read it finished layers ago
        if (colq == 0) write_pe16<0, kDirOctaves>(aux_row, r8, dir);
        else if (colq == 1) write_pe16<16, kDirOctaves>(aux_row, r8, dir);  // columns 27..31: K padding
        act_ready();
        // viewdirs (ReLU) then the f32 rgb head (network.rs:222-223): 32 columns per warp
        acc_wait();
        uint32_t v[32];
        tc_ld32(tmem_base + ((quad * 32u) << 16) + 32u * (uint32_t)colq, v);
        tc_wait_ld();
        float r0 = 0.f, r1 = 0.f, r2 = 0.f;
        if (colq == 0) { r0 = C.b_rgb[0]; r1 = C.b_rgb[1]; r2 = C.b_rgb[2]; }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int k = colq * 32 + j;
          const float c = fmaxf(__uint_as_float(v[j]) + C.bias[9][k], 0.0f);
          r0 = fmaf(c, C.w_rgb[k * 3 + 0], r0);
          r1 = fmaf(c, C.w_rgb[k * 3 + 1], r1);
          r2 = fmaf(c, C.w_rgb[k * 3 + 2], r2);
        }
        tc_fence_before();
        if (colq != 0) {
          scratch[0] = r0;
          scratch[1] = r1;
          scratch[2] = r2;
        }
        named_bar_sync(row_bar, 128);
        if (colq == 0) {
#pragma unroll
          for (uint32_t k = 1; k < 4; ++k) {
            const float* sc = scratch_of(k);
            r0 += sc[0];
            r1 += sc[1];
            r2 += sc[2];
          }
          if (valid) {
            dst.rgb[g * 3 + 0] = 1.0f / (1.0f + expf(-r0));  // network.rs:165
            dst.rgb[g * 3 + 1] = 1.0f / (1.0f + expf(-r1));
            dst.rgb[g * 3 + 2] = 1.0f / (1.0f + expf(-r2));
          }
        }
      }
      named_bar_sync(row_bar, 128);  // scratch is inside the aux tile the next unit's PE overwrites
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's shared memory and TMEM stay alive until the pair's last MMA has been consumed
  if (warp == kWarpAlloc) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
  }
}

// f32 -> tf32 on the host: round to nearest, ties away from zero, on the 13 dropped mantissa bits (what cvt.rna does);
// infinities and NaN pass through.
inline float host_tf32(float w) {
  uint32_t u;
  std::memcpy(&u, &w, 4);
  if ((u & 0x7f800000u) != 0x7f800000u) u = (u + 0x1000u) & 0xffffe000u;
  float r;
  std::memcpy(&r, &u, 4);
  return r;
}

// Appends the [n x 32] SWIZZLE_128B image of W rows [k0, k0 + krows): the B operand element (n, k) is W[k0 + k][n]
// (W is [in][out] row-major); rows beyond krows are zero padding.  Rows 0..n/2-1 come first, so each CTA of a pair
// copies its own contiguous half.
void append_chunk_tf32(std::vector<uint8_t>& img, const std::vector<float>& W, int out_dim, int k0, int krows, int n) {
  const size_t base = img.size();
  img.resize(base + (size_t)n * 128, 0);
  for (int row = 0; row < n; ++row) {
    for (int k = 0; k < krows; ++k) {
      const float w = host_tf32(W[(size_t)(k0 + k) * out_dim + row]);
      std::memcpy(&img[base + sw128_offset_tf32((uint32_t)row, (uint32_t)k)], &w, 4);
    }
  }
}

}  // namespace

void build_tc_tf32_image(const TensorMap& t, std::vector<uint8_t>* out) {
  std::vector<uint8_t>& img = *out;
  img.clear();
  img.reserve(2400 * 1024);
  static const char* const names[10] = {"dense0", "dense1", "dense2", "dense3", "dense4",
                                        "dense5", "dense6", "dense7", "bottleneck", "viewdirs"};
  for (int l = 0; l < kNumLayers; ++l) {
    const std::vector<float>& W = t.at(std::string(names[l]) + "_kernel").data;
    const int n = layer_n(l);
    auto rows = [&](int k0, int count) {  // `count` rows of W from k0 on, as K-blocks of 32
      for (int k = 0; k < count; k += 32) append_chunk_tf32(img, W, n, k0 + k, count - k < 32 ? count - k : 32, n);
    };
    if (l == 0) {
      rows(0, kPosEnc);  // PE10 rows 0..62
    } else if (l == 5) {
      rows(0, kPosEnc);  // rows 0..62 multiply h0 (concat_rows(h0, h4), network.rs:210)
      rows(kPosEnc, kHidden);
    } else if (l == 9) {
      rows(0, kHidden);        // bottleneck FIRST (network.rs:220)
      rows(kHidden, kDirEnc);  // then PE4(dir) rows 256..282
    } else {
      rows(0, kHidden);
    }
  }
}

size_t tc_tf32_image_bytes() {
  size_t n = 0;
  for (int l = 0; l < kNumLayers; ++l) n += (size_t)layer_kblocks(l) * layer_n(l) * 128;
  return n;
}

cudaError_t init_mlp_tc_tf32() {  // once per process and device, at context creation
  return cudaFuncSetAttribute(mlp_tc_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
}

cudaError_t launch_mlp_tc_tf32(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                               cudaStream_t s) {
  if (src.B == 0) return cudaSuccess;
  if (src.index != nullptr || src.sub_n != 0 || dst.sigma_floor < 0.0f) return cudaErrorInvalidValue;
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_units = (n_tiles + 1) / 2;  // a CTA pair works on two tiles at a time
  const int64_t max_pairs = sm_count / 2;
  const int grid = 2 * (int)(n_units < max_pairs ? n_units : max_pairs);
  mlp_tc_tf32_kernel<<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst);
  return cudaGetLastError();
}

}  // namespace nerf
