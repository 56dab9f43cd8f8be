// mlp_tc_x3.cu — Network::forward_batch (src/network.rs:197-237) on the tensor cores at (almost) f32 accuracy:
// the "fp16x3" mode.  Every operand of the 256-wide layers is split into two fp16 numbers, x = hi + lo with
// hi = fp16(x) and lo = fp16(x - hi) (22 significand bits together), and every product A.B is evaluated as
//   Ahi.Bhi + Alo.Bhi + Ahi.Blo        (Alo.Blo is below 2^-22 relative and dropped)
// by three tcgen05 kind::f16 MMAs with f32 accumulation in TMEM.  That is 3x the tensor work of the single-pass
// kernel (mlp_tc.cu) for operand error ~2^-22 instead of 2^-11, which no single-pass 16/19-bit format (bf16, fp16,
// tf32) can reach when sigma is O(100).  Two details keep the result at that level:
//   - lo is stored scaled by 2^11 (lo' = fp16((x - hi) * 2048)): an unscaled residual of a typical weight (~0.05) is
//     an fp16 subnormal and would lose most of its bits;
//   - the two small terms accumulate in their own TMEM columns (D2 = sum Alo'.Bhi + Ahi.Blo') and are added to
//     D1 = sum Ahi.Bhi once, as D1 + 2^-11 * D2, by the epilogue: the tensor core truncates its f32 accumulator at
//     every step, so the large accumulator takes 16 steps per 256-deep layer instead of 48.
//
// Structure: the same CTA pair (cta_group::2, M = 256 across the two SMs of a TPC, weights split by N between the
// CTAs, 1-D bulk-TMA weight ring, f32 bias / sigma head / rgb head in the epilogue) as mlp_tc.cu, with one change of
// shape forced by shared memory: an activation tile now needs a hi and a lo image (2 x 64 KiB per CTA), so a pair
// holds ONE 128-sample tile per CTA instead of two slots, and a layer's MMAs and its drain alternate instead of
// overlapping.  With three MMA passes per K-block (6 144 tensor cycles per 256-wide layer) against a ~1 400-cycle
// drain the tensor pipe still is the bound.
//   shared memory: act hi [128 x 256] | act lo | aux hi [128 x 64] | aux lo | 4 x 16 KiB weight ring | barriers
//   weight stream: per layer, per 64-row K-block: hi chunk, lo chunk (each the SWIZZLE_128B image of [N x 64])
//   issue order per K-block kb: wait hi chunk -> Ahi.Bhi, Alo.Bhi -> release; wait lo chunk -> Ahi.Blo -> release
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
constexpr uint32_t kStageBytes = 128 * 128;      // this CTA's N-half of one (K-block, hi|lo) chunk
constexpr uint32_t kKBlockBytes = kTileM * 128;  // one [128 x 64] K-block of an A tile
constexpr uint32_t kActBytes = 4 * kKBlockBytes; // [128 x 256]
constexpr uint32_t kAuxBytes = kKBlockBytes;     // [128 x 64]: PE10(p), later PE4(dir)
constexpr uint32_t kOffActHi = 0, kOffActLo = kActBytes;
constexpr uint32_t kOffAuxHi = 2 * kActBytes, kOffAuxLo = kOffAuxHi + kAuxBytes;
constexpr uint32_t kOffStage = kOffAuxLo + kAuxBytes;
constexpr uint32_t kOffBar = kOffStage + kStages * kStageBytes;
// barrier block, 8 B each: full[4] pfull[4] empty[4] acc act | tmem slot
constexpr uint32_t kBarFull = 0, kBarPFull = 32, kBarEmpty = 64, kBarAcc = 96, kBarAct = 104, kTmemSlotOff = 128;
constexpr uint32_t kSmemBytes = kOffBar + 256;
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
static_assert(kOffActLo - kOffActHi == kActBytes && kOffAuxLo - kOffAuxHi == kAuxBytes,
              "the epilogue and the MMA issuer address a lo image as its hi image + one tile");
constexpr int kEpiRegs = 112, kCtlRegs = 32;
constexpr int kNumLayers = 10;  // dense0..7, bottleneck, viewdirs

// 64-row K-blocks of a layer's input in consumption order: dense0 = PE10(p) alone, dense5 = PE10(p) then h4
// (concat_rows(h0, h4), network.rs:210), viewdirs = bottleneck then PE4(dir) (network.rs:220).
__host__ __device__ constexpr int layer_kblocks(int l) { return l == 0 ? 1 : ((l == 5 || l == 9) ? 5 : 4); }
__host__ __device__ constexpr int layer_n(int l) { return l == 9 ? 128 : 256; }

__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kCtlRegs)); }
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kEpiRegs)); }
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

constexpr float kLoScale = 2048.0f, kLoUnscale = 1.0f / 2048.0f;  // lo' = (x - hi) * 2^11, see the file header

// (a, b) -> packed fp16 pair hi and the packed fp16 pair of the scaled residuals (a - hi.x) * 2^11, (b - hi.y) * 2^11
// (the subtraction and the scaling are exact in f32).
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
  hi = pack2<true, false>(a, b);
  const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hi));
  lo = pack2<true, false>((a - hf.x) * kLoScale, (b - hf.y) * kLoScale);
}
// One 32-column block of a row's accumulator: D1 + 2^-11 * D2 (D2 sits 256 TMEM columns after D1).
__device__ __forceinline__ void ld_acc32(uint32_t taddr, uint32_t (&v)[32]) {
  uint32_t w[32];
  tc_ld32(taddr, v);
  tc_ld32(taddr + 256u, w);
  tc_wait_ld();
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(fmaf(__uint_as_float(w[j]), kLoUnscale, __uint_as_float(v[j])));
}

// One 32-column accumulator block of a row: f32 bias, optional sigma partial on the unrounded values, optional
// ReLU, then the hi / lo images of the next layer's A operand (16-byte chunks chunk0..chunk0+3 of the K-block row).
template <bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_block(uint32_t (&v)[32], const float* __restrict__ bias,
                                          const float* __restrict__ walpha, float& sig, uint8_t* row_hi,
                                          const uint32_t (&swz)[8], int chunk0) {
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
    for (int i = 0; i < 4; ++i) {
      uint4 h, l;
      split2(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]), h.x, l.x);
      split2(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]), h.y, l.y);
      split2(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]), h.z, l.z);
      split2(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]), h.w, l.w);
      *reinterpret_cast<uint4*>(row_hi + swz[chunk0 + i]) = h;
      *reinterpret_cast<uint4*>(row_hi + kActBytes + swz[chunk0 + i]) = l;  // the lo tile sits kActBytes further
    }
  }
}

// The warp's 64 columns of one layer; bias / walpha already point at the warp's first column.
template <bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_cols(uint32_t taddr, const float* __restrict__ bias,
                                         const float* __restrict__ walpha, float& sig, uint8_t* row_hi,
                                         const uint32_t (&swz)[8]) {
  uint32_t v[32];
  ld_acc32(taddr, v);
  epi_block<kRelu, kSigma, kStore>(v, bias, walpha, sig, row_hi, swz, 0);
  ld_acc32(taddr + 32, v);
  epi_block<kRelu, kSigma, kStore>(v, bias + 32, walpha + 32, sig, row_hi, swz, 4);
}

// positional_encoding_* (network.rs:263-330): columns [kCol0, kCol0 + 16) of one sample's encoding as hi / lo
// 16-byte chunks of its aux-tile row.  Column layout: 0..2 = v, 3 + 6o + a = sin(2^o v_a), 6 + 6o + a = cos(2^o v_a),
// zero padding after 3 + 6 * kOctaves.  Every octave is evaluated directly (|err| ~2e-7; the single-pass kernel's
// angle doubling would add ~1e-6, visible at this mode's accuracy); the argument 2^o * v is exact, as in the
// reference (f doubles per octave, network.rs:287).
template <int kCol0, int kOctaves>
__device__ __forceinline__ void write_pe16(uint8_t* aux_row_hi, uint32_t r8, const float (&v)[3]) {
  constexpr int kEnd = 3 + 6 * kOctaves;
  float out[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) out[j] = 0.0f;
#pragma unroll
  for (int o = 0; o < kOctaves; ++o) {
    // columns of octave o: [3 + 6o, 9 + 6o)
    if (3 + 6 * o + 5 >= kCol0 && 3 + 6 * o <= kCol0 + 15) {
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
  (void)kEnd;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint4 qh, ql;
    split2(out[8 * h + 0], out[8 * h + 1], qh.x, ql.x);
    split2(out[8 * h + 2], out[8 * h + 3], qh.y, ql.y);
    split2(out[8 * h + 4], out[8 * h + 5], qh.z, ql.z);
    split2(out[8 * h + 6], out[8 * h + 7], qh.w, ql.w);
    const uint32_t off = (((uint32_t)(kCol0 / 8 + h)) ^ r8) << 4;
    *reinterpret_cast<uint4*>(aux_row_hi + off) = qh;
    *reinterpret_cast<uint4*>(aux_row_hi + kAuxBytes + off) = ql;
  }
}

// ------------------------------------------------------------------------------------------ the kernel
// kGather: refinement launch over a compacted sample list (SampleSrc::index / count), a separate instantiation so that
// the plain kernel keeps its register allocation (no spills at the 112-register epilogue cap).
template <bool kGather>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_tc_x3_kernel(const __grid_constant__ TcConsts C, const uint8_t* __restrict__ chunks, const SampleSrc src,
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
  // refinement launches (SampleSrc::index) learn their size from the list the selection kernel has just built
  const int64_t B = kGather ? (int64_t)__ldg(src.count) : src.B;
  const int64_t n_tiles = (B + kTileM - 1) / kTileM;
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
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // barriers of both CTAs initialised before anyone arrives remotely
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  int n_fills_unit = 0;  // ring-stage fills per unit: (hi, lo) per K-block
  for (int l = 0; l < n_layers; ++l) n_fills_unit += 2 * layer_kblocks(l);

  if (warp == kWarpTma) {
    reg_dec();
    // ===================================================== TMA producer: this CTA's half of every weight chunk
    uint32_t stage = 0, phase = 0;
    for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
      const uint8_t* src_ptr = chunks;
      for (int l = 0; l < n_layers; ++l) {
        const uint32_t bytes = (uint32_t)layer_n(l) * 64u;  // [N/2 x 64] fp16
        for (int c = 0; c < 2 * layer_kblocks(l); ++c) {
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
      const uint64_t act_desc0 = make_desc(base + kOffActHi), aux_desc0 = make_desc(base + kOffAuxHi);
      const uint64_t stage_desc0 = make_desc(base + kOffStage);
      uint32_t g = 0;        // ring fill counter
      uint32_t act_par = 0;  // parity of the "layer input ready" barrier
      for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
#pragma unroll 1
        for (int l = 0; l < n_layers; ++l) {
          const int nkb = layer_kblocks(l);
          const uint32_t idesc = make_idesc(1, layer_n(l));
          const int aux_kb = (l == 0 || l == 5) ? 0 : (l == 9 ? 4 : -1);  // the K-block that is a positional encoding
          const int kb_off = (l == 5) ? 1 : 0;
          mbar_wait(bar_act, act_par);  // layer input written and accumulator drained (both CTAs)
          act_par ^= 1u;
#pragma unroll 1
          for (int kb = 0; kb < nkb; ++kb) {
            const bool is_aux = kb == aux_kb;
            const uint64_t a_hi = is_aux ? aux_desc0 : act_desc0 + (uint32_t)(kb - kb_off) * (kKBlockBytes >> 4);
            const uint64_t a_lo = a_hi + (is_aux ? (kAuxBytes >> 4) : (kActBytes >> 4));
            const int ksteps = (is_aux && l == 9) ? 2 : 4;  // PE4(dir) is K = 32: aux columns 32..63 are scratch by then
            // ---- hi weights: D1 += Ahi.Bhi, D2 += Alo'.Bhi
            uint32_t stage = g & (kStages - 1);
            mbar_wait(bar_pfull + 8 * stage, (g / kStages) & 1u);
            tc_fence_after();
            uint64_t bd = stage_desc0 + stage * (kStageBytes >> 4);
            if (elect_one()) {
#pragma unroll 1
              for (int j = 0; j < ksteps; ++j)
                tc_mma2_f16(tmem_base, a_hi + 2u * (uint32_t)j, bd + 2u * (uint32_t)j, idesc, (kb | j) != 0 ? 1u : 0u);
#pragma unroll 1
              for (int j = 0; j < ksteps; ++j)
                tc_mma2_f16(tmem_base + 256u, a_lo + 2u * (uint32_t)j, bd + 2u * (uint32_t)j, idesc,
                            (kb | j) != 0 ? 1u : 0u);
              tc_commit2(bar_empty + 8 * stage);
            }
            __syncwarp();
            ++g;
            // ---- lo weights: D2 += Ahi.Blo'
            stage = g & (kStages - 1);
            mbar_wait(bar_pfull + 8 * stage, (g / kStages) & 1u);
            tc_fence_after();
            bd = stage_desc0 + stage * (kStageBytes >> 4);
            if (elect_one()) {
#pragma unroll 1
              for (int j = 0; j < ksteps; ++j)
                tc_mma2_f16(tmem_base + 256u, a_hi + 2u * (uint32_t)j, bd + 2u * (uint32_t)j, idesc, 1u);
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
    const uint32_t taddr0 = tmem_base + ((quad * 32u) << 16) + (uint32_t)colq * 64u;
    // this row of K-block colq of the hi A tile
    uint8_t* const kblock_row = smem + kOffActHi + (uint32_t)colq * kKBlockBytes + (row >> 3) * 1024u + r8 * 128u;
    uint8_t* const aux_row = smem + kOffAuxHi + (row >> 3) * 1024u + r8 * 128u;
    uint32_t swz[8];  // byte offset of 16-byte chunk c inside the row's 128 bytes (SWIZZLE_128B)
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c) swz[c] = (c ^ r8) << 4;
    // f32 scratch for combining the four column quarters of a row: quarter k >= 1 owns the 16-byte chunk at aux-hi
    // column 32 + 8k = {rgb partials, sigma partial}; aux columns 32..63 are free whenever it is used
    const uint32_t scratch_off = sw128_offset(row, 32u + 8u * (uint32_t)colq);
    float* const scratch = reinterpret_cast<float*>(smem + kOffAuxHi + scratch_off);
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
    int32_t g_pre = 0;  // kGather: the sample id `pre` belongs to (the launch slot otherwise, recomputed where needed)
    int32_t idx_pre = 0;  // index-list entry of this row's slot in the unit the NEXT prefetch fetches: loaded one unit
                          // earlier, so that the dependent loads of dirs / t never wait for it inside a drain
    auto load_index = [&](int64_t unit) {
      const int64_t gi = (2 * unit + (int64_t)rank) * kTileM + row;
      if (kGather && gi < B) idx_pre = __ldg(src.index + gi);
    };
    auto prefetch = [&](int64_t unit) {
      const int64_t gi = (2 * unit + (int64_t)rank) * kTileM + row;
      pre = make_float4(0.f, 0.f, 0.f, 0.f);
      if (gi < B) {
        if (src.mode == 0) {
          pre = make_float4(src.points[gi], src.points[B + gi], src.points[2 * B + gi], 0.f);
        } else if (kGather) {
          g_pre = idx_pre;
          const int32_t ray = g_pre / src.S;
          pre = make_float4(src.dirs[ray * 3 + 0], src.dirs[ray * 3 + 1], src.dirs[ray * 3 + 2], src.t[g_pre]);
        } else {
          const int64_t ray = gi / src.S;
          pre = make_float4(src.dirs[ray * 3 + 0], src.dirs[ray * 3 + 1], src.dirs[ray * 3 + 2], src.t[gi]);
        }
      }
      if (kGather) load_index(unit + n_pairs);
    };
    if (kGather) load_index(pair_id);
    prefetch(pair_id);
    for (int64_t unit = pair_id; unit < n_units; unit += n_pairs) {
      const int64_t slot = (2 * unit + (int64_t)rank) * kTileM + row;
      const int64_t g = kGather ? (int64_t)g_pre : slot;
      const bool valid = slot < B;  // tiles past the end run the same barrier protocol on zeros
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
        for (uint32_t k = 1; k < 4; ++k)
          v += reinterpret_cast<const float*>(smem + kOffAuxHi + sw128_offset(row, 32u + 8u * k))[3];
        if (valid) dst.sigma[g] = fmaxf(v, 0.0f);
      }
      if (want_rgb) {
        // bottleneck, ActivationType::None (network.rs:218)
        acc_wait();
        epi_cols<false, false, true>(taddr0, C.bias[8] + colq * 64, C.w_alpha, sig, kblock_row, swz);
        // PE4(dir) replaces PE10(p) in the aux tiles (columns 0..31): the skip MMA that read them finished layers ago
        if (colq == 0) write_pe16<0, kDirOctaves>(aux_row, r8, dir);
        else if (colq == 1) write_pe16<16, kDirOctaves>(aux_row, r8, dir);  // columns 27..31: K padding
        act_ready();
        // viewdirs (ReLU) then the f32 rgb head (network.rs:222-223): 32 columns per warp
        acc_wait();
        uint32_t v[32];
        ld_acc32(tmem_base + ((quad * 32u) << 16) + 32u * (uint32_t)colq, v);
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
            const float* sc = reinterpret_cast<const float*>(smem + kOffAuxHi + sw128_offset(row, 32u + 8u * k));
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
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

// Appends the [n x 64] SWIZZLE_128B images (hi, then lo) of W rows [k0, k0 + krows): the B operand element (n, k)
// is W[k0 + k][n] (W is [in][out] row-major); rows beyond krows are zero padding.
void append_split_chunks(std::vector<uint8_t>& img, const std::vector<float>& W, int out_dim, int k0, int krows,
                         int n) {
  const size_t base = img.size();
  img.resize(base + 2 * (size_t)n * 128, 0);
  for (int row = 0; row < n; ++row) {
    for (int k = 0; k < krows; ++k) {
      const float w = W[(size_t)(k0 + k) * out_dim + row];
      const __half hi = __float2half_rn(w);
      const __half lo = __float2half_rn((w - __half2float(hi)) * kLoScale);
      const uint16_t h = __half_as_ushort(hi), l = __half_as_ushort(lo);
      const uint32_t off = sw128_offset((uint32_t)row, (uint32_t)k);
      std::memcpy(&img[base + off], &h, 2);
      std::memcpy(&img[base + (size_t)n * 128 + off], &l, 2);
    }
  }
}

}  // namespace

void build_tc_x3_image(const TensorMap& t, std::vector<uint8_t>* out) {
  std::vector<uint8_t>& img = *out;
  img.clear();
  img.reserve(2600 * 1024);
  static const char* const names[10] = {"dense0", "dense1", "dense2", "dense3", "dense4",
                                        "dense5", "dense6", "dense7", "bottleneck", "viewdirs"};
  for (int l = 0; l < kNumLayers; ++l) {
    const std::vector<float>& W = t.at(std::string(names[l]) + "_kernel").data;
    const int n = layer_n(l);
    if (l == 0) {
      append_split_chunks(img, W, n, 0, kPosEnc, n);  // PE10 rows 0..62
    } else if (l == 5) {
      append_split_chunks(img, W, n, 0, kPosEnc, n);  // rows 0..62 multiply h0 (concat_rows(h0, h4), network.rs:210)
      for (int c = 0; c < 4; ++c) append_split_chunks(img, W, n, kPosEnc + 64 * c, 64, n);
    } else if (l == 9) {
      for (int c = 0; c < 4; ++c) append_split_chunks(img, W, n, 64 * c, 64, n);  // bottleneck FIRST (network.rs:220)
      append_split_chunks(img, W, n, kHidden, kDirEnc, n);                        // then PE4(dir) rows 256..282
    } else {
      for (int c = 0; c < 4; ++c) append_split_chunks(img, W, n, 64 * c, 64, n);
    }
  }
}

cudaError_t launch_mlp_tc_x3(const TcWeights& w, const SampleSrc& src, const SampleDst& dst, int sm_count,
                             cudaStream_t s) {
  if (src.B == 0) return cudaSuccess;  // with a device-resident count, B is the upper bound of it
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_units = (n_tiles + 1) / 2;  // a CTA pair works on two tiles at a time
  const int64_t max_pairs = sm_count / 2;
  const int grid = 2 * (int)(n_units < max_pairs ? n_units : max_pairs);
  if (src.index != nullptr) {
    if (src.mode != 1 || src.count == nullptr) return cudaErrorInvalidValue;
    mlp_tc_x3_kernel<true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst);
  } else {
    mlp_tc_x3_kernel<false><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst);
  }
  return cudaGetLastError();
}

cudaError_t init_mlp_tc_x3() {  // once per process and device, at context creation
  cudaError_t e = cudaFuncSetAttribute(mlp_tc_x3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(mlp_tc_x3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
}

}  // namespace nerf
