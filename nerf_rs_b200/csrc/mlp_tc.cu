// mlp_tc.cu — Network::forward_batch (src/network.rs:197-237) as ONE fused persistent tcgen05 kernel.
//
// A cluster of two CTAs (one CTA per SM, the two SMs of a TPC) works as a pair with tcgen05.mma.cta_group::2:
// every MMA is M = 256 (128 samples on the TMEM lanes of each CTA) x N = 256 x K = 16, kind::f16, f32 accumulate in
// TMEM.  A pair processes "quads" of four 128-sample tiles as two slots (slot X: TMEM columns 0..255, slot Y:
// 256..511); each slot spans both CTAs.
//   - A (activations) never leaves the SM: the epilogue reads D with tcgen05.ld, applies ReLU, rounds to fp16/bf16
//     and writes the next layer's A operand straight into its own CTA's shared memory in the K-major
//     SWIZZLE_128B layout the MMA descriptor expects.
//   - B (weights) is split by N between the two CTAs: each CTA streams only its [N/2 x 64] half of every chunk
//     (16 KiB) from L2 with 1-D bulk TMA copies into a four-deep ring.  Both slots consume a chunk before it is
//     released, so L2 -> SMEM traffic is one weight pass per four tiles.
//   - Bias never goes through the tensor core: the epilogue adds it in f32 to the accumulator values it has just
//     read, from the kernel's constant bank (TcConsts is a __grid_constant__ parameter), before ReLU and rounding.
//   - The issue order is static: slot X's layer l, then slot Y's layer l, then X's layer l + 1, ...  The single
//     MMA thread blocks on one barrier at a time and runs ahead of the tensor pipe, so while slot X's epilogue
//     drains layer l the pipe is busy with Y's layer l (already queued) and vice versa.  A 256-wide layer is four
//     chunks = the whole ring; the only five-chunk layer (dense5) is issued X0-3, Y0, X4, Y1-4.
//   - concat_rows(h0, h4) (network.rs:210) is split-K: the PE tile (K = 64) and h4 (K = 256) accumulate into
//     the same D.  concat_rows(bottleneck, PE4(dir)) (network.rs:220) likewise (K = 256 + 32).
//   - sigma (256 -> 1) and rgb (128 -> 3) heads are f32 dot products in the epilogue on the f32 accumulators.
// Warp roles (both CTAs): warps 0..15 = epilogue (warp % 4 = TMEM lane quadrant, warp / 4 = column quarter; all
// sixteen drain slot X's and slot Y's accumulators alternately), warp 16 = TMA producer of the CTA's weight halves,
// warp 17 = MMA issuer (leader CTA only, one elected lane), warp 18 = TMEM allocator, warp 19 = relay ("my half of
// chunk g has landed" -> leader).  The control warps have the highest warp ids (measured placement-neutral, DESIGN.md
// 4.1); setmaxnreg gives them 32 registers and the epilogue warps 112.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "mlp_tc.h"
#include "tc_ptx.cuh"

namespace nerf {
namespace {

constexpr int kThreads = 640;  // 16 epilogue warps + 4 control warps
// The control warps sit above the epilogue warps (one setmaxnreg warp group of their own); their position among the
// warp ids measured neutral for issue latency (scripts/sched_probe.cu).
constexpr int kEpiWarps = 16, kWarpTma = 16, kWarpMma = 17, kWarpAlloc = 18, kWarpRelay = 19;
constexpr int kStages = 4;
constexpr uint32_t kStageBytes = 128 * 128;      // this CTA's N-half of a chunk: [128 x 64] 16-bit
constexpr uint32_t kKBlockBytes = kTileM * 128;  // one [128 x 64] K-block of an A tile
constexpr uint32_t kActBytes = 4 * kKBlockBytes; // [128 x 256]
constexpr uint32_t kAuxBytes = kKBlockBytes;     // [128 x 64]: PE10(p) for layers 0/5, later PE4(dir)
constexpr uint32_t kOffAct = 0;
constexpr uint32_t kOffAux = kOffAct + 2 * kActBytes;
constexpr uint32_t kOffStage = kOffAux + 2 * kAuxBytes;
constexpr uint32_t kOffBar = kOffStage + kStages * kStageBytes;
// barrier block, 8 B each: full[4] pfull[4] empty[4] acc[2] act[2] | tmem slot | flag[2] | flag words [2][8]
constexpr uint32_t kBarFull = 0, kBarPFull = 32, kBarEmpty = 64, kBarAcc = 96, kBarAct = 112, kTmemSlotOff = 128;
// kSkipEmpty: "does slot s need its colour branch" - one word per (CTA, row quadrant), written with st.async into the
// leader CTA, whose flag[s] barrier counts the 8 x 4 bytes
constexpr uint32_t kBarFlag = 136, kFlagWords = 160, kFlagBytes = 8 * 4;
constexpr uint32_t kSmemBytes = kOffBar + 256;  // 224.25 KiB
static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
constexpr int kLead = 4;  // chunks of a layer slot X issues before slot Y starts the same layer
constexpr int kEpiRegs = 112, kCtlRegs = 32;  // 512 * 112 + 128 * 32 = 61440 <= 65536 registers per SM

// Layer table: chunks per layer in consumption order.
constexpr int kNumLayers = 10;  // dense0..7, bottleneck, viewdirs
// A ring chunk is one [N/2 x 64] weight half per CTA (16 KiB at N = 256).  dense0 is the PE chunk alone, dense5 the
// PE chunk + four K-blocks of h4; viewdirs (N = 128, 8 KiB halves) packs two K-blocks of the bottleneck per ring
// stage and ends with the PE4(dir) chunk (K = 32).
__host__ __device__ constexpr int layer_chunks(int l) { return l == 0 ? 1 : (l == 5 ? 5 : (l == 9 ? 3 : 4)); }
__host__ __device__ constexpr int layer_n(int l) { return l == 9 ? 128 : 256; }



// Developer timeline (NERF_B200_TRACE=<file>): CTA 0 records (tag, clock64) pairs per role; null in production.
constexpr int kTraceCap = 2048;
struct Tracer {
  unsigned long long* p;
  int n;
  __device__ __forceinline__ void rec(uint32_t tag) {
    if (p != nullptr && n < kTraceCap) {
      p[2 * n] = tag;
      p[2 * n + 1] = (unsigned long long)clock64();
      ++n;
    }
  }
};
// ------------------------------------------------------------------------------------------ epilogue pieces
// All 16 epilogue warps drain whichever slot's accumulator completes next (X and Y alternate): warp (quad, colq)
// owns TMEM lanes 32*quad..+31 (one sample row per thread) and columns 64*colq..+63 of both slots = K-block colq
// of the next layer's A tile, as two 32-column blocks.
// One 32-column block: f32 bias, optional sigma partial on the unrounded f32 values, optional ReLU + 16-bit store.
// `bias` and `walpha` point into the kernel's constant bank (warp-uniform operands, no loads).
template <bool kFp16, bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_block(uint32_t (&v)[32], const float* __restrict__ bias,
                                          const float* __restrict__ walpha, float& sig, uint8_t* kblock_row,
                                          const uint32_t (&swz)[8], int chunk0) {
#ifndef NERF_EXP_NOBIAS
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + bias[j]);
#endif
  if constexpr (kSigma) {
#pragma unroll
    for (int j = 0; j < 32; ++j) sig = fmaf(fmaxf(__uint_as_float(v[j]), 0.0f), walpha[j], sig);
  }
  if constexpr (kStore) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint4 q;
      q.x = pack2<kFp16, kRelu>(__uint_as_float(v[i * 8 + 0]), __uint_as_float(v[i * 8 + 1]));
      q.y = pack2<kFp16, kRelu>(__uint_as_float(v[i * 8 + 2]), __uint_as_float(v[i * 8 + 3]));
      q.z = pack2<kFp16, kRelu>(__uint_as_float(v[i * 8 + 4]), __uint_as_float(v[i * 8 + 5]));
      q.w = pack2<kFp16, kRelu>(__uint_as_float(v[i * 8 + 6]), __uint_as_float(v[i * 8 + 7]));
#ifdef NERF_EXP_NOSTS
      if (q.x == 0x12345678u)
#endif
      *reinterpret_cast<uint4*>(kblock_row + swz[chunk0 + i]) = q;
    }
  }
}

// The warp's 64 columns of one slot-layer; bias / walpha already point at the warp's first column.
template <bool kFp16, bool kRelu, bool kSigma, bool kStore>
__device__ __forceinline__ void epi_cols(Tracer& tr, uint32_t taddr, const float* __restrict__ bias, const float* __restrict__ walpha,
                                         float& sig, uint8_t* kblock_row, const uint32_t (&swz)[8]) {
  uint32_t va[32], vb[32];
#ifdef NERF_EXP_NOLDTM
#pragma unroll
  for (int j = 0; j < 32; ++j) va[j] = vb[j] = taddr + j;
#else
  tc_ld32(taddr, va);
  tc_ld32(taddr + 32, vb);
  tc_wait_ld();
#endif
  epi_block<kFp16, kRelu, kSigma, kStore>(va, bias, walpha, sig, kblock_row, swz, 0);
  epi_block<kFp16, kRelu, kSigma, kStore>(vb, bias + 32, walpha + 32, sig, kblock_row, swz, 4);
}

// positional_encoding_* (network.rs:263-330): columns [kCol0, kCol0 + 16) of one sample's encoding with kOctaves
// octaves, as two 16-byte chunks of row `row` of an aux tile.  Column layout: 0..2 = v, 3 + 6o + a = sin(2^o v_a),
// 6 + 6o + a = cos(2^o v_a), zero padding after 3 + 6 * kOctaves.  The lowest octave the 16 columns touch is
// evaluated with fast_sincos; the (at most three) following ones by angle doubling, sin 2x = 2 sin x cos x and
// cos 2x = 1 - 2 sin^2 x, which adds ~1e-6 absolute error - far below the operand's 16-bit rounding.
template <bool kFp16, int kCol0, int kOctaves>
__device__ __forceinline__ void write_pe16(uint8_t* aux_row, uint32_t r8, const float (&v)[3]) {
  constexpr int kEnd = 3 + 6 * kOctaves;
  constexpr int kLo = kCol0 < 3 ? 0 : (kCol0 - 3) / 6;
  constexpr int kHiCol = (kCol0 + 15 < kEnd ? kCol0 + 15 : kEnd - 1);
  constexpr int kHi = kHiCol < 3 ? -1 : (kHiCol - 3) / 6;
  constexpr int kN = kHi >= kLo ? kHi - kLo + 1 : 0;
  float sn[kN > 0 ? kN : 1][3], cs[kN > 0 ? kN : 1][3];
  if constexpr (kN > 0) {
#pragma unroll
    for (int a = 0; a < 3; ++a) fast_sincos((float)(1 << kLo) * v[a], sn[0][a], cs[0][a]);  // f doubles per octave (network.rs:287)
#pragma unroll
    for (int i = 1; i < kN; ++i) {
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        sn[i][a] = 2.0f * sn[i - 1][a] * cs[i - 1][a];
        cs[i][a] = fmaf(-2.0f * sn[i - 1][a], sn[i - 1][a], 1.0f);
      }
    }
  }
  float out[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const int col = kCol0 + j;
    if (col < 3) out[j] = v[col];
    else if (col >= kEnd) out[j] = 0.0f;
    else {
      const int o = (col - 3) / 6 - kLo, k = (col - 3) % 6;
      out[j] = k < 3 ? sn[o][k] : cs[o][k - 3];
    }
  }
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint4 q;
    q.x = pack2<kFp16, false>(out[8 * h + 0], out[8 * h + 1]);
    q.y = pack2<kFp16, false>(out[8 * h + 2], out[8 * h + 3]);
    q.z = pack2<kFp16, false>(out[8 * h + 4], out[8 * h + 5]);
    q.w = pack2<kFp16, false>(out[8 * h + 6], out[8 * h + 7]);
    *reinterpret_cast<uint4*>(aux_row + ((((uint32_t)(kCol0 / 8 + h)) ^ r8) << 4)) = q;
  }
}
// Register rebalancing (setmaxnreg): the launch gives each of the 640 threads 96 registers; the control warp group
// (warps 16..19) shrinks to kCtlRegs and the four epilogue warp groups grow to kEpiRegs, which is what the epilogue
// needs to keep two 32-column accumulator blocks plus the prefetched inputs in registers.  A spill here is an L2
// round trip (227 KB of the SM's array is shared memory, so local memory does not stay in L1).  Each role's branch
// executes its own instruction: ptxas takes a region's register limit from the setmaxnreg that dominates it, and
// after a merge of branches with different limits it assumes the smallest.
__device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kCtlRegs)); }
__device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kEpiRegs)); }
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}


// ------------------------------------------------------------------------------------------ the kernel
// kRawSigma: store the density head's pre-activation alpha(h8) instead of relu(alpha(h8)) (first pass of the two-pass
// coarse densities).  A template parameter, not a run-time flag: reading one more kernel parameter in the epilogue
// made ptxas spill two registers of the 112-register epilogue warps.
// kSkipEmpty (render launches with colour): a sample with sigma = 0 has alpha = 0 and weight exactly 0 in
// compute_weights / integrate_ray (lib.rs:176-195,250-283), so its colour cannot reach the pixel, and 73 % of the rays
// of the lego frame have no density at all.  The MMA issuer skips the bottleneck and viewdirs MMAs of a slot in which no
// row of either CTA has density; the weight chunks still stream and are released unread, and the epilogue drains the
// stale accumulator (finite values: the colour stored for such samples is arbitrary but finite, and never used).
// forward_batch never skips: the reference returns the colour of every sample.
// kStoreH8 (first phase of the two-phase fine pass): the launch ends at the density head like a density-only launch,
// but dense7's activations are stored to the slot's A tile as usual and, once a row's density is known, quarter 0 of
// the row block copies the rows WITH density to global memory (512 B each, compacted through one atomicAdd per 32 rows)
// and stores rgb = 0 for the rows without - the colour branch then runs over the compacted rows only
// (mlp_colour_kernel below), instead of over every slot that holds at least one such row.
template <bool kFp16, bool kRawSigma, bool kSkipEmpty, bool kStoreH8 = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_tc_kernel(const __grid_constant__ TcConsts C, const uint8_t* __restrict__ chunks, const SampleSrc src,
              const SampleDst dst, unsigned long long* __restrict__ trace) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0u) __trap();  // SWIZZLE_128B atoms need 1 KiB alignment; the kernel has no static shared memory
  const uint32_t bar0 = base + kOffBar;
  const uint32_t bar_full = bar0 + kBarFull, bar_pfull = bar0 + kBarPFull, bar_empty = bar0 + kBarEmpty,
                 bar_acc = bar0 + kBarAcc, bar_act = bar0 + kBarAct;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kOffBar + kTmemSlotOff);

  // Broadcast through a shuffle so that the compiler knows the warp index is warp-uniform: role branches then
  // are uniform branches and the MMA issuer's descriptors live in uniform registers (no R2UR per MMA).
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();       // 0 = leader (issues the pair's MMAs), 1 = peer
  const int64_t pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const bool want_rgb = !kStoreH8 && dst.rgb != nullptr;
  const int n_layers = want_rgb ? kNumLayers : 8;
  // kRawSigma launches (coarse first pass) cover a slice of every ray's samples, optionally of listed rays only, whose
  // number the previous stage left on the device
  const int64_t B = (kRawSigma && src.count != nullptr) ? (int64_t)__ldg(src.count) * src.sub_n : src.B;
  const int64_t n_tiles = (B + kTileM - 1) / kTileM;
  const int64_t n_quads = (n_tiles + 3) / 4;  // quad u = tiles 4u..4u+3: slot s of CTA r owns tile 4u + 2s + r

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar_full + 8 * s, 1);   // this CTA's bulk copy (expect_tx)
      mbar_init(bar_pfull + 8 * s, 2);  // leader only: both CTAs' relays
      mbar_init(bar_empty + 8 * s, 2);  // slot X's and slot Y's commit
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar_acc + 8 * s, 1);
      mbar_init(bar_act + 8 * s, 32);   // leader only: one elected arrive per epilogue warp, both CTAs
      mbar_init(bar0 + kBarFlag + 8 * s, 1);  // leader only: the MMA issuer's own arrive.expect_tx; the stores complete the bytes
    }
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

  int n_chunks_quad = 0;
  for (int l = 0; l < n_layers; ++l) n_chunks_quad += layer_chunks(l);

  if (warp == kWarpTma) {
    reg_dec();
    // ===================================================== TMA producer: this CTA's half of every weight chunk
    Tracer tr{(trace != nullptr && blockIdx.x == 0 && lane == 0) ? trace + 1 * 2 * kTraceCap : nullptr, 0};
    uint32_t stage = 0, phase = 0;
    for (int64_t quad = pair_id; quad < n_quads; quad += n_pairs) {
      const uint8_t* src_ptr = chunks;
      for (int l = 0; l < n_layers; ++l) {
        const uint32_t bytes = (uint32_t)layer_n(l) * 64u;  // [N/2 x 64] 16-bit
        for (int c = 0; c < layer_chunks(l); ++c) {
          const uint32_t ncopy = (l == 9 && c < 2) ? 2u : 1u;  // viewdirs: two 8 KiB K-blocks share a stage
          tr.rec(0x100 | l);
          mbar_wait(bar_empty + 8 * stage, phase ^ 1u);
          tr.rec(0x200 | l);
          if (elect_one()) {
            mbar_arrive_expect_tx(bar_full + 8 * stage, bytes * ncopy);
            const uint32_t dst_s = base + kOffStage + stage * kStageBytes;
            tma_bulk_g2s(dst_s, src_ptr + rank * bytes, bytes, bar_full + 8 * stage);
            if (ncopy == 2) tma_bulk_g2s(dst_s + bytes, src_ptr + (2 + rank) * bytes, bytes, bar_full + 8 * stage);
          }
          __syncwarp();
          src_ptr += 2 * bytes * ncopy;
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == kWarpAlloc) {
    reg_dec();  // all four warps of the group must execute it
  } else if (warp == kWarpRelay) {
    reg_dec();
    // ===================================================== relay: tell the leader that this CTA's half has landed
    Tracer tr{(trace != nullptr && blockIdx.x == 0 && lane == 0) ? trace + 6 * 2 * kTraceCap : nullptr, 0};
    uint32_t stage = 0, phase = 0;
    for (int64_t quad = pair_id; quad < n_quads; quad += n_pairs) {
      for (int c = 0; c < n_chunks_quad; ++c) {
        mbar_wait(bar_full + 8 * stage, phase);
        tr.rec(0x300 | stage);
        if (elect_one()) mbar_arrive_cluster(bar_pfull + 8 * stage, 0u);
        __syncwarp();
        if (tr.p != nullptr) {  // timeline only: when did the peer's half land too
          mbar_wait(bar_pfull + 8 * stage, phase);
          tr.rec(0x400 | stage);
        }
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == kWarpMma) {
    reg_dec();
    // ===================================================== MMA issuer (leader CTA): one thread drives both tensor cores
    if (rank == 0) {
      Tracer tr{(trace != nullptr && blockIdx.x == 0 && lane == 0) ? trace : nullptr, 0};
      const uint64_t act_desc0 = make_desc(base + kOffAct), aux_desc0 = make_desc(base + kOffAux);
      const uint64_t stage_desc0 = make_desc(base + kOffStage);
      uint32_t g0 = 0;       // ring index of the current layer's first chunk (the same for both slots)
      // bit s: parity of slot s's "layer input ready" barrier; kSkipEmpty: bit 2 + s = slot s runs bottleneck + viewdirs
      // in this quad, bit 4 + s = parity of slot s's colour-flag barrier (one word: uniform registers are scarce here)
      uint32_t act_par = 0;
      for (int64_t quad = pair_id; quad < n_quads; quad += n_pairs) {
        if (kSkipEmpty) act_par |= 12u;
#pragma unroll 1
        for (int l = 0; l < n_layers; ++l) {
          const int nch = layer_chunks(l);
          const uint32_t idesc = make_idesc(kFp16 ? 1 : 0, layer_n(l));
          // the chunk that reads the aux tile (PE10(p) for dense0 / dense5, PE4(dir) last for viewdirs); the others
          // read K-block c - kb_off of the activations
          const int aux_c = (l == 0 || l == 5) ? 0 : (l == 9 ? 2 : -1);
          const int kb_off = (l == 5) ? 1 : 0;
          const int kpc = (l == 9) ? 2 : 1;  // K-blocks per ring chunk
          // Static interleave (X leads Y by kLead chunks): X[0, kLead), Y[0, nch - kLead), X[kLead, nch), Y[rest]:
          // a stage is released early enough for the next layer's load to land, and each slot's epilogue is covered
          // by the other slot's queued MMAs.
          const int nseg = (nch > kLead) ? 4 : 2;
#pragma unroll 1
          for (int seg = 0; seg < nseg; ++seg) {
            const int s = seg & 1;
            int cb = 0, ce = nch;
            if (nseg == 4) {
              const int split = s ? nch - kLead : kLead;
              if (seg < 2) ce = split; else cb = split;
            }
            const uint32_t d_tmem = tmem_base + (uint32_t)s * 256u;
            const uint64_t auxd = aux_desc0 + (uint32_t)s * (kAuxBytes >> 4);
            const uint64_t actd = act_desc0 + (uint32_t)s * (kActBytes >> 4) - (uint32_t)kb_off * (kKBlockBytes >> 4);
#pragma unroll 1
            for (int c = cb; c < ce; ++c) {
              const uint32_t g = g0 + (uint32_t)c, stage = g & (kStages - 1);
              if (c == 0) {  // layer input written and accumulator drained (both CTAs)
                mbar_wait(bar_act + 8 * s, (act_par >> s) & 1u);
                act_par ^= 1u << s;
                if (kSkipEmpty && l == 8) {  // the eight (CTA, row quadrant) verdicts on slot s's densities
                  const uint32_t bar_flag = base + kOffBar + kBarFlag + 8u * (uint32_t)s;
                  if (elect_one()) mbar_arrive_expect_tx(bar_flag, kFlagBytes);
                  __syncwarp();
                  mbar_wait_cluster(bar_flag, (act_par >> (4 + s)) & 1u);
                  act_par ^= 16u << s;
                  uint32_t any = 0;
#pragma unroll
                  for (uint32_t w = 0; w < 8; ++w)
                    any |= ld_shared_u32(base + kOffBar + kFlagWords + (uint32_t)s * kFlagBytes + 4u * w);
                  if (any == 0u) act_par &= ~(4u << s);
                }
              }
              const bool run = !kSkipEmpty || ((act_par >> (2 + s)) & 1u) != 0u;
              if (s == 0) mbar_wait(bar_pfull + 8 * stage, (g / kStages) & 1u);  // X is every chunk's first consumer
              tc_fence_after();
              tr.rec(0x100 | (l << 4) | c | (s << 12));
              const uint64_t bd = stage_desc0 + stage * (kStageBytes >> 4);
              if (elect_one()) {
                if (!run) {
                  // nothing to multiply for this slot: the commits below still release the chunk and the epilogue
                } else if (c == aux_c) {
                  tc_mma2_f16(d_tmem, auxd, bd, idesc, c != 0 ? 1u : 0u);
                  tc_mma2_f16(d_tmem, auxd + 2, bd + 2, idesc, 1u);
                  if (l != 9) {  // PE4(dir) is K = 32: aux columns 32..63 hold f32 scratch by then
                    tc_mma2_f16(d_tmem, auxd + 4, bd + 4, idesc, 1u);
                    tc_mma2_f16(d_tmem, auxd + 6, bd + 6, idesc, 1u);
                  }
                } else {
#pragma unroll 1
                  for (int j = 0; j < kpc; ++j) {
                    const uint64_t ad = actd + (uint32_t)(c * kpc + j) * (kKBlockBytes >> 4);
                    const uint64_t bj = bd + (uint32_t)j * (8192u >> 4);
                    tc_mma2_f16(d_tmem, ad, bj, idesc, (c | j) != 0 ? 1u : 0u);
                    tc_mma2_f16(d_tmem, ad + 2, bj + 2, idesc, 1u);
                    tc_mma2_f16(d_tmem, ad + 4, bj + 4, idesc, 1u);
                    tc_mma2_f16(d_tmem, ad + 6, bj + 6, idesc, 1u);
                  }
                }
                if (c == nch - 1) tc_commit2(bar_acc + 8 * s);  // this layer's accumulator complete -> both epilogues
                tc_commit2(bar_empty + 8 * stage);               // this slot is done with the chunk (both CTAs' rings)
              }
              __syncwarp();
            }
          }
          g0 += (uint32_t)nch;
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
    const uint32_t taddr0 = tmem_base + ((quad * 32u) << 16) + (uint32_t)colq * 64u;  // + 256 * slot
    // this row of K-block colq of slot 0's A tile; slot 1's is kActBytes further
    uint8_t* const kblock_row0 = smem + kOffAct + (uint32_t)colq * kKBlockBytes + (row >> 3) * 1024u + r8 * 128u;
    uint32_t swz[8];  // byte offset of 16-byte chunk c inside the row's 128 bytes (SWIZZLE_128B)
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c) swz[c] = (c ^ r8) << 4;
    // f32 scratch for combining the four column quarters of a row: quarter k >= 1 owns the 16-byte chunk at aux
    // column 32 + 8k = {rgb partials, sigma partial}; aux columns 32..63 are free whenever it is used
    const uint32_t scratch_off = sw128_offset(row, 32u + 8u * (uint32_t)colq);
    const uint32_t row_bar = 1u + quad;  // named barrier of the 4 warps (128 threads) that share a row
    uint32_t acc_par = 0;                // bit s: parity of slot s's "accumulator complete" barrier
    Tracer tr{(trace != nullptr && blockIdx.x == 0 && quad == 0 && lane == 0) ? trace + (2 + colq) * 2 * kTraceCap : nullptr, 0};

    auto acc_wait = [&](int s) {
      mbar_wait(bar_acc + 8 * s, (acc_par >> s) & 1u);
      acc_par ^= 1u << s;
      tc_fence_after();
    };
    auto act_ready = [&](int s) {  // this warp's part of slot s's next A operand is in shared memory, its columns drained
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (elect_one()) mbar_arrive_cluster(bar_act + 8 * s, 0u);  // one arrive per warp on the leader's barrier (elected:
      //                                                               keeping `lane` alive for this costs a register)
    };

    // The next quad's inputs are fetched two layers before they are needed: {p.xyz, -} for explicit points,
    // {dir.xyz, t} for samples along rays.  Tiles past the end run the same barrier protocol on zeros.
    float4 pre[2];
    // kRawSigma: sample id (ray * S + s) of the prefetched row, and the ray-list entry of the row's slot one quad
    // further on (loaded a quad early so that the dependent loads of dirs / t never wait for it inside a drain)
    int32_t g_pre[2] = {0, 0}, idx_pre[2] = {0, 0};
    auto load_index = [&](int64_t qd) {
      if (kRawSigma && src.index != nullptr) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const int64_t gi = (4 * qd + 2 * s + (int64_t)rank) * kTileM + row;
          if (gi < B) idx_pre[s] = __ldg(src.index + gi / src.sub_n);
        }
      }
    };
    auto prefetch = [&](int64_t qd) {
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const int64_t gi = (4 * qd + 2 * s + (int64_t)rank) * kTileM + row;
        pre[s] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gi < B) {
          if (kRawSigma) {
            const int32_t r = (int32_t)(gi / src.sub_n);
            const int32_t ray = src.index != nullptr ? idx_pre[s] : r;
            g_pre[s] = ray * src.S + src.sub_s0 + (int32_t)(gi - (int64_t)r * src.sub_n);
            pre[s] = make_float4(src.dirs[ray * 3 + 0], src.dirs[ray * 3 + 1], src.dirs[ray * 3 + 2], src.t[g_pre[s]]);
          } else if (src.mode == 0) {
            pre[s] = make_float4(src.points[gi], src.points[B + gi], src.points[2 * B + gi], 0.f);
          } else {
            const int64_t ray = gi / src.S;
            pre[s] = make_float4(src.dirs[ray * 3 + 0], src.dirs[ray * 3 + 1], src.dirs[ray * 3 + 2], src.t[gi]);
          }
        }
      }
      load_index(qd + n_pairs);
    };
    load_index(pair_id);
    prefetch(pair_id);
    for (int64_t qd = pair_id; qd < n_quads; qd += n_pairs) {
      int64_t g[2];
      bool valid[2];
      float dir[2][3];
      tr.rec(0x500);
#pragma unroll
      for (int s = 0; s < 2; ++s) {
        const int64_t tile = 4 * qd + 2 * s + (int64_t)rank;
        g[s] = tile * kTileM + row;
        valid[s] = g[s] < B;
        if (kRawSigma) g[s] = g_pre[s];  // where this row's density goes: ray * S + sample
        float p[3] = {pre[s].x, pre[s].y, pre[s].z};
        dir[s][0] = dir[s][1] = dir[s][2] = 0.f;
        if (src.mode == 0) {
          if (valid[s]) {  // not needed before the bottleneck layer: the load stays in flight
            dir[s][0] = src.dirs[g[s] * 3 + 0];
            dir[s][1] = src.dirs[g[s] * 3 + 1];
            dir[s][2] = src.dirs[g[s] * 3 + 2];
          }
        } else {
          dir[s][0] = pre[s].x;
          dir[s][1] = pre[s].y;
          dir[s][2] = pre[s].z;
          if (valid[s]) {
            p[0] = __fadd_rn(src.origin[0], __fmul_rn(dir[s][0], pre[s].w));  // origin + dir_hat * t (lib.rs:396,436)
            p[1] = __fadd_rn(src.origin[1], __fmul_rn(dir[s][1], pre[s].w));
            p[2] = __fadd_rn(src.origin[2], __fmul_rn(dir[s][2], pre[s].w));
          }
        }
        // PE10(p): layer-0 / skip operand (network.rs:204); each column quarter writes its 16 of the 64 columns
        uint8_t* aux_row = smem + kOffAux + s * kAuxBytes + (row >> 3) * 1024u + r8 * 128u;
        if (colq == 0) write_pe16<kFp16, 0, kPosOctaves>(aux_row, r8, p);
        else if (colq == 1) write_pe16<kFp16, 16, kPosOctaves>(aux_row, r8, p);
        else if (colq == 2) write_pe16<kFp16, 32, kPosOctaves>(aux_row, r8, p);
        else write_pe16<kFp16, 48, kPosOctaves>(aux_row, r8, p);
        act_ready(s);
      }
      tr.rec(0x400);

      float sig[2];
      sig[0] = sig[1] = colq == 0 ? C.b_alpha : 0.0f;
#pragma unroll 1
      for (int l = 0; l < 7; ++l) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          tr.rec(0x100 | l | (s << 12));
          acc_wait(s);
          tr.rec(0x200 | l | (s << 12));
          epi_cols<kFp16, true, false, true>(tr, taddr0 + 256u * s, C.bias[l] + colq * 64, C.w_alpha, sig[s],
                                             kblock_row0 + s * kActBytes, swz);
          act_ready(s);
          tr.rec(0x300 | l | (s << 12));
        }
      }
      if (qd + n_pairs < n_quads) prefetch(qd + n_pairs);
#pragma unroll
      for (int s = 0; s < 2; ++s) {  // dense7 -> h8; sigma = relu(alpha(h8)) (network.rs:216)
        acc_wait(s);
        if (want_rgb) {
          epi_cols<kFp16, true, true, true>(tr, taddr0 + 256u * s, C.bias[7] + colq * 64, C.w_alpha + colq * 64, sig[s],
                                            kblock_row0 + s * kActBytes, swz);
          act_ready(s);
        } else {
          // kStoreH8: h8 goes to the (free) A tile so that rows with density can be copied out below
          epi_cols<kFp16, true, true, kStoreH8>(tr, taddr0 + 256u * s, C.bias[7] + colq * 64, C.w_alpha + colq * 64,
                                                sig[s], kblock_row0 + s * kActBytes, swz);
          tc_fence_before();
        }
        float* scratch = reinterpret_cast<float*>(smem + kOffAux + s * kAuxBytes + scratch_off);
        if (kSkipEmpty || colq != 0) scratch[3] = sig[s];
      }
      named_bar_sync(row_bar, 128);
      bool need[2] = {true, true};  // kSkipEmpty: some row of this warp has density in slot s (warp-uniform)
      if (kSkipEmpty || colq == 0) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {
          const uint8_t* aux = smem + kOffAux + s * kAuxBytes;
          float v = kSkipEmpty ? reinterpret_cast<const float*>(aux + sw128_offset(row, 32u))[3] : sig[s];
#pragma unroll
          for (uint32_t k = 1; k < 4; ++k) v += reinterpret_cast<const float*>(aux + sw128_offset(row, 32u + 8u * k))[3];
          if (colq == 0 && valid[s]) dst.sigma[g[s]] = kRawSigma ? v : fmaxf(v, 0.0f);
          if constexpr (kStoreH8) {  // colq == 0 here (kSkipEmpty is off in this instantiation)
            const bool has = valid[s] && v > 0.0f;
            const uint32_t m = __ballot_sync(0xffffffffu, has);
            if (valid[s] && !has) {  // weight exactly 0 in integrate_ray: the colour is never looked at; keep it finite
              dst.rgb[g[s] * 3 + 0] = 0.0f;
              dst.rgb[g[s] * 3 + 1] = 0.0f;
              dst.rgb[g[s] * 3 + 2] = 0.0f;
            }
            if (m != 0u) {
              uint32_t pos0 = 0;
              if (lane == 0) pos0 = atomicAdd(dst.h8_count, (uint32_t)__popc(m));
              pos0 = __shfl_sync(0xffffffffu, pos0, 0);
              if (has)  // what the colour branch needs besides h8: the ray's direction and where the colour goes
                dst.h8_meta[pos0 + __popc(m & ((1u << lane) - 1u))] =
                    make_float4(dir[s][0], dir[s][1], dir[s][2], __int_as_float((int32_t)g[s]));
              // A row is 4 K-blocks x 8 chunks of 16 bytes = one chunk per lane.  Stored row p goes to row p % 128 of tile
              // p / 128 of dst.h8, each tile being the exact SWIZZLE_128B shared-memory image of a [128 x 256] A operand
              // (64 KiB), so that the colour kernel brings a tile in with bulk copies and no register staging.
              const uint8_t* act = smem + kOffAct + s * kActBytes + (uint32_t)(lane >> 3) * kKBlockBytes;
              uint8_t* const out0 = reinterpret_cast<uint8_t*>(dst.h8) + (uint32_t)(lane >> 3) * kKBlockBytes;
              uint32_t p = pos0;
              for (uint32_t mm = m; mm != 0u; mm &= mm - 1u, ++p) {
                const uint32_t rr = quad * 32u + (uint32_t)(__ffs(mm) - 1);
                const uint4 val = *reinterpret_cast<const uint4*>(act + (rr >> 3) * 1024u + (rr & 7u) * 128u +
                                                                  ((((uint32_t)lane & 7u) ^ (rr & 7u)) << 4));
                *reinterpret_cast<uint4*>(out0 + (size_t)(p >> 7) * kActBytes + ((p & 127u) >> 3) * 1024u + (p & 7u) * 128u +
                                          ((((uint32_t)lane & 7u) ^ (p & 7u)) << 4)) = val;
              }
            }
          }
          if constexpr (kSkipEmpty) {
            // does any of this warp's 32 rows need slot s's colour?  Every quarter of a row block reaches the same
            // verdict; quarter 0 sends it to the leader, one word per (CTA, row quadrant)
            need[s] = __any_sync(0xffffffffu, valid[s] && v > 0.0f);
            if (colq == 0 && elect_one())
              st_async_u32(base + kOffBar + kFlagWords + (uint32_t)s * kFlagBytes + 4u * (rank * 4u + quad),
                           need[s] ? 1u : 0u, base + kOffBar + kBarFlag + 8u * (uint32_t)s, 0u);
          }
        }
      }
      if (want_rgb) {
#pragma unroll
        for (int s = 0; s < 2; ++s) {  // bottleneck, ActivationType::None (network.rs:218)
          acc_wait(s);
          if (!kSkipEmpty || need[s]) {
            epi_cols<kFp16, false, false, true>(tr, taddr0 + 256u * s, C.bias[8] + colq * 64, C.w_alpha, sig[s],
                                                kblock_row0 + s * kActBytes, swz);
            // PE4(dir) replaces PE10(p) in the aux tile (columns 0..31): the skip MMA that read it finished layers ago
            uint8_t* aux_row = smem + kOffAux + s * kAuxBytes + (row >> 3) * 1024u + r8 * 128u;
            if (colq == 0) write_pe16<kFp16, 0, kDirOctaves>(aux_row, r8, dir[s]);
            else if (colq == 1) write_pe16<kFp16, 16, kDirOctaves>(aux_row, r8, dir[s]);  // columns 27..31: K padding
          }
          act_ready(s);
        }
#pragma unroll
        for (int s = 0; s < 2; ++s) {  // viewdirs (ReLU) then the f32 rgb head (network.rs:222-223): 32 columns per warp
          acc_wait(s);
          if (kSkipEmpty && !need[s]) {  // nothing was multiplied for rows without density (or nothing worth reading)
            tc_fence_before();
            continue;
          }
          uint32_t v[32];
          tc_ld32(tmem_base + ((quad * 32u) << 16) + 256u * s + 32u * (uint32_t)colq, v);
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
          float* scratch = reinterpret_cast<float*>(smem + kOffAux + s * kAuxBytes + scratch_off);
          if (colq != 0) {
            scratch[0] = r0;
            scratch[1] = r1;
            scratch[2] = r2;
          } else {
            sig[s] = r0;  // quarter 0 keeps its partials in registers until the others have written theirs
            dir[s][0] = r1;
            dir[s][1] = r2;
          }
        }
        named_bar_sync(row_bar, 128);
        if (colq == 0) {
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            if (kSkipEmpty && !need[s]) {  // no density in these 32 rows: their colour is never looked at; keep it finite
              if (valid[s]) {
                dst.rgb[g[s] * 3 + 0] = 0.0f;
                dst.rgb[g[s] * 3 + 1] = 0.0f;
                dst.rgb[g[s] * 3 + 2] = 0.0f;
              }
              continue;
            }
            const uint8_t* aux = smem + kOffAux + s * kAuxBytes;
            float r0 = sig[s], r1 = dir[s][0], r2 = dir[s][1];
#pragma unroll
            for (uint32_t k = 1; k < 4; ++k) {
              const float* sc = reinterpret_cast<const float*>(aux + sw128_offset(row, 32u + 8u * k));
              r0 += sc[0];
              r1 += sc[1];
              r2 += sc[2];
            }
            if (valid[s]) {
              dst.rgb[g[s] * 3 + 0] = 1.0f / (1.0f + expf(-r0));  // network.rs:165
              dst.rgb[g[s] * 3 + 1] = 1.0f / (1.0f + expf(-r1));
              dst.rgb[g[s] * 3 + 2] = 1.0f / (1.0f + expf(-r2));
            }
          }
        }
      }
      named_bar_sync(row_bar, 128);  // scratch is inside the aux tiles the next quad's PE overwrites
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's shared memory and TMEM stay alive until the pair's last MMA has been consumed
  if (warp == kWarpAlloc) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}


// ------------------------------------------------------------------------------------------ colour branch alone
// Second phase of the two-phase fine pass: bottleneck (256 -> 256, no activation, network.rs:218), viewdirs
// (concat_rows(bottleneck, PE4(dir)) -> 128, ReLU, :220-222) and the f32 rgb head + sigmoid (:223) over the compacted
// rows the first phase stored (h8 = relu(dense7), 256 16-bit values per sample, as [128 x 256] tile images).  Same
// operands, same MMA sequence per layer and the same f32 reductions as the fused kernel above, so a sample's colour is
// bit-identical to the fused kernel's; rows of an MMA tile are independent, so it does not matter which samples share
// a tile.
//   A CTA pair works on 256 rows at a time (cta_group::2, M = 256).  Both layers' weight halves are resident in shared
// memory (64 + 40 KiB per CTA, bulk-copied once from the fused kernel's chunk stream).  A unit's rows come in with bulk
// copies straight into the one A tile.  The bottleneck's output does NOT go back into that tile: it stays in TENSOR
// MEMORY as the viewdirs layer's A operand (tcgen05.mma with A from TMEM: row = lane, two 16-bit elements per 32-bit
// column), which frees the A tile as soon as the bottleneck MMAs of unit u have completed - the rows of unit u + 1
// arrive while unit u's viewdirs layer runs, and the bottleneck MMAs of unit u + 1 run under unit u's rgb head.
// TMEM: bottleneck accumulator columns 0..255, viewdirs accumulator 256..383, the bottleneck's 256 16-bit outputs
// packed in columns 384..511.  PE4(dir) tiles alternate between two aux buffers.
//   How it got here (800x800 frame, ncu; every version bit-identical to the fused kernel): one A tile that the
// bottleneck was drained back into, every step of a unit after the other, rows through registers - 6.2 ms, tensor
// pipe 31 %; A from TMEM - 5.1 ms (a timeline showed the nine loads of a unit taking 1 500 - 5 700 cycles to ISSUE:
// 64 KiB per SM and unit is more than the L1 miss path sustains, coalesced or not); rows by bulk copy - 3.8 ms, tensor
// pipe 48 %, period 6 000 cycles for 3 750 of MMA.  Under the frame's power cap the last step is worth less than
// under ncu: what the cap prices is energy, and a kernel that idles costs little of it.
constexpr int kColThreads = 17 * 32;  // 16 worker warps (mapped like the fused epilogue) + 1 control warp
constexpr uint32_t kCW8Bytes = 4 * 16384, kCW9Bytes = 5 * 8192;
// byte offsets of the bottleneck and viewdirs chunks in the chunk stream (build_tc_image): 30 K-block images of 32 KiB
constexpr size_t kChunkOff8 = 30 * 32768, kChunkOff9 = kChunkOff8 + 4 * 32768;
constexpr uint32_t kTOffAct = 0, kTOffAux = kTOffAct + kActBytes, kTOffW8 = kTOffAux + 2 * kAuxBytes;
constexpr uint32_t kTOffW9 = kTOffW8 + kCW8Bytes, kTOffBar = kTOffW9 + kCW9Bytes;
constexpr uint32_t kTBarW = 0, kTBarIn8 = 8, kTBarIn9 = 16, kTBarAcc8 = 24, kTBarAcc9 = 32, kTTmemSlot = 40, kTBarRows = 48;
constexpr uint32_t kTSmemBytes = kTOffBar + 64;  // 200.06 KiB
constexpr uint32_t kTColA9 = 384;

template <bool kFp16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kColThreads, 1)
mlp_colour_kernel(const __grid_constant__ TcConsts C, const uint8_t* __restrict__ chunks,
                     const uint16_t* __restrict__ h8, const float4* __restrict__ meta,
                     const uint32_t* __restrict__ count, float* __restrict__ rgb) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  const uint32_t base = smem_u32(smem_raw);
  if ((base & 1023u) != 0u) __trap();
  const uint32_t bar0 = base + kTOffBar;
  const uint32_t bar_w = bar0 + kTBarW, bar_in8 = bar0 + kTBarIn8, bar_in9 = bar0 + kTBarIn9, bar_acc8 = bar0 + kTBarAcc8,
                 bar_acc9 = bar0 + kTBarAcc9, bar_rows = bar0 + kTBarRows;
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kTOffBar + kTTmemSlot);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int64_t pair_id = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int64_t n = (int64_t)__ldg(count);
  const int64_t n_units = (n + 2 * kTileM - 1) / (2 * kTileM);

  if (threadIdx.x == 0) {
    mbar_init(bar_w, 1);
    mbar_init(bar_in8, 32);  // leader only: one elected arrive per worker warp, both CTAs
    mbar_init(bar_in9, 32);
    mbar_init(bar_acc8, 1);
    mbar_init(bar_acc9, 1);
    mbar_init(bar_rows, 1);  // this CTA's tile of rows (bulk copies, expect_tx)
    fence_barrier_init();
  }
  if (warp == 16) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(base + kTOffBar + kTTmemSlot),
                 "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(bar_w, kCW8Bytes + kCW9Bytes);
    for (uint32_t c = 0; c < 4; ++c)
      tma_bulk_g2s(base + kTOffW8 + c * 16384u, chunks + kChunkOff8 + c * 32768u + rank * 16384u, 16384u, bar_w);
    for (uint32_t c = 0; c < 5; ++c)
      tma_bulk_g2s(base + kTOffW9 + c * 8192u, chunks + kChunkOff9 + c * 16384u + rank * 8192u, 8192u, bar_w);
  }
  mbar_wait(bar_w, 0u);
  cluster_sync_all();

  if (warp == 16) {
    // ===================================================== MMA issuer (leader CTA)
    if (rank == 0) {
      const uint64_t act_desc0 = make_desc(base + kTOffAct), aux_desc0 = make_desc(base + kTOffAux);
      const uint64_t w8_desc0 = make_desc(base + kTOffW8), w9_desc0 = make_desc(base + kTOffW9);
      const uint32_t idesc8 = make_idesc(kFp16 ? 1 : 0, 256), idesc9 = make_idesc(kFp16 ? 1 : 0, 128);
      uint32_t par = 0, buf = 0;
      for (int64_t u = pair_id; u < n_units; u += n_pairs) {
        mbar_wait(bar_in8, par);  // rows and PE4(dir) of unit u in shared memory; bottleneck accumulator drained
        tc_fence_after();
        if (elect_one()) {
#pragma unroll 1
          for (uint32_t c = 0; c < 4; ++c) {
            const uint64_t ad = act_desc0 + c * (kKBlockBytes >> 4), bd = w8_desc0 + c * (16384u >> 4);
            tc_mma2_f16(tmem_base, ad, bd, idesc8, c != 0 ? 1u : 0u);
            tc_mma2_f16(tmem_base, ad + 2, bd + 2, idesc8, 1u);
            tc_mma2_f16(tmem_base, ad + 4, bd + 4, idesc8, 1u);
            tc_mma2_f16(tmem_base, ad + 6, bd + 6, idesc8, 1u);
          }
          tc_commit2(bar_acc8);
        }
        __syncwarp();
        mbar_wait(bar_in9, par);  // the bottleneck's output of unit u is in tensor memory; viewdirs accumulator read
        tc_fence_after();
        if (elect_one()) {
          const uint32_t d1 = tmem_base + 256u, a9 = tmem_base + kTColA9;
#pragma unroll 1
          for (uint32_t c = 0; c < 4; ++c) {  // the bottleneck's four K-blocks first, PE4(dir) last (network.rs:220)
            const uint64_t bd = w9_desc0 + c * (8192u >> 4);
            tc_mma2_f16_ts(d1, a9 + c * 32u, bd, idesc9, c != 0 ? 1u : 0u);
            tc_mma2_f16_ts(d1, a9 + c * 32u + 8u, bd + 2, idesc9, 1u);
            tc_mma2_f16_ts(d1, a9 + c * 32u + 16u, bd + 4, idesc9, 1u);
            tc_mma2_f16_ts(d1, a9 + c * 32u + 24u, bd + 6, idesc9, 1u);
          }
          const uint64_t bd = w9_desc0 + 4u * (8192u >> 4), auxd = aux_desc0 + buf * (kAuxBytes >> 4);
          tc_mma2_f16(d1, auxd, bd, idesc9, 1u);  // K = 32: aux columns 32..63 hold f32 scratch
          tc_mma2_f16(d1, auxd + 2, bd + 2, idesc9, 1u);
          tc_commit2(bar_acc9);
        }
        __syncwarp();
        par ^= 1u;
        buf ^= 1u;
      }
    }
  } else {
    // ===================================================== workers: one thread per (row, column quarter)
    const int colq = warp >> 2;
    const uint32_t quad = (uint32_t)(warp & 3);
    const uint32_t row = quad * 32u + (uint32_t)lane;
    const uint32_t r8 = row & 7u;
    const uint32_t aux_row_off = kTOffAux + (row >> 3) * 1024u + r8 * 128u;
    const uint32_t scratch_off = kTOffAux + sw128_offset(row, 32u + 8u * (uint32_t)colq);
    const uint32_t row_bar = 1u + quad;
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16);
    uint32_t par = 0, buf = 0;

    // The rows of a unit come in with four 16 KiB bulk copies straight into the A tile (the first phase stored them as
    // the tile's shared-memory image); warp 0 issues them as soon as the tile is free and waits for them before it tells
    // the issuer.  (Through registers - 16 bytes per lane, ld.global + st.shared - the loads of a unit took 1 500 to
    // 5 700 cycles to ISSUE: 64 KiB per SM and unit is more than the L1 miss path sustains.)  {ray direction, sample id}
    // of the row is prefetched into registers.
    float4 meta_f = make_float4(0.f, 0.f, 0.f, 0.f);
    bool valid_f = false;
    uint32_t rows_par = 0;
    auto copy_rows = [&](int64_t u) {  // warp 0 only
      if (warp == 0 && elect_one()) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(h8) + (size_t)(2 * u + (int64_t)rank) * kActBytes;
        mbar_arrive_expect_tx(bar_rows, kActBytes);
#pragma unroll
        for (uint32_t c = 0; c < 4; ++c)
          tma_bulk_g2s(base + kTOffAct + c * kKBlockBytes, src + c * kKBlockBytes, kKBlockBytes, bar_rows);
      }
      __syncwarp();
    };
    auto fetch = [&](int64_t u) {
      const int64_t i = (2 * u + (int64_t)rank) * kTileM + row;
      meta_f = make_float4(0.f, 0.f, 0.f, 0.f);
      valid_f = u < n_units && i < n;
      if (valid_f && colq < 2) meta_f = __ldg(meta + i);  // quarter 0 stores the colour, quarters 0 and 1 write PE4(dir)
    };
    // PE4(dir) of the fetched unit -> aux buffer `b`; warp 0 also waits for the unit's rows; tells the issuer
    auto stage = [&](uint32_t b) {
      uint8_t* aux_row = smem + aux_row_off + b * kAuxBytes;
      const float dir_f[3] = {meta_f.x, meta_f.y, meta_f.z};
      if (colq == 0) write_pe16<kFp16, 0, kDirOctaves>(aux_row, r8, dir_f);
      else if (colq == 1) write_pe16<kFp16, 16, kDirOctaves>(aux_row, r8, dir_f);
      if (warp == 0) {
        mbar_wait(bar_rows, rows_par);
        rows_par ^= 1u;
      }
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (elect_one()) mbar_arrive_cluster(bar_in8, 0u);
    };

    int64_t g_cur = 0;
    bool valid_cur = false;
    if (pair_id < n_units) {
      copy_rows(pair_id);
      fetch(pair_id);
      g_cur = (int64_t)__float_as_int(meta_f.w);
      valid_cur = valid_f;
      stage(0u);
      fetch(pair_id + n_pairs);
    }
    for (int64_t u = pair_id; u < n_units; u += n_pairs) {
      // bottleneck, ActivationType::None: accumulator + bias -> 16 bits -> tensor memory, 32 columns at a time
      mbar_wait(bar_acc8, par);
      tc_fence_after();
#pragma unroll
      for (int hb = 0; hb < 2; ++hb) {
        uint32_t v[32], pk[16];
        tc_ld32(taddr + (uint32_t)colq * 64u + 32u * hb, v);
        tc_wait_ld();
        const float* bias = C.bias[8] + colq * 64 + 32 * hb;
#pragma unroll
        for (int j = 0; j < 16; ++j)
          pk[j] = pack2<kFp16, false>(__uint_as_float(v[2 * j]) + bias[2 * j], __uint_as_float(v[2 * j + 1]) + bias[2 * j + 1]);
        tc_st16(taddr + kTColA9 + (uint32_t)colq * 32u + 16u * hb, pk);
      }
      tc_wait_st();
      tc_fence_before();
      __syncwarp();
      if (elect_one()) mbar_arrive_cluster(bar_in9, 0u);
      // the A tile is free (the bottleneck MMAs have completed): the next unit's rows go in while viewdirs(u) runs
      const bool more = u + n_pairs < n_units;
      const int64_t g_next = (int64_t)__float_as_int(meta_f.w);
      const bool valid_next = valid_f;
      if (more) {
        copy_rows(u + n_pairs);
        stage(buf ^ 1u);
        fetch(u + 2 * n_pairs);
      }

      // viewdirs (ReLU) and the f32 rgb head: 32 columns per warp, as in the fused kernel
      mbar_wait(bar_acc9, par);
      tc_fence_after();
      uint32_t v[32];
      tc_ld32(taddr + 256u + 32u * (uint32_t)colq, v);
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
        float* scratch = reinterpret_cast<float*>(smem + scratch_off + buf * kAuxBytes);
        scratch[0] = r0;
        scratch[1] = r1;
        scratch[2] = r2;
      }
      named_bar_sync(row_bar, 128);
      if (colq == 0) {
        const uint8_t* aux = smem + kTOffAux + buf * kAuxBytes;
#pragma unroll
        for (uint32_t k = 1; k < 4; ++k) {
          const float* sc = reinterpret_cast<const float*>(aux + sw128_offset(row, 32u + 8u * k));
          r0 += sc[0];
          r1 += sc[1];
          r2 += sc[2];
        }
        if (valid_cur) {
          rgb[g_cur * 3 + 0] = 1.0f / (1.0f + expf(-r0));  // network.rs:165
          rgb[g_cur * 3 + 1] = 1.0f / (1.0f + expf(-r1));
          rgb[g_cur * 3 + 2] = 1.0f / (1.0f + expf(-r2));
        }
      }
      named_bar_sync(row_bar, 128);
      g_cur = g_next;
      valid_cur = valid_next;
      par ^= 1u;
      buf ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 16) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
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
  for (int row = 0; row < n; ++row) {
    for (int k = 0; k < krows; ++k) {
      const uint16_t h = host_cvt(W[(size_t)(k0 + k) * out_dim + row], fmt);
      std::memcpy(&img[base + sw128_offset((uint32_t)row, (uint32_t)k)], &h, 2);
    }
  }
}

}  // namespace

static const char* const kLayerNames[10] = {"dense0", "dense1", "dense2", "dense3", "dense4",
                                            "dense5", "dense6", "dense7", "bottleneck", "viewdirs"};

void fill_tc_consts(const TensorMap& t, TcConsts* c) {
  std::memset(c, 0, sizeof(*c));
  for (int l = 0; l < kNumLayers; ++l) {
    const std::vector<float>& b = t.at(std::string(kLayerNames[l]) + "_bias").data;
    for (int o = 0; o < layer_n(l); ++o) c->bias[l][o] = b[o];  // stays f32, added in the epilogue
  }
  const std::vector<float>& wa = t.at("alpha_kernel").data;
  const std::vector<float>& wr = t.at("rgb_kernel").data;
  for (int k = 0; k < 256; ++k) c->w_alpha[k] = wa[k];
  for (int k = 0; k < 128 * 3; ++k) c->w_rgb[k] = wr[k];
  for (int k = 0; k < 3; ++k) c->b_rgb[k] = t.at("rgb_bias").data[k];
  c->b_alpha = t.at("alpha_bias").data[0];
}

void build_tc_image(const TensorMap& t, int fmt, std::vector<uint8_t>* out) {
  if (fmt == 2) {
    build_tc_x3_image(t, out);
    return;
  }
  if (fmt == 3) {
    build_tc_tf32_image(t, out);
    return;
  }
  std::vector<uint8_t>& img = *out;
  img.clear();
  img.reserve(1500 * 1024);
  for (int l = 0; l < kNumLayers; ++l) {
    const std::vector<float>& W = t.at(std::string(kLayerNames[l]) + "_kernel").data;
    const int n = layer_n(l);
    // K-block images in consumption order
    if (l == 0) {
      append_chunk(img, W, n, 0, kPosEnc, n, fmt);  // PE10 rows 0..62
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
}

size_t tc_image_bytes(int fmt) {
  if (fmt == 3) return tc_tf32_image_bytes();
  size_t n = 0;
  for (int l = 0; l < kNumLayers; ++l) {
    const int kblocks = l == 0 ? 1 : ((l == 5 || l == 9) ? 5 : 4);  // 64-row K-blocks, encodings padded to 64
    n += (size_t)kblocks * layer_n(l) * 128;
  }
  return fmt == 2 ? 2 * n : n;
}

int upload_tc_weights(const uint8_t* img, size_t bytes, int fmt, const TensorMap& t, TcWeights* out) {
  fill_tc_consts(t, &out->consts);
  out->fmt = fmt;
  out->bytes = bytes;
  cudaError_t e = cudaMalloc(&out->d_chunks, bytes);
  if (e == cudaSuccess) e = cudaMemcpy(out->d_chunks, img, bytes, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error(std::string("upload tensor-core weight tiles: ") + cudaGetErrorString(e));
    return NERF_E_CUDA;
  }
  return NERF_OK;
}

int build_tc_weights(const TensorMap& t, int fmt, TcWeights* out) {
  std::vector<uint8_t> img;
  build_tc_image(t, fmt, &img);
  return upload_tc_weights(img.data(), img.size(), fmt, t, out);
}

void free_tc_weights(TcWeights* w) {
  if (w->d_chunks) cudaFree(w->d_chunks);
  w->d_chunks = nullptr;
  w->bytes = 0;
}

cudaError_t init_mlp_tc() {  // once per process and device, at context creation
  const void* kernels[7] = {(const void*)mlp_tc_kernel<true, true, false>, (const void*)mlp_tc_kernel<true, false, true>,
                            (const void*)mlp_tc_kernel<true, false, false>, (const void*)mlp_tc_kernel<false, false, true>,
                            (const void*)mlp_tc_kernel<false, false, false>,
                            (const void*)mlp_tc_kernel<true, false, false, true>,
                            (const void*)mlp_tc_kernel<false, false, false, true>};
  for (const void* k : kernels) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
  }
  for (const void* k : {(const void*)mlp_colour_kernel<true>, (const void*)mlp_colour_kernel<false>}) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTSmemBytes);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_mlp_tc(const TcWeights& w, const SampleSrc& src_in, const SampleDst& dst, int sm_count,
                          cudaStream_t s) {
  SampleSrc src = src_in;
  if (dst.sigma_floor < 0.0f) {  // raw-sigma launches address samples as (ray, slice): default = the whole ray
    if (src.mode != 1) return cudaErrorInvalidValue;
    if (src.sub_n <= 0) { src.sub_s0 = 0; src.sub_n = src.S; }
  } else if (src.index != nullptr || src.sub_n != 0) {
    return cudaErrorInvalidValue;  // ray lists and slices exist in the raw-sigma instantiation only
  }
  if (src.B == 0) return cudaSuccess;  // with a device-resident count, B is its upper bound
  const int64_t n_tiles = (src.B + kTileM - 1) / kTileM;
  const int64_t n_quads = (n_tiles + 3) / 4;  // a CTA pair works on four tiles at a time
  const int64_t max_pairs = sm_count / 2;
  const int grid = 2 * (int)(n_quads < max_pairs ? n_quads : max_pairs);
  cudaError_t e;
  unsigned long long* trace = nullptr;
  const char* trace_path = std::getenv("NERF_B200_TRACE");  // developer timeline, off in production
  const size_t trace_bytes = 7 * 2 * kTraceCap * sizeof(unsigned long long);
  if (trace_path != nullptr) {
    if ((e = cudaMalloc(&trace, trace_bytes)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(trace, 0, trace_bytes, s)) != cudaSuccess) {
      cudaFree(trace);
      return e;
    }
  }
  const bool raw = dst.sigma_floor < 0.0f;
  if (raw && (w.fmt != 1 || dst.rgb != nullptr)) return cudaErrorInvalidValue;  // only the fp16 density-only pass has it
  const bool skip = dst.skip_empty_rgb != 0 && dst.rgb != nullptr;
  const bool h8 = dst.h8 != nullptr;  // first phase of the two-phase fine pass
  if (h8 && (raw || dst.rgb == nullptr || dst.h8_meta == nullptr || dst.h8_count == nullptr || src.B > 0x7fffffff))
    return cudaErrorInvalidValue;
  const int which = raw ? 0 : (h8 ? (w.fmt == 1 ? 5 : 6) : (w.fmt == 1 ? (skip ? 1 : 2) : (skip ? 3 : 4)));
  switch (which) {
    case 5: mlp_tc_kernel<true, false, false, true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    case 6: mlp_tc_kernel<false, false, false, true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    case 0: mlp_tc_kernel<true, true, false><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    case 1: mlp_tc_kernel<true, false, true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    case 2: mlp_tc_kernel<true, false, false><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    case 3: mlp_tc_kernel<false, false, true><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
    default: mlp_tc_kernel<false, false, false><<<grid, kThreads, kSmemBytes, s>>>(w.consts, w.d_chunks, src, dst, trace); break;
  }
  e = cudaGetLastError();
  if (trace != nullptr) {
    std::vector<unsigned long long> h(trace_bytes / 8);
    cudaStreamSynchronize(s);
    cudaMemcpy(h.data(), trace, trace_bytes, cudaMemcpyDeviceToHost);
    cudaFree(trace);
    if (FILE* f = std::fopen(trace_path, "ab")) {
      std::fwrite(h.data(), 8, h.size(), f);
      std::fclose(f);
    }
  }
  return e;
}

cudaError_t launch_mlp_colour(const TcWeights& w, const uint16_t* d_h8, const float4* d_meta, const uint32_t* d_count,
                              float* d_rgb, int sm_count, cudaStream_t s) {
  if (w.fmt != 0 && w.fmt != 1) return cudaErrorInvalidValue;
  const int grid = 2 * (sm_count / 2);  // the number of rows is on the device: a full persistent grid
  if (w.fmt == 1)
    mlp_colour_kernel<true><<<grid, kColThreads, kTSmemBytes, s>>>(w.consts, w.d_chunks, d_h8, d_meta, d_count, d_rgb);
  else
    mlp_colour_kernel<false><<<grid, kColThreads, kTSmemBytes, s>>>(w.consts, w.d_chunks, d_h8, d_meta, d_count, d_rgb);
  return cudaGetLastError();
}

}  // namespace nerf
