// Developer probe 4: does the tensor core's shared-memory operand fetch (SS tcgen05.mma) compete with LSU traffic
// (STS.128 / broadcast LDS.128) for the shared-memory crossbar?  One CTA per SM: warp 0 issues N = 256 MMAs over
// rotating operands, `workers` warps stream stores or loads.  Reports cycles per MMA and LSU bytes per clock, each
// alone and together.  nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  }
}
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
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
__host__ __device__ constexpr uint32_t make_idesc(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

constexpr uint32_t kA = 0, kB = 65536, kW = 65536 + 65536, kT = kW + 32768, kTotal = kT + 32768;

// lsu_mode: 0 none, 1 STS.128 (conflict-free), 2 LDS.128 broadcast, 3 LDS.128 conflict-free
__global__ void __launch_bounds__(576, 1) probe(int n_mma, int lsu_mode, int lsu_iters, long long* out,
                                                const uint8_t* gbuf = nullptr, int n_copies = 0, int workers = 0) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = smem_u32(smem);
  __shared__ uint64_t bar;
  __shared__ uint64_t tbar[2];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (int)(kTotal / 4); i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    mbar_init(smem_u32(&tbar[0]), 1);
    mbar_init(smem_u32(&tbar[1]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_slot;
  if (warp == 0) {
    if (lane == 0 && n_mma > 0) {
      const uint32_t idesc = make_idesc(256);
      const uint64_t ad0 = make_desc(base + kA), bd0 = make_desc(base + kB);
      const long long t0 = clock64();
      for (int i = 0; i < n_mma; i += 4) {
        const uint32_t blk = (uint32_t)(i >> 2);
        const uint64_t ad = ad0 + (blk & 3u) * (16384u >> 4), bd = bd0 + (blk & 1u) * (32768u >> 4);
        const uint32_t d = tmem_base + (lsu_mode == 4 ? 0u : ((blk >> 2) & 1u) * 256u);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc_mma_f16(d, ad + 2 * ks, bd + 2 * ks, idesc, 1u);
      }
      tc_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), 0);
      const long long t1 = clock64();
      if (blockIdx.x == 0) out[0] = t1 - t0;
    }
  } else if (n_copies > 0 && warp == workers + 1) {
    // 16 KiB bulk copies (two in flight) of an L2-resident buffer, as the weight ring does
    if (lane == 0) {
      size_t off = 0;
      const long long t0 = clock64();
      for (int i = 0; i < n_copies + 2; ++i) {
        const int st = i & 1;
        if (i >= 2) mbar_wait(smem_u32(&tbar[st]), ((i >> 1) - 1) & 1);
        if (i < n_copies) {
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&tbar[st])), "r"(16384u) : "memory");
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                           base + kT + st * 16384u),
                       "l"(gbuf + off), "r"(16384u), "r"(smem_u32(&tbar[st]))
                       : "memory");
          off = (off + 16384) % (1280 * 1024);
        }
      }
      const long long t1 = clock64();
      if (blockIdx.x == 0) out[40] = t1 - t0;
    }
  } else if (lsu_mode != 0 && warp <= workers) {
    const uint32_t w = (uint32_t)(warp - 1);
    // each warp owns a 512 B row group per iteration; 16 B per lane: conflict-free 128-bit accesses
    const uint32_t my = base + kW + ((w * 2048u) & 32767u) + (uint32_t)lane * 16u;
    const uint32_t bc = base + kW + ((w * 2048u) & 32767u);
    uint32_t a = lane, b = warp, c = 3, d = 4;
    const long long t0 = clock64();
    if (lsu_mode == 1) {
      for (int i = 0; i < lsu_iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(my + 512u * j), "r"(a), "r"(b), "r"(c), "r"(d)
                       : "memory");
      }
    } else if (lsu_mode == 4) {  // tcgen05.ld of the other accumulator (columns 256..511), this warp's lane quadrant
      const uint32_t taddr = tmem_base + 256u + (((uint32_t)(warp & 3) * 32u) << 16);
      for (int i = 0; i < lsu_iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t v[16];
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
              : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
              : "r"(taddr + 64u * (uint32_t)(j & 3) + 16u * (uint32_t)(i & 3)));
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          a ^= v[0] ^ v[5]; b ^= v[9]; c ^= v[15]; d ^= v[3];
        }
      }
    } else {
      const uint32_t addr = lsu_mode == 2 ? bc : my;
      for (int i = 0; i < lsu_iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t x, y, z, q;
          asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(q) : "r"(addr + 512u * j));
          a ^= x; b ^= y; c ^= z; d ^= q;
        }
      }
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0 && lane == 0) out[1 + w] = t1 - t0;
    if (a == 0xdeadbeef && b == c && d == 7) out[63] = a;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int main() {
  long long* d;
  cudaMalloc(&d, 64 * 8);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTotal);
  const int n_mma = 2048;  // 2048 x 128 clk = 262k clk
  const char* names[5] = {"none", "STS.128", "LDS.128 broadcast", "LDS.128 distinct", "LDTM.x16"};
  for (int workers : {8, 16}) {
    for (int mode : {1, 2, 4}) {
      // per-warp instruction count sized for ~250k clk when the LSU has the crossbar alone (4 wavefronts per STS/LDS.128,
      // 1 for a broadcast)
      const int wf = mode == 2 ? 1 : (mode == 4 ? 8 : 4);
      const int iters = 250000 / (workers * wf * 4);
      for (int with_mma : {0, 1}) {
        cudaMemset(d, 0, 64 * 8);
        probe<<<148, 32 * (1 + workers), kTotal>>>(with_mma ? n_mma : 0, mode, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[64];
        cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        long long worst = 0;
        for (int w = 0; w < workers; ++w) worst = h[1 + w] > worst ? h[1 + w] : worst;
        const double bytes = (double)workers * iters * 4 * 512;
        printf("%2d warps %-18s %s: LSU %.1f B/clk (%.0f wavefronts/kclk)", workers, names[mode],
               with_mma ? "+ MMA" : "alone", mode == 2 ? bytes / 32 / worst : (mode == 4 ? bytes * 4 / worst : bytes / worst),
               1000.0 * workers * iters * 4 * wf / worst);
        if (with_mma) printf(", %.1f clk/MMA", (double)h[0] / n_mma);
        printf(" %s\n", e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
    }
  }
  // three-way: MMA operand reads + STS stream + TMA copies into shared memory
  uint8_t* gbuf;
  cudaMalloc(&gbuf, 1280 * 1024);
  cudaMemset(gbuf, 0, 1280 * 1024);
  for (int with_mma : {0, 1})
    for (int with_sts : {0, 1}) {
      const int workers = 16, n_copies = 600;
      const int iters = with_sts ? 250000 / (workers * 4 * 4) : 0;
      cudaMemset(d, 0, 64 * 8);
      probe<<<148, 32 * (2 + workers), kTotal>>>(with_mma ? n_mma : 0, with_sts ? 1 : 0, iters, d, gbuf, n_copies, workers);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[64];
      cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
      printf("TMA 16 KiB copies x2 in flight%s%s: %.0f clk per copy", with_mma ? " + MMA" : "", with_sts ? " + 16 warps STS.128" : "",
             (double)h[40] / n_copies);
      if (with_mma) printf(", %.1f clk/MMA", (double)h[0] / n_mma);
      if (with_sts) printf(", STS %.1f B/clk", (double)workers * iters * 4 * 512 / h[1]);
      printf(" %s\n", e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  cudaMemset(d, 0, 64 * 8);
  probe<<<148, 32, kTotal>>>(n_mma, 0, 0, d);
  cudaDeviceSynchronize();
  long long h0;
  cudaMemcpy(&h0, d, 8, cudaMemcpyDeviceToHost);
  printf("MMA alone: %.1f clk/MMA\n", (double)h0 / n_mma);
  return 0;
}
