// Developer probe 3: tcgen05.mma.cta_group::2 (M = 256 across a CTA pair, N = 256, B split by N between the CTAs):
// numerics against a host matmul and cycles per MMA.  nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
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
    if (spins > (1u << 22)) __trap();
  }
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// commit to the barrier at the same offset in both CTAs of the pair
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}
__device__ __forceinline__ void tc_mma2_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {  // bf16 x bf16 -> f32
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__host__ __device__ inline uint32_t sw128_offset(uint32_t row, uint32_t k) {
  const uint32_t r8 = row & 7u;
  return (row >> 3) * 1024u + r8 * 128u + ((((k >> 3) ^ r8) & 7u) << 4) + (k & 7u) * 2u;
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

__host__ __device__ inline float a_val(int m, int k) { return (float)((m * 7 + k * 3) % 17 - 8) * 0.125f; }
__host__ __device__ inline float b_val(int n, int k) { return (float)((n * 5 + k) % 13 - 6) * 0.25f; }

// D[256 x 256] = A[256 x 64] * B[256 x 64]^T; CTA r holds A rows 128r.. and B rows (N) 128r..
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(160, 1) probe(float* out, long long* cyc, int timing_mmas, int variant) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw);
  __shared__ uint64_t bar[4];
  __shared__ uint32_t tmem_slot;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1;
  // A tile at +0 (16 KiB), B half at +16 KiB
  for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) {
    const int row = i >> 6, k = i & 63;
    *reinterpret_cast<__nv_bfloat16*>(smem + sw128_offset(row, k)) = __float2bfloat16(a_val(128 * rank + row, k));
    *reinterpret_cast<__nv_bfloat16*>(smem + 16384 + sw128_offset(row, k)) = __float2bfloat16(b_val(128 * rank + row, k));
  }
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar[0]), 1);
    mbar_init(smem_u32(&bar[1]), 1);
    mbar_init(smem_u32(&bar[2]), 1);
    mbar_init(smem_u32(&bar[3]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_slot;
  const uint32_t idesc = make_idesc(256, 256);

  if (warp == 4 && rank == 0) {  // the leader CTA issues for the pair
    const uint64_t ad = make_desc(base), bd = make_desc(base + 16384);
    if (elect_one()) {
      tc_mma2_f16(tmem_base, ad, bd, idesc, 0u);
      tc_mma2_f16(tmem_base, ad + 2, bd + 2, idesc, 1u);
      tc_mma2_f16(tmem_base, ad + 4, bd + 4, idesc, 1u);
      tc_mma2_f16(tmem_base, ad + 6, bd + 6, idesc, 1u);
      tc_commit2(smem_u32(&bar[0]));
    }
    __syncwarp();
  }
  if (warp < 4) {  // both CTAs drain their own 128 rows
    mbar_wait(smem_u32(&bar[0]), 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int row = warp * 32 + lane;
    for (int cb = 0; cb < 8; ++cb) {
      uint32_t v[32];
      tc_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + cb * 32, v);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (cluster_id == 0)
        for (int j = 0; j < 32; ++j) out[(size_t)(128 * rank + row) * 256 + cb * 32 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  cluster_sync();
  // timing: a stream of accumulating MMAs into the second half... (same accumulator; values irrelevant)
  if (warp == 4 && rank == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint64_t ad = make_desc(base), bd = make_desc(base + 16384);
    const long long t0 = clock64();
    if (variant == 9) {  // issue-queue depth: time stamp after every issue of 24 back-to-back MMAs
      if (elect_one()) {
        long long ts[24];
#pragma unroll
        for (int i = 0; i < 24; ++i) {
          tc_mma2_f16(tmem_base, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
          ts[i] = clock64();
        }
        tc_commit2(smem_u32(&bar[1]));
        if (blockIdx.x == 0)
          for (int i = 0; i < 24; ++i) cyc[1 + i] = ts[i] - t0;
      }
    } else if (elect_one()) {
      for (int i = 0; i < timing_mmas; i += 4) {
        const uint32_t blk = (uint32_t)(i >> 2);
        const uint32_t d = tmem_base + ((blk >> 2) & 1u) * 256u;
        // variant >= 1: operands rotate over 64 KiB of A (base + 32 KiB..) and a four-stage B ring (base + 96 KiB..)
        const uint64_t a2 = variant >= 1 ? make_desc(base + 32768 + (blk & 3u) * 16384u) : ad;
        const uint64_t b2 = variant >= 1 ? make_desc(base + 98304 + ((blk * 5u) & 3u) * 16384u) : bd;
        tc_mma2_f16(d, a2, b2, idesc, 1u);
        tc_mma2_f16(d, a2 + 2, b2 + 2, idesc, 1u);
        tc_mma2_f16(d, a2 + 4, b2 + 4, idesc, 1u);
        tc_mma2_f16(d, a2 + 6, b2 + 6, idesc, 1u);
        if (variant >= 2) tc_commit2(smem_u32(&bar[2]));  // nobody waits on these: cost of the commit itself
        if (variant >= 3) tc_commit2(smem_u32(&bar[3]));
      }
      tc_commit2(smem_u32(&bar[1]));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar[1]), 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && lane == 0) cyc[0] = t1 - t0;
  } else if (warp == 4) {
    mbar_wait(smem_u32(&bar[1]), 0);  // the peer must stay resident while the pair's MMAs read its shared memory
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync();
  if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int main() {
  float* d_out;
  long long* d_cyc;
  cudaMalloc(&d_out, 256 * 256 * 4);
  cudaMalloc(&d_cyc, 8 * 32);
  cudaMemset(d_out, 0, 256 * 256 * 4);
  const int smem = 98304 + 65536 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int n = 512;
  probe<<<148, 160, smem>>>(d_out, d_cyc, n, 0);
  cudaError_t e = cudaDeviceSynchronize();
  printf("launch: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> h(256 * 256);
  long long cyc;
  cudaMemcpy(h.data(), d_out, h.size() * 4, cudaMemcpyDeviceToHost);
  cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
  double worst = 0;
  int bad = 0;
  for (int m = 0; m < 256; ++m)
    for (int nn = 0; nn < 256; ++nn) {
      float ref = 0;
      for (int k = 0; k < 64; ++k) ref += a_val(m, k) * b_val(nn, k);
      const double d = std::abs((double)h[m * 256 + nn] - ref);
      if (d > worst) worst = d;
      if (d > 1e-3 && bad++ < 8) printf("mismatch D[%d][%d] = %f, want %f\n", m, nn, h[m * 256 + nn], ref);
    }
  printf("cta_group::2 M256 N256 K64: max |err| = %g, mismatches = %d; %.1f clk per MMA (%d MMAs, ideal 128)\n", worst, bad,
         (double)cyc / n, n);
  for (int variant = 0; variant < 4; ++variant)
    for (int nn : {512, 16384}) {
      probe<<<148, 160, smem>>>(d_out, d_cyc, nn, variant);
      e = cudaDeviceSynchronize();
      cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost);
      printf("variant %d (%s), %5d MMAs: %.1f clk per MMA %s\n", variant,
             variant == 0 ? "fixed operands" : variant == 1 ? "rotating operands" : variant == 2 ? "rotating + commit per 4" : "rotating + 2 commits per 4",
             nn, (double)cyc / nn, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  probe<<<148, 160, smem>>>(d_out, d_cyc, 24, 9);
  cudaDeviceSynchronize();
  long long ts[32];
  cudaMemcpy(ts, d_cyc, sizeof(ts), cudaMemcpyDeviceToHost);
  printf("clock after issuing MMA i (24 back to back, total %lld):", ts[0]);
  for (int i = 0; i < 24; ++i) printf(" %lld", ts[1 + i]);
  printf("\n");
  return 0;
}
