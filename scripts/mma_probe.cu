// Developer probe: cycles per tcgen05.mma (kind::f16, M128, K16, SS) as a function of N and of how many
// independent accumulators the issue stream rotates over.  nvcc -gencode arch=compute_100a,code=sm_100a.
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

// mode: number of independent accumulators rotated over (1, 2, 4); ksteps: K16 steps per smem tile visit
__global__ void __launch_bounds__(128, 1) probe(int n, int n_acc, int total, int same_ab, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (16384 + 32768 * 2) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(smem_raw + (base - raw))[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
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
  if (threadIdx.x == 0) {
    const uint32_t idesc = make_idesc(n);
    const uint32_t a0 = base, b0 = base + 16384;
    uint32_t par = 0;
    const uint32_t stride = 512u / n_acc;
    const int sh = n_acc == 1 ? 0 : (n_acc == 2 ? 1 : 2);
    for (int rep = 0; rep < 3; ++rep) {
      const long long t0 = clock64();
      if (same_ab == 2) {
        // lean issue: descriptors built once, only the low word advances per K16 step
        const uint64_t ad = make_desc(a0), bd = make_desc(b0);
        for (int i = 0; i < total; i += 4) {
          const uint32_t d = tmem_base + (uint32_t)((i >> 2) & (n_acc - 1)) * stride;
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) tc_mma_f16(d, ad + 2 * ks, bd + 2 * ks, idesc, 1u);
        }
      } else {
#pragma unroll 4
      for (int i = 0; i < total; ++i) {
        const int acc = i & (n_acc - 1);
        const int ks = same_ab ? 0 : (i >> sh) & 3;
        const uint32_t boff = same_ab ? 0 : (uint32_t)((i >> (sh + 2)) & 1) * 32768u;
        tc_mma_f16(tmem_base + (uint32_t)acc * stride, make_desc(a0 + ks * 32), make_desc(b0 + boff + ks * 32),
                   idesc, 1u);
      }
      }
      const long long t1 = clock64();
      tc_commit(smem_u32(&bar));
      mbar_wait(smem_u32(&bar), par);
      par ^= 1u;
      const long long t2 = clock64();
      if (blockIdx.x == 0) {
        out[rep * 2] = t1 - t0;
        out[rep * 2 + 1] = t2 - t0;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  const int smem = 16384 + 65536 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int total = 256;
  for (int grid : {148})
    for (int same_ab : {2, 0})
      for (int n : {64, 128, 256})
        for (int n_acc : {1, 2, 4}) {
          if (n * n_acc > 512) continue;
          probe<<<grid, 128, smem>>>(n, n_acc, total, same_ab, d);
          cudaError_t e = cudaDeviceSynchronize();
          long long h[6];
          cudaMemcpy(h, d, 48, cudaMemcpyDeviceToHost);
          printf("grid %3d same_ab %d N %3d accs %d: issue %.1f clk/mma, complete %.1f clk/mma (ideal %d) %s\n", grid,
                 same_ab, n, n_acc, (double)h[4] / total, (double)h[5] / total, n / 2,
                 e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
  return 0;
}
