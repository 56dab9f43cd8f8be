// Developer probe: sustained per-SM throughput and latency of 16 KiB 1-D bulk copies (L2 -> shared memory) when all
// 148 SMs stream the same L2-resident buffer (the weight-ring access pattern) or disjoint buffers, with 1..4 copies
// in flight per SM.  nvcc -gencode arch=compute_100a,code=sm_100a.
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
__device__ __forceinline__ void expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

constexpr uint32_t kChunk = 16384;

__global__ void __launch_bounds__(544, 1) probe(const uint8_t* buf, size_t buf_bytes, size_t per_sm_offset, int depth,
                                                int n_copies, int sts_iters, int fence_every, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[4];
  const uint32_t base = smem_u32(smem);
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint8_t* src = buf + (size_t)blockIdx.x * per_sm_offset;
    const size_t span = per_sm_offset ? per_sm_offset : buf_bytes;
    size_t off = 0;
    const long long t0 = clock64();
    long long lat_sum = 0;
    for (int i = 0; i < n_copies + depth; ++i) {
      const int st = i % depth;
      if (i >= depth) {  // wait for the copy that used this stage
        mbar_wait(smem_u32(&bars[st]), ((i / depth) - 1) & 1);
      }
      if (i < n_copies) {
        expect_tx(smem_u32(&bars[st]), kChunk);
        bulk_g2s(base + st * kChunk, src + off, kChunk, smem_u32(&bars[st]));
        off += kChunk;
        if (off + kChunk > span) off = 0;
      }
      if (depth == 1 && i >= 1) lat_sum = clock64() - t0;
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = lat_sum;
    }
  } else if (threadIdx.x >= 32 && sts_iters > 0) {
    // epilogue-like store stream: conflict-free STS.128 into a separate 32 KiB region
    const uint32_t w = threadIdx.x / 32 - 1, lane = threadIdx.x & 31;
    const uint32_t my = base + 4 * kChunk + ((w * 2048u) & 32767u) + lane * 16u;
    const long long t0 = clock64();
    for (int i = 0; i < sts_iters; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(my + 512u * j), "r"(w), "r"(lane), "r"(i), "r"(j) : "memory");
      if (fence_every > 0 && (i % fence_every) == fence_every - 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    const long long t1 = clock64();
    if (blockIdx.x == 0 && threadIdx.x == 32) out[2] = t1 - t0;
  }
}

int main() {
  const size_t buf_bytes = 148ull * 65536;  // 9.25 MiB: L2 resident
  uint8_t* buf;
  long long* d;
  cudaMalloc(&buf, buf_bytes);
  cudaMemset(buf, 1, buf_bytes);
  cudaMalloc(&d, 64);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * kChunk + 32768);
  const int n = 512;
  for (int grid : {1, 148})
    for (int shared_buf : {1, 0})
      for (int depth : {1, 2, 4}) {
        // shared: every SM streams the same 1.25 MiB (the weight image); disjoint: 64 KiB of its own
        const size_t span = shared_buf ? 1280 * 1024 : 0;
        for (int rep = 0; rep < 2; ++rep)
          probe<<<grid, 32, 4 * kChunk + 32768>>>(buf, shared_buf ? span : buf_bytes, shared_buf ? 0 : 65536, depth, n, 0, 0, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("grid %3d %-8s depth %d: %.0f clk per 16 KiB copy = %.1f B/clk/SM (%.0f B/clk chip) %s\n", grid,
               shared_buf ? "shared" : "disjoint", depth, (double)h[0] / n, (double)kChunk * n / h[0],
               (double)kChunk * n / h[0] * grid, e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
  // the same copies while 16 warps stream STS.128 (sized to outlast the copies)
  for (int fence_every : {0, 4, 1})
  for (int depth : {1, 4}) {
    const int sts_iters = 2 * n * 400 / (16 * 4 * 4);
    for (int rep = 0; rep < 2; ++rep)
      probe<<<148, 544, 4 * kChunk + 32768>>>(buf, 1280 * 1024, 0, depth, n, sts_iters, fence_every, d);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[3];
    cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
    printf("fence.proxy.async every %d x4 STS; grid 148 shared depth %d + 16 warps STS.128: %.0f clk per 16 KiB copy = %.1f B/clk/SM; STS %.1f B/clk %s\n", fence_every, depth,
           (double)h[0] / n, (double)kChunk * n / h[0], 16.0 * sts_iters * 4 * 512 / h[2],
           e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
  return 0;
}
