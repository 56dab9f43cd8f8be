// Developer probe 2: what slows a stream of M128 N128 K16 tcgen05.mma down inside a busy CTA?
// One issuing thread (warp 17) + 16 "background" warps whose activity is selected by a bit mask:
//   bit0: issuer commits to an mbarrier after every 4 MMAs           bit1: issuer also waits on the commit of 4 groups ago
//   bit2: background warps poll an mbarrier that never completes       bit3: background warps loop on tcgen05.ld (32x32b.x32)
//   bit4: background warps loop on 16-byte st.shared (swizzled rows)   bit5: a TMA warp streams 16 KiB bulk copies into smem
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try(bar, parity)) {
  }
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
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

constexpr int kA = 0;                // A tile: 4 K-blocks x 16 KiB
constexpr int kB = 65536;            // B ring: 4 x 16 KiB
constexpr int kScratch = 131072;     // st.shared / TMA target: 64 KiB
constexpr int kSmem = 196608 + 1024;

__global__ void __launch_bounds__(640, 1) probe(int n, int mask, int groups, const uint8_t* gsrc, long long* out,
                                                uint32_t* sink, int fill) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw);
  __shared__ uint64_t bars[8];  // 0..3 commit ring, 4 never completes, 5 TMA, 6 done flag
  __shared__ uint32_t tmem_slot;
  __shared__ volatile int stop;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 196608 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = fill == 0 ? 0u : (fill == 1 ? 0x3c003c00u : ((i * 2654435761u) & 0x3fff3fffu));
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(smem_u32(&bars[i]), 1);
    stop = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 18) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_slot;

  if (warp == 17) {
    // warp-uniform issue, as in the product kernel: all lanes run the loop, one elected lane issues
    const uint32_t idesc = make_idesc(n);
    uint32_t parbits = 0;
    const long long t0 = clock64();
    if (mask & 256) {
      if (elect_one()) {
    for (int g = 0; g < groups; ++g) {
      const int st = g & 3;
      if ((mask & 2) && g >= 4) {
        mbar_wait(smem_u32(&bars[st]), (parbits >> st) & 1u);
        parbits ^= 1u << st;
      }
      if (mask & 64) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (mask & 128) {
        out[8 + (g & 7)] = clock64();  // tracer-like global store + clock read
      }
      const uint64_t ad = make_desc(base + kA + (g & 3) * 16384), bd = make_desc(base + kB + st * 16384);
      const uint32_t d = tmem_base + (uint32_t)((g >> 2) & (512 / n - 1)) * (uint32_t)n;
      {
        tc_mma_f16(d, ad, bd, idesc, 1u);
        tc_mma_f16(d, ad + 2, bd + 2, idesc, 1u);
        tc_mma_f16(d, ad + 4, bd + 4, idesc, 1u);
        tc_mma_f16(d, ad + 6, bd + 6, idesc, 1u);
        if (mask & 1) tc_commit(smem_u32(&bars[st]));
      }
    }
      }
      __syncwarp();
    } else {
    for (int g = 0; g < groups; ++g) {
      const int st = g & 3;
      if ((mask & 2) && g >= 4) {
        mbar_wait(smem_u32(&bars[st]), (parbits >> st) & 1u);
        parbits ^= 1u << st;
      }
      if (mask & 64) asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (mask & 128) {
        out[8 + (g & 7)] = clock64();  // tracer-like global store + clock read
      }
      const uint64_t ad = make_desc(base + kA + (g & 3) * 16384), bd = make_desc(base + kB + st * 16384);
      const uint32_t d = tmem_base + (uint32_t)((g >> 2) & (512 / n - 1)) * (uint32_t)n;
      if (elect_one()) {
        tc_mma_f16(d, ad, bd, idesc, 1u);
        tc_mma_f16(d, ad + 2, bd + 2, idesc, 1u);
        tc_mma_f16(d, ad + 4, bd + 4, idesc, 1u);
        tc_mma_f16(d, ad + 6, bd + 6, idesc, 1u);
        if (mask & 1) tc_commit(smem_u32(&bars[st]));
      }
      __syncwarp();
    }
    }
    const long long t1 = clock64();
    if (elect_one()) tc_commit(smem_u32(&bars[6]));
    __syncwarp();
    mbar_wait(smem_u32(&bars[6]), 0);
    const long long t2 = clock64();
    stop = 1;
    if (blockIdx.x == 0 && lane == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  } else if (warp == 16) {
    if ((mask & 32) && lane == 0) {
      uint32_t par = 0;
      while (!stop) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[5])), "r"(16384u)
                     : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         base + kScratch),
                     "l"(gsrc + (size_t)(blockIdx.x & 63) * 16384), "r"(16384u), "r"(smem_u32(&bars[5]))
                     : "memory");
        mbar_wait(smem_u32(&bars[5]), par);
        par ^= 1u;
      }
    }
  } else if (warp < 16) {
    uint32_t acc = 0;
    const uint32_t quad = warp & 3;
    if (mask & 4) {
      while (!stop) acc += mbar_try(smem_u32(&bars[4]), 0) ? 1u : 0u;
    } else if (mask & 8) {
      uint32_t v[32];
      const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + (uint32_t)(warp >> 2) * 32u;
      while (!stop) {
        tc_ld32(taddr, v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc += v[lane & 31];
      }
    } else if (mask & 16) {
      const uint32_t row = quad * 32 + lane, r8 = row & 7;
      uint8_t* p = smem + kScratch + (warp >> 2) * 16384 + (row >> 3) * 1024 + r8 * 128;
      while (!stop) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p + ((i ^ r8) << 4))), "r"(acc), "r"(i),
                       "r"(2), "r"(3)
                       : "memory");
        ++acc;
      }
    }
    if (acc == 0xdeadbeef) sink[0] = acc;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 18) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

int main() {
  long long* d;
  uint32_t* sink;
  uint8_t* g;
  cudaMalloc(&d, 256);
  cudaMalloc(&sink, 64);
  cudaMalloc(&g, 64 * 16384);
  cudaMemset(g, 0, 64 * 16384);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  const int groups = 256;
  const int masks[] = {0, 3, 256, 256 | 1, 256 | 3};
  for (int grid : {148})
  for (int fill : {2})
  for (int n : {128, 256})
    for (int mask : masks) {
      printf("grid %3d fill %d ", grid, fill);
      probe<<<grid, 640, kSmem>>>(n, mask, groups, g, d, sink, fill);
      cudaError_t e = cudaDeviceSynchronize();
      long long h[2];
      cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      printf("N %3d mask %2d: issue %.1f clk/mma, complete %.1f clk/mma (ideal %d) %s\n", n, mask,
             (double)h[0] / (groups * 4), (double)h[1] / (groups * 4), n / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
  return 0;
}
