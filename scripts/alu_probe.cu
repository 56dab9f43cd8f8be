// Developer probe: issue cost per warp-instruction (clocks per SM sub-partition) of the epilogue's arithmetic:
// cvt.rn[.relu].f16x2.f32 / bf16x2 (F2FP), FADD, HFMA2, LOP3, with 1..4 warps per sub-partition.
// nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int kOp>
__global__ void __launch_bounds__(512, 1) probe(int iters, float seed, long long* out, uint32_t* sink) {
  float x[16];
  uint32_t r[8];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = seed * (float)(threadIdx.x + i);
#pragma unroll
  for (int i = 0; i < 8; ++i) r[i] = threadIdx.x * 7 + i;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if constexpr (kOp == 0) {
        asm volatile("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(x[2 * i + 1]), "f"(x[2 * i]));
        x[2 * i] = __uint_as_float(r[i]);  // dependent chain per register pair; 8 chains per warp
      } else if constexpr (kOp == 1) {
        asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(x[2 * i + 1]), "f"(x[2 * i]));
        x[2 * i] = __uint_as_float(r[i]);
      } else if constexpr (kOp == 2) {
        asm volatile("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(x[2 * i + 1]), "f"(x[2 * i]));
        x[2 * i] = __uint_as_float(r[i]);
      } else if constexpr (kOp == 3) {
        asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(seed));
        asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i + 8]) : "f"(seed));
      } else if constexpr (kOp == 4) {
        asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(r[i]) : "r"(0x3c003c00u), "r"(0x00010001u));
      } else if constexpr (kOp == 5) {
        asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(0x55u), "r"(r[(i + 1) & 7]));
      } else if constexpr (kOp == 6) {
        asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(r[i]) : "r"(__float_as_uint(x[2 * i])), "r"(__float_as_uint(x[2 * i + 1])));
        x[2 * i] = __uint_as_float(r[i]);
      }
    }
  }
  const long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) acc ^= r[i] ^ __float_as_uint(x[i]) ^ __float_as_uint(x[i + 8]);
  if (acc == 0xdeadbeefu) sink[0] = acc;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
}

template <int kOp>
void run(const char* name, int per_iter, long long* d, uint32_t* sink) {
  for (int warps_per_smsp : {1, 2, 4}) {
    const int iters = 4096;
    probe<kOp><<<148, 128 * warps_per_smsp>>>(iters, 1.0009765625f, d, sink);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    printf("%-28s %d warps/SMSP: %.2f clk per warp-instruction per SMSP\n", name, warps_per_smsp,
           (double)h / ((double)iters * per_iter * warps_per_smsp));
  }
}

int main() {
  long long* d;
  uint32_t* sink;
  cudaMalloc(&d, 8);
  cudaMalloc(&sink, 4);
  run<0>("cvt.rn.relu.f16x2.f32", 8, d, sink);
  run<1>("cvt.rn.f16x2.f32", 8, d, sink);
  run<2>("cvt.rn.relu.bf16x2.f32", 8, d, sink);
  run<3>("add.f32", 16, d, sink);
  run<4>("fma.rn.f16x2", 8, d, sink);
  run<5>("lop3", 8, d, sink);
  run<6>("prmt", 8, d, sink);
  return 0;
}
