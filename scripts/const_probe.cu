// Developer probe: cost of reading per-layer bias vectors from a large __grid_constant__ kernel parameter the way the
// MLP epilogue does (64 floats per warp per layer, layer index at run time), as a function of how many KB of the
// constant bank the loop cycles through.  nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct Consts {
  alignas(16) float bias[12][256];
};

template <int kVec>
__global__ void __launch_bounds__(512, 1) probe(const __grid_constant__ Consts C, int iters, int layers, long long* out, float* sink) {
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  (void)warp;
  float x[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) x[j] = (float)(threadIdx.x + j);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll 1
    for (int l = 0; l < layers; ++l) {
#pragma unroll
      for (int blk = 0; blk < 4; ++blk) {
        const float* b = C.bias[l] + (it & 3) * 64 + blk * 16;
        if constexpr (kVec == 0) {  // no constants at all: the FADD floor
#pragma unroll
          for (int j = 0; j < 16; ++j) x[j] += 1.5f;
        } else if constexpr (kVec == 1) {
#pragma unroll
          for (int j = 0; j < 16; ++j) x[j] += b[j];
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 v = reinterpret_cast<const float4*>(b)[j];
            x[4 * j] += v.x;
            x[4 * j + 1] += v.y;
            x[4 * j + 2] += v.z;
            x[4 * j + 3] += v.w;
          }
        }
      }
    }
  }
  const long long t1 = clock64();
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < 16; ++j) acc += x[j];
  if (acc == 12345.678f) sink[0] = acc;
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
}

template <int kVec>
void run(const Consts& h, long long* d, float* sink) {
  for (int threads : {128, 512})
    for (int layers : {1, 12}) {
      const int iters = 240 / layers * 20;
      probe<kVec><<<148, threads>>>(h, iters, layers, d, sink);
      cudaError_t e = cudaDeviceSynchronize();
      long long t;
      cudaMemcpy(&t, d, 8, cudaMemcpyDeviceToHost);
      printf("%d-float loads, %3d threads, cycling %2d KB of constants: %.0f clk per 64-float bias block per warp %s\n", kVec,
             threads, layers, (double)t / ((double)iters * layers), e == cudaSuccess ? "" : cudaGetErrorString(e));
    }
}

int main() {
  long long* d;
  float* sink;
  cudaMalloc(&d, 8);
  cudaMalloc(&sink, 4);
  Consts h;
  for (int l = 0; l < 12; ++l)
    for (int j = 0; j < 256; ++j) h.bias[l][j] = 1e-3f * (l + j);
  run<0>(h, d, sink);
  run<1>(h, d, sink);
  run<4>(h, d, sink);
  return 0;
}
