// Microbenchmark: fp64 FMA throughput and dependent-issue latency of one SM (B200, sm_100a).
//   throughput: W warps per SM, each thread runs K independent DFMA chains -> FMA/clk/SM
//   latency   : one warp, one dependent chain                              -> cycles per DFMA
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak.bin fp64_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void dfma_kernel(int iters, double a, double b, long long* cycles, double* sink) {
  double x[K];
#pragma unroll
  for (int k = 0; k < K; ++k) x[k] = threadIdx.x * 1e-3 + k;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int k = 0; k < K; ++k) x[k] = fma(x[k], a, b);
    }
  }
  long long t1 = clock64();
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) s += x[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int K>
void run(int warps, int iters, long long* cyc, double* sink) {
  dfma_kernel<K><<<148, warps * 32>>>(iters, 0.999999, 1e-9, cyc, sink);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0; for (int i = 0; i < 148; ++i) mean += h[i]; mean /= 148;
  double fma = (double)warps * 32 * K * 8 * iters;
  printf("K=%d chains/thread, %2d warps/SM: %.0f cycles, %.1f DFMA/clk/SM (%.2f cycles per warp-DFMA per SMSP), chain latency <= %.1f\n",
         K, warps, mean, fma / mean, mean / ((double)warps / 4 * K * 8 * iters) , mean / (8.0 * iters));
}

int main() {
  long long* cyc; double* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 8);
  const int iters = 2000;
  run<1>(1, iters, cyc, sink);     // latency: one warp, one chain
  run<1>(4, iters, cyc, sink);
  run<2>(4, iters, cyc, sink);
  run<4>(4, iters, cyc, sink);
  run<8>(4, iters, cyc, sink);
  run<8>(8, iters, cyc, sink);
  run<8>(16, iters, cyc, sink);
  run<8>(32, iters, cyc, sink);
  run<2>(8, iters, cyc, sink);
  run<4>(8, iters, cyc, sink);
  return 0;
}
