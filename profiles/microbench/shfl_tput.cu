// Microbenchmark: warp-shuffle throughput and latency per SM (B200): W warps per SM, each issuing
// K independent chains of __shfl_up_sync on 32-bit values (a double shuffle is two of these).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o shfl_tput.bin shfl_tput.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int K>
__global__ void shfl_kernel(int iters, long long* cycles, unsigned* sink) {
  unsigned x[K];
#pragma unroll
  for (int k = 0; k < K; ++k) x[k] = threadIdx.x * 2654435761u + k;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
#pragma unroll
      for (int k = 0; k < K; ++k) x[k] = __shfl_up_sync(0xffffffffu, x[k], 1) + 1u;
    }
  }
  long long t1 = clock64();
  __syncthreads();
  unsigned s = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) s += x[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int K>
void run(int warps, int iters, long long* cyc, unsigned* sink) {
  shfl_kernel<K><<<148, warps * 32>>>(iters, cyc, sink);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0; for (int i = 0; i < 148; ++i) mean += h[i]; mean /= 148;
  double n = (double)warps * K * 8 * iters;
  printf("K=%d chains, %2d warps/SM: %.2f warp-SHFL/clk/SM, %.1f cycles per dependent SHFL(+IADD)\n", K, warps,
         n / mean, mean / (8.0 * iters));
}

int main() {
  long long* cyc; unsigned* sink;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 4);
  const int iters = 2000;
  run<1>(1, iters, cyc, sink);
  run<1>(4, iters, cyc, sink);
  run<4>(4, iters, cyc, sink);
  run<8>(4, iters, cyc, sink);
  run<4>(8, iters, cyc, sink);
  run<8>(8, iters, cyc, sink);
  run<8>(16, iters, cyc, sink);
  run<8>(32, iters, cyc, sink);
  return 0;
}
