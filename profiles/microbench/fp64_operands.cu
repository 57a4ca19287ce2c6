// Microbenchmark: does the fp64 pipe of one SM (B200, sm_100a) keep its 64 DFMA/clk/SM when every
// DFMA reads THREE DISTINCT 64-bit register operands (as the fused solver's sweeps do), instead
// of the x = fma(x, a, b) with loop-invariant a, b that fp64_peak.cu times?
//   K independent chains per thread; operands come from two 16-entry register arrays and change
//   with every instruction, so the operand-reuse cache cannot help.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_operands.bin fp64_operands.cu
#include <cstdio>
#include <cuda_runtime.h>

// DISTINCT: 3 = all three operands change with every instruction; 2 = the addend is loop-invariant;
// 1 = only the chain value changes (what fp64_peak.cu times)
template <int K, int DISTINCT>
__global__ void __launch_bounds__(256, 1) dfma_kernel(int iters, const double* __restrict__ in, long long* cycles, double* sink) {
  double x[K], a[16], b[16];
#pragma unroll
  for (int k = 0; k < K; ++k) x[k] = threadIdx.x * 1e-3 + k;
#pragma unroll
  for (int j = 0; j < 16; ++j) { a[j] = in[j] + threadIdx.x * 1e-9; b[j] = in[16 + j] - threadIdx.x * 1e-9; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
#pragma unroll
      for (int k = 0; k < K; ++k) {
        if (DISTINCT == 3) x[k] = fma(a[(r + 5 * k) & 15], x[k], b[(3 * r + k) & 15]);
        else if (DISTINCT == 2) x[k] = fma(a[(r + 5 * k) & 15], x[k], b[0]);
        else x[k] = fma(a[0], x[k], b[0]);
      }
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

template <int K, int D>
void run(int warps, int iters, const double* in, long long* cyc, double* sink) {
  dfma_kernel<K, D><<<148, warps * 32>>>(iters, in, cyc, sink);
  cudaDeviceSynchronize();
  long long h[148];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  double mean = 0; for (int i = 0; i < 148; ++i) mean += h[i]; mean /= 148;
  double fma = (double)warps * 32 * K * 16 * iters;
  printf("%d changing register operands per DFMA, K=%d chains/thread, %2d warps/SM: %.1f DFMA/clk/SM\n", D, K, warps, fma / mean);
}

int main() {
  long long* cyc; double* sink; double* in;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 1024 * 8); cudaMalloc(&in, 32 * 8);
  double h[32]; for (int i = 0; i < 32; ++i) h[i] = (i < 16) ? 0.999 + 1e-5 * i : 1e-9 * i;
  cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
  const int iters = 1000;
  run<4, 1>(4, iters, in, cyc, sink);
  run<4, 2>(4, iters, in, cyc, sink);
  run<4, 3>(4, iters, in, cyc, sink);
  run<8, 3>(4, iters, in, cyc, sink);
  run<2, 1>(8, iters, in, cyc, sink);
  run<2, 2>(8, iters, in, cyc, sink);
  run<2, 3>(8, iters, in, cyc, sink);
  run<4, 2>(8, iters, in, cyc, sink);
  run<4, 3>(8, iters, in, cyc, sink);
  run<8, 3>(8, iters, in, cyc, sink);
  return 0;
}
