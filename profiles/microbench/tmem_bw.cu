// Microbenchmark: is tensor memory (TMEM) usable as read-mostly coefficient storage next to
// shared memory?  Measures, per SM, bytes/clk of (a) tcgen05.ld 32x32b.x32 by 1..4 warps (one
// per lane quadrant), (b) conflict-free LDS.64 by the same warps, (c) both interleaved.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bw.bin tmem_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]),"=r"(r[16]),"=r"(r[17]),"=r"(r[18]),"=r"(r[19]),"=r"(r[20]),"=r"(r[21]),"=r"(r[22]),"=r"(r[23]),"=r"(r[24]),"=r"(r[25]),"=r"(r[26]),"=r"(r[27]),"=r"(r[28]),"=r"(r[29]),"=r"(r[30]),"=r"(r[31]) : "r"(taddr));
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%32], {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31};"
               :: "r"(r[0]),"r"(r[1]),"r"(r[2]),"r"(r[3]),"r"(r[4]),"r"(r[5]),"r"(r[6]),"r"(r[7]),"r"(r[8]),"r"(r[9]),"r"(r[10]),"r"(r[11]),"r"(r[12]),"r"(r[13]),"r"(r[14]),"r"(r[15]),"r"(r[16]),"r"(r[17]),"r"(r[18]),"r"(r[19]),"r"(r[20]),"r"(r[21]),"r"(r[22]),"r"(r[23]),"r"(r[24]),"r"(r[25]),"r"(r[26]),"r"(r[27]),"r"(r[28]),"r"(r[29]),"r"(r[30]),"r"(r[31]), "r"(taddr));
}

// mode 0: TMEM only, 1: smem only, 2: both
__global__ void __launch_bounds__(256) bw_kernel(int mode, int active_warps, int iters, long long* cycles,
                                                 unsigned* sink, double* check) {
  extern __shared__ double sm[];
  __shared__ uint32_t tmem_base;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base);
    uint32_t ncols = 512;
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(dst), "r"(ncols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t base = tmem_base;
  const int quad = warp & 3;                               // lane quadrant this warp may touch
  const int half = warp >> 2;                              // warps 4..7 use the upper 256 columns
  const uint32_t my = base + ((uint32_t)(quad * 32) << 16) + half * 256;
  // fill: column c of lane l holds a recognisable value
  uint32_t r[32];
  for (int blk = 0; blk < 8; ++blk) {
#pragma unroll
    for (int i = 0; i < 32; ++i) r[i] = (warp << 24) | (lane << 16) | (blk * 32 + i);
    tmem_st32(my + blk * 32, r);
  }
  asm volatile("tcgen05.wait::st.sync.aligned;");
  for (int i = threadIdx.x; i < 8 * 1024 * 2; i += blockDim.x) sm[i] = (double)i;
  __syncthreads();
  // read-back check of one block (correctness of addressing)
  tmem_ld32(my + 3 * 32, r);
  asm volatile("tcgen05.wait::ld.sync.aligned;");
  if (blockIdx.x == 0) check[threadIdx.x] = (double)r[5];   // expect (warp<<24)|(lane<<16)|(3*32+5)
  __syncthreads();

  unsigned acc = 0; double dacc = 0.0;
  const double* mysm = sm + warp * 2048 + lane;
  long long t0 = clock64();
  if (warp < active_warps) {
    for (int it = 0; it < iters; ++it) {
      if (mode != 1) {
#pragma unroll
        for (int blk = 0; blk < 8; ++blk) {                 // 8 x 32 columns = 256 columns = 32 KB/warp
          tmem_ld32(my + blk * 32, r);
          asm volatile("tcgen05.wait::ld.sync.aligned;");
#pragma unroll
          for (int i = 0; i < 32; ++i) acc += r[i];
        }
      }
      if (mode != 0) {
#pragma unroll
        for (int k = 0; k < 64; ++k) dacc += mysm[k * 32];  // 64 x 256 B = 16 KB/warp
      }
    }
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + (unsigned)dacc;
  __syncthreads();
  if (warp == 0) {
    uint32_t ncols = 512;
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(base), "r"(ncols));
  }
}

int main() {
  long long* cyc; unsigned* sink; double* check;
  cudaMalloc(&cyc, 148 * 8); cudaMalloc(&sink, 148 * 256 * 4); cudaMalloc(&check, 256 * 8);
  cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * 1024 * 2 * 8);
  const int iters = 200;
  double h[256];
  for (int mode = 0; mode < 3; ++mode)
    for (int aw = 1; aw <= 8; aw *= 2) {
      bw_kernel<<<148, 256, 8 * 1024 * 2 * 8>>>(mode, aw, iters, cyc, sink, check);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
      long long hc[148]; cudaMemcpy(hc, cyc, sizeof(hc), cudaMemcpyDeviceToHost);
      cudaMemcpy(h, check, sizeof(h), cudaMemcpyDeviceToHost);
      double mean = 0; for (int i = 0; i < 148; ++i) mean += hc[i]; mean /= 148;
      double tm = (mode != 1) ? (double)aw * iters * 8 * 32 * 32 * 4 : 0;   // bytes per SM
      double sb = (mode != 0) ? (double)aw * iters * 64 * 256 : 0;
      printf("mode=%d (%s) warps=%d cycles=%.0f  tmem %.1f B/clk/SM  smem %.1f B/clk/SM  check[37]=%.0f (expect %d)\n",
             mode, mode == 0 ? "tmem" : mode == 1 ? "smem" : "both", aw, mean, tm / mean, sb / mean, h[37],
             (1 << 24) | (5 << 16) | (3 * 32 + 5));
    }
  return 0;
}
