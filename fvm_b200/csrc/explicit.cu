// theta = 0 (explicit Euler): A = I, so a step of the reference's loop [fvm_classes.py:408-411]
// is just  w <- (I + dt*alpha*M) w + beta.  One thread per (problem, cell); reads the assembled
// diagonals of M, so it streams 40 B per cell-update.  The implicit entry points require
// theta in (0, 1]; the host layer routes theta == 0 here.
#include "fvm_device.cuh"

namespace fvm {

__global__ void __launch_bounds__(256) explicit_step_kernel(
    int64_t J, int64_t P, const double* __restrict__ lo, const double* __restrict__ di,
    const double* __restrict__ up, int64_t m_ps, int64_t m_cs, const double* __restrict__ dt, int alpha_left,
    int alpha_right, const double* __restrict__ beta_left, const double* __restrict__ beta_right,
    const double* __restrict__ w_in, double* __restrict__ w_out, int64_t w_ps, int64_t w_cs, int cell_fast) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= J * P) return;
  int64_t p, j;
  if (cell_fast) { p = tid / J; j = tid - p * J; }
  else           { j = tid / P; p = tid - j * P; }
  const int64_t o = p * m_ps + j * m_cs;
  const double* W_ = w_in + p * w_ps;
  const double s = mul(dt[p], alpha_of(j, J, alpha_left, alpha_right));   // (k*(1-theta))*alpha_j with theta = 0
  double r = fma(mul(s, di[o]), W_[j * w_cs], W_[j * w_cs]);
  if (j > 0) r = fma(mul(s, lo[o]), W_[(j - 1) * w_cs], r);
  if (j < J - 1) r = fma(mul(s, up[o]), W_[(j + 1) * w_cs], r);
  if (j == 0 && beta_left) r += beta_left[p];
  if (j == J - 1 && beta_right) r += beta_right[p];
  w_out[p * w_ps + j * w_cs] = r;
}

}  // namespace fvm

extern "C" int fvm_explicit_step(int64_t J, int64_t P, const double* lo, const double* di, const double* up,
                                 int64_t m_ps, int64_t m_cs, const double* dt, int32_t alpha_left,
                                 int32_t alpha_right, const double* beta_left, const double* beta_right,
                                 const double* w_in, double* w_out, int64_t w_ps, int64_t w_cs, void* stream) {
  if (!lo || !di || !up || !dt || !w_in || !w_out || w_in == w_out) return FVM_ERR_INVALID_ARG;
  if (J < 3 || P < 1 || ((alpha_left | alpha_right) & ~1)) return FVM_ERR_INVALID_ARG;
  const int64_t blocks = (J * P + 255) / 256;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  fvm::explicit_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      J, P, lo, di, up, m_ps, m_cs, dt, alpha_left, alpha_right, beta_left, beta_right, w_in, w_out, w_ps, w_cs,
      w_cs <= w_ps ? 1 : 0);
  return (int)cudaGetLastError();
}
