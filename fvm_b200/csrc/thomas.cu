// Variant 1: thread-per-system Thomas, state and factors streamed from HBM.
//
// One thread owns one system and walks its cells; with the interleaved layout
// (cell stride = P) the 32 systems of a warp sit side by side, so every load and
// store of the warp is one fully used 256-byte segment.  The recurrences carry only
// one scalar (g or x); everything else in an iteration is independent of it, so the
// loads of U rows are issued as a batch ahead of the dependent FMA chain.
//
// HBM bytes per cell-update (algorithmic): forward reads w, fp, fq, fb and writes g
// (40 B), backward reads g, fq and writes w (24 B): 64 B.
#include "fvm_device.cuh"

namespace fvm {

// TV: time-varying boundary data.  The boundary term of step s is bt_left[p] * scale_left[s*P + p]
// (likewise on the right): beta is linear in the boundary flux (fvm_classes.py:244-263), so the
// caller passes the factors of a unit flux and the flux history as the scales.
template <int U, bool TV>
__global__ void __launch_bounds__(128) thomas_steps_kernel(
    int64_t J, int64_t P, const double* __restrict__ fp, const double* __restrict__ fq,
    const double* __restrict__ fb, int64_t f_ps, int64_t f_cs, double cth,
    const double* __restrict__ bt_left, const double* __restrict__ bt_right, double* w,
    int64_t w_ps, int64_t w_cs, int64_t nsteps, const double* __restrict__ scale_left,
    const double* __restrict__ scale_right) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double* P_ = fp + p * f_ps;
  const double* Q_ = fq + p * f_ps;
  const double* B_ = fb + p * f_ps;
  double* W_ = w + p * w_ps;
  const double bl0 = bt_left ? bt_left[p] : 0.0;
  const double br0 = bt_right ? bt_right[p] : 0.0;

  for (int64_t s = 0; s < nsteps; ++s) {
    double bl = bl0, br = br0;
    if (TV) {
      if (scale_left) bl *= scale_left[s * P + p];
      if (scale_right) br *= scale_right[s * P + p];
    }
    // forward: r~_j = fb_j w_j - c (fp_j w_{j-1} + fq_j w_{j+1}) ; g_j = r~_j - fp_j g_{j-1}
    double wm = 0.0, w0 = W_[0], g = 0.0;
    int64_t j = 0;
    for (; j + U <= J; j += U) {
      double pj[U], qj[U], bj[U], wn[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t jj = j + u;
        pj[u] = P_[jj * f_cs]; qj[u] = Q_[jj * f_cs]; bj[u] = B_[jj * f_cs];
        wn[u] = (jj + 1 < J) ? W_[(jj + 1) * w_cs] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t jj = j + u;
        double t = fma(qj[u], wn[u], pj[u] * wm);
        double r = fma(bj[u], w0, -cth * t);
        if (jj == 0) r += bl;
        if (jj == J - 1) r += br;
        g = fma(-pj[u], g, r);
        W_[jj * w_cs] = g;
        wm = w0; w0 = wn[u];
      }
    }
    for (; j < J; ++j) {
      const double pj = P_[j * f_cs], qj = Q_[j * f_cs], bj = B_[j * f_cs];
      const double wn = (j + 1 < J) ? W_[(j + 1) * w_cs] : 0.0;
      double t = fma(qj, wn, pj * wm);
      double r = fma(bj, w0, -cth * t);
      if (j == 0) r += bl;
      if (j == J - 1) r += br;
      g = fma(-pj, g, r);
      W_[j * w_cs] = g;
      wm = w0; w0 = wn;
    }
    // backward: x_j = g_j - fq_j x_{j+1}
    double x = g;
    j = J - 2;
    for (; j - (U - 1) >= 0; j -= U) {
      double qj[U], gj[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { qj[u] = Q_[(j - u) * f_cs]; gj[u] = W_[(j - u) * w_cs]; }
#pragma unroll
      for (int u = 0; u < U; ++u) { x = fma(-qj[u], x, gj[u]); W_[(j - u) * w_cs] = x; }
    }
    for (; j >= 0; --j) { x = fma(-Q_[j * f_cs], x, W_[j * w_cs]); W_[j * w_cs] = x; }
  }
}

}  // namespace fvm

extern "C" int fvm_theta_step_thomas(int64_t J, int64_t P, const double* fp, const double* fq,
                                     const double* fb, int64_t f_ps, int64_t f_cs, double theta,
                                     const double* bt_left, const double* bt_right, double* w,
                                     int64_t w_ps, int64_t w_cs, int64_t nsteps, void* stream) {
  if (!fp || !fq || !fb || !w) return FVM_ERR_INVALID_ARG;
  if (J < 3 || P < 1 || nsteps < 0 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (nsteps == 0) return FVM_OK;
  const int64_t blocks = (P + 127) / 128;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  const double cth = (1.0 - theta) / theta;
  fvm::thomas_steps_kernel<8, false><<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(
      J, P, fp, fq, fb, f_ps, f_cs, cth, bt_left, bt_right, w, w_ps, w_cs, nsteps, nullptr, nullptr);
  return (int)cudaGetLastError();
}

extern "C" int fvm_theta_step_thomas_tv(int64_t J, int64_t P, const double* fp, const double* fq,
                                        const double* fb, int64_t f_ps, int64_t f_cs, double theta,
                                        const double* bt_left, const double* bt_right,
                                        const double* scale_left, const double* scale_right, double* w,
                                        int64_t w_ps, int64_t w_cs, int64_t nsteps, void* stream) {
  if (!fp || !fq || !fb || !w) return FVM_ERR_INVALID_ARG;
  if (J < 3 || P < 1 || nsteps < 0 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if ((scale_left && !bt_left) || (scale_right && !bt_right)) return FVM_ERR_INVALID_ARG;
  if (nsteps == 0) return FVM_OK;
  const int64_t blocks = (P + 127) / 128;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  const double cth = (1.0 - theta) / theta;
  fvm::thomas_steps_kernel<8, true><<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(
      J, P, fp, fq, fb, f_ps, f_cs, cth, bt_left, bt_right, w, w_ps, w_cs, nsteps, scale_left, scale_right);
  return (int)cudaGetLastError();
}
