// Variant 1: thread-per-system Thomas, state and factors streamed from HBM.
//
// One thread owns one system and walks its cells; with the interleaved layout
// (cell stride = P) the 32 systems of a warp sit side by side, so every load and
// store of the warp is one fully used 256-byte segment.  The recurrences carry only
// one scalar (h or z); everything else in an iteration is independent of it, so the
// loads of U rows are issued as a batch ahead of the dependent FMA chain.
//
// The step has NO right-hand side:  B = I/theta - c A  (c = (1-theta)/theta) gives
// A^-1 B = (1/theta) A^-1 - c I.  With the factors of fvm_theta_factor, l = fp, s = fq,
// kappa = fb (= 1/(theta pivot)) and z = yhat - c w:
//   forward   h_j = w_j + beta_j' - l_j h_{j-1} ;  t_j = h_j kappa_j - c w_j - c s_j w_{j+1}   (t over w)
//   backward  z_j = t_j - s_j z_{j+1}                                                           (z over t)
// A right-hand side  B w  would cancel terms of size dt*d/h^2 against each other at every step;
// this form does not, and its error over 10^4 steps is 100x smaller (tests/test_gpu_long_run.py).
//
// HBM bytes per cell-update as this kernel moves them: forward reads w, l, kappa, s and writes t
// (40 B), backward reads t, s and writes w (24 B): 64 B.  The algorithmic figure of SURVEY 8(d)
// for three streamed per-cell arrays is 40 B (what a kernel that kept t on chip would move).
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
  const double* L_ = fp + p * f_ps;
  const double* S_ = fq + p * f_ps;
  const double* K_ = fb + p * f_ps;
  double* W_ = w + p * w_ps;
  const double bl0 = bt_left ? bt_left[p] : 0.0;
  const double br0 = bt_right ? bt_right[p] : 0.0;

  for (int64_t s = 0; s < nsteps; ++s) {
    double bl = bl0, br = br0;
    if (TV) {
      if (scale_left) bl *= scale_left[s * P + p];
      if (scale_right) br *= scale_right[s * P + p];
    }
    double w0 = W_[0], h = 0.0;
    int64_t j = 0;
    for (; j + U <= J; j += U) {
      double lj[U], sj[U], kj[U], wn[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t jj = j + u;
        lj[u] = L_[jj * f_cs]; sj[u] = S_[jj * f_cs]; kj[u] = K_[jj * f_cs];
        wn[u] = (jj + 1 < J) ? W_[(jj + 1) * w_cs] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t jj = j + u;
        double r = w0;
        if (jj == 0) r += bl;
        if (jj == J - 1) r += br;
        h = fma(-lj[u], h, r);
        const double cw = cth * w0, cwn = cth * wn[u];
        W_[jj * w_cs] = fma(-sj[u], cwn, fma(h, kj[u], -cw));
        w0 = wn[u];
      }
    }
    for (; j < J; ++j) {
      const double lj = L_[j * f_cs], sj = S_[j * f_cs], kj = K_[j * f_cs];
      const double wn = (j + 1 < J) ? W_[(j + 1) * w_cs] : 0.0;
      double r = w0;
      if (j == 0) r += bl;
      if (j == J - 1) r += br;
      h = fma(-lj, h, r);
      W_[j * w_cs] = fma(-sj, cth * wn, fma(h, kj, -cth * w0));
      w0 = wn;
    }
    // backward: z_j = t_j - s_j z_{j+1}   (s_{J-1} = 0: z_{J-1} = t_{J-1} is already in place)
    double z = W_[(J - 1) * w_cs];
    j = J - 2;
    for (; j - (U - 1) >= 0; j -= U) {
      double sj[U], tj[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { sj[u] = S_[(j - u) * f_cs]; tj[u] = W_[(j - u) * w_cs]; }
#pragma unroll
      for (int u = 0; u < U; ++u) { z = fma(-sj[u], z, tj[u]); W_[(j - u) * w_cs] = z; }
    }
    for (; j >= 0; --j) { z = fma(-S_[j * f_cs], z, W_[j * w_cs]); W_[j * w_cs] = z; }
  }
}

// ------------------------------------------------------------------------------------------
// Tiny meshes (J <= JR): the whole system lives in the thread's registers for all `nsteps` steps of
// the launch.  Per step only the three factor arrays are read (24 B per cell-update; s a second time
// in the backward sweep), and they are served by L1 after the first step (see the persistent loop);
// the state is read once at the start and written once at the end.
// Same arithmetic, operation for operation, as thomas_steps_kernel.
template <int JR>
__global__ void __launch_bounds__(128) thomas_reg_kernel(
    int J, int64_t P, const double* __restrict__ fp, const double* __restrict__ fq,
    const double* __restrict__ fb, int64_t f_ps, int64_t f_cs, double cth,
    const double* __restrict__ bt_left, const double* __restrict__ bt_right, double* __restrict__ w,
    int64_t w_ps, int64_t w_cs, int64_t nsteps) {
  // Persistent: the grid holds as many CTAs as keep the factors of the systems in flight inside the
  // SM's L1 (3*J*8 B per system: 128 systems of 64 cells = 192 KB), and every CTA walks over groups
  // of systems; the factors of a group then come from HBM once per LAUNCH, not once per step.
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
  const double* L_ = fp + p * f_ps;
  const double* S_ = fq + p * f_ps;
  const double* K_ = fb + p * f_ps;
  double* W_ = w + p * w_ps;
  const double bl = bt_left ? bt_left[p] : 0.0;
  const double br = bt_right ? bt_right[p] : 0.0;
  double x[JR];
#pragma unroll
  for (int j = 0; j < JR; ++j) x[j] = (j < J) ? W_[j * w_cs] : 0.0;
  const int64_t g8 = 8 * f_cs;
  for (int64_t s = 0; s < nsteps; ++s) {
    // Running pointers, passed through an empty asm at every group of 8 cells: the 3*JR addresses are
    // loop invariants that the compiler would otherwise compute once and keep in registers across the
    // steps (spilling the state), and loads are not hoisted beyond their group.
    const double *lp = L_, *sp = S_, *kp = K_;
    double h = 0.0;
#pragma unroll
    for (int g = 0; g < JR / 8; ++g) {
      asm volatile("" : "+l"(lp), "+l"(sp), "+l"(kp) :: "memory");
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = 8 * g + u;
        if (j < J) {
          const double lj = lp[u * f_cs], sj = sp[u * f_cs], kj = kp[u * f_cs];
          double r = x[j];
          if (j == 0) r += bl;
          if (j == J - 1) r += br;
          h = fma(-lj, h, r);
          const double wn = (j + 1 < JR) ? x[(j + 1) % JR] : 0.0;             // x[j+1] is 0 beyond the mesh
          x[j] = fma(-sj, cth * wn, fma(h, kj, -(cth * x[j])));
        }
      }
      lp += g8; sp += g8; kp += g8;
    }
    double z = 0.0;
#pragma unroll
    for (int g = JR / 8 - 1; g >= 0; --g) {
      sp -= g8;
      asm volatile("" : "+l"(sp) :: "memory");
#pragma unroll
      for (int u = 7; u >= 0; --u) {
        const int j = 8 * g + u;
        if (j < J) {
          z = fma(-sp[u * f_cs], z, x[j]);   // s_{J-1} = 0: the first link returns t_{J-1}
          x[j] = z;
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < JR; ++j)
    if (j < J) W_[j * w_cs] = x[j];
  }
}

template <int JR>
static void launch_thomas_reg(int64_t J, int64_t P, const double* fp, const double* fq, const double* fb,
                              int64_t f_ps, int64_t f_cs, double cth, const double* bt_left, const double* bt_right,
                              double* w, int64_t w_ps, int64_t w_cs, int64_t nsteps, cudaStream_t s) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int64_t blocks = (P + 127) / 128;
  const int64_t resident = (int64_t)sms * (JR <= 32 ? 2 : 1);
  if (blocks > resident) blocks = resident;
  thomas_reg_kernel<JR><<<(unsigned)blocks, 128, 0, s>>>((int)J, P, fp, fq, fb, f_ps, f_cs, cth, bt_left,
                                                                   bt_right, w, w_ps, w_cs, nsteps);
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
  if (J <= 64 && nsteps >= 4) {                  // tiny meshes, several steps: the system stays in registers and its
                                                 // factors in L1 for the whole launch (a single step streams better)
    if (J <= 32)
      fvm::launch_thomas_reg<32>(J, P, fp, fq, fb, f_ps, f_cs, cth, bt_left, bt_right, w, w_ps, w_cs, nsteps,
                                 (cudaStream_t)stream);
    else
      fvm::launch_thomas_reg<64>(J, P, fp, fq, fb, f_ps, f_cs, cth, bt_left, bt_right, w, w_ps, w_cs, nsteps,
                                 (cudaStream_t)stream);
    return (int)cudaGetLastError();
  }
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
