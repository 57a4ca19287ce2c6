// fvm_assemble + fvm_theta_factor: once-per-model work (not on the per-step path).
#include "fvm_device.cuh"
#include "dd.cuh"

namespace fvm {

// One thread per (problem, cell).  The fast index follows the output's unit stride so
// that the stores coalesce for both the problem-major and the interleaved layout.
__global__ void __launch_bounds__(256) assemble_kernel(fvm_problem pr, double* lo, double* di,
                                                       double* up, double* kappa, double* D,
                                                       int64_t out_ps, int64_t out_cs,
                                                       double* beta_left, double* beta_right,
                                                       int cell_fast) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n = pr.J * pr.P;
  if (tid >= n) return;
  int64_t p, j;
  if (cell_fast) { p = tid / pr.J; j = tid - p * pr.J; }
  else           { j = tid / pr.P; p = tid - j * pr.P; }
  double kap, Dc;
  Row r = assemble_row(pr, p, j, kap, Dc);
  const int64_t o = p * out_ps + j * out_cs;
  lo[o] = r.lo; di[o] = r.di; up[o] = r.up;
  if (kappa) kappa[o] = kap;
  if (D) D[o] = Dc;
  // beta_vector(): Robin entries flux/h at the ends, 0 for Dirichlet [fvm_classes.py:182,188,223,230]
  if (j == 0 && beta_left) {
    double b = 0.0;
    if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
      b = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
    beta_left[p] = b;
  }
  if (j == pr.J - 1 && beta_right) {
    double b = 0.0;
    if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
      b = dvd(-pr.right_flux[p], sub(pr.faces[pr.J], pr.faces[pr.J - 1]));
    beta_right[p] = b;
  }
}

// One thread per problem: the factorisation recurrence is sequential in j.
__global__ void __launch_bounds__(128) theta_factor_kernel(
    int64_t J, int64_t P, const double* __restrict__ lo, const double* __restrict__ di,
    const double* __restrict__ up, int64_t m_ps, int64_t m_cs, const double* __restrict__ dt,
    double theta, int alpha_left, int alpha_right, const double* __restrict__ beta_left,
    const double* __restrict__ beta_right, double* __restrict__ fp, double* __restrict__ fq,
    double* __restrict__ fb, int64_t f_ps, int64_t f_cs, double* __restrict__ bt_left,
    double* __restrict__ bt_right) {
  const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double dt_theta = mul(dt[p], theta);
  const double rth = 1.0 / theta;
  // The pivot recurrence runs in double-double and every factor is rounded ONCE: the step
  // kernel applies the same factors at every step, so a chain's correlated rounding error would
  // grow linearly with the step count (dd.cuh).
  dd s_prev = dd_make(0.0), inv_prev = dd_make(0.0);
  for (int64_t j = 0; j < J; ++j) {
    Row m;
    const int64_t o = p * m_ps + j * m_cs;
    m.lo = lo[o]; m.di = di[o]; m.up = up[o];
    Row a = theta_lhs_row(m, dt_theta, alpha_of(j, J, alpha_left, alpha_right));
    const dd inv = dd_recip(dd_add_d(dd_neg(dd_mul_d(s_prev, a.lo)), a.di));
    const int64_t f = p * f_ps + j * f_cs;
    fp[f] = dd_round(dd_mul_d(inv_prev, a.lo));      // l_j     = A_j,j-1 / pivot_{j-1}
    fb[f] = dd_round(dd_mul_d(inv, rth));            // kappa_j = 1 / (theta pivot_j)
    s_prev = dd_mul_d(inv, a.up);
    fq[f] = dd_round(s_prev);                        // s_j     = A_j,j+1 / pivot_j
    inv_prev = inv;
  }
  // beta enters the elimination un-normalised: A yhat = w/theta + beta, h = theta * (that)
  if (bt_left) bt_left[p] = beta_left ? theta * beta_left[p] : 0.0;
  if (bt_right) bt_right[p] = beta_right ? theta * beta_right[p] : 0.0;
}

static int check_problem(const fvm_problem* pr) {
  if (!pr || !pr->faces || !pr->a || !pr->d) return FVM_ERR_INVALID_ARG;
  if (pr->J < 3 || pr->P < 1) return FVM_ERR_INVALID_ARG;
  if (pr->scheme < FVM_CENTRAL || pr->scheme > FVM_EXPONENTIAL) return FVM_ERR_INVALID_ARG;
  if ((pr->left_kind | pr->right_kind) & ~1) return FVM_ERR_INVALID_ARG;
  if ((pr->alpha_left | pr->alpha_right) & ~1) return FVM_ERR_INVALID_ARG;
  return FVM_OK;
}

}  // namespace fvm

extern "C" int fvm_assemble(const fvm_problem* prob, double* lo, double* di, double* up,
                            double* kappa, double* D, int64_t out_ps, int64_t out_cs,
                            double* beta_left, double* beta_right, void* stream) {
  int st = fvm::check_problem(prob);
  if (st != FVM_OK) return st;
  if (!lo || !di || !up) return FVM_ERR_INVALID_ARG;
  const int64_t n = prob->J * prob->P;
  const int64_t blocks = (n + 255) / 256;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  const int cell_fast = (out_cs <= out_ps) ? 1 : 0;
  fvm::assemble_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      *prob, lo, di, up, kappa, D, out_ps, out_cs, beta_left, beta_right, cell_fast);
  return (int)cudaGetLastError();
}

extern "C" int fvm_theta_factor(int64_t J, int64_t P, const double* lo, const double* di,
                                const double* up, int64_t m_ps, int64_t m_cs, const double* dt,
                                double theta, int32_t alpha_left, int32_t alpha_right,
                                const double* beta_left, const double* beta_right, double* fp,
                                double* fq, double* fb, int64_t f_ps, int64_t f_cs,
                                double* bt_left, double* bt_right, void* stream) {
  if (!lo || !di || !up || !dt || !fp || !fq || !fb) return FVM_ERR_INVALID_ARG;
  if (J < 3 || P < 1 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if ((alpha_left | alpha_right) & ~1) return FVM_ERR_INVALID_ARG;
  const int64_t blocks = (P + 127) / 128;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  fvm::theta_factor_kernel<<<(unsigned)blocks, 128, 0, (cudaStream_t)stream>>>(
      J, P, lo, di, up, m_ps, m_cs, dt, theta, alpha_left, alpha_right, beta_left, beta_right,
      fp, fq, fb, f_ps, f_cs, bt_left, bt_right);
  return (int)cudaGetLastError();
}
