// Device-side building blocks shared by every kernel: exactly-rounded fp64 helpers,
// the modified diffusivity, and one row of M = coefficient_matrix().
//
// The row arithmetic keeps the reference's operation order and uses the *_rn
// intrinsics (which nvcc never contracts into FMAs), so central/upwind rows are
// bit-identical to numpy's [fvm_classes.py:81-87,149-177].
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/fvm_b200.h"

namespace fvm {

__device__ __forceinline__ double mul(double x, double y) { return __dmul_rn(x, y); }
__device__ __forceinline__ double add(double x, double y) { return __dadd_rn(x, y); }
__device__ __forceinline__ double sub(double x, double y) { return __dsub_rn(x, y); }
__device__ __forceinline__ double dvd(double x, double y) { return __ddiv_rn(x, y); }

// kappa = coth(mu/2) - 2/mu, the exponential-fitting weight [fvm_classes.py:117-120],
// evaluated without overflow (the reference's form is inf/inf for mu >= 709.8) and
// without cancellation (series below |mu| = 0.5).  Diagnostic output only: D below does
// not go through kappa.
__device__ __forceinline__ double kappa_exponential(double mu) {
  if (mu == 0.0) return 0.0;                     // :118
  if (isinf(mu)) return mu > 0 ? 1.0 : -1.0;     // :119-120
  double x = 0.5 * mu;
  if (fabs(mu) < 0.5) {
    double x2 = x * x;
    double s = -3617.0 / 325641566250.0;
    s = fma(s, x2, 4.0 / 18243225.0);
    s = fma(s, x2, -1382.0 / 638512875.0);
    s = fma(s, x2, 2.0 / 93555.0);
    s = fma(s, x2, -1.0 / 4725.0);
    s = fma(s, x2, 2.0 / 945.0);
    s = fma(s, x2, -1.0 / 45.0);
    s = fma(s, x2, 1.0 / 3.0);
    return s * x;
  }
  return 1.0 / tanh(x) - 1.0 / x;
}

// kappa [fvm_classes.py:116-126] and D = d + a h kappa / 2 [:131] of one cell.
// central / upwind follow the reference bit for bit; exponential uses
// D = d * (B(mu) + mu/2), B(mu) = mu / expm1(mu), B(0) = 1.
__device__ __forceinline__ void cell_kappa_D(int scheme, double a, double d, double h,
                                             double& kappa, double& D) {
  if (scheme == FVM_EXPONENTIAL) {
    double mu = dvd(mul(a, h), d);               // peclet_number(), :135-136
    kappa = kappa_exponential(mu);
    if (mu == 0.0) {
      D = d;
    } else if (isinf(mu)) {
      D = add(d, mul(mul(mul(0.5, a), h), kappa));
    } else {
      double B = mu / expm1(mu);                 // expm1 -> inf gives B = 0 (pure upwind limit)
      D = d * (B + 0.5 * mu);
    }
  } else {
    kappa = 0.0;                                 // :126
    if (scheme == FVM_UPWIND) kappa = (a > 0.0) ? 1.0 : ((a < 0.0) ? -1.0 : 0.0);   // :122-124
    D = add(d, mul(mul(mul(0.5, a), h), kappa)); // :131, left to right
  }
}

struct Row { double lo, di, up; };

// Row j of M.  f[0..3] = faces[j-1], faces[j], faces[j+1], faces[j+2] (entries outside
// the mesh are never used); am/a0/ap and Dm/D0/Dp are a and the MODIFIED diffusivity at
// cells j-1, j, j+1.  kind = 0 interior, 1 first row, 2 last row.
// Face interpolation is CellVariable.m/p [fvm_classes.py:83,87]; rows are :151-153
// (interior), :158-159 / :169-170 (Robin) and :200-201 / :214-215 (Dirichlet).
__device__ __forceinline__ Row matrix_row(const double f[4], double am, double a0, double ap,
                                          double Dm, double D0, double Dp, bool first, bool last,
                                          int left_kind, int right_kind) {
  Row r;
  if (first && left_kind == FVM_BC_DIRICHLET) { r.lo = 0.0; r.di = 0.0; r.up = 1.0; return r; }
  if (last && right_kind == FVM_BC_DIRICHLET) { r.lo = 0.0; r.di = 1.0; r.up = 0.0; return r; }
  const double h0 = sub(f[2], f[1]);
  const double c0 = mul(0.5, add(f[1], f[2]));
  const double rh = dvd(1.0, h0);
  double l_adv = 0.0, l_dif = 0.0, l_advd = 0.0;   // left-face terms
  double r_adv = 0.0, r_dif = 0.0, r_advd = 0.0;   // right-face terms
  if (!first) {
    const double hm = sub(f[1], f[0]);
    const double cm = mul(0.5, add(f[0], f[1]));
    const double dx = sub(c0, cm);                                  // Mesh.hm(j), :51
    const double two_dx = mul(2.0, dx);
    const double af = add(mul(dvd(h0, two_dx), am), mul(dvd(hm, two_dx), a0));   // a.m(j), :83
    const double Df = add(mul(dvd(h0, two_dx), Dm), mul(dvd(hm, two_dx), D0));   // d.m(j)
    l_adv = dvd(mul(af, h0), two_dx);     // a.m(j)*h(j)/(2 hm(j))      -> lo
    l_advd = dvd(mul(af, hm), two_dx);    // a.m(j)*h(j-1)/(2 hm(j))    -> di
    l_dif = dvd(Df, dx);                  // d.m(j)/hm(j)
  }
  if (!last) {
    const double hp = sub(f[3], f[2]);
    const double cp = mul(0.5, add(f[2], f[3]));
    const double dx = sub(cp, c0);                                  // Mesh.hp(j), :57
    const double two_dx = mul(2.0, dx);
    const double af = add(mul(dvd(hp, two_dx), a0), mul(dvd(h0, two_dx), ap));   // a.p(j), :87
    const double Df = add(mul(dvd(hp, two_dx), D0), mul(dvd(h0, two_dx), Dp));   // d.p(j)
    r_adv = dvd(mul(-af, h0), two_dx);    // -a.p(j)*h(j)/(2 hp(j))     -> up
    r_advd = dvd(mul(af, hp), two_dx);    // a.p(j)*h(j+1)/(2 hp(j))    -> di
    r_dif = dvd(Df, dx);                  // d.p(j)/hp(j)
  }
  if (first) {                             // Robin, :158-159
    r.lo = 0.0;
    r.di = mul(rh, sub(-r_advd, r_dif));
    r.up = mul(rh, add(r_adv, r_dif));
  } else if (last) {                       // Robin, :169-170
    r.lo = mul(rh, add(l_adv, l_dif));
    r.di = mul(rh, sub(l_advd, l_dif));
    r.up = 0.0;
  } else {                                 // interior, :151-153
    r.lo = mul(rh, add(l_adv, l_dif));
    r.di = mul(rh, sub(sub(sub(l_advd, r_advd), l_dif), r_dif));
    r.up = mul(rh, add(r_adv, r_dif));
  }
  return r;
}

// Modified diffusivity of cell j of problem p (0 outside the mesh, where it is never used).
__device__ __forceinline__ double cell_D(const fvm_problem& pr, int64_t p, int64_t j) {
  if (j < 0 || j >= pr.J) return 0.0;
  double kap, D;
  cell_kappa_D(pr.scheme, pr.a[p * pr.a_ps + j * pr.a_cs], pr.d[p * pr.d_ps + j * pr.d_cs],
               sub(pr.faces[j + 1], pr.faces[j]), kap, D);
  return D;
}

// Row j of M given the modified diffusivities of cells j-1, j, j+1 (a sweep over consecutive
// cells keeps them in a sliding window instead of evaluating expm1 three times per row).
__device__ __forceinline__ Row assemble_row_D(const fvm_problem& pr, int64_t p, int64_t j,
                                              double Dm, double D0, double Dp) {
  const int64_t J = pr.J;
  const bool first = (j == 0), last = (j == J - 1);
  double f[4];
  f[0] = first ? 0.0 : pr.faces[j - 1];
  f[1] = pr.faces[j];
  f[2] = pr.faces[j + 1];
  f[3] = last ? 0.0 : pr.faces[j + 2];
  const double* ap_ = pr.a + p * pr.a_ps;
  const double a0 = ap_[j * pr.a_cs];
  const double am = first ? 0.0 : ap_[(j - 1) * pr.a_cs];
  const double ap = last ? 0.0 : ap_[(j + 1) * pr.a_cs];
  Row r = matrix_row(f, am, a0, ap, Dm, D0, Dp, first, last, pr.left_kind, pr.right_kind);
  if (pr.reaction != nullptr) {
    const bool dirichlet_row = (first && pr.left_kind == FVM_BC_DIRICHLET) ||
                               (last && pr.right_kind == FVM_BC_DIRICHLET);
    if (!dirichlet_row) r.di = sub(r.di, pr.reaction[p * pr.r_ps + j * pr.r_cs]);
  }
  return r;
}

// Loads what row j of problem p needs and returns the row of M (with the optional
// reaction term folded into the diagonal); also kappa and D of cell j.
__device__ __forceinline__ Row assemble_row(const fvm_problem& pr, int64_t p, int64_t j,
                                            double& kappa0, double& D0) {
  const int64_t J = pr.J;
  const bool first = (j == 0), last = (j == J - 1);
  double f[4];
  f[0] = first ? 0.0 : pr.faces[j - 1];
  f[1] = pr.faces[j];
  f[2] = pr.faces[j + 1];
  f[3] = last ? 0.0 : pr.faces[j + 2];
  const double* ap_ = pr.a + p * pr.a_ps;
  const double* dp_ = pr.d + p * pr.d_ps;
  const double a0 = ap_[j * pr.a_cs];
  const double d0 = dp_[j * pr.d_cs];
  cell_kappa_D(pr.scheme, a0, d0, sub(f[2], f[1]), kappa0, D0);
  double am = 0.0, Dm = 0.0, ap = 0.0, Dp = 0.0, kk;
  if (!first) {
    am = ap_[(j - 1) * pr.a_cs];
    cell_kappa_D(pr.scheme, am, dp_[(j - 1) * pr.d_cs], sub(f[1], f[0]), kk, Dm);
  }
  if (!last) {
    ap = ap_[(j + 1) * pr.a_cs];
    cell_kappa_D(pr.scheme, ap, dp_[(j + 1) * pr.d_cs], sub(f[3], f[2]), kk, Dp);
  }
  Row r = matrix_row(f, am, a0, ap, Dm, D0, Dp, first, last, pr.left_kind, pr.right_kind);
  if (pr.reaction != nullptr) {
    const bool dirichlet_row = (first && pr.left_kind == FVM_BC_DIRICHLET) ||
                               (last && pr.right_kind == FVM_BC_DIRICHLET);
    if (!dirichlet_row) r.di = sub(r.di, pr.reaction[p * pr.r_ps + j * pr.r_cs]);
  }
  return r;
}

// Row j of A = I - (dt*theta*alpha_j) M, rounded like scipy rounds it
// [fvm_classes.py:376]: s = (dt*theta)*alpha_j; off-diagonals -(s*m); diagonal 1 - s*m.
__device__ __forceinline__ Row theta_lhs_row(const Row& m, double dt_theta, double alpha) {
  const double s = mul(dt_theta, alpha);
  Row a;
  a.lo = -mul(s, m.lo);
  a.di = sub(1.0, mul(s, m.di));
  a.up = -mul(s, m.up);
  return a;
}

__device__ __forceinline__ double alpha_of(int64_t j, int64_t J, int alpha_left, int alpha_right) {
  if (j == 0) return (double)alpha_left;
  if (j == J - 1) return (double)alpha_right;
  return 1.0;
}

}  // namespace fvm
