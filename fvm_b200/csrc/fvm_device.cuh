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

// What the face between cells f and f+1 contributes to the two rows it touches
// [fvm_classes.py:151-153 through CellVariable.m/p, :83,87]:
//   A_hi = a_face*h(f+1)/(2 dx)   (lo of row f+1; enters di of row f with a minus sign)
//   A_lo = a_face*h(f)  /(2 dx)   (di of row f+1; -A_lo is the advective part of up of row f)
//   Dd   = D_face/dx
// with dx = cells[f+1]-cells[f] (Mesh.hp(f) = Mesh.hm(f+1)) and the face values interpolated
// as h(f+1)/(2dx)*v[f] + h(f)/(2dx)*v[f+1].  Every operation is rounded as numpy rounds it.
struct Face { double A_hi, A_lo, Dd; };

// ff[0..2] = faces[f], faces[f+1], faces[f+2]; (a0, D0) belong to cell f, (a1, D1) to cell f+1.
__device__ __forceinline__ Face face_terms(const double ff[3], double a0, double a1, double D0, double D1) {
  const double h0 = sub(ff[1], ff[0]), h1 = sub(ff[2], ff[1]);
  const double c0 = mul(0.5, add(ff[0], ff[1])), c1 = mul(0.5, add(ff[1], ff[2]));
  const double dx = sub(c1, c0);
  const double two_dx = mul(2.0, dx);
  const double wl = dvd(h1, two_dx), wr = dvd(h0, two_dx);
  const double af = add(mul(wl, a0), mul(wr, a1));
  const double Df = add(mul(wl, D0), mul(wr, D1));
  Face F;
  F.A_hi = dvd(mul(af, h1), two_dx);
  F.A_lo = dvd(mul(af, h0), two_dx);
  F.Dd = dvd(Df, dx);
  return F;
}

// Row j of M from its two faces (L = face j-1, R = face j; the one outside the mesh is ignored).
// rh = 1/h(j).  Interior rows :151-153, Robin :158-159 / :169-170, Dirichlet :200-201 / :214-215.
__device__ __forceinline__ Row row_from_faces(double rh, const Face& L, const Face& R, bool first, bool last,
                                              int left_kind, int right_kind) {
  Row r;
  if (first && left_kind == FVM_BC_DIRICHLET) { r.lo = 0.0; r.di = 0.0; r.up = 1.0; return r; }
  if (last && right_kind == FVM_BC_DIRICHLET) { r.lo = 0.0; r.di = 1.0; r.up = 0.0; return r; }
  if (first) {
    r.lo = 0.0;
    r.di = mul(rh, sub(-R.A_hi, R.Dd));
    r.up = mul(rh, add(-R.A_lo, R.Dd));
  } else if (last) {
    r.lo = mul(rh, add(L.A_hi, L.Dd));
    r.di = mul(rh, sub(L.A_lo, L.Dd));
    r.up = 0.0;
  } else {
    r.lo = mul(rh, add(L.A_hi, L.Dd));
    r.di = mul(rh, sub(sub(sub(L.A_lo, R.A_hi), L.Dd), R.Dd));
    r.up = mul(rh, add(-R.A_lo, R.Dd));
  }
  return r;
}

// Row j of M.  f[0..3] = faces[j-1], faces[j], faces[j+1], faces[j+2] (entries outside
// the mesh are never used); am/a0/ap and Dm/D0/Dp are a and the MODIFIED diffusivity at
// cells j-1, j, j+1.
__device__ __forceinline__ Row matrix_row(const double f[4], double am, double a0, double ap,
                                          double Dm, double D0, double Dp, bool first, bool last,
                                          int left_kind, int right_kind) {
  Face L, R;
  L.A_hi = L.A_lo = L.Dd = 0.0; R = L;
  if (!first) L = face_terms(f, am, a0, Dm, D0);
  if (!last) R = face_terms(f + 1, a0, ap, D0, Dp);
  return row_from_faces(dvd(1.0, sub(f[2], f[1])), L, R, first, last, left_kind, right_kind);
}

// The face between cells f and f+1 of problem p (zeros outside the mesh), for sweeps that keep
// faces in a sliding window: D0, D1 = modified diffusivities of cells f and f+1.
__device__ __forceinline__ Face problem_face(const fvm_problem& pr, int64_t p, int64_t f, double D0, double D1) {
  Face F; F.A_hi = F.A_lo = F.Dd = 0.0;
  if (f < 0 || f + 1 >= pr.J) return F;
  const double ff[3] = {pr.faces[f], pr.faces[f + 1], pr.faces[f + 2]};
  const double* ap_ = pr.a + p * pr.a_ps;
  return face_terms(ff, ap_[f * pr.a_cs], ap_[(f + 1) * pr.a_cs], D0, D1);
}

// Row j of problem p from its two faces, with the optional reaction term.
__device__ __forceinline__ Row problem_row(const fvm_problem& pr, int64_t p, int64_t j, const Face& L, const Face& R) {
  const bool first = (j == 0), last = (j == pr.J - 1);
  Row r = row_from_faces(dvd(1.0, sub(pr.faces[j + 1], pr.faces[j])), L, R, first, last, pr.left_kind, pr.right_kind);
  if (pr.reaction != nullptr) {
    const bool dirichlet_row = (first && pr.left_kind == FVM_BC_DIRICHLET) ||
                               (last && pr.right_kind == FVM_BC_DIRICHLET);
    if (!dirichlet_row) r.di = sub(r.di, pr.reaction[p * pr.r_ps + j * pr.r_cs]);
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
