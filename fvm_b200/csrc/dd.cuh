// Double-double arithmetic for the set-up phases (factorisation chains, spikes, separator
// systems).  A dd is an unevaluated sum hi + lo with |lo| <= ulp(hi)/2, i.e. about 106 bits.
//
// Why: a theta step applies the SAME prefactorised coefficients ten thousand times.  Their
// rounding errors are therefore systematic, and a coefficient that came out of an fp64 chain
// (pivot recurrence, products of 16-31 elimination factors, Schur complements of the separator
// system) carries a correlated error of several ulp that the undamped modes of Crank-Nicolson
// accumulate linearly in the step count: cfg 3's stiffest problems end 1e-10 from the exact
// solution of the reference's own fp64 operators after 10,000 steps.  With every chain carried
// in dd and every stored coefficient rounded ONCE, the same run ends ~1e-12 from it
// (tests/test_gpu_long_run.py).  The time loop itself stays plain fp64.
//
// "Sloppy" variants are used throughout: the relative error of each operation is a few 1e-32,
// twenty orders below what matters here, and no operand is ever inf/nan on the valid path.
#pragma once
#include <cuda_runtime.h>

namespace fvm {

struct dd { double hi, lo; };

#define FVM_HD __host__ __device__ __forceinline__

FVM_HD dd dd_make(double hi, double lo = 0.0) { dd r; r.hi = hi; r.lo = lo; return r; }
FVM_HD double dd_round(const dd& a) { return a.hi + a.lo; }       // nearest double (hi + lo is normalised)
FVM_HD dd dd_neg(const dd& a) { return dd_make(-a.hi, -a.lo); }

FVM_HD double dd_fma(double a, double b, double c) {
#ifdef __CUDA_ARCH__
  return __fma_rn(a, b, c);
#else
  return __builtin_fma(a, b, c);
#endif
}
// the *_rn forms keep nvcc from contracting the error-free transformations into FMAs
FVM_HD double dd_addr(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dadd_rn(a, b);
#else
  volatile double s = a + b; return s;
#endif
}
FVM_HD double dd_subr(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dsub_rn(a, b);
#else
  volatile double s = a - b; return s;
#endif
}
FVM_HD double dd_mulr(double a, double b) {
#ifdef __CUDA_ARCH__
  return __dmul_rn(a, b);
#else
  volatile double s = a * b; return s;
#endif
}

// s + e = a + b exactly (Knuth)
FVM_HD dd two_sum(double a, double b) {
  const double s = dd_addr(a, b);
  const double bb = dd_subr(s, a);
  const double e = dd_addr(dd_subr(a, dd_subr(s, bb)), dd_subr(b, bb));
  return dd_make(s, e);
}
// s + e = a + b exactly when |a| >= |b| (Dekker)
FVM_HD dd quick_two_sum(double a, double b) {
  const double s = dd_addr(a, b);
  return dd_make(s, dd_subr(b, dd_subr(s, a)));
}
// p + e = a * b exactly
FVM_HD dd two_prod(double a, double b) {
  const double p = dd_mulr(a, b);
  return dd_make(p, dd_fma(a, b, -p));
}

FVM_HD dd dd_add(const dd& a, const dd& b) {
  dd s = two_sum(a.hi, b.hi);
  s.lo = dd_addr(s.lo, dd_addr(a.lo, b.lo));
  return quick_two_sum(s.hi, s.lo);
}
FVM_HD dd dd_add_d(const dd& a, double b) {
  dd s = two_sum(a.hi, b);
  s.lo = dd_addr(s.lo, a.lo);
  return quick_two_sum(s.hi, s.lo);
}
FVM_HD dd dd_sub(const dd& a, const dd& b) { return dd_add(a, dd_neg(b)); }

FVM_HD dd dd_mul(const dd& a, const dd& b) {
  dd p = two_prod(a.hi, b.hi);
  p.lo = dd_fma(a.hi, b.lo, dd_fma(a.lo, b.hi, p.lo));
  return quick_two_sum(p.hi, p.lo);
}
FVM_HD dd dd_mul_d(const dd& a, double b) {
  dd p = two_prod(a.hi, b);
  p.lo = dd_fma(a.lo, b, p.lo);
  return quick_two_sum(p.hi, p.lo);
}
FVM_HD dd dd_mul_dd(double a, double b) { return two_prod(a, b); }

// 1 / a: fp64 reciprocal refined by one Newton step carried in dd
FVM_HD dd dd_recip(const dd& a) {
  const double x0 = 1.0 / a.hi;
  // r = 1 - a * x0 in dd; x = x0 + x0 * r
  dd ax = dd_mul_d(a, x0);
  dd r = dd_add_d(dd_neg(ax), 1.0);
  dd corr = dd_mul_d(r, x0);           // |r| ~ 1e-16: the neglected x0 r^2 is ~1e-32 relative
  return dd_add_d(corr, x0);
}
FVM_HD dd dd_div(const dd& a, const dd& b) { return dd_mul(a, dd_recip(b)); }

// a - b * c, the link of every elimination / substitution chain
FVM_HD dd dd_nmul_add(const dd& b, const dd& c, const dd& a) { return dd_sub(a, dd_mul(b, c)); }
FVM_HD dd dd_nmul_add_d(const dd& b, double c, double a) { return dd_add_d(dd_neg(dd_mul_d(b, c)), a); }

#ifdef __CUDACC__
__device__ __forceinline__ dd dd_shfl_up(const dd& x, int s) {
  return dd_make(__shfl_up_sync(0xffffffffu, x.hi, s), __shfl_up_sync(0xffffffffu, x.lo, s));
}
__device__ __forceinline__ dd dd_shfl_dn(const dd& x, int s) {
  return dd_make(__shfl_down_sync(0xffffffffu, x.hi, s), __shfl_down_sync(0xffffffffu, x.lo, s));
}
#endif

}  // namespace fvm
