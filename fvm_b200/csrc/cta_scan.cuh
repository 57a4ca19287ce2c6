// CTA-wide first-order linear recurrences (the two sweeps of a Thomas solve with pre-factored
// coefficients) by a scan of affine maps: every thread owns CPT consecutive positions, reduces
// them to one map x -> A*x + B, the maps are scanned across the CTA (shuffles inside a warp,
// shared memory across warps), and a second pass applies the true carry.
#pragma once
#include <cuda_runtime.h>

namespace fvm {

constexpr unsigned kAll = 0xffffffffu;

struct Affine { double A, B; };   // x -> A*x + B

// later o earlier
__device__ __forceinline__ Affine compose(const Affine& later, const Affine& earlier) {
  Affine r;
  r.A = later.A * earlier.A;
  r.B = fma(later.A, earlier.B, later.B);
  return r;
}

// Exclusive scan of affine maps over the threads of a CTA, in thread order (REVERSE = false)
// or in reverse thread order (REVERSE = true).  Returns the VALUE entering this thread when the
// value entering the first thread of the order is `carry_in`; *total receives the map of the
// whole CTA (valid in every thread).  sm: 2*32 doubles of shared memory.
template <int TPB, bool REVERSE>
__device__ __forceinline__ double cta_scan_affine(Affine m, double carry_in, double* sm, Affine* total) {
  constexpr int NW = TPB / 32;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pos = REVERSE ? 31 - lane : lane;              // position inside the warp's order
  Affine inc = m;                                          // inclusive scan inside the warp
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    Affine o;
    o.A = REVERSE ? __shfl_down_sync(kAll, inc.A, d) : __shfl_up_sync(kAll, inc.A, d);
    o.B = REVERSE ? __shfl_down_sync(kAll, inc.B, d) : __shfl_up_sync(kAll, inc.B, d);
    if (pos >= d) inc = compose(inc, o);
  }
  Affine exc;                                              // exclusive inside the warp
  exc.A = REVERSE ? __shfl_down_sync(kAll, inc.A, 1) : __shfl_up_sync(kAll, inc.A, 1);
  exc.B = REVERSE ? __shfl_down_sync(kAll, inc.B, 1) : __shfl_up_sync(kAll, inc.B, 1);
  if (pos == 0) { exc.A = 1.0; exc.B = 0.0; }
  if (NW == 1) {
    if (total) {
      total->A = __shfl_sync(kAll, inc.A, REVERSE ? 0 : 31);
      total->B = __shfl_sync(kAll, inc.B, REVERSE ? 0 : 31);
    }
    return fma(exc.A, carry_in, exc.B);
  }
  __syncthreads();                                         // sm may still be in use by a previous scan
  if (pos == 31) { sm[warp] = inc.A; sm[32 + warp] = inc.B; }
  __syncthreads();
  if (warp == 0) {                                         // scan of the warp totals
    const int wi = REVERSE ? NW - 1 - lane : lane;         // warp index at this lane's position
    Affine t; t.A = 1.0; t.B = 0.0;
    if (lane < NW) { t.A = sm[wi]; t.B = sm[32 + wi]; }
    Affine ti = t;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      Affine o;
      o.A = __shfl_up_sync(kAll, ti.A, d);
      o.B = __shfl_up_sync(kAll, ti.B, d);
      if (lane >= d) ti = compose(ti, o);
    }
    Affine te;
    te.A = __shfl_up_sync(kAll, ti.A, 1);
    te.B = __shfl_up_sync(kAll, ti.B, 1);
    if (lane == 0) { te.A = 1.0; te.B = 0.0; }
    __syncwarp();
    if (lane < NW) { sm[wi] = te.A; sm[32 + wi] = te.B; }  // exclusive prefix of warp wi
    if (lane == NW - 1) { sm[64] = ti.A; sm[65] = ti.B; }  // map of the whole CTA
  }
  __syncthreads();
  Affine wp; wp.A = sm[warp]; wp.B = sm[32 + warp];
  if (total) { total->A = sm[64]; total->B = sm[65]; }
  const double into_warp = fma(wp.A, carry_in, wp.B);
  return fma(exc.A, into_warp, exc.B);
}

// g_c <- r_c - p(c) * g_{c-1} along the CTA's TPB*CPT positions (thread-blocked ownership:
// thread t owns positions [t*CPT, (t+1)*CPT)).  p(c) returns the coefficient of the thread's
// c-th cell (from registers or shared memory); it is evaluated in both passes.
template <int TPB, int CPT, class PF>
__device__ __forceinline__ void cta_forward(double (&g)[CPT], PF p, double* sm, double carry_in, Affine* total) {
  Affine m; m.A = 1.0; m.B = 0.0;
#pragma unroll
  for (int c = 0; c < CPT; ++c) { const double pc = p(c); m.B = fma(-pc, m.B, g[c]); m.A = -pc * m.A; }
  double prev = cta_scan_affine<TPB, false>(m, carry_in, sm, total);
#pragma unroll
  for (int c = 0; c < CPT; ++c) { prev = fma(-p(c), prev, g[c]); g[c] = prev; }
}

// x_c <- g_c - q_c * x_{c+1}, from the last position to the first, carry-in (x beyond the end).
template <int TPB, int CPT, class QF>
__device__ __forceinline__ void cta_backward(double (&x)[CPT], QF q, double* sm, double carry_in, Affine* total) {
  Affine m; m.A = 1.0; m.B = 0.0;
#pragma unroll
  for (int c = CPT - 1; c >= 0; --c) { const double qc = q(c); m.B = fma(-qc, m.B, x[c]); m.A = -qc * m.A; }
  double nxt = cta_scan_affine<TPB, true>(m, carry_in, sm, total);
#pragma unroll
  for (int c = CPT - 1; c >= 0; --c) { nxt = fma(-q(c), nxt, x[c]); x[c] = nxt; }
}


}  // namespace fvm
