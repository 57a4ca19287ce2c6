// Variant 2 for 1024 < J <= 8192: one CTA per system, everything on chip between checkpoints.
//
// The warp-per-system kernels (fused.cu) stop at 1024 cells (32 lanes x 32 cells of registers).
// Here a CTA of TPB threads owns one system; thread t keeps its CPT consecutive cells of the state
// in registers for the whole launch, and the un-normalised LU factors of A (l, kappa, s: the same
// step WITHOUT a right-hand side as the other kernels, A^-1 B = A^-1/theta - c I) live in shared
// memory in tiled order ([c][t]: conflict-free), for CPT = 8 also in registers.  A step is
//   h_c = g_c + theta*beta_c - l_c h_{c-1}        forward recurrence over all TPB*CPT cells
//   u_c = h_c - s"_c u_{c+1}                       backward recurrence on u = yhat / kappa,
//                                                  s"_c = s_c kappa_{c+1} / kappa_c
//   w'_c = kappa_c u_c - c g_c
// (five operations per cell, and no cell needs its neighbour's old value) with both recurrences as
// CTA-wide scans of affine maps x -> A x + B.  The A halves of those maps are products of l (of s") and
// do not change from step to step: they are scanned ONCE at set-up (in double-double, rounded once), so
// a step only scans the B halves -- one shuffle and one FMA per level, one barrier per scan (every warp
// folds the warp totals itself), two barriers per step in all.
//
// Set-up inside the launch: rows of A (fvm_device.cuh), then the pivot recurrence
// s_j = A_j,j+1 / (A_jj - A_j,j-1 s_{j-1}) in double-double.  It is a Moebius recurrence, i.e. a product
// of 2x2 matrices [[0, up], [-lo, di]] acting on (numerator, denominator): every thread multiplies the
// matrices of its own cells, the products are scanned across the CTA (normalised by powers of two, which
// is exact), and each thread then runs the recurrence over its own cells from its exact starting value.
// HBM traffic: J doubles in, J out, J per checkpoint.
#include "fvm_device.cuh"
#include "dd.cuh"

namespace fvm {

namespace {

constexpr unsigned kAllLanes = 0xffffffffu;

struct M2 { dd a, b, c, d; };                       // [[a, b], [c, d]]

__device__ __forceinline__ dd dd_scale(const dd& x, double p2) { return dd_make(x.hi * p2, x.lo * p2); }   // p2 = 2^k: exact

__device__ __forceinline__ void m2_normalise(M2& m) {
  const double mx = fmax(fmax(fabs(m.a.hi), fabs(m.b.hi)), fmax(fabs(m.c.hi), fabs(m.d.hi)));
  if (mx == 0.0) return;
  int e;
  frexp(mx, &e);
  const double p2 = ldexp(1.0, -e);
  m.a = dd_scale(m.a, p2); m.b = dd_scale(m.b, p2); m.c = dd_scale(m.c, p2); m.d = dd_scale(m.d, p2);
}

// later x earlier
__device__ __forceinline__ M2 m2_mul(const M2& l, const M2& e) {
  M2 r;
  r.a = dd_add(dd_mul(l.a, e.a), dd_mul(l.b, e.c));
  r.b = dd_add(dd_mul(l.a, e.b), dd_mul(l.b, e.d));
  r.c = dd_add(dd_mul(l.c, e.a), dd_mul(l.d, e.c));
  r.d = dd_add(dd_mul(l.c, e.b), dd_mul(l.d, e.d));
  m2_normalise(r);
  return r;
}

__device__ __forceinline__ M2 m2_shfl_up(const M2& m, int d) {
  M2 r;
  r.a = dd_shfl_up(m.a, d); r.b = dd_shfl_up(m.b, d); r.c = dd_shfl_up(m.c, d); r.d = dd_shfl_up(m.d, d);
  return r;
}

__device__ __forceinline__ M2 m2_identity() {
  M2 r; r.a = dd_make(1.0); r.b = dd_make(0.0); r.c = dd_make(0.0); r.d = dd_make(1.0);
  return r;
}

__device__ __forceinline__ void m2_store(double* s, const M2& m) {
  s[0] = m.a.hi; s[1] = m.a.lo; s[2] = m.b.hi; s[3] = m.b.lo; s[4] = m.c.hi; s[5] = m.c.lo; s[6] = m.d.hi; s[7] = m.d.lo;
}
__device__ __forceinline__ M2 m2_load(const double* s) {
  M2 m;
  m.a = dd_make(s[0], s[1]); m.b = dd_make(s[2], s[3]); m.c = dd_make(s[4], s[5]); m.d = dd_make(s[6], s[7]);
  return m;
}

// Step-invariant halves of one direction's scan, per thread.
struct ScanConsts {
  double lvl[5];      // this lane's accumulated A at the start of level d of the in-warp scan
  double exc;         // product of the A's of the lanes before this one (in scan order)
};

// In-warp scan of the A halves (thread totals a_t), in double-double; REVERSE: scan order = descending lanes.
// Returns the warp's total (valid in the last lane of the order).
template <bool REVERSE>
__device__ __forceinline__ dd scan_consts(dd a_t, ScanConsts& K) {
  const int lane = threadIdx.x & 31;
  const int pos = REVERSE ? 31 - lane : lane;
  dd inc = a_t;
  int li = 0;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1, ++li) {
    K.lvl[li] = dd_round(inc);
    const dd o = REVERSE ? dd_shfl_dn(inc, d) : dd_shfl_up(inc, d);
    if (pos >= d) inc = dd_mul(inc, o);
  }
  const dd e = REVERSE ? dd_shfl_dn(inc, 1) : dd_shfl_up(inc, 1);
  K.exc = (pos == 0) ? 1.0 : dd_round(e);
  return inc;
}

// One step's scan of the B halves: b_t = this thread's value with a zero carry-in.  Returns the value
// entering this thread.  tot: shared array of NW doubles for this direction (warp totals of B);
// k2: the step-invariant halves of the SECOND level (the scan over the warp totals, lane i standing for
// warp i, identity beyond NW), [6][32] in shared memory: every warp scans the totals itself, in five
// shuffle levels instead of a serial fold over up to NW-1 warps.  ONE barrier.
template <int NW, bool REVERSE>
__device__ __forceinline__ double scan_values(double b_t, const ScanConsts& K, double* tot, const double* k2) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pos = REVERSE ? 31 - lane : lane;
  double inc = b_t;
  int li = 0;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1, ++li) {
    const double o = REVERSE ? __shfl_down_sync(kAllLanes, inc, d) : __shfl_up_sync(kAllLanes, inc, d);
    if (pos >= d) inc = fma(K.lvl[li], o, inc);
  }
  double exc = REVERSE ? __shfl_down_sync(kAllLanes, inc, 1) : __shfl_up_sync(kAllLanes, inc, 1);
  if (pos == 0) exc = 0.0;
  if (pos == 31) tot[warp] = inc;
  __syncthreads();
  double inc2 = (lane < NW) ? tot[lane] : 0.0;
  li = 0;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1, ++li) {
    const double o = REVERSE ? __shfl_down_sync(kAllLanes, inc2, d) : __shfl_up_sync(kAllLanes, inc2, d);
    if (pos >= d) inc2 = fma(k2[li * 32 + lane], o, inc2);
  }
  // value entering warp `warp` = inclusive result of the warp before it in scan order
  const int src = REVERSE ? warp + 1 : warp - 1;
  double cw = __shfl_sync(kAllLanes, inc2, src & 31);
  if (REVERSE ? (warp == NW - 1) : (warp == 0)) cw = 0.0;
  return fma(K.exc, cw, exc);
}

}  // namespace

template <int TPB, int CPT, int MODE>
__global__ void __launch_bounds__(TPB, CPT == 16 ? (TPB == 256 ? 2 : (TPB == 128 ? 4 : 1)) : 1)
fused_cta_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
                 int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt) {
  constexpr int T = TPB * CPT;
  constexpr int NW = TPB / 32;
  constexpr bool REGC = CPT <= 8;                   // l and s also in registers
  extern __shared__ double smem[];
  double* sl = smem;                 // [c*TPB + t]: rows of A (lo, di, up) during set-up, then l, kappa, s
  double* sk = sl + T;
  double* ss = sk + T;
  double* totf = ss + T;             // [NW] warp totals of the forward scan (B halves), per step
  double* totb = totf + NW;          // [NW] ... of the backward scan
  double* awf = totb + NW;           // [NW] warp totals of the A halves, fixed
  double* awb = awf + NW;            // [NW]
  double* k2f = awb + NW;            // [6][32] second-level scan constants, forward
  double* k2b = k2f + 6 * 32;        // [6][32] ... backward
  double* sbeta = k2b + 6 * 32;      // [2] theta*beta[0], theta*beta[J-1]
  double* sscr = sbeta + 2;          // [max(8*NW, 2*TPB)] set-up scratch: matrix totals of the warps, then 1/pivot of every thread's last cell
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
  auto cmul = [&](double x) { return (MODE == 1) ? x : cth * x; };
  (void)rth;

  for (int64_t p = blockIdx.x; p < pr.P; p += gridDim.x) {
    // ---------------------------------------------------------------- set-up: rows of A
    const int64_t jl0 = (int64_t)tid * CPT;
    {
      const double dt_theta = mul(dt[p], theta);
      double D0 = cell_D(pr, p, jl0);
      Face FL = problem_face(pr, p, jl0 - 1, cell_D(pr, p, jl0 - 1), D0);
#pragma unroll 1
      for (int c = 0; c < CPT; ++c) {
        const int64_t j = jl0 + c;
        const double Dp = cell_D(pr, p, j + 1);
        const Face FR = problem_face(pr, p, j, D0, Dp);
        Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;          // padding rows (j >= J) are identity
        if (j < J) {
          Row m = problem_row(pr, p, j, FL, FR);
          a = theta_lhs_row(m, dt_theta, alpha_of(j, J, pr.alpha_left, pr.alpha_right));
        }
        D0 = Dp; FL = FR;
        const int o = c * TPB + tid;
        sl[o] = a.lo; sk[o] = a.di; ss[o] = a.up;
      }
      if (tid == 0) {
        double beta_l = 0.0, beta_r = 0.0;
        if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
          beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
        if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
          beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
        sbeta[0] = theta * beta_l; sbeta[1] = theta * beta_r;
      }
    }
    // ---------------------------------------------------------------- factorisation: Moebius scan in dd
    dd s_in;
    {
      M2 P = m2_identity();
#pragma unroll 1
      for (int c = 0; c < CPT; ++c) {                        // P <- [[0, up], [-lo, di]] x P
        const int o = c * TPB + tid;
        const double lo = sl[o], di = sk[o], up = ss[o];
        M2 n;
        n.a = dd_mul_d(P.c, up); n.b = dd_mul_d(P.d, up);
        n.c = dd_sub(dd_mul_d(P.c, di), dd_mul_d(P.a, lo));
        n.d = dd_sub(dd_mul_d(P.d, di), dd_mul_d(P.b, lo));
        m2_normalise(n);
        P = n;
      }
      M2 inc = P;                                            // inclusive scan inside the warp
#pragma unroll 1
      for (int d = 1; d < 32; d <<= 1) {
        const M2 o = m2_shfl_up(inc, d);
        if (lane >= d) inc = m2_mul(inc, o);
      }
      M2 exc = m2_shfl_up(inc, 1);
      if (lane == 0) exc = m2_identity();
      if (lane == 31) m2_store(sscr + 8 * warp, inc);
      __syncthreads();
      M2 pre = m2_identity();                                // product of the warps before this one
      for (int i = 0; i < warp; ++i) pre = m2_mul(m2_load(sscr + 8 * i), pre);
      const M2 all = m2_mul(exc, pre);                       // everything before this thread
      // the recurrence starts from s_{-1} = 0, i.e. from the vector (0, 1): second column
      s_in = (tid == 0) ? dd_make(0.0) : dd_div(all.b, all.d);
      __syncthreads();                                       // sscr is reused below
    }
    double lreg[REGC ? CPT : 1], sreg[REGC ? CPT : 1];
    double lo_first;
    dd a_f = dd_make(1.0), a_b = dd_make(1.0);               // products of -l (forward map) and -s (backward map) over this thread's cells
    {
      dd s_prev = s_in, inv_prev = dd_make(0.0);
#pragma unroll 1
      for (int c = 0; c < CPT; ++c) {
        const int o = c * TPB + tid;
        const double lo = sl[o], di = sk[o], up = ss[o];
        if (c == 0) lo_first = lo;
        const dd inv = dd_recip(dd_add_d(dd_neg(dd_mul_d(s_prev, lo)), di));
        s_prev = dd_mul_d(inv, up);
        const double l = (c == 0) ? 0.0 : dd_round(dd_mul_d(inv_prev, lo));     // l of the first cell: after the barrier
        sl[o] = l; sk[o] = dd_round(dd_mul_d(inv, rth)); ss[o] = dd_round(s_prev);
        inv_prev = inv;
      }
      sscr[2 * tid] = inv_prev.hi; sscr[2 * tid + 1] = inv_prev.lo;
      __syncthreads();
      if (tid > 0) sl[tid] = dd_round(dd_mul_d(dd_make(sscr[2 * (tid - 1)], sscr[2 * (tid - 1) + 1]), lo_first));
      // s -> s" = s kappa_next / kappa (the substitution runs on u = yhat / kappa), from the rounded factors
      // in double-double, rounded once; kappa_next of a thread's last cell is the next thread's first
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        const int o = c * TPB + tid;
        const bool last = (c == CPT - 1);
        const double kn = !last ? sk[o + TPB] : (tid < TPB - 1 ? sk[tid + 1] : 0.0);
        ss[o] = dd_round(dd_div(two_prod(ss[o], kn), dd_make(sk[o])));
      }
      // products of the ROUNDED factors, as the sweeps apply them
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        const int o = c * TPB + tid;
        const double l = sl[o], s = ss[o];
        if (REGC) { lreg[c] = l; sreg[c] = s; }
        a_f = dd_neg(dd_mul_d(a_f, l));
        a_b = dd_neg(dd_mul_d(a_b, s));
      }
    }
    ScanConsts Kf, Kb;
    {
      const dd tf = scan_consts<false>(a_f, Kf);
      const dd tb = scan_consts<true>(a_b, Kb);
      if (lane == 31) awf[warp] = dd_round(tf);
      if (lane == 0) awb[warp] = dd_round(tb);
    }
    __syncthreads();
    if (warp == 0) {                                         // second level: scan over the warps' A totals
      ScanConsts K2;
      scan_consts<false>(dd_make(lane < NW ? awf[lane] : 1.0), K2);
#pragma unroll
      for (int i = 0; i < 5; ++i) k2f[i * 32 + lane] = K2.lvl[i];
      scan_consts<true>(dd_make(lane < NW ? awb[lane] : 1.0), K2);
#pragma unroll
      for (int i = 0; i < 5; ++i) k2b[i * 32 + lane] = K2.lvl[i];
    }
    __syncthreads();
    const double bfirst = sbeta[0], blast = sbeta[1];
    const int c_last = (J - 1 >= jl0 && J - 1 < jl0 + CPT) ? (int)(J - 1 - jl0) : -1;
    auto L = [&](int c) { return REGC ? lreg[c] : sl[c * TPB + tid]; };
    auto S = [&](int c) { return REGC ? sreg[c] : ss[c * TPB + tid]; };

    // ---------------------------------------------------------------- state in
    double w[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t j = jl0 + c;
      w[c] = (j < J) ? w_io[p * J + j] : 0.0;
    }

    // ---------------------------------------------------------------- time loop
    int64_t until_ckpt = ckpt_every, n_ckpt = 0;
    for (int64_t n = 0; n < nsteps; ++n) {
      double h[CPT];
      double prev = 0.0;
#pragma unroll
      for (int c = 0; c < CPT; ++c) {                        // forward recurrence with a zero carry-in
        double rc = w[c];
        if (c == c_last) rc += blast;
        if (c == 0 && tid == 0) rc += bfirst;
        h[c] = rc;
        prev = fma(-L(c), prev, rc);
      }
      prev = scan_values<NW, false>(prev, Kf, totf, k2f);
#pragma unroll
      for (int c = 0; c < CPT; ++c) { prev = fma(-L(c), prev, h[c]); h[c] = prev; }
      double nxt = 0.0;
#pragma unroll
      for (int c = CPT - 1; c >= 0; --c) nxt = fma(-S(c), nxt, h[c]);     // backward recurrence on u, zero carry-in
      nxt = scan_values<NW, true>(nxt, Kb, totb, k2b);
#pragma unroll
      for (int c = CPT - 1; c >= 0; --c) {
        nxt = fma(-S(c), nxt, h[c]);
        const double kap = sk[c * TPB + tid];
        w[c] = (MODE == 2) ? kap * nxt : fma(kap, nxt, -cmul(w[c]));
      }

      if (ckpt != nullptr && ckpt_every > 0 && --until_ckpt == 0) {
        until_ckpt = ckpt_every;
        double* dst = ckpt + (n_ckpt * pr.P + p) * J;
        ++n_ckpt;
#pragma unroll
        for (int c = 0; c < CPT; ++c) if (jl0 + c < J) dst[jl0 + c] = w[c];
      }
    }
#pragma unroll
    for (int c = 0; c < CPT; ++c) if (jl0 + c < J) w_io[p * J + jl0 + c] = w[c];
    __syncthreads();                                         // shared memory is reused by the next system
  }
}

template <int TPB, int CPT, int MODE>
static int launch_cta(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps, int64_t ckpt_every,
                      double* w, double* ckpt, cudaStream_t stream) {
  constexpr int NW = TPB / 32;
  constexpr size_t scr = (8 * NW > 2 * TPB) ? 8 * NW : 2 * TPB;
  constexpr size_t smem = (size_t)(3 * TPB * CPT + 4 * NW + 12 * 32 + 2 + scr) * sizeof(double);
  int dev = 0, sms = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  auto kern = fused_cta_kernel<TPB, CPT, MODE>;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, TPB, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) per_sm = 1;
  int64_t blocks = pr.P;
  if (blocks > (int64_t)sms * per_sm) blocks = (int64_t)sms * per_sm;    // persistent over the systems
  kern<<<(unsigned)blocks, TPB, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt);
  return (int)cudaGetLastError();
}

template <int TPB, int CPT>
static int dispatch_cta(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps, int64_t ckpt_every,
                        double* w, double* ckpt, cudaStream_t stream) {
  if (theta == 0.5) return launch_cta<TPB, CPT, 1>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (theta == 1.0) return launch_cta<TPB, CPT, 2>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  return launch_cta<TPB, CPT, 0>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
}

// Called by fvm_theta_run_fused (fused.cu) for 1024 < J <= 8192.
int run_fused_cta(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps, int64_t ckpt_every,
                  double* w, double* ckpt, cudaStream_t stream) {
  // A step is a chain of latencies (two scans, two barriers): with more systems than SMs, CTAs of half the
  // threads and twice the cells per thread share an SM (four of 128 threads, two of 256) and fill each other's
  // waits; with fewer systems, the wider CTA finishes a system sooner.
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return (int)cudaGetLastError();
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (pr.J <= 2048) {
    if (pr.P > sms) return dispatch_cta<128, 16>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    return dispatch_cta<256, 8>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  }
  if (pr.J <= 4096) {
    if (pr.P > sms) return dispatch_cta<256, 16>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    return dispatch_cta<512, 8>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  }
  return dispatch_cta<512, 16>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
}

}  // namespace fvm
