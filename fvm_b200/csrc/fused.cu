// Variant 2: one warp per system, everything on chip between checkpoints.
//
// Partition: lane l owns the C contiguous cells [l*C, (l+1)*C); the first N = C-1 are
// "interior" cells, the last one is the lane's SEPARATOR.  With the separators known the
// 32 interior blocks decouple (SPIKE).  A theta step has NO right-hand side: B = I/theta - c A
// gives A^-1 B = A^-1/theta - c I, so with the un-normalised LU factors (l, kappa, s) of a lane's
// block T, all factored ONCE per launch (in double-double, dd.cuh, each rounded once):
//   1. sweeps on the block          4 operations / cell for Crank-Nicolson
//   2. separator system (32 rows)   precomputed parallel cyclic reduction, 5 levels of 2 shuffles + 2 FMA
//   3. spike correction             2 FMA / cell   (V, U = T^-1 of the couplings to the two separators)
// State w[C] lives in registers for the whole launch; where the coefficients live decides the speed:
//   * fused_kernel           (J <= 512, and the slow path of a right flux at a lane-interior cell):
//                            s in registers, l, kappa, V, U in shared memory.
//   * fused_tmem_pipe_kernel (512 < J <= 1024, the default of cfg 3): s in registers, (l, kappa) and
//                            the delayed-correction vectors (P1, P2) in TENSOR MEMORY, which is idle
//                            in this workload -- there is no MMA -- and whose tcgen05.ld path is
//                            separate from, and wider than, the 128 B/clk shared-memory pipe.
// HBM traffic: J doubles in, J doubles out, J per checkpoint.  Assembly (fvm_device.cuh) and the
// factorisation run inside the same launch, so a launch needs only the mesh, the per-problem scalars
// and the state.
#include "fvm_device.cuh"
#include "tmem.cuh"
#include "dd.cuh"
#include <type_traits>

namespace fvm {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kFusedMaxC = 32;
constexpr int kFusedThreads = 256;   // upper bound used for register allocation

__device__ __forceinline__ double shfl_up(double x, int s) { return __shfl_up_sync(kFull, x, s); }
__device__ __forceinline__ double shfl_dn(double x, int s) { return __shfl_down_sync(kFull, x, s); }

// Per-lane constants of one system of the shared-memory kernel, fixed for the whole launch.
struct Reduced {
  double al[5], ga[5];           // PCR elimination factors of the separator system, per level
  double invD;                   // 1 / final PCR diagonal
  double slo, sup;               // separator's row of A (off-diagonals)
  double bfirst, bint;           // theta*beta[0] / theta*beta[J-1] where they enter the elimination (interior cells)
  double bsep, bsep_r;           // beta at the separator: left-flux part (cell 0), right-flux part (cell J-1)
  int rb_c;                      // RB only: position of cell J-1 inside its lane
};

// PCR factorisation of the 32-unknown separator system (RL, RD, RU per lane) in double-double: the
// elimination factors of each of the 5 levels and the final diagonal are kept (each rounded once), so
// that a step only applies them to its right-hand side (pcr_solve).  K: any struct with al, ga, invD.
template <class KT>
__device__ __forceinline__ void pcr_factor_dd(int lane, dd RL, dd RD, dd RU, KT& K) {
#pragma unroll                                      // K.al[l], K.ga[l] with a run-time l would put K in local memory
  for (int l = 0; l < 5; ++l) {
    const int s = 1 << l;
    const bool hm = lane >= s, hp = lane + s < 32;
    const dd Dm = dd_shfl_up(RD, s), Dp = dd_shfl_dn(RD, s);
    const dd Lm = dd_shfl_up(RL, s), Um = dd_shfl_up(RU, s);
    const dd Lp = dd_shfl_dn(RL, s), Up = dd_shfl_dn(RU, s);
    dd a_ = dd_make(0.0), g_ = dd_make(0.0);
    if (hm) a_ = dd_neg(dd_div(RL, Dm));
    if (hp) g_ = dd_neg(dd_div(RU, Dp));
    K.al[l] = dd_round(a_); K.ga[l] = dd_round(g_);
    RD = dd_add(RD, dd_add(dd_mul(a_, Um), dd_mul(g_, Lp)));
    RL = dd_mul(a_, Lm);
    RU = dd_mul(g_, Up);
  }
  K.invD = dd_round(dd_recip(RD));
}

// Set-up of one system of the shared-memory kernel, in double-double.
// Phase 1 (top-down over the lane's C cells): rows of A, un-normalised LU of the interior block
// (l -> sl, kappa -> sk, s -> ss with its low word in tlo), elimination of the left spike with the
// UNROUNDED factors (g -> sv, low word in glo), the separator's row.  Phase 2 (bottom-up): both
// spikes V -> sv, U -> su of the exact block, the separator system and its PCR factors.
template <int C, bool RB>
__device__ __forceinline__ void setup_system(const fvm_problem& pr, int64_t p, int lane, double dt_p, double theta,
                                             double* sl, double* sk, double* ss, double* sv, double* su,
                                             double* tlo, double* glo, Reduced& R) {
  constexpr int N = C - 1;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta;
  const double dt_theta = mul(dt_p, theta);
  double beta_l = 0.0, beta_r = 0.0;
  if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
    beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
  if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
    beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
  R.slo = 0.0; R.sup = 0.0;
  R.bfirst = 0.0; R.bint = 0.0; R.bsep = 0.0; R.bsep_r = 0.0;
  R.rb_c = RB ? (int)((J - 1) % C) : -1;
  double sdi = 1.0;
  const int64_t j0 = (int64_t)lane * C;
  // sliding window over faces: one expm1 and one face (5 divisions) per row
  double D0 = cell_D(pr, p, j0);
  Face FL = problem_face(pr, p, j0 - 1, cell_D(pr, p, j0 - 1), D0);
  dd s_prev = dd_make(0.0), inv_prev = dd_make(0.0), hv = dd_make(0.0);
  double up_last = 0.0;                              // A_{N-1,N}: coupling of the last interior cell to the separator
#pragma unroll 1
  for (int c = 0; c < C; ++c) {
    const int64_t j = j0 + c;
    const double Dp = cell_D(pr, p, j + 1);
    const Face FR = problem_face(pr, p, j, D0, Dp);
    Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;      // padding rows (j >= J) are identity
    if (j < J) {
      Row m = problem_row(pr, p, j, FL, FR);
      a = theta_lhs_row(m, dt_theta, alpha_of(j, J, pr.alpha_left, pr.alpha_right));
    }
    D0 = Dp; FL = FR;
    if (c < N) {
      const dd inv = dd_recip(dd_add_d(dd_neg(dd_mul_d(s_prev, a.lo)), a.di));
      const dd l = dd_mul_d(inv_prev, a.lo);
      const dd kap = dd_mul_d(inv, rth);
      s_prev = dd_mul_d(inv, a.up);
      inv_prev = inv;
      hv = (c == 0) ? dd_make(a.lo) : dd_neg(dd_mul(hv, l));
      const dd g = dd_mul(hv, kap);
      sl[c * 32] = dd_round(l); sk[c * 32] = dd_round(kap);
      ss[c * 32] = s_prev.hi; tlo[c * 32] = s_prev.lo;
      sv[c * 32] = g.hi; glo[c * 32] = g.lo;
      up_last = a.up;
      if (j == 0) R.bfirst = theta * beta_l;
      if (RB && j == J - 1) R.bint = theta * beta_r;
    } else {
      R.slo = a.lo; sdi = a.di; R.sup = a.up;
      if (j == 0) R.bsep += beta_l;
      if (j == J - 1) R.bsep_r += beta_r;
    }
  }
  dd RL = dd_make(R.slo), RD = dd_make(sdi), RU = dd_make(R.sup);
  if (N > 0) {
    // substitution of both spikes, in yhat units (Vhat = V / theta): U starts from the last interior cell
    dd vn = dd_make(sv[(N - 1) * 32], glo[(N - 1) * 32]);
    dd un = dd_mul_d(dd_mul_d(inv_prev, rth), up_last);
    sv[(N - 1) * 32] = dd_round(dd_mul_d(vn, theta)); su[(N - 1) * 32] = dd_round(dd_mul_d(un, theta));
    const dd vlast = dd_mul_d(vn, theta), ulast = dd_mul_d(un, theta);
#pragma unroll 1
    for (int c = N - 2; c >= 0; --c) {
      const dd sc = dd_make(ss[c * 32], tlo[c * 32]);
      vn = dd_sub(dd_make(sv[c * 32], glo[c * 32]), dd_mul(vn, sc));
      un = dd_neg(dd_mul(un, sc));
      sv[c * 32] = dd_round(dd_mul_d(vn, theta)); su[c * 32] = dd_round(dd_mul_d(un, theta));
    }
    const dd v0 = dd_mul_d(vn, theta), u0 = dd_mul_d(un, theta);
    const dd v0n = dd_shfl_dn(v0, 1), u0n = dd_shfl_dn(u0, 1);        // next lane's first interior cell
    RL = dd_neg(dd_mul_d(vlast, R.slo));
    RD = dd_sub(dd_add_d(dd_neg(dd_mul_d(ulast, R.slo)), sdi), dd_mul_d(v0n, R.sup));   // sup == 0 on lane 31
    RU = dd_neg(dd_mul_d(u0n, R.sup));
  }
  pcr_factor_dd(lane, RL, RD, RU, R);
}

__device__ __forceinline__ double pcr_solve(double rr, const Reduced& R) {
#pragma unroll
  for (int l = 0; l < 5; ++l) {
    const int s = 1 << l;
    const double rm = shfl_up(rr, s), rp = shfl_dn(rr, s);
    rr = fma(R.ga[l], rp, fma(R.al[l], rm, rr));
  }
  return rr * R.invD;
}

template <int C>
__device__ __forceinline__ void load_state(const double* __restrict__ src, int64_t J, int lane, double (&w)[C]) {
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int64_t j = (int64_t)lane * C + c;
    w[c] = (j < J) ? src[j] : 0.0;
  }
}
template <int C>
__device__ __forceinline__ void store_state(double* __restrict__ dst, int64_t J, int lane, const double (&w)[C]) {
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const int64_t j = (int64_t)lane * C + c;
    if (j < J) dst[j] = w[c];
  }
}

// What changes from step to step (nothing of this is in the reference, whose beta and k
// are constants of a model: fvm_classes.py:221,228 only mention the case).  All pointers may be NULL.
//  * boundary fluxes: step s uses  flux * scale_left[s*sc_ss + p*sc_ps]  (beta is linear in the flux,
//    fvm_classes.py:244-263, so prob.left_flux / right_flux hold the UNIT or base flux);
//  * time step: the run is cut into `n_runs` runs of steps [run_starts[r], run_starts[r+1]) that share one
//    time step  dt_hist[run_starts[r]*dt_ss + p*dt_ps];  the kernel re-factors at the start of every run
//    (set-up in the same launch, state parked in w_io meanwhile).
__device__ const double kUnitScale = 1.0;

struct FusedExtras {
  const double* scale_left; const double* scale_right; int64_t sc_ps, sc_ss;
  const double* dt_hist; int64_t dt_ps, dt_ss;
  const int64_t* run_starts; int64_t n_runs;
};

// ------------------------------------------------------------------------------------------
// Shared-memory variant (J <= 512, and the RB slow path of J <= 1024).  Same formulation as the
// tensor-memory kernel below -- no right-hand side, separator rows in yhat, set-up in double-double,
// spikes of the exact block -- without its twisted factorisation and delayed correction: per step
//   h_c = g_c + theta*beta_c - l_c h_{c-1};  u_c = h_c - s"_c u_{c+1};  z_c = kappa_c u_c - c g_c
//   R yhat_s = xs/theta + beta_s - slo ghat_last - sup ghat_first(next lane),  ghat = z + c g
//   x'_c = z_c - V_c yhat_s(l-1) - U_c yhat_s(l);   x'_s = yhat_s - c xs
// l, kappa, V, U in shared memory ([coef][c][lane], lane-consecutive doubles: conflict-free), s in
// registers (QREG).  RB: beta[J-1] is nonzero AND cell J-1 is an interior cell of its lane.
template <int C, int MODE, bool RB, bool QREG, bool EXT>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
             int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt, FusedExtras ex) {
  constexpr int N = C - 1;
  constexpr int NS = (N > 0 ? N : 1);
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  constexpr int NARR = 7;                            // l, kappa, V, U, and during set-up s, low words of s and g
  double* sl = smem + (size_t)warp * (NARR * NS * 32) + lane;   // [c*32] indexing below
  double* sk = sl + NS * 32;
  double* sv = sk + NS * 32;
  double* su = sv + NS * 32;
  double* ss = su + NS * 32;
  double* tlo = ss + NS * 32;
  double* glo = tlo + NS * 32;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
  auto cmul = [&](double x) { return (MODE == 1) ? x : cth * x; };      // c * x (MODE 2 never asks)

  const int64_t n_ck_total = (ckpt != nullptr && ckpt_every > 0) ? nsteps / ckpt_every : 0;
  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
   const int64_t n_runs = (EXT && ex.run_starts != nullptr) ? ex.n_runs : 1;
   int64_t s_glob = 0;
   for (int64_t run = 0; run < n_runs; ++run) {
    int64_t run_len = nsteps;
    double dt_run = dt[p];
    if (EXT && ex.run_starts != nullptr) {
      const int64_t s0 = ex.run_starts[run];
      run_len = ex.run_starts[run + 1] - s0;
      dt_run = ex.dt_hist[s0 * ex.dt_ss + p * ex.dt_ps];
    }
    Reduced R;
    setup_system<C, RB>(pr, p, lane, dt_run, theta, sl, sk, ss, sv, su, tlo, glo, R);
    // s" = s kappa_next / kappa: the substitution runs on u = yhat / kappa (see the tensor-memory kernel, 2.)
    double qr[NS];
#pragma unroll
    for (int c = 0; c < NS; ++c)
      qr[c] = (c + 1 < N) ? dd_round(dd_div(two_prod(ss[c * 32], sk[(c + 1) * 32]), dd_make(sk[c * 32]))) : 0.0;
    __syncwarp();

    double w[C];
    load_state<C>(w_io + p * J, J, lane, w);

    for (int64_t n = 0; n < run_len; ++n, ++s_glob) {
      double bfirst = R.bfirst, bint = R.bint, bsep = R.bsep + R.bsep_r;
      if (EXT) {
        const double sc_l = ex.scale_left ? ex.scale_left[s_glob * ex.sc_ss + p * ex.sc_ps] : 1.0;
        const double sc_r = ex.scale_right ? ex.scale_right[s_glob * ex.sc_ss + p * ex.sc_ps] : 1.0;
        bfirst *= sc_l; bint *= sc_r; bsep = R.bsep * sc_l + R.bsep_r * sc_r;
      }
      const double xs_old = w[C - 1];
      double rr = fma(xs_old, rth, bsep);
      if (N > 0) {
        const double g0_old = w[0], gl_old = w[N - 1];
        double hh[NS];
        double h = 0.0;
#pragma unroll
        for (int c = 0; c < N; ++c) {
          double r = w[c];
          if (c == 0) r += bfirst;
          if (RB) { if (c == R.rb_c) r += bint; }
          h = (c == 0) ? r : fma(-sl[c * 32], h, r);
          hh[c] = h;
        }
        double u = 0.0;
#pragma unroll
        for (int c = N - 1; c >= 0; --c) {
          u = (c == N - 1) ? hh[c] : fma(-qr[c], u, hh[c]);
          w[c] = (MODE == 2) ? sk[c * 32] * u : fma(sk[c * 32], u, -cmul(w[c]));    // z = kappa u - c g
        }
        // ghat = z + c g: the block solution of yhat with the separators at zero
        const double gh_last = (MODE == 2) ? w[N - 1] : w[N - 1] + cmul(gl_old);
        const double gh_first = (MODE == 2) ? w[0] : w[0] + cmul(g0_old);
        const double gh_first_next = shfl_dn(gh_first, 1);        // lane 31: sup == 0
        rr = fma(-R.sup, gh_first_next, fma(-R.slo, gh_last, rr));
      }
      const double ys = pcr_solve(rr, R);
      if (N > 0) {
        const double ysl = shfl_up(ys, 1);          // lane 0: V == 0, value unused
#pragma unroll
        for (int c = 0; c < N; ++c) w[c] = fma(-su[c * 32], ys, fma(-sv[c * 32], ysl, w[c]));
      }
      w[C - 1] = (MODE == 2) ? ys : ys - cmul(xs_old);

      const int64_t done = s_glob + 1;
      if (n_ck_total > 0 && done % ckpt_every == 0 && done / ckpt_every <= n_ck_total)
        store_state<C>(ckpt + ((done / ckpt_every - 1) * pr.P + p) * J, J, lane, w);
    }
    store_state<C>(w_io + p * J, J, lane, w);
    __syncwarp();
   }
  }
}

// ------------------------------------------------------------------------------------------
// Tensor-memory kernel, C = 32 (512 < J <= 1024), 8 warps per CTA, one CTA per SM: the default
// for cfg 3.  Coefficients live in TENSOR MEMORY: TMEM (256 KB/SM) is idle in this workload --
// there is no MMA -- and its tcgen05.ld path is separate from, and wider than, the 128 B/clk
// shared-memory pipe (profiles/microbench/r01_tmem_bw_b200.txt).
//
// 1. TWISTED factorisation of each lane's interior block T: cells 0..14 are eliminated top-down,
//    cells 30..16 bottom-up, the sweeps meet AT cell 15 (its pivot takes a term from either side)
//    and substitution runs outwards from there: two independent chains of half the length.
// 2. NO RIGHT-HAND SIDE.  B = I/theta - c A (c = (1-theta)/theta) gives A^-1 B = (1/theta) A^-1 - c I.
//    With the un-normalised factors l_c = A_c,prev / pivot_prev, kappa_c = 1 / (theta pivot_c),
//    s_c = A_c,next / pivot_c, yhat = (1/theta) T^-1 g, and the substitution run on u = yhat / kappa
//    (so that its inhomogeneous term is h itself):
//        elimination   h_c = g_c - l_c h_prev                          (1 FMA, the chain; h kept beside g)
//        substitution  u_c = h_c - s"_c u_next ,  s"_c = s_c kappa_next / kappa_c     (1 FMA, the chain)
//                      z_c = kappa_c u_c - c g_c                       (1 FMA off the chain, written over g_c)
//    3 operations per cell for Crank-Nicolson and implicit Euler, 4 in general (c g_c needs a product).
//    The price is a second register array (h) where round 1 kept s; the five coefficient arrays
//    (l, kappa, s", P1, P2) are split between tensor memory (the four the substitution reads, 128
//    doubles per lane: all of it) and shared memory (l).  Writing t_c = h_c kappa_c - c g_c - c s_c g_next
//    over g_c, as before, needs no second array but one more operation per cell: measured on cfg 3,
//    the operation count is what the step time follows (profiles/microbench/r02_theta_modes_b200.txt).
// 3. DELAYED CORRECTION.  The separator PCR is a pure latency chain.  The step is linear with
//    constant coefficients, so the kernel keeps, between steps, the UNCORRECTED block solutions g
//    (true state x = g - V xsl - U xs) and advances
//        g' = (1/theta) T^-1 g - c g  +  P1 xsl  +  P2 xs ,   P1 = -(1/theta) T^-1 V,  P2 = -(1/theta) T^-1 U ,
//    so the PCR of step n runs interleaved with the elimination chunks of step n+1 and the
//    correction rides on the substitution links.  V and U stay in shared memory for the points
//    where the true state is needed (checkpoints, end of the launch).
// 4. SEPARATOR ROWS in yhat as well:  A yhat = x/theta + beta,  x' = yhat - c x.  Row of separator s:
//        R yhat_s = xs/theta + beta_s - slo ghat_last - sup ghat_first(next lane),   ghat = g' + c g_old
//    (ghat = (1/theta) T^-1 x_int is the uncorrected block solution of yhat), then x'_s = yhat_s - c xs.
//    Nothing of size dt*d/h^2 is multiplied by a state value and cancelled, and the true end values
//    of the blocks are not needed.
// 5. SET-UP IN DOUBLE-DOUBLE (dd.cuh).  The loop applies the same rounded constants 10^4 times, so
//    their errors are systematic.  (a) The pivot chain runs in dd and l, kappa, s, delta are
//    rounded ONCE.  (b) The spikes V, U, the separator system built from their ends and its PCR
//    factors come from the UNROUNDED factors, chains and intermediates in dd, and are rounded once
//    at the end: the spikes carry the separator values (the large part of the solution of a stiff
//    block) into the interior, so they have to be those of the exact block.  Spikes "consistent"
//    with the rounded factors were tried and are 100x worse (3.5e-11 against 2e-13 after 10,000
//    steps on cfg 3's stiffest problems): the rounding of l, kappa, s then acts on terms of size
//    (dt d/h^2) x instead of on the small uncorrected block solutions.  (c) P1, P2 = -Lhat(V, U)
//    use the loop's own rounded factors with dd chains.
//
// TMEM layout per warp (256 columns of its lane quadrant) DURING SET-UP (pipe_setup):
//   columns [0,128)  : slot i < 15 = (l, kappa) of left cell i, then of right cell 30-i  (8 columns);
//                      slot 15 = (l_down, kappa) and (l_up, -) of the middle cell
//   columns [128,256): slot 0 = (P1, P2) of the middle cell; slot j = (P1, P2) of left cell 15-j,
//                      then of right cell 15+j                                         (substitution order)
// and IN THE LOOP (pipe_relayout), in the order the substitution consumes it, one .x16 chunk per slot:
//   slot j (16 columns) = (kappa, P1, P2, s") of left cell 15-j, then of right cell 15+j; slot 0 = the
//   middle cell.  l moves to shared memory (the slots s occupied; l_up of the middle cell in slot 0,
//   whose own l is never read).  Chunks are double-buffered.
// MODE 0: any theta in (0,1];  1: Crank-Nicolson;  2: implicit Euler.
struct TmemChunk {
  static constexpr int REGS = 16;
  static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[16]) { tmem_ld_x16(t, r); }
  static __device__ __forceinline__ void wait(uint32_t (&r)[16]) { tmem_wait_ld_x16(r); }
};

// Issue a chunk load and thread a value through an empty volatile asm behind it: whatever is
// computed from `dep` afterwards is ordered after the ISSUE of the load.  This pins a prefetch
// ahead of the arithmetic it is meant to overlap; the compiler otherwise sinks the load down to
// its wait to save registers, and every chunk then pays the tensor-memory latency.
__device__ __forceinline__ void chunk_ld_after(uint32_t t, uint32_t (&r)[16], double& dep) {
  TmemChunk::ld(t, r);
  asm volatile("" : "+d"(dep));
}

// k-th double of a chunk (two 32-bit columns each)
__device__ __forceinline__ double chunk_dbl(const uint32_t* r, int k) { return __hiloint2double((int)r[2 * k + 1], (int)r[2 * k]); }

constexpr int kPipeN = 31, kPipeMid = 15;           // interior cells per lane; the cell the two sweeps meet at
// column offsets of cell c inside the (l, kappa) region and inside the (P1, P2) region
__device__ __forceinline__ uint32_t pipe_col_lk(int c) {
  return (uint32_t)(c <= kPipeMid ? 8 * c : 8 * (kPipeN - 1 - c) + 4);
}
__device__ __forceinline__ uint32_t pipe_col_pp(int c) {
  return (uint32_t)(c <= kPipeMid ? 8 * (kPipeMid - c) : 8 * (c - kPipeMid) + 4);
}

// Per-lane constants of a step (kept in the shared-memory slots that s occupies during set-up).
struct PipeConsts {
  double al[5], ga[5], invD;     // PCR factors of the separator system
  double slo, sup;               // separator's row of A (off-diagonals)
  double bsep, bfirst;           // beta at the separator; theta * beta at global cell 0
  double om_l, om_r;             // (1 - theta) * coupling of the block's end cells to the separators
};

// Set-up of one system (see 5. above).  On return: (l, kappa) and (P1, P2) are in tensor memory,
// s in sS, V and U in sv / su (natural cell order), the step constants in K.
__device__ __forceinline__ void pipe_setup(const fvm_problem& pr, int64_t p, int lane, double dt_p, double theta,
                                           uint32_t t_lk, uint32_t t_pp, double* sS, double* sv, double* su,
                                           PipeConsts& K) {
  constexpr int C = 32, N = kPipeN, MID = kPipeMid;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta;
  double sdi = 1.0;
  dd vm, um;                                      // yhat-units spikes at the middle cell
  double l_dn, l_up, kap_m;                       // the middle cell's factors
  {
    const double dt_theta = mul(dt_p, theta);
    double beta_l = 0.0, beta_r = 0.0;
    if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
      beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
    if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
      beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
    K.bfirst = 0.0; K.bsep = 0.0; K.om_l = 0.0; K.om_r = 0.0;
    const int64_t j0 = (int64_t)lane * C;
    Row asep;
    auto row_of = [&](int c, const Face& FL, const Face& FR) {
      const int64_t j = j0 + c;
      Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;    // padding rows (j >= J) are identity
      if (j < J) {
        Row m = problem_row(pr, p, j, FL, FR);
        a = theta_lhs_row(m, dt_theta, alpha_of(j, J, pr.alpha_left, pr.alpha_right));
      }
      return a;
    };
    // sliding windows over faces, one per sweep direction: one expm1 and one face (5 divisions)
    // per row instead of three and two
    double lD0 = cell_D(pr, p, j0);
    Face lFL = problem_face(pr, p, j0 - 1, cell_D(pr, p, j0 - 1), lD0);
    double rD0 = cell_D(pr, p, j0 + C - 1);                                 // the separator's cell
    Face rFR = problem_face(pr, p, j0 + C - 1, rD0, cell_D(pr, p, j0 + C));
    {
      const double Dm = cell_D(pr, p, j0 + C - 2);
      const Face F = problem_face(pr, p, j0 + C - 2, Dm, rD0);
      asep = row_of(C - 1, F, rFR);                                         // the separator's row
      rFR = F; rD0 = Dm;
    }
    dd sl_prev = dd_make(0.0), sr_prev = dd_make(0.0), invl_prev = dd_make(0.0), invr_prev = dd_make(0.0);
    dd hv = dd_make(0.0), hu = dd_make(0.0);
#pragma unroll 1
    for (int i = 0; i < MID; ++i) {
      {                                            // left cell i: top-down elimination
        const double lDp = cell_D(pr, p, j0 + i + 1);
        const Face lFR = problem_face(pr, p, j0 + i, lD0, lDp);
        const Row a = row_of(i, lFL, lFR);
        lD0 = lDp; lFL = lFR;
        const dd inv = dd_recip(dd_add_d(dd_neg(dd_mul_d(sl_prev, a.lo)), a.di));
        const dd l = dd_mul_d(invl_prev, a.lo);
        const dd kap = dd_mul_d(inv, rth);
        sl_prev = dd_mul_d(inv, a.up);
        invl_prev = inv;
        // left spike with the UNROUNDED factors: h_0 = A_0,-1, h_i = -l_i h_{i-1}, g_i = h_i kappa_i
        hv = (i == 0) ? dd_make(a.lo) : dd_neg(dd_mul(hv, l));
        const dd g = dd_mul(hv, kap);
        sS[i * 32] = dd_round(sl_prev); sv[i * 32] = g.hi; su[i * 32] = 0.0;
        tmem_st_2d(t_lk + pipe_col_lk(i), dd_round(l), dd_round(kap));
        tmem_st_2d(t_pp + pipe_col_pp(i), g.lo, sl_prev.lo);   // staging: low words of the dd values g, s
        if (i == 0) K.om_l = (1.0 - theta) * a.lo;
        if (j0 + i == 0) K.bfirst = theta * beta_l;
      }
      {                                            // right cell 30-i: bottom-up elimination
        const int c = N - 1 - i;
        const double rDm = cell_D(pr, p, j0 + c - 1);
        const Face rFL = problem_face(pr, p, j0 + c - 1, rDm, rD0);
        const Row a = row_of(c, rFL, rFR);
        rD0 = rDm; rFR = rFL;
        const dd inv = dd_recip(dd_add_d(dd_neg(dd_mul_d(sr_prev, a.up)), a.di));
        const dd l = dd_mul_d(invr_prev, a.up);
        const dd kap = dd_mul_d(inv, rth);
        sr_prev = dd_mul_d(inv, a.lo);
        invr_prev = inv;
        hu = (i == 0) ? dd_make(a.up) : dd_neg(dd_mul(hu, l));
        const dd g = dd_mul(hu, kap);
        sS[c * 32] = dd_round(sr_prev); su[c * 32] = g.hi; sv[c * 32] = 0.0;
        tmem_st_2d(t_lk + pipe_col_lk(c), dd_round(l), dd_round(kap));
        tmem_st_2d(t_pp + pipe_col_pp(c), g.lo, sr_prev.lo);
        if (i == 0) K.om_r = (1.0 - theta) * a.up;
      }
    }
    {
      // the middle cell takes both sweeps: pivot = A_mm - A_m,m-1 s_{m-1} - A_m,m+1 s_{m+1}.  (A 2x2 solve
      // between the last cells of the two chains would divide by 1 - s s', which is ~2/sqrt(dt d/h^2) for
      // stiff cells and amplifies the rounding of the factors by its inverse at every step.)
      const Row a = row_of(MID, lFL, rFR);
      const dd piv = dd_sub(dd_add_d(dd_neg(dd_mul_d(sl_prev, a.lo)), a.di), dd_mul_d(sr_prev, a.up));
      const dd inv = dd_recip(piv);
      const dd ld = dd_mul_d(invl_prev, a.lo), lu = dd_mul_d(invr_prev, a.up), km = dd_mul_d(inv, rth);
      l_dn = dd_round(ld); l_up = dd_round(lu); kap_m = dd_round(km);
      vm = dd_mul(dd_neg(dd_mul(hv, ld)), km);
      um = dd_mul(dd_neg(dd_mul(hu, lu)), km);
      sS[MID * 32] = 0.0;
      tmem_st_2d(t_lk + pipe_col_lk(MID), l_dn, kap_m);
      tmem_st_2d(t_lk + pipe_col_lk(MID) + 4, l_up, 0.0);
    }
    K.slo = asep.lo; sdi = asep.di; K.sup = asep.up;
    if (j0 + C - 1 == J - 1) K.bsep += beta_r;
  }
  tmem_wait_st();
  // ---- spikes of the exact block:  Vhat = (1/theta) T^-1 (A_0,-1 e_0), Uhat likewise;  V = theta Vhat
  dd V0, U0, VL, UL;
  {
    sv[MID * 32] = dd_round(dd_mul_d(vm, theta)); su[MID * 32] = dd_round(dd_mul_d(um, theta));
    dd lv = vm, lu = um;
    double glo, slo_;
#pragma unroll 1
    for (int c = MID - 1; c >= 0; --c) {
      tmem_ld_2d(t_pp + pipe_col_pp(c), glo, slo_);
      const dd sc = dd_make(sS[c * 32], slo_);
      lv = dd_sub(dd_make(sv[c * 32], glo), dd_mul(lv, sc));
      lu = dd_neg(dd_mul(lu, sc));
      sv[c * 32] = dd_round(dd_mul_d(lv, theta)); su[c * 32] = dd_round(dd_mul_d(lu, theta));
    }
    dd rv = vm, ru = um;
#pragma unroll 1
    for (int c = MID + 1; c < N; ++c) {
      tmem_ld_2d(t_pp + pipe_col_pp(c), glo, slo_);
      const dd sc = dd_make(sS[c * 32], slo_);
      rv = dd_neg(dd_mul(rv, sc));
      ru = dd_sub(dd_make(su[c * 32], glo), dd_mul(ru, sc));
      sv[c * 32] = dd_round(dd_mul_d(rv, theta)); su[c * 32] = dd_round(dd_mul_d(ru, theta));
    }
    V0 = dd_mul_d(lv, theta); U0 = dd_mul_d(lu, theta); VL = dd_mul_d(rv, theta); UL = dd_mul_d(ru, theta);
  }
  // ---- separator system  R = A_ss - A_si (spike ends)  and its PCR factors
  {
    const dd v0n = dd_shfl_dn(V0, 1), u0n = dd_shfl_dn(U0, 1);        // next lane's first interior cell
    const dd RL = dd_neg(dd_mul_d(VL, K.slo));
    const dd RD = dd_sub(dd_add_d(dd_neg(dd_mul_d(UL, K.slo)), sdi), dd_mul_d(v0n, K.sup));   // sup == 0 on lane 31
    const dd RU = dd_neg(dd_mul_d(u0n, K.sup));
    pcr_factor_dd(lane, RL, RD, RU, K);
  }
  // ---- P1 = -Lhat(V), P2 = -Lhat(U): the loop's own sweeps with right-hand sides -V, -U, chains in dd,
  //      in place in tensor memory
  {
    double l, kap;
    dd h1 = dd_make(-sv[0]), h2 = dd_make(-su[0]);
    tmem_st_2d(t_pp + pipe_col_pp(0), h1.hi, h2.hi);
#pragma unroll 1
    for (int c = 1; c < MID; ++c) {                // eliminate downwards
      tmem_ld_2d(t_lk + pipe_col_lk(c), l, kap);
      h1 = dd_add_d(dd_neg(dd_mul_d(h1, l)), -sv[c * 32]); h2 = dd_add_d(dd_neg(dd_mul_d(h2, l)), -su[c * 32]);
      tmem_st_2d(t_pp + pipe_col_pp(c), h1.hi, h2.hi);
    }
    const dd hl1 = h1, hl2 = h2;                   // eliminated values at cell MID-1
    h1 = dd_make(-sv[(N - 1) * 32]); h2 = dd_make(-su[(N - 1) * 32]);
    tmem_st_2d(t_pp + pipe_col_pp(N - 1), h1.hi, h2.hi);
#pragma unroll 1
    for (int c = N - 2; c > MID; --c) {            // eliminate upwards
      tmem_ld_2d(t_lk + pipe_col_lk(c), l, kap);
      h1 = dd_add_d(dd_neg(dd_mul_d(h1, l)), -sv[c * 32]); h2 = dd_add_d(dd_neg(dd_mul_d(h2, l)), -su[c * 32]);
      tmem_st_2d(t_pp + pipe_col_pp(c), h1.hi, h2.hi);
    }
    // the middle cell
    const dd m1 = dd_mul_d(dd_sub(dd_add_d(dd_neg(dd_mul_d(hl1, l_dn)), -sv[MID * 32]), dd_mul_d(h1, l_up)), kap_m);
    const dd m2 = dd_mul_d(dd_sub(dd_add_d(dd_neg(dd_mul_d(hl2, l_dn)), -su[MID * 32]), dd_mul_d(h2, l_up)), kap_m);
    tmem_wait_st();
    dd p1 = m1, p2 = m2;
    double x1, x2;
#pragma unroll 1
    for (int c = MID - 1; c >= 0; --c) {           // substitute outwards
      tmem_ld_2d(t_pp + pipe_col_pp(c), x1, x2);
      tmem_ld_2d(t_lk + pipe_col_lk(c), l, kap);
      const double sc = sS[c * 32];
      p1 = dd_sub(two_prod(x1, kap), dd_mul_d(p1, sc)); p2 = dd_sub(two_prod(x2, kap), dd_mul_d(p2, sc));
      tmem_st_2d(t_pp + pipe_col_pp(c), dd_round(p1), dd_round(p2));
    }
    tmem_st_2d(t_pp + pipe_col_pp(MID), dd_round(m1), dd_round(m2));
    p1 = m1; p2 = m2;
#pragma unroll 1
    for (int c = MID + 1; c < N; ++c) {
      tmem_ld_2d(t_pp + pipe_col_pp(c), x1, x2);
      tmem_ld_2d(t_lk + pipe_col_lk(c), l, kap);
      const double sc = sS[c * 32];
      p1 = dd_sub(two_prod(x1, kap), dd_mul_d(p1, sc)); p2 = dd_sub(two_prod(x2, kap), dd_mul_d(p2, sc));
      tmem_st_2d(t_pp + pipe_col_pp(c), dd_round(p1), dd_round(p2));
    }
    tmem_wait_st();
  }
}

// From the set-up layout to the loop's (see above).  In: (l, kappa) and (P1, P2) in tensor memory, s in sS.
// Out: (kappa, P1, P2, s") per substitution slot in tensor memory, l in sS.  s" = s kappa_next / kappa with
// `next` the neighbour towards the middle cell, from the rounded factors in double-double, rounded once.
// New slot j overwrites old columns that have been read by then: [0,128) is read in full first, and the
// old (P1, P2) slots 2(j-8), 2(j-8)+1 that new slot j >= 8 covers are <= j.
__device__ __forceinline__ void pipe_relayout(uint32_t t_lk, uint32_t t_pp, double* sS) {
  constexpr int N = kPipeN, MID = kPipeMid;
  // fully unrolled: lA, kA stay in registers (with run-time indices they go to local memory, 1 GB of
  // DRAM writes per cfg 3 launch)
  double lA[N], kA[N], l_up, unused;
#pragma unroll
  for (int c = 0; c < N; ++c) tmem_ld_2d(t_lk + pipe_col_lk(c), lA[c], kA[c]);
  tmem_ld_2d(t_lk + pipe_col_lk(MID) + 4, l_up, unused);
#pragma unroll
  for (int j = 0; j <= MID; ++j) {
    const int cl = MID - j, cr = MID + j;
    double p1l, p2l, p1r = 0.0, p2r = 0.0, s2l = 0.0, s2r = 0.0;
    tmem_ld_2d(t_pp + pipe_col_pp(cl), p1l, p2l);
    if (j > 0) {
      tmem_ld_2d(t_pp + pipe_col_pp(cr), p1r, p2r);
      s2l = dd_round(dd_div(two_prod(sS[cl * 32], kA[cl + 1]), dd_make(kA[cl])));
      s2r = dd_round(dd_div(two_prod(sS[cr * 32], kA[cr - 1]), dd_make(kA[cr])));
    }
    const uint32_t t = t_lk + 16 * j;
    tmem_st_2d(t, kA[cl], p1l); tmem_st_2d(t + 4, p2l, s2l);
    tmem_st_2d(t + 8, kA[cr], p1r); tmem_st_2d(t + 12, p2r, s2r);
    tmem_wait_st();
  }
  __syncwarp();
#pragma unroll
  for (int c = 1; c < N; ++c) sS[c * 32] = lA[c];
  sS[0] = l_up;                                     // cell 0 starts its chain: its own l is never read
  __syncwarp();
}

// EXT: 0 = constant coefficients; 1 = per-step boundary-flux scales; 2 = scales and runs of different time
// steps (the run loop costs the steady-state loop its uniform registers, ~8 %, so it has its own instance).
template <int MODE, int EXT_>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_tmem_pipe_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
                       int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt, FusedExtras ex) {
  constexpr int C = 32, N = kPipeN, MID = kPipeMid;
  constexpr bool EXT = EXT_ >= 1, RUNS = EXT_ == 2;
  constexpr int IT = 2;                              // slots per elimination group: a slot = a left and a right cell
  constexpr int NCH = (MID + 1) / IT;                // 16 slots: 15 cell pairs and the middle cell
  using TC = TmemChunk;
  // shared-memory slots (per lane) of the constants of a step
  enum { K_OML = 0, K_OMR = 1, K_AL = 2, K_GA = 7, K_INVD = 12, K_SLO = 13, K_SUP = 14, K_BSEP = 15, K_BFIRST = 16,
         K_SLOTS = 17 };
  extern __shared__ double smem[];
  __shared__ uint32_t tmem_base_slot;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  double* sS = smem + (size_t)warp * ((3 * N + K_SLOTS) * 32) + lane;   // s during set-up, then l; V; U; constants
  double* sv = sS + N * 32;
  double* su = sv + N * 32;
  double* sK = su + N * 32;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;

  if (warp == 0) tmem_alloc_512(&tmem_base_slot);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t t_lk = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 256;
  const uint32_t t_pp = t_lk + 128;                  // set-up only: (P1, P2) pairs in substitution order

  const int64_t n_ck_total = (ckpt != nullptr && ckpt_every > 0) ? nsteps / ckpt_every : 0;
  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
   const int64_t n_runs = (RUNS && ex.run_starts != nullptr) ? ex.n_runs : 1;
   int64_t s_glob = 0;                               // steps done so far
   for (int64_t run = 0; run < n_runs; ++run) {
    int64_t run_len = nsteps;
    double dt_run = dt[p];
    if (RUNS && ex.run_starts != nullptr) {
      const int64_t s0 = ex.run_starts[run];
      run_len = ex.run_starts[run + 1] - s0;
      dt_run = ex.dt_hist[s0 * ex.dt_ss + p * ex.dt_ps];
    }
    // ------------------------------------------------------------------ set-up (once per run)
    {
      PipeConsts Kc;
      pipe_setup(pr, p, lane, dt_run, theta, t_lk, t_pp, sS, sv, su, Kc);
      pipe_relayout(t_lk, t_pp, sS);
      // the per-lane constants of a step live in shared memory, so that the registers they would
      // occupy can hold the second tensor-memory chunk buffer
      sK[K_OML * 32] = Kc.om_l; sK[K_OMR * 32] = Kc.om_r;
#pragma unroll
      for (int l2 = 0; l2 < 5; ++l2) { sK[(K_AL + l2) * 32] = Kc.al[l2]; sK[(K_GA + l2) * 32] = Kc.ga[l2]; }
      sK[K_INVD * 32] = Kc.invD; sK[K_SLO * 32] = Kc.slo; sK[K_SUP * 32] = Kc.sup;
      sK[K_BSEP * 32] = Kc.bsep; sK[K_BFIRST * 32] = Kc.bfirst;
      __syncwarp();
    }
    auto K = [&](int slot) { return sK[slot * 32]; };
    // l of cell c (c = MID: l_down; slot 0: l_up of the middle cell): the left chain's and the first NRL of the
    // right chain's from registers, the others from shared memory (all from shared memory, the pipe that also
    // carries the shuffles is 70 % busy and the step 2 % slower; NRL >= 4 spills inside the loop)
    constexpr int NRL = 0;
    double lreg[MID + 1 + NRL];
#pragma unroll
    for (int c = 1; c <= MID + NRL; ++c) lreg[c] = sS[c * 32];
    auto Lc = [&](int c) { return (c >= 1 && c <= MID + NRL) ? lreg[c] : sS[c * 32]; };
    // x' = yhat - c x at a separator
    auto from_yhat = [&](double ys, double x_old) {
      return (MODE == 2) ? ys : ((MODE == 1) ? ys - x_old : fma(-cth, x_old, ys));
    };
    // g + c g_old: the uncorrected block solution of yhat
    auto ghat = [&](double g_new, double g_old) {
      return (MODE == 2) ? g_new : ((MODE == 1) ? g_new + g_old : fma(cth, g_old, g_new));
    };

    double w[C];
    load_state<C>(w_io + p * J, J, lane, w);

    // ------------------------------------------------------------------ time loop
    uint32_t ra[TC::REGS], rb[TC::REGS];
    double hh[N];                                    // h of the step being swept (hh[MID] is unused)
    // One step on w[0..N-1] in place; the cells outside the block that the end cells couple to
    // contribute oldL0 / oldR0 (first step of a segment only).  With PIPE the five PCR levels of
    // the previous step's separator system (rr in) are issued between the elimination groups, the
    // new separator values xs (and xsl = the left neighbour's) follow as soon as the last level is
    // done, and the substitution applies the delayed correction cell by cell.  base returns the
    // part of the next reduced right-hand side that does not depend on this step's sweeps.
    auto sweeps = [&](auto pipe_tag, double oldL0, double oldR0, double rr, double& xs, double& base, double sc) {
      constexpr bool PIPE = decltype(pipe_tag)::value;
      double hL = 0.0, hR = 0.0, uM = 0.0;
      double xsl = 0.0;
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        // the first two substitution chunks are fetched behind the last elimination group
        if (k == NCH - 1) { chunk_ld_after(t_lk, ra, hL); chunk_ld_after(t_lk + 16, rb, hR); }
        // One PCR level per group.  Warps issue in order, so the level's shuffles are issued HERE and
        // their consumer sits after the group's cell work: the ~26-cycle shuffle latency is covered.
        double rm = 0.0, rp = 0.0, fa = 0.0, fg = 0.0;
        const bool lvl = PIPE && k < 5;
        if (lvl) { rm = shfl_up(rr, 1 << k); rp = shfl_dn(rr, 1 << k); fa = K(K_AL + k); fg = K(K_GA + k); }
        if (PIPE && k == 5) { xs = from_yhat(rr * K(K_INVD), xs); xsl = shfl_up(xs, 1); }
        if (PIPE && k == 6) base = fma(xs, rth, EXT ? K(K_BSEP) * sc : K(K_BSEP));
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const int i = k * IT + ii;
          if (i < MID) {
            {
              const int c = i;                        // left cell, eliminated top-down
              const double gc = w[c];
              if (i == 0) {
                hL = EXT ? fma(K(K_BFIRST), sc, gc) : gc + K(K_BFIRST);
                if (!PIPE) hL = fma(-K(K_OML), oldL0, hL);
              } else {
                hL = fma(-Lc(c), hL, gc);
              }
              hh[c] = hL;
              if (MODE == 0) w[c] = cth * gc;
            }
            {
              const int c = N - 1 - i;                // right cell, eliminated bottom-up
              const double gc = w[c];
              if (i == 0) {
                hR = gc;
                if (!PIPE) hR = fma(-K(K_OMR), oldR0, hR);
              } else {
                hR = fma(-Lc(c), hR, gc);
              }
              hh[c] = hR;
              if (MODE == 0) w[c] = cth * gc;
            }
          } else {
            // the middle cell takes both sweeps: h = g - l_dn h_{m-1} - l_up h_{m+1}
            const double gc = w[MID];
            uM = fma(-Lc(0), hR, fma(-Lc(MID), hL, gc));
            if (MODE == 0) w[MID] = cth * gc;
          }
        }
        if (lvl) rr = fma(fg, rp, fma(fa, rm, rr));
      }
      // substitution outwards from the middle on u = yhat / kappa, one chunk (kappa, P1, P2, s" of a left
      // and a right cell) per link; z = kappa u - c g and the delayed correction follow off the chain
      double uL = uM, uR = uM;
#pragma unroll
      for (int jj = 0; jj < NCH; ++jj) {              // two links (slots 2jj in ra, 2jj+1 in rb) per group
        TC::wait(ra);
        // chains first, then kappa u - c g, then the xsl products of the group, then the xs products:
        // consecutive FMAs that share an operand can take it from the operand-reuse cache instead of
        // the register file, whose bandwidth is what bounds these loops (fp64_operands.cu)
        double al[IT], ar[IT];
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const uint32_t* cur = ii ? rb : ra;
          const int j = jj * IT + ii;
          if (j == 0) { al[ii] = uM; ar[ii] = 0.0; continue; }
          uL = fma(-chunk_dbl(cur, 3), uL, hh[MID - j]); al[ii] = uL;
          uR = fma(-chunk_dbl(cur, 7), uR, hh[MID + j]); ar[ii] = uR;
        }
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const uint32_t* cur = ii ? rb : ra;
          const int j = jj * IT + ii;
          al[ii] = (MODE == 2) ? chunk_dbl(cur, 0) * al[ii] : fma(chunk_dbl(cur, 0), al[ii], -w[MID - j]);
          if (j > 0) ar[ii] = (MODE == 2) ? chunk_dbl(cur, 4) * ar[ii] : fma(chunk_dbl(cur, 4), ar[ii], -w[MID + j]);
        }
        if (PIPE) {
#pragma unroll
          for (int ii = 0; ii < IT; ++ii) {
            const uint32_t* cur = ii ? rb : ra;
            al[ii] = fma(chunk_dbl(cur, 1), xsl, al[ii]);
            if (jj * IT + ii > 0) ar[ii] = fma(chunk_dbl(cur, 5), xsl, ar[ii]);
          }
#pragma unroll
          for (int ii = 0; ii < IT; ++ii) {
            const uint32_t* cur = ii ? rb : ra;
            al[ii] = fma(chunk_dbl(cur, 2), xs, al[ii]);
            if (jj * IT + ii > 0) ar[ii] = fma(chunk_dbl(cur, 6), xs, ar[ii]);
          }
        }
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const int j = jj * IT + ii;
          w[MID - j] = al[ii];
          if (j > 0) w[MID + j] = ar[ii];
        }
        if (jj + 1 < NCH) {
          chunk_ld_after(t_lk + 16 * (IT * jj + 2), ra, uL);
          chunk_ld_after(t_lk + 16 * (IT * jj + 3), rb, uR);
        }
      }
    };
    // reduced right-hand side of the step just swept:  base - slo ghat_last - sup ghat_first(next lane)
    auto reduced_rhs = [&](double base, double g0_old, double gl_old) {
      const double e_first_next = shfl_dn(ghat(w[0], g0_old), 1);     // lane 31: sup == 0
      return fma(-K(K_SUP), e_first_next, fma(-K(K_SLO), ghat(w[N - 1], gl_old), base));
    };

    // Boundary-flux scale of step s, loaded one step ahead of its use.  A lane needs at most one of
    // the two histories: beta[0] enters at lane 0 only (K_BFIRST is 0 elsewhere), beta[J-1] at the
    // lane whose separator is cell J-1 (K_BSEP is 0 elsewhere, and that lane is not lane 0: J > 512).
    const double* hist = nullptr;
    if (EXT) {
      hist = (lane == 0) ? ex.scale_left : ex.scale_right;
      if (hist != nullptr) hist += p * ex.sc_ps;
    }
    const int64_t hist_step = (EXT && hist != nullptr) ? ex.sc_ss : 0;
    if (EXT && hist == nullptr) hist = &kUnitScale;  // no history at this end: the scale stays 1
    if (EXT) hist += s_glob * hist_step;              // -> the scale of the next step to be taken
    const double* const hist_last = EXT ? hist + (nsteps - 1 - s_glob) * hist_step : nullptr;
    auto advance = [&]() { if (hist != hist_last) hist += hist_step; };
    int64_t remaining = run_len;
    while (remaining > 0) {
      // steps until the true state is needed again: the next checkpoint or the end of the run
      int64_t L = remaining;
      if (n_ck_total > 0) {
        const int64_t to_ck = ckpt_every - s_glob % ckpt_every;
        if (s_glob / ckpt_every < n_ck_total && to_ck < L) L = to_ck;
      }
      double sc = 1.0, sc_n = 1.0;
      if (EXT) { sc = *hist; advance(); sc_n = *hist; }
      // -------- first step of the segment: sweeps on the true state
      double xs = w[C - 1];
      double wl = shfl_up(xs, 1);
      if (lane == 0) wl = 0.0;
      // the true state in uncorrected form: g_old = x + V xsl + U xs (only its end cells are needed)
      double g0_old = fma(su[0], xs, fma(sv[0], wl, w[0]));
      double gl_old = fma(su[(N - 1) * 32], xs, fma(sv[(N - 1) * 32], wl, w[N - 1]));
      double base = fma(xs, rth, EXT ? K(K_BSEP) * sc : K(K_BSEP));
      sweeps(std::false_type{}, wl, xs, 0.0, xs, base, sc);
      double rr = reduced_rhs(base, g0_old, gl_old);
      // -------- pipelined steps: PCR of step n overlaps the sweeps of step n+1
      for (int64_t i = 1; i < L; ++i) {
        if (EXT) { sc = sc_n; advance(); sc_n = *hist; }
        g0_old = w[0]; gl_old = w[N - 1];
        sweeps(std::true_type{}, 0.0, 0.0, rr, xs, base, sc);
        rr = reduced_rhs(base, g0_old, gl_old);
      }
      // -------- end of the segment: finish the last step and materialise the true state
      {
        double ys = rr;
#pragma unroll
        for (int l = 0; l < 5; ++l) {
          const double rm = shfl_up(ys, 1 << l), rp = shfl_dn(ys, 1 << l);
          ys = fma(K(K_GA + l), rp, fma(K(K_AL + l), rm, ys));
        }
        xs = from_yhat(ys * K(K_INVD), xs);
        const double xsl_end = shfl_up(xs, 1);      // lane 0: V == 0, value unused
#pragma unroll
        for (int c = 0; c < N; ++c) w[c] = fma(-su[c * 32], xs, fma(-sv[c * 32], xsl_end, w[c]));
        w[C - 1] = xs;
      }
      remaining -= L;
      s_glob += L;
      if (n_ck_total > 0 && s_glob % ckpt_every == 0 && s_glob / ckpt_every <= n_ck_total)
        store_state<C>(ckpt + ((s_glob / ckpt_every - 1) * pr.P + p) * J, J, lane, w);
    }
    store_state<C>(w_io + p * J, J, lane, w);
    __syncwarp();
   }
  }

  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc_512(tmem_base);
}

// ------------------------------------------------------------------------------------------
static int device_limits(int& sms, int& max_smem) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return 0;
}

template <int C, int MODE, bool RB, bool QREG, bool EXT>
static int launch_fused(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                        int64_t ckpt_every, double* w, double* ckpt, const FusedExtras& ex, cudaStream_t stream) {
  constexpr int NS = (C > 1 ? C - 1 : 1);
  constexpr size_t per_warp = (size_t)7 * NS * 32 * sizeof(double);   // l, kappa, V, U + three set-up staging arrays
  int sms = 0, max_smem = 0;
  int st = device_limits(sms, max_smem);
  if (st) return st;
  int warps = (int)((size_t)max_smem / per_warp);
  if (warps > kFusedThreads / 32) warps = kFusedThreads / 32;
  if (warps < 1) return FVM_ERR_UNSUPPORTED;
  if ((int64_t)warps > pr.P) warps = (int)pr.P;
  const size_t smem = per_warp * warps;
  auto kern = fused_kernel<C, MODE, RB, QREG, EXT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, warps * 32, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) per_sm = 1;
  int64_t blocks = (pr.P + warps - 1) / warps;
  const int64_t resident = (int64_t)sms * per_sm;
  if (blocks > resident) blocks = resident;      // persistent: each warp strides over problems
  kern<<<(unsigned)blocks, warps * 32, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex);
  return (int)cudaGetLastError();
}

template <int MODE, int EXT>
static int launch_fused_tmem(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                             int64_t ckpt_every, double* w, double* ckpt, const FusedExtras& ex, cudaStream_t stream) {
  constexpr int warps = kFusedThreads / 32;
  constexpr size_t smem = (size_t)warps * (3 * 31 + 17) * 32 * sizeof(double);   // s/l, V, U and the constants
  int sms = 0, max_smem = 0;
  int st = device_limits(sms, max_smem);
  if (st) return st;
  int64_t blocks = (pr.P + warps - 1) / warps;
  if (blocks > sms) blocks = sms;                // one CTA per SM: it holds the SM's whole TMEM
  auto kern = fused_tmem_pipe_kernel<MODE, EXT>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<(unsigned)blocks, warps * 32, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex);
  return (int)cudaGetLastError();
}

template <int C, bool EXT>
static int dispatch_mode(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                         int64_t ckpt_every, double* w, double* ckpt, const FusedExtras& ex, cudaStream_t stream) {
  const bool rb = pr.right_kind == FVM_BC_ROBIN && pr.right_flux != nullptr &&
                  ((pr.J - 1) % C) != C - 1;
  constexpr bool Q = true;
  if (rb) return launch_fused<C, 0, true, Q, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
  if (C == 32) {
    if (EXT && ex.run_starts != nullptr) {
      if (theta == 0.5) return launch_fused_tmem<1, 2>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
      if (theta == 1.0) return launch_fused_tmem<2, 2>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
      return launch_fused_tmem<0, 2>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
    }
    constexpr int E = EXT ? 1 : 0;
    if (theta == 0.5) return launch_fused_tmem<1, E>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
    if (theta == 1.0) return launch_fused_tmem<2, E>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
    return launch_fused_tmem<0, E>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
  }
  if (theta == 0.5) return launch_fused<C, 1, false, Q, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
  if (theta == 1.0) return launch_fused<C, 2, false, Q, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
  return launch_fused<C, 0, false, Q, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, stream);
}

template <bool EXT>
static int dispatch_cells(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                          int64_t ckpt_every, double* w, double* ckpt, const FusedExtras& ex, cudaStream_t s) {
  const int64_t J = pr.J;
  if (J <= 32) return dispatch_mode<1, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
  if (J <= 64) return dispatch_mode<2, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
  if (J <= 128) return dispatch_mode<4, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
  if (J <= 256) return dispatch_mode<8, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
  if (J <= 512) return dispatch_mode<16, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
  return dispatch_mode<32, EXT>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, ex, s);
}

// fused_cta.cu
int run_fused_cta(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps, int64_t ckpt_every,
                  double* w, double* ckpt, cudaStream_t stream);

}  // namespace fvm

extern "C" int64_t fvm_fused_max_cells(void) { return 8192; }

static int check_fused_args(const fvm_problem* prob, const double* dt, double theta, int64_t nsteps,
                            int64_t checkpoint_every, const double* w) {
  if (!prob || !prob->faces || !prob->a || !prob->d || !dt || !w) return FVM_ERR_INVALID_ARG;
  if (prob->J < 3 || prob->P < 1 || nsteps < 0 || checkpoint_every < 0) return FVM_ERR_INVALID_ARG;
  if (!(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (prob->scheme < FVM_CENTRAL || prob->scheme > FVM_EXPONENTIAL) return FVM_ERR_INVALID_ARG;
  if ((prob->left_kind | prob->right_kind | prob->alpha_left | prob->alpha_right) & ~1)
    return FVM_ERR_INVALID_ARG;
  return FVM_OK;
}

extern "C" int fvm_theta_run_fused(const fvm_problem* prob, const double* dt, double theta,
                                   int64_t nsteps, int64_t checkpoint_every, double* w,
                                   double* ckpt, void* stream) {
  int st = check_fused_args(prob, dt, theta, nsteps, checkpoint_every, w);
  if (st) return st;
  if (prob->J > fvm_fused_max_cells()) return FVM_ERR_UNSUPPORTED;
  if (nsteps == 0) return FVM_OK;
  cudaStream_t s = (cudaStream_t)stream;
  using namespace fvm;
  const FusedExtras none = {nullptr, nullptr, 0, 0, nullptr, 0, 0, nullptr, 0};
  if (prob->J <= 32 * kFusedMaxC) return dispatch_cells<false>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, none, s);
  return run_fused_cta(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
}

extern "C" int64_t fvm_fused_ex_max_cells(void) { return 32 * fvm::kFusedMaxC; }

extern "C" int fvm_theta_run_fused_ex(const fvm_problem* prob, const double* dt, double theta, int64_t nsteps,
                                      int64_t checkpoint_every, double* w, double* ckpt,
                                      const double* scale_left, const double* scale_right, int64_t scale_ps,
                                      int64_t scale_ss, const double* dt_hist, int64_t dt_ps, int64_t dt_ss,
                                      const int64_t* run_starts, int64_t n_runs, void* stream) {
  int st = check_fused_args(prob, dt, theta, nsteps, checkpoint_every, w);
  if (st) return st;
  if ((dt_hist != nullptr) != (run_starts != nullptr)) return FVM_ERR_INVALID_ARG;
  if (run_starts != nullptr && n_runs < 1) return FVM_ERR_INVALID_ARG;
  if (scale_left && !(prob->left_kind == FVM_BC_ROBIN && prob->left_flux)) return FVM_ERR_INVALID_ARG;
  if (scale_right && !(prob->right_kind == FVM_BC_ROBIN && prob->right_flux)) return FVM_ERR_INVALID_ARG;
  if (prob->J > fvm_fused_ex_max_cells()) return FVM_ERR_UNSUPPORTED;
  if (nsteps == 0) return FVM_OK;
  using namespace fvm;
  const FusedExtras ex = {scale_left, scale_right, scale_ps, scale_ss, dt_hist, dt_ps, dt_ss, run_starts, n_runs};
  return dispatch_cells<true>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, ex, (cudaStream_t)stream);
}
