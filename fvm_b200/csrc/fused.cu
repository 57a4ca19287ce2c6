// Variant 2: one warp per system, everything on chip between checkpoints.
//
// Partition: lane l owns the C contiguous cells [l*C, (l+1)*C); the first N = C-1 are
// "interior" cells, the last one is the lane's SEPARATOR.  With the separators known the
// 32 interior blocks decouple, so (per step, all with coefficients factored ONCE per launch):
//   1. r~  = inv*(B w)                      3 FMA / cell   (B's off-diagonals are -c*p, -c*q)
//   2. local forward / backward sweep       2 FMA / cell   g = T^-1 r   (T = interior block of A)
//   3. reduced system on the 32 separators  precomputed parallel cyclic reduction, 5 levels
//                                           of 2 shuffles + 2 FMA, no divisions
//   4. x = g - v*xs(l-1) - u*xs(l)          2 FMA / cell   (v, u = left/right spikes T^-1 e)
// State w[C] lives in registers for the whole launch.  Five coefficients per interior cell
// (p, q, b, v, u) must be read every step; where they live decides the speed, because a
// step is bound by on-chip bandwidth, not arithmetic:
//   * fused_kernel      (any J <= 1024): q in registers, p/b/v/u in shared memory
//                       ([coef][c][lane], lane-consecutive doubles: conflict-free).
//   * fused_tmem_kernel (512 < J <= 1024): q in registers, p/b/v/u in TENSOR MEMORY.
//                       TMEM (256 KB/SM) is idle in this workload -- there is no MMA -- and
//                       its tcgen05.ld path is separate from, and wider than, the 128 B/clk
//                       shared-memory pipe (profiles/microbench/r01_tmem_bw_b200.txt).  Each
//                       warp owns 256 columns of its lane quadrant: columns [0,128) hold
//                       (p_c, b_c) pairs, [128,256) hold (v_c, u_c) pairs, 4 columns per cell.
//                       Shared memory is only a staging area during set-up.
// HBM traffic: J doubles in, J doubles out, J per checkpoint.
//
// Assembly (fvm_device.cuh) and factorisation run inside the same launch, one lane per
// owned cell, so a launch needs only the mesh, the per-problem scalars and the state.
#include "fvm_device.cuh"
#include "tmem.cuh"
#include <cstdlib>
#include <type_traits>

namespace fvm {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kFusedMaxC = 32;
constexpr int kFusedThreads = 256;   // upper bound used for register allocation

__device__ __forceinline__ double shfl_up(double x, int s) { return __shfl_up_sync(kFull, x, s); }
__device__ __forceinline__ double shfl_dn(double x, int s) { return __shfl_down_sync(kFull, x, s); }

// Per-lane constants of one system, fixed for the whole launch.
struct Reduced {
  double al[5], ga[5];           // PCR elimination factors of the separator system, per level
  double invD;                   // 1 / final PCR diagonal
  double slo, sup;               // separator's row of A (off-diagonals)
  double sblo, sbdi, sbup;       // separator's row of B
  double bfirst, bint, bsep;     // beta entries as they enter the sweeps
  int rb_c;                      // RB only: position of cell J-1 inside its lane
};

// Where phase 1 puts (p_c, b_c).
struct PutSmem {
  double* sp; double* sb;
  __device__ __forceinline__ void operator()(int c, double p, double b) const { sp[c * 32] = p; sb[c * 32] = b; }
};
struct PutTmem {
  uint32_t t_pb;
  __device__ __forceinline__ void operator()(int c, double p, double b) const { tmem_st_2d(t_pb + 4 * c, p, b); }
};

// Phase 1: assemble this lane's rows of A, factor the interior block (p, q, b), run the
// forward sweep of the left spike (into sv) and keep the separator's row.
template <int C, bool RB, class Put>
__device__ __forceinline__ void setup_rows(const fvm_problem& pr, int64_t p, int lane, double dt_p,
                                           double theta, const Put& put, double* sq, double* sv,
                                           Reduced& R, double& sdi) {
  constexpr int N = C - 1;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
  const double dt_theta = mul(dt_p, theta);
  double beta_l = 0.0, beta_r = 0.0;
  if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
    beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
  if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
    beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
  R.slo = 0.0; sdi = 1.0; R.sup = 0.0;
  R.bfirst = 0.0; R.bint = 0.0; R.bsep = 0.0;
  R.rb_c = RB ? (int)((J - 1) % C) : -1;
  double q_prev = 0.0, gv_prev = 0.0;
  const int64_t j0 = (int64_t)lane * C;
  // sliding window over faces: one expm1 and one face (5 divisions) per row
  double D0 = cell_D(pr, p, j0);
  Face FL = problem_face(pr, p, j0 - 1, cell_D(pr, p, j0 - 1), D0);
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
      const double inv = 1.0 / (a.di - a.lo * q_prev);
      const double pc = a.lo * inv;
      q_prev = a.up * inv;
      gv_prev = (c == 0) ? pc : -pc * gv_prev;      // forward sweep of the left spike
      put(c, pc, inv * (rth - cth * a.di));
      sq[c * 32] = q_prev;
      sv[c * 32] = gv_prev;
      if (j == 0) R.bfirst = beta_l * inv;
      if (RB && j == J - 1) R.bint = beta_r * inv;
    } else {
      R.slo = a.lo; sdi = a.di; R.sup = a.up;
      if (j == 0) R.bsep += beta_l;
      if (j == J - 1) R.bsep += beta_r;
    }
  }
  R.sblo = -cth * R.slo; R.sbdi = rth - cth * sdi; R.sbup = -cth * R.sup;
}

// Parallel cyclic reduction of the 32-unknown separator system (RL, RD, RU per lane): the
// elimination factors of each of the 5 levels and the final diagonal are kept, so that a
// step only applies them to its right-hand side (pcr_solve).
__device__ __forceinline__ void pcr_factor(int lane, double RL, double RD, double RU, Reduced& R) {
#pragma unroll
  for (int l = 0; l < 5; ++l) {
    const int s = 1 << l;
    const bool hm = lane >= s, hp = lane + s < 32;
    const double Dm = shfl_up(RD, s), Dp = shfl_dn(RD, s);
    const double Lm = shfl_up(RL, s), Um = shfl_up(RU, s);
    const double Lp = shfl_dn(RL, s), Up = shfl_dn(RU, s);
    const double a_ = hm ? -RL / Dm : 0.0;
    const double g_ = hp ? -RU / Dp : 0.0;
    R.al[l] = a_; R.ga[l] = g_;
    RD = RD + (hm ? a_ * Um : 0.0) + (hp ? g_ * Lp : 0.0);
    RL = hm ? a_ * Lm : 0.0;
    RU = hp ? g_ * Up : 0.0;
  }
  R.invD = 1.0 / RD;
}

// Phase 2: backward sweeps of both spikes (sv: gv -> v; su <- u, reading q from sq, which
// may alias su), the separator system and its PCR factors.
template <int C>
__device__ __forceinline__ void setup_reduced(int lane, const double* sq, double* sv, double* su,
                                              double sdi, Reduced& R) {
  constexpr int N = C - 1;
  double RL = R.slo, RD = sdi, RU = R.sup;
  if (N > 0) {
    double vn = sv[(N - 1) * 32];
    double un = sq[(N - 1) * 32];
    su[(N - 1) * 32] = un;
    const double vlast = vn, ulast = un;
#pragma unroll 1
    for (int c = N - 2; c >= 0; --c) {
      const double qc = sq[c * 32];
      vn = sv[c * 32] - qc * vn;
      un = -qc * un;
      sv[c * 32] = vn;
      su[c * 32] = un;
    }
    const double v0n = shfl_dn(vn, 1), u0n = shfl_dn(un, 1);   // next lane's first interior cell
    RL = -R.slo * vlast;
    RD = sdi - R.slo * ulast - R.sup * v0n;         // sup == 0 on lane 31
    RU = -R.sup * u0n;
  }
  pcr_factor(lane, RL, RD, RU, R);
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

// r~ of one interior cell.  MODE 0: any theta in (0,1];  1: Crank-Nicolson (c = 1);  2: implicit (c = 0).
template <int MODE>
__device__ __forceinline__ double rhs_cell(double pc, double qc, double bc, double wm, double wc,
                                           double wp, double cth) {
  if (MODE == 2) return bc * wc;
  const double t = fma(qc, wp, pc * wm);
  return (MODE == 1) ? fma(bc, wc, -t) : fma(-cth, t, bc * wc);
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

// ------------------------------------------------------------------------------------------
// Shared-memory variant.  RB: beta[J-1] is nonzero AND cell J-1 is an interior cell of its
// lane (slow path).  QREG: q in registers (4 instead of 6 shared-memory reads per cell-update).
template <int C, int MODE, bool RB, bool QREG>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
             int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt) {
  constexpr int N = C - 1;
  constexpr int NS = (N > 0 ? N : 1);
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  constexpr int NARR = QREG ? 4 : 5;
  double* sp = smem + (size_t)warp * (NARR * NS * 32) + lane;   // [c*32] indexing below
  double* sb = sp + NS * 32;
  double* sv = sb + NS * 32;
  double* su = sv + NS * 32;
  double* sq = QREG ? su : su + NS * 32;   // QREG: q is staged in u's slots during set-up only
  const int64_t J = pr.J;
  const double cth = (1.0 - theta) / theta;

  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
    Reduced R;
    double sdi;
    setup_rows<C, RB>(pr, p, lane, dt[p], theta, PutSmem{sp, sb}, sq, sv, R, sdi);
    double qr[NS];
    if (QREG) {
#pragma unroll
      for (int c = 0; c < NS; ++c) qr[c] = (c < N) ? sq[c * 32] : 0.0;
    }
    setup_reduced<C>(lane, sq, sv, su, sdi, R);
    __syncwarp();

    double w[C];
    load_state<C>(w_io + p * J, J, lane, w);

    int64_t until_ckpt = ckpt_every, n_ckpt = 0;
    for (int64_t n = 0; n < nsteps; ++n) {
      double wl = shfl_up(w[C - 1], 1);
      double wr = shfl_dn(w[0], 1);
      if (lane == 0) wl = 0.0;
      if (lane == 31) wr = 0.0;
      double rr;
      if (N > 0) {
        double wm = wl, g = 0.0;
#pragma unroll
        for (int c = 0; c < N; ++c) {
          const double pc = sp[c * 32];
          const double wc = w[c];
          double r = rhs_cell<MODE>(pc, QREG ? qr[c] : sq[c * 32], sb[c * 32], wm, wc, w[c + 1], cth);
          if (c == 0) r += R.bfirst;
          if (RB) { if (c == R.rb_c) r += R.bint; }
          g = (c == 0) ? r : fma(-pc, g, r);
          w[c] = g;
          wm = wc;
        }
        const double rs = fma(R.sbup, wr, fma(R.sbdi, w[C - 1], fma(R.sblo, wm, R.bsep)));
#pragma unroll
        for (int c = N - 2; c >= 0; --c) w[c] = fma(-(QREG ? qr[c] : sq[c * 32]), w[c + 1], w[c]);
        const double g0n = shfl_dn(w[0], 1);
        rr = fma(-R.sup, g0n, fma(-R.slo, w[N - 1], rs));
      } else {
        rr = fma(R.sbup, wr, fma(R.sbdi, w[0], fma(R.sblo, wl, R.bsep)));
      }
      const double xs = pcr_solve(rr, R);
      if (N > 0) {
        const double xsl = shfl_up(xs, 1);          // lane 0: v == 0, value unused
#pragma unroll
        for (int c = 0; c < N; ++c) w[c] = fma(-su[c * 32], xs, fma(-sv[c * 32], xsl, w[c]));
      }
      w[C - 1] = xs;

      if (ckpt != nullptr && ckpt_every > 0 && --until_ckpt == 0) {
        until_ckpt = ckpt_every;
        store_state<C>(ckpt + (n_ckpt * pr.P + p) * J, J, lane, w);
        ++n_ckpt;
      }
    }
    store_state<C>(w_io + p * J, J, lane, w);
    __syncwarp();
  }
}

// Set-up shared by the tensor-memory kernels (C = 32, N = 31 interior cells per lane, twisted at
// M = 16): assembles this lane's rows of A, factors the interior block from both ends (e, b ->
// tensor memory at t_eb; s -> sS and the registers sr), computes both spikes (sv = V, su = U in
// shared memory, natural cell order), the separator system and its PCR factors.
//
// LK (the 4-operation Crank-Nicolson sweep of fused_tmem_pipe_kernel<3, IT>): tensor memory gets
// (l, kappa) instead of (e, b) -- l = the un-normalised elimination factor A_c,prev / pivot_prev,
// kappa = 1 / (theta pivot_c) -- R.bfirst = theta beta_0 (un-normalised), and om_l / om_r return
// (1 - theta) times the coupling of the block's end cells to the two separators.
template <bool LK>
__device__ __forceinline__ void tmem_setup_common(const fvm_problem& pr, int64_t p, int lane, double dt_p, double theta,
                                                  uint32_t t_eb, double* sS, double* sv, double* su, Reduced& R,
                                                  double (&sr)[31], double& delta, double& om_l, double& om_r) {
  constexpr int C = 32, N = 31, M = 16;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
    double sdi = 1.0;
    {
      const double dt_theta = mul(dt_p, theta);
      double beta_l = 0.0, beta_r = 0.0;
      if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
        beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
      if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
        beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
      R.bfirst = 0.0; R.bint = 0.0; R.bsep = 0.0; R.rb_c = -1;
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
      double sl_prev = 0.0, sr_prev = 0.0, gv_prev = 0.0, gu_prev = 0.0;
      double invl_prev = 0.0, invr_prev = 0.0;
      om_l = 0.0; om_r = 0.0;
#pragma unroll 1
      for (int i = 0; i < M; ++i) {
        {                                            // left cell i: top-down elimination
          const double lDp = cell_D(pr, p, j0 + i + 1);
          const Face lFR = problem_face(pr, p, j0 + i, lD0, lDp);
          const Row a = row_of(i, lFL, lFR);
          lD0 = lDp; lFL = lFR;
          const double inv = 1.0 / (a.di - a.lo * sl_prev);
          const double e = a.lo * inv;
          sl_prev = a.up * inv;
          gv_prev = (i == 0) ? e : -e * gv_prev;     // elimination sweep of the left spike
          sS[i * 32] = sl_prev; sv[i * 32] = gv_prev; su[i * 32] = 0.0;
          if (LK) {
            tmem_st_2d(t_eb + 8 * i, a.lo * invl_prev, inv * rth);
            if (i == 0) om_l = (1.0 - theta) * a.lo;
            if ((int64_t)lane * C + i == 0) R.bfirst = theta * beta_l;
            invl_prev = inv;
          } else {
            tmem_st_2d(t_eb + 8 * i, e, inv * (rth - cth * a.di));
            if ((int64_t)lane * C + i == 0) R.bfirst = beta_l * inv;
          }
        }
        if (i < N - M) {                             // right cell 30-i: bottom-up elimination
          const int c = N - 1 - i;
          const double rDm = cell_D(pr, p, j0 + c - 1);
          const Face rFL = problem_face(pr, p, j0 + c - 1, rDm, rD0);
          const Row a = row_of(c, rFL, rFR);
          rD0 = rDm; rFR = rFL;
          const double inv = 1.0 / (a.di - a.up * sr_prev);
          const double e = a.up * inv;
          sr_prev = a.lo * inv;
          gu_prev = (i == 0) ? e : -e * gu_prev;     // elimination sweep of the right spike
          sS[c * 32] = sr_prev; su[c * 32] = gu_prev; sv[c * 32] = 0.0;
          if (LK) {
            tmem_st_2d(t_eb + 8 * i + 4, a.up * invr_prev, inv * rth);
            if (i == 0) om_r = (1.0 - theta) * a.up;
            invr_prev = inv;
          } else {
            tmem_st_2d(t_eb + 8 * i + 4, e, inv * (rth - cth * a.di));
          }
        }
      }
      R.slo = asep.lo; sdi = asep.di; R.sup = asep.up;
      const int64_t js = (int64_t)lane * C + C - 1;
      if (js == J - 1) R.bsep += beta_r;
      R.sblo = -cth * R.slo; R.sbdi = rth - cth * sdi; R.sbup = -cth * R.sup;
    }
#pragma unroll
    for (int c = 0; c < N; ++c) sr[c] = sS[c * 32];
    delta = 1.0 / (1.0 - sr[M] * sr[M - 1]);
    {
      // substitution sweeps of both spikes, outwards from the meeting point
      double xv = (sv[M * 32] - sr[M] * sv[(M - 1) * 32]) * delta;
      double xu = (su[M * 32] - sr[M] * su[(M - 1) * 32]) * delta;
      sv[M * 32] = xv; su[M * 32] = xu;
      double lv = xv, lu = xu;
#pragma unroll 1
      for (int c = M - 1; c >= 0; --c) {
        const double sc = sS[c * 32];
        lv = sv[c * 32] - sc * lv; lu = su[c * 32] - sc * lu;
        sv[c * 32] = lv; su[c * 32] = lu;
      }
      double rv = xv, ru = xu;
#pragma unroll 1
      for (int c = M + 1; c < N; ++c) {
        const double sc = sS[c * 32];
        rv = sv[c * 32] - sc * rv; ru = su[c * 32] - sc * ru;
        sv[c * 32] = rv; su[c * 32] = ru;
      }
      // lv, lu = spikes at cell 0; rv, ru = spikes at cell N-1
      const double v0n = shfl_dn(lv, 1), u0n = shfl_dn(lu, 1);
      const double RL = -R.slo * rv;
      const double RD = sdi - R.slo * ru - R.sup * v0n;          // sup == 0 on lane 31
      const double RU = -R.sup * u0n;
      pcr_factor(lane, RL, RD, RU, R);
    }
}

// ------------------------------------------------------------------------------------------
// Tensor-memory variant, C = 32 (512 < J <= 1024), 8 warps per CTA, one CTA per SM.
//
// A step of the shared-memory kernel is bound by the latency of its dependent-FMA chains
// (31 links forward, 30 backward, ~9 cycles each; profiles/r01_fused_v3_tmem.summary.txt), not
// by the fp64 pipe.  This kernel therefore factors each lane's interior block T with a TWISTED
// factorisation: cells 0..15 are eliminated top-down, cells 30..16 bottom-up, the two sweeps
// meet between cells 15 and 16, and substitution runs outwards from there.  Same flops, same
// five coefficients per cell, but two independent chains of half the length: 16 + 1 + 16
// dependent links per step instead of 61.  Per cell:  e = elimination coefficient (times the
// neighbour the sweep comes from), s = substitution coefficient (times the neighbour the
// sweep moves towards), b = diagonal of inv*B; inv*B's off-diagonals are -c*e and -c*s.
//
// TMEM layout per warp (256 columns of its lane quadrant), in the order the loop consumes it:
//   columns [0,128)  : slot i = (e, b) of left cell i, then (e, b) of right cell 30-i   (8 columns)
//   columns [128,256): slot i = (v, u) of left cell i, then (v, u) of right cell 30-i
// s lives in registers (it is read twice per step).  IT = slots per tcgen05.ld (2 -> .x16,
// 4 -> .x32); loads are double-buffered so the next chunk is in flight while one is consumed.
template <int IT> struct TmemChunk;
template <> struct TmemChunk<4> {
  static constexpr int REGS = 32;
  static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[32]) { tmem_ld_x32(t, r); }
  static __device__ __forceinline__ void wait(uint32_t (&r)[32]) { tmem_wait_ld_x32(r); }
};
template <> struct TmemChunk<1> {
  static constexpr int REGS = 8;
  static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[8]) { tmem_ld_x8(t, r); }
  static __device__ __forceinline__ void wait(uint32_t (&r)[8]) { tmem_wait_ld_x8(r); }
};
template <> struct TmemChunk<2> {
  static constexpr int REGS = 16;
  static __device__ __forceinline__ void ld(uint32_t t, uint32_t (&r)[16]) { tmem_ld_x16(t, r); }
  static __device__ __forceinline__ void wait(uint32_t (&r)[16]) { tmem_wait_ld_x16(r); }
};

// Issue a chunk load and thread a value through an empty volatile asm behind it: whatever is
// computed from `dep` afterwards is ordered after the ISSUE of the load.  This pins a prefetch
// ahead of the arithmetic it is meant to overlap; the compiler otherwise sinks the load down to
// its wait to save registers, and every chunk then pays the tensor-memory latency.
template <class TC, int REGS>
__device__ __forceinline__ void chunk_ld_after(uint32_t t, uint32_t (&r)[REGS], double& dep) {
  TC::ld(t, r);
  asm volatile("" : "+d"(dep));
}

// k-th double of a chunk (two 32-bit columns each)
__device__ __forceinline__ double chunk_dbl(const uint32_t* r, int k) { return __hiloint2double((int)r[2 * k + 1], (int)r[2 * k]); }

template <int MODE, int IT>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_tmem_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
                  int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt) {
  constexpr int C = 32, N = 31, M = 16;              // left cells 0..M-1, right cells N-1..M
  constexpr int NCH = M / IT;                        // chunks per region
  using TC = TmemChunk<IT>;
  extern __shared__ double smem[];
  __shared__ uint32_t tmem_base_slot;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  double* sS = smem + (size_t)warp * (3 * N * 32) + lane;        // set-up staging: s, v, u
  double* sv = sS + N * 32;
  double* su = sv + N * 32;
  const int64_t J = pr.J;
  const double cth = (1.0 - theta) / theta;

  if (warp == 0) tmem_alloc_512(&tmem_base_slot);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t t_eb = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 256;
  const uint32_t t_vu = t_eb + 128;

  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
    // ------------------------------------------------------------------ set-up (once)
    Reduced R;
    double delta;
    double sr[N];                                    // s for the whole launch
    double om_unused_l, om_unused_r;
    tmem_setup_common<false>(pr, p, lane, dt[p], theta, t_eb, sS, sv, su, R, sr, delta, om_unused_l, om_unused_r);
    {
#pragma unroll 1
      for (int i = 0; i < M; ++i) {
        tmem_st_2d(t_vu + 8 * i, sv[i * 32], su[i * 32]);
        if (i < N - M) tmem_st_2d(t_vu + 8 * i + 4, sv[(N - 1 - i) * 32], su[(N - 1 - i) * 32]);
      }
    }
    tmem_wait_st();
    __syncwarp();

    double w[C];
    load_state<C>(w_io + p * J, J, lane, w);

    // ------------------------------------------------------------------ time loop
    uint32_t ra[TC::REGS], rb[TC::REGS];
    int64_t until_ckpt = ckpt_every, n_ckpt = 0;
    TC::ld(t_eb, ra);                               // first (e, b) chunk of the first step
    for (int64_t n = 0; n < nsteps; ++n) {
      double wl = shfl_up(w[C - 1], 1);
      double wr = shfl_dn(w[0], 1);
      if (lane == 0) wl = 0.0;
      if (lane == 31) wr = 0.0;
      const double w_m_old = w[M], w_last_old = w[N - 1], w_sep = w[C - 1];
      double oldL = wl, gL = 0.0;                   // left sweep comes from cell -1 (previous separator)
      double oldR = w_sep, gR = 0.0;                // right sweep comes from cell 31 (own separator)
      // elimination: chunk k is consumed from one buffer while chunk k+1 (or the first (v, u)
      // chunk of the correction) is already on its way into the other
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        uint32_t* cur = (k & 1) ? rb : ra;
        if (k & 1) { TC::wait(rb); TC::ld(k + 1 < NCH ? t_eb + 8 * IT * (k + 1) : t_vu, ra); }
        else       { TC::wait(ra); TC::ld(k + 1 < NCH ? t_eb + 8 * IT * (k + 1) : t_vu, rb); }
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const int i = k * IT + ii;
          {
            const int c = i;
            const double e = chunk_dbl(cur, 4 * ii), b = chunk_dbl(cur, 4 * ii + 1);
            const double wc = w[c];
            const double wn = (c == M - 1) ? w_m_old : w[c + 1];
            double r = rhs_cell<MODE>(e, sr[c], b, oldL, wc, wn, cth);
            if (c == 0) r += R.bfirst;
            gL = (i == 0) ? r : fma(-e, gL, r);
            w[c] = gL;
            oldL = wc;
          }
          if (i < N - M) {
            const int c = N - 1 - i;
            const double e = chunk_dbl(cur, 4 * ii + 2), b = chunk_dbl(cur, 4 * ii + 3);
            const double wc = w[c];
            double r = rhs_cell<MODE>(e, sr[c], b, oldR, wc, w[c - 1], cth);
            gR = (i == 0) ? r : fma(-e, gR, r);
            w[c] = gR;
            oldR = wc;
          }
        }
      }
      const double rs = fma(R.sbup, wr, fma(R.sbdi, w_sep, fma(R.sblo, w_last_old, R.bsep)));
      // the sweeps meet between cells M-1 and M; substitute outwards
      w[M] = fma(-sr[M], w[M - 1], w[M]) * delta;
#pragma unroll
      for (int j = 0; j < M; ++j) {
        const int cl = M - 1 - j, cr = M + 1 + j;
        w[cl] = fma(-sr[cl], w[cl + 1], w[cl]);
        if (cr < N) w[cr] = fma(-sr[cr], w[cr - 1], w[cr]);
      }
      const double g0n = shfl_dn(w[0], 1);
      const double rr = fma(-R.sup, g0n, fma(-R.slo, w[N - 1], rs));
      const double xs = pcr_solve(rr, R);
      const double xsl = shfl_up(xs, 1);            // lane 0: v == 0, value unused
      // correction; NCH is even, so the first (v, u) chunk sits in ra again
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        uint32_t* cur = (k & 1) ? rb : ra;
        if (k & 1) { TC::wait(rb); TC::ld(k + 1 < NCH ? t_vu + 8 * IT * (k + 1) : t_eb, ra); }
        else       { TC::wait(ra); TC::ld(k + 1 < NCH ? t_vu + 8 * IT * (k + 1) : t_eb, rb); }
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const int i = k * IT + ii;
          w[i] = fma(-chunk_dbl(cur, 4 * ii + 1), xs, fma(-chunk_dbl(cur, 4 * ii), xsl, w[i]));
          if (i < N - M) {
            const int c = N - 1 - i;
            w[c] = fma(-chunk_dbl(cur, 4 * ii + 3), xs, fma(-chunk_dbl(cur, 4 * ii + 2), xsl, w[c]));
          }
        }
      }
      w[C - 1] = xs;

      if (ckpt != nullptr && ckpt_every > 0 && --until_ckpt == 0) {
        until_ckpt = ckpt_every;
        store_state<C>(ckpt + (n_ckpt * pr.P + p) * J, J, lane, w);
        ++n_ckpt;
      }
    }
    TC::wait(ra);                                   // drain the prefetch of the step that never comes
    store_state<C>(w_io + p * J, J, lane, w);
    __syncwarp();
  }

  tmem_fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc_512(tmem_base);
}

// ------------------------------------------------------------------------------------------
// Tensor-memory variant with a DELAYED CORRECTION, the default for 512 < J <= 1024.
//
// In fused_tmem_kernel a step is  sweeps -> separator PCR -> correction, and the PCR (five
// levels of shuffle + FMA, a pure latency chain that leaves the fp64 pipe idle) cannot overlap
// anything, because the next sweeps need the corrected state.  But the whole step is linear with
// constant coefficients.  Keep, between steps, the UNCORRECTED block solutions g and the reduced
// right-hand side rr; the true state is x = g - V xsl - U xs with the separator values
// xs = R^-1 rr still to come.  Then
//     g' = T^-1 (B x)_interior = T^-1 B_ii g  +  P1 xsl  +  P2 xs ,
// and the sweeps T^-1 B_ii g do not need xs: they run while the PCR of the PREVIOUS step is in
// flight, its levels interleaved with the elimination chunks.  V and U stay in shared memory for
// the points where the true state is needed (checkpoints, the end of the launch), which cut the
// run into pipelined segments (scratch/proto_pipe.py).
//
// The sweeps use  B = I/theta - c A  once more:  A^-1 B = (1/theta) A^-1 - c I, so on the block
// interior  T^-1 B_ii g = (1/theta) T^-1 g - c g  -- there is no right-hand side to build -- and
//     P1 = -(1/theta) T^-1 V ,   P2 = -(1/theta) T^-1 U
// (T^-1 b_left = -c V cancels the -c V of T^-1 B_ii V).  With the un-normalised factors
// l_c = A_c,prev / pivot_prev, kappa_c = 1 / (theta pivot_c) and z = y - c g:
//     elimination   h_c = g_c - l_c h_prev                          (1 FMA, the chain)
//                   t_c = h_c kappa_c - c g_c - c s_c g_next        (off the chain, written over g_c)
//     substitution  z_c = t_c - s_c z_next                          (1 FMA, the chain)
// which is 4 operations per cell for Crank-Nicolson (c = 1), 3 for implicit Euler (c = 0) and 5
// in general (scratch/proto_cn4.py).  The correction rides on the substitution links: P1, P2 are
// stored in the order the substitution visits the cells (slot j = left cell M-1-j, right cell M+j),
// the chain runs on the uncorrected values and each link is followed by its two correction FMAs.
//
// MODE 0: any theta in (0,1];  1: Crank-Nicolson;  2: implicit Euler.
template <int MODE, int IT>
__global__ void __launch_bounds__(kFusedThreads, 1)
fused_tmem_pipe_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
                       int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt) {
  constexpr int C = 32, N = 31, M = 16;
  constexpr int NCH = M / IT;
  using TC = TmemChunk<IT>;
  // shared-memory slots (per lane) of the constants of a step, in the place s occupied during set-up
  enum { K_OML = 0, K_OMR = 1, K_AL = 2, K_GA = 7, K_INVD = 12, K_SLO = 13, K_SUP = 14, K_SBLO = 15, K_SBDI = 16,
         K_SBUP = 17, K_BSEP = 18, K_V0 = 19, K_U0 = 20, K_VL = 21, K_UL = 22, K_BFIRST = 23 };
  extern __shared__ double smem[];
  __shared__ uint32_t tmem_base_slot;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int warps = blockDim.x >> 5;
  double* sS = smem + (size_t)warp * (3 * N * 32) + lane;        // s during set-up, then the constants; V; U
  double* sv = sS + N * 32;
  double* su = sv + N * 32;
  const int64_t J = pr.J;
  const double cth = (1.0 - theta) / theta;

  if (warp == 0) tmem_alloc_512(&tmem_base_slot);
  tmem_fence_before_sync();
  __syncthreads();
  tmem_fence_after_sync();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t t_lk = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 256;
  const uint32_t t_pp = t_lk + 128;                  // (P1, P2) pairs in substitution order

  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
    // ------------------------------------------------------------------ set-up (once)
    double delta;
    double sr[N];
    {
      Reduced R;
      double om_l, om_r;
      tmem_setup_common<true>(pr, p, lane, dt[p], theta, t_lk, sS, sv, su, R, sr, delta, om_l, om_r);
      // P1, P2: twisted solve with right-hand sides -V, -U, in place in tensor memory
      auto tcol = [&](int c) { return (uint32_t)(c < M ? 8 * c : 8 * (N - 1 - c) + 4); };
      auto tpp = [&](int c) { return (uint32_t)(c < M ? 8 * (M - 1 - c) : 8 * (c - M) + 4); };
      tmem_wait_st();
      double l, kap, h1 = -sv[0], h2 = -su[0];
      tmem_st_2d(t_pp + tpp(0), h1, h2);
#pragma unroll 1
      for (int c = 1; c < M; ++c) {                  // eliminate downwards
        tmem_ld_2d(t_lk + tcol(c), l, kap);
        h1 = fma(-l, h1, -sv[c * 32]); h2 = fma(-l, h2, -su[c * 32]);
        tmem_st_2d(t_pp + tpp(c), h1, h2);
      }
      const double hl1 = h1, hl2 = h2;               // eliminated values at cell M-1
      h1 = -sv[(N - 1) * 32]; h2 = -su[(N - 1) * 32];
      tmem_st_2d(t_pp + tpp(N - 1), h1, h2);
#pragma unroll 1
      for (int c = N - 2; c >= M; --c) {             // eliminate upwards
        tmem_ld_2d(t_lk + tcol(c), l, kap);
        h1 = fma(-l, h1, -sv[c * 32]); h2 = fma(-l, h2, -su[c * 32]);
        tmem_st_2d(t_pp + tpp(c), h1, h2);
      }
      double kap_l, kap_r;                           // meet between cells M-1 and M
      tmem_ld_2d(t_lk + tcol(M - 1), l, kap_l);
      tmem_ld_2d(t_lk + tcol(M), l, kap_r);
      const double m1 = fma(-sS[M * 32], hl1 * kap_l, h1 * kap_r) * delta;
      const double m2 = fma(-sS[M * 32], hl2 * kap_l, h2 * kap_r) * delta;
      tmem_wait_st();
      double p1 = m1, p2 = m2, x1, x2;
#pragma unroll 1
      for (int c = M - 1; c >= 0; --c) {             // substitute outwards
        tmem_ld_2d(t_pp + tpp(c), x1, x2);
        tmem_ld_2d(t_lk + tcol(c), l, kap);
        const double sc = sS[c * 32];
        p1 = fma(-sc, p1, x1 * kap); p2 = fma(-sc, p2, x2 * kap);
        tmem_st_2d(t_pp + tpp(c), p1, p2);
      }
      tmem_st_2d(t_pp + tpp(M), m1, m2);
      p1 = m1; p2 = m2;
#pragma unroll 1
      for (int c = M + 1; c < N; ++c) {
        tmem_ld_2d(t_pp + tpp(c), x1, x2);
        tmem_ld_2d(t_lk + tcol(c), l, kap);
        const double sc = sS[c * 32];
        p1 = fma(-sc, p1, x1 * kap); p2 = fma(-sc, p2, x2 * kap);
        tmem_st_2d(t_pp + tpp(c), p1, p2);
      }
      tmem_wait_st();
      __syncwarp();
      // s is in registers now: its shared-memory slots take the per-lane constants of a step, so
      // that the registers they would occupy can hold the second tensor-memory chunk buffer
      const double V0 = sv[0], U0 = su[0], VL = sv[(N - 1) * 32], UL = su[(N - 1) * 32];
      sS[K_OML * 32] = om_l; sS[K_OMR * 32] = om_r;
#pragma unroll
      for (int l2 = 0; l2 < 5; ++l2) { sS[(K_AL + l2) * 32] = R.al[l2]; sS[(K_GA + l2) * 32] = R.ga[l2]; }
      sS[K_INVD * 32] = R.invD; sS[K_SLO * 32] = R.slo; sS[K_SUP * 32] = R.sup; sS[K_SBLO * 32] = R.sblo;
      sS[K_SBDI * 32] = R.sbdi; sS[K_SBUP * 32] = R.sbup; sS[K_BSEP * 32] = R.bsep; sS[K_V0 * 32] = V0;
      sS[K_U0 * 32] = U0; sS[K_VL * 32] = VL; sS[K_UL * 32] = UL; sS[K_BFIRST * 32] = R.bfirst;
    }
    auto K = [&](int slot) { return sS[slot * 32]; };

    double w[C];
    load_state<C>(w_io + p * J, J, lane, w);

    // ------------------------------------------------------------------ time loop
    uint32_t ra[TC::REGS], rb[TC::REGS];
    // t_c of one cell (see the header): MODE 1 takes g, g_next as they are, MODE 0 their c-multiples
    auto tval = [&](double hk_minus, double s, double cg_next) {
      return (MODE == 2) ? hk_minus : fma(-s, cg_next, hk_minus);
    };
    // One step on w[0..N-1] in place; the cells outside the block that the end cells couple to
    // contribute oldL0 / oldR0 (first step of a segment only).  With PIPE the five PCR levels of
    // the previous step's separator system (rr in, xs out) are issued between the elimination
    // chunks, the step-end chain of the PREVIOUS step (xs -> shuffle -> true end values -> shuffle
    // -> rs) follows as soon as the last level is done (EARLY), and the substitution applies the
    // delayed correction link by link.
    auto sweeps = [&](auto pipe_tag, double oldL0, double oldR0, uint32_t next_region, double& rr,
                      double g0_old, double gl_old, double& xsl, double& rs) {
      constexpr bool PIPE = decltype(pipe_tag)::value;
      constexpr bool EARLY = PIPE && NCH >= 8;
      double x_first = 0.0, x_last = 0.0, x_first_next = 0.0;
      const double w_m_old = w[M];
      double hL = 0.0, hR = 0.0, hk_l = 0.0, hk_r = 0.0;
      double cgL = 0.0, cgR = 0.0;                    // MODE 0: c * g of the cell the chain is at
      int level = 0;
#pragma unroll
      for (int k = 0; k < NCH; ++k) {
        uint32_t* cur = (k & 1) ? rb : ra;
        // prefetch of the next chunk, pinned ahead of this chunk's arithmetic (chunk_ld_after)
        if (k == 0) { TC::wait(ra); chunk_ld_after<TC>(t_lk + 8 * IT, rb, w[0]); }
        else if (k & 1) { TC::wait(rb); chunk_ld_after<TC>(k + 1 < NCH ? t_lk + 8 * IT * (k + 1) : next_region, ra, hL); }
        else            { TC::wait(ra); chunk_ld_after<TC>(k + 1 < NCH ? t_lk + 8 * IT * (k + 1) : next_region, rb, hL); }
        // One PCR level per chunk.  Warps issue in order, so the level's shuffles are issued HERE and
        // their consumer sits after the chunk's cell work: the ~26-cycle shuffle latency is covered.
        double rm = 0.0, rp = 0.0, fa = 0.0, fg = 0.0;
        const bool lvl = PIPE && k < 5;
        if (lvl) { rm = shfl_up(rr, 1 << k); rp = shfl_dn(rr, 1 << k); fa = K(K_AL + k); fg = K(K_GA + k); }
        if (EARLY && k == 5) { rr *= K(K_INVD); xsl = shfl_up(rr, 1); }      // rr now holds xs
        if (EARLY && k == 6) {
          x_last = fma(-K(K_UL), rr, fma(-K(K_VL), xsl, gl_old));            // true values of the finished step
          x_first = fma(-K(K_U0), rr, fma(-K(K_V0), xsl, g0_old));
          x_first_next = shfl_dn(x_first, 1);
        }
        if (EARLY && k == 7) {
          if (lane == 31) x_first_next = 0.0;
          rs = fma(K(K_SBUP), x_first_next, fma(K(K_SBDI), rr, fma(K(K_SBLO), x_last, K(K_BSEP))));
        }
#pragma unroll
        for (int ii = 0; ii < IT; ++ii) {
          const int i = k * IT + ii;
          {
            const int c = i;                          // left cell, eliminated top-down
            const double l = chunk_dbl(cur, 4 * ii), kap = chunk_dbl(cur, 4 * ii + 1);
            const double gc = w[c];
            const double gn = (c == M - 1) ? w_m_old : w[c + 1];
            if (i == 0) {
              hL = gc + K(K_BFIRST);
              if (!PIPE) hL = fma(-K(K_OML), oldL0, hL);
              if (MODE == 0) cgL = cth * gc;
            } else {
              hL = fma(-l, hL, gc);
            }
            const double cg = (MODE == 0) ? cgL : gc, cgn = (MODE == 0) ? cth * gn : gn;
            if (c == M - 1) { hk_l = hL * kap; w[c] = tval(MODE == 2 ? hk_l : hk_l - cg, sr[c], cgn); }
            else w[c] = tval(MODE == 2 ? hL * kap : fma(hL, kap, -cg), sr[c], cgn);
            cgL = cgn;
          }
          if (i < N - M) {
            const int c = N - 1 - i;                  // right cell, eliminated bottom-up
            const double l = chunk_dbl(cur, 4 * ii + 2), kap = chunk_dbl(cur, 4 * ii + 3);
            const double gc = w[c];
            const double gp = w[c - 1];
            if (i == 0) {
              hR = gc;
              if (!PIPE) hR = fma(-K(K_OMR), oldR0, hR);
              if (MODE == 0) cgR = cth * gc;
            } else {
              hR = fma(-l, hR, gc);
            }
            const double cg = (MODE == 0) ? cgR : gc, cgp = (MODE == 0) ? cth * gp : gp;
            if (c == M) { hk_r = hR * kap; w[c] = tval(MODE == 2 ? hk_r : hk_r - cg, sr[c], cgp); }
            else w[c] = tval(MODE == 2 ? hR * kap : fma(hR, kap, -cg), sr[c], cgp);
            cgR = cgp;
          }
        }
        if (lvl) { rr = fma(fg, rp, fma(fa, rm, rr)); level = k + 1; }
      }
      if (PIPE) {
#pragma unroll
        for (int l = 0; l < 5; ++l) {
          if (l >= level) {
            const int s = 1 << l;
            const double rm = shfl_up(rr, s), rp = shfl_dn(rr, s);
            rr = fma(K(K_GA + l), rp, fma(K(K_AL + l), rm, rr));
          }
        }
        if (!EARLY) {
          rr *= K(K_INVD);                           // rr now holds xs
          xsl = shfl_up(rr, 1);
          x_last = fma(-K(K_UL), rr, fma(-K(K_VL), xsl, gl_old));
          x_first = fma(-K(K_U0), rr, fma(-K(K_V0), xsl, g0_old));
          x_first_next = shfl_dn(x_first, 1);
          if (lane == 31) x_first_next = 0.0;
          rs = fma(K(K_SBUP), x_first_next, fma(K(K_SBDI), rr, fma(K(K_SBLO), x_last, K(K_BSEP))));
        }
      }
      // meeting point: z_M = (hk_r - s_M hk_l) delta - c g_M
      double zM = fma(-sr[M], hk_l, hk_r) * delta;
      if (MODE == 1) zM -= w_m_old;
      if (MODE == 0) zM = fma(-cth, w_m_old, zM);
      if (PIPE) {
        // substitution outwards with the delayed correction riding on every link: the chain runs on
        // the uncorrected values zL / zR, the corrected ones go to w
        const double xs = rr;
        double zL = zM, zR = zM;
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
          uint32_t* cur = (k & 1) ? rb : ra;
          if (k & 1) { TC::wait(rb); chunk_ld_after<TC>(k + 1 < NCH ? t_pp + 8 * IT * (k + 1) : t_lk, ra, zL); }
          else       { TC::wait(ra); chunk_ld_after<TC>(k + 1 < NCH ? t_pp + 8 * IT * (k + 1) : t_lk, rb, zL); }
          // links first, then the xsl products of the whole chunk, then the xs products: consecutive
          // FMAs that share an operand can take it from the operand-reuse cache instead of the
          // register file, whose bandwidth is what bounds these loops (fp64_operands.cu)
          double zl[IT], zr[IT];
#pragma unroll
          for (int ii = 0; ii < IT; ++ii) {
            const int j = k * IT + ii;
            const int cl = M - 1 - j, cr = M + j;
            zL = fma(-sr[cl], zL, w[cl]);
            zl[ii] = zL;
            if (j < N - M) {
              if (j > 0) zR = fma(-sr[cr], zR, w[cr]);
              zr[ii] = zR;
            }
          }
#pragma unroll
          for (int ii = 0; ii < IT; ++ii) {
            const int j = k * IT + ii;
            zl[ii] = fma(chunk_dbl(cur, 4 * ii), xsl, zl[ii]);
            if (j < N - M) zr[ii] = fma(chunk_dbl(cur, 4 * ii + 2), xsl, zr[ii]);
          }
#pragma unroll
          for (int ii = 0; ii < IT; ++ii) {
            const int j = k * IT + ii;
            w[M - 1 - j] = fma(chunk_dbl(cur, 4 * ii + 1), xs, zl[ii]);
            if (j < N - M) w[M + j] = fma(chunk_dbl(cur, 4 * ii + 3), xs, zr[ii]);
          }
        }
      } else {
        w[M] = zM;
#pragma unroll
        for (int j = 0; j < M; ++j) {
          const int cl = M - 1 - j, cr = M + 1 + j;
          w[cl] = fma(-sr[cl], w[cl + 1], w[cl]);
          if (cr < N) w[cr] = fma(-sr[cr], w[cr - 1], w[cr]);
        }
      }
    };

    int64_t remaining = nsteps, n_ckpt = 0;
    TC::ld(t_lk, ra);                               // first (l, kappa) chunk
    while (remaining > 0) {
      const bool ck = ckpt != nullptr && ckpt_every > 0 && remaining >= ckpt_every;
      const int64_t L = ck ? ckpt_every : remaining; // steps until the true state is needed again
      // -------- first step of the segment: sweeps on the true state
      double wl = shfl_up(w[C - 1], 1);
      double wr = shfl_dn(w[0], 1);
      if (lane == 0) wl = 0.0;
      if (lane == 31) wr = 0.0;
      double rs = fma(K(K_SBUP), wr, fma(K(K_SBDI), w[C - 1], fma(K(K_SBLO), w[N - 1], K(K_BSEP))));
      double rr = 0.0, xsl = 0.0;
      sweeps(std::false_type{}, wl, w[C - 1], t_lk, rr, 0.0, 0.0, xsl, rs);
      rr = fma(-K(K_SUP), shfl_dn(w[0], 1), fma(-K(K_SLO), w[N - 1], rs));
      // -------- pipelined steps: PCR of step n overlaps the sweeps of step n+1
      for (int64_t i = 1; i < L; ++i) {
        sweeps(std::true_type{}, 0.0, 0.0, t_pp, rr, w[0], w[N - 1], xsl, rs);
        const double g0n = shfl_dn(w[0], 1);        // lane 31: sup == 0
        rr = fma(-K(K_SUP), g0n, fma(-K(K_SLO), w[N - 1], rs));
      }
      // -------- end of the segment: finish the last step and materialise the true state
      {
        double xs = rr;
#pragma unroll
        for (int l = 0; l < 5; ++l) {
          const double rm = shfl_up(xs, 1 << l), rp = shfl_dn(xs, 1 << l);
          xs = fma(K(K_GA + l), rp, fma(K(K_AL + l), rm, xs));
        }
        xs *= K(K_INVD);
        const double xsl_end = shfl_up(xs, 1);      // lane 0: V == 0, value unused
#pragma unroll
        for (int c = 0; c < N; ++c) w[c] = fma(-su[c * 32], xs, fma(-sv[c * 32], xsl_end, w[c]));
        w[C - 1] = xs;
      }
      remaining -= L;
      if (ck) {
        store_state<C>(ckpt + (n_ckpt * pr.P + p) * J, J, lane, w);
        ++n_ckpt;
      }
    }
    TC::wait(ra);                                   // drain the prefetch of the step that never comes
    store_state<C>(w_io + p * J, J, lane, w);
    __syncwarp();
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

template <int C, int MODE, bool RB, bool QREG>
static int launch_fused(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                        int64_t ckpt_every, double* w, double* ckpt, cudaStream_t stream) {
  constexpr int NS = (C > 1 ? C - 1 : 1);
  constexpr size_t per_warp = (size_t)(QREG ? 4 : 5) * NS * 32 * sizeof(double);
  int sms = 0, max_smem = 0;
  int st = device_limits(sms, max_smem);
  if (st) return st;
  int warps = (int)((size_t)max_smem / per_warp);
  if (warps > kFusedThreads / 32) warps = kFusedThreads / 32;
  if (warps < 1) return FVM_ERR_UNSUPPORTED;
  if ((int64_t)warps > pr.P) warps = (int)pr.P;
  const size_t smem = per_warp * warps;
  auto kern = fused_kernel<C, MODE, RB, QREG>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  int per_sm = 1;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, warps * 32, smem);
  if (e != cudaSuccess) return (int)e;
  if (per_sm < 1) per_sm = 1;
  int64_t blocks = (pr.P + warps - 1) / warps;
  const int64_t resident = (int64_t)sms * per_sm;
  if (blocks > resident) blocks = resident;      // persistent: each warp strides over problems
  kern<<<(unsigned)blocks, warps * 32, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt);
  return (int)cudaGetLastError();
}

template <int MODE, int IT, bool PIPE>
static int launch_fused_tmem(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                             int64_t ckpt_every, double* w, double* ckpt, cudaStream_t stream) {
  constexpr int warps = kFusedThreads / 32;
  constexpr size_t smem = (size_t)warps * 3 * 31 * 32 * sizeof(double);
  int sms = 0, max_smem = 0;
  int st = device_limits(sms, max_smem);
  if (st) return st;
  int64_t blocks = (pr.P + warps - 1) / warps;
  if (blocks > sms) blocks = sms;                // one CTA per SM: it holds the SM's whole TMEM
  int launch_warps = warps;
  const char* lw = getenv("FVM_FUSED_WARPS");    // experiment: 4 = one warp per scheduler
  if (lw && lw[0] == '4') launch_warps = 4;
  cudaError_t e;
  if constexpr (PIPE) {
    auto kern = fused_tmem_pipe_kernel<MODE, IT>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<(unsigned)blocks, launch_warps * 32, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt);
  } else {
    auto kern = fused_tmem_kernel<MODE, IT>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<(unsigned)blocks, launch_warps * 32, smem, stream>>>(pr, dt, theta, nsteps, ckpt_every, w, ckpt);
  }
  return (int)cudaGetLastError();
}

// FVM_FUSED_STORE=smem forces the shared-memory kernel (A/B comparison and profiling).
static bool tmem_enabled() {
  const char* s = getenv("FVM_FUSED_STORE");
  return !(s && s[0] == 's');
}

template <int C>
static int dispatch_mode(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                         int64_t ckpt_every, double* w, double* ckpt, cudaStream_t stream) {
  const bool rb = pr.right_kind == FVM_BC_ROBIN && pr.right_flux != nullptr &&
                  ((pr.J - 1) % C) != C - 1;
  constexpr bool Q = true;
  if (rb) return launch_fused<C, 0, true, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (C == 32 && tmem_enabled()) {
    // FVM_FUSED_PIPE=0 selects the kernel without the delayed correction (A/B runs and profiling)
    const char* pp = getenv("FVM_FUSED_PIPE");
    if (pp && pp[0] == '0') {
      if (theta == 0.5) return launch_fused_tmem<1, 4, false>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
      if (theta == 1.0) return launch_fused_tmem<2, 4, false>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
      return launch_fused_tmem<0, 4, false>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    }
    const char* its = getenv("FVM_TMEM_IT");     // 1: .x8 chunks (experiments); default 2: .x16
    if (its && its[0] == '1') {
      if (theta == 0.5) return launch_fused_tmem<1, 1, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
      if (theta == 1.0) return launch_fused_tmem<2, 1, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
      return launch_fused_tmem<0, 1, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    }
    if (theta == 0.5) return launch_fused_tmem<1, 2, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    if (theta == 1.0) return launch_fused_tmem<2, 2, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
    return launch_fused_tmem<0, 2, true>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  }
  if (theta == 0.5) return launch_fused<C, 1, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (theta == 1.0) return launch_fused<C, 2, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  return launch_fused<C, 0, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
}

// fused_cta.cu
int run_fused_cta(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps, int64_t ckpt_every,
                  double* w, double* ckpt, cudaStream_t stream);

}  // namespace fvm

extern "C" int64_t fvm_fused_max_cells(void) { return 8192; }

extern "C" int fvm_theta_run_fused(const fvm_problem* prob, const double* dt, double theta,
                                   int64_t nsteps, int64_t checkpoint_every, double* w,
                                   double* ckpt, void* stream) {
  if (!prob || !prob->faces || !prob->a || !prob->d || !dt || !w) return FVM_ERR_INVALID_ARG;
  if (prob->J < 3 || prob->P < 1 || nsteps < 0 || checkpoint_every < 0) return FVM_ERR_INVALID_ARG;
  if (!(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (prob->scheme < FVM_CENTRAL || prob->scheme > FVM_EXPONENTIAL) return FVM_ERR_INVALID_ARG;
  if ((prob->left_kind | prob->right_kind | prob->alpha_left | prob->alpha_right) & ~1)
    return FVM_ERR_INVALID_ARG;
  if (prob->J > fvm_fused_max_cells()) return FVM_ERR_UNSUPPORTED;
  if (nsteps == 0) return FVM_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int64_t J = prob->J;
  using namespace fvm;
  if (J <= 32) return dispatch_mode<1>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  if (J <= 64) return dispatch_mode<2>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  if (J <= 128) return dispatch_mode<4>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  if (J <= 256) return dispatch_mode<8>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  if (J <= 512) return dispatch_mode<16>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  if (J <= 32 * kFusedMaxC) return dispatch_mode<32>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
  return run_fused_cta(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
}
