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
// State w[C] lives in registers for the whole launch; the five per-cell coefficients
// (p, q, b, v, u) live in shared memory as [coef][c][lane] (lane-consecutive doubles:
// conflict-free).  HBM traffic: J doubles in, J doubles out, J per checkpoint.
//
// Assembly (fvm_device.cuh) and factorisation run inside the same launch, one lane per
// owned cell, so a launch needs only the mesh, the per-problem scalars and the state.
#include "fvm_device.cuh"

namespace fvm {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kFusedMaxC = 32;
constexpr int kFusedThreads = 256;   // upper bound used for register allocation

__device__ __forceinline__ double shfl_up(double x, int s) { return __shfl_up_sync(kFull, x, s); }
__device__ __forceinline__ double shfl_dn(double x, int s) { return __shfl_down_sync(kFull, x, s); }

// MODE 0: any theta in (0,1];  1: Crank-Nicolson (c = (1-theta)/theta = 1);  2: implicit (c = 0).
// RB: beta[J-1] is nonzero AND cell J-1 is an interior cell of its lane (slow path).
// QREG: q (the one coefficient a step reads twice) is held in registers instead of shared
//       memory: 4 instead of 6 shared-memory reads per cell-update, and 4/5 of the footprint.
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
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;

  for (int64_t p = (int64_t)blockIdx.x * warps + warp; p < pr.P; p += (int64_t)gridDim.x * warps) {
    // ------------------------------------------------------------------ setup (once)
    const double dt_theta = mul(dt[p], theta);
    double beta_l = 0.0, beta_r = 0.0;
    if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
      beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
    if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
      beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
    double slo = 0.0, sdi = 1.0, sup = 0.0;         // separator's row of A
    double bfirst = 0.0, bint = 0.0, bsep = 0.0;    // beta entries as they enter the sweeps
    int rb_c = -1;
    {
      double q_prev = 0.0, gv_prev = 0.0;
#pragma unroll 1
      for (int c = 0; c < C; ++c) {
        const int64_t j = (int64_t)lane * C + c;
        Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;  // padding rows (j >= J) are identity
        if (j < J) {
          double kap, Dc;
          Row m = assemble_row(pr, p, j, kap, Dc);
          a = theta_lhs_row(m, dt_theta, alpha_of(j, J, pr.alpha_left, pr.alpha_right));
        }
        if (c < N) {
          const double inv = 1.0 / (a.di - a.lo * q_prev);
          const double pc = a.lo * inv;
          q_prev = a.up * inv;
          gv_prev = (c == 0) ? pc : -pc * gv_prev;  // forward sweep of the left spike
          sp[c * 32] = pc;
          sq[c * 32] = q_prev;
          sb[c * 32] = inv * (rth - cth * a.di);
          sv[c * 32] = gv_prev;
          if (j == 0) bfirst = beta_l * inv;
          if (RB && j == J - 1) bint = beta_r * inv;
        } else {
          slo = a.lo; sdi = a.di; sup = a.up;
          if (j == 0) bsep += beta_l;
          if (j == J - 1) bsep += beta_r;
        }
      }
      if (RB) rb_c = (int)((J - 1) % C);
    }
    double qr[NS];                                  // QREG: q for the whole launch
    if (QREG) {
#pragma unroll
      for (int c = 0; c < NS; ++c) qr[c] = (c < N) ? sq[c * 32] : 0.0;
    }
    double RL = slo, RD = sdi, RU = sup;            // reduced (separator) system
    if (N > 0) {
      double vn = sv[(N - 1) * 32];                 // backward sweeps of both spikes
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
      RL = -slo * vlast;
      RD = sdi - slo * ulast - sup * v0n;           // sup == 0 on lane 31
      RU = -sup * u0n;
    }
    double al[5], ga[5];                            // PCR elimination factors per level
#pragma unroll
    for (int l = 0; l < 5; ++l) {
      const int s = 1 << l;
      const bool hm = lane >= s, hp = lane + s < 32;
      const double Dm = shfl_up(RD, s), Dp = shfl_dn(RD, s);
      const double Lm = shfl_up(RL, s), Um = shfl_up(RU, s);
      const double Lp = shfl_dn(RL, s), Up = shfl_dn(RU, s);
      const double a_ = hm ? -RL / Dm : 0.0;
      const double g_ = hp ? -RU / Dp : 0.0;
      al[l] = a_; ga[l] = g_;
      RD = RD + (hm ? a_ * Um : 0.0) + (hp ? g_ * Lp : 0.0);
      RL = hm ? a_ * Lm : 0.0;
      RU = hp ? g_ * Up : 0.0;
    }
    const double invD = 1.0 / RD;
    const double sblo = -cth * slo, sbdi = rth - cth * sdi, sbup = -cth * sup;
    __syncwarp();

    // ------------------------------------------------------------------ state in
    double w[C];
    {
      const double* src = w_io + p * J;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int64_t j = (int64_t)lane * C + c;
        w[c] = (j < J) ? src[j] : 0.0;
      }
    }

    // ------------------------------------------------------------------ time loop
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
          const double pc = sp[c * 32], bc = sb[c * 32];
          const double wc = w[c];
          double r;
          if (MODE == 2) {
            r = bc * wc;
          } else {
            const double t = fma(QREG ? qr[c] : sq[c * 32], w[c + 1], pc * wm);
            r = (MODE == 1) ? fma(bc, wc, -t) : fma(-cth, t, bc * wc);
          }
          if (c == 0) r += bfirst;
          if (RB) { if (c == rb_c) r += bint; }
          g = (c == 0) ? r : fma(-pc, g, r);
          w[c] = g;
          wm = wc;
        }
        const double rs = fma(sbup, wr, fma(sbdi, w[C - 1], fma(sblo, wm, bsep)));
#pragma unroll
        for (int c = N - 2; c >= 0; --c) w[c] = fma(-(QREG ? qr[c] : sq[c * 32]), w[c + 1], w[c]);
        const double g0n = shfl_dn(w[0], 1);
        rr = fma(-sup, g0n, fma(-slo, w[N - 1], rs));
      } else {
        rr = fma(sbup, wr, fma(sbdi, w[0], fma(sblo, wl, bsep)));
      }
#pragma unroll
      for (int l = 0; l < 5; ++l) {
        const int s = 1 << l;
        const double rm = shfl_up(rr, s), rp = shfl_dn(rr, s);
        rr = fma(ga[l], rp, fma(al[l], rm, rr));
      }
      const double xs = rr * invD;
      if (N > 0) {
        const double xsl = shfl_up(xs, 1);          // lane 0: v == 0, value unused
#pragma unroll
        for (int c = 0; c < N; ++c) w[c] = fma(-su[c * 32], xs, fma(-sv[c * 32], xsl, w[c]));
      }
      w[C - 1] = xs;

      if (ckpt != nullptr && ckpt_every > 0 && --until_ckpt == 0) {
        until_ckpt = ckpt_every;
        double* dst = ckpt + (n_ckpt * pr.P + p) * J;
        ++n_ckpt;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const int64_t j = (int64_t)lane * C + c;
          if (j < J) dst[j] = w[c];
        }
      }
    }

    // ------------------------------------------------------------------ state out
    {
      double* dst = w_io + p * J;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const int64_t j = (int64_t)lane * C + c;
        if (j < J) dst[j] = w[c];
      }
    }
    __syncwarp();
  }
}

template <int C, int MODE, bool RB, bool QREG>
static int launch_fused(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                        int64_t ckpt_every, double* w, double* ckpt, cudaStream_t stream) {
  constexpr int NS = (C > 1 ? C - 1 : 1);
  constexpr size_t per_warp = (size_t)(QREG ? 4 : 5) * NS * 32 * sizeof(double);
  int dev = 0, sms = 0, max_smem = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  int warps = (int)((size_t)max_smem / per_warp);
  if (warps > kFusedThreads / 32) warps = kFusedThreads / 32;
  if (warps < 1) return FVM_ERR_UNSUPPORTED;
  if ((int64_t)warps > pr.P) warps = (int)pr.P;
  const size_t smem = per_warp * warps;
  auto kern = fused_kernel<C, MODE, RB, QREG>;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
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

template <int C>
static int dispatch_mode(const fvm_problem& pr, const double* dt, double theta, int64_t nsteps,
                         int64_t ckpt_every, double* w, double* ckpt, cudaStream_t stream) {
  const bool rb = pr.right_kind == FVM_BC_ROBIN && pr.right_flux != nullptr &&
                  ((pr.J - 1) % C) != C - 1;
  constexpr bool Q = true;
  if (rb) return launch_fused<C, 0, true, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (theta == 0.5) return launch_fused<C, 1, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (theta == 1.0) return launch_fused<C, 2, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  return launch_fused<C, 0, false, Q>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
}

}  // namespace fvm

extern "C" int64_t fvm_fused_max_cells(void) { return 32 * fvm::kFusedMaxC; }

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
  return dispatch_mode<32>(*prob, dt, theta, nsteps, checkpoint_every, w, ckpt, s);
}
