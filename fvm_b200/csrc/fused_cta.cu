// Variant 2 for 1024 < J <= 8192: one CTA per system, everything on chip between checkpoints.
//
// The warp-per-system kernels (fused.cu) stop at 1024 cells (32 lanes x 32 cells of registers).
// Here a CTA of TPB threads owns one system; thread t keeps its CPT consecutive cells of the state
// in registers for the whole launch, the three factored coefficient arrays (p, q, b: global LU of
// A, as in variant 1) live in shared memory in tiled order ([c][t]: conflict-free), and each step
// is  r~ = inv*(B w)  followed by the two Thomas sweeps as CTA-wide affine-map scans
// (cta_scan.cuh).  Neighbouring threads exchange one edge value per side through shared memory.
// Assembly and the (sequential) factorisation run inside the launch.  HBM traffic: J doubles in,
// J out, J per checkpoint.
#include "fvm_device.cuh"
#include "cta_scan.cuh"

namespace fvm {

template <int TPB, int CPT, int MODE>
__global__ void __launch_bounds__(TPB, 1)
fused_cta_kernel(fvm_problem pr, const double* __restrict__ dt, double theta, int64_t nsteps,
                 int64_t ckpt_every, double* __restrict__ w_io, double* __restrict__ ckpt) {
  constexpr int T = TPB * CPT;
  extern __shared__ double smem[];
  double* sp = smem;                 // [c*TPB + t]
  double* sq = sp + T;
  double* sb = sq + T;
  double* sfirst = sb + T;           // [TPB] first cell of every thread
  double* slast = sfirst + TPB;      // [TPB] last cell of every thread
  double* sscan = slast + TPB;       // [66] scratch of the scans
  double* sbeta = sscan + 66;        // [2] beta[0]*inv_0, beta[J-1]*inv_{J-1}
  const int tid = threadIdx.x;
  const int64_t J = pr.J;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
  auto at = [&](int r) { return (r % CPT) * TPB + r / CPT; };

  for (int64_t p = blockIdx.x; p < pr.P; p += gridDim.x) {
    // ---------------------------------------------------------------- set-up: rows of A
    {
      const double dt_theta = mul(dt[p], theta);
      const int64_t j0 = (int64_t)tid * CPT;
      double D0 = cell_D(pr, p, j0);
      Face FL = problem_face(pr, p, j0 - 1, cell_D(pr, p, j0 - 1), D0);
#pragma unroll 1
      for (int c = 0; c < CPT; ++c) {
        const int64_t j = j0 + c;
        const double Dp = cell_D(pr, p, j + 1);
        const Face FR = problem_face(pr, p, j, D0, Dp);
        Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;          // padding rows (j >= J) are identity
        if (j < J) {
          Row m = problem_row(pr, p, j, FL, FR);
          a = theta_lhs_row(m, dt_theta, alpha_of(j, J, pr.alpha_left, pr.alpha_right));
        }
        D0 = Dp; FL = FR;
        const int o = c * TPB + tid;
        sp[o] = a.lo; sb[o] = a.di; sq[o] = a.up;
      }
    }
    __syncthreads();
    // ---------------------------------------------------------------- factorisation (sequential)
    if (tid == 0) {
      double beta_l = 0.0, beta_r = 0.0;
      if (pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
        beta_l = dvd(pr.left_flux[p], sub(pr.faces[1], pr.faces[0]));
      if (pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
        beta_r = dvd(-pr.right_flux[p], sub(pr.faces[J], pr.faces[J - 1]));
      double q_prev = 0.0;
      for (int r = 0; r < T; ++r) {
        const int o = at(r);
        const double lo = sp[o], di = sb[o], up = sq[o];
        const double inv = 1.0 / (di - lo * q_prev);
        q_prev = up * inv;
        sp[o] = lo * inv; sq[o] = q_prev; sb[o] = inv * (rth - cth * di);
        if (r == 0) sbeta[0] = beta_l * inv;
        if (r == J - 1) sbeta[1] = beta_r * inv;
      }
    }
    __syncthreads();
    const double bfirst = sbeta[0], blast = sbeta[1];
    const int64_t jl0 = (int64_t)tid * CPT;
    const int c_last = (J - 1 >= jl0 && J - 1 < jl0 + CPT) ? (int)(J - 1 - jl0) : -1;

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
      sfirst[tid] = w[0];
      slast[tid] = w[CPT - 1];
      __syncthreads();
      double wm = (tid > 0) ? slast[tid - 1] : 0.0;          // left neighbour of this thread's first cell
      const double wr = (tid < TPB - 1) ? sfirst[tid + 1] : 0.0;
      double g[CPT];
#pragma unroll
      for (int c = 0; c < CPT; ++c) {
        const int o = c * TPB + tid;
        const double wc = w[c];
        const double wp = (c + 1 < CPT) ? w[c + 1] : wr;
        double r;
        if (MODE == 2) {
          r = sb[o] * wc;
        } else {
          const double t = fma(sq[o], wp, sp[o] * wm);
          r = (MODE == 1) ? fma(sb[o], wc, -t) : fma(-cth, t, sb[o] * wc);
        }
        if (c == c_last) r += blast;
        g[c] = r;
        wm = wc;
      }
      if (tid == 0) g[0] += bfirst;
      cta_forward<TPB, CPT>(g, [&](int c) { return sp[c * TPB + tid]; }, sscan, 0.0, nullptr);
      cta_backward<TPB, CPT>(g, [&](int c) { return sq[c * TPB + tid]; }, sscan, 0.0, nullptr);
#pragma unroll
      for (int c = 0; c < CPT; ++c) w[c] = g[c];

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
  constexpr size_t smem = (size_t)(3 * TPB * CPT + 2 * TPB + 66 + 2) * sizeof(double);
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
  if (pr.J <= 2048) return dispatch_cta<256, 8>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  if (pr.J <= 4096) return dispatch_cta<512, 8>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
  return dispatch_cta<512, 16>(pr, dt, theta, nsteps, ckpt_every, w, ckpt, stream);
}

}  // namespace fvm
