// Variant 3: partitioned SPIKE / Thomas for LONG systems, streaming from HBM.
//
// A system of J cells is cut into tiles of T = TPB*CPT consecutive cells; the LAST cell of
// every tile is a separator.  With the separators known the tile interiors decouple:
//
//   set-up (once per model)
//     spike_assemble_kernel : rows of A = I - dt*theta*alpha*M for every local cell
//     spike_tile_factor_kernel : per tile, LU of the interior block (p, q, b as in the other
//                             variants, factorisation restarted at the tile's first cell) and
//                             the tile spikes Vt = T^-1(lo_first e_first), Wt = T^-1(up_last e_last)
//     (host)                  reduced tridiagonal over ALL tile separators of ALL ranks,
//     reduced_factor_kernel   factored once
//   every step
//     tile_solve_kernel     : one CTA per tile: r~ = inv*(B w); tile-local Thomas, parallelised
//                             over the CTA by a scan of affine maps g_out = A*g_in + B (each
//                             thread owns CPT consecutive cells: one sequential pass for its
//                             (A, B), a shuffle/shared-memory scan, one pass with the true carry);
//                             same backwards.  Reads w, p, q, b and writes g: 40 B per cell.
//     reduced_rhs_kernel    : rr_k = rs_k - lo_k g_last(k) - up_k g_first(k+1)
//     (host, multi-GPU)       all-gather of rr over the ranks (NCCL), nT doubles per rank
//     reduced_solve_kernel  : Thomas on the factored separator system, one CTA, affine scans
//     correct_kernel        : x = g - Vt xs(k-1) - Wt xs(k), only over the extent where a spike
//                             exceeds `spike_tol` (spikes decay away from the tile edges).  In a
//                             time loop this pass is skipped: the next step's tile_solve_kernel
//                             applies the correction while it loads the tile (LAZY), and only the
//                             separators and ghosts are written between steps.
//   multi-GPU, rank level   : rank_pack_kernel / rank_top_kernel around ONE all-gather per step
//
// All indices are LOCAL to the rank's slab unless they say "global"; the slab starts at global
// cell j_offset = tile_offset*T.  State arrays carry one ghost cell on each side
// (w[-1], w[n_local]) that holds the neighbouring rank's edge value.
//
// LAYOUT.  Thread t of a tile's CTA owns the CPT consecutive cells [t*CPT, (t+1)*CPT) (the
// recurrences need a blocked ownership), but every per-cell array of the library (p, q, b, Vt,
// Wt and the state w) stores cell r = t*CPT + c of a tile at  c*TPB + t  ("tiled" order), so that
// a warp's load of its c-th cells is two full 128-byte lines.  With the natural order the same
// loads touch 32 lines each and the kernel is bound by L1 tag look-ups at 30 % of HBM speed
// (profiles/r01_spike_launches_natural.csv).  The first and last positions of a tile are its
// first and last cell in both orders, so ghosts and tile neighbours need no special casing.
#include "fvm_device.cuh"
#include "cta_scan.cuh"

namespace fvm {

// Position of cell r (0..T-1) of a tile in the tiled order.
__device__ __forceinline__ int tiled_pos(int r, int TPB, int CPT) { return (r % CPT) * TPB + r / CPT; }

// Per-tile constants, SEP_STRIDE doubles per tile.
enum { SEP_LO = 0, SEP_UP, SEP_BLO, SEP_BDI, SEP_BUP, SEP_BETA, SEP_RL, SEP_RDA, SEP_V0, SEP_W0, SEP_STRIDE };

// ------------------------------------------------------------------------------------------ set-up
// Rows of A for the local cells; lo -> p, di -> b, up -> q (transformed in place by the next kernel).
__global__ void __launch_bounds__(256) spike_assemble_kernel(fvm_problem pr, const double* __restrict__ dt,
                                                             double theta, int64_t j_offset, int64_t n_local,
                                                             int TPB, int CPT, double* __restrict__ p,
                                                             double* __restrict__ q, double* __restrict__ b) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (i >= n_local) return;
  const int64_t j = j_offset + i;
  Row a; a.lo = 0.0; a.di = 1.0; a.up = 0.0;             // cells past the end of the mesh: identity
  if (j < pr.J) {
    double kap, D;
    Row m = assemble_row(pr, pi, j, kap, D);
    a = theta_lhs_row(m, mul(dt[pi], theta), alpha_of(j, pr.J, pr.alpha_left, pr.alpha_right));
  }
  const int T = TPB * CPT;
  const int64_t o = pi * n_local + (i / T) * T + tiled_pos((int)(i % T), TPB, CPT);
  p[o] = a.lo; b[o] = a.di; q[o] = a.up;
}

// One thread per tile: LU of the interior block restarted at the tile's first cell, the tile
// spikes, the separator's rows and the tile's contribution to the reduced system.
__global__ void __launch_bounds__(128) spike_tile_factor_kernel(
    fvm_problem pr, double theta, int64_t tile_offset, int64_t n_tiles, int TPB, int CPT, double* __restrict__ p,
    double* __restrict__ q, double* __restrict__ b, double* __restrict__ Vt, double* __restrict__ Wt,
    double* __restrict__ sep, double* __restrict__ bends, int* __restrict__ extents, double spike_tol) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (k >= n_tiles) return;
  const double rth = 1.0 / theta, cth = (1.0 - theta) / theta;
  const int T = TPB * CPT;
  const int64_t base = (pi * n_tiles + k) * T;
  const int64_t jg0 = (tile_offset + k) * T;               // global index of the tile's first cell
  auto at = [&](int r) { return base + tiled_pos(r, TPB, CPT); };
  // beta[0] / beta[J-1]: only the tile that HOLDS the end cell may touch the end faces.  With the
  // mesh split over ranks `faces` is the caller's pointer shifted by the slab offset, and a rank
  // that does not own cell 0 (or J-1) has no memory behind faces[0..1] (faces[J-1..J]).
  double beta_l = 0.0, beta_r = 0.0;
  if (jg0 == 0 && pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
    beta_l = dvd(pr.left_flux[pi], sub(pr.faces[1], pr.faces[0]));
  if (pr.J - 1 >= jg0 && pr.J - 1 < jg0 + T && pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
    beta_r = dvd(-pr.right_flux[pi], sub(pr.faces[pr.J], pr.faces[pr.J - 1]));
  const int N = T - 1;                                     // interior cells
  double q_prev = 0.0, gv = 0.0;
  for (int c = 0; c < N; ++c) {
    const int64_t o = at(c);
    const double lo = p[o], di = b[o], up = q[o];
    const double inv = 1.0 / (di - lo * q_prev);
    const double pc = lo * inv;
    q_prev = up * inv;
    gv = (c == 0) ? pc : -pc * gv;                         // forward sweep of the left spike
    p[o] = pc; q[o] = q_prev; b[o] = inv * (rth - cth * di);
    Vt[o] = gv;
    if (jg0 + c == 0) bends[2 * pi] = beta_l * inv;        // beta[0] as it enters the sweep
    if (jg0 + c == pr.J - 1) bends[2 * pi + 1] = beta_r * inv;   // beta[J-1] when cell J-1 is a tile-interior cell
  }
  double* s = sep + (pi * n_tiles + k) * SEP_STRIDE;
  {
    const int64_t oN = at(N);
    const double lo = p[oN], di = b[oN], up = q[oN];
    s[SEP_LO] = lo; s[SEP_UP] = up;
    s[SEP_BLO] = -cth * lo; s[SEP_BDI] = rth - cth * di; s[SEP_BUP] = -cth * up;
    double bs = 0.0;
    if (jg0 + N == 0) bs += beta_l;
    if (jg0 + N == pr.J - 1) bs += beta_r;
    s[SEP_BETA] = bs;
    p[oN] = 0.0; q[oN] = 0.0; b[oN] = 0.0;                 // separator: not part of the tile solve
    Vt[oN] = 0.0; Wt[oN] = 0.0;
    // backward sweeps of both spikes
    double vn = Vt[at(N - 1)], un = q[at(N - 1)];
    Wt[at(N - 1)] = un;
    const double vlast = vn, ulast = un;
    int eL = (fabs(vn) > spike_tol) ? N : 0, eR = 1;
    for (int c = N - 2; c >= 0; --c) {
      const int64_t o = at(c);
      const double qc = q[o];
      vn = Vt[o] - qc * vn;
      un = -qc * un;
      Vt[o] = vn; Wt[o] = un;
      if (fabs(vn) > spike_tol && c + 1 > eL) eL = c + 1;  // cells [0, eL) need the left correction
      if (fabs(un) > spike_tol) eR = N - c;                // cells [N-eR, N) need the right correction
    }
    s[SEP_RL] = -lo * vlast;
    s[SEP_RDA] = di - lo * ulast;                          // the -up*V0(k+1) term is added by the host
    s[SEP_V0] = vn; s[SEP_W0] = un;
    atomicMax(&extents[0], eL);
    atomicMax(&extents[1], eR);
  }
}

// Thomas factorisation of the reduced tridiagonal (one thread per problem; n = global tiles).
__global__ void reduced_factor_kernel(int64_t n, int64_t P, const double* __restrict__ RL,
                                      const double* __restrict__ RD, const double* __restrict__ RU,
                                      double* __restrict__ rp, double* __restrict__ rq, double* __restrict__ rinv) {
  const int64_t pi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pi >= P) return;
  double q_prev = 0.0;
  for (int64_t k = 0; k < n; ++k) {
    const int64_t o = pi * n + k;
    const double inv = 1.0 / (RD[o] - RL[o] * q_prev);
    rp[o] = RL[o] * inv;
    q_prev = RU[o] * inv;
    rq[o] = q_prev;
    rinv[o] = inv;
  }
}

// ------------------------------------------------------------------------------------------ per step
// LAZY correction: w_in holds the previous step's UNCORRECTED tile solutions g (true values only at
// the separators and in the ghosts); the correction x = g - Vt xs(k-1) - Wt xs(k) is applied while
// the tile is loaded, by the threads whose cells lie inside the spike extents (the first tl and the
// last tr threads of the CTA).  This replaces correct_kernel's separate read-modify-write pass:
// Vt, Wt are still read (16 B per corrected cell) but g is not re-read and x is never written.
struct LazyCorrection {
  const double* Vt; const double* Wt; const double* xs;   // xs indexed by tile_offset + k, per problem n_tiles_global
  int64_t tile_offset, n_tiles_global;
  int tl, tr;
};

template <int TPB, int CPT, int MODE, bool LAZY>
__global__ void __launch_bounds__(TPB, (TPB == 256 ? 2 : 1)) tile_solve_kernel(
    int64_t n_tiles, const double* __restrict__ p, const double* __restrict__ q,
    const double* __restrict__ b, const double* __restrict__ sep, const double* __restrict__ bends,
    int has_first_cell, int64_t jl_last, double cth, const double* __restrict__ w_in,
    double* __restrict__ w_out, double* __restrict__ rs, int64_t w_stride, LazyCorrection lz) {
  constexpr int T = TPB * CPT;
  __shared__ double sm[66];
  __shared__ double sedge[LAZY ? 2 * TPB : 1];
  const int64_t k = blockIdx.x, pi = blockIdx.y;
  const int tid = threadIdx.x;
  const int64_t base = (pi * n_tiles + k) * T + tid;                 // into p, q, b (tiled: + c*TPB)
  const int64_t wbase = pi * w_stride + k * T;                       // tile's first position in w (ghosted)
  double w[CPT + 2], pc[CPT], qc[CPT], g[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    pc[c] = p[base + c * TPB]; qc[c] = q[base + c * TPB]; g[c] = b[base + c * TPB];
    w[c + 1] = w_in[wbase + c * TPB + tid];
  }
  if (!LAZY) {
    // left neighbour of this thread's first cell = last cell of thread tid-1 (or the position just
    // before the tile: previous tile's separator / the left ghost); right neighbour likewise
    w[0] = w_in[tid == 0 ? wbase - 1 : wbase + (CPT - 1) * TPB + tid - 1];
    w[CPT + 1] = w_in[tid == TPB - 1 ? wbase + T : wbase + tid + 1];
  } else {
    const int64_t kg = lz.tile_offset + k;
    const double* X_ = lz.xs + pi * lz.n_tiles_global;
    const double xk = X_[kg];
    const double xl = kg > 0 ? X_[kg - 1] : 0.0;
    if (tid < lz.tl || tid >= TPB - lz.tr) {               // (the separator's Vt, Wt are 0: it stays xs)
#pragma unroll
      for (int c = 0; c < CPT; ++c)
        w[c + 1] = fma(-lz.Wt[base + c * TPB], xk, fma(-lz.Vt[base + c * TPB], xl, w[c + 1]));
    }
    sedge[tid] = w[1];
    sedge[TPB + tid] = w[CPT];
    __syncthreads();
    w[0] = tid == 0 ? w_in[wbase - 1] : sedge[TPB + tid - 1];   // before the tile: a separator or the ghost, both true
    if (tid < TPB - 1) {
      w[CPT + 1] = sedge[tid + 1];
    } else if (k + 1 < n_tiles) {                          // next tile's first cell, corrected with ITS separators
      const int64_t on = (pi * n_tiles + k + 1) * T;
      w[CPT + 1] = fma(-lz.Wt[on], X_[kg + 1], fma(-lz.Vt[on], xk, w_in[wbase + T]));
    } else {
      w[CPT + 1] = w_in[wbase + T];                        // right ghost: already a true value
    }
  }
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    if (MODE == 2) {
      g[c] = g[c] * w[c + 1];
    } else {
      const double t = fma(qc[c], w[c + 2], pc[c] * w[c]);
      g[c] = (MODE == 1) ? fma(g[c], w[c + 1], -t) : fma(-cth, t, g[c] * w[c + 1]);
    }
  }
  if (has_first_cell && k == 0 && threadIdx.x == 0) g[0] += bends[2 * pi];
  {                                                        // beta[J-1] at a tile-interior cell (rare)
    const int64_t l0 = k * T + (int64_t)tid * CPT;
    if (jl_last >= l0 && jl_last < l0 + CPT) {
      const int cl = (int)(jl_last - l0);
#pragma unroll
      for (int c = 0; c < CPT; ++c) if (c == cl) g[c] += bends[2 * pi + 1];
    }
  }
  if (threadIdx.x == TPB - 1) {                            // the separator's right-hand side
    const double* s = sep + (pi * n_tiles + k) * SEP_STRIDE;
    rs[pi * n_tiles + k] = fma(s[SEP_BUP], w[CPT + 1], fma(s[SEP_BDI], w[CPT], fma(s[SEP_BLO], w[CPT - 1], s[SEP_BETA])));
  }
  cta_forward<TPB, CPT>(g, [&](int c) { return pc[c]; }, sm, 0.0, nullptr);
  cta_backward<TPB, CPT>(g, [&](int c) { return qc[c]; }, sm, 0.0, nullptr);
#pragma unroll
  for (int c = 0; c < CPT; ++c) w_out[wbase + c * TPB + tid] = g[c];
}

// rr_k = rs_k - lo_k g_last(k) - up_k g_first(k+1); the last local tile leaves the g_first term
// to whoever knows the next rank's first cell (g_first_out[pi] = this rank's own first value).
__global__ void __launch_bounds__(256) reduced_rhs_kernel(int64_t n_tiles, int TPB, int CPT, const double* __restrict__ sep,
                                                          const double* __restrict__ rs, const double* __restrict__ g,
                                                          int64_t w_stride, double* __restrict__ rr,
                                                          int64_t rr_stride, double* __restrict__ g_first_out) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (k >= n_tiles) return;
  const int T = TPB * CPT;
  const double* s = sep + (pi * n_tiles + k) * SEP_STRIDE;
  const double* gp = g + pi * w_stride + k * T;
  double r = fma(-s[SEP_LO], gp[tiled_pos(T - 2, TPB, CPT)], rs[pi * n_tiles + k]);
  if (k + 1 < n_tiles) r = fma(-s[SEP_UP], gp[T], r);
  rr[pi * rr_stride + k] = r;
  if (k == 0) g_first_out[pi] = gp[0];
}

// Thomas solve of the factored reduced system by ONE CTA per problem, in chunks of TPB*CPT
// unknowns with the carry handed from chunk to chunk.
template <int TPB, int CPT>
__global__ void __launch_bounds__(TPB) reduced_solve_kernel(int64_t n, const double* __restrict__ rp,
                                                            const double* __restrict__ rq,
                                                            const double* __restrict__ rinv,
                                                            const double* __restrict__ rr, double* __restrict__ xs) {
  constexpr int CH = TPB * CPT;
  __shared__ double sm[66];
  const int64_t pi = blockIdx.x;
  const double* P_ = rp + pi * n; const double* Q_ = rq + pi * n; const double* I_ = rinv + pi * n;
  const double* R_ = rr + pi * n; double* X_ = xs + pi * n;
  double carry = 0.0;
  for (int64_t c0 = 0; c0 < n; c0 += CH) {                 // forward: g_k = rr_k*inv_k - p_k g_{k-1}
    double g[CPT], pc[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t k = c0 + (int64_t)threadIdx.x * CPT + c;
      const bool in = k < n;
      pc[c] = in ? P_[k] : 0.0;
      g[c] = in ? R_[k] * I_[k] : 0.0;
    }
    Affine tot;
    cta_forward<TPB, CPT>(g, [&](int c) { return pc[c]; }, sm, carry, &tot);
    carry = fma(tot.A, carry, tot.B);
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t k = c0 + (int64_t)threadIdx.x * CPT + c;
      if (k < n) X_[k] = g[c];
    }
  }
  __syncthreads();
  carry = 0.0;
  const int64_t nch = (n + CH - 1) / CH;
  for (int64_t ch = nch - 1; ch >= 0; --ch) {              // backward: x_k = g_k - q_k x_{k+1}
    const int64_t c0 = ch * CH;
    double x[CPT], qc[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t k = c0 + (int64_t)threadIdx.x * CPT + c;
      const bool in = k < n;
      qc[c] = in ? Q_[k] : 0.0;
      x[c] = in ? X_[k] : 0.0;
    }
    Affine tot;
    cta_backward<TPB, CPT>(x, [&](int c) { return qc[c]; }, sm, carry, &tot);
    carry = fma(tot.A, carry, tot.B);
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t k = c0 + (int64_t)threadIdx.x * CPT + c;
      if (k < n) X_[k] = x[c];
    }
  }
}

// x = g - Vt*xs(k-1) - Wt*xs(k) over the spike extents, x_sep = xs(k); also refreshes the two
// ghost cells of the local state.  xs is indexed GLOBALLY (tile_offset + k).  The extents are
// rounded up to whole threads' chunks: tl threads from the left edge, tr from the right; items
// are enumerated in tiled order so that the accesses coalesce.
__global__ void __launch_bounds__(256) correct_kernel(int64_t n_tiles, int TPB, int CPT, int tl, int tr,
                                                      int64_t tile_offset, int64_t n_tiles_global,
                                                      const double* __restrict__ Vt, const double* __restrict__ Wt,
                                                      const double* __restrict__ xs, double* __restrict__ w,
                                                      int64_t w_stride, const double* __restrict__ next_first) {
  const int T = TPB * CPT;
  const int per_tile = (tl + tr) * CPT + 1;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (idx >= n_tiles * per_tile) return;
  const int64_t k = idx / per_tile;
  const int m = (int)(idx - k * per_tile);
  const double* X_ = xs + pi * n_tiles_global;
  const int64_t kg = tile_offset + k;
  const double xk = X_[kg];
  const double xl = (kg > 0) ? X_[kg - 1] : 0.0;
  double* wp = w + pi * w_stride + k * T;
  if (m == (tl + tr) * CPT) {                              // the separator itself, and the ghosts
    wp[T - 1] = xk;
    if (k == 0) wp[-1] = xl;                               // left ghost = previous rank's last separator
    if (k == n_tiles - 1) wp[T] = next_first ? next_first[pi] : 0.0;   // right ghost = next rank's first cell
    return;
  }
  int t, c;
  if (m < tl * CPT) { c = m / tl; t = m % tl; }
  else { const int m2 = m - tl * CPT; c = m2 / tr; t = TPB - tr + m2 % tr; }
  if (t == TPB - 1 && c == CPT - 1) return;                // the separator is written above
  const int pos = c * TPB + t;
  const int64_t o = (pi * n_tiles + k) * T + pos;
  wp[pos] = fma(-Wt[o], xk, fma(-Vt[o], xl, wp[pos]));
}

}  // namespace fvm

// ------------------------------------------------------------------------------------------ C-ABI
namespace {
constexpr int kTileCfgs = 3;      // 0: 512 x 16 = 8192 cells (production), 1: 64 x 4 (tests), 2: 256 x 16 = 4096 cells, two CTAs per SM
inline int tile_threads(int cfg) { return cfg == 0 ? 512 : (cfg == 1 ? 64 : 256); }
inline int tile_cpt(int cfg) { return cfg == 1 ? 4 : 16; }
inline int tile_cells(int cfg) { return tile_threads(cfg) * tile_cpt(cfg); }
}

extern "C" int64_t fvm_spike_tile_cells(int32_t config) {
  return (config >= 0 && config < kTileCfgs) ? tile_cells(config) : 0;
}

extern "C" int64_t fvm_spike_tile_threads(int32_t config) {
  return (config >= 0 && config < kTileCfgs) ? tile_threads(config) : 0;
}

extern "C" int fvm_spike_setup(const fvm_problem* prob, const double* dt, double theta, int64_t tile_offset,
                               int64_t n_tiles, int32_t config, double* p, double* q, double* b, double* Vt,
                               double* Wt, double* sep, double* bends, int32_t* extents, double spike_tol,
                               void* stream) {
  if (!prob || !prob->faces || !prob->a || !prob->d || !dt || !p || !q || !b || !Vt || !Wt || !sep || !bends ||
      !extents)
    return FVM_ERR_INVALID_ARG;
  if (prob->J < 3 || prob->P < 1 || n_tiles < 1 || tile_offset < 0 || !(theta > 0.0 && theta <= 1.0))
    return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs) return FVM_ERR_INVALID_ARG;
  if (prob->P > 65535) return FVM_ERR_UNSUPPORTED;
  const int T = tile_cells(config);
  const int64_t n_local = n_tiles * T;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(extents, 0, 2 * sizeof(int32_t), s);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemsetAsync(bends, 0, 2 * prob->P * sizeof(double), s);
  if (e != cudaSuccess) return (int)e;
  dim3 g1((unsigned)((n_local + 255) / 256), (unsigned)prob->P);
  fvm::spike_assemble_kernel<<<g1, 256, 0, s>>>(*prob, dt, theta, tile_offset * T, n_local, tile_threads(config),
                                                tile_cpt(config), p, q, b);
  dim3 g2((unsigned)((n_tiles + 127) / 128), (unsigned)prob->P);
  fvm::spike_tile_factor_kernel<<<g2, 128, 0, s>>>(*prob, theta, tile_offset, n_tiles, tile_threads(config),
                                                   tile_cpt(config), p, q, b, Vt, Wt, sep, bends, (int*)extents,
                                                   spike_tol);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_reduced_factor(int64_t n, int64_t P, const double* RL, const double* RD,
                                        const double* RU, double* rp, double* rq, double* rinv, void* stream) {
  if (!RL || !RD || !RU || !rp || !rq || !rinv || n < 1 || P < 1) return FVM_ERR_INVALID_ARG;
  fvm::reduced_factor_kernel<<<(unsigned)((P + 63) / 64), 64, 0, (cudaStream_t)stream>>>(n, P, RL, RD, RU, rp, rq,
                                                                                       rinv);
  return (int)cudaGetLastError();
}

template <int TPB, int CPT>
static int launch_tile_solve(int64_t n_tiles, int64_t P, const double* p, const double* q, const double* b,
                             const double* sep, const double* bfirst, int has_first, int64_t jl_last, double theta,
                             const double* w_in, double* w_out, double* rs, int64_t w_stride, cudaStream_t s,
                             const fvm::LazyCorrection* lazy = nullptr) {
  dim3 grid((unsigned)n_tiles, (unsigned)P);
  const double cth = (1.0 - theta) / theta;
  if (lazy) {
    if (theta == 0.5)
      fvm::tile_solve_kernel<TPB, CPT, 1, true><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, *lazy);
    else if (theta == 1.0)
      fvm::tile_solve_kernel<TPB, CPT, 2, true><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, *lazy);
    else
      fvm::tile_solve_kernel<TPB, CPT, 0, true><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, *lazy);
    return (int)cudaGetLastError();
  }
  const fvm::LazyCorrection none = {nullptr, nullptr, nullptr, 0, 0, 0, 0};
  if (theta == 0.5)
    fvm::tile_solve_kernel<TPB, CPT, 1, false><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, none);
  else if (theta == 1.0)
    fvm::tile_solve_kernel<TPB, CPT, 2, false><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, none);
  else
    fvm::tile_solve_kernel<TPB, CPT, 0, false><<<grid, TPB, 0, s>>>(n_tiles, p, q, b, sep, bfirst, has_first, jl_last,
                                                                   cth, w_in, w_out, rs, w_stride, none);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_local(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                               const double* b, const double* sep, const double* bends, int32_t has_first_cell,
                               int64_t last_cell_local, double theta, const double* w_in, double* w_out,
                               int64_t w_stride, double* rs,
                               double* rr, int64_t rr_stride, double* g_first, void* stream) {
  if (!p || !q || !b || !sep || !bends || !w_in || !w_out || !rs || !rr || !g_first) return FVM_ERR_INVALID_ARG;
  if (n_tiles < 1 || P < 1 || P > 65535 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs || n_tiles > 0x7fffffffLL) return FVM_ERR_INVALID_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  int st = config == 0 ? launch_tile_solve<512, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s)
       : config == 1 ? launch_tile_solve<64, 4>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s)
                     : launch_tile_solve<256, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s);
  if (st) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, tile_threads(config), tile_cpt(config), sep, rs, w_out, w_stride,
                                            rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_reduce_solve(int64_t n_global, int64_t P, const double* rp, const double* rq,
                                      const double* rinv, const double* rr, double* xs, void* stream) {
  if (!rp || !rq || !rinv || !rr || !xs || n_global < 1 || P < 1) return FVM_ERR_INVALID_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (n_global <= 1024)
    fvm::reduced_solve_kernel<128, 8><<<(unsigned)P, 128, 0, s>>>(n_global, rp, rq, rinv, rr, xs);
  else
    fvm::reduced_solve_kernel<1024, 8><<<(unsigned)P, 1024, 0, s>>>(n_global, rp, rq, rinv, rr, xs);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_finalize(int64_t n_tiles, int64_t P, int32_t config, const int32_t* extents_host,
                                  int64_t tile_offset, int64_t n_tiles_global, const double* Vt, const double* Wt,
                                  const double* xs, double* w, int64_t w_stride, const double* next_first,
                                  void* stream) {
  if (!extents_host || !Vt || !Wt || !xs || !w || n_tiles < 1 || P < 1 || P > 65535) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs) return FVM_ERR_INVALID_ARG;
  const int T = tile_cells(config), TPB = tile_threads(config), CPT = tile_cpt(config);
  const int eL = extents_host[0], eR = extents_host[1];
  if (eL < 0 || eR < 0 || eL > T - 1 || eR > T - 1) return FVM_ERR_INVALID_ARG;
  int tl = (eL + CPT - 1) / CPT;                           // whole threads' chunks from the left edge
  int tr = TPB - (T - 1 - eR) / CPT;                       // ... and from the right edge
  if (tl > TPB) tl = TPB;
  if (tr > TPB - tl) tr = TPB - tl;                        // the two ranges never overlap
  const int64_t total = n_tiles * (int64_t)((tl + tr) * CPT + 1);
  dim3 g((unsigned)((total + 255) / 256), (unsigned)P);
  fvm::correct_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(n_tiles, TPB, CPT, tl, tr, tile_offset, n_tiles_global, Vt,
                                                          Wt, xs, w, w_stride, next_first);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Lazy correction: fvm_spike_local_lazy + fvm_spike_separators replace fvm_spike_local +
// fvm_spike_finalize in a time loop.  The state buffer then holds the uncorrected tile solutions
// between steps (true values at the separators and ghosts); fvm_spike_finalize materialises the true
// state when it is needed (output, checkpoints).
static void extent_threads(int config, const int32_t* extents_host, int& tl, int& tr) {
  const int T = tile_cells(config), TPB = tile_threads(config), CPT = tile_cpt(config);
  tl = (extents_host[0] + CPT - 1) / CPT;                  // whole threads' chunks from the left edge
  tr = TPB - (T - 1 - extents_host[1]) / CPT;              // ... and from the right edge
  if (tl > TPB) tl = TPB;
  if (tr > TPB - tl) tr = TPB - tl;                        // the two ranges never overlap
}

extern "C" int fvm_spike_local_lazy(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                                    const double* b, const double* sep, const double* bends,
                                    int32_t has_first_cell, int64_t last_cell_local, double theta,
                                    const double* w_in, double* w_out, int64_t w_stride, double* rs, double* rr,
                                    int64_t rr_stride, double* g_first, const double* Vt, const double* Wt,
                                    const double* xs, int64_t tile_offset, int64_t n_tiles_global,
                                    const int32_t* extents_host, void* stream) {
  if (!p || !q || !b || !sep || !bends || !w_in || !w_out || !rs || !rr || !g_first || !Vt || !Wt || !xs ||
      !extents_host)
    return FVM_ERR_INVALID_ARG;
  if (n_tiles < 1 || P < 1 || P > 65535 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs || n_tiles > 0x7fffffffLL) return FVM_ERR_INVALID_ARG;
  const int T = tile_cells(config);
  if (extents_host[0] < 0 || extents_host[1] < 0 || extents_host[0] > T - 1 || extents_host[1] > T - 1)
    return FVM_ERR_INVALID_ARG;
  fvm::LazyCorrection lz;
  lz.Vt = Vt; lz.Wt = Wt; lz.xs = xs; lz.tile_offset = tile_offset; lz.n_tiles_global = n_tiles_global;
  extent_threads(config, extents_host, lz.tl, lz.tr);
  cudaStream_t s = (cudaStream_t)stream;
  int st = config == 0 ? launch_tile_solve<512, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s, &lz)
       : config == 1 ? launch_tile_solve<64, 4>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s, &lz)
                     : launch_tile_solve<256, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta,
                                            w_in, w_out, rs, w_stride, s, &lz);
  if (st) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, tile_threads(config), tile_cpt(config), sep, rs, w_out, w_stride,
                                            rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_separators(int64_t n_tiles, int64_t P, int32_t config, int64_t tile_offset,
                                    int64_t n_tiles_global, const double* xs, double* w, int64_t w_stride,
                                    const double* next_first, void* stream) {
  if (!xs || !w || n_tiles < 1 || P < 1 || P > 65535) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs) return FVM_ERR_INVALID_ARG;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::correct_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(n_tiles, tile_threads(config), tile_cpt(config), 0, 0,
                                                          tile_offset, n_tiles_global, nullptr, nullptr, xs, w, w_stride,
                                                          next_first);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// fvm_spike_local_static: fvm_spike_local for callers that cannot provide the `rs` work array.
//
// Between its two kernels fvm_spike_local parks the separators' right-hand sides in `rs`, an
// array the caller must allocate.  A time loop that lives inside a larger application often
// cannot do that: the application may have handed every byte of device memory to its own pool
// when it started, or the step may sit in a latency-critical section where cudaMalloc is not
// acceptable because it synchronises the whole device.  For such callers the scratch lives in
// the library image itself, in the fixed-size device array below: every GPU that loads
// libfvm_b200.so carries these 512 KiB until the process exits, and a call needs
// P * n_tiles <= FVM_SPIKE_STATIC_MAX_TILES.
namespace {
constexpr int64_t kStaticMaxTiles = 65536;
static __device__ __align__(256) double g_spike_rs_scratch[kStaticMaxTiles];

// The array's name is only a handle in host code; the address a kernel can use has to be asked
// for while the device that will run the kernel is current.  It is asked for on EVERY call and
// never cached: the lookup is a host-side table access once the library's module is loaded on
// the device, and a cached pointer would go stale after cudaDeviceReset() or under a second
// context on the same device.  The first call on a device loads the module (so it should come
// before any stream capture); the device is the CURRENT one, which must be the stream's.
int static_scratch(double** out) {
  double* ptr = nullptr;
  cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_spike_rs_scratch);
  if (e != cudaSuccess) return (int)e;
  if (ptr == nullptr) return (int)cudaErrorInvalidSymbol;
  *out = ptr;
  return FVM_OK;
}
}  // namespace

extern "C" int64_t fvm_spike_static_max_tiles(void) { return kStaticMaxTiles; }

extern "C" int fvm_spike_local_static(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                                      const double* b, const double* sep, const double* bends,
                                      int32_t has_first_cell, int64_t last_cell_local, double theta,
                                      const double* w_in, double* w_out, int64_t w_stride, double* rr,
                                      int64_t rr_stride, double* g_first, void* stream) {
  if (!p || !q || !b || !sep || !bends || !w_in || !w_out || !rr || !g_first) return FVM_ERR_INVALID_ARG;
  if (n_tiles < 1 || P < 1 || P > 65535 || !(theta > 0.0 && theta <= 1.0)) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs || n_tiles > 0x7fffffffLL) return FVM_ERR_INVALID_ARG;
  if (n_tiles > kStaticMaxTiles || P * n_tiles > kStaticMaxTiles) return FVM_ERR_SCRATCH_LIMIT;
  double* rs = nullptr;
  int st = static_scratch(&rs);
  if (st) return st;
  // Nothing in the scratch has to start at zero: the first kernel writes every entry
  // [0, P*n_tiles) before the second one reads it; what earlier calls left behind is overwritten.
  cudaStream_t s = (cudaStream_t)stream;
  st = config == 0 ? launch_tile_solve<512, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s)
       : config == 1 ? launch_tile_solve<64, 4>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s)
                     : launch_tile_solve<256, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s);
  if (st) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, tile_threads(config), tile_cpt(config), sep, rs, w_out, w_stride,
                                            rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Rank-level SPIKE over the separator system (partitioned.py: _prepare_rank_level): the two small
// kernels around the per-step all-gather.
namespace fvm {

// What a rank contributes to the gather, 4 doubles per problem: first and last value of its block
// solution gr, its separator's right-hand side (the local solve passes it through: that row is
// decoupled) and the first cell of its slab.
__global__ void rank_pack_kernel(int64_t P, int64_t nt, const double* __restrict__ gr,
                                 const double* __restrict__ g_first, double* __restrict__ packet) {
  const int64_t pi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pi >= P) return;
  const double* g = gr + pi * nt;
  double* o = packet + pi * 4;
  o[0] = g[0]; o[1] = g[nt - 2]; o[2] = g[nt - 1]; o[3] = g_first[pi];
}

// After the gather (allp[rank][problem][4]): right-hand side and solution of the `world`-unknown
// top system (dense inverse), this rank's separator values xs_ext[1..nt] with the previous rank's
// separator in xs_ext[0], and the new value of the next rank's first cell (this rank's right ghost).
// cst[problem][7][world]: sup of every rank's last tile, the separator rows' couplings (lo, up), the
// rank spikes at every rank's first tile (vf, wf) and the tile spikes (V0, W0) of every rank's first tile.
__global__ void __launch_bounds__(256) rank_top_kernel(int64_t P, int64_t nt, int world, int rank,
                                                       const double* __restrict__ allp,
                                                       const double* __restrict__ top_inv,
                                                       const double* __restrict__ cst,
                                                       const double* __restrict__ gr, const double* __restrict__ Vr,
                                                       const double* __restrict__ Wr, double* __restrict__ xs_ext,
                                                       double* __restrict__ next_first) {
  __shared__ double rhs[64], X[64];
  const int64_t pi = blockIdx.x;
  const int t = threadIdx.x;
  const double* c = cst + pi * 7 * world;
  auto pk = [&](int r, int i) { return allp[((int64_t)r * P + pi) * 4 + i]; };
  if (t < world) {
    const bool has_next = t + 1 < world;
    const double gfirst_n = has_next ? pk(t + 1, 0) : 0.0, cell_n = has_next ? pk(t + 1, 3) : 0.0;
    rhs[t] = pk(t, 2) - c[0 * world + t] * cell_n - c[1 * world + t] * pk(t, 1) - c[2 * world + t] * gfirst_n;
  }
  __syncthreads();
  if (t < world) {
    const double* row = top_inv + (pi * world + t) * world;
    double acc = 0.0;
    for (int j = 0; j < world; ++j) acc = fma(row[j], rhs[j], acc);
    X[t] = acc;
  }
  __syncthreads();
  const double Xr = X[rank], Xl = rank > 0 ? X[rank - 1] : 0.0;
  double* xe = xs_ext + pi * (nt + 1);
  for (int64_t k = t; k < nt; k += blockDim.x) {
    const int64_t o = pi * nt + k;
    xe[1 + k] = (k == nt - 1) ? Xr : fma(-Wr[o], Xr, fma(-Vr[o], Xl, gr[o]));
  }
  if (t == 0) {
    xe[0] = Xl;
    if (rank + 1 < world) {
      const int n = rank + 1;
      const double xs_n0 = fma(-c[4 * world + n], X[n], fma(-c[3 * world + n], Xr, pk(n, 0)));   // next rank's first separator
      next_first[pi] = fma(-c[6 * world + n], xs_n0, fma(-c[5 * world + n], Xr, pk(n, 3)));
    }
  }
}

}  // namespace fvm

extern "C" int fvm_spike_rank_pack(int64_t P, int64_t nt, const double* gr, const double* g_first, double* packet,
                                   void* stream) {
  if (!gr || !g_first || !packet || P < 1 || nt < 2) return FVM_ERR_INVALID_ARG;
  fvm::rank_pack_kernel<<<(unsigned)((P + 127) / 128), 128, 0, (cudaStream_t)stream>>>(P, nt, gr, g_first, packet);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_rank_top(int64_t P, int64_t nt, int32_t world, int32_t rank, const double* allp,
                                  const double* top_inv, const double* cst, const double* gr, const double* Vr,
                                  const double* Wr, double* xs_ext, double* next_first, void* stream) {
  if (!allp || !top_inv || !cst || !gr || !Vr || !Wr || !xs_ext || !next_first) return FVM_ERR_INVALID_ARG;
  if (P < 1 || P > 0x7fffffffLL || nt < 2 || world < 2 || world > 64 || rank < 0 || rank >= world)
    return FVM_ERR_INVALID_ARG;
  fvm::rank_top_kernel<<<(unsigned)P, 256, 0, (cudaStream_t)stream>>>(P, nt, world, rank, allp, top_inv, cst, gr, Vr,
                                                                     Wr, xs_ext, next_first);
  return (int)cudaGetLastError();
}
