// Variant 3: partitioned SPIKE / Thomas for LONG systems, streaming from HBM.
//
// A system of J cells is cut into tiles of T = TPB*CPT consecutive cells; the LAST cell of
// every tile is a separator.  With the separators known the tile interiors decouple:
//
//   set-up (once per model)
//     spike_assemble_kernel : rows of A = I - dt*theta*alpha*M for every local cell
//     spike_tile_factor_kernel : per tile, un-normalised LU of the interior block (l, kappa, s as in
//                             the other variants, restarted at the tile's first cell) and the tile
//                             spikes Vt = T^-1(lo_first e_first), Wt = T^-1(up_last e_last); every
//                             chain in double-double, every stored number rounded once
//     (host)                  reduced tridiagonal over ALL tile separators of ALL ranks,
//     reduced_factor_kernel   factored once (double-double chain)
//   every step (no right-hand side: w' = A^-1(w/theta + beta) - c w, c = (1-theta)/theta)
//     tile_solve_kernel     : one CTA per tile: tile-local solve of A g = w/theta + beta, parallelised
//                             over the CTA by a scan of affine maps g_out = A*g_in + B (each
//                             thread owns CPT consecutive cells: one sequential pass for its
//                             (A, B), a shuffle/shared-memory scan, one pass with the true carry);
//                             same backwards; writes z = g - c w.  Reads w, l, kappa, s and writes
//                             z: 40 B per cell.  No cell needs its neighbours' old values.
//     reduced_rhs_kernel    : rr_k = w_s/theta + beta_s - lo_k g_last(k) - up_k g_first(k+1)
//     (host, multi-GPU)       all-gather over the ranks (NCCL)
//     reduced_solve_kernel  : Thomas on the factored separator system, one CTA, affine scans: the
//                             separators' values ys of y = A^-1(w/theta + beta)
//     correct_kernel        : w' = z - Vt ys(k-1) - Wt ys(k), only over the extent where a spike
//                             exceeds `spike_tol` (spikes decay away from the tile edges; at the
//                             separator Vt = 0, Wt = -1).  In a time loop this pass is skipped: the
//                             next step's tile_solve_kernel applies the correction while it loads
//                             the tile (LAZY), and nothing is written between steps.
//   multi-GPU, rank level   : rank_pack_kernel / rank_top_kernel around ONE all-gather per step
//
// All indices are LOCAL to the rank's slab unless they say "global"; the slab starts at global
// cell j_offset = tile_offset*T.
//
// LAYOUT.  Thread t of a tile's CTA owns the CPT consecutive cells [t*CPT, (t+1)*CPT) (the
// recurrences need a blocked ownership), but every per-cell array of the library (p, q, b, Vt,
// Wt and the state w) stores cell r = t*CPT + c of a tile at  c*TPB + t  ("tiled" order), so that
// a warp's load of its c-th cells is two full 128-byte lines.  With the natural order the same
// loads touch 32 lines each and the kernel is bound by L1 tag look-ups at 30 % of HBM speed
// (profiles/r01_spike_launches_natural.csv).  The first and last positions of a tile are its
// first and last cell in both orders.
#include "fvm_device.cuh"
#include "cta_scan.cuh"
#include "dd.cuh"
#include <new>

namespace fvm {

// Position of cell r (0..T-1) of a tile in the tiled order.
__device__ __forceinline__ int tiled_pos(int r, int TPB, int CPT) { return (r % CPT) * TPB + r / CPT; }

// Per-tile constants, SEP_STRIDE doubles per tile.
enum { SEP_LO = 0, SEP_UP, SEP_RL, SEP_RDA, SEP_V0, SEP_W0, SEP_STRIDE };

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

// One thread per tile: un-normalised LU of the interior block restarted at the tile's first cell
// (l_c = lo_c / pivot_{c-1}, kappa_c = 1 / (theta pivot_c), s_c = up_c / pivot_c; DESIGN.md section 4,
// the same factors as every other variant), the tile spikes, the separator's rows and the tile's
// contribution to the reduced system.  Every chain -- the pivots from the top, the pivots from the
// bottom, the products that make the spikes -- is carried in double-double and every stored number is
// rounded ONCE: a theta step applies these numbers ten thousand times, so a rounding error that a
// chain has amplified is a systematic error of the run (dd.cuh).  In: p, b, q = lo, di, up of A.
// Out: p = l, b = kappa, q = s, Vt, Wt; at the separator l = kappa = s = 0, Vt = 0 and Wt = -1, so
// that the correction  x = z - Vt xs(k-1) - Wt xs(k)  ADDS the separator's own value there.
__global__ void __launch_bounds__(128) spike_tile_factor_kernel(
    fvm_problem pr, double theta, int64_t tile_offset, int64_t n_tiles, int TPB, int CPT, double* __restrict__ p,
    double* __restrict__ q, double* __restrict__ b, double* __restrict__ Vt, double* __restrict__ Wt,
    double* __restrict__ sep, double* __restrict__ bends, int* __restrict__ extents, double spike_tol) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (k >= n_tiles) return;
  const int T = TPB * CPT;
  const int64_t base = (pi * n_tiles + k) * T;
  const int64_t jg0 = (tile_offset + k) * T;               // global index of the tile's first cell
  auto at = [&](int r) { return base + tiled_pos(r, TPB, CPT); };
  const int N = T - 1;                                     // interior cells
  // theta * beta of the two end cells of the mesh, added to the state before the forward sweep.  Only
  // the tile that HOLDS the end cell may touch the end faces: with the mesh split over ranks `faces` is
  // the caller's pointer shifted by the slab offset, and a rank that does not own cell 0 (or J-1) has
  // no memory behind faces[0..1] (faces[J-1..J]).
  if (jg0 == 0 && pr.left_kind == FVM_BC_ROBIN && pr.left_flux)
    bends[2 * pi] = mul(theta, dvd(pr.left_flux[pi], sub(pr.faces[1], pr.faces[0])));
  if (pr.J - 1 >= jg0 && pr.J - 1 < jg0 + T && pr.right_kind == FVM_BC_ROBIN && pr.right_flux)
    bends[2 * pi + 1] = mul(theta, dvd(-pr.right_flux[pi], sub(pr.faces[pr.J], pr.faces[pr.J - 1])));
  const int64_t oN = at(N);
  const double loN = p[oN], diN = b[oN], upN = q[oN];
  // pass 1, bottom-up: pivots of the UL factorisation, pv_c = di_c - up_c lo_{c+1} / pv_{c+1}; the left
  // spike V = T^-1 (lo_0 e_0) is then a running product of m_c = lo_c / pv_c.  m is parked in Vt (hi), Wt (lo).
  {
    dd m = dd_make(0.0);
    for (int c = N - 1; c >= 0; --c) {
      const int64_t o = at(c);
      const dd pv = dd_sub(dd_make(b[o]), dd_mul_d(m, q[o]));
      m = dd_div(dd_make(p[o]), pv);
      Vt[o] = m.hi; Wt[o] = m.lo;
    }
  }
  // pass 2, top-down: pivots of the LU factorisation, the stored factors, and V_c = -m_c V_{c-1}.
  // s_c is needed once more in double-double (pass 3): its low part is parked in Wt.
  const dd dth = dd_make(theta);
  dd sprev = dd_make(0.0), rprev = dd_make(0.0), V = dd_make(0.0), Vlast = dd_make(0.0);
  double V0 = 0.0;
  int eL = 0, eR = 1;
  for (int c = 0; c < N; ++c) {
    const int64_t o = at(c);
    const double lo = p[o], di = b[o], up = q[o];
    const dd piv = dd_sub(dd_make(di), dd_mul_d(sprev, lo));
    const dd l = dd_mul_d(rprev, lo);                      // c = 0: couples to the previous separator -> spike
    rprev = dd_recip(piv);
    sprev = dd_mul_d(rprev, up);
    const dd m = dd_make(Vt[o], Wt[o]);
    V = (c == 0) ? m : dd_neg(dd_mul(m, V));
    const double Vr = dd_round(V);
    if (c == 0) V0 = Vr;
    if (fabs(Vr) > spike_tol) eL = c + 1;                  // cells [0, eL) need the left correction
    p[o] = (c == 0) ? 0.0 : dd_round(l);
    b[o] = dd_round(dd_div(rprev, dth));
    q[o] = sprev.hi;
    Vt[o] = Vr; Wt[o] = sprev.lo;
  }
  Vlast = V;
  // pass 3, bottom-up: right spike W = T^-1 (up_{N-1} e_{N-1}): W_{N-1} = s_{N-1}, W_c = -s_c W_{c+1}
  dd Wd = dd_make(0.0), Wlast = dd_make(0.0);
  for (int c = N - 1; c >= 0; --c) {
    const int64_t o = at(c);
    const dd sc = dd_make(q[o], Wt[o]);
    Wd = (c == N - 1) ? sc : dd_neg(dd_mul(sc, Wd));
    if (c == N - 1) Wlast = Wd;
    const double Wr = dd_round(Wd);
    if (fabs(Wr) > spike_tol) eR = N - c;                  // cells [N-eR, N) need the right correction
    Wt[o] = Wr;
  }
  q[at(N - 1)] = 0.0;                                      // s_{N-1} couples to the separator -> spike, not sweep
  p[oN] = 0.0; q[oN] = 0.0; b[oN] = 0.0; Vt[oN] = 0.0; Wt[oN] = -1.0;
  double* s = sep + (pi * n_tiles + k) * SEP_STRIDE;
  s[SEP_LO] = loN; s[SEP_UP] = upN;
  s[SEP_RL] = -dd_round(dd_mul_d(Vlast, loN));
  s[SEP_RDA] = dd_round(dd_sub(dd_make(diN), dd_mul_d(Wlast, loN)));   // the -up*V0(k+1) term is added by the host
  s[SEP_V0] = V0; s[SEP_W0] = dd_round(Wd);
  atomicMax(&extents[0], eL);
  atomicMax(&extents[1], eR);
}

// Thomas factorisation of the reduced tridiagonal (one thread per problem; n = global tiles): pivot
// chain in double-double, stored factors rounded once.
__global__ void reduced_factor_kernel(int64_t n, int64_t P, const double* __restrict__ RL,
                                      const double* __restrict__ RD, const double* __restrict__ RU,
                                      double* __restrict__ rp, double* __restrict__ rq, double* __restrict__ rinv) {
  const int64_t pi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (pi >= P) return;
  dd q_prev = dd_make(0.0);
  for (int64_t k = 0; k < n; ++k) {
    const int64_t o = pi * n + k;
    const dd inv = dd_recip(dd_sub(dd_make(RD[o]), dd_mul_d(q_prev, RL[o])));
    rp[o] = dd_round(dd_mul_d(inv, RL[o]));
    q_prev = dd_mul_d(inv, RU[o]);
    rq[o] = q_prev.hi;
    rinv[o] = inv.hi;
  }
}

// ------------------------------------------------------------------------------------------ per step
// One theta step without a right-hand side: with B = I/theta - cA, c = (1-theta)/theta,
//   w' = A^-1 (w/theta + beta) - c w,
// and per tile, with the un-normalised factors:  h_c = w_c + theta beta_c - l_c h_{c-1}  (forwards),
// g_c = h_c kappa_c - s_c g_{c+1}  (backwards),  z_c = g_c - c w_c.  z is what the kernel writes: the
// new state up to the spike correction  w' = z - Vt ys(k-1) - Wt ys(k), where ys are the separators'
// values of y = A^-1(...), the solution of the reduced system.  No neighbour of a cell enters, so a
// tile needs nothing from the tiles next to it.
//
// LAZY correction: w_in holds the previous step's UNCORRECTED z; the correction is applied while the
// tile is loaded, by the threads whose cells lie inside the spike extents (the first tl and the last
// tr >= 1 threads of the CTA; the separator is the last cell of the last thread, its Wt is -1).  This
// replaces correct_kernel's separate read-modify-write pass: Vt, Wt are still read (16 B per corrected
// cell) but z is not re-read and the corrected state is never written.
struct LazyCorrection {
  const double* Vt; const double* Wt; const double* xs;   // xs indexed by tile_offset + k, per problem n_tiles_global
  int64_t tile_offset, n_tiles_global;
  int tl, tr;
};

__device__ __forceinline__ void cp_async8(double* smem_dst, const double* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
               : "memory");
}

// MODE 0: general theta, 1: theta = 1/2 (c = 1), 2: theta = 1 (c = 0).  Dynamic shared memory: s of the
// tile (T doubles, filled by cp.async so that it never occupies registers across the forward scan) and,
// unless c = 0, the loaded state (T doubles); every thread reads back only what it parked itself.
template <int TPB, int CPT, int MODE, bool LAZY>
__global__ void __launch_bounds__(TPB, (TPB == 256 ? 2 : 1)) tile_solve_kernel(
    int64_t n_tiles, const double* __restrict__ fl, const double* __restrict__ fs,
    const double* __restrict__ fk, const double* __restrict__ sep, const double* __restrict__ bends,
    int has_first_cell, int64_t jl_last, double rth, double cth, const double* __restrict__ w_in,
    double* __restrict__ w_out, double* __restrict__ rs, int64_t w_stride, LazyCorrection lz) {
  constexpr int T = TPB * CPT;
  extern __shared__ double park[];                         // [0, T): s; [T, 2T): w
  __shared__ double sm[66];
  const int64_t k = blockIdx.x, pi = blockIdx.y;
  const int tid = threadIdx.x;
  const int64_t base = (pi * n_tiles + k) * T + tid;                 // into the factors (tiled: + c*TPB)
  const int64_t wbase = pi * w_stride + k * T + tid;                 // into the state
#pragma unroll
  for (int c = 0; c < CPT; ++c) cp_async8(park + c * TPB + tid, fs + base + c * TPB);
  asm volatile("cp.async.commit_group;" ::: "memory");
  double h[CPT], lc[CPT], kc[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    h[c] = w_in[wbase + c * TPB]; lc[c] = fl[base + c * TPB]; kc[c] = fk[base + c * TPB];
  }
  if (LAZY) {
    if (tid < lz.tl || tid >= TPB - lz.tr) {
      const int64_t kg = lz.tile_offset + k;
      const double* X_ = lz.xs + pi * lz.n_tiles_global;
      const double xk = X_[kg];
      const double xl = kg > 0 ? X_[kg - 1] : 0.0;
#pragma unroll
      for (int c = 0; c < CPT; ++c)
        h[c] = fma(-lz.Wt[base + c * TPB], xk, fma(-lz.Vt[base + c * TPB], xl, h[c]));
    }
  }
  if (MODE != 2) {
#pragma unroll
    for (int c = 0; c < CPT; ++c) park[T + c * TPB + tid] = h[c];
  }
  if (has_first_cell && k == 0 && tid == 0) h[0] += bends[2 * pi];
  {                                                        // theta*beta[J-1] (its cell is wherever the mesh ends)
    const int64_t l0 = k * T + (int64_t)tid * CPT;
    if (jl_last >= l0 && jl_last < l0 + CPT) {
      const int cl = (int)(jl_last - l0);
#pragma unroll
      for (int c = 0; c < CPT; ++c) if (c == cl) h[c] += bends[2 * pi + 1];
    }
  }
  cta_forward<TPB, CPT>(h, [&](int c) { return lc[c]; }, sm, 0.0, nullptr);
  const double h_sep = h[CPT - 1];                         // last thread: w_s + theta*beta_s (l = 0 there)
#pragma unroll
  for (int c = 0; c < CPT; ++c) h[c] *= kc[c];
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  cta_backward<TPB, CPT>(h, [&](int c) { return park[c * TPB + tid]; }, sm, 0.0, nullptr);
  // what the reduced system needs of this tile: the separator row's right-hand side less its coupling
  // to the tile's last interior cell, and the tile's first value (coupled to the previous separator)
  double* r2 = rs + 2 * (pi * n_tiles + k);
  if (tid == TPB - 1) r2[0] = fma(-sep[(pi * n_tiles + k) * SEP_STRIDE + SEP_LO], h[CPT - 2], h_sep * rth);
  if (tid == 0) r2[1] = h[0];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    double z = h[c];
    if (MODE == 1) z = h[c] - park[T + c * TPB + tid];
    if (MODE == 0) z = fma(-cth, park[T + c * TPB + tid], h[c]);
    w_out[wbase + c * TPB] = z;
  }
}

// rr_k = rs_k - up_k g_first(k+1); the last local tile leaves that term to whoever knows the next
// rank's first value (g_first_out[pi] = this rank's own first value).
__global__ void __launch_bounds__(256) reduced_rhs_kernel(int64_t n_tiles, const double* __restrict__ sep,
                                                          const double* __restrict__ rs, double* __restrict__ rr,
                                                          int64_t rr_stride, double* __restrict__ g_first_out) {
  const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t pi = blockIdx.y;
  if (k >= n_tiles) return;
  const double* r2 = rs + 2 * (pi * n_tiles + k);
  double r = r2[0];
  if (k + 1 < n_tiles) r = fma(-sep[(pi * n_tiles + k) * SEP_STRIDE + SEP_UP], r2[3], r);
  rr[pi * rr_stride + k] = r;
  if (k == 0) g_first_out[pi] = r2[1];
}

// FUSED form of reduced_solve_kernel (fvm_spike_steps): the right-hand side is built while it is loaded
// (what reduced_rhs_kernel does) and the block's contribution to the gather is written at the end (what
// rank_pack_kernel does), so a step has two launches fewer.  nt: tiles per problem; B: blocks per problem.
struct FusedRhs {
  const double* rs; const double* sep; int64_t nt, B; double* g_first; double* packet;
};

// Thomas solve of the factored reduced system by ONE CTA per problem (per block of it), in chunks of
// TPB*CPT unknowns with the carry handed from chunk to chunk.
template <int TPB, int CPT, bool FUSED>
__global__ void __launch_bounds__(TPB) reduced_solve_kernel(int64_t n, const double* __restrict__ rp,
                                                            const double* __restrict__ rq,
                                                            const double* __restrict__ rinv,
                                                            const double* __restrict__ rr, double* __restrict__ xs,
                                                            FusedRhs fr) {
  constexpr int CH = TPB * CPT;
  __shared__ double sm[66];
  const int64_t pi = blockIdx.x;
  const double* P_ = rp + pi * n; const double* Q_ = rq + pi * n; const double* I_ = rinv + pi * n;
  const double* R_ = rr + pi * n; double* X_ = xs + pi * n;
  const int64_t prob = FUSED ? pi / fr.B : 0, kt0 = FUSED ? (pi % fr.B) * n : 0;
  // rr_k = rs_k - up_k g_first(k+1), the term left out at the last tile of the slab (reduced_rhs_kernel)
  auto rhs = [&](int64_t k) {
    if (!FUSED) return R_[k];
    const int64_t t = prob * fr.nt + kt0 + k;
    double r = fr.rs[2 * t];
    if (kt0 + k + 1 < fr.nt) r = fma(-fr.sep[t * SEP_STRIDE + SEP_UP], fr.rs[2 * t + 3], r);
    return r;
  };
  double carry = 0.0;
  for (int64_t c0 = 0; c0 < n; c0 += CH) {                 // forward: g_k = rr_k*inv_k - p_k g_{k-1}
    double g[CPT], pc[CPT];
#pragma unroll
    for (int c = 0; c < CPT; ++c) {
      const int64_t k = c0 + (int64_t)threadIdx.x * CPT + c;
      const bool in = k < n;
      pc[c] = in ? P_[k] : 0.0;
      g[c] = in ? rhs(k) * I_[k] : 0.0;
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
  if (FUSED) {
    __syncthreads();
    if (threadIdx.x == 0) {
      const double gf = fr.rs[2 * prob * fr.nt + 1];
      if (kt0 == 0) fr.g_first[prob] = gf;
      if (fr.packet) {
        double* o = fr.packet + pi * 4;
        o[0] = X_[0]; o[1] = X_[n - 2]; o[2] = X_[n - 1]; o[3] = gf;
      }
    }
  }
}

// w = z - Vt*xs(k-1) - Wt*xs(k) over the spike extents (the separator is the last cell of the right
// extent: Vt = 0, Wt = -1 there).  xs is indexed GLOBALLY (tile_offset + k).  The extents are rounded up
// to whole threads' chunks: tl threads from the left edge, tr >= 1 from the right; items are enumerated
// in tiled order so that the accesses coalesce.  NOT idempotent: once per step.
__global__ void __launch_bounds__(256) correct_kernel(int64_t n_tiles, int TPB, int CPT, int tl, int tr,
                                                      int64_t tile_offset, int64_t n_tiles_global,
                                                      const double* __restrict__ Vt, const double* __restrict__ Wt,
                                                      const double* __restrict__ xs, double* __restrict__ w,
                                                      int64_t w_stride) {
  const int T = TPB * CPT;
  const int per_tile = (tl + tr) * CPT;
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
  int t, c;
  if (m < tl * CPT) { c = m / tl; t = m % tl; }
  else { const int m2 = m - tl * CPT; c = m2 / tr; t = TPB - tr + m2 % tr; }
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

template <int TPB, int CPT, int MODE, bool LAZY>
static int launch_tile_solve_inst(dim3 grid, cudaStream_t s, int64_t n_tiles, const double* p, const double* q,
                                  const double* b, const double* sep, const double* bends, int has_first,
                                  int64_t jl_last, double theta, const double* w_in, double* w_out, double* rs,
                                  int64_t w_stride, const fvm::LazyCorrection& lz) {
  auto kern = fvm::tile_solve_kernel<TPB, CPT, MODE, LAZY>;
  const size_t smem = (size_t)TPB * CPT * sizeof(double) * (MODE == 2 ? 1 : 2);
  if (smem > 48 * 1024) {                                  // per call, not cached: valid on whichever device is current
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
  }
  kern<<<grid, TPB, smem, s>>>(n_tiles, p, q, b, sep, bends, has_first, jl_last, 1.0 / theta, (1.0 - theta) / theta,
                               w_in, w_out, rs, w_stride, lz);
  return (int)cudaGetLastError();
}

template <int TPB, int CPT>
static int launch_tile_solve(int64_t n_tiles, int64_t P, const double* p, const double* q, const double* b,
                             const double* sep, const double* bfirst, int has_first, int64_t jl_last, double theta,
                             const double* w_in, double* w_out, double* rs, int64_t w_stride, cudaStream_t s,
                             const fvm::LazyCorrection* lazy = nullptr) {
  dim3 grid((unsigned)n_tiles, (unsigned)P);
  const fvm::LazyCorrection none = {nullptr, nullptr, nullptr, 0, 0, 0, 0};
#define FVM_TILE_INST(MODE, LAZY, LZ)                                                                          \
  launch_tile_solve_inst<TPB, CPT, MODE, LAZY>(grid, s, n_tiles, p, q, b, sep, bfirst, has_first, jl_last, theta, w_in, \
                                               w_out, rs, w_stride, LZ)
  if (lazy) return theta == 0.5 ? FVM_TILE_INST(1, true, *lazy) : theta == 1.0 ? FVM_TILE_INST(2, true, *lazy)
                                                                               : FVM_TILE_INST(0, true, *lazy);
  return theta == 0.5 ? FVM_TILE_INST(1, false, none) : theta == 1.0 ? FVM_TILE_INST(2, false, none)
                                                                     : FVM_TILE_INST(0, false, none);
#undef FVM_TILE_INST
}

// with_rhs = false: the tile solves only; the caller builds the reduced right-hand side itself (fvm_spike_steps:
// inside the fused separator solve)
static int spike_local_impl(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                            const double* b, const double* sep, const double* bends, int32_t has_first_cell,
                            int64_t last_cell_local, double theta, const double* w_in, double* w_out,
                            int64_t w_stride, double* rs, double* rr, int64_t rr_stride, double* g_first,
                            void* stream, bool with_rhs) {
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
  if (st || !with_rhs) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, sep, rs, rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_local(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                               const double* b, const double* sep, const double* bends, int32_t has_first_cell,
                               int64_t last_cell_local, double theta, const double* w_in, double* w_out,
                               int64_t w_stride, double* rs,
                               double* rr, int64_t rr_stride, double* g_first, void* stream) {
  return spike_local_impl(n_tiles, P, config, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in, w_out,
                          w_stride, rs, rr, rr_stride, g_first, stream, true);
}

extern "C" int fvm_spike_reduce_solve(int64_t n_global, int64_t P, const double* rp, const double* rq,
                                      const double* rinv, const double* rr, double* xs, void* stream) {
  if (!rp || !rq || !rinv || !rr || !xs || n_global < 1 || P < 1) return FVM_ERR_INVALID_ARG;
  cudaStream_t s = (cudaStream_t)stream;
  if (n_global <= 1024)
    fvm::reduced_solve_kernel<128, 8, false><<<(unsigned)P, 128, 0, s>>>(n_global, rp, rq, rinv, rr, xs, fvm::FusedRhs{});
  else
    fvm::reduced_solve_kernel<1024, 8, false><<<(unsigned)P, 1024, 0, s>>>(n_global, rp, rq, rinv, rr, xs, fvm::FusedRhs{});
  return (int)cudaGetLastError();
}

// whole threads' chunks of the spike extents: tl threads from the left edge, tr >= 1 from the right
// (the separator is the last cell of the last thread); the two ranges never overlap
static void extent_threads(int config, const int32_t* extents_host, int& tl, int& tr) {
  const int T = tile_cells(config), TPB = tile_threads(config), CPT = tile_cpt(config);
  tl = (extents_host[0] + CPT - 1) / CPT;
  tr = TPB - (T - 1 - extents_host[1]) / CPT;
  if (tr < 1) tr = 1;
  if (tr > TPB) tr = TPB;
  if (tl > TPB - tr) tl = TPB - tr;
}

extern "C" int fvm_spike_finalize(int64_t n_tiles, int64_t P, int32_t config, const int32_t* extents_host,
                                  int64_t tile_offset, int64_t n_tiles_global, const double* Vt, const double* Wt,
                                  const double* xs, double* w, int64_t w_stride, void* stream) {
  if (!extents_host || !Vt || !Wt || !xs || !w || n_tiles < 1 || P < 1 || P > 65535) return FVM_ERR_INVALID_ARG;
  if (config < 0 || config >= kTileCfgs) return FVM_ERR_INVALID_ARG;
  const int T = tile_cells(config), TPB = tile_threads(config), CPT = tile_cpt(config);
  if (extents_host[0] < 0 || extents_host[1] < 0 || extents_host[0] > T - 1 || extents_host[1] > T - 1)
    return FVM_ERR_INVALID_ARG;
  int tl, tr;
  extent_threads(config, extents_host, tl, tr);
  const int64_t total = n_tiles * (int64_t)((tl + tr) * CPT);
  dim3 g((unsigned)((total + 255) / 256), (unsigned)P);
  fvm::correct_kernel<<<g, 256, 0, (cudaStream_t)stream>>>(n_tiles, TPB, CPT, tl, tr, tile_offset, n_tiles_global, Vt,
                                                          Wt, xs, w, w_stride);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Lazy correction: fvm_spike_local_lazy replaces fvm_spike_local + fvm_spike_finalize in a time loop.
// The state buffer then holds the uncorrected z between steps and the separator values of the step
// that produced it stay in `xs`; fvm_spike_finalize materialises the true state when it is needed
// (output, checkpoints).
static int spike_local_lazy_impl(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                                 const double* b, const double* sep, const double* bends,
                                 int32_t has_first_cell, int64_t last_cell_local, double theta,
                                 const double* w_in, double* w_out, int64_t w_stride, double* rs, double* rr,
                                 int64_t rr_stride, double* g_first, const double* Vt, const double* Wt,
                                 const double* xs, int64_t tile_offset, int64_t n_tiles_global,
                                 const int32_t* extents_host, void* stream, bool with_rhs) {
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
  if (st || !with_rhs) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, sep, rs, rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_local_lazy(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                                    const double* b, const double* sep, const double* bends,
                                    int32_t has_first_cell, int64_t last_cell_local, double theta,
                                    const double* w_in, double* w_out, int64_t w_stride, double* rs, double* rr,
                                    int64_t rr_stride, double* g_first, const double* Vt, const double* Wt,
                                    const double* xs, int64_t tile_offset, int64_t n_tiles_global,
                                    const int32_t* extents_host, void* stream) {
  return spike_local_lazy_impl(n_tiles, P, config, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                               w_out, w_stride, rs, rr, rr_stride, g_first, Vt, Wt, xs, tile_offset, n_tiles_global,
                               extents_host, stream, true);
}

// ------------------------------------------------------------------------------------------
// fvm_spike_local_static: fvm_spike_local for callers that cannot provide the `rs` work array.
//
// Between its two kernels fvm_spike_local parks two numbers per tile (the separator's right-hand
// side and the tile's first value) in `rs`, an array the caller must allocate.  A time loop that lives inside a larger application often
// cannot do that: the application may have handed every byte of device memory to its own pool
// when it started, or the step may sit in a latency-critical section where cudaMalloc is not
// acceptable because it synchronises the whole device.  For such callers the scratch lives in
// the library image itself, in the fixed-size device array below: every GPU that loads
// libfvm_b200.so carries these 512 KiB until the process exits, and a call needs
// P * n_tiles <= FVM_SPIKE_STATIC_MAX_TILES.
namespace {
constexpr int64_t kStaticMaxTiles = 32768;
static __device__ __align__(256) double g_spike_rs_scratch[2 * kStaticMaxTiles];

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
  // [0, 2*P*n_tiles) before the second one reads it; what earlier calls left behind is overwritten.
  cudaStream_t s = (cudaStream_t)stream;
  st = config == 0 ? launch_tile_solve<512, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s)
       : config == 1 ? launch_tile_solve<64, 4>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s)
                     : launch_tile_solve<256, 16>(n_tiles, P, p, q, b, sep, bends, has_first_cell, last_cell_local, theta, w_in,
                                        w_out, rs, w_stride, s);
  if (st) return st;
  dim3 g((unsigned)((n_tiles + 255) / 256), (unsigned)P);
  fvm::reduced_rhs_kernel<<<g, 256, 0, s>>>(n_tiles, sep, rs, rr, rr_stride, g_first);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// SPIKE once more over the separator system (partitioned.py: _prepare_block_level).  The reduced
// tridiagonal (one unknown per tile) is cut into blocks of m consecutive unknowns -- B blocks per rank,
// world*B in all -- whose last unknown is the BLOCK separator.  The blocks are solved in parallel
// (fvm_spike_reduce_solve with the blocks as its batch: one CTA each instead of one CTA for the whole
// system), the block separators satisfy a (world*B)-unknown tridiagonal that every CTA of the last
// kernel solves for itself, and across GPUs a step exchanges 4 doubles per block.  The two small
// kernels around the per-step all-gather:
namespace fvm {

// What a block contributes to the gather, 4 doubles: first and last interior value of its block
// solution gr, its separator's right-hand side (the block solve passes it through: that row is
// decoupled) and the first tile value of the rank's slab (used for block 0 only).
__global__ void rank_pack_kernel(int64_t P, int64_t B, int64_t m, const double* __restrict__ gr,
                                 const double* __restrict__ g_first, double* __restrict__ packet) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P * B) return;
  const double* g = gr + i * m;
  double* o = packet + i * 4;
  o[0] = g[0]; o[1] = g[m - 2]; o[2] = g[m - 1]; o[3] = g_first[i / B];
}

// After the gather (allp[rank][problem][block][4]): right-hand side and solution X of the top system
// (factors tp, tq, tinv [problem][world*B], affine scans over the CTA), then this rank's separator
// values xs_ext[1..nt] with the previous rank's last separator in xs_ext[0]:
//   block separator = X(block),  others = gr - Vr X(block-1) - Wr X(block).
// cst[problem][3][world*B]: coupling of a rank's last separator to the next rank's first tile value
// (0 for the other blocks), and the block separator rows' couplings (lo, up) to the neighbouring block
// interiors.  grid = (problems, shares): every CTA solves the small top system for itself and then
// writes its share of the nt separators.
constexpr int kTopTPB = 256, kTopCPT = 4;                  // world*B <= 1024

__global__ void __launch_bounds__(kTopTPB) rank_top_kernel(int64_t P, int64_t B, int64_t m, int world, int rank,
                                                           const double* __restrict__ allp,
                                                           const double* __restrict__ tp, const double* __restrict__ tq,
                                                           const double* __restrict__ tinv,
                                                           const double* __restrict__ cst,
                                                           const double* __restrict__ gr, const double* __restrict__ Vr,
                                                           const double* __restrict__ Wr, double* __restrict__ xs_ext) {
  __shared__ double X[kTopTPB * kTopCPT];
  __shared__ double sm[66];
  const int64_t pi = blockIdx.x;
  const int t = threadIdx.x;
  const int NB = world * (int)B;
  const double* c = cst + pi * 3 * NB;
  auto pk = [&](int blk, int i) { return allp[((((int64_t)(blk / B)) * P + pi) * B + blk % B) * 4 + i]; };
  double g[kTopCPT], pc[kTopCPT], qc[kTopCPT];
#pragma unroll
  for (int j = 0; j < kTopCPT; ++j) {
    const int blk = t * kTopCPT + j;
    g[j] = 0.0; pc[j] = 0.0; qc[j] = 0.0;
    if (blk < NB) {
      double r = pk(blk, 2) - c[NB + blk] * pk(blk, 1);
      if (blk + 1 < NB) r -= c[blk] * pk(blk + 1, 3) + c[2 * NB + blk] * pk(blk + 1, 0);
      g[j] = r * tinv[pi * NB + blk];
      pc[j] = tp[pi * NB + blk]; qc[j] = tq[pi * NB + blk];
    }
  }
  cta_forward<kTopTPB, kTopCPT>(g, [&](int j) { return pc[j]; }, sm, 0.0, nullptr);
  cta_backward<kTopTPB, kTopCPT>(g, [&](int j) { return qc[j]; }, sm, 0.0, nullptr);
#pragma unroll
  for (int j = 0; j < kTopCPT; ++j) X[t * kTopCPT + j] = g[j];
  __syncthreads();
  const int64_t nt = B * m;
  const int64_t share = (nt + gridDim.y - 1) / gridDim.y;
  const int64_t k0 = (int64_t)blockIdx.y * share, k1 = min(nt, k0 + share);
  double* xe = xs_ext + pi * (nt + 1);
  for (int64_t k = k0 + t; k < k1; k += kTopTPB) {
    const int blk = rank * (int)B + (int)(k / m);
    const double Xr = X[blk], Xl = blk > 0 ? X[blk - 1] : 0.0;
    const int64_t o = pi * nt + k;
    xe[1 + k] = (k % m == m - 1) ? Xr : fma(-Wr[o], Xr, fma(-Vr[o], Xl, gr[o]));
  }
  if (blockIdx.y == 0 && t == 0) xe[0] = rank > 0 ? X[rank * (int)B - 1] : 0.0;
}

}  // namespace fvm

extern "C" int fvm_spike_rank_pack(int64_t P, int64_t B, int64_t m, const double* gr, const double* g_first,
                                   double* packet, void* stream) {
  if (!gr || !g_first || !packet || P < 1 || B < 1 || m < 2) return FVM_ERR_INVALID_ARG;
  fvm::rank_pack_kernel<<<(unsigned)((P * B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(P, B, m, gr, g_first, packet);
  return (int)cudaGetLastError();
}

extern "C" int fvm_spike_rank_top(int64_t P, int64_t B, int64_t m, int32_t world, int32_t rank, const double* allp,
                                  const double* tp, const double* tq, const double* tinv, const double* cst,
                                  const double* gr, const double* Vr, const double* Wr, double* xs_ext,
                                  void* stream) {
  if (!allp || !tp || !tq || !tinv || !cst || !gr || !Vr || !Wr || !xs_ext) return FVM_ERR_INVALID_ARG;
  if (P < 1 || P > 0x7fffffffLL || B < 1 || m < 2 || world < 1 || rank < 0 || rank >= world)
    return FVM_ERR_INVALID_ARG;
  if ((int64_t)world * B > fvm::kTopTPB * fvm::kTopCPT) return FVM_ERR_UNSUPPORTED;
  int64_t shares = (B * m + 511) / 512;                    // ~512 separators per CTA: two passes of its 256 threads
  if (shares > 256) shares = 256;                          // (the redundant top solve is cheap; the pass over the
                                                           // separators is a chain of load latencies)
  dim3 grid((unsigned)P, (unsigned)shares);
  fvm::rank_top_kernel<<<grid, fvm::kTopTPB, 0, (cudaStream_t)stream>>>(P, B, m, world, rank, allp, tp, tq, tinv, cst,
                                                                       gr, Vr, Wr, xs_ext);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// fvm_spike_steps: a whole time loop in one host call (see the header).  Only sequences the entry
// points above on the caller's stream; nothing is allocated, nothing synchronises.
extern "C" int fvm_spike_steps(const fvm_spike_plan* pl, double* w0, double* w1, int32_t cur, int32_t pending,
                               int64_t nsteps, fvm_allgather_fn allgather, void* comm, void* stream) {
  if (!pl || !w0 || !w1 || nsteps < 0 || cur < 0 || cur > 1) return FVM_ERR_INVALID_ARG;
  if (pl->world < 1 || pl->rank < 0 || pl->rank >= pl->world || pl->B < 1) return FVM_ERR_INVALID_ARG;
  const bool blocks = pl->B > 1 || pl->world > 1;
  if (blocks && (pl->B * pl->m != pl->n_tiles || pl->m < 2)) return FVM_ERR_INVALID_ARG;
  if (blocks && (!pl->Vr || !pl->Wr || !pl->tp || !pl->tq || !pl->tinv || !pl->cst || !pl->gr || !pl->packet || !pl->allp))
    return FVM_ERR_INVALID_ARG;
  if (pl->world > 1 && !allgather) return FVM_ERR_INVALID_ARG;
  if (!pl->xs || !pl->rp || !pl->rq || !pl->rinv) return FVM_ERR_INVALID_ARG;
  double* w[2] = {w0, w1};
  const int64_t xs_off = blocks ? 1 : 0, xs_n = blocks ? pl->n_tiles + 1 : pl->n_tiles;
  for (int64_t i = 0; i < nsteps; ++i) {
    const double* w_in = w[cur];
    double* w_out = w[1 - cur];
    int st;
    if (pending)
      st = spike_local_lazy_impl(pl->n_tiles, pl->P, pl->config, pl->p, pl->q, pl->b, pl->sep, pl->bends,
                                 pl->has_first_cell, pl->last_cell_local, pl->theta, w_in, w_out, pl->w_stride, pl->rs,
                                 pl->rr, pl->n_tiles, pl->g_first, pl->Vt, pl->Wt, pl->xs, xs_off, xs_n, pl->extents,
                                 stream, false);
    else
      st = spike_local_impl(pl->n_tiles, pl->P, pl->config, pl->p, pl->q, pl->b, pl->sep, pl->bends, pl->has_first_cell,
                            pl->last_cell_local, pl->theta, w_in, w_out, pl->w_stride, pl->rs, pl->rr, pl->n_tiles,
                            pl->g_first, stream, false);
    if (st) return st;
    // the separator solve builds its right-hand side while loading it and writes the blocks' packets itself:
    // reduced_rhs_kernel and rank_pack_kernel, fused (same arithmetic, same bits)
    {
      const int64_t n = blocks ? pl->m : pl->n_tiles, nb = blocks ? pl->P * pl->B : pl->P;
      fvm::FusedRhs fr;
      fr.rs = pl->rs; fr.sep = pl->sep; fr.nt = pl->n_tiles; fr.B = blocks ? pl->B : 1;
      fr.g_first = pl->g_first; fr.packet = blocks ? pl->packet : nullptr;
      double* sol = blocks ? pl->gr : pl->xs;
      if (n <= 1024)
        fvm::reduced_solve_kernel<128, 8, true><<<(unsigned)nb, 128, 0, (cudaStream_t)stream>>>(n, pl->rp, pl->rq, pl->rinv,
                                                                                             pl->rr, sol, fr);
      else
        fvm::reduced_solve_kernel<1024, 8, true><<<(unsigned)nb, 1024, 0, (cudaStream_t)stream>>>(n, pl->rp, pl->rq,
                                                                                               pl->rinv, pl->rr, sol, fr);
      st = (int)cudaGetLastError();
      if (st) return st;
    }
    if (blocks) {
      const double* allp = pl->packet;
      if (pl->world > 1) {
        const int rc = allgather(pl->packet, pl->allp, (size_t)(pl->P * pl->B * 4), 8 /* ncclDouble */, comm, stream);
        if (rc != 0) return FVM_ERR_COLLECTIVE;
        allp = pl->allp;
      }
      st = fvm_spike_rank_top(pl->P, pl->B, pl->m, pl->world, pl->rank, allp, pl->tp, pl->tq, pl->tinv, pl->cst, pl->gr,
                              pl->Vr, pl->Wr, pl->xs, stream);
      if (st) return st;
    }
    if (pl->lazy) {
      pending = 1;
    } else {
      st = fvm_spike_finalize(pl->n_tiles, pl->P, pl->config, pl->extents, xs_off, xs_n, pl->Vt, pl->Wt, pl->xs, w_out,
                              pl->w_stride, stream);
      if (st) return st;
    }
    cur = 1 - cur;
  }
  return FVM_OK;
}

// ------------------------------------------------------------------------------------------
// fvm_spike_graph_*: two lazy steps of fvm_spike_steps captured once and replayed (see the header).
struct fvm_spike_graph {
  cudaGraph_t graph;
  cudaGraphExec_t exec;
};

extern "C" int fvm_spike_graph_create(const fvm_spike_plan* pl, double* w0, double* w1, void* stream,
                                      fvm_spike_graph** out) {
  if (!pl || !w0 || !w1 || !out || !stream) return FVM_ERR_INVALID_ARG;     // the legacy default stream cannot capture
  if (pl->world != 1 || !pl->lazy) return FVM_ERR_UNSUPPORTED;
  *out = nullptr;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
  if (e != cudaSuccess) return (int)e;
  const int st = fvm_spike_steps(pl, w0, w1, 0, 1, 2, nullptr, nullptr, stream);
  cudaGraph_t graph = nullptr;
  e = cudaStreamEndCapture(s, &graph);                   // always end the capture, also after a failed step
  if (st) { if (graph) cudaGraphDestroy(graph); return st; }
  if (e != cudaSuccess) return (int)e;
  cudaGraphExec_t exec = nullptr;
  e = cudaGraphInstantiate(&exec, graph, 0);
  if (e != cudaSuccess) { cudaGraphDestroy(graph); return (int)e; }
  fvm_spike_graph* g = new (std::nothrow) fvm_spike_graph;
  if (!g) { cudaGraphExecDestroy(exec); cudaGraphDestroy(graph); return (int)cudaErrorMemoryAllocation; }
  g->graph = graph; g->exec = exec;
  *out = g;
  return FVM_OK;
}

extern "C" int fvm_spike_graph_run(fvm_spike_graph* g, int64_t n_pairs, void* stream) {
  if (!g || n_pairs < 0) return FVM_ERR_INVALID_ARG;
  for (int64_t i = 0; i < n_pairs; ++i) {
    const cudaError_t e = cudaGraphLaunch(g->exec, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
  }
  return FVM_OK;
}

extern "C" void fvm_spike_graph_destroy(fvm_spike_graph* g) {
  if (!g) return;
  cudaGraphExecDestroy(g->exec);
  cudaGraphDestroy(g->graph);
  delete g;
}
