/* fvm_b200 -- C-ABI of the B200-native FVM theta-step hot path.
 *
 * The reference (danieljfarrell/FVM) is pure Python: it has no FFI of its own.
 * Each entry point below replaces a piece of the reference's Python/scipy path
 * (cited as fvm_classes.py:LINE) and is what a binding for that path would call;
 * INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer to fp64 data unless stated otherwise;
 *  - the caller owns every buffer; the library allocates nothing, keeps no state
 *    between calls and reads no environment variable (two exceptions, stated at their
 *    declarations: fvm_spike_local_static works in a fixed device array inside the
 *    library image, whose contents persist from call to call, and the handle of
 *    fvm_spike_graph_create owns a CUDA graph until it is destroyed); all work is
 *    enqueued on `stream` (a CUstream / cudaStream_t handle passed as void*;
 *    NULL = the legacy default stream) and no call synchronises the host;
 *  - strided arrays: element (problem p, cell j) of X lives at X[p*X_ps + j*X_cs]
 *    (strides in elements).  Problem-major [P][J] is (ps=J, cs=1); interleaved
 *    [J][P] is (ps=1, cs=P); a per-problem scalar is (ps=1, cs=0); one array
 *    shared by all problems is (ps=0, cs=1);
 *  - return value: 0 = FVM_OK, < 0 = FVM_ERR_* (bad argument, nothing was
 *    launched), > 0 = a cudaError_t from the launch.  Nothing throws.
 */
#ifndef FVM_B200_H
#define FVM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FVM_OK 0
#define FVM_ERR_INVALID_ARG (-1)  /* NULL pointer, J < 3, P < 1, theta outside (0, 1], bad enum */
#define FVM_ERR_UNSUPPORTED (-2)  /* size outside what this entry point handles               */
#define FVM_ERR_SCRATCH_LIMIT (-3) /* fvm_spike_local_static: P*n_tiles exceeds the built-in scratch */
#define FVM_ERR_COLLECTIVE (-4)   /* fvm_spike_steps: the caller's all-gather returned an error    */

/* discretisation of the advective flux [fvm_classes.py:116-128] */
#define FVM_CENTRAL 0
#define FVM_UPWIND 1
#define FVM_EXPONENTIAL 2

/* which row an end of the mesh gets [fvm_classes.py:156-177 Robin, :191-216 Dirichlet] */
#define FVM_BC_DIRICHLET 0
#define FVM_BC_ROBIN 1

/* One batch of P advection-diffusion problems on ONE shared mesh.
 * Mirrors the constructor + set_boundary_conditions of AdvectionDiffusionModel
 * [fvm_classes.py:92-98,141-147].  Host struct holding device pointers. */
typedef struct fvm_problem {
  int64_t J;              /* cells (>= 3: the reference fails below that, fvm_classes.py:292) */
  int64_t P;              /* independent problems                                            */
  const double* faces;    /* [J+1] face positions, shared by all problems [:38]               */
  const double* a;        /* advection velocity  a[p*a_ps + j*a_cs]        [:96]              */
  int64_t a_ps, a_cs;
  const double* d;        /* diffusion coefficient (unmodified)            [:97]              */
  int64_t d_ps, d_cs;
  const double* reaction; /* optional decay rate r >= 0 (dw/dt = ... - r w); NULL = none.
                             Not in the reference: -r is added to the diagonal
                             of M on rows that are not Dirichlet rows.                          */
  int64_t r_ps, r_cs;
  int32_t scheme;         /* FVM_CENTRAL | FVM_UPWIND | FVM_EXPONENTIAL                       */
  int32_t left_kind;      /* FVM_BC_DIRICHLET | FVM_BC_ROBIN: row 0 of M                      */
  int32_t right_kind;     /* ... row J-1 of M                                                 */
  int32_t alpha_left;     /* end entries (0 or 1) of the alpha mask [:237-241].  The reference */
  int32_t alpha_right;    /* sets BOTH from `left_flux is None`; the host layer reproduces that */
  const double* left_flux;  /* [P] Robin flux at the left end  (beta[0]   =  flux/h[0],   :182); NULL = 0 */
  const double* right_flux; /* [P] Robin flux at the right end (beta[J-1] = -flux/h[J-1], :188); NULL = 0 */
} fvm_problem;

/* Library version (major*10000 + minor*100 + patch). */
int fvm_version(void);

/* Text for a status code returned by any entry point (static storage). */
const char* fvm_status_string(int status);

/* kappa, modified diffusivity D, and the three diagonals of M = coefficient_matrix().
 * Replaces AdvectionDiffusionModel.__init__ (kappa, D = d + a h kappa / 2,
 * fvm_classes.py:116-131), _interior_matrix_elements (:149-154), the Robin and
 * Dirichlet row builders (:156-231), beta_vector (:244-263) and coefficient_matrix
 * (:265-310).  M[j][j-1] = lo, M[j][j] = di, M[j][j+1] = up (lo at j=0 and up at
 * j=J-1 are written as 0).  central/upwind results are bit-identical to the
 * reference's numpy arithmetic; the exponential scheme evaluates
 * D = d (B(mu) + mu/2), B = mu/expm1(mu) (cancellation- and overflow-safe; the
 * reference's formula, :117, returns NaN for mu >= 709.8).
 * lo/di/up/kappa/D use strides (out_ps, out_cs); kappa and D may be NULL.
 * beta_left/beta_right: [P], beta[0] and beta[J-1] (beta is 0 elsewhere); may be NULL. */
int fvm_assemble(const fvm_problem* prob, double* lo, double* di, double* up, double* kappa,
                 double* D, int64_t out_ps, int64_t out_cs, double* beta_left,
                 double* beta_right, void* stream);

/* Factor the theta-method operators once per model.  Replaces
 *   A = I - k*theta*alpha*M ;  B = I + k*(1-theta)*alpha*M      [fvm_classes.py:373-377]
 * and the per-step SuperLU factorisation inside spsolve (:411).
 * In : M diagonals (strides m_ps, m_cs), dt[P] (the reference's `k`), theta in (0, 1],
 *      alpha end entries, beta_left/right [P] (may be NULL = 0).
 * Out: fp, fq, fb (strides f_ps, f_cs), bt_left/bt_right [P] -- the un-normalised LU factors
 *      of A, which carry both operators because B = I/theta - c A with c = (1-theta)/theta:
 *        pivot_j = A_jj - fp_j * A_j-1,j ;  fp_j = l_j = A_j,j-1 / pivot_{j-1} ;
 *        fq_j = s_j = A_j,j+1 / pivot_j ;   fb_j = kappa_j = 1 / (theta * pivot_j) ;
 *        bt = theta * beta at the two ends.
 *      The pivot recurrence is evaluated in double-double and each factor is rounded once:
 *      the step kernels apply the same factors at every step, so chain round-off would
 *      otherwise grow linearly with the number of steps. */
int fvm_theta_factor(int64_t J, int64_t P, const double* lo, const double* di, const double* up,
                     int64_t m_ps, int64_t m_cs, const double* dt, double theta,
                     int32_t alpha_left, int32_t alpha_right, const double* beta_left,
                     const double* beta_right, double* fp, double* fq, double* fb,
                     int64_t f_ps, int64_t f_cs, double* bt_left, double* bt_right, void* stream);

/* Variant 1: `nsteps` theta steps, one thread per system, state updated IN PLACE.
 * Replaces the loop body  rhs = B*w + beta ; w = spsolve(A, rhs)  [fvm_classes.py:408-411],
 * evaluated as  w' = (1/theta) A^-1 (w + theta*beta) - c w  (no right-hand side is formed).
 * Streams fp/fq/fb and w from HBM every step; use the interleaved layout (ps=1, cs=P)
 * for both so that a warp's accesses coalesce.  Any J >= 3. */
int fvm_theta_step_thomas(int64_t J, int64_t P, const double* fp, const double* fq,
                          const double* fb, int64_t f_ps, int64_t f_cs, double theta,
                          const double* bt_left, const double* bt_right, double* w, int64_t w_ps,
                          int64_t w_cs, int64_t nsteps, void* stream);

/* Variant 1 with TIME-VARYING boundary fluxes.  No counterpart in the reference, whose beta is
 * constant in time; fvm_classes.py:221,228 ("unless BCs are time varying") only hints at it
 * (SURVEY.md 8(f)2).  Step s (0-based) uses  beta_left * scale_left[s*P + p]  and
 * beta_right * scale_right[s*P + p]:  beta is linear in the boundary flux [fvm_classes.py:
 * 244-263], so with bt_left/bt_right factored for a UNIT flux the scales are the flux values at
 * step s.  A NULL scale array leaves that end constant (scale 1).  Everything else as in
 * fvm_theta_step_thomas. */
int fvm_theta_step_thomas_tv(int64_t J, int64_t P, const double* fp, const double* fq,
                             const double* fb, int64_t f_ps, int64_t f_cs, double theta,
                             const double* bt_left, const double* bt_right,
                             const double* scale_left, const double* scale_right, double* w,
                             int64_t w_ps, int64_t w_cs, int64_t nsteps, void* stream);

/* theta = 0 (explicit Euler, A = I): one step  w_out = (I + dt*alpha*M) w_in + beta  of the same
 * loop [fvm_classes.py:408-411 with theta = 0 at :376-377].  M diagonals from fvm_assemble
 * (strides m_ps, m_cs); w_in and w_out must be different buffers with strides (w_ps, w_cs);
 * beta_left/right [P] may be NULL. */
int fvm_explicit_step(int64_t J, int64_t P, const double* lo, const double* di, const double* up,
                      int64_t m_ps, int64_t m_cs, const double* dt, int32_t alpha_left,
                      int32_t alpha_right, const double* beta_left, const double* beta_right,
                      const double* w_in, double* w_out, int64_t w_ps, int64_t w_cs, void* stream);

/* Diagnostics of a batch, out[p*5 + 0..4] = min and max over the cells of the Peclet number
 * a h / d (on the unmodified d, as AdvectionDiffusionModel.__init__ checks it,
 * fvm_classes.py:103-107,135-136), min and max of the CFL number a dt / h (:110-114,138-139;
 * dt may be NULL: 0), and the "area" sum_j w_j h_j the demo prints every step (:420-421; w may be
 * NULL: 0).  w is strided like every state array. */
int fvm_diagnostics(const fvm_problem* prob, const double* dt, const double* w, int64_t w_ps,
                    int64_t w_cs, double* out, void* stream);

/* Mesh validation for meshes too large for the reference's host-side duplicate check
 * (Mesh.__init__, fvm_classes.py:36-37): adds to counts[0] the number of zero-width cells
 * (adjacent duplicate faces) and to counts[1] the number of negative-width cells (unsorted faces,
 * which the reference accepts) among the n_cells cells of faces[0..n_cells].  counts: two device
 * uint64, cleared by the caller. */
int fvm_mesh_check(const double* faces, int64_t n_cells, uint64_t* counts, void* stream);

/* Distance between two states a and b of a batch on one mesh (steady-state detection: w_n
 * against an earlier w; comparison with an analytic solution, fvm_examples.py:71).
 * out[p*4 + 0..3] = max_j |a_j - b_j| (NaN if any difference is NaN), sum_j |a_j - b_j| h_j,
 * max_j |b_j|, sum_j b_j h_j.  a and b strided like every state array. */
int fvm_state_compare(int64_t J, int64_t P, const double* faces, const double* a, int64_t a_ps,
                      int64_t a_cs, const double* b, int64_t b_ps, int64_t b_cs, double* out,
                      void* stream);

/* Largest J the fused on-chip variant accepts. */
int64_t fvm_fused_max_cells(void);

/* Variant 2: assembly + factorisation + `nsteps` theta steps in ONE launch with the state in
 * registers for the whole launch: one warp per system up to 1024 cells (coefficients in shared
 * or tensor memory), one CTA per system up to 8192 cells (coefficients in shared memory); HBM is
 * touched when the state is loaded, at checkpoints, and when it is stored.  Replaces the whole
 * of fvm_classes.py:367-377,408-411 for a batch.  J <= fvm_fused_max_cells().
 * w: [P][J] problem-major, in/out.  dt: [P].
 * checkpoint_every > 0: after every that many steps the state is also written to
 * ckpt[(n-1)*P*J + p*J + j], n = 1..nsteps/checkpoint_every (ckpt may be NULL: none). */
int fvm_theta_run_fused(const fvm_problem* prob, const double* dt, double theta, int64_t nsteps,
                        int64_t checkpoint_every, double* w, double* ckpt, void* stream);

/* Variant 2 with what changes from step to step.  NO COUNTERPART IN THE REFERENCE,
 * whose beta and k are constants of a model; fvm_classes.py:221,228 ("unless BCs are time varying")
 * only hint at it.  Same on-chip kernels and arguments as fvm_theta_run_fused, for J <=
 * fvm_fused_ex_max_cells() (1024: the warp-per-system kernels), plus:
 *  - scale_left / scale_right (may be NULL): step s (0-based) uses the boundary flux
 *    prob->left_flux[p] * scale_left[s*scale_ss + p*scale_ps]  (likewise on the right).  beta is
 *    linear in the flux [fvm_classes.py:244-263], so with a unit flux in prob the scales ARE the
 *    flux history; scale_ps = 0 shares one history between all problems.  A scaled end must be a
 *    Robin end with a non-NULL flux array in prob.
 *  - dt_hist, run_starts, n_runs (all NULL/0, or all given): the nsteps steps are cut into n_runs
 *    runs [run_starts[r], run_starts[r+1]) (run_starts: DEVICE array of n_runs+1 int64, first 0, last
 *    nsteps); all steps of run r use the time step dt_hist[run_starts[r]*dt_ss + p*dt_ps] and the
 *    kernel re-factors the operators at the start of every run, inside the launch (`dt` is then
 *    unused but must be valid).  Checkpoints count steps over the whole launch. */
int64_t fvm_fused_ex_max_cells(void);
int fvm_theta_run_fused_ex(const fvm_problem* prob, const double* dt, double theta, int64_t nsteps,
                           int64_t checkpoint_every, double* w, double* ckpt,
                           const double* scale_left, const double* scale_right, int64_t scale_ps,
                           int64_t scale_ss, const double* dt_hist, int64_t dt_ps, int64_t dt_ss,
                           const int64_t* run_starts, int64_t n_runs, void* stream);

/* ---- Variant 3: partitioned SPIKE/Thomas for LONG systems (one huge mesh, or a batch of long
 * ones), optionally with the cells split in contiguous slabs over several GPUs.  Replaces the
 * same loop (fvm_classes.py:373-377,408-411) where J is far beyond what fits on chip; the
 * reference itself cannot assemble such sizes (.todok() at :303).
 *
 * A slab of n_tiles tiles of T = fvm_spike_tile_cells(config) cells starts at global tile
 * `tile_offset`.  The last cell of a tile is a separator; the separators of ALL tiles of ALL
 * slabs form one tridiagonal "reduced" system (n_global unknowns per problem) that every rank
 * solves redundantly after an all-gather of its right-hand side (the host layer owns the NCCL
 * call).  A step has no right-hand side: with B = I/theta - cA, c = (1-theta)/theta, it is
 * w' = A^-1(w/theta + beta) - c w, so no cell needs its neighbours' old values, a tile needs nothing
 * from the tiles next to it and the state carries no ghost cells: `w` points at the slab's first
 * cell and w_stride >= n_tiles*T separates problems.
 * Work arrays (all caller-owned): p, q, b, Vt, Wt [P][n_tiles*T]; sep [P][n_tiles][6];
 * bends [P][2]; rs [P][n_tiles][2]; rr [P][n_tiles]; g_first [P]; xs, rp, rq, rinv [P][n_global]. */

/* Cells per tile of a tile configuration, 0 if the configuration does not exist.
 *   0: 8192 cells (512 threads x 16 cells, one CTA per SM)
 *   1: 256 cells (small tiles: tests reach many tiles on small meshes)
 *   2: 4096 cells (256 threads x 16 cells, two CTAs per SM: the scans of one tile overlap the loads
 *      of the other; 16 % faster per step on cfg 5 and the host layer's default) */
int64_t fvm_spike_tile_cells(int32_t config);

/* Threads TPB of the CTA that owns a tile.  Every per-cell array of variant 3 (p, q, b, Vt, Wt
 * and the state w) is stored TILED: cell r of a tile (thread r/CPT, its cell r%CPT, with
 * CPT = tile_cells/TPB) sits at position (r % CPT)*TPB + r/CPT of the tile, so that the CTA's
 * loads coalesce.  Callers convert their state once on the way in and once on the way out. */
int64_t fvm_spike_tile_threads(int32_t config);

/* Once per model: rows of A = I - dt*theta*alpha*M for the slab (prob->faces / a / d are indexed
 * by GLOBAL cell, so pass base pointers shifted by the slab offset), tile-local un-normalised LU
 * restarted at every tile's first cell (p = l, q = s, b = kappa as fvm_theta_factor defines them; all
 * three 0 at the separator), tile spikes (Vt, Wt; 0 and -1 at the separator), per-tile separator
 * rows and reduced-system contributions (sep: lo, up, RL, RD less the next tile's term, V0, W0),
 * theta*beta of the mesh's two end cells (bends) and, in extents[0..1] (device int32), how many
 * cells from the left/right tile edge carry a spike larger than spike_tol.  Every elimination chain
 * runs in double-double and every stored number is rounded once. */
int fvm_spike_setup(const fvm_problem* prob, const double* dt, double theta, int64_t tile_offset,
                    int64_t n_tiles, int32_t config, double* p, double* q, double* b, double* Vt,
                    double* Wt, double* sep, double* bends, int32_t* extents, double spike_tol,
                    void* stream);

/* Once per model: Thomas factors (rp, rq, rinv) of the reduced tridiagonal (RL, RD, RU), n
 * unknowns per problem (pivot chain in double-double). */
int fvm_spike_reduced_factor(int64_t n, int64_t P, const double* RL, const double* RD,
                             const double* RU, double* rp, double* rq, double* rinv, void* stream);

/* Step, part 1: the tile-local solves g of A g = w/theta + beta (w_in -> w_out = z = g - c w, the new
 * state up to the spike correction of part 3), the reduced right-hand side rr (the last tile of the
 * slab leaves out the term that needs the next slab's first value) and g_first = g at the slab's
 * first cell.  rs is scratch between the call's two kernels.
 * has_first_cell: the slab holds global cell 0; last_cell_local: local index of global cell J-1
 * if this slab holds it and it carries a nonzero beta, else -1. */
int fvm_spike_local(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                    const double* b, const double* sep, const double* bends,
                    int32_t has_first_cell, int64_t last_cell_local, double theta,
                    const double* w_in, double* w_out, int64_t w_stride, double* rs, double* rr,
                    int64_t rr_stride, double* g_first, void* stream);

/* Step, part 2: solve the factored reduced system, rr -> xs (the separators' values of
 * y = A^-1(w/theta + beta)), n_global unknowns per problem. */
int fvm_spike_reduce_solve(int64_t n_global, int64_t P, const double* rp, const double* rq,
                           const double* rinv, const double* rr, double* xs, void* stream);

/* Step, part 3: w' = z - Vt*xs(k-1) - Wt*xs(k) over the spike extents (extents_host: the two
 * HOST integers read back from fvm_spike_setup), the separators included (Vt = 0, Wt = -1 there).
 * In place on w, and NOT idempotent: exactly once per fvm_spike_local. */
int fvm_spike_finalize(int64_t n_tiles, int64_t P, int32_t config, const int32_t* extents_host,
                       int64_t tile_offset, int64_t n_tiles_global, const double* Vt,
                       const double* Wt, const double* xs, double* w, int64_t w_stride, void* stream);

/* Lazy correction for time loops: fvm_spike_local_lazy in place of fvm_spike_local +
 * fvm_spike_finalize.  The state then holds the UNCORRECTED z between steps, and the next step's tile
 * solve applies  w = z - Vt xs(k-1) - Wt xs(k)  while it loads the tile (xs = the previous step's
 * separator values, indexed like in fvm_spike_finalize; it must stay untouched until then).  This
 * saves the separate correction pass: 16 B instead of 32 B of HBM traffic per corrected cell.  The
 * FIRST step after a true state was stored uses fvm_spike_local; fvm_spike_finalize turns the buffer
 * into the true state whenever it is needed. */
int fvm_spike_local_lazy(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                         const double* b, const double* sep, const double* bends,
                         int32_t has_first_cell, int64_t last_cell_local, double theta,
                         const double* w_in, double* w_out, int64_t w_stride, double* rs, double* rr,
                         int64_t rr_stride, double* g_first, const double* Vt, const double* Wt,
                         const double* xs, int64_t tile_offset, int64_t n_tiles_global,
                         const int32_t* extents_host, void* stream);

/* SPIKE once more, over the separator system (no counterpart in the reference).  The reduced
 * tridiagonal is cut into blocks of m consecutive unknowns, B blocks per rank (world*B <= 1024 in all),
 * the last unknown of a block being the block's separator.  The caller factors the blocks with their
 * couplings cut and the separator rows replaced by identity (fvm_spike_reduced_factor with n = m and the
 * blocks as the batch), solves them every step with fvm_spike_reduce_solve (n = m, batch P*B: one CTA
 * per block instead of one CTA for the whole system), then:
 *  fvm_spike_rank_pack: packet[P][B][4] = what each block contributes to the per-step all-gather
 *    (first and last interior value of the block solution gr, its separator's right-hand side, the
 *    slab's first tile value);
 *  fvm_spike_rank_top: from the gathered packets allp[world][P][B][4], the Thomas factors
 *    tp, tq, tinv [P][world*B] of the block-separator system and the constants cst[P][3][world*B]
 *    (rows: coupling of a rank's last separator to the next rank's first tile value, 0 for the other
 *    blocks; lo, up of the block separators' rows) it writes this rank's separator values
 *    xs_ext[P][B*m+1] (entry 0 = the previous rank's last separator).
 * On one GPU (world = 1) there is no gather and `allp` is the packet itself. */
int fvm_spike_rank_pack(int64_t P, int64_t B, int64_t m, const double* gr, const double* g_first,
                        double* packet, void* stream);
int fvm_spike_rank_top(int64_t P, int64_t B, int64_t m, int32_t world, int32_t rank, const double* allp,
                       const double* tp, const double* tq, const double* tinv, const double* cst,
                       const double* gr, const double* Vr, const double* Wr, double* xs_ext, void* stream);

/* A whole time loop of variant 3 in ONE host call: nsteps theta steps, each the sequence
 *   fvm_spike_local(_lazy) -> fvm_spike_reduce_solve (blocks) -> fvm_spike_rank_pack -> all-gather ->
 *   fvm_spike_rank_top [-> fvm_spike_finalize when lazy == 0]
 * enqueued on `stream` with nothing allocated and no host synchronisation (without blocks -- B == 1 and
 * world == 1 -- a step is fvm_spike_local(_lazy) -> fvm_spike_reduce_solve [-> fvm_spike_finalize]).  Replaces
 * the body of the reference's loop (fvm_classes.py:408-411) for long systems; the per-step host cost of the
 * separate entry points (five calls) is what bounds meshes of 10^5..10^6 cells, whose step is a few
 * microseconds of GPU time.
 * The all-gather across GPUs is the caller's: `allgather` has the signature of ncclAllGather
 * (sendbuff, recvbuff, sendcount, datatype, comm, stream) and is called with datatype 8 (ncclDouble /
 * ncclFloat64) -- pass the address of ncclAllGather and an ncclComm_t whose ranks are the slabs in order;
 * the library itself does not link NCCL.  NULL when world == 1.
 * The state ping-pongs between w0 and w1: step i reads buffer (cur + i) % 2 and writes the other.
 * `pending` != 0 says the buffer read first holds an uncorrected state left by a previous lazy call. */
typedef int (*fvm_allgather_fn)(const void* sendbuff, void* recvbuff, size_t sendcount, int datatype,
                                void* comm, void* stream);

typedef struct fvm_spike_plan {
  int64_t n_tiles, P;             /* this rank's tiles, problems                                     */
  int32_t config, lazy;           /* tile configuration; lazy != 0: correction inside the next solve  */
  int32_t has_first_cell;         /* as in fvm_spike_local                                            */
  int32_t world, rank;            /* ranks the cells are split over, this rank                        */
  int32_t extents[2];             /* spike extents (host values read back from fvm_spike_setup)       */
  int64_t last_cell_local;        /* as in fvm_spike_local                                            */
  int64_t B, m;                   /* blocks of the separator system per rank, unknowns per block;
                                     B == 1 && world == 1: no block level, xs has n_tiles entries     */
  double theta;
  int64_t w_stride;
  const double *p, *q, *b, *Vt, *Wt, *sep, *bends;     /* from fvm_spike_setup                        */
  const double *rp, *rq, *rinv;   /* factors of the blocks (of the whole separator system without blocks) */
  const double *Vr, *Wr;          /* block spikes [P][n_tiles]                  (block level only)     */
  const double *tp, *tq, *tinv, *cst;   /* top system, see fvm_spike_rank_top   (block level only)     */
  double *rs, *rr, *g_first;      /* work arrays of fvm_spike_local                                    */
  double *gr, *packet, *allp;     /* [P][n_tiles], [P][B][4], [world][P][B][4]  (block level only)     */
  double *xs;                     /* separator values: [P][n_tiles+1] with blocks (entry 0 = previous
                                     rank's last separator), else [P][n_tiles]                         */
} fvm_spike_plan;

int fvm_spike_steps(const fvm_spike_plan* plan, double* w0, double* w1, int32_t cur, int32_t pending,
                    int64_t nsteps, fvm_allgather_fn allgather, void* comm, void* stream);

/* The same loop as a CUDA graph, for meshes whose step is launch-bound (10^5..10^6 cells: a few
 * microseconds of GPU time in three to five launches).  fvm_spike_graph_create captures TWO consecutive lazy
 * steps (w0 -> w1 -> w0) of fvm_spike_steps on `stream` -- which must not be the legacy default stream -- and
 * instantiates them; fvm_spike_graph_run replays the pair n_pairs times.  Requirements: plan->world == 1,
 * plan->lazy != 0, and w0 must hold a PENDING (uncorrected) state when the graph runs, i.e. at least one
 * fvm_spike_steps call came first and left the state in w0.  The handle owns host memory and a CUDA graph:
 * this and fvm_spike_local_static are the two places where the library keeps something between calls;
 * fvm_spike_graph_destroy releases it.  A plan, its arrays and w0, w1 must outlive the handle unchanged. */
typedef struct fvm_spike_graph fvm_spike_graph;
int fvm_spike_graph_create(const fvm_spike_plan* plan, double* w0, double* w1, void* stream,
                           fvm_spike_graph** out);
int fvm_spike_graph_run(fvm_spike_graph* graph, int64_t n_pairs, void* stream);
void fvm_spike_graph_destroy(fvm_spike_graph* graph);

/* Largest P * n_tiles fvm_spike_local_static accepts (32768 tiles, two doubles each = 512 KiB of
 * scratch). */
int64_t fvm_spike_static_max_tiles(void);

/* fvm_spike_local without the `rs` work array.  NO COUNTERPART IN THE REFERENCE: it exists for
 * callers of this library that cannot allocate device memory for a step -- an application that
 * gave all of the GPU's memory to its own pool at start-up, or a time loop in a latency-critical
 * section where cudaMalloc (which synchronises the device) must not be called.  The scratch that
 * fvm_spike_local takes from the caller lives in a fixed device array inside the library
 * instead: 512 KiB that every GPU which loads the library holds until the process exits.
 *  - Same arguments, results and status codes as fvm_spike_local, minus `rs`; allocates nothing.
 *  - P * n_tiles must not exceed fvm_spike_static_max_tiles(); beyond that the call returns
 *    FVM_ERR_SCRATCH_LIMIT and launches nothing (use fvm_spike_local with your own `rs`).
 *  - Every call looks up the device address of the built-in array for the CURRENT device, which
 *    must therefore be the device `stream` belongs to (cudaSetDevice before the call).  Nothing is
 *    remembered between calls, so cudaDeviceReset() or a second context on the device is harmless.
 *    The first call on a device makes the CUDA runtime load the library's module there: make that
 *    first call before any stream capture begins.
 *  - All calls on a device share the one array, and it keeps what the last call wrote.  Nothing in
 *    it needs clearing (each call overwrites the part it uses before reading it), but two calls
 *    whose kernels overlap on the GPU -- different streams, or several host threads -- race on it:
 *    keep calls on one device in stream order. */
int fvm_spike_local_static(int64_t n_tiles, int64_t P, int32_t config, const double* p, const double* q,
                           const double* b, const double* sep, const double* bends,
                           int32_t has_first_cell, int64_t last_cell_local, double theta,
                           const double* w_in, double* w_out, int64_t w_stride, double* rr,
                           int64_t rr_stride, double* g_first, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FVM_B200_H */
