/* nsb200.h -- C ABI of the B200-native Krylov solve path (libnsb200.so).
 *
 * The reference (matschiner/navier-stokes-solver) is pure Python over NGSolve and has no FFI of its
 * own; its boundary for this path is the Python signature of each solver plus NGSolve's duck-typed
 * operator/vector protocol.  This header is the C boundary a binding for that path
 * has to bind: every entry point names the reference interface (file:line under the reference tree)
 * it replaces.  The Python mirror of the reference signatures lives in
 * navier-stokes-solver_b200/{la,solvers}.py and calls these functions through ctypes.
 *
 * Conventions
 *   - all functions return NSB_OK (0) or a negative error code; nsb_last_error() gives the text
 *     of the last failure on the calling thread.  No C++ exception crosses this boundary.
 *   - "const double* / double*" vector arguments may be HOST or DEVICE pointers (resolved through
 *     unified virtual addressing).  Host buffers are copied in at the start of a call and results
 *     are copied back into the same caller buffer before the call returns (the reference mutates
 *     `sol` in place: minres.py:50,118; bramble_pasciak_cg.py:131; solvers/bramblepasciak_new.py:137,228).
 *   - matrices are fp64 CSR with int32 column indices; row pointers may be given as int32 or int64.
 *   - one CUDA stream per context; a context is used by one host thread at a time (the reference's
 *     solvers are single-caller, SURVEY.md 8(b) "Threading").
 *   - distributed contexts (one process per GPU) row-partition every operator; vectors passed to the
 *     solvers then hold the rank's OWNED rows only.
 */
#ifndef NSB200_H
#define NSB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NSB_API __attribute__((visibility("default")))
#else
#define NSB_API
#endif

#define NSB_OK 0
#define NSB_ERR_INVALID -1   /* bad argument */
#define NSB_ERR_CUDA -2      /* CUDA runtime failure */
#define NSB_ERR_NCCL -3      /* NCCL failure / NCCL not loadable */
#define NSB_ERR_BREAKDOWN -4 /* non-finite or zero pivot in a recurrence */
#define NSB_ERR_UNSUPPORTED -5

/* solver exit status (written to *status) */
#define NSB_CONVERGED 0        /* left through the relative-tolerance break */
#define NSB_MAXSTEPS 1         /* loop condition ended the iteration: step cap  (minres.py:96,145; bramble_pasciak_cg.py:144; bramblepasciak_new.py:248) */
#define NSB_ABSTOL 2           /* MinRes only: loop condition ended on the absolute test ResNorm > tol (minres.py:96) */
#define NSB_ZERO_RESIDUAL 3    /* BramblePasciakCG early return when <w,d> == 0 (bramblepasciak_new.py:191-192) */
#define NSB_BREAKDOWN 4        /* a reduced scalar became non-finite; iteration stopped */

typedef struct nsb_ctx_s* nsb_ctx;
typedef struct nsb_mat_s* nsb_mat;
typedef struct nsb_pre_s* nsb_pre;
typedef struct nsb_halo_s* nsb_halo;
typedef struct nsb_cond_s* nsb_cond;
typedef struct nsb_simple_s* nsb_simple;

/* ---- library / context -------------------------------------------------------------------
 * Infrastructure with no counterpart in the reference (one Python process over NGSolve, no FFI): what SURVEY.md 8(b) "What a
 * C-ABI replacement must export" lists as items 1 (context: device, stream) and 5 (last-error string), plus the diagnostics
 * bench.py reads (launch counter, synchronisation statistics, CUDA-graph switch). */
NSB_API int nsb_version(void);
NSB_API const char* nsb_last_error(void);

/* Single-GPU context on CUDA device `device`. */
NSB_API int nsb_ctx_create(int device, nsb_ctx* out);
/* Row-partitioned context: rank `rank` of `nranks`, NCCL communicator bootstrapped from a 128-byte
 * unique id obtained on rank 0 with nsb_nccl_unique_id() and broadcast by the caller
 * (torch.distributed in the Python host).  No reference counterpart (the reference is
 * single-process: SURVEY.md 2, 8(e)). */
NSB_API int nsb_nccl_unique_id(char out[128]);
NSB_API int nsb_ctx_create_dist(int device, int rank, int nranks, const char unique_id[128], nsb_ctx* out);
NSB_API int nsb_ctx_destroy(nsb_ctx ctx);
NSB_API int nsb_ctx_sync(nsb_ctx ctx);
/* raw cudaStream_t of the context (for callers that order their own device work against it) */
NSB_API void* nsb_ctx_stream(nsb_ctx ctx);
/* number of kernels this context has launched so far (bench.py's gpu_launches claim) */
NSB_API int64_t nsb_ctx_launch_count(nsb_ctx ctx);
/* diagnostic, row-partitioned contexts: nanoseconds the exchange kernels spent polling peer flags, and how often
 * ([0,1] scalar all-reduce, [2,3] halo push waiting for "consumed", [4,5] SpMV waiting for the halo, [6,7] wait/copy
 * kernel).  All zero unless the library was built with -DNSB_SYNC_STATS (tools/sync_stats.sh). */
NSB_API int nsb_ctx_sync_stats(nsb_ctx ctx, uint64_t out[16], int reset);
/* capture the per-iteration kernel sequence in CUDA graphs (default on for single-GPU contexts and for
 * row-partitioned contexts whose exchanges run over NVLink peer memory; NSB_P2P=0 forces the NCCL path) */
NSB_API int nsb_ctx_set_graphs(nsb_ctx ctx, int enabled);

/* ---- operators: replaces NGSolve SparseMatrix / BilinearForm.mat (run.py:106,
 *      stokes_hcurldiv.py:76, solvers/bramblepasciak_new.py:108-109) -------------------------- */
NSB_API int nsb_csr_create(nsb_ctx ctx, int64_t rows, int64_t cols, int64_t nnz, const void* rowptr,
                   int rowptr_bits /*32|64*/, const int32_t* colidx, const double* values, nsb_mat* out);
/* explicit transpose, built once on the device (matB.CreateTranspose(), bramblepasciak_new.py:198; b.T, run.py:45) */
NSB_API int nsb_csr_transpose(nsb_ctx ctx, nsb_mat a, nsb_mat* out);
/* K = [[A, BT], [B, C]] as ONE merged CSR so A and B^T are applied in a single pass
 * (BlockMatrix([[a, b.T], [b, None]]) run.py:45; bramble_pasciak_cg.py:77-78).  BT may be NULL (built by
 * transposition); C may be NULL. */
NSB_API int nsb_saddle_create(nsb_ctx ctx, nsb_mat A, nsb_mat B, nsb_mat BT, nsb_mat C, nsb_mat* out);
NSB_API int nsb_mat_destroy(nsb_mat m);
NSB_API int nsb_mat_info(nsb_mat m, int64_t* rows, int64_t* cols, int64_t* nnz);
/* (no reference counterpart: NGSolve hides its storage)  storage actually used by the SpMV kernels: rows_per_chunk > 0 when the SELL-LX companion (rows sorted by
 * length inside chunks, sliced-ELL pair-blocks, x staged in shared memory, 16-bit chunk-local column indices)
 * is active, else the CSR-vector kernel runs.  total_distinct = sum over chunks of distinct columns (4 B each
 * of extra index traffic per SpMV); stored_nnz = nonzeros including slice padding; lanes_per_row = lanes that
 * share a row in the active kernel; column_runs > 0 when the distinct-column lists are stored as that many runs
 * of consecutive columns (8 B per run instead of 4 B per column). */
NSB_API int nsb_mat_format(nsb_mat m, int* rows_per_chunk, int* chunks, int* max_distinct, int64_t* total_distinct,
                           int* lanes_per_row, int64_t* stored_nnz, int64_t* column_runs);
/* (no reference counterpart)  Row panels: rows [bounds[k], bounds[k+1]) get their own SELL-LX shape (chunk size, lanes per row).  For operators whose
 * row blocks differ strongly in length, e.g. the stacked local rows [[A, B^T], [B, 0]] of a row-partitioned saddle-point
 * operator (nsb_saddle_create does this by itself).  Rebuilds the SpMV storage. */
NSB_API int nsb_mat_set_row_panels(nsb_ctx ctx, nsb_mat m, int nbounds, const int64_t* bounds);
/* main diagonal of a square (or owned-square) operator, length rows: what Preconditioner(bf, 'local') inverts (run.py:62) */
NSB_API int nsb_mat_diagonal(nsb_ctx ctx, nsb_mat m, double* out);
/* y = M x   (the `mat * vec` primitive, e.g. minres.py:97).  x has `cols`, y has `rows` entries. */
NSB_API int nsb_spmv(nsb_ctx ctx, nsb_mat m, const double* x, double* y);
/* device-timed SpMV micro-benchmark on resident vectors: `reps` launches after `warmup`, average
 * milliseconds per launch by CUDA events on the context's stream. */
NSB_API int nsb_spmv_bench(nsb_ctx ctx, nsb_mat m, int warmup, int reps, double* ms_per_launch);

/* Row-partition support: attach the halo plan of a row-partitioned operator whose column indices
 * are renumbered [owned | halo].  send_idx lists, per neighbour, the owned entries to ship;
 * the halo tail is filled neighbour by neighbour in `neighbors` order.   */
NSB_API int nsb_halo_create(nsb_ctx ctx, int64_t n_owned, int64_t n_halo, int n_neighbors, const int32_t* neighbors,
                    const int64_t* send_counts, const int32_t* send_idx, const int64_t* recv_counts, nsb_halo* out);
NSB_API int nsb_halo_destroy(nsb_halo h);
NSB_API int nsb_mat_set_halo(nsb_mat m, nsb_halo h);

/* ---- preconditioners: replaces Preconditioner(bf, 'local') and block smoothers
 *      (run.py:61-62, stokes_hcurldiv.py:53, NavierStokesSIMPLE_iterative.py:199,360-373) -------- */
/* y = d .* x with caller-supplied d (already inverted; zero on constrained dofs): Preconditioner(bf, 'local') restricted to
 * FreeDofs (run.py:61-62), the diagonal taken on the host */
NSB_API int nsb_pre_create_diag(nsb_ctx ctx, int64_t n, const double* d, nsb_pre* out);
/* the same from the device operator's diagonal (run.py:62, stokes_hcurldiv.py:53); free_mask (n bytes, 1 = free: fes.FreeDofs()) may be NULL */
NSB_API int nsb_pre_create_jacobi(nsb_ctx ctx, nsb_mat m, const uint8_t* free_mask, nsb_pre* out);
/* BlockMatrix([[preA, None], [None, preS]]) (run.py:46).  Two diagonal preconditioners are merged into one diagonal; with a
 * block smoother / auxiliary-space composite in either block the result applies them block by block and borrows the two handles
 * (they must outlive it). */
NSB_API int nsb_pre_create_blockdiag2(nsb_ctx ctx, nsb_pre a, nsb_pre b, nsb_pre* out);
/* Block smoother: a.mat.CreateBlockSmoother(blocks) with the facet blocks of NavierStokesSIMPLE_iterative.py:360-373.
 * Disjoint dof blocks (at most 64 dofs each) in CSR-like form, listed colour by colour: blocks [color_ptr[c], color_ptr[c+1])
 * are mutually uncoupled through A, so one colour is one parallel launch and a sweep equals the sequential block
 * Gauss-Seidel sweep in the given block order.  ncolors == 0: additive use only.  As a preconditioner (nsb_pre_apply,
 * nsb_bpcg preA, nsb_cg pre) the handle acts as  jacobi * x  (gauss_seidel == 0; MypreA with GS = False without the
 * auxiliary-space term, :383)  or as  y = 0; Smooth(y, x); SmoothBack(y, x)  (gauss_seidel != 0; :375-381).
 * Row-partitioned context: A holds this rank's rows (columns [owned | halo]); the blocks list owned dofs only (local numbering; a
 * block is never cut by the partition) under a colouring of the GLOBAL block graph -- every rank passes the same ncolors, a colour
 * may be empty on a rank -- and a sweep exchanges the iterate's halo before every colour, which makes it the sequential sweep in the
 * global colour-major order.  (Row-partitioned use of the smoother and of the composite below has not run on a device: DESIGN.md 7.) */
NSB_API int nsb_pre_create_blocks(nsb_ctx ctx, nsb_mat A, int64_t nblocks, const int64_t* block_ptr, const int32_t* block_dofs,
                                  int ncolors, const int64_t* color_ptr, int gauss_seidel, nsb_pre* out);
/* jacobi.Smooth(y, x) (backward == 0) / jacobi.SmoothBack(y, x) (backward != 0): one block Gauss-Seidel sweep on y for the
 * right-hand side x (:377,:381) */
NSB_API int nsb_pre_smooth(nsb_ctx ctx, nsb_pre p, const double* x, double* y, int backward);
/* Auxiliary-space preconditioner MypreA (templates/NavierStokesSIMPLE_iterative.py:364-391):
 *   gauss_seidel != 0 (:375-381):  y = 0; jacobi.Smooth(y, x); t = x - A y; y += transform coarse transform^T t; jacobi.SmoothBack(y, x)
 *   gauss_seidel == 0 (:383):      y = (transform coarse transform^T + jacobi) x
 * smoother: a handle from nsb_pre_create_blocks on A (its matrix is the A of the residual); transform (n x n_c) and its
 * explicit transpose as device matrices; coarse: any preconditioner handle of size n_c (the reference puts component-wise
 * h1amg there, which is NGSolve-internal; a Jacobi or block smoother of the H1 matrix, or another composite, goes here).
 * The handles passed in are not owned and must outlive the result.
 * Row-partitioned context: transform = this rank's velocity rows with columns over a halo plan of the auxiliary space, transform_t =
 * this rank's auxiliary rows of the global transpose with columns over a halo plan of the velocity space, coarse of the owned
 * auxiliary size; the composite exchanges the input of each of its SpMVs. */
NSB_API int nsb_pre_create_auxspace(nsb_ctx ctx, nsb_pre smoother, nsb_mat transform, nsb_mat transform_t, nsb_pre coarse,
                                    int gauss_seidel, nsb_pre* out);
NSB_API int nsb_pre_destroy(nsb_pre p);
NSB_API int nsb_pre_apply(nsb_ctx ctx, nsb_pre p, const double* x, double* y);

/* ---- solvers ------------------------------------------------------------------------------- */
/* MinRes(mat, rhs, pre, sol, maxsteps, printrates, initialize, tol)   minres.py:12-149.
 * history: capacity maxsteps+1, history[0] = 1.0, one relative entry per iteration (minres.py:95,125).
 * iterations = number of entries written minus one.  pre may be NULL (minres.py:68,101). */
NSB_API int nsb_minres(nsb_ctx ctx, nsb_mat K, nsb_pre pre, const double* rhs, double* sol, int maxsteps, double tol,
               int initialize, double* history, int* iterations, double* seconds, int* status);

/* Bramble-Pasciak CG engine serving both reference variants:
 *   mode 2: BramblePasciakCG(blfA, blfB, matC, f, g, preA_unscaled, preM, sol, tol, maxsteps, printrates,
 *           initialize, rel_err)      solvers/bramblepasciak_new.py:24-253   (update, then test the previous error)
 *   mode 1: bramble_pasciak_cg(a, b, c, pre_a, pre_schur, upper_rhs, lower_rhs, solution, tolerance,
 *           max_steps, print_rates)   bramble_pasciak_cg.py:65-148            (test, then update)
 * k_scale is the factor k = 1/lambda_min + 1e-3 (bramble_pasciak_cg.py:71; bramblepasciak_new.py:118).
 * sol holds (u, p) contiguously.  history[i] = sqrt|<w,d>| before update i (capacity maxsteps+1);
 * updates = number of solution updates performed; err0 = history[0]. */
NSB_API int nsb_bpcg(nsb_ctx ctx, nsb_mat A, nsb_mat B, nsb_mat BT, nsb_pre preA, nsb_pre preM, double k_scale,
             const double* f, const double* g, double* sol, int maxsteps, double tol, int initialize, int rel_err,
             int mode, double* history, int* updates, double* seconds, int* status);

/* Statically condensed velocity operator of BramblePasciakCG (blfA.condense == True, bramblepasciak_new.py:8-21,84-103):
 *   matA = (I - HT)(A + inner_matrix)(I - H)     with A = blfA.mat (the condensed matrix), H = blfA.harmonic_extension,
 *   HT = blfA.harmonic_extension_trans, inner_solve = blfA.inner_solve, inner_matrix = blfA.inner_matrix, all exported as
 *   CSR of the full velocity size.  The handle borrows the five operators (they must outlive it).
 *   Row-partitioned context: each operator holds this rank's rows with columns numbered [owned | halo] over ONE velocity halo
 *   plan (nsb_mat_set_halo with the same nsb_halo on each; B of nsb_bpcg_condensed shares it); an operator that references owned
 *   columns only (element-local inner_solve / inner_matrix) may come with n_owned columns and no plan.  The chains exchange the
 *   halo of every SpMV input themselves.  (Row-partitioned use has not run on a device yet: DESIGN.md 7.) */
NSB_API int nsb_cond_create(nsb_ctx ctx, nsb_mat A, nsb_mat H, nsb_mat HT, nsb_mat inner_solve, nsb_mat inner_matrix, nsb_cond* out);
NSB_API int nsb_cond_destroy(nsb_cond k);
/* y = matA x  (myAmatrix.Mult, bramblepasciak_new.py:90-91) */
NSB_API int nsb_cond_apply(nsb_ctx ctx, nsb_cond k, const double* x, double* y);
/* BramblePasciakCG with blfA.condense == True: matA as above, and every application of the A-preconditioner goes through
 * harmonic_extension(f, blfA, k * preA_unscaled) (bramblepasciak_new.py:8-21).  Same outputs as nsb_bpcg mode 2. */
NSB_API int nsb_bpcg_condensed(nsb_ctx ctx, nsb_cond cond, nsb_mat B, nsb_mat BT, nsb_pre preA, nsb_pre preM, double k_scale,
                               const double* f, const double* g, double* sol, int maxsteps, double tol, int initialize,
                               int rel_err, double* history, int* updates, double* seconds, int* status);
/* EigenValues_Preconditioner(mat=matA, pre=preA_unscaled) with the condensed matA (bramblepasciak_new.py:104-115) */
NSB_API int nsb_lanczos_eigs_condensed(nsb_ctx ctx, nsb_cond cond, nsb_pre pre, const double* start, int maxsteps, double tol,
                                       double* lam_min, double* lam_max, int* steps);

/* Preconditioned CG with NGSolve CGSolver's stopping rule sqrt<Cr,r> <= tol * initial
 * (CGSolver(mstar, pre, precision=1e-4) NavierStokesSIMPLE_iterative.py:93,128; stands in for the direct
 * heat.mat.Inverse of heat.py:72).  history: absolute errors, capacity maxsteps+1. */
NSB_API int nsb_cg(nsb_ctx ctx, nsb_mat A, nsb_pre pre, const double* rhs, double* sol, int maxsteps, double tol,
           int initialize, double* history, int* iterations, double* seconds, int* status);

/* Extreme eigenvalues of pre*A by CG-Lanczos (EigenValues_Preconditioner(mat, pre, tol):
 * bramble_pasciak_cg.py:70, bramblepasciak_new.py:115).  start (length rows) may be NULL: the
 * deterministic vector 1 + 0.5 sin(i+1) is used.  Stops when lambda_min moved < tol relative. */
NSB_API int nsb_lanczos_eigs(nsb_ctx ctx, nsb_mat A, nsb_pre pre, const double* start, int maxsteps, double tol,
                             double* lam_min, double* lam_max, int* steps);

/* orthonormalize(basis, tries=3)   orthonormalization.py:5-16: modified Gram-Schmidt, `tries` sweeps,
 * in place.  basis = m vectors of length n stored one after another (stride n). */
NSB_API int nsb_orthonormalize(nsb_ctx ctx, double* basis, int64_t n, int m, int tries);

/* orthonormalize(basis, tries) as nsb_orthonormalize, for callers that must not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The call allocates nothing -- no cudaMalloc / cudaFree / stream-ordered allocation: its scratch (Gram-Schmidt state, per-block
 * partial sums, arrival ticket) is a file-scope device array of 96 KiB that every GPU loading the library holds until exit.
 * Limits (larger requests are refused with NSB_ERR_INVALID): m <= NSB_NOALLOC_MAX_VECTORS, n <= NSB_NOALLOC_MAX_LENGTH, basis
 * device-resident (m vectors of length n, stride n), single-GPU context.  All calls on a device share the array and it keeps the
 * last call's contents (the call clears what must start at zero): calls that overlap on the GPU race -- one call at a time per device. */
#define NSB_NOALLOC_MAX_VECTORS 8
#define NSB_NOALLOC_MAX_LENGTH ((int64_t)1 << 31)
NSB_API int nsb_orthonormalize_noalloc(nsb_ctx ctx, double* basis, int64_t n, int m, int tries);

/* Device-resident vector helpers for the time-stepper chains (DoTimeStep / Project,
 * templates/NavierStokesSIMPLE_iterative.py:427-443): all pointers must be device pointers.
 *   nsb_vec_axpby:        y = a x + b y          (temp.data += ..., gfu.vec.data += timestep * temp2.data, :432-438)
 *   nsb_vec_gather_axpy:  y[i] += a x[idx[i]]    (vel.data -= mapV * tmp_projection with mapV = PermutationMatrix, :138-144,:442) */
NSB_API int nsb_vec_axpby(nsb_ctx ctx, double* y, double a, const double* x, double b, int64_t n);
NSB_API int nsb_vec_gather_axpy(nsb_ctx ctx, double* y, double a, const double* x, const int32_t* idx, int64_t n);

/* One time step of the SIMPLE-type splitting, NavierStokes.DoTimeStep + NavierStokes.Project
 * (templates/NavierStokesSIMPLE_iterative.py:427-443), as ONE call on a handle that owns its work vectors:
 *     temp = conv + f - a gfu ; temp2 = invmstar temp ; Project(temp2[0:n_vel]) ; gfu += timestep temp2
 *     Project(vel): tmp = invproj (bproj vel) ; gfup = tmp[p_begin:p_end] ; vel[i] -= tmp[map_ind[i]]
 * with the operator compositions of :93-98 and :128-131
 *     invmstar = (I + mstar_H) CGSolver(mstar, pre_mstar, mstar_precision) (I + mstar_HT) + mstar_inner_solve
 *     invproj  = (I + aproj_H) CGSolver(aproj, pre_aproj, proj_precision)  (I + aproj_HT) + aproj_inner_solve
 * (the three condensation operators of a form may all be NULL: the CG solve alone).  a, mstar: X.ndof x X.ndof; bproj:
 * Xproj.ndof x n_vel; aproj and its condensation operators: Xproj.ndof x Xproj.ndof; map_ind (n_vel entries, host or device):
 * the index list of PermutationMatrix(Xproj.ndof, ind) (:138-144).  CGSolver semantics as nsb_cg (start from zero, stop at
 * sqrt<Cr,r> <= precision * initial or after maxsteps iterations; NGSolve's defaults are 1e-8 and 200).  Row-partitioned contexts:
 * every operator carries its own halo plan, vectors hold owned rows, map_ind indexes [owned | halo of aproj].  The handle borrows
 * all operators and preconditioners (they must outlive it). */
NSB_API int nsb_simple_create(nsb_ctx ctx, nsb_mat a, nsb_mat mstar, nsb_pre pre_mstar, nsb_mat mstar_H, nsb_mat mstar_HT,
                              nsb_mat mstar_inner_solve, nsb_mat bproj, nsb_mat aproj, nsb_pre pre_aproj, nsb_mat aproj_H,
                              nsb_mat aproj_HT, nsb_mat aproj_inner_solve, const int32_t* map_ind, int64_t n_vel, int64_t p_begin,
                              int64_t p_end, double timestep, double mstar_precision, int mstar_maxsteps, double proj_precision,
                              int proj_maxsteps, nsb_simple* out);
NSB_API int nsb_simple_destroy(nsb_simple s);
/* DoTimeStep (:427-438).  gfu (X.ndof; host or device) is updated in place; f (f.vec) and conv (the vector conv_operator * gfu.vec,
 * evaluated by the caller: the matrix-free convection form stays in NGSolve) may be NULL; gfup (p_end - p_begin entries) may be NULL.
 * cg_iterations[0..1]: iterations of the mstar and of the projection solve; seconds: the whole step by CUDA events. */
NSB_API int nsb_simple_step(nsb_ctx ctx, nsb_simple s, double* gfu, const double* f, const double* conv, double* gfup,
                            int* cg_iterations, double* seconds);
/* Project(vel) alone (:440-443; SolveInitial calls it on gfu.vec[0:V.ndof], :415,:426): vel (n_vel entries) corrected in place. */
NSB_API int nsb_simple_project(nsb_ctx ctx, nsb_simple s, double* vel, double* gfup, int* cg_iterations);

/* One pass of the time-loop body of heat.py:81-146 ("exponential integrator" in an m-dimensional subspace):
 * basis u, then m-1 implicit Euler sub-steps u -= dt H^-1 K u with dt = time_step/m (heat.py:95-98; H, the caller's
 * heat.mat = M + time_step K of heat.py:51-52, is solved by Jacobi-CG in place of the direct heat.mat.Inverse of heat.py:72), orthonormalize (:100), projected
 * m x m diffusion and mass matrices (:109-118), Gauss collocation step of the projected ODE on the host
 * (runge_kutta_method.py:26-59; rk_a is stages x stages row-major, rk_b has `stages` entries) and the lift back
 * (:140-142).  u (length rows; host or device) is replaced by the new temperature.  Khat_out / Mhat_out
 * (m*m, row-major, host) may be NULL. */
NSB_API int nsb_heat_step(nsb_ctx ctx, nsb_mat K, nsb_mat M, nsb_mat H, nsb_pre preH, double* u, double time_step, int m,
                          int stages, const double* rk_a, const double* rk_b, double cg_tol, int cg_maxsteps,
                          int* cg_iterations, double* Khat_out, double* Mhat_out);

#ifdef __cplusplus
}
#endif
#endif /* NSB200_H */
