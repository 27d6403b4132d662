/* krylov_omp.c -- CPU oracle, C + OpenMP restatement of the reference's Krylov hot path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs (and its parity objects, where this is the checker of the device result) may load this library; the
 * product (libnsb200.so) never does.
 *
 * Purpose: the reference runs its loops line by line through NGSolve, which executes every vector
 * expression and every sparse MultAdd as one multi-threaded primitive inside `with TaskManager():`
 * (run.py:239, NavierStokesSIMPLE_test.py:17).  NGSolve is absent here, so this file restates the loops the
 * same way -- one OpenMP-parallel primitive per reference line, no fusion -- to time "the reference's CPU
 * path on all host cores".  The numpy oracle (krylov_oracle.py) is the single-threaded twin; both are pinned
 * against tests/golden/*.npz, which come from the reference's own files run unmodified (make_golden.py).
 * Parity below the loops (NGSolve's own kernels) is unpinned: see DESIGN.md section 4.
 *
 * Matrices: CSR, int64 row pointers, int32 column indices, fp64 values.  Diagonal preconditioners.
 */
#include <math.h>
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int64_t rows, cols;
    const int64_t* rowptr;
    const int32_t* col;
    const double* val;
} csr_t;

int ko_num_threads(void) { return omp_get_max_threads(); }

/* y = M x   (NGSolve SparseMatrix::Mult; every `mat * vec` of the loops) */
static void spmv(const csr_t* M, const double* x, double* y) {
#pragma omp parallel for schedule(dynamic, 512)
    for (int64_t r = 0; r < M->rows; ++r) {
        double s = 0.0;
        for (int64_t p = M->rowptr[r]; p < M->rowptr[r + 1]; ++p) s += M->val[p] * x[M->col[p]];
        y[r] = s;
    }
}

void ko_spmv(int64_t rows, int64_t cols, const int64_t* rowptr, const int32_t* col, const double* val, const double* x,
             double* y) {
    csr_t M = {rows, cols, rowptr, col, val};
    spmv(&M, x, y);
}

/* InnerProduct.  Deterministic for a given thread count: static blocks, 8 interleaved partial sums per
 * block (what a SIMD ddot does; a single sequential accumulator loses enough digits on the ill-conditioned
 * small cases to shift late iteration counts), block results added in thread order. */
static double dot(const double* a, const double* b, int64_t n) {
    double part[256];
    int nt = 1;
#pragma omp parallel
    {
        const int t = omp_get_thread_num(), T = omp_get_num_threads();
#pragma omp single
        nt = T;
        const int64_t lo = n * t / T, hi = n * (t + 1) / T;
        double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        int64_t i = lo;
        for (; i + 8 <= hi; i += 8)
            for (int k = 0; k < 8; ++k) acc[k] += a[i + k] * b[i + k];
        for (int k = 0; i < hi; ++i, ++k) acc[k] += a[i] * b[i];
        if (t < 256) part[t] = ((acc[0] + acc[4]) + (acc[2] + acc[6])) + ((acc[1] + acc[5]) + (acc[3] + acc[7]));
    }
    double s = 0.0;
    for (int t = 0; t < nt && t < 256; ++t) s += part[t];
    return s;
}
/* y = a*x + b*y' style primitives: one pass per reference statement */
static void lin2(double* y, double a, const double* x1, double b, const double* x2, int64_t n) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) y[i] = a * x1[i] + b * x2[i];
}
static void lin3(double* y, double a, const double* x1, double b, const double* x2, double c, const double* x3, int64_t n) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) y[i] = a * x1[i] + b * x2[i] + c * x3[i];
}
static void axpy(double* y, double a, const double* x, int64_t n) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) y[i] += a * x[i];
}
static void scal_copy(double* y, double a, const double* x, int64_t n) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) y[i] = a * x[i];
}
static void diagmul(double* y, double s, const double* d, const double* x, int64_t n) { /* 'local' preconditioner */
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) y[i] = d ? s * (d[i] * x[i]) : s * x[i];
}
static double* vec(int64_t n) { return (double*)calloc((size_t)(n > 0 ? n : 1), sizeof(double)); }

/* ------------------------------------------------------------------------------------------------
 * BramblePasciakCG, solvers/bramblepasciak_new.py:24-253, condense = False.
 * history[i] = err printed at loop index i (:243-245); returns `it` (:253); *seconds = loop time (:195-196,251).
 * sol = (u0 | u1), mutated in place (:137). */
int ko_bpcg_v2(int64_t n_u, int64_t n_p, const int64_t* Ap, const int32_t* Ac, const double* Av, const int64_t* Bp,
               const int32_t* Bc, const double* Bv, const int64_t* BTp, const int32_t* BTc, const double* BTv,
               const double* dA, const double* dM, double k, const double* f, const double* g, double* sol, double tol,
               int maxsteps, int initialize, int rel_err, double* history, int* nhist, double* err0_out,
               double* seconds) {
    csr_t A = {n_u, n_u, Ap, Ac, Av}, B = {n_p, n_u, Bp, Bc, Bv}, BT = {n_u, n_p, BTp, BTc, BTv};
    double *u0 = sol, *u1 = sol + n_u;
    double *tmp0 = vec(n_u), *tmp1 = vec(n_u), *tmp2 = vec(n_u), *tmp4 = vec(n_u), *tmp3 = vec(n_p), *tmp5 = vec(n_p);
    double *f_new = vec(n_u), *g_new = vec(n_p), *d0 = vec(n_u), *d1 = vec(n_p), *w0 = vec(n_u), *w1 = vec(n_p);
    double *s0 = vec(n_u), *s1 = vec(n_p), *v0 = vec(n_u), *z0 = vec(n_u), *z_old0 = vec(n_u), *As0 = vec(n_u);
    double *Bs1 = vec(n_u), *pr0 = vec(n_u), *pr1 = vec(n_p), *pm = vec(n_p);
    diagmul(tmp0, k, dA, f, n_u);                         /* :128-129  tmp0 = preA f */
    spmv(&A, tmp0, f_new);                                /* :130 */
    axpy(f_new, -1.0, f, n_u);
    spmv(&B, tmp0, g_new);                                /* :133 */
    axpy(g_new, -1.0, g, n_p);
    if (initialize) memset(sol, 0, sizeof(double) * (size_t)(n_u + n_p)); /* :139 */
    spmv(&A, u0, tmp0);                                   /* :160 */
    spmv(&BT, u1, Bs1);
    axpy(tmp0, 1.0, Bs1, n_u);
    diagmul(tmp1, k, dA, tmp0, n_u);                      /* :162 */
    spmv(&A, tmp1, tmp2);                                 /* :163 */
    lin2(tmp4, 1.0, tmp1, -1.0, u0, n_u);                 /* :165 */
    spmv(&B, tmp4, tmp3);                                 /* :166 */
    lin3(d0, 1.0, f_new, -1.0, tmp2, 1.0, tmp0, n_u);     /* :169 */
    lin2(d1, 1.0, g_new, -1.0, tmp3, n_p);                /* :170 */
    diagmul(pr0, k, dA, f, n_u);                          /* :173 */
    spmv(&B, pr0, tmp5);                                  /* :177 */
    axpy(tmp5, -1.0, g, n_p);
    diagmul(pr1, 1.0, dM, tmp5, n_p);                     /* :179 */
    lin2(w0, 1.0, pr0, -1.0, tmp1, n_u);                  /* :182 */
    diagmul(pm, 1.0, dM, tmp3, n_p);                      /* :183 */
    lin2(w1, 1.0, pr1, -1.0, pm, n_p);
    double wdn = dot(w0, d0, n_u) + dot(w1, d1, n_p);     /* :185 */
    double err0 = sqrt(fabs(wdn));                        /* :187 */
    if (err0_out) *err0_out = err0;
    memcpy(s0, w0, sizeof(double) * (size_t)n_u);         /* :189 */
    memcpy(s1, w1, sizeof(double) * (size_t)n_p);
    int it = 0, nh = 0;
    double alpha = 0.0, beta = 0.0;
    double t0 = omp_get_wtime();
    if (wdn != 0.0) {                                     /* :191-192 */
        for (it = 0; it < maxsteps; ++it) {               /* :200 */
            if (it == 0) {
                spmv(&A, s0, As0);                        /* :202-203 */
                memcpy(z0, As0, sizeof(double) * (size_t)n_u);
            } else {
                lin3(As0, beta, As0, 1.0, z_old0, -alpha, tmp2, n_u); /* :205 */
            }
            spmv(&BT, s1, Bs1);                           /* :206 */
            lin2(tmp0, 1.0, As0, 1.0, Bs1, n_u);          /* :207 */
            diagmul(tmp1, k, dA, tmp0, n_u);              /* :209 */
            spmv(&A, tmp1, tmp2);                         /* :210 */
            lin2(tmp4, 1.0, tmp1, -1.0, s0, n_u);         /* :212 */
            spmv(&B, tmp4, tmp3);                         /* :213 */
            memcpy(z_old0, z0, sizeof(double) * (size_t)n_u); /* :215 */
            lin2(v0, 1.0, tmp2, -1.0, tmp0, n_u);         /* :218 */
            const double wd = wdn;                        /* :221 */
            const double as_s = dot(s0, v0, n_u) + dot(s1, tmp3, n_p); /* :222 */
            alpha = wd / as_s;                            /* :226 */
            axpy(u0, alpha, s0, n_u);                     /* :228 */
            axpy(u1, alpha, s1, n_p);
            axpy(d0, -alpha, v0, n_u);                    /* :229 */
            axpy(d1, -alpha, tmp3, n_p);
            axpy(w0, -alpha, tmp1, n_u);                  /* :232 */
            diagmul(pm, 1.0, dM, tmp3, n_p);              /* :233 */
            axpy(w1, -alpha, pm, n_p);
            wdn = dot(w0, d0, n_u) + dot(w1, d1, n_p);    /* :235 */
            beta = wdn / wd;                              /* :236 */
            axpy(z0, -alpha, tmp2, n_u);                  /* :238 */
            lin2(s0, beta, s0, 1.0, w0, n_u);             /* :240-241 */
            lin2(s1, beta, s1, 1.0, w1, n_p);
            const double err = sqrt(fabs(wd));            /* :243 */
            history[nh++] = err;
            if (err < tol * (rel_err ? err0 : 1.0)) break; /* :246 */
        }
        if (it == maxsteps) it = maxsteps - 1;            /* Python's loop variable after exhaustion */
    }
    if (seconds) *seconds = omp_get_wtime() - t0;
    if (nhist) *nhist = nh;
    double* all[] = {tmp0, tmp1, tmp2, tmp4, tmp3, tmp5, f_new, g_new, d0, d1, w0, w1, s0, s1, v0, z0, z_old0, As0, Bs1, pr0, pr1, pm};
    for (size_t i = 0; i < sizeof(all) / sizeof(all[0]); ++i) free(all[i]);
    return it;
}

/* ------------------------------------------------------------------------------------------------
 * MinRes, minres.py:12-149, on K = [[A, B^T], [B, 0]] (run.py:45) applied block-wise (NGSolve BlockMatrix)
 * with the diagonal preconditioner C = diag(dA, dM) (run.py:46).  errors[0] = 1.0, one entry per iteration.
 * Returns the number of iterations; *converged = left through the relative break (:126). */
static void apply_K(const csr_t* A, const csr_t* B, const csr_t* BT, const double* x, double* y, double* scratch) {
    const int64_t n_u = A->rows;
    spmv(A, x, y);
    spmv(BT, x + n_u, scratch);
    axpy(y, 1.0, scratch, n_u);
    spmv(B, x, y + n_u);
}

int ko_minres(int64_t n_u, int64_t n_p, const int64_t* Ap, const int32_t* Ac, const double* Av, const int64_t* Bp,
              const int32_t* Bc, const double* Bv, const int64_t* BTp, const int32_t* BTc, const double* BTv,
              const double* dC, const double* rhs, double* u, double tol, int maxsteps, int initialize, double* errors,
              int* converged, double* seconds) {
    csr_t A = {n_u, n_u, Ap, Ac, Av}, B = {n_p, n_u, Bp, Bc, Bv}, BT = {n_u, n_p, BTp, BTc, BTv};
    const int64_t n = n_u + n_p;
    double *v = vec(n), *v_old = vec(n), *v_new = vec(n), *w = vec(n), *w_old = vec(n), *w_new = vec(n);
    double *z = vec(n), *z_new = vec(n), *mz = vec(n), *scratch = vec(n_u);
    if (initialize) {                                      /* :62-66 */
        memset(u, 0, sizeof(double) * (size_t)n);
        memcpy(v, rhs, sizeof(double) * (size_t)n);
    } else {
        apply_K(&A, &B, &BT, u, mz, scratch);
        lin2(v, 1.0, rhs, -1.0, mz, n);
    }
    diagmul(z, 1.0, dC, v, n);                             /* :68 */
    double gamma = sqrt(dot(z, v, n)), gamma_new = 0.0;    /* :71 */
    scal_copy(z, 1.0 / gamma, z, n);                       /* :73-74 */
    scal_copy(v, 1.0 / gamma, v, n);
    double res = gamma, res_old = gamma;                   /* :76-78 */
    const double err0 = gamma;
    double eta_old = gamma, c_old = 1.0, c = 1.0, s_new = 0.0, s = 0.0, s_old = 0.0;
    int k = 1, ne = 0;
    errors[ne++] = 1.0;                                    /* :95 */
    *converged = 0;
    const double t0 = omp_get_wtime();
    while (k < maxsteps + 1 && res > tol) {                /* :96 */
        apply_K(&A, &B, &BT, z, mz, scratch);              /* :97 */
        const double delta = dot(mz, z, n);                /* :98 */
        lin3(v_new, 1.0, mz, -delta, v, -gamma, v_old, n); /* :99 */
        diagmul(z_new, 1.0, dC, v_new, n);                 /* :101 */
        gamma_new = sqrt(dot(z_new, v_new, n));            /* :103 */
        scal_copy(z_new, 1.0 / gamma_new, z_new, n);       /* :104-105 */
        scal_copy(v_new, 1.0 / gamma_new, v_new, n);
        const double a0 = c * delta - c_old * s * gamma;   /* :107-110 */
        const double a1 = sqrt(a0 * a0 + gamma_new * gamma_new);
        const double a2 = s * delta + c_old * c * gamma;
        const double a3 = s_old * gamma;
        const double c_new = a0 / a1;                      /* :112-113 */
        s_new = gamma_new / a1;
        lin3(w_new, 1.0, z, -a3, w_old, -a2, w, n);        /* :115 */
        scal_copy(w_new, 1.0 / a1, w_new, n);              /* :116 */
        axpy(u, c_new * eta_old, w_new, n);                /* :118 */
        const double eta = -s_new * eta_old;               /* :119 */
        res = fabs(s_new) * res_old;                       /* :122 */
        errors[ne++] = res / err0;                         /* :125 */
        if (res < tol * err0) {                            /* :126 */
            *converged = 1;
            break;
        }
        ++k;
        double* t;
        t = v_old; v_old = v; v = v_new; v_new = t;        /* :131-133 */
        t = w_old; w_old = w; w = w_new; w_new = t;
        t = z; z = z_new; z_new = t;
        eta_old = eta;
        s_old = s; s = s_new;
        c_old = c; c = c_new;
        gamma = gamma_new;
        res_old = res;
    }
    if (seconds) *seconds = omp_get_wtime() - t0;
    double* all[] = {v, v_old, v_new, w, w_old, w_new, z, z_new, mz, scratch};
    for (size_t i = 0; i < sizeof(all) / sizeof(all[0]); ++i) free(all[i]);
    return ne - 1;
}

/* ------------------------------------------------------------------------------------------------
 * Thread count.  torch.distributed.run exports OMP_NUM_THREADS=1 into every rank; the CPU baseline of bench.py
 * states its thread count and sets it here, whatever the launcher exported. */
void ko_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
}

/* ------------------------------------------------------------------------------------------------
 * NGSolve CGSolver as the reference uses it (templates/NavierStokesSIMPLE_iterative.py:93,128): preconditioned CG
 * from u = 0, err_k = sqrt<C r_k, r_k>, stop when err_k <= tol * err_0 or after maxsteps iterations; diagonal
 * preconditioner (the reference's bddc / h1amg are NGSolve-internal).  One threaded primitive per
 * statement of krylov_oracle.pcg.  Returns the number of iterations. */
static int pcg(const csr_t* A, const double* dinv, const double* rhs, double* u, double tol, int maxsteps, double* d, double* w,
               double* s, double* q) {
    const int64_t n = A->rows;
    memset(u, 0, sizeof(double) * (size_t)n);
    memcpy(d, rhs, sizeof(double) * (size_t)n);
    diagmul(w, 1.0, dinv, d, n);
    memcpy(s, w, sizeof(double) * (size_t)n);
    double wdn = dot(w, d, n);
    const double err0 = sqrt(fabs(wdn));
    if (err0 == 0.0) return 0;
    int it = 0;
    for (it = 1; it <= maxsteps; ++it) {
        spmv(A, s, q);
        const double wd = wdn;
        const double alpha = wd / dot(s, q, n);
        axpy(u, alpha, s, n);
        axpy(d, -alpha, q, n);
        diagmul(w, 1.0, dinv, d, n);
        wdn = dot(w, d, n);
        lin2(s, wdn / wd, s, 1.0, w, n);
        if (sqrt(fabs(wdn)) <= tol * err0) break;
    }
    return it > maxsteps ? maxsteps : it;
}

int ko_pcg(const csr_t* A, const double* dinv, const double* rhs, double* u, double tol, int maxsteps) {
    const int64_t n = A->rows;
    double *d = vec(n), *w = vec(n), *s = vec(n), *q = vec(n);
    const int it = pcg(A, dinv, rhs, u, tol, maxsteps, d, w, s, q);
    free(d); free(w); free(s); free(q);
    return it;
}

/* ext @ inv1 @ extT + inner_solve  (:94-98, :129-131) with inv1 = Jacobi-CG; H == NULL: inv1 alone.  y = result. */
static int extended_inverse(const csr_t* M, const double* dinv, const csr_t* H, const csr_t* HT, const csr_t* inner,
                            const double* x, double* y, double tol, int maxsteps, double* w[6]) {
    const int64_t n = M->rows;
    if (!H) return pcg(M, dinv, x, y, tol, maxsteps, w[0], w[1], w[2], w[3]);
    double *r = w[4], *t = w[5];
    spmv(HT, x, r);                                       /* r = (I + HT) x */
    axpy(r, 1.0, x, n);
    const int it = pcg(M, dinv, r, y, tol, maxsteps, w[0], w[1], w[2], w[3]);
    spmv(H, y, t);                                        /* y = (I + H) y */
    axpy(y, 1.0, t, n);
    spmv(inner, x, t);                                    /* y += inner_solve x */
    axpy(y, 1.0, t, n);
    return it;
}

/* NavierStokes.DoTimeStep + Project, templates/NavierStokesSIMPLE_iterative.py:427-443, one threaded primitive per
 * statement; gfu (length n_x) is updated in place, gfup receives tmp_projection[p_begin:p_end].  conv: the vector
 * conv_operator * gfu.vec (the matrix-free convection stays with the caller), may be NULL; f may be NULL.
 * ops = {a, mstar, mH, mHT, m_inner, bproj, aproj, pH, pHT, p_inner}; absent ones have rows < 0. */
int ko_simple_step(const csr_t* ops, const double* dm, const double* dp, const int32_t* mapV, int64_t n_vel, int64_t p_begin,
                   int64_t p_end, double dt, double prec_m, int max_m, double prec_p, int max_p, double* gfu, const double* f,
                   const double* conv, double* gfup, int* its, double* seconds) {
    const csr_t *a = &ops[0], *mstar = &ops[1], *mH = ops[2].rows >= 0 ? &ops[2] : NULL, *mHT = &ops[3], *minner = &ops[4];
    const csr_t *bproj = &ops[5], *aproj = &ops[6], *pH = ops[7].rows >= 0 ? &ops[7] : NULL, *pHT = &ops[8], *pinner = &ops[9];
    const int64_t n = a->rows, np_ = aproj->rows, nw = n > np_ ? n : np_;
    double* w[6];
    for (int i = 0; i < 6; ++i) w[i] = vec(nw);
    double *temp = vec(n), *temp2 = vec(n), *rp = vec(np_), *tmp = vec(np_);
    const double t0 = omp_get_wtime();
    spmv(a, gfu, temp);                                   /* :432-434  temp = conv + f - a gfu */
    scal_copy(temp, -1.0, temp, n);
    if (conv) axpy(temp, 1.0, conv, n);
    if (f) axpy(temp, 1.0, f, n);
    its[0] = extended_inverse(mstar, dm, mH, mHT, minner, temp, temp2, prec_m, max_m, w);   /* :436 */
    spmv(bproj, temp2, rp);                               /* :441 (bproj has n_vel columns: reads temp2[0:n_vel]) */
    its[1] = extended_inverse(aproj, dp, pH, pHT, pinner, rp, tmp, prec_p, max_p, w);
    if (gfup) memcpy(gfup, tmp + p_begin, sizeof(double) * (size_t)(p_end - p_begin));   /* :442 */
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n_vel; ++i) temp2[i] -= tmp[mapV[i]];                     /* :443 */
    axpy(gfu, dt, temp2, n);                              /* :438 */
    if (seconds) *seconds = omp_get_wtime() - t0;
    for (int i = 0; i < 6; ++i) free(w[i]);
    free(temp); free(temp2); free(rp); free(tmp);
    return 0;
}

/* orthonormalize(basis, tries), orthonormalization.py:5-16: modified Gram-Schmidt, m vectors of length n (stride n) */
void ko_orthonormalize(double* Q, int64_t n, int m, int tries) {
    for (int t = 0; t < tries; ++t)
        for (int j = 0; j < m; ++j) {
            double* bj = Q + (int64_t)j * n;
            for (int i = 0; i < j; ++i) {
                const double* bi = Q + (int64_t)i * n;
                axpy(bj, -(dot(bi, bj, n) / dot(bi, bi, n)), bi, n);   /* :11-12 */
            }
            scal_copy(bj, 1.0 / sqrt(dot(bj, bj, n)), bj, n);          /* :13 */
        }
}
